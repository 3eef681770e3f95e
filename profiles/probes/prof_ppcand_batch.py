"""Probe: the PointPillars candidate pass at batch 16 / 64 (62 / 247 MB of logits) - run under ncu (-k regex:pp_candidates)."""
import os
import sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from triton_client_b200 import synth  # noqa: E402
from triton_client_b200.gpu import Engine  # noqa: E402
KITTI = dict(anchor_range=[0, -39.68, -3, 69.12, 39.68, 1], anchor_sizes=[[3.9, 1.6, 1.56], [0.8, 0.6, 1.73], [1.76, 0.6, 1.73]],
             anchor_bottoms=[-1.78, -0.6, -0.6], anchor_rotations=[[0, 1.57]] * 3)
eng = Engine(0, workspace_bytes=2 << 30)
frames = [synth.pp_head_outputs(4100 + i) for i in range(2)]
for b in (16, 64):
    cls, box, dirp = (torch.from_numpy(np.stack([frames[i % 2][j] for i in range(b)])).cuda() for j in range(3))
    for _ in range(2):
        eng.pp_anchor_post(cls, box, dirp, **KITTI)
    torch.cuda.synchronize()
    print("batch", b, "logit bytes", cls.numel() * 4)
