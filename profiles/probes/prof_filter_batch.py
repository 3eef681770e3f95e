"""Probe: the YOLO streaming filter at batch 64 / 256 (45 / 181 MB of predictions) - run under ncu (-k regex:yolo_filter)."""
import os
import sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from triton_client_b200 import synth  # noqa: E402
from triton_client_b200.gpu import Engine  # noqa: E402
eng = Engine(0, workspace_bytes=2 << 30)
base = torch.from_numpy(synth.yolo_output(64, 25200, 2, 2002)).cuda()
for b in (64, 256):
    pred = base.repeat(b // 64, 1, 1).contiguous()
    for _ in range(2):
        eng.yolo_nms(pred, conf_thres=0.3)
    torch.cuda.synchronize()
    print("batch", b, "bytes", pred.numel() * 4)
