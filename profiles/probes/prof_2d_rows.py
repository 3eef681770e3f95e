"""Workload for an ncu launch list of the detectron / YOLOv4 rows (same inputs as profiles/bench_rows.py)."""
import sys
import numpy as np
import torch
sys.path.insert(0, '/root/repo')
from oracle import detectron_post
from triton_client_b200 import synth
from triton_client_b200.gpu import Engine
eng = Engine(0, workspace_bytes=1 << 30)
lg, dl = synth.detectron_head_outputs(7)
lgd, dld = [torch.from_numpy(t).cuda() for t in lg], [torch.from_numpy(t).cuda() for t in dl]
cells = detectron_post.retinanet_cell_anchors()
for _ in range(3):
    eng.detectron_post(lgd, dld, (8, 16, 32, 64, 128), cells, (800, 1344))
rng = np.random.default_rng(3)
heads = [rng.normal(0, 1.5, (1, 255, s, s)).astype(np.float32) for s in (64, 32, 16)]
anchors = [12, 16, 19, 36, 40, 28, 36, 75, 76, 55, 72, 146, 142, 110, 192, 243, 459, 401]
per = [[a / st for a in anchors[6 * i:6 * i + 6]] for i, st in enumerate((8, 16, 32))]
hd = [torch.from_numpy(h).cuda() for h in heads]
for _ in range(3):
    b4, c4 = eng.yolo_head_decode_v4(hd, per, 80)
    eng.yolo4_post(b4, c4, 0.4, 0.6)
torch.cuda.synchronize()
