"""ncu target: one batched step of the detectron / PointPillars / CenterPoint workloads of bench.py."""
import os
import sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from triton_client_b200.gpu import Engine  # noqa: E402
dev = torch.device("cuda", 0)
for key in (sys.argv[1:] or ["detectron", "pointpillars", "centerpoint"]):
    eng = Engine(0, workspace_bytes=3 << 30)
    wl = bench.make_workload(key, eng, dev, 0, 1)
    for i in range(3):
        wl.step(i)
    torch.cuda.synchronize()
    del wl, eng
    torch.cuda.empty_cache()
