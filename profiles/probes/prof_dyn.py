"""Probe: DynamicPillarVFE decoration (tcb_dynamic_pillar) on C5 clouds at batch 1 and 16, CUDA events (queued back to back)."""
import os
import sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from triton_client_b200 import synth  # noqa: E402
from triton_client_b200.gpu import Engine  # noqa: E402
eng = Engine(0, workspace_bytes=2 << 30)
NUSC = ([-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8])
clouds = [synth.lidar_multisweep(synth.seed_for(5, r)) for r in range(4)]
for b in (1, 16):
    rows = np.concatenate([np.concatenate([np.full((len(clouds[s % 4]), 1), s, np.float32), clouds[s % 4]], 1) for s in range(b)])
    pts = torch.from_numpy(rows).cuda()
    n, ncol = pts.shape
    for _ in range(3):
        eng.dynamic_pillar(pts, NUSC[0], NUSC[1], batch_size=b)
    torch.cuda.synchronize()
    reps = 30
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        eng.dynamic_pillar(pts, NUSC[0], NUSC[1], batch_size=b)
    e.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(e) / reps
    alg = n * (ncol * 4 + (ncol - 1 + 6) * 4 + 8)
    print(f"dynamic pillar b={b} n={n}: {ms*1e3:.1f} us, {alg/1e6:.1f} MB algorithmic, {alg/ms/1e6:.0f} GB/s ({alg/ms/1e6/6556.5:.3f})")
