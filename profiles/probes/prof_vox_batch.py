"""Probe: batched voxelisation + decoration (C4 KITTI pillars, C5 nuScenes pillars) at batch 1/16, CUDA events.
Usage: python profiles/probes/prof_vox_batch.py [reps]   (under ncu for the per-kernel list)"""
import os
import sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from triton_client_b200 import synth  # noqa: E402
from triton_client_b200.gpu import Engine  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
eng = Engine(0, workspace_bytes=1 << 30)
KITTI = ([0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4])
NUSC = ([-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8])


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


for name, cfg, p, mv, mk in [("C4", KITTI, 32, 12000, lambda s: synth.lidar_sweep(synth.seed_for(4, s))),
                             ("C5", NUSC, 20, 60000, lambda s: synth.lidar_multisweep(synth.seed_for(5, s)))]:
    for b in (1, 16):
        sweeps = [mk(s) for s in range(min(b, 4))]
        sweeps = [sweeps[i % len(sweeps)] for i in range(b)]
        offs = np.concatenate([[0], np.cumsum([len(s) for s in sweeps])]).astype(int).tolist()
        pts = torch.from_numpy(np.concatenate(sweeps)).cuda()
        f = pts.shape[1]
        res = {}

        def vox():
            res["v"] = eng.voxelize(pts, cfg[0], cfg[1], p, mv, overflow="continue" if name == "C4" else "break", offsets=offs)
        t_v = timed(vox, reps)
        v, c, n, m = res["v"]
        M = m.cpu().numpy()
        bytes_v = 4 * pts.shape[0] * f + int(M.sum()) * (4 * p * f + 20)
        t_d = timed(lambda: eng.pillar_decorate(v, c, n, cfg[1], cfg[0], m=m), reps)
        bytes_d = int(M.sum()) * (4 * p * f + 20 + 4 * p * (f + 6))
        print(f"{name} b={b:2d} n={pts.shape[0]} M={M[:3].tolist()} vox {t_v*1e3:8.1f} us  {bytes_v/t_v/1e6:7.1f} GB/s ({bytes_v/t_v/1e6/6556.5:.3f}) | "
              f"decorate {t_d*1e3:7.1f} us {bytes_d/t_d/1e6:7.1f} GB/s ({bytes_d/t_d/1e6/6556.5:.3f})", flush=True)
