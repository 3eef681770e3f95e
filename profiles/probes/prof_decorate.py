"""ncu target: one batched decoration (and voxelisation) launch per config. Usage: ncu ... python profiles/probes/prof_decorate.py"""
import os
import sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from triton_client_b200 import synth  # noqa: E402
from triton_client_b200.gpu import Engine  # noqa: E402
eng = Engine(0, workspace_bytes=1 << 30)
KITTI = ([0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4])
NUSC = ([-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8])
for cfg, p, mv, mk, ov in [(KITTI, 32, 12000, lambda s: synth.lidar_sweep(synth.seed_for(4, s)), "continue"),
                           (NUSC, 20, 60000, lambda s: synth.lidar_multisweep(synth.seed_for(5, s)), "break")]:
    sweeps = [mk(s % 4) for s in range(16)]
    offs = np.concatenate([[0], np.cumsum([len(s) for s in sweeps])]).astype(int).tolist()
    pts = torch.from_numpy(np.concatenate(sweeps)).cuda()
    for _ in range(2):
        v, c, n, m = eng.voxelize(pts, cfg[0], cfg[1], p, mv, overflow=ov, offsets=offs)
        f = eng.pillar_decorate(v, c, n, cfg[1], cfg[0], m=m)
    torch.cuda.synchronize()
