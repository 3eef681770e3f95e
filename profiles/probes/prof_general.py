"""Workload for ncu: the reference's own resize, 64 frames of 1280x720 stretched to 640x640 (row a1-a3)."""
import sys
import numpy as np
import torch
sys.path.insert(0, '/root/repo')
from triton_client_b200 import synth
from triton_client_b200.gpu import Engine
eng = Engine(0)
base = synth.image_smooth(1080, 1920, 1)
frames = torch.from_numpy(np.stack([np.roll(base, 17 * i, axis=1) for i in range(32)])).cuda()
out = torch.empty((32, 3, 640, 640), dtype=torch.float32, device="cuda")
for _ in range(4):
    eng.preprocess_images(frames, 640, 640, mode="stretch", out=out)
torch.cuda.synchronize()
