import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from triton_client_b200 import synth
from triton_client_b200.gpu import Engine
eng = Engine(0, workspace_bytes=1 << 30)
PP = dict(anchor_range=[0, -39.68, -3, 69.12, 39.68, 1], anchor_sizes=[[3.9, 1.6, 1.56], [0.8, 0.6, 1.73], [1.76, 0.6, 1.73]],
          anchor_bottoms=[-1.78, -0.6, -0.6], anchor_rotations=[[0, 1.57]] * 3)
x = [torch.from_numpy(t).cuda() for t in synth.pp_head_outputs(4000)]
for _ in range(3):
    out = eng.pp_anchor_post(*x, **PP)
torch.cuda.synchronize()
print("kept", int(out[3]))
heads = synth.centerpoint_head_outputs(5000)
hd = [{k: torch.from_numpy(t).cuda() for k, t in h.items()} for h in heads]
CLASS_MAP = [[0], [1, 2], [3, 4], [5], [6, 7], [8, 9]]
for _ in range(2):
    eng.centerpoint_post(hd, CLASS_MAP, 4, [0.2, 0.2, 8], [-51.2, -51.2], [-61.2, -61.2, -10, 61.2, 61.2, 10])
pts = torch.from_numpy(synth.lidar_sweep(11)).cuda()
for _ in range(2):
    eng.voxelize(pts, [0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4], 32, 40000, range_mask=True)
torch.cuda.synchronize()
