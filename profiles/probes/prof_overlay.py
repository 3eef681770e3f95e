"""Workload for ncu: the overlay row of profiles/bench_rows.py (64 frames of 1280x720, detections of the YOLO post)."""
import sys
import torch
sys.path.insert(0, '/root/repo')
from triton_client_b200 import synth
from triton_client_b200.gpu import Engine
eng = Engine(0, workspace_bytes=1 << 30)
pred = torch.from_numpy(synth.yolo_output(64, 25200, 2, 5)).cuda()
det, cnt, _ = eng.yolo_nms(pred, conf_thres=0.3)
frames = torch.from_numpy(synth.image_smooth(720, 1280, 3)).cuda().unsqueeze(0).repeat(64, 1, 1, 1).contiguous()
for _ in range(3):
    eng.draw_detections(frames, det, cnt, ["crop", "weed"], (0, 255, 0), [640, 640], math_mode=1)
torch.cuda.synchronize()
print("detections", int(cnt.sum()))
for kw in (dict(), dict(font_scale=1.2, text_thickness=1, text_dy=0, thickness=1, labels="name"), dict(labels=False)):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(20):
        eng.draw_detections(frames, det, cnt, ["crop", "weed"], (0, 255, 0), [640, 640], math_mode=1, **kw)
    ev[1].record()
    torch.cuda.synchronize()
    print(kw, "ms per call of 64 frames:", ev[0].elapsed_time(ev[1]) / 20)
