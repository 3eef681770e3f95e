"""Probe: batched JPEG decode of 64 camera frames (1280x720, q90, 4:2:0), CUDA events + host time; under ncu for the kernel list."""
import os
import sys
import time
import cv2
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from triton_client_b200 import synth  # noqa: E402
from triton_client_b200.gpu import Engine  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
eng = Engine(0, workspace_bytes=2 << 30)
for q, rst in ((90, 0), (90, 80), (50, 0)):
    frames = [np.roll(synth.image_smooth(720, 1280, 7300 + i % 3), 31 * i, axis=1) for i in range(64)]
    streams = [cv2.imencode(".jpg", f, [cv2.IMWRITE_JPEG_QUALITY, q] + ([cv2.IMWRITE_JPEG_RST_INTERVAL, rst] if rst else []))[1].tobytes() for f in frames]
    out = eng.jpeg_decode(streams)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for _ in range(reps):
        eng.jpeg_decode(streams, out=out)
    b.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / reps
    t0 = time.perf_counter()
    for s in streams[:8]:
        cv2.imdecode(np.frombuffer(s, np.uint8), cv2.IMREAD_ANYCOLOR)
    cpu = (time.perf_counter() - t0) / 8
    kb = sum(len(s) for s in streams) / 64 / 1024
    print(f"q{q} rst{rst}: {kb:.0f} KB/frame; GPU {a.elapsed_time(b)/reps:.3f} ms per 64 frames (wall {wall*1e3:.3f} ms) = {64/wall:.0f} frames/s; cv2.imdecode {cpu*1e3:.2f} ms/frame on one core", flush=True)
