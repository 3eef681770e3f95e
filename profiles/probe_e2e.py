"""Where the end-to-end step of bench.py spends its time: raw PCIe copies of the step's bytes, alone and together, then
the two drop-in calls alone. Run on the GPU box: python profiles/probe_e2e.py"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from triton_client_b200 import synth  # noqa: E402
from triton_client_b200.clients import Yolov5client  # noqa: E402
from triton_client_b200.clients.wire import HostResponse  # noqa: E402
from triton_client_b200.gpu import Engine  # noqa: E402

B, H, W, S = 64, 720, 1280, 640


def timed(fn, n=10):
    fn(); fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / n * 1e3


def main():
    dev = torch.device("cuda", 0)
    up_h = torch.empty(B * H * W * 3, dtype=torch.uint8).pin_memory()
    dn_h = torch.empty(B * 3 * S * S, dtype=torch.float32).pin_memory()
    up_d, dn_d = torch.empty_like(up_h, device=dev), torch.empty_like(dn_h, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def h2d():
        up_d.copy_(up_h, non_blocking=True)

    def d2h():
        dn_h.copy_(dn_d, non_blocking=True)

    def both():
        with torch.cuda.stream(s1):
            up_d.copy_(up_h, non_blocking=True)
        with torch.cuda.stream(s2):
            dn_h.copy_(dn_d, non_blocking=True)

    a, b, c = timed(h2d), timed(d2h), timed(both)
    print(f"H2D {up_h.numel() / 1e6:.0f} MB: {a:.3f} ms = {up_h.numel() / a / 1e6:.1f} GB/s")
    print(f"D2H {dn_h.numel() * 4 / 1e6:.0f} MB: {b:.3f} ms = {dn_h.numel() * 4 / b / 1e6:.1f} GB/s")
    print(f"both directions at once: {c:.3f} ms")

    eng = Engine(0, workspace_bytes=512 << 20)
    client = Yolov5client(engine=eng)
    pre, post = client.get_preprocess(), client.get_postprocess()
    frames = torch.from_numpy(np.stack([synth.image_noise(H, W, k) for k in range(B)])).pin_memory()
    preds = torch.from_numpy(synth.yolo_output(B, 25200, 2, 5)).pin_memory()
    for ch in (4, 8, 16, 32):
        t = timed(lambda: pre.image_adjust_batch(frames, (S, S), mode="letterbox", copy=False, chunks=ch))
        print(f"image_adjust_batch chunks={ch}: {t:.3f} ms")
    t = timed(lambda: post.extract_boxes(HostResponse([preds]), conf_thres=0.3, iou_thres=0.45))
    print(f"extract_boxes: {t:.3f} ms")


if __name__ == "__main__":
    main()
