#!/usr/bin/env python
"""bench.py — frames/sec of the client pre+post hot path on B200 (BASELINE.json metric).

A step = one batch of the N=1 workload, BASELINE.json configs[1] ("YOLOv5nCROP batch 64"): 64 synthetic
1280x720 RGB frames letterboxed to 640x640 (u8 HWC -> f32 NCHW /255) + decode/filter/class-aware NMS of a
synthetic 64x25200x7 model output (conf 0.3 as the reference's callers pass, iou 0.45, max_det 300).

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this framework (CUDA, C ABI)
  python bench.py --impl reference [--steps K] [--warmup W]      # the reference's CPU path on host cores

Under torchrun (N>1) every rank runs its own shard of the frame stream (weak scaling, no collective
on the data path); one NCCL gather of the last batch's detections closes the timed region.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH, SRC_H, SRC_W, OUT, NROWS, NC = 64, 720, 1280, 640, 25200, 2
CONF, IOU, MAX_DET = 0.3, 0.45, 300
RING = 3  # distinct input batches cycled through so no step re-reads what the previous one left in L2
WORKLOAD = "configs[1]: YOLOv5nCROP batch 64, 1280x720 frames letterboxed to 640 + decode/class-aware NMS of 64x25200x7"


def ncu_traffic():
    """DRAM bytes per launch of the roofline kernel from the committed ncu capture (profiles/r01_pre_kernel.json)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r01_pre_kernel.json")))["dram_traffic_bytes_per_launch"]
    except Exception:
        return None


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """Polls SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.sm_max = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.005)

    def result(self):
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def make_inputs(rank):
    """RING distinct host batches (frames u8 NHWC + model output f32), seeded per SURVEY §8d."""
    from triton_client_b200 import synth
    frames, preds = [], []
    for r in range(RING):
        base = synth.image_smooth(SRC_H, SRC_W, synth.seed_for(2, 100 * rank + r))
        noise = synth.image_noise(SRC_H, SRC_W, synth.seed_for(2, 100 * rank + r + 50))
        f = np.empty((BATCH, SRC_H, SRC_W, 3), dtype=np.uint8)
        for i in range(BATCH):  # every frame distinct: shifted smooth scene, every 4th one pure noise
            f[i] = np.roll(noise if i % 4 == 3 else base, (13 * i, 29 * i), axis=(0, 1))
        frames.append(f)
        preds.append(synth.yolo_output(BATCH, NROWS, NC, synth.seed_for(2, 1000 + 100 * rank + r)))
    return frames, preds


def cpu_step_frames(frames, preds, idxs, as_written=True):
    """The reference's CPU path (oracle port) over the given frame indices: letterbox + image_adjust + extract_boxes."""
    from oracle import pre2d, yolo_post
    ndet = 0
    for i in idxs:
        x, _ = pre2d.letterbox(frames[i], OUT, OUT)
        t = pre2d.image_adjust_yolov5(x)
        assert t.shape == (1, 3, OUT, OUT)
        res = yolo_post.extract_boxes(yolo_post.FakeResponse([preds[i:i + 1]]), conf_thres=CONF, iou_thres=IOU,
                                      max_det=MAX_DET, as_written=as_written)
        ndet += sum(int(r.shape[0]) for r in res) if isinstance(res, list) else 0
    return ndet


def _cpu_worker(args):
    seed_rank, idxs, as_written = args
    import torch
    torch.set_num_threads(1)
    try:
        import cv2
        cv2.setNumThreads(1)
    except Exception:
        pass
    frames, preds = _CPU_DATA
    return cpu_step_frames(frames, preds, idxs, as_written)


_CPU_DATA = None


def cpu_baseline(sample_frames=8):
    """Oracle port, one process / one thread, on a bounded sample of the workload's frames."""
    global _CPU_DATA
    import torch
    from triton_client_b200 import synth
    torch.set_num_threads(1)
    try:
        import cv2
        cv2.setNumThreads(1)
    except Exception:
        pass
    frames = np.stack([synth.image_smooth(SRC_H, SRC_W, synth.seed_for(2, i)) for i in range(2)] * (sample_frames // 2))
    preds = synth.yolo_output(sample_frames, NROWS, NC, synth.seed_for(2, 7))
    cpu_step_frames(frames, preds, [0], True)  # warm-up
    t0 = time.perf_counter()
    cpu_step_frames(frames, preds, range(sample_frames), True)
    t_written = time.perf_counter() - t0
    t0 = time.perf_counter()
    cpu_step_frames(frames, preds, range(sample_frames), False)
    t_fast = time.perf_counter() - t0
    return {"value": sample_frames / t_written, "unit": "frames/s", "cores": 1, "kind": "port",
            "sample": f"{sample_frames} frames of the workload (letterbox+image_adjust+extract_boxes, struct.unpack wire parsing as the reference writes it)",
            "value_without_wire_parsing": sample_frames / t_fast}


def run_reference(args):
    """--impl reference: the reference's CPU implementation (oracle port; the reference is Python and
    cannot travel to the GPU box) on all host cores, one process per core, frames sharded."""
    global _CPU_DATA
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    from triton_client_b200 import synth
    cores = os.cpu_count() or 1
    # frames per step: a bounded sample of the 64-frame batch, sized so K steps end within ~2 minutes
    # (the as-written path costs ~0.2 s/frame/core)
    budget_frames = int(120.0 / 0.2 * cores / max(args.steps + args.warmup, 1))
    per_step = max(1, min(BATCH, budget_frames))
    frames = np.stack([synth.image_smooth(SRC_H, SRC_W, synth.seed_for(2, i % 4)) for i in range(per_step)])
    preds = synth.yolo_output(per_step, NROWS, NC, synth.seed_for(2, 7))
    _CPU_DATA = (frames, preds)
    nproc = min(cores, per_step)
    shards = [list(range(i, per_step, nproc)) for i in range(nproc)]
    ctx = mp.get_context("fork")
    with ctx.Pool(nproc) as pool:
        for _ in range(args.warmup):
            pool.map(_cpu_worker, [(0, s[:1], True) for s in shards])
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_cpu_worker, [(0, s, True) for s in shards])
        dt = time.perf_counter() - t0
    fps = per_step * args.steps / dt
    line = {"impl": "reference", "metric": "frames/sec pre+post", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": f"{per_step} of the batch's 64 frames per step"},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": nproc, "kind": "port",
                             "sample": f"{per_step} frames/step, {nproc} processes, reference path as written (struct.unpack wire parsing)"},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def latency_c1(eng, dev, frames=1000):
    """p50 per-frame latency at batch 1 (BASELINE.json configs[0]: 640x640 frame -> image_adjust, 1x25200x85 -> decode/NMS),
    device-resident inputs cycled over 8 distinct frames, one CUDA-event pair per frame and stage; also replayed from a
    CUDA graph of the two calls (launch overhead removed)."""
    import torch
    from triton_client_b200 import synth
    imgs = [torch.from_numpy(synth.image_smooth(640, 640, 900 + i)).to(dev).unsqueeze(0) for i in range(8)]
    preds = [torch.from_numpy(synth.yolo_output(1, NROWS, 80, 950 + i)).to(dev) for i in range(8)]
    out = torch.empty((1, 3, 640, 640), dtype=torch.float32, device=dev)
    det = torch.zeros((1, MAX_DET, 6), dtype=torch.float64, device=dev)
    cnt = torch.zeros((2, 1), dtype=torch.int32, device=dev)

    def one(i):
        eng.preprocess_images(imgs[i % 8], 640, 640, mode="stretch", div=255.0, out=out)
        eng.yolo_nms(preds[i % 8], conf_thres=CONF, iou_thres=IOU, max_det=MAX_DET, out=det, counts=cnt)

    for i in range(20):
        one(i)
    torch.cuda.synchronize()
    pre_t, post_t, tot_t = [], [], []
    for i in range(frames):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        eng.preprocess_images(imgs[i % 8], 640, 640, mode="stretch", div=255.0, out=out)
        e[1].record()
        eng.yolo_nms(preds[i % 8], conf_thres=CONF, iou_thres=IOU, max_det=MAX_DET, out=det, counts=cnt)
        e[2].record()
        e[2].synchronize()
        pre_t.append(e[0].elapsed_time(e[1]))
        post_t.append(e[1].elapsed_time(e[2]))
        tot_t.append(e[0].elapsed_time(e[2]))
    res = {"config": "configs[0]: YOLOv5n COCO 640x640 batch 1 (image_adjust + decode/NMS of 1x25200x85)", "frames": frames,
           "pre_ms": statistics.median(pre_t), "post_ms": statistics.median(post_t), "pre_post_ms": statistics.median(tot_t),
           "graph_pre_post_ms": None}
    try:   # the same two calls captured once and replayed: what a steady-state caller pays without launch overhead
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream(dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):
            one(0)
        torch.cuda.current_stream(dev).wait_stream(s)
        with torch.cuda.graph(g, stream=s):
            one(0)
        torch.cuda.synchronize()
        ts = []
        for i in range(frames):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            g.replay()
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b))
        res["graph_pre_post_ms"] = statistics.median(ts)
    except Exception as ex:  # graph capture is an extra, never a reason to lose the bench line
        res["graph_error"] = str(ex)[:200]
    return res


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from triton_client_b200.gpu import Engine
    from triton_client_b200.clients import Yolov5client

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from triton_client_b200.shard import bind_to_gpu_numa_node
    numa = bind_to_gpu_numa_node(local)      # before any pinned buffer is allocated
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    eng = Engine(local, workspace_bytes=256 << 20)
    frames_h, preds_h = make_inputs(rank)
    frames_d = [torch.from_numpy(f).to(dev) for f in frames_h]
    preds_d = [torch.from_numpy(p).to(dev) for p in preds_h]
    pre_out = torch.empty((BATCH, 3, OUT, OUT), dtype=torch.float32, device=dev)
    det = torch.zeros((BATCH, MAX_DET, 6), dtype=torch.float64, device=dev)
    cnt = torch.zeros((2, BATCH), dtype=torch.int32, device=dev)

    def step(i, ev=None):
        r = i % RING
        if ev:
            ev[0].record()
        eng.preprocess_images(frames_d[r], OUT, OUT, mode="letterbox", div=255.0, pad_value=114, out=pre_out)
        if ev:
            ev[1].record()
        eng.yolo_nms(preds_d[r], conf_thres=CONF, iou_thres=IOU, max_det=MAX_DET, out=det, counts=cnt)
        if ev:
            ev[2].record()

    def gather_results():
        """Final result collection: counts + padded detections of the last batch to every rank (KBs)."""
        if world == 1:
            return
        c = [torch.empty_like(cnt[0]) for _ in range(world)]
        dist.all_gather(c, cnt[0].contiguous())
        d = [torch.empty_like(det) for _ in range(world)]
        dist.all_gather(d, det)

    for i in range(max(args.warmup, 3)):
        step(i)
    gather_results()
    torch.cuda.synchronize()

    # ---- device-resident throughput (inputs already in HBM) ------------------------------------------
    # per-kernel events on every `--event-every`-th step only: an event record between two kernels is a stream operation
    # of its own and takes away the programmatic launch edge between them
    ev_every = max(1, args.event_every)
    evs = {i: [torch.cuda.Event(enable_timing=True) for _ in range(3)] for i in range(min(ev_every // 2, args.steps - 1), args.steps, ev_every)}
    sampler = ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = eng.launches
    sampler.start()
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_start.record()
    for i in range(args.steps):
        step(i, evs.get(i))
    gather_results()
    t_end.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler.stop_flag = True
    sampler.join()
    launches = eng.launches - l0
    ms = torch.tensor([t_start.elapsed_time(t_end)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    pre_ms = statistics.mean(e[0].elapsed_time(e[1]) for e in evs.values())
    post_ms = statistics.mean(e[1].elapsed_time(e[2]) for e in evs.values())

    # ---- the same step with the two halves on two streams (what a frame pipeline does: the preprocess of the next
    #      batch runs while the previous batch's model output is post-processed). Extra figure; `value` stays sequential
    #      so that the roofline kernel is timed alone. Two contexts: a context is one frame in flight. ----------------
    eng2 = Engine(local, workspace_bytes=256 << 20)
    s_pre, s_post = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def step2(i):
        r = i % RING
        with torch.cuda.stream(s_post):
            eng2.yolo_nms(preds_d[r], conf_thres=CONF, iou_thres=IOU, max_det=MAX_DET, out=det, counts=cnt)
        with torch.cuda.stream(s_pre):
            eng.preprocess_images(frames_d[r], OUT, OUT, mode="letterbox", div=255.0, pad_value=114, out=pre_out)

    cur = torch.cuda.current_stream(dev)
    for i in range(3):
        step2(i)
    torch.cuda.synchronize()
    o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    o0.record()
    s_pre.wait_event(o0)
    s_post.wait_event(o0)
    for i in range(args.steps):
        step2(i)
    cur.wait_stream(s_pre)
    cur.wait_stream(s_post)
    o1.record()
    torch.cuda.synchronize()
    ov = torch.tensor([o0.elapsed_time(o1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ov, op=dist.ReduceOp.MAX)
    overlap_ms = float(ov.item())

    # ---- end to end through the drop-in client classes: host buffers in, host results out ------------
    client = Yolov5client(engine=eng)
    pre, post = client.get_preprocess(), client.get_postprocess()
    from triton_client_b200.clients.wire import HostResponse
    pin_frames = [torch.from_numpy(f).pin_memory() for f in frames_h]
    pin_preds = [torch.from_numpy(p).pin_memory() for p in preds_h]
    e2e_steps = max(3, min(args.steps, 20))

    def e2e_issue(i):
        r = i % RING
        # both halves of a step are queued before either is waited for: the response's upload follows the frames' uploads on
        # the link while the preprocess download is still running
        hx = pre.image_adjust_batch_async(pin_frames[r], (OUT, OUT), mode="letterbox", copy=False)  # H2D + kernel + D2H (host f32 NCHW)
        pending = post.extract_boxes_async(HostResponse([pin_preds[r]]), conf_thres=CONF, iou_thres=IOU)  # H2D + kernels + D2H
        return hx, pending

    def e2e_collect(h):
        dets = h[1].result()
        x = h[0].result()
        assert isinstance(dets, list) and len(dets) == BATCH and x.shape == (BATCH, 3, OUT, OUT), dets
        return x, dets

    def e2e_run(k):
        """k steps, software-pipelined by one: step i+1 is queued before step i's host results are collected, so the
        link never idles between steps; every step's inputs go up and its results come down inside the timed region."""
        prev = None
        for i in range(k):
            cur_h = e2e_issue(i)
            if prev is not None:
                e2e_collect(prev)
            prev = cur_h
        return e2e_collect(prev)

    e2e_run(3)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    x, dets = e2e_run(e2e_steps)
    torch.cuda.synchronize()
    e2e_wall = time.perf_counter() - t0
    # the pipelined host results are the device-resident ones (checked outside the timed region)
    r_last = (e2e_steps - 1) % RING
    eng.preprocess_images(frames_d[r_last], OUT, OUT, mode="letterbox", div=255.0, pad_value=114, out=pre_out)
    eng.yolo_nms(preds_d[r_last], conf_thres=CONF, iou_thres=IOU, max_det=MAX_DET, out=det, counts=cnt)
    torch.cuda.synchronize()
    assert torch.equal(torch.from_numpy(x).to(dev), pre_out), "e2e preprocess result differs from the device-resident one"
    assert [len(d) for d in dets] == cnt[0].tolist() and all(torch.equal(d.to(dev), det[b, :len(d)]) for b, d in enumerate(dets)), "e2e detections differ"
    e2e_s = torch.tensor([e2e_wall], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_fps = BATCH * world * e2e_steps / float(e2e_s.item())
    h2d = frames_h[0].nbytes + preds_h[0].nbytes
    d2h = BATCH * 3 * OUT * OUT * 4 + det.numel() * 8 + cnt.numel() * 4

    latency = latency_c1(eng, dev) if rank == 0 else None
    if rank == 0:
        peak, which = peaks()
        # letterboxed C2 frame: the kernel reads the 3*h*w source once and writes 12*S*S output once
        pre_bytes = BATCH * (3 * SRC_H * SRC_W + 12 * OUT * OUT)
        achieved = pre_bytes / (pre_ms * 1e-3) / 1e9
        line = {
            "metric": "frames/sec pre+post", "value": BATCH * world * args.steps / (total_ms * 1e-3), "unit": "frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8->f32 (pre), f64 (NMS)",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_per_gpu": BATCH, "conf_thres": CONF, "iou_thres": IOU, "max_det": MAX_DET,
                       "l2": f"inputs cycle through {RING} distinct batches of 222 MB each (> 126 MB L2); no flush needed",
                       "parallelism": f"frame stream sharded over {world} rank(s), no data-path collective", "host_numa": numa},
            "kernel_ms": {"preprocess": pre_ms, "yolo_post": post_ms, "timed_on_steps": len(evs)},
            "roofline": {"bound": "hbm", "kernel": "preprocess2_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "peak_source": which + " copy bandwidth (MEASURED_PEAKS.json hbm_gbs)",
                         "bytes_per_launch": pre_bytes, "traffic": ncu_traffic()},
            "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "api": "Yolov5preprocess.image_adjust_batch_async + Yolov5postprocess.extract_boxes_async + .result(), two steps in flight"},
            "gpu_launches": launches, "clocks": sampler.result(), "p50_latency": latency,
            "two_streams": {"value": BATCH * world * args.steps / (overlap_ms * 1e-3), "unit": "frames/s", "ms_per_step": overlap_ms / args.steps,
                            "what": "the same steps with preprocess and YOLO post on two streams (two contexts), device-resident"},
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--event-every", type=int, default=8, help="record the per-kernel CUDA events on every n-th timed step")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps == 200:
            args.steps = 3
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
