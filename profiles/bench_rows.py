#!/usr/bin/env python
"""Per-row measurement of the hot path: for every BASELINE.json config, the device time of each stage
(CUDA events, inputs resident in HBM, distinct input sets cycled so nothing is served from L2), the algorithmic
bytes of the stage, the achieved GB/s against MEASURED_PEAKS.json, and the oracle's CPU time for the same stage on
the same synthetic input (one process; the reported baseline, not the target).

    python profiles/bench_rows.py [--iters 30] [--no-cpu] [--only substring]

Prints one JSON line per row and a markdown table at the end (profiles/r01_rows.md is a committed copy).
Also used, with --iters 2 --no-cpu, as the workload for the ncu launch list."""
import argparse
import json
import os
import statistics
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from triton_client_b200 import synth  # noqa: E402
from triton_client_b200.gpu import Engine  # noqa: E402

KITTI = ([0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4])
SHIPPED = ([0, -40, -3, 70.4, 40, 1], [0.05, 0.05, 0.1])
NUSC = ([-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8])
PP = dict(anchor_range=[0, -39.68, -3, 69.12, 39.68, 1], anchor_sizes=[[3.9, 1.6, 1.56], [0.8, 0.6, 1.73], [1.76, 0.6, 1.73]],
          anchor_bottoms=[-1.78, -0.6, -0.6], anchor_rotations=[[0, 1.57]] * 3)
CLASS_MAP = [[0], [1, 2], [3, 4], [5], [6, 7], [8, 9]]
RING = 3


def peak():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        return 6650.0


REP = 8  # launches per timed sample, enqueued back to back so that host launch latency is not what gets timed


def gpu_time(fn, iters):
    """fn(i) launches the stage on input set i % RING; returns (p50 ms, mean ms) per launch."""
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    ts = []
    for i in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn(i)                      # keeps the queue busy while the timed launches are being enqueued
        a.record()
        for k in range(REP):
            fn(i * REP + k + 1)
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b) / REP)
    return statistics.median(ts), statistics.mean(ts)


def cpu_time(fn, reps=3):
    fn()
    ts = []
    for _ in range(reps):
        t = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t)
    return 1e3 * statistics.median(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    eng = Engine(0, workspace_bytes=1 << 30)
    pk = peak()
    rows = []

    def row(name, config, gpu_fn, bytes_alg, frames, cpu_fn=None, note=""):
        if args.only and args.only not in name:
            return
        p50, mean = gpu_time(gpu_fn, args.iters)
        r = {"row": name, "config": config, "gpu_ms_p50": round(p50, 4), "gpu_ms_mean": round(mean, 4), "frames_per_launch": frames,
             "algorithmic_bytes": int(bytes_alg), "achieved_gbs": round(bytes_alg / (p50 * 1e-3) / 1e9, 1),
             "frac_of_measured_hbm": round(bytes_alg / (p50 * 1e-3) / 1e9 / pk, 3), "note": note}
        if cpu_fn is not None and not args.no_cpu:
            r["cpu_ms_per_frame"] = round(cpu_time(cpu_fn) / 1.0, 3)
            r["gpu_ms_per_frame"] = round(p50 / frames, 5)
        rows.append(r)
        print(json.dumps(r), flush=True)

    from oracle import box3d, detectron_post, pre2d, yolo_head, yolo_post
    from oracle import voxelize as ov
    torch.set_num_threads(1)

    # ---- 2D preprocess (a1-a7) -------------------------------------------------------------------------
    def pre_row(name, config, n, h, w, oh, ow, mode, div=255.0):
        bases = [synth.image_smooth(h, w, 10 * r + 1) for r in range(RING)]
        frames = [torch.from_numpy(np.stack([np.roll(bases[r], 17 * i, axis=1) for i in range(n)])).cuda() for r in range(RING)]
        out = torch.empty((n, 3, oh, ow), dtype=torch.float32, device="cuda")
        host = frames[0][0].cpu().numpy()
        row(name, config, lambda i: eng.preprocess_images(frames[i % RING], oh, ow, mode=mode, div=div, out=out), n * (3 * h * w + 12 * oh * ow), n,
            lambda: pre2d.preprocess_frame(host, oh, ow, mode=mode, div=div))

    pre_row("a1-a3 stretch+adjust", "C2 64x1280x720->640x640", 64, 720, 1280, 640, 640, "stretch")
    pre_row("a7 letterbox+adjust", "C2 64x1280x720->640x640", 64, 720, 1280, 640, 640, "letterbox")
    pre_row("a1 stretch, general scales", "32x1920x1080->640x640 (two-pass kernel: 3:1 and 1.69:1)", 32, 1080, 1920, 640, 640, "stretch")
    pre_row("a3 adjust only", "C1 1x640x640", 1, 640, 640, 640, 640, "stretch")
    pre_row("a3 adjust only (batch 64)", "C1-shape 64x640x640", 64, 640, 640, 640, 640, "stretch")
    pre_row("a4 FCOS adjust", "C3 1x800x1333", 1, 800, 1333, 800, 1333, "stretch", 1.0)

    # ---- YOLO post (c1-c2) -----------------------------------------------------------------------------
    def yolo_row(name, config, b, n, nc):
        preds = [torch.from_numpy(synth.yolo_output(b, n, nc, 50 + r)).cuda() for r in range(RING)]
        out = torch.zeros((b, 300, 6), dtype=torch.float64, device="cuda")
        cnt = torch.zeros((2, b), dtype=torch.int32, device="cuda")
        host = preds[0][:1].cpu().numpy()
        row(name, config, lambda i: eng.yolo_nms(preds[i % RING], conf_thres=0.3, out=out, counts=cnt), b * (4 * n * (5 + nc) + 48 * 300 + 4), b,
            lambda: yolo_post.extract_boxes(host, conf_thres=0.3, as_written=True),
            "CPU = reference as written (struct.unpack wire parsing)")

    yolo_row("c1-c2 YOLOv5 post", "C1 1x25200x85", 1, 25200, 80)
    yolo_row("c1-c2 YOLOv5 post", "C2 64x25200x7", 64, 25200, 2)
    yolo_row("c1-c2 YOLOv5 post", "shipped 1x16128x7", 1, 16128, 2)

    # ---- YOLO head decode + YOLOv4 post (c4, c6) -------------------------------------------------------------
    anchors = [12, 16, 19, 36, 40, 28, 36, 75, 76, 55, 72, 146, 142, 110, 192, 243, 459, 401]
    heads_h, per = [], []
    rng = np.random.default_rng(1)
    for mask, stride, hw in [([0, 1, 2], 8, 64), ([3, 4, 5], 16, 32), ([6, 7, 8], 32, 16)]:
        heads_h.append(rng.normal(-2, 1.5, (1, 255, hw, hw)).astype(np.float32))
        per.append([anchors[2 * k + j] / stride for k in mask for j in (0, 1)])
    heads_d = [torch.from_numpy(h).cuda() for h in heads_h]
    row("c6 YOLOv4 head decode", "3 heads 1x255x{64,32,16}^2 -> 16128 rows", lambda i: eng.yolo_head_decode_v4(heads_d, per, 80),
        sum(h.size for h in heads_h) * 4 + 16128 * 84 * 4, 1,
        lambda: yolo_head.get_region_boxes([yolo_head.yolo_forward_dynamic(h, 80, a, 3) for h, a in zip(heads_h, per)]))
    b4, c4 = eng.yolo_head_decode_v4(heads_d, per, 80)
    b4h, c4h = b4.cpu().numpy(), c4.cpu().numpy()
    row("c4 YOLOv4 post (per-class NMS)", "1x16128x80", lambda i: eng.yolo4_post(b4, c4, 0.4, 0.6), 16128 * 84 * 4, 1,
        lambda: yolo_head.extract_boxes_onnx(0.4, 0.6, [b4h, c4h]))

    # ---- detectron (c7) ----------------------------------------------------------------------------------------
    cells = detectron_post.retinanet_cell_anchors()
    if not args.only or args.only in "c7 RetinaNet dense inference":
        for tag, bgd in (("", (-7.0, 1.2)), (" (stress: 2 % of the logits pass the threshold)", (-6.0, 1.5))):
            lg, dl = synth.detectron_head_outputs(3001, background=bgd)
            lgd, dld = [torch.from_numpy(t).cuda() for t in lg], [torch.from_numpy(t).cuda() for t in dl]
            ncand = int(sum((t > np.log(0.05 / 0.95)).sum() for t in lg))
            row("c7 RetinaNet dense inference", f"C3 201600 anchors x 80 classes, {ncand} candidates{tag}",
                lambda i: eng.detectron_post(lgd, dld, (8, 16, 32, 64, 128), cells, (800, 1333)),
                201600 * 84 * 4, 1, lambda: detectron_post.dense_inference(lg, dl, (8, 16, 32, 64, 128), cells, (800, 1333)))

    # ---- voxelise + decorate (b1-b7) ----------------------------------------------------------------------------
    def vox_row(name, config, pts_list, cfg, p, mv, overflow, f=None, **kw):
        dev = [torch.from_numpy(x).cuda() for x in pts_list]
        n, fdim = pts_list[0].shape
        v, c, nn_, m = eng.voxelize(dev[0], cfg[0], cfg[1], p, mv, overflow=overflow, **kw)
        mm = int(m.item())
        row(name, f"{config} (N={n}, M={mm})", lambda i: eng.voxelize(dev[i % len(dev)], cfg[0], cfg[1], p, mv, overflow=overflow, **kw),
            4 * n * fdim + mm * (4 * p * fdim + 20), 1, lambda: ov.points_to_voxel(pts_list[0], cfg[0], cfg[1], p, mv, overflow),
            "7 kernels + 4 memsets; output buffers allocated per call by the wrapper")
        return v, c, nn_, m

    c4pts = [synth.lidar_sweep(synth.seed_for(4, r)) for r in range(RING)]
    v, c, nn_, m = vox_row("b3-b5 voxelise", "C4 0.16 m pillars, P=32, cap 12000", c4pts, KITTI, 32, 12000, "continue", range_mask=True)
    vox_row("b3-b5 voxelise", "C4 file cap 40000", c4pts, KITTI, 32, 40000, "continue", range_mask=True)
    vox_row("b3-b5 voxelise", "C4' shipped 0.05 m, P=5, cap 40000", c4pts, SHIPPED, 5, 40000, "continue", range_mask=True)
    c5pts = [synth.lidar_multisweep(synth.seed_for(5, r)) for r in range(RING)]
    vox_row("b4-b5 voxelise (det3d, break)", "C5 0.2 m, P=20, cap 60000", c5pts, NUSC, 20, 60000, "break")
    mm = int(m.item())
    vh, ch, nh = v[:mm].cpu().numpy(), c[:mm].cpu().numpy(), nn_[:mm].cpu().numpy()
    row("b6 pillar decoration", f"C4 M={mm}, P=32", lambda i: eng.pillar_decorate(v, c, nn_, KITTI[1], KITTI[0], m=m),
        mm * (4 * 32 * 4 + 20) + mm * 4 * 32 * 10, 1, lambda: ov.pillar_decorate(vh, ch, nh, KITTI[1], np.array(KITTI[0], np.float32)))
    dyn = [torch.from_numpy(np.concatenate([np.zeros((len(x), 1), np.float32), x], 1)).cuda() for x in c5pts]
    dynh = dyn[0].cpu().numpy()
    row("b7 dynamic pillar VFE", f"C5 N={len(c5pts[0])}", lambda i: eng.dynamic_pillar(dyn[i % RING], NUSC[0], NUSC[1]),
        len(c5pts[0]) * (6 * 4 + 11 * 4 + 8), 1, lambda: ov.dynamic_pillar_vfe(dynh, NUSC[1], NUSC[0], [512, 512, 1]))

    # ---- 3D post (d2-d6) ----------------------------------------------------------------------------------------
    pps = [synth.pp_head_outputs(4000 + r) for r in range(RING)]
    ppd = [[torch.from_numpy(t).cuda() for t in x] for x in pps]
    row("d2-d5 PointPillars anchor post", "C4 321408 anchors, 3 classes", lambda i: eng.pp_anchor_post(*ppd[i % RING], **PP),
        (18 + 42 + 12) * 248 * 216 * 4, 1, lambda: box3d.pp_anchor_post(*pps[0], **PP), "latency-bound: 1 CTA orders + NMS")
    cps = [synth.centerpoint_head_outputs(5000 + r) for r in range(RING)]
    cpd = [[{k: torch.from_numpy(t).cuda() for k, t in hd.items()} for hd in x] for x in cps]
    row("d6 CenterPoint head post", "C5 6 heads 128x128, K=500", lambda i: eng.centerpoint_post(cpd[i % RING], CLASS_MAP, 4, NUSC[1], NUSC[0], [-61.2, -61.2, -10, 61.2, 61.2, 10]),
        (10 + 60) * 128 * 128 * 4, 1, lambda: box3d.centerhead_generate_predicted_boxes(cps[0], CLASS_MAP, NUSC[0], NUSC[1], 4, [-61.2, -61.2, -10, 61.2, 61.2, 10]),
        "latency-bound: 6 CTAs")
    bx = synth.random_boxes7(4096, 7, extent=70)
    sc = np.random.default_rng(3).uniform(0, 1, 4096).astype(np.float32)
    bxd, scd = torch.from_numpy(bx).cuda(), torch.from_numpy(sc).cuda()
    row("d5 rotated NMS", "4096 boxes, thr 0.01, post 500", lambda i: eng.rotated_nms(bxd, scd, 0.01, 4096, 500), 4096 * 32, 1,
        lambda: box3d.nms_gpu(bx, sc, 0.01), "SM-issue/latency-bound")
    row("d5 rotated IoU matrix", "1000x1000", lambda i: eng.rotated_iou(bxd[:1000], bxd[1000:2000]), 1000 * 1000 * 4 + 2000 * 28, 1,
        lambda: box3d.boxes_iou_bev(bx[:1000], bx[1000:2000]), "compute-bound (polygon clipping)")

    # ---- after the path: box rescale + overlay (8f rank 2) ---------------------------------------------------------
    if not args.only or args.only in "f2 box rescale + overlay":
        from oracle import overlay
        predo = torch.from_numpy(synth.yolo_output(64, 25200, 2, 5)).cuda()
        deto, cnto, _ = eng.yolo_nms(predo, conf_thres=0.3)
        base = torch.from_numpy(synth.image_smooth(720, 1280, 3)).cuda()
        ring = [base.unsqueeze(0).repeat(64, 1, 1, 1).contiguous() for _ in range(RING)]
        host_frame, host_det = base.cpu().numpy(), deto[0, :int(cnto[0])].cpu().numpy()
        eng.draw_detections(ring[0], deto, cnto, ["crop", "weed"], (0, 255, 0), [640, 640], math_mode=1)
        painted = int((ring[0] != base.unsqueeze(0)).any(-1).sum())

        def cv2_loop():   # the reference's own host loop (ros_inference.py:148-169) on one frame
            import cv2
            img = host_frame.copy()
            for row_ in host_det:
                b = overlay.scale_box(row_[:4].astype(np.float32), (720, 1280), [640, 640], f64=True)
                cv2.rectangle(img, (int(b[0]), int(b[1])), (int(b[2]), int(b[3])), (0, 255, 0), 2)
                cv2.putText(img, overlay.label_text(row_[4], ["crop", "weed"][int(row_[5])]), (int(b[0]), int(b[1] - 10)),
                            cv2.FONT_HERSHEY_SIMPLEX, 0.5, (0, 255, 0), 2)
        row("f2 box rescale + overlay", f"C2 64x1280x720, {int(cnto.sum())} detections", lambda i: eng.draw_detections(ring[i % RING], deto, cnto, ["crop", "weed"], (0, 255, 0), [640, 640], math_mode=1),
            int(cnto.sum()) * 48 + painted * 3, 64, cv2_loop, "CPU column: the reference's cv2 loop on one frame; launch-bound (two kernels)")

    # ---- ROS message rows (8f rank 3) ----------------------------------------------------------------------------------
    if not args.only or args.only in "g1 PointCloud2 unpack" or args.only in "g2 BoundingBoxArray fields":
        from oracle import pointcloud2 as opc2
        sweep = synth.lidar_sweep(11)
        sweep[::37, 0] = np.nan
        fields = [opc2.make_field("x", 0, 7), opc2.make_field("y", 4, 7), opc2.make_field("z", 8, 7), opc2.make_field("intensity", 16, 7),
                  opc2.make_field("ring", 20, 4)]
        npts = len(sweep)
        raw = np.zeros((npts, 32), np.uint8)
        raw[:, 0:12] = sweep[:, :3].copy().view(np.uint8).reshape(npts, 12)
        raw[:, 16:20] = sweep[:, 3:4].copy().view(np.uint8).reshape(npts, 4)
        from types import SimpleNamespace
        msg = SimpleNamespace(width=npts // 64, height=64, point_step=32, row_step=(npts // 64) * 32, data=raw.tobytes(), is_bigendian=False, fields=fields)
        devs = [torch.from_numpy(raw.reshape(-1).copy()).cuda() for _ in range(RING)]
        row("g1 PointCloud2 unpack", f"C4 {npts} points x 32 B, 4 fields, skip_nans", lambda i: eng.pointcloud2_unpack(devs[i % RING], msg.width, 64, 32, msg.row_step, [0, 4, 8, 16], [7, 7, 7, 7]),
            npts * 32 + npts * 16, 1, lambda: np.array(list(opc2.read_points(msg, field_names=("x", "y", "z", "intensity"), skip_nans=True))),
            "CPU column: read_points' per-point struct.unpack generator; launch-bound (4 kernels)")
        bxj = synth.random_boxes7(500, 5, extent=40)
        scj = np.random.default_rng(5).uniform(0, 1, 500).astype(np.float32)
        lbj = np.random.default_rng(6).integers(1, 4, 500)
        bj, sj, lj = torch.from_numpy(bxj).cuda(), torch.from_numpy(scj).cuda(), torch.from_numpy(lbj).cuda()
        row("g2 BoundingBoxArray fields", "500 detections", lambda i: eng.boxes_to_jsk(bj, sj, lj), 500 * 40 + 100 * 96, 1,
            lambda: opc2.jsk_rows(bxj, scj, lbj), "launch-bound (1 CTA)")

    # ---- evaluation math (8f rank 4) -------------------------------------------------------------------------------------
    if not args.only or args.only in "h1 evaluation metrics":
        from oracle import eval_metrics as oem
        ev = [synth.eval_frame(synth.seed_for(12, k), n_gt=40, nc=2, false_pos=12) for k in range(64)]
        gcnt, pcnt = [len(g) for g, _ in ev], [len(p) for _, p in ev]
        gd = torch.from_numpy(np.concatenate([g for g, _ in ev])).cuda()
        pd = torch.from_numpy(np.concatenate([p for _, p in ev])).cuda()
        row("h1 evaluation metrics", f"64 frames, {sum(gcnt)} labels, {sum(pcnt)} predictions, 2 classes",
            lambda i: eng.calculate_metrics(gd, gcnt, pd, pcnt, max_classes=2), (sum(gcnt) + sum(pcnt)) * 48 + sum(pcnt) * 10 + 64 * 2 * 13 * 8, 64,
            lambda: oem.calculate_metrics([ev[0][0]], [ev[0][1]]),
            "CPU column: the oracle on one frame (the reference's torch+numpy version: about 1 ms per frame); latency-bound (two launches, one CTA per frame)")

    print("\n| row | config | GPU p50 ms | GPU ms/frame | CPU oracle ms/frame | alg. bytes | GB/s | frac of measured HBM |")
    print("|---|---|---:|---:|---:|---:|---:|---:|")
    for r in rows:
        print(f"| {r['row']} | {r['config']} | {r['gpu_ms_p50']} | {r.get('gpu_ms_per_frame', '')} | {r.get('cpu_ms_per_frame', '')} | "
              f"{r['algorithmic_bytes']} | {r['achieved_gbs']} | {r['frac_of_measured_hbm']} |")


if __name__ == "__main__":
    main()
