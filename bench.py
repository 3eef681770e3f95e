#!/usr/bin/env python
"""bench.py — frames/sec of the client pre+post hot path on B200 (BASELINE.json metric).

Default workload = BASELINE.json configs[1] ("YOLOv5nCROP batch 64"): a step is 64 synthetic 1280x720 RGB frames
letterboxed to 640x640 (u8 HWC -> f32 NCHW /255) + decode/filter/class-aware NMS of a synthetic 64x25200x7 model output
(conf 0.3 as the reference's callers pass, iou 0.45, max_det 300). `--workload centerpoint` (alias c5) makes the step
16 nuScenes-like 10-sweep clouds voxelised + 16 CenterPoint head outputs decoded (BASELINE.json configs[4], the stream
north_star shards over 8 GPUs).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME]     # this framework (CUDA, C ABI)
  python bench.py --impl reference [--steps K] [--warmup W]                   # the reference's CPU path on host cores

The K steps are repeated R times back to back inside one timed region (R chosen so that the region lasts >= 0.5 s);
`ms_per_step` is the mean over those R*K steps, `ms_per_step_min` the fastest of the R rounds. Under torchrun (N>1)
every rank runs its own shard of the frame stream (shard.shard_frames; weak scaling, no collective on the data path);
the final result gather (shard.gather_detections over NCCL) is timed apart as `comm_ms`. At N=1 the line also carries
`configs`: every BASELINE.json config measured the same way (c0..c4 = configs[0..4]) with the CPU oracle beside it.
"""
import argparse
import json
import math
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONF, IOU, MAX_DET = 0.3, 0.45, 300
KITTI = ([0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4])
NUSC = ([-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8])
PP = dict(anchor_range=[0, -39.68, -3, 69.12, 39.68, 1], anchor_sizes=[[3.9, 1.6, 1.56], [0.8, 0.6, 1.73], [1.76, 0.6, 1.73]],
          anchor_bottoms=[-1.78, -0.6, -0.6], anchor_rotations=[[0, 1.57]] * 3)
CP_MAP = [[0], [1, 2], [3, 4], [5], [6, 7], [8, 9]]
CP_LIMIT = [-61.2, -61.2, -10.0, 61.2, 61.2, 10.0]
DET_STRIDES = (8, 16, 32, 64, 128)
ALIASES = {"c0": "yolo1", "c1": "yolo64", "c2": "detectron", "c3": "pointpillars", "c4": "centerpoint", "c5": "centerpoint",
           "C1": "yolo1", "C2": "yolo64", "C3": "detectron", "C4": "pointpillars", "C5": "centerpoint"}
CONFIG_KEYS = ["yolo1", "yolo64", "detectron", "pointpillars", "centerpoint"]     # c0..c4 = BASELINE.json configs[0..4]


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


def ncu_traffic():
    """DRAM bytes per launch of the roofline kernel from the committed ncu capture (builder-run, not measured here)."""
    for name in ("r02_pre_kernel.json", "r01_pre_kernel.json"):
        try:
            return json.load(open(os.path.join(ROOT, "profiles", name)))["dram_traffic_bytes_per_launch"], "ncu --set full, builder-run (profiles/%s)" % name
        except Exception:
            pass
    return None, None


class ClockSampler(threading.Thread):
    """Polls SM clock + throttle reasons through NVML while a timed region runs."""

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

    def finish(self):
        self.stop_flag = True
        self.join()
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


# =====================================================================================================================
# Workloads. A workload owns RING distinct device-resident input sets (cycled so that no step re-reads what the previous
# one left in the 126 MB L2), launches one step through the C ABI, names its stages with their algorithmic bytes
# (every input read once + every output written once), and can run the CPU oracle on its frames.
# =====================================================================================================================
class Workload:
    key = name = survey = ""
    batch = 1
    ring = 3
    stage_names = ()

    def __init__(self, eng, dev, frame_ids):
        self.eng, self.dev, self.ids = eng, dev, list(frame_ids)

    def stage_bytes(self):                      # algorithmic bytes per step and stage (may use measured counts, e.g. M)
        raise NotImplementedError

    def step(self, i, ev=None):                 # ev: len(stage_names) + 1 events recorded at the stage boundaries
        raise NotImplementedError

    def results(self):                          # (rows [frames, max_det, C], counts [frames]) of the last step
        raise NotImplementedError

    def cpu_frames(self, idxs, as_written=True):   # the oracle on host frames idxs of input set 0; returns a detection count
        raise NotImplementedError

    def l2_note(self):
        return ""


def _rec(ev, k):
    if ev is not None:
        ev[k].record()


class Yolo(Workload):
    """BASELINE configs[0] (batch 1, 640x640 frame already at model size, 1x25200x85) and configs[1] (batch 64, 1280x720
    letterboxed to 640, 64x25200x7)."""
    stage_names = ("preprocess", "yolo_post")

    def __init__(self, eng, dev, frame_ids, batch, src_hw, nc, mode):
        import torch
        from triton_client_b200 import synth
        super().__init__(eng, dev, frame_ids)
        self.batch, self.src_hw, self.nc, self.mode, self.out, self.nrows = batch, src_hw, nc, mode, 640, 25200
        self.ring = 3 if batch > 1 else 16
        h, w = src_hw
        self.frames_h, self.preds_h = [], []
        cfg = 2 if batch > 1 else 1
        for r in range(self.ring):
            f = np.empty((batch, h, w, 3), dtype=np.uint8)
            base = noise = None
            for i in range(batch):          # every frame distinct: shifted smooth scene, every 4th one pure noise; seeded by global frame id
                g = self.ids[i]
                if base is None or batch == 1:
                    base = synth.image_smooth(h, w, synth.seed_for(cfg, 100 * r + g % 7))
                    noise = synth.image_noise(h, w, synth.seed_for(cfg, 100 * r + 50 + g % 7))
                f[i] = np.roll(noise if g % 4 == 3 else base, (13 * g, 29 * g), axis=(0, 1))
            self.frames_h.append(f)
            self.preds_h.append(synth.yolo_output(batch, self.nrows, nc, synth.seed_for(cfg, 1000 + 100 * r + self.ids[0])))
        self.frames_d = [torch.from_numpy(f).to(dev) for f in self.frames_h]
        self.preds_d = [torch.from_numpy(p).to(dev) for p in self.preds_h]
        self.pre_out = torch.empty((batch, 3, self.out, self.out), dtype=torch.float32, device=dev)
        self.det = torch.zeros((batch, MAX_DET, 6), dtype=torch.float64, device=dev)
        self.cnt = torch.zeros((2, batch), dtype=torch.int32, device=dev)

    def stage_bytes(self):
        h, w = self.src_hw
        return {"preprocess": self.batch * (3 * h * w + 12 * self.out * self.out),
                "yolo_post": self.batch * (4 * self.nrows * (5 + self.nc) + 48 * MAX_DET + 4)}

    def step(self, i, ev=None):
        r = i % self.ring
        _rec(ev, 0)
        self.eng.preprocess_images(self.frames_d[r], self.out, self.out, mode=self.mode, div=255.0, pad_value=114, out=self.pre_out)
        _rec(ev, 1)
        self.eng.yolo_nms(self.preds_d[r], conf_thres=CONF, iou_thres=IOU, max_det=MAX_DET, out=self.det, counts=self.cnt)
        _rec(ev, 2)

    def results(self):
        return self.det, self.cnt[0]

    def cpu_frames(self, idxs, as_written=True):
        from oracle import pre2d, yolo_post
        frames, preds = self.frames_h[0], self.preds_h[0]
        ndet = 0
        for i in idxs:
            x = pre2d.letterbox(frames[i], self.out, self.out)[0] if self.mode == "letterbox" else frames[i]
            t = pre2d.image_adjust_yolov5(x)
            assert t.shape == (1, 3, self.out, self.out)
            res = yolo_post.extract_boxes(yolo_post.FakeResponse([preds[i:i + 1]]), conf_thres=CONF, iou_thres=IOU, max_det=MAX_DET, as_written=as_written)
            ndet += sum(int(r.shape[0]) for r in res) if isinstance(res, list) else 0
        return ndet

    def l2_note(self):
        mb = (self.frames_h[0].nbytes + self.preds_h[0].nbytes) / 1e6
        return f"inputs cycle through {self.ring} distinct sets of {mb:.0f} MB each ({self.ring * mb:.0f} MB > 126 MB L2); no flush needed"


class Yolo1(Yolo):
    key, name, survey = "yolo1", "configs[0]: YOLOv5n COCO 640x640 batch 1: image_adjust + decode/NMS of 1x25200x85", "C1"
    batch = 1

    def __init__(self, eng, dev, frame_ids):
        super().__init__(eng, dev, frame_ids, 1, (640, 640), 80, "stretch")


class Yolo64(Yolo):
    key, name, survey = "yolo64", "configs[1]: YOLOv5nCROP batch 64, 1280x720 frames letterboxed to 640 + decode/class-aware NMS of 64x25200x7", "C2"
    batch = 64

    def __init__(self, eng, dev, frame_ids):
        super().__init__(eng, dev, frame_ids, 64, (720, 1280), 2, "letterbox")


class Detectron(Workload):
    """BASELINE configs[2]: FCOS/RetinaNet at 800x1333: FCOSpreprocess.image_adjust (u8 HWC -> f32 NCHW, unscaled) +
    detectron2 dense inference (per-level threshold, top-k 1000, NMS over 80 classes) of 201 600 anchors per image."""
    key, name, survey = "detectron", "configs[2]: RetinaNet 800x1333 batch 16: image_adjust + per-level threshold/top-k 1000/NMS over 80 classes (201 600 anchors per image)", "C3"
    batch, ring = 16, 1
    stage_names = ("preprocess", "detectron_post")

    def __init__(self, eng, dev, frame_ids):
        import torch
        from oracle import detectron_post as dp
        from triton_client_b200 import synth
        super().__init__(eng, dev, frame_ids)
        self.cells = dp.retinanet_cell_anchors()
        self.hw = (800, 1333)
        base = [synth.detectron_head_outputs(synth.seed_for(3, k)) for k in range(2)]     # two host frames; the batch = device-side shifts of them
        self.host_heads = base
        self.logits, self.deltas = [], []
        for l in range(5):
            lg = torch.stack([torch.roll(torch.from_numpy(base[i % 2][0][l]).to(dev), shifts=i // 2, dims=-1) for i in range(self.batch)])
            dl = torch.stack([torch.roll(torch.from_numpy(base[i % 2][1][l]).to(dev), shifts=i // 2, dims=-1) for i in range(self.batch)])
            self.logits.append(lg.contiguous())
            self.deltas.append(dl.contiguous())
        self.frames_h = np.stack([np.roll(synth.image_smooth(800, 1333, synth.seed_for(3, 10 + g % 5)), 11 * g, axis=1) for g in self.ids[:self.batch]])
        self.frames_d = torch.from_numpy(self.frames_h).to(dev)
        self.pre_out = torch.empty((self.batch, 3, 800, 1333), dtype=torch.float32, device=dev)
        self.res = None

    def stage_bytes(self):
        return {"preprocess": self.batch * (3 * 800 * 1333 + 12 * 800 * 1333),
                "detectron_post": self.batch * (201600 * (80 + 4) * 4 + 100 * 24)}

    def step(self, i, ev=None):
        _rec(ev, 0)
        self.eng.preprocess_images(self.frames_d, 800, 1333, mode="stretch", div=1.0, out=self.pre_out)
        _rec(ev, 1)
        self.res = self.eng.detectron_post(self.logits, self.deltas, DET_STRIDES, self.cells, self.hw)
        _rec(ev, 2)

    def results(self):
        import torch
        ob, osc, oc, cnt = self.res
        return torch.cat([ob, osc[..., None], oc[..., None].float()], -1), cnt

    def cpu_frames(self, idxs, as_written=True):
        from oracle import detectron_post as dp, pre2d
        n = 0
        for i in idxs:
            pre2d.preprocess_frame(self.frames_h[i % self.batch], 800, 1333, div=1.0)
            lg, dl = self.host_heads[i % 2]
            n += len(dp.dense_inference(lg, dl, DET_STRIDES, self.cells, self.hw)[1])
        return n

    def l2_note(self):
        return "one input set: a step reads 1.19 GB of head outputs and frames (>> 126 MB L2); no flush needed"


def _concat_sweeps(sweeps):
    offs = np.concatenate([[0], np.cumsum([len(s) for s in sweeps])]).astype(int).tolist()
    return np.concatenate(sweeps).astype(np.float32), offs


class PointPillars(Workload):
    """BASELINE configs[3]: PointPillars KITTI: ~120 k-point sweeps -> 0.16 m pillars (12 000 voxels x 32 points), pillar
    decoration, 3-class anchor decode + rotated BEV NMS (321 408 anchors per frame)."""
    key, name, survey = "pointpillars", "configs[3]: PointPillars KITTI batch 16: voxelise (0.16 m pillars, 12000 x 32) + pillar decoration + anchor decode/rotated NMS of 321 408 anchors", "C4"
    batch, ring = 16, 2
    stage_names = ("voxelize", "decorate", "pp_post")

    def __init__(self, eng, dev, frame_ids):
        import torch
        from triton_client_b200 import synth
        super().__init__(eng, dev, frame_ids)
        self.sweeps_h, self.pts, self.offs, self.heads, self.heads_h = [], [], [], [], []
        for r in range(self.ring):
            sweeps = [synth.lidar_sweep(synth.seed_for(4, 100 * r + g)) for g in self.ids[:self.batch]]
            p, o = _concat_sweeps(sweeps)
            self.sweeps_h.append(sweeps)
            self.pts.append(torch.from_numpy(p).to(dev))
            self.offs.append(o)
            hs = [synth.pp_head_outputs(synth.seed_for(4, 500 + 100 * r + g % 4)) for g in range(4)]   # four host frames, shifted along x on the device
            self.heads_h.append(hs)
            self.heads.append(tuple(torch.stack([torch.roll(torch.from_numpy(hs[i % 4][j]).to(dev), shifts=i // 4, dims=1) for i in range(self.batch)]).contiguous()
                                    for j in range(3)))
        self.vox = self.res = None
        self.n_pts = [int(o[-1]) for o in self.offs]

    def stage_bytes(self):
        m = int(self.vox[3].sum().item()) if self.vox is not None else 12000 * self.batch
        n = self.n_pts[0]
        return {"voxelize": 16 * n + m * (4 * 32 * 4 + 20), "decorate": m * (4 * 32 * 4 + 20 + 4 * 32 * 10),
                "pp_post": self.batch * ((18 + 42 + 12) * 248 * 216 * 4 + 500 * 36)}

    def step(self, i, ev=None):
        r = i % self.ring
        _rec(ev, 0)
        self.vox = self.eng.voxelize(self.pts[r], KITTI[0], KITTI[1], 32, 12000, offsets=self.offs[r], out=self.vox)
        _rec(ev, 1)
        v, c, n, m = self.vox
        self.feats = self.eng.pillar_decorate(v, c, n, KITTI[1], KITTI[0], m=m, out=getattr(self, "feats", None))
        _rec(ev, 2)
        self.res = self.eng.pp_anchor_post(*self.heads[r], **PP, out=self.res)
        _rec(ev, 3)

    def results(self):
        import torch
        ob, osc, ol, cnt = self.res
        return torch.cat([ob, osc[..., None], ol[..., None].float()], -1), cnt

    def cpu_frames(self, idxs, as_written=True):
        from oracle import box3d, voxelize as ov
        n = 0
        for i in idxs:
            v = ov.points_to_voxel(self.sweeps_h[0][i % self.batch], KITTI[0], KITTI[1], 32, 12000)
            ov.pillar_decorate(v[0], np.pad(v[1], ((0, 0), (1, 0))), v[2], KITTI[1], np.array(KITTI[0], np.float32))
            n += len(box3d.pp_anchor_post(*self.heads_h[0][i % 4], **PP)["pred_scores"])
        return n

    def l2_note(self):
        return "inputs cycle through 2 distinct sets of ~270 MB each (sweeps + head outputs) and a step writes ~340 MB of voxels and features; no flush needed"


class CenterPoint(Workload):
    """BASELINE configs[4]: CenterPoint-PP nuScenes: ~300 k-point 10-sweep clouds -> 0.2 m pillars (det3d voxel generator,
    'break' at max_voxels, 20 points x 5 features), 6-head heatmap top-K decode + rotated NMS."""
    key, name, survey = "centerpoint", "configs[4]: CenterPoint-PP nuScenes batch 16: voxelise ~300k-point 10-sweep clouds (0.2 m, 60000 x 20 x 5) + 6-head heatmap top-500 decode/rotated NMS", "C5"
    batch, ring = 16, 2
    stage_names = ("voxelize", "centerpoint_post")

    def __init__(self, eng, dev, frame_ids):
        import torch
        from triton_client_b200 import synth
        super().__init__(eng, dev, frame_ids)
        self.sweeps_h, self.pts, self.offs, self.heads, self.heads_h = [], [], [], [], []
        for r in range(self.ring):
            sweeps = [synth.lidar_multisweep(synth.seed_for(5, 100 * r + g % 8)) for g in self.ids[:4]]     # four host clouds, rotated about z per frame below
            clouds = []
            for i in range(self.batch):
                c = sweeps[i % 4].copy()
                a = 0.05 * (i // 4) + 0.003 * self.ids[i]
                x, y = c[:, 0].copy(), c[:, 1].copy()
                c[:, 0], c[:, 1] = np.float32(math.cos(a)) * x - np.float32(math.sin(a)) * y, np.float32(math.sin(a)) * x + np.float32(math.cos(a)) * y
                clouds.append(c)
            p, o = _concat_sweeps(clouds)
            self.sweeps_h.append(clouds)
            self.pts.append(torch.from_numpy(p).to(dev))
            self.offs.append(o)
            hs = [synth.centerpoint_head_outputs(synth.seed_for(5, 500 + 100 * r + g)) for g in range(4)]
            self.heads_h.append(hs)
            self.heads.append([{k: torch.stack([torch.roll(torch.from_numpy(hs[i % 4][h][k]).to(dev), shifts=i // 4, dims=-1) for i in range(self.batch)]).contiguous()
                                for k in hs[0][h]} for h in range(6)])
        self.vox = self.res = None
        self.n_pts = [int(o[-1]) for o in self.offs]

    def stage_bytes(self):
        m = int(self.vox[3].sum().item()) if self.vox is not None else 21000 * self.batch
        return {"voxelize": 20 * self.n_pts[0] + m * (4 * 20 * 5 + 20), "centerpoint_post": self.batch * (70 * 128 * 128 * 4 + 6 * 83 * 44)}

    def step(self, i, ev=None):
        r = i % self.ring
        _rec(ev, 0)
        self.vox = self.eng.voxelize(self.pts[r], NUSC[0], NUSC[1], 20, 60000, overflow="break", offsets=self.offs[r], out=self.vox)
        _rec(ev, 1)
        self.res = self.eng.centerpoint_post(self.heads[r], CP_MAP, 4, NUSC[1], NUSC[0], CP_LIMIT, out=self.res)
        _rec(ev, 2)

    def results(self):
        import torch
        ob, osc, ol, cnt = self.res
        return torch.cat([ob, osc[..., None], ol[..., None].float()], -1), cnt

    def cpu_frames(self, idxs, as_written=True):
        from oracle import box3d, voxelize as ov
        n = 0
        for i in idxs:
            ov.points_to_voxel(self.sweeps_h[0][i % self.batch], NUSC[0], NUSC[1], 20, 60000, "break")
            n += len(box3d.centerhead_generate_predicted_boxes(self.heads_h[0][i % 4], CP_MAP, NUSC[0], NUSC[1], 4, CP_LIMIT)["pred_scores"])
        return n

    def l2_note(self):
        return "inputs cycle through 2 distinct sets of ~170 MB each (clouds + head outputs) and a step writes ~140 MB of voxels; no flush needed"


class JpegDecode(Workload):
    """the CompressedImage branch of the 2D callback (communicator/ros_inference.py:123-131):
    cv2.imdecode(buf, IMREAD_ANYCOLOR)[..., ::-1] of every frame -> tcb_jpeg_decode of the batch. The entry point takes HOST
    streams, so a step holds the host marker parse, the upload of the scan bytes and the three kernels."""
    key, survey = "jpeg", "f3"
    name = "next row f3: 64 CompressedImage frames (1280x720 baseline JPEG, quality 90, 4:2:0) -> RGB u8 frames on the device (host marker parse + scan upload + Huffman / IDCT / colour kernels)"
    batch, ring = 64, 2
    stage_names = ("jpeg_decode",)
    cpu_kind, cpu_what = "reference", "cv2.imdecode + channel flip, the reference's own call"

    def __init__(self, eng, dev, frame_ids):
        import cv2
        import torch
        from triton_client_b200 import synth
        super().__init__(eng, dev, frame_ids)
        base = [cv2.imencode(".jpg", synth.image_smooth(720, 1280, synth.seed_for(2, 40 + k)), [cv2.IMWRITE_JPEG_QUALITY, 90])[1].tobytes() for k in range(6)]
        self.sets = [[base[(i + 3 * r) % 6] for i in range(self.batch)] for r in range(self.ring)]
        self.out = torch.empty((self.batch, 720, 1280, 3), dtype=torch.uint8, device=dev)

    def stage_bytes(self):
        return {"jpeg_decode": sum(len(s) for s in self.sets[0]) + self.batch * 720 * 1280 * 3}

    def step(self, i, ev=None):
        _rec(ev, 0)
        self.eng.jpeg_decode(self.sets[i % self.ring], out=self.out)
        _rec(ev, 1)

    def cpu_frames(self, idxs, as_written=True):
        import cv2
        n = 0
        for i in idxs:
            n += int(cv2.imdecode(np.frombuffer(self.sets[0][i % self.batch], np.uint8), cv2.IMREAD_ANYCOLOR)[..., ::-1].shape[0] > 0)
        return n

    def l2_note(self):
        return "a step writes 177 MB of frames (> 126 MB L2) from ~16 MB of compressed bytes that arrive from the host inside the step"


WORKLOADS = {c.key: c for c in (Yolo1, Yolo64, Detectron, PointPillars, CenterPoint, JpegDecode)}


def make_workload(key, eng, dev, rank, world):
    from triton_client_b200.shard import shard_frames
    cls = WORKLOADS[key]
    ids = shard_frames(cls.batch * world, rank, world, "interleave")     # this rank's frames of the global stream
    return cls(eng, dev, ids)


def workload_config(w_cls, world):
    """The `config` object of the JSON line — identical for the GPU arm and the reference arm."""
    return {"workload": w_cls.name, "batch_per_gpu": w_cls.batch, "conf_thres": CONF, "iou_thres": IOU, "max_det": MAX_DET,
            "parallelism": f"frame stream sharded over {world} rank(s) (shard.shard_frames), no data-path collective"}


# =====================================================================================================================
# timing
# =====================================================================================================================
def timed_region(torch, dist, world, dev, wl, steps, warmup, min_seconds, event_every=8, want_samples=16):
    """W warm-up steps, then R rounds of exactly `steps` steps back to back between two events (barrier + synchronize on
    both sides, max over ranks). Per-stage events are recorded on every `event_every`-th step only (an event record
    between two kernels takes away the programmatic launch edge between them). Returns a dict."""
    for i in range(max(warmup, 3)):
        wl.step(i)
    torch.cuda.synchronize()
    # pilot round: how many rounds make the region >= min_seconds (the same number on every rank)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        wl.step(i)
    b.record()
    torch.cuda.synchronize()
    pilot = torch.tensor([a.elapsed_time(b)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(pilot, op=dist.ReduceOp.MAX)
    rounds = max(1, min(20000, int(math.ceil(min_seconds * 1e3 / max(float(pilot.item()), 1e-3)))))
    total_steps = rounds * steps
    ns = len(wl.stage_names)
    n_ev = min(max(want_samples, 1), max(1, total_steps // max(1, event_every)))
    ev_steps = {min(total_steps - 1, event_every // 2 + k * event_every): [torch.cuda.Event(enable_timing=True) for _ in range(ns + 1)] for k in range(n_ev)}
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(rounds + 1)]
    sampler = ClockSampler(dev.index or 0)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = wl.eng.launches
    sampler.start()
    marks[0].record()
    k = 0
    for r in range(rounds):
        for _ in range(steps):
            wl.step(k, ev_steps.get(k))
            k += 1
        marks[r + 1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.finish()
    launches = wl.eng.launches - l0
    per_round = [marks[r].elapsed_time(marks[r + 1]) for r in range(rounds)]
    tot = torch.tensor([marks[0].elapsed_time(marks[rounds]), min(per_round)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    stage_ms = {}
    for j, nme in enumerate(wl.stage_names):
        stage_ms[nme] = statistics.mean(e[j].elapsed_time(e[j + 1]) for e in ev_steps.values())
    return {"total_ms": float(tot[0].item()), "rounds": rounds, "timed_steps": total_steps, "ms_per_step": float(tot[0].item()) / total_steps,
            "ms_per_step_min": float(tot[1].item()) / steps, "stage_ms": stage_ms, "stage_samples": len(ev_steps), "launches": launches,
            "clocks": clocks}


def stage_report(wl, stage_ms, peak):
    by = wl.stage_bytes()
    return {n: {"ms": stage_ms[n], "algorithmic_bytes": int(by[n]), "gbs": by[n] / (stage_ms[n] * 1e-3) / 1e9,
                "frac_of_hbm_peak": by[n] / (stage_ms[n] * 1e-3) / 1e9 / peak} for n in wl.stage_names}


def cpu_port(wl, budget_s=6.0, max_frames=8):
    """The oracle port, one process / one thread, on a bounded sample of the workload's own frames."""
    import torch
    torch.set_num_threads(1)
    try:
        import cv2
        cv2.setNumThreads(1)
    except Exception:
        pass
    t0 = time.perf_counter()
    wl.cpu_frames([0], True)                                   # warm-up, also the per-frame cost estimate
    per = time.perf_counter() - t0
    n = int(max(1, min(max_frames, wl.batch, budget_s / max(per, 1e-3))))
    t0 = time.perf_counter()
    wl.cpu_frames(range(n), True)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": "frames/s", "cores": 1, "kind": getattr(wl, "cpu_kind", "port"),
            "sample": f"{n} frame(s) of the workload, {getattr(wl, 'cpu_what', 'oracle port as the reference writes the path')}"}, n


def latency_c1(wl, frames=1000):
    """p50 per-frame latency of the batch-1 workload (BASELINE configs[0]): one CUDA-event pair per frame and stage; also
    replayed from a CUDA graph of the two calls (launch overhead removed)."""
    import torch
    dev = wl.dev
    for i in range(20):
        wl.step(i)
    torch.cuda.synchronize()
    pre_t, post_t, tot_t = [], [], []
    for i in range(frames):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        wl.step(i, e)
        e[2].synchronize()
        pre_t.append(e[0].elapsed_time(e[1]))
        post_t.append(e[1].elapsed_time(e[2]))
        tot_t.append(e[0].elapsed_time(e[2]))
    res = {"frames": frames, "pre_ms": statistics.median(pre_t), "post_ms": statistics.median(post_t), "pre_post_ms": statistics.median(tot_t),
           "graph_pre_post_ms": None}
    try:   # the same two calls captured once and replayed: what a steady-state caller pays without launch overhead
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream(dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):
            wl.step(0)
        torch.cuda.current_stream(dev).wait_stream(s)
        with torch.cuda.graph(g, stream=s):
            wl.step(0)
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


def measure_config(torch, key, dev, peak, steps, min_seconds, with_cpu):
    """One BASELINE config, single GPU: the same timed region as the main line, with its own clock samples."""
    from triton_client_b200.gpu import Engine
    eng = Engine(dev.index or 0, workspace_bytes=(3 << 30) if key == "detectron" else ((2 << 30) if key == "jpeg" else (1 << 30)))   # detectron: room for every logit of 16 images as a candidate
    wl = make_workload(key, eng, dev, 0, 1)
    t = timed_region(torch, None, 1, dev, wl, steps, 3, min_seconds)
    out = {"workload": wl.name, "survey_config": wl.survey, "batch": wl.batch, "frames_per_s": wl.batch * 1e3 / t["ms_per_step"],
           "ms_per_step": t["ms_per_step"], "ms_per_frame": t["ms_per_step"] / wl.batch, "timed_steps": t["timed_steps"],
           "stages": stage_report(wl, t["stage_ms"], peak), "stage_samples": t["stage_samples"], "gpu_launches_per_step": t["launches"] / t["timed_steps"],
           "l2": wl.l2_note(), "clocks": t["clocks"]}
    if key == "yolo1":
        out["p50_latency"] = latency_c1(wl)
    if with_cpu:
        out["cpu_baseline"], _ = cpu_port(wl, budget_s=4.0, max_frames=4)
    del wl, eng
    torch.cuda.empty_cache()
    return out


# =====================================================================================================================
# end to end through the drop-in classes (host buffers in, host results out)
# =====================================================================================================================
def e2e_yolo(torch, dist, world, dev, wl, eng, steps):
    from triton_client_b200.clients import Yolov5client
    from triton_client_b200.clients.wire import HostResponse
    client = Yolov5client(engine=eng)
    pre, post = client.get_preprocess(), client.get_postprocess()
    pin_frames = [torch.from_numpy(f).pin_memory() for f in wl.frames_h]
    pin_preds = [torch.from_numpy(p).pin_memory() for p in wl.preds_h]
    B, OUT = wl.batch, wl.out

    def issue(i):
        r = i % wl.ring
        # both halves of a step are queued before either is waited for: the response's upload follows the frames' uploads on
        # the link while the preprocess download is still running
        hx = pre.image_adjust_batch_async(pin_frames[r], (OUT, OUT), mode=wl.mode, copy=False)            # H2D + kernel + D2H (host f32 NCHW)
        pending = post.extract_boxes_async(HostResponse([pin_preds[r]]), conf_thres=CONF, iou_thres=IOU)  # H2D + kernels + D2H
        return hx, pending

    def collect(h):
        dets = h[1].result()
        x = h[0].result()
        assert isinstance(dets, list) and len(dets) == B and x.shape == (B, 3, OUT, OUT), dets
        return x, dets

    def run(k):
        """k steps, software-pipelined by one: step i+1 is queued before step i's host results are collected, so the
        link never idles between steps; every step's inputs go up and its results come down inside the timed region."""
        prev = None
        for i in range(k):
            cur_h = issue(i)
            if prev is not None:
                collect(prev)
            prev = cur_h
        return collect(prev)

    run(3)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    x, dets = run(steps)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    # the pipelined host results are the device-resident ones (checked outside the timed region)
    wl.step(steps - 1)
    torch.cuda.synchronize()
    assert torch.equal(torch.from_numpy(x).to(dev), wl.pre_out), "e2e preprocess result differs from the device-resident one"
    assert [len(d) for d in dets] == wl.cnt[0].tolist() and all(torch.equal(d.to(dev), wl.det[b, :len(d)]) for b, d in enumerate(dets)), "e2e detections differ"
    h2d = wl.frames_h[0].nbytes + wl.preds_h[0].nbytes
    d2h = getattr(pre, "last_d2h_bytes", B * 3 * OUT * OUT * 4) + wl.det.numel() * 8 + wl.cnt.numel() * 4
    return wall, h2d, d2h, "Yolov5preprocess.image_adjust_batch_async + Yolov5postprocess.extract_boxes_async + .result(), two steps in flight"


def e2e_centerpoint(torch, dist, world, dev, wl, eng, steps):
    from triton_client_b200.clients import PointPillarPostprocess, det3DPreprocess
    pre = det3DPreprocess(engine=eng, f64_math=False, max_voxel_num=[60000, 60000])
    post = PointPillarPostprocess(engine=eng)

    def run(k):
        n = 0
        for i in range(k):
            r = i % wl.ring
            for f in range(wl.batch):
                v, c, nn = pre.filter_pc(wl.sweeps_h[r][f][:, :4])                # host cloud in, host voxels out
                d = post.decode_centerpoint(wl.heads_h[r][f % 4])                 # host head outputs in, host boxes out
                n += len(d["pred_scores"]) + len(nn)
        return n

    run(1)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    run(steps)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    h2d = sum(s[:, :4].nbytes for s in wl.sweeps_h[0]) + wl.batch * 70 * 128 * 128 * 4
    m = int(wl.vox[3].sum().item()) if wl.vox is not None else 0
    d2h = m * (20 * 5 * 4 + 20) + wl.batch * 6 * 83 * 44
    return wall, h2d, d2h, "det3DPreprocess.filter_pc + PointPillarPostprocess.decode_centerpoint per frame (host arrays in and out)"


def pcie_probe(torch, dist, world, dev, mb=128, reps=6):
    """What the host link gives every rank when all ranks copy pinned memory both ways at once (the e2e path's ceiling;
    profiles/probe_pcie_nranks.py is the stand-alone version with the per-direction figures). GB/s per rank and direction."""
    up_h, dn_h = torch.empty(mb << 20, dtype=torch.uint8).pin_memory(), torch.empty(mb << 20, dtype=torch.uint8).pin_memory()
    up_d, dn_d = torch.empty(mb << 20, dtype=torch.uint8, device=dev), torch.zeros(mb << 20, dtype=torch.uint8, device=dev)
    s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def once(n):
        for _ in range(n):
            with torch.cuda.stream(s_up):
                up_d.copy_(up_h, non_blocking=True)
            with torch.cuda.stream(s_dn):
                dn_h.copy_(dn_d, non_blocking=True)
    once(1)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    once(reps)
    torch.cuda.synchronize()
    t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return reps * (mb << 20) / float(t.item()) / 1e9


# =====================================================================================================================
# arms
# =====================================================================================================================
_CPU_WL = None


def _cpu_worker(idxs):
    import torch
    torch.set_num_threads(1)
    try:
        import cv2
        cv2.setNumThreads(1)
    except Exception:
        pass
    return _CPU_WL.cpu_frames(idxs, True)


class _HostOnly:
    launches = 0


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port; the reference is Python and cannot
    travel to the GPU box) on all host cores, one process per core, on the same input batches as the GPU arm."""
    global _CPU_WL
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    import torch
    key = ALIASES.get(args.workload, args.workload)
    cls = WORKLOADS[key]
    cores = os.cpu_count() or 1
    os.environ.setdefault("CUDA_VISIBLE_DEVICES", "")
    torch.set_num_threads(1)     # no OpenMP pool in the parent: the workers below are forked
    try:
        import cv2
        cv2.setNumThreads(1)
    except Exception:
        pass
    # the GPU arm's rank-0 inputs (same classes, same seeds), kept on the host
    from triton_client_b200.shard import shard_frames
    _CPU_WL = cls(_HostOnly(), torch.device("cpu"), shard_frames(cls.batch, 0, 1))
    t0 = time.perf_counter()
    _CPU_WL.cpu_frames([0], True)
    per_frame = time.perf_counter() - t0
    # frames per step: a bounded sample of the batch, sized so that K steps end within ~2 minutes
    budget_frames = int(120.0 / max(per_frame, 1e-3) * cores / max(args.steps + args.warmup, 1))
    per_step = max(1, min(cls.batch, budget_frames))
    nproc = min(cores, per_step)
    shards = [list(range(i, per_step, nproc)) for i in range(nproc)]
    ctx = mp.get_context("fork")
    with ctx.Pool(nproc) as pool:
        for _ in range(args.warmup):
            pool.map(_cpu_worker, [s[:1] for s in shards])
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_cpu_worker, shards)
        dt = time.perf_counter() - t0
    fps = per_step * args.steps / dt
    line = {"impl": "reference", "metric": "frames/sec pre+post", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(cls, args.gpus),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": nproc, "kind": "port",
                             "sample": f"the first {per_step} of the GPU arm's {cls.batch} frames per step (same seeds), {nproc} processes, reference path as written"},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from triton_client_b200.gpu import Engine
    from triton_client_b200.shard import bind_to_gpu_numa_node, gather_detections

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(local)      # before any pinned buffer is allocated
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    key = ALIASES.get(args.workload, args.workload)
    peak, which = peaks()

    eng = Engine(local, workspace_bytes=(256 << 20) if key.startswith("yolo") else ((3 << 30) if key == "detectron" else (1 << 30)))
    wl = make_workload(key, eng, dev, rank, world)

    # ---- device-resident throughput (inputs already in HBM) ------------------------------------------
    t = timed_region(torch, dist, world, dev, wl, args.steps, args.warmup, args.min_seconds, max(1, args.event_every))

    # ---- final result collection: counts + padded rows of the last step to every rank, in global frame order (KBs);
    #      timed apart: it closes a run, it is not part of a step ------------------------------------------------------
    rows, counts = wl.results()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gather_detections(rows, counts, wl.ids, wl.batch * world)      # first call: NCCL connection set-up
    torch.cuda.synchronize()
    c0.record()
    all_rows, all_counts = gather_detections(rows, counts, wl.ids, wl.batch * world)
    c1.record()
    torch.cuda.synchronize()
    comm = torch.tensor([c0.elapsed_time(c1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(comm, op=dist.ReduceOp.MAX)
    assert all_rows.shape[0] == wl.batch * world and torch.equal(all_counts[wl.ids], counts)

    # ---- the same steps with the two halves on two streams (what a frame pipeline does: the preprocess of the next
    #      batch runs while the previous batch's model output is post-processed). Extra figure; `value` stays sequential
    #      so that the roofline kernel is timed alone. Two contexts: a context is one frame in flight. ----------------
    two = None
    if key == "yolo64":
        eng2 = Engine(local, workspace_bytes=256 << 20)
        s_pre, s_post = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

        def step2(i):
            r = i % wl.ring
            with torch.cuda.stream(s_post):
                eng2.yolo_nms(wl.preds_d[r], conf_thres=CONF, iou_thres=IOU, max_det=MAX_DET, out=wl.det, counts=wl.cnt)
            with torch.cuda.stream(s_pre):
                eng.preprocess_images(wl.frames_d[r], wl.out, wl.out, mode=wl.mode, div=255.0, pad_value=114, out=wl.pre_out)

        cur = torch.cuda.current_stream(dev)
        for i in range(3):
            step2(i)
        torch.cuda.synchronize()
        n2 = t["timed_steps"]
        o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        o0.record()
        s_pre.wait_event(o0)
        s_post.wait_event(o0)
        for i in range(n2):
            step2(i)
        cur.wait_stream(s_pre)
        cur.wait_stream(s_post)
        o1.record()
        torch.cuda.synchronize()
        ov = torch.tensor([o0.elapsed_time(o1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ov, op=dist.ReduceOp.MAX)
        two = {"value": wl.batch * world * n2 / (float(ov.item()) * 1e-3), "unit": "frames/s", "ms_per_step": float(ov.item()) / n2,
               "what": "the same steps with preprocess and YOLO post on two streams (two contexts), device-resident"}

    # ---- end to end through the drop-in client classes: host buffers in, host results out ------------
    e2e_steps = max(3, min(args.steps, 20)) if key == "yolo64" else 2
    e2e_fn = e2e_yolo if key.startswith("yolo") else (e2e_centerpoint if key == "centerpoint" else None)
    e2e = None
    if e2e_fn is not None:
        wall, h2d, d2h, api = e2e_fn(torch, dist, world, dev, wl, eng, e2e_steps)
        e2e_s = torch.tensor([wall], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
        e2e = {"value": wl.batch * world * e2e_steps / float(e2e_s.item()), "unit": "frames/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "steps": e2e_steps, "api": api}
        link = pcie_probe(torch, dist, world, dev)
        step_s = wl.batch * world / e2e["value"]
        busy = max(h2d, d2h) / step_s / 1e9
        e2e["roofline"] = {"bound": "pcie", "achieved": busy, "peak": link, "unit": "GB/s per rank on the busier direction", "frac": busy / link,
                           "peak_source": f"pinned 128 MB copies both ways at once on all {world} rank(s), measured in this run"}

    if rank == 0:
        stages = stage_report(wl, t["stage_ms"], peak)
        dom = max(wl.stage_names, key=lambda n: stages[n]["algorithmic_bytes"])     # the stage that moves the most bytes
        traffic, traffic_src = ncu_traffic() if key == "yolo64" else (None, None)
        kernel_name = {"preprocess": "preprocess2_kernel", "voxelize": "tcb_voxelize (7 kernels: init, cell, mark, scan, assign, scatter, place)"}.get(dom, dom)
        cfg = workload_config(type(wl), world)
        line = {
            "metric": "frames/sec pre+post", "value": wl.batch * world * 1e3 / t["ms_per_step"], "unit": "frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": t["ms_per_step"],
            "ms_per_step_min": t["ms_per_step_min"], "rounds": t["rounds"], "timed_steps": t["timed_steps"], "timed_region_ms": t["total_ms"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8->f32 (pre), f64 (NMS)" if key.startswith("yolo") else "f32", "data": "synthetic",
            "config": cfg, "l2": wl.l2_note(), "host_numa": numa,
            "kernel_ms": {n: stages[n]["ms"] for n in wl.stage_names} | {"timed_on_steps": t["stage_samples"]},
            "stages": stages,
            "roofline": {"bound": "hbm", "kernel": kernel_name, "achieved": stages[dom]["gbs"], "peak": peak, "unit": "GB/s",
                         "frac": stages[dom]["frac_of_hbm_peak"], "peak_source": which + " copy bandwidth (MEASURED_PEAKS.json hbm_gbs)",
                         "bytes_per_launch": stages[dom]["algorithmic_bytes"], "traffic": traffic, "traffic_source": traffic_src},
            "comm_ms": float(comm.item()), "comm": "shard.gather_detections (all_gather of frame ids, counts, padded rows), once per run, outside the timed steps",
            "gpu_launches": t["launches"], "clocks": t["clocks"],
        }
        if e2e is not None:
            line["e2e"] = e2e
        if two is not None:
            line["two_streams"] = two
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"], _ = cpu_port(wl, budget_s=12.0, max_frames=8)
        if world == 1 and not args.no_configs:
            del wl
            torch.cuda.empty_cache()
            cfgs = {}
            for k, ck in enumerate(CONFIG_KEYS):
                try:
                    cfgs[f"c{k}"] = measure_config(torch, ck, dev, peak, 4 if ck != "yolo1" else 50, args.config_seconds, not args.no_cpu_baseline)
                except Exception as ex:       # a config that cannot run must not cost the headline line
                    cfgs[f"c{k}"] = {"workload": WORKLOADS[ck].name, "error": str(ex)[:300]}
            try:      # SURVEY 8(f) rank 3, beside the five BASELINE configs
                cfgs["f3_jpeg"] = measure_config(torch, "jpeg", dev, peak, 4, args.config_seconds, not args.no_cpu_baseline)
            except Exception as ex:
                cfgs["f3_jpeg"] = {"workload": WORKLOADS["jpeg"].name, "error": str(ex)[:300]}
            line["configs"] = cfgs
            if "p50_latency" in cfgs.get("c0", {}):
                line["p50_latency"] = dict(cfgs["c0"]["p50_latency"], config=WORKLOADS["yolo1"].name)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="yolo64", choices=sorted((set(WORKLOADS) - {"jpeg"}) | set(ALIASES)))
    ap.add_argument("--min-seconds", type=float, default=0.5, help="the K steps are repeated until the timed region lasts this long")
    ap.add_argument("--config-seconds", type=float, default=0.25, help="timed region of each entry of `configs`")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-config section (c0..c4) of the N=1 line")
    ap.add_argument("--event-every", type=int, default=8, help="record the per-stage CUDA events on every n-th timed step")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps == 20 and "--steps" not in sys.argv:
            args.steps = 3
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
