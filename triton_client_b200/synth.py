"""Seeded synthetic inputs for every BASELINE.json config.

The Triton model call is replaced by synthetic model outputs of the exact named shapes; camera frames
and LiDAR sweeps are synthetic too (no datasets in the image). All generators are
numpy.random.default_rng(seed) with seed = 1000*config + frame_idx.
"""
import numpy as np


def seed_for(config, frame):
    return 1000 * int(config) + int(frame)


# ---- C1/C2 camera frames -------------------------------------------------------------------------
def image_noise(h, w, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)


def image_smooth(h, w, seed):
    """Sum of 8 random 2-D sinusoids per channel + sigma=8 noise, clipped to u8."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    out = np.empty((h, w, 3), dtype=np.float32)
    for c in range(3):
        acc = np.full((h, w), 128.0, dtype=np.float32)
        for _ in range(8):
            fx, fy = rng.uniform(-0.05, 0.05, 2)
            ph, amp = rng.uniform(0, 2 * np.pi), rng.uniform(5, 25)
            acc += amp * np.sin(fx * xx + fy * yy + ph)
        out[:, :, c] = acc
    out += rng.normal(0, 8, out.shape).astype(np.float32)
    return np.clip(np.rint(out), 0, 255).astype(np.uint8)


# ---- C1/C2 decoded YOLOv5 head output [B, N, 5+nc] ------------------------------------------------
def yolo_output(batch, n, nc, seed, img=640, objects=40, rows_per_object=30, tie_frac=0.05):
    """40 'objects' per frame each spawning 30 jittered rows with high objectness; background rows have
    obj in U(0,0.05); 5 % of all scores are quantised to 1/1024 to force score ties."""
    rng = np.random.default_rng(seed)
    out = np.empty((batch, n, 5 + nc), dtype=np.float32)
    out[:, :, 0:2] = rng.uniform(0, img, (batch, n, 2))
    out[:, :, 2:4] = np.exp(rng.normal(np.log(60.0), 0.7, (batch, n, 2)))
    out[:, :, 4] = rng.uniform(0, 0.05, (batch, n))
    out[:, :, 5:] = rng.uniform(0, 0.05, (batch, n, nc))
    k = objects * rows_per_object
    for b in range(batch):
        cxy = rng.uniform(0, img, (objects, 2))
        wh = np.exp(rng.normal(np.log(60.0), 0.7, (objects, 2)))
        cls = rng.integers(0, nc, objects)
        rows = rng.choice(n, size=min(k, n), replace=False)
        obj_id = np.arange(len(rows)) % objects
        out[b, rows, 0:2] = cxy[obj_id] + rng.normal(0, 4.0, (len(rows), 2))
        out[b, rows, 2:4] = wh[obj_id] * np.exp(rng.normal(0, 0.05, (len(rows), 2)))
        out[b, rows, 4] = rng.uniform(0.3, 1.0, len(rows))
        out[b, rows, 5 + cls[obj_id]] = rng.uniform(0.5, 1.0, len(rows))
    q = rng.random((batch, n)) < tie_frac
    out[:, :, 4] = np.where(q, np.round(out[:, :, 4] * 1024) / 1024, out[:, :, 4])
    qc = rng.random((batch, n, nc)) < tie_frac
    out[:, :, 5:] = np.where(qc, np.round(out[:, :, 5:] * 1024) / 1024, out[:, :, 5:])
    return out.astype(np.float32)


# ---- C4/C5 LiDAR sweeps ---------------------------------------------------------------------------
def lidar_sweep(seed, beams=64, azimuths=1875, n_boxes=30, dropout=0.1, max_range=75.0, sensor_h=1.73, full_circle=False):
    """Spinning-LiDAR-like sweep: each beam/azimuth ray hits the ground plane (z=-sensor_h) or one of
    n_boxes random upright boxes; 10 % dropout. Returns (N,4) f32 x,y,z,intensity (C4: 64x1875=120k rays)."""
    rng = np.random.default_rng(seed)
    elev = np.deg2rad(np.linspace(-24.8, 2.0, beams))[:, None]
    span = np.pi if not full_circle else 2 * np.pi
    az = (np.linspace(-span / 2, span / 2, azimuths, endpoint=False) + rng.uniform(0, 1e-3))[None, :]
    dx, dy, dz = np.cos(elev) * np.cos(az), np.cos(elev) * np.sin(az), np.sin(elev) * np.ones_like(az)
    with np.errstate(divide="ignore", invalid="ignore"):
        t = np.where(dz < -1e-3, -sensor_h / dz, np.inf)
    t = np.minimum(t, max_range)
    # boxes: centre (cx, cy), half sizes, height
    bc = rng.uniform([5, -30], [60, 30], (n_boxes, 2)) if not full_circle else rng.uniform(-45, 45, (n_boxes, 2))
    bh = rng.uniform([0.8, 0.8], [2.5, 1.2], (n_boxes, 2))
    bz = rng.uniform(1.4, 3.0, n_boxes)
    for k in range(n_boxes):  # slab test in xy, then check the hit height
        with np.errstate(divide="ignore", invalid="ignore"):
            tx1, tx2 = (bc[k, 0] - bh[k, 0]) / dx, (bc[k, 0] + bh[k, 0]) / dx
            ty1, ty2 = (bc[k, 1] - bh[k, 1]) / dy, (bc[k, 1] + bh[k, 1]) / dy
        tn = np.maximum(np.minimum(tx1, tx2), np.minimum(ty1, ty2))
        tf = np.minimum(np.maximum(tx1, tx2), np.maximum(ty1, ty2))
        hit = (tn < tf) & (tn > 0) & (tn < t) & ((tn * dz) < (bz[k] - sensor_h))
        t = np.where(hit, tn, t)
    t = t * (1 + rng.normal(0, 0.002, t.shape))
    keep = (rng.random(t.shape) > dropout) & np.isfinite(t)
    pts = np.stack([t * dx, t * dy, t * dz, rng.random(t.shape)], axis=-1)[keep]
    return pts.astype(np.float32)


def lidar_multisweep(seed, sweeps=10, beams=32, azimuths=940):
    """C5: 10 accumulated nuScenes-like sweeps (32 beams x 940 az ~ 30k each), 5th feature = time lag."""
    out = []
    for s in range(sweeps):
        p = lidar_sweep(seed * 100 + s, beams=beams, azimuths=azimuths, n_boxes=25, dropout=0.0, max_range=60.0,
                        sensor_h=1.84, full_circle=True)
        p[:, 3] *= 255.0
        out.append(np.concatenate([p, np.full((len(p), 1), 0.05 * s, np.float32)], axis=1))
    return np.concatenate(out, axis=0).astype(np.float32)


# ---- C4 PointPillars head outputs / C5 CenterPoint head outputs ------------------------------------
def _quantise(x, step=1.0 / 64):
    """Logits on a 1/64 grid: two distinct logits then differ by >= 1e-5 relative after the sigmoid for every
    |x| < 9, so no two libm implementations of exp() can order a pair of scores differently."""
    return (np.round(x / step) * step).astype(np.float32)


def pp_head_outputs(seed, h=248, w=216, num_class=3, num_rot=2, planted=50):
    """cls (h,w,A*nc) ~ N(-5,1.2) with planted object clusters (+6), box (h,w,A*7) ~ N(0,0.3), dir (h,w,A*2) ~ N(0,1)."""
    rng = np.random.default_rng(seed)
    a = num_class * num_rot
    cls = rng.normal(-5.0, 1.2, (h, w, a, num_class))
    for _ in range(planted):
        y, x = rng.integers(2, h - 2), rng.integers(2, w - 2)
        k, c = rng.integers(0, a), rng.integers(0, num_class)
        cls[y - 1:y + 2, x - 1:x + 2, k, c] += rng.uniform(4, 8)
    box = rng.normal(0, 0.3, (h, w, a * 7)).astype(np.float32)
    dirp = _quantise(rng.normal(0, 1, (h, w, a * 2)))
    return _quantise(np.clip(cls, -8.9, 8.9)).reshape(h, w, a * num_class), box, dirp


def centerpoint_head_outputs(seed, classes_per_head=(1, 2, 2, 1, 2, 2), h=128, w=128, peaks=80):
    """Per head: hm logits N(-6,1) + Gaussian peaks (sigma 1.5 px, amp +8), center U(0,1), center_z N(-1,.5),
    dim (log) N(.7,.3), rot unit vectors, vel N(0,2). NCHW planes of one frame."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    heads = []
    for c in classes_per_head:
        hm = rng.normal(-6, 1, (c, h, w))
        for _ in range(max(1, peaks // len(classes_per_head))):
            k, py, px = rng.integers(0, c), rng.uniform(0, h), rng.uniform(0, w)
            hm[k] += 8.0 * np.exp(-((yy - py) ** 2 + (xx - px) ** 2) / (2 * 1.5 ** 2))
        ang = rng.uniform(-np.pi, np.pi, (h, w))
        heads.append({
            "hm": _quantise(np.clip(hm, -8.9, 8.9)),
            "center": rng.uniform(0, 1, (2, h, w)).astype(np.float32),
            "center_z": rng.normal(-1, 0.5, (1, h, w)).astype(np.float32),
            "dim": rng.normal(0.7, 0.3, (3, h, w)).astype(np.float32),
            "rot": np.stack([np.cos(ang), np.sin(ang)]).astype(np.float32),
            "vel": rng.normal(0, 2, (2, h, w)).astype(np.float32),
        })
    return heads


def random_boxes7(n, seed, extent=40.0):
    rng = np.random.default_rng(seed)
    b = np.zeros((n, 7), np.float32)
    b[:, 0:2] = rng.uniform(0, extent, (n, 2))
    b[:, 2] = rng.uniform(-2, 0, n)
    b[:, 3:6] = rng.uniform(0.5, 5, (n, 3))
    b[:, 6] = rng.uniform(-4, 4, n)
    return b


# ---- C3 FCOS / RetinaNet head outputs at 800x1333 (padded 800x1344) --------------------------------
def detectron_head_outputs(seed, anchors_per_loc=9, num_classes=80, image_hw=(800, 1344), strides=(8, 16, 32, 64, 128), planted=60, fcos=False,
                           background=(-6.0, 1.5)):
    """Per level l: logits (A*K,H_l,W_l) ~ N(background) with planted positives N(+2,1), deltas (A*4,H_l,W_l) ~ N(0,0.4);
    FCOS adds centerness (1,H_l,W_l) ~ N(0,1). Logits on a 1/64 grid (see _quantise). The default background leaves 2 % of
    all logits above the 0.05 score threshold (250 k candidates on the stride-8 level of C3: a stress case); (-7, 1.2)
    leaves 0.04 % (4 k there), the order of magnitude a trained focal-loss head produces."""
    rng = np.random.default_rng(seed)
    logits, deltas, ctr = [], [], []
    for s in strides:
        h, w = -(-image_hw[0] // s), -(-image_hw[1] // s)
        lg = rng.normal(background[0], background[1], (anchors_per_loc * num_classes, h, w))
        for _ in range(planted // len(strides) + 1):
            lg[rng.integers(0, lg.shape[0]), rng.integers(0, h), rng.integers(0, w)] = rng.normal(2, 1)
        logits.append(_quantise(np.clip(lg, -8.9, 8.9)))
        deltas.append(rng.normal(0, 0.4, (anchors_per_loc * 4, h, w)).astype(np.float32))
        ctr.append(_quantise(rng.normal(0, 1, (1, h, w))))
    return (logits, deltas, ctr) if fcos else (logits, deltas)


# ---- evaluation frames (§8f rank 4): ground truths + predictions as EvaluateInference collects them ----------------
def eval_frame(seed, n_gt=24, nc=2, img_wh=(1280, 720), miss=0.15, swap=0.1, dup=0.3, false_pos=8, jitter=6.0):
    """One frame of evaluation input: `gts` (G,6) float64 rows [x1,y1,x2,y2,score,cls] (gt_callback after xywh2xyxy,
    communicator/evaluate_inference.py:363-374) and `preds` (D,6) float64 rows [x1,y1,x2,y2,conf,cls]
    (_scale_box_array of extract_boxes' rows, :237-256). Each ground truth is found (jittered box) unless missed, sometimes
    twice, sometimes with the wrong class; false positives are added; rows come in descending confidence like NMS
    output. Coordinates carry f32-derived fractions so IoUs are generic (no exact ties)."""
    rng = np.random.default_rng(seed)
    w, h = img_wh
    cxy = rng.uniform([0, 0], [w, h], (n_gt, 2))
    wh = np.exp(rng.normal(np.log(70.0), 0.5, (n_gt, 2)))
    cls = rng.integers(0, nc, n_gt).astype(np.float64)
    gts = np.concatenate([cxy - wh / 2, cxy + wh / 2, np.ones((n_gt, 1)), cls[:, None]], 1)
    rows = []
    for g in range(n_gt):
        if rng.uniform() < miss:
            continue
        for _ in range(2 if rng.uniform() < dup else 1):
            b = gts[g, :4] + rng.normal(0, jitter, 4)
            c = cls[g] if rng.uniform() >= swap else float((int(cls[g]) + 1) % max(nc, 2))
            rows.append([*b, rng.uniform(0.3, 1.0), c])
    for _ in range(false_pos):
        c0 = rng.uniform([0, 0], [w, h])
        s = np.exp(rng.normal(np.log(70.0), 0.5, 2))
        rows.append([*(c0 - s / 2), *(c0 + s / 2), rng.uniform(0.3, 0.7), float(rng.integers(0, nc))])
    preds = np.array(rows, np.float64).reshape(-1, 6)
    preds[:, :4] = preds[:, :4].astype(np.float32).astype(np.float64) * (w / 640.0)   # f32 boxes scaled in float64
    gts[:, :4] *= w / 640.0
    preds = preds[np.argsort(-preds[:, 4], kind="stable")]
    return gts, preds
