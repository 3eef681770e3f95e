"""Generates tests/golden/*.npz by running the UNMODIFIED reference (/root/reference) on small seeded
inputs. Run in the build container: `python -m oracle.make_golden`. The fixtures are committed; the
GPU box (where /root/reference does not exist) checks oracle and CUDA outputs against them.
Test infrastructure only.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_harness  # noqa: E402
from oracle.yolo_post import FakeResponse  # noqa: E402
from triton_client_b200 import synth  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def save(name, **arrays):
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(f"wrote {path} ({os.path.getsize(path)} B)")


def golden_pre2d():
    """Reference 2D preprocess: cv2.resize (communicator/ros_inference.py:140) + Yolov5preprocess /
    FCOSpreprocess.image_adjust, on small frames (kept small so the fixture stays a few KB)."""
    import cv2
    c = ref_harness.clients()
    ypre = c.Yolov5client().get_preprocess()
    fpre = c.FCOS_client().get_preprocess()
    out = {}
    cases = [(36, 64, 32, 32), (45, 80, 40, 40), (30, 40, 64, 48), (64, 64, 32, 32), (33, 51, 33, 51)]
    for k, (h, w, oh, ow) in enumerate(cases):
        img = synth.image_noise(h, w, synth.seed_for(1, k))
        resized = cv2.resize(img, (ow, oh))
        out[f"img{k}"] = img
        out[f"shape{k}"] = np.array([h, w, oh, ow])
        out[f"yolo{k}"] = ypre.image_adjust(resized)
        if k == 0:
            out[f"fcos{k}"] = fpre.image_adjust(resized)
    save("pre2d", **out)


def golden_yolo_post():
    """Reference Yolov5postprocess.extract_boxes on small synthetic responses (wire bytes included)."""
    c = ref_harness.clients()
    post = c.Yolov5client().get_postprocess()
    out = {}
    cases = [dict(b=2, n=600, nc=2, conf=0.3, iou=0.45, classes=None, agnostic=False, max_det=300),
             dict(b=1, n=300, nc=80, conf=0.3, iou=0.45, classes=None, agnostic=False, max_det=300),
             dict(b=2, n=500, nc=5, conf=0.25, iou=0.6, classes=[0, 3], agnostic=False, max_det=300),
             dict(b=1, n=700, nc=3, conf=0.3, iou=0.3, classes=None, agnostic=True, max_det=20)]
    for k, cs in enumerate(cases):
        pred = synth.yolo_output(cs["b"], cs["n"], cs["nc"], synth.seed_for(2, k), objects=8, rows_per_object=12, tie_frac=0.3)
        res = post.extract_boxes(FakeResponse([pred]), conf_thres=cs["conf"], iou_thres=cs["iou"], classes=cs["classes"],
                                 agnostic=cs["agnostic"], max_det=cs["max_det"])
        assert isinstance(res, list), res
        out[f"pred{k}"] = pred
        out[f"cfg{k}"] = np.array([cs["conf"], cs["iou"], float(cs["agnostic"]), cs["max_det"]])
        out[f"classes{k}"] = np.array(cs["classes"] if cs["classes"] is not None else [], dtype=np.int64)
        for bi, t in enumerate(res):
            out[f"out{k}_{bi}"] = t.numpy().astype(np.float64)
    # an all-background response: the reference returns the AssertionError object
    bg = np.zeros((1, 64, 7), dtype=np.float32)
    res = post.extract_boxes(FakeResponse([bg]), conf_thres=0.3)
    assert isinstance(res, AssertionError)
    out["empty_msg"] = np.array(str(res))
    save("yolo_post", **out)


def golden_pillar_vfe():
    """Reference PillarVFE.forward decoration (pillar_vfe.py:94-118, PFN layers removed) on voxels produced
    by the oracle voxeliser from a small synthetic sweep; KITTI pillar config and the nuScenes one."""
    from oracle import voxelize as ov
    out = {}
    cfgs = [([0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4], 32, 4), ([-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8], 20, 5)]
    for k, (rng, vs, p, f) in enumerate(cfgs):
        pts = synth.lidar_sweep(synth.seed_for(4, k), beams=16, azimuths=200, full_circle=(k == 1))
        if f == 5:
            pts = np.concatenate([pts, np.zeros((len(pts), 1), np.float32)], 1)
        voxels, coords, num = ov.points_to_voxel(pts, rng, vs, p, 400, "continue")
        coords4 = np.pad(coords, ((0, 0), (1, 0)))
        vfe = ref_harness.pillar_vfe(vs, np.array(rng, np.float32) if k == 0 else rng, f)
        bd = {"voxels": torch.from_numpy(voxels), "voxel_num_points": torch.from_numpy(num), "voxel_coords": torch.from_numpy(coords4)}
        feats = vfe.forward(bd)["pillar_features"].numpy()
        assert feats.shape == (voxels.shape[0], p, f + 6), feats.shape
        out[f"points{k}"] = pts
        out[f"range{k}"] = np.array(rng, np.float64)
        out[f"vsize{k}"] = np.array(vs, np.float64)
        out[f"maxpts{k}"] = np.array(p)
        out[f"feats{k}"] = feats
    save("pillar_vfe", **out)


def golden_box3d():
    """Reference 3D decode pieces run unmodified on CPU: ResidualCoder.decode_torch (box_coder_utils.py:45-77),
    limit_period (common_utils.py:21-24), AnchorGenerator.generate_anchors (anchor_generator.py:17-60, with
    Tensor.cuda() patched to a no-op: there is no GPU in the build container) and decode_bbox_from_heatmap
    (centernet_utils.py:154-216) on tie-free heatmaps."""
    rng = np.random.default_rng(77)
    out = {}
    bc = ref_harness.box_coder_utils()
    anchors = torch.from_numpy(synth.random_boxes7(300, 1))
    enc = torch.from_numpy(rng.normal(0, 0.4, (300, 7)).astype(np.float32))
    out["dec_anchors"], out["dec_enc"] = anchors.numpy(), enc.numpy()
    out["dec_out"] = bc.ResidualCoder().decode_torch(enc, anchors).numpy()
    cu = ref_harness.common_utils()
    val = torch.from_numpy(rng.uniform(-10, 10, 500).astype(np.float32))
    out["lp_val"] = val.numpy()
    out["lp_out"] = cu.limit_period(val - 0.78539, 0.0, 2 * np.pi / 2).numpy()
    # anchors
    ag = ref_harness.load_file("ref_anchor_generator", "utils/export_models/OpenPCDet/pcdet/models/dense_heads/target_assigner/anchor_generator.py")
    cfg = [{"anchor_sizes": [[3.9, 1.6, 1.56]], "anchor_rotations": [0, 1.57], "anchor_bottom_heights": [-1.78]},
           {"anchor_sizes": [[0.8, 0.6, 1.73]], "anchor_rotations": [0, 1.57], "anchor_bottom_heights": [-0.6]},
           {"anchor_sizes": [[1.76, 0.6, 1.73]], "anchor_rotations": [0, 1.57], "anchor_bottom_heights": [-0.6]}]
    pc_range = np.array([0, -39.68, -3, 69.12, 39.68, 1], np.float32)
    saved = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        gen = ag.AnchorGenerator(anchor_range=pc_range, anchor_generator_config=cfg)
        grid = [np.array([216, 248])] * 3
        alist, _ = gen.generate_anchors(grid)
    finally:
        torch.Tensor.cuda = saved
    full = torch.cat(alist, dim=-3).view(-1, 7).numpy()
    assert full.shape == (321408, 7)
    pick = rng.choice(full.shape[0], 4000, replace=False)
    pick.sort()
    out["anchor_idx"], out["anchor_val"] = pick.astype(np.int64), full[pick]
    out["anchor_first_last"] = np.stack([full[0], full[-1]])
    # centerpoint decode (K small, continuous logits -> no ties)
    cn = ref_harness.centernet_utils()
    h = w = 24
    hm = torch.from_numpy(rng.normal(-3, 2, (1, 2, h, w)).astype(np.float32)).sigmoid()
    parts = {"center": rng.uniform(0, 1, (1, 2, h, w)), "center_z": rng.normal(-1, .5, (1, 1, h, w)), "dim": np.exp(rng.normal(.7, .3, (1, 3, h, w))),
             "rot_cos": rng.normal(0, 1, (1, 1, h, w)), "rot_sin": rng.normal(0, 1, (1, 1, h, w)), "vel": rng.normal(0, 2, (1, 2, h, w))}
    parts = {k: torch.from_numpy(v.astype(np.float32)) for k, v in parts.items()}
    limit = torch.tensor([-61.2, -61.2, -10.0, 61.2, 61.2, 10.0]).float()
    res = cn.decode_bbox_from_heatmap(heatmap=hm, rot_cos=parts["rot_cos"], rot_sin=parts["rot_sin"], center=parts["center"],
                                      center_z=parts["center_z"], dim=parts["dim"], vel=parts["vel"],
                                      point_cloud_range=[-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], voxel_size=[0.2, 0.2, 8], feature_map_stride=4,
                                      K=40, score_thresh=0.1, post_center_limit_range=limit)[0]
    out["cp_hm"] = hm.numpy()
    for k, v in parts.items():
        out["cp_" + k] = v.numpy()
    out["cp_boxes"], out["cp_scores"], out["cp_labels"] = res["pred_boxes"].numpy(), res["pred_scores"].numpy(), res["pred_labels"].numpy()
    save("box3d", **out)


def golden_yolo4():
    """Reference YOLOv4 path run unmodified: tools/yolo_layer.py YoloLayer.forward (-> yolo_forward_dynamic) on small
    synthetic heads, and utils/postprocess.py extract_boxes_onnx on its output (scores made tie-free)."""
    rng = np.random.default_rng(404)
    yl = ref_harness.yolo_layer()
    up = ref_harness.utils_postprocess()
    nc, anchors = 4, [12, 16, 19, 36, 40, 28, 36, 75, 76, 55, 72, 146, 142, 110, 192, 243, 459, 401]
    out = {"anchors": np.array(anchors)}
    bc = []
    for hi, (mask, stride, hw) in enumerate([([0, 1, 2], 8, 12), ([3, 4, 5], 16, 6), ([6, 7, 8], 32, 3)]):
        head = rng.normal(0, 1.5, (2, 3 * (5 + nc), hw, hw)).astype(np.float32)
        layer = yl.YoloLayer(anchor_mask=mask, num_classes=nc, anchors=anchors, num_anchors=9, stride=stride)
        layer.eval()
        b, c = layer(torch.from_numpy(head))
        out[f"head{hi}"] = head
        out[f"boxes{hi}"], out[f"confs{hi}"] = b.numpy(), c.numpy()
        bc.append((b, c))
    boxes, confs = torch.cat([b for b, _ in bc], 1).numpy(), torch.cat([c for _, c in bc], 1).numpy()
    res = up.extract_boxes_onnx(0.2, 0.4, [boxes, confs])
    for i, rows in enumerate(res):
        out[f"post{i}"] = np.array(rows, dtype=np.float32).reshape(-1, 7)
    save("yolo4", **out)


def overlay_pred(seed, n, nc, size, objects, margin):
    """Small YOLO response with `objects` planted boxes whose centres keep `margin` model pixels from the edge."""
    rng = np.random.default_rng(seed)
    out = np.zeros((1, n, 5 + nc), np.float32)
    out[0, :, 0:2] = rng.uniform(0, size, (n, 2))
    out[0, :, 2:4] = rng.uniform(8, 40, (n, 2))
    out[0, :, 4] = rng.uniform(0, 0.05, n)
    out[0, :, 5:] = rng.uniform(0, 0.05, (n, nc))
    cxy = rng.uniform(margin, size - margin, (objects, 2))
    wh = rng.uniform(12, 2 * margin if margin > 12 else 60, (objects, 2))
    cls = rng.integers(0, nc, objects)
    rows = rng.choice(n, size=objects * 5, replace=False)
    oid = np.arange(len(rows)) % objects
    out[0, rows, 0:2] = cxy[oid] + rng.normal(0, 1.5, (len(rows), 2))
    out[0, rows, 2:4] = wh[oid] * np.exp(rng.normal(0, 0.03, (len(rows), 2)))
    out[0, rows, 4] = rng.uniform(0.4, 1.0, len(rows))
    out[0, rows, 5 + cls[oid]] = rng.uniform(0.6, 1.0, len(rows))
    return out


def golden_overlay():
    """The unmodified RosInference._callback (communicator/ros_inference.py:119-175) run end to end on one frame:
    cv2.resize -> image_adjust -> [fake Triton response] -> extract_boxes -> _scale_boxes -> cv2.rectangle/putText.
    ROS and gRPC objects are fakes that only carry data. Stored per case: the frame's seed/shape, the response, the
    detections the reference extracted, the request bytes' checksum and the set of pixels it painted.
    numpy here is %s: np.float32 scalar * Python number stays float32 (the oracle's f64=False mode)."""
    import types
    import zlib
    comm = ref_harness.communicator()
    c = ref_harness.clients()
    client = c.Yolov5client()
    out = {}
    cases = [dict(h=180, w=320, size=160, n=300, objects=7, margin=40),     # every label inside the frame
             dict(h=240, w=200, size=128, n=300, objects=9, margin=0)]      # boxes and labels cut by the frame edge
    for k, cs in enumerate(cases):
        frame = synth.image_noise(cs["h"], cs["w"], synth.seed_for(7, k))
        pred = overlay_pred(synth.seed_for(8, k), cs["n"], 2, cs["size"], cs["objects"], cs["margin"])

        class Req:
            def __init__(self):
                self.inputs, self.raw_input_contents = [], []

            def ClearField(self, name):
                getattr(self, name).clear()

        chan = types.SimpleNamespace(input=types.SimpleNamespace(shape=[3, cs["size"], cs["size"]]), request=Req(),
                                     do_inference=lambda pred=pred: FakeResponse([pred]))
        published = []
        node = object.__new__(comm.RosInference)
        node.channel, node.client = chan, client
        node.br = types.SimpleNamespace(imgmsg_to_cv2=lambda msg, desired_encoding: msg.frame,
                                        cv2_to_imgmsg=lambda img, encoding: types.SimpleNamespace(img=img, header=None))
        node.color, node.topic_exists, node.image = (0, 255, 0), True, None
        node.client_postprocess, node.client_preprocess = client.get_postprocess(), client.get_preprocess()
        node.class_names = ["crop", "weed"]
        node.input_size = [cs["size"], cs["size"]]
        node.detection = types.SimpleNamespace(publish=published.append)
        node._callback(types.SimpleNamespace(format="rgb8", frame=frame, header=None))
        annotated = published[0].img
        assert annotated.shape == frame.shape and isinstance(node.prediction, list)
        painted = (annotated != frame).any(-1)
        assert (annotated[painted] == np.array(node.color, np.uint8)).all()
        out[f"cfg{k}"] = np.array([cs["h"], cs["w"], cs["size"], synth.seed_for(7, k)])
        out[f"frame_crc{k}"] = np.array(zlib.crc32(frame.tobytes()))
        out[f"pred{k}"] = pred
        out[f"det{k}"] = node.prediction[0].numpy().astype(np.float64)
        out[f"request_crc{k}"] = np.array(zlib.crc32(chan.request.raw_input_contents[0]))
        out[f"painted{k}"] = np.packbits(painted)
        print(f"overlay case {k}: {len(node.prediction[0])} detections, {int(painted.sum())} pixels painted")
    out["numpy_version"] = np.array(np.__version__)
    save("overlay", **out)


def golden_plot_boxes():
    """Yolov5postprocess.plot_boxes_cv2 (clients/postprocess/yolov5_postprocess.py:127-169) run unmodified: normalised
    boxes (rows [x1, y1, x2, y2, conf, conf, class] as extract_boxes_onnx builds them: np.float32 values, np.int64 class),
    per-class colours from get_color, putText(class name, scale 1.2, thickness 1), rectangle thickness 1. Boxes reach
    over every frame edge, so labels and rectangles are clipped. numpy here is %s (np.float32 * int stays float32)."""
    import contextlib
    import io
    c = ref_harness.clients()
    post = c.Yolov5client().get_postprocess()
    names = ["crop", "weed", "person", "traffic light", "Qjg|y_{"]
    out = {"names": np.array(names), "numpy_version": np.array(np.__version__)}
    for k, (h, w, n) in enumerate([(180, 320, 9), (97, 131, 14)]):
        rng = np.random.default_rng(600 + k)
        img = synth.image_noise(h, w, synth.seed_for(9, k))
        xy = rng.uniform(-0.15, 0.95, (n, 2))
        wh = rng.uniform(0.05, 0.5, (n, 2))
        boxes = np.concatenate([xy, xy + wh, rng.uniform(0.3, 1, (n, 2))], 1).astype(np.float32)
        cls = rng.integers(0, len(names), n)
        rows = [[b[0], b[1], b[2], b[3], b[4], b[5], np.int64(ci)] for b, ci in zip(boxes, cls)]
        with contextlib.redirect_stdout(io.StringIO()):
            res = post.plot_boxes_cv2(img, rows, class_names=names)
        assert res.shape == img.shape and (res != img).any()
        out[f"img{k}"], out[f"boxes{k}"], out[f"cls{k}"], out[f"out{k}"] = img, boxes, cls.astype(np.int64), res
    save("plot_boxes", **out)


golden_plot_boxes.__doc__ = golden_plot_boxes.__doc__ % np.__version__

EVAL_CASES = [dict(n_gt=24, nc=2), dict(n_gt=40, nc=2, miss=0.3, false_pos=20), dict(n_gt=6, nc=3, dup=0.6),
              dict(n_gt=12, nc=2, miss=1.0, false_pos=5),              # nothing matches: the frame is skipped
              dict(n_gt=60, nc=5, dup=0.5, swap=0.2, false_pos=30), dict(n_gt=1, nc=1, miss=0.0, false_pos=0, dup=0.0, swap=0.0),
              dict(n_gt=30, nc=2, jitter=14.0)]


def golden_eval():
    """The unmodified EvaluateInference.calculate_metrics / ap_per_class / compute_ap
    (communicator/evaluate_inference.py:131-218, 400-446) on seeded frames of ground truths + predictions. The node is
    built without its ROS/Prometheus set-up (object.__new__); the four Summary objects are recorders."""
    import types
    comm = ref_harness.communicator()
    c = ref_harness.clients()
    sys.modules["rospy"].loginfo = lambda *a, **k: None
    node = object.__new__(comm.EvaluateInference)
    node.client_postprocess = c.Yolov5client().get_postprocess()
    node.img_processed = node.gt_processed = True
    rec = {k: [] for k in ("p", "r", "f1", "ap_class")}
    for k in rec:
        setattr(node, k + "_summary", types.SimpleNamespace(observe=rec[k].append))
    frames = [synth.eval_frame(synth.seed_for(9, k), **cs) for k, cs in enumerate(EVAL_CASES)]
    node.all_groundtruths = [{"header": k, "ground_truths": g} for k, (g, _) in enumerate(frames)]
    node.all_predictions = [{"header": k, "prediction": p} for k, (_, p) in enumerate(frames)]
    stats = node.calculate_metrics()
    out, si, pos = {}, 0, {k: 0 for k in rec}
    for k, (g, p) in enumerate(frames):
        out[f"gts{k}"], out[f"preds{k}"] = g, p
        matched = si < len(stats) and np.shares_memory(stats[si][0][1], p)   # frames without a match leave no entry
        out[f"matched{k}"] = np.array(bool(matched))
        if not matched:
            continue
        correct = stats[si][0][0].numpy()
        si += 1
        pp, rr, ap, f1, cls = node.ap_per_class(correct, p[:, 4], p[:, 5], g[:, 5])
        nc = len(cls)
        for name, v in (("p", pp), ("r", rr), ("f1", f1), ("ap_class", cls)):      # what calculate_metrics itself observed
            got = np.array(rec[name][pos[name]:pos[name] + nc])
            pos[name] += nc
            assert np.array_equal(got, v), name
        out[f"correct{k}"], out[f"p{k}"], out[f"r{k}"], out[f"ap{k}"], out[f"f1{k}"], out[f"cls{k}"] = correct, pp, rr, ap, f1, cls
        print(f"eval frame {k}: G={len(g)} D={len(p)} TP@0.5={int(correct[:, 0].sum())} classes={cls.tolist()} mAP50={ap[:, 0].mean():.4f}")
    assert si == len(stats)
    save("eval", **out)


def spread_scores(logits, rng, thresh=0.1, rel_gap=4e-6, fresh=lambda rng, n: rng.normal(-5.0, 1.2, n)):
    """Re-draws logits until the sigmoid scores >= thresh (as torch computes them) are pairwise more than rel_gap apart,
    so that neither torch.topk's unspecified tie order nor the ulps between two exp() implementations can change an
    order the golden depends on."""
    x = logits.reshape(-1).astype(np.float32)
    for _ in range(200):
        s = torch.sigmoid(torch.from_numpy(x)).numpy()
        cand = np.nonzero(s >= thresh * 0.98)[0]
        o = cand[np.argsort(s[cand], kind="stable")]
        close = np.nonzero(np.diff(s[o]) <= rel_gap * s[o][1:])[0]
        if close.size == 0:
            return x.reshape(logits.shape)
        x[o[close]] = fresh(rng, close.size).astype(np.float32)
    raise RuntimeError("could not separate the scores")


def golden_utils_preprocess():
    """utils/preprocess.py::image_adjust (:127-175) run unmodified. cv2 path (pil=False, the reachable one): a PNG on
    disk -> imread -> resize(w,h) -> BGR2RGB -> CHW -> /255, for every `scaling` value (the cv2 path ignores it).
    PIL path (:133-167): unreachable as shipped, because `Image` is rebound to sensor_msgs.msg.Image (:7 vs :10); it is
    run here with that one global bound back to PIL.Image — the module's own first import — on frames already at
    (w,h), so that Pillow's resize is the identity and what is pinned is the scaling arithmetic of :147-159 and the
    NCHW/NHWC choice of :161-165."""
    import tempfile
    import types
    import cv2
    import PIL.Image
    ref_harness.communicator()                       # installs the rospy / sensor_msgs / cv_bridge / tritonclient placeholders
    up = ref_harness.load_file("ref_utils_preprocess", "utils/preprocess.py")
    up.mc.ModelInput = types.SimpleNamespace(FORMAT_NCHW=2, FORMAT_NHWC=1, FORMAT_NONE=0)
    out = {}
    cases = [(synth.image_smooth(180, 320, synth.seed_for(1, 50)), 160, 96), (synth.image_noise(97, 131, synth.seed_for(1, 51)), 64, 48),
             (synth.image_noise(120, 200, synth.seed_for(1, 52)), 100, 60)]
    with tempfile.TemporaryDirectory() as d:
        for k, (bgr, w, h) in enumerate(cases):
            path = os.path.join(d, f"f{k}.png")
            cv2.imwrite(path, bgr)
            res = [up.image_adjust(path, None, "FP32", 3, h, w, sc, pil=False) for sc in ("NONE", "INCEPTION", "VGG", "COCO")]
            assert all(np.array_equal(res[0], r) for r in res)
            out[f"cv_bgr{k}"], out[f"cv_wh{k}"], out[f"cv_out{k}"] = bgr, np.array([w, h]), res[0]
    saved = up.Image
    up.Image = PIL.Image
    try:
        rgb = synth.image_noise(48, 64, synth.seed_for(1, 53))
        out["pil_rgb"] = rgb
        for sc in ("NONE", "INCEPTION", "VGG", "COCO"):
            for fmt, tag in ((2, "nchw"), (1, "nhwc")):
                r = up.image_adjust(PIL.Image.fromarray(rgb), fmt, "FP32", 3, 48, 64, sc, pil=True)
                assert r.dtype == np.float32
                out[f"pil_{sc}_{tag}"] = r
    finally:
        up.Image = saved
    save("utils_preprocess", **out)


def golden_dynamic_vfe():
    """DynamicPillarVFE.forward (OPCD/pcdet/models/backbones_3d/vfe/dynamic_pillar_vfe.py:90-142) run unmodified up to the
    PFN layers (pfn_layers emptied, as golden_pillar_vfe does) on a batch of two small nuScenes-like clouds; torch_scatter
    is the stand-in of ref_harness.install_torch_scatter, `.cuda()` is a no-op, and torch.unique is wrapped only to record
    the inverse index the function keeps local."""
    ref_harness.install_torch_scatter()
    mod = ref_harness.pcdet("models.backbones_3d.vfe.dynamic_pillar_vfe")
    rng_, vs = [-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8.0]
    a = synth.lidar_multisweep(synth.seed_for(5, 60), sweeps=2, beams=8, azimuths=300)
    b = synth.lidar_multisweep(synth.seed_for(5, 61), sweeps=3, beams=6, azimuths=250)
    pts = np.concatenate([np.concatenate([np.zeros((len(a), 1), np.float32), a], 1), np.concatenate([np.ones((len(b), 1), np.float32), b], 1)])
    pts[5, 1] = 1e3                                   # out of range rows
    pts[77, 2] = -51.3
    cfg = ref_harness._Cfg(USE_NORM=True, WITH_DISTANCE=False, USE_ABSLOTE_XYZ=True, NUM_FILTERS=[64])
    with ref_harness.no_cuda():
        vfe = mod.DynamicPillarVFE(cfg, 5, vs, [512, 512, 1], rng_)
        vfe.pfn_layers = torch.nn.ModuleList([])
        seen = {}
        real_unique = torch.unique

        def spy(*a, **k):
            r = real_unique(*a, **k)
            seen["inv"] = r[1]
            return r
        torch.unique = spy
        try:
            bd = vfe.forward({"points": torch.from_numpy(pts)})
        finally:
            torch.unique = real_unique
    feats = bd["pillar_features"].numpy()
    assert feats.shape[1] == 11 and feats.shape[0] < len(pts)
    save("dynamic_vfe", points=pts, range=np.array(rng_, np.float64), vsize=np.array(vs, np.float64), grid=np.array([512, 512, 1]),
         feats=feats, unq_inv=seen["inv"].numpy(), voxel_coords=bd["voxel_coords"].numpy())


def golden_pcdet_post():
    """The pcdet glue that runs server-side behind the 3D clients, unmodified, on batches of two frames:
    (pp)  AnchorHeadTemplate.generate_anchors + generate_predicted_boxes (anchor_head_template.py:37-52, 225-272) ->
          Detector3DTemplate.post_processing (detector3d_template.py:178-284) -> model_nms_utils.class_agnostic_nms (:6-25)
          -> iou3d_nms_utils.nms_gpu (:84-99), with the compiled extension replaced by ref_harness.install_iou3d_nms_cuda
          (the reference's own iou_bev + its host sweep); the pointpillar.yaml head on a 48x48 map;
    (cp)  CenterHead.generate_predicted_boxes (center_head.py:253-304): two heads (1 + 2 classes) on a 32x32 map;
    (nms) class_agnostic_nms alone, NMS_TYPE nms_gpu and nms_normal_gpu, with and without score_thresh."""
    C = ref_harness._Cfg
    ref_harness.install_iou3d_nms_cuda()
    rng = np.random.default_rng(4040)
    out = {}
    with ref_harness.no_cuda():
        aht = ref_harness.pcdet("models.dense_heads.anchor_head_template")
        det = ref_harness.pcdet("models.detectors.detector3d_template")
        ch = ref_harness.pcdet("models.dense_heads.center_head")
        nmsu = ref_harness.pcdet("models.model_utils.model_nms_utils")
        bcu = ref_harness.pcdet("utils.box_coder_utils")
        # ---- (pp)
        B, H, W = 2, 48, 48
        pc_range = np.array([0, -7.68, -3, 15.36, 7.68, 1], np.float32)
        gen_cfg = [C(class_name=n, anchor_sizes=[s], anchor_rotations=[0, 1.57], anchor_bottom_heights=[b], align_center=False, feature_map_stride=2,
                     matched_threshold=0.6, unmatched_threshold=0.45)
                   for n, s, b in (("Car", [3.9, 1.6, 1.56], -1.78), ("Pedestrian", [0.8, 0.6, 1.73], -0.6), ("Cyclist", [1.76, 0.6, 1.73], -0.6))]
        head = aht.AnchorHeadTemplate.__new__(aht.AnchorHeadTemplate)
        torch.nn.Module.__init__(head)
        head.model_cfg = C(DIR_OFFSET=0.78539, DIR_LIMIT_OFFSET=0.0, NUM_DIR_BINS=2)
        head.use_multihead = False
        head.box_coder = bcu.ResidualCoder()
        anchors, head.num_anchors_per_location = aht.AnchorHeadTemplate.generate_anchors(gen_cfg, grid_size=np.array([96, 96, 1]),
                                                                                         point_cloud_range=pc_range, anchor_ndim=7)
        head.anchors = [x.cuda() for x in anchors]
        cls = rng.normal(-5.0, 1.2, (B, H, W, 6, 3))
        for b in range(B):
            for _ in range(30):
                y, x, k, c = rng.integers(2, H - 2), rng.integers(2, W - 2), rng.integers(0, 6), rng.integers(0, 3)
                cls[b, y - 1:y + 2, x - 1:x + 2, k, c] += rng.uniform(4, 8)
        cls = spread_scores(np.clip(cls, -8.9, 6.0), rng).reshape(B, H, W, 18)
        box = rng.normal(0, 0.3, (B, H, W, 42)).astype(np.float32)
        dirp = synth._quantise(rng.normal(0, 1, (B, H, W, 12)))
        bc, bb = head.generate_predicted_boxes(B, torch.from_numpy(cls), torch.from_numpy(box), torch.from_numpy(dirp))
        d = det.Detector3DTemplate.__new__(det.Detector3DTemplate)
        torch.nn.Module.__init__(d)
        d.num_class = 3
        d.model_cfg = C(POST_PROCESSING=C(RECALL_THRESH_LIST=[0.3, 0.5, 0.7], SCORE_THRESH=0.1, OUTPUT_RAW_SCORE=False, EVAL_METRIC="kitti",
                                          NMS_CONFIG=C(MULTI_CLASSES_NMS=False, NMS_TYPE="nms_gpu", NMS_THRESH=0.01, NMS_PRE_MAXSIZE=4096, NMS_POST_MAXSIZE=500)))
        preds, _ = d.post_processing({"batch_size": B, "batch_cls_preds": bc, "batch_box_preds": bb, "cls_preds_normalized": False})
        out["pp_range"], out["pp_cls"], out["pp_box"], out["pp_dir"] = pc_range, cls, box, dirp
        for b in range(B):
            assert 20 < preds[b]["pred_boxes"].shape[0] <= 500
            out[f"pp_boxes{b}"], out[f"pp_scores{b}"], out[f"pp_labels{b}"] = (preds[b][k].numpy() for k in ("pred_boxes", "pred_scores", "pred_labels"))
        # ---- (cp)
        hd = ch.CenterHead.__new__(ch.CenterHead)
        torch.nn.Module.__init__(hd)
        hd.model_cfg = C(POST_PROCESSING=C(SCORE_THRESH=0.1, POST_CENTER_LIMIT_RANGE=[-61.2, -61.2, -10.0, 61.2, 61.2, 10.0], MAX_OBJ_PER_SAMPLE=60,
                                           NMS_CONFIG=C(NMS_TYPE="nms_gpu", NMS_THRESH=0.2, NMS_PRE_MAXSIZE=1000, NMS_POST_MAXSIZE=30)))
        hd.separate_head_cfg = C(HEAD_ORDER=["center", "center_z", "dim", "rot", "vel"])
        hd.point_cloud_range, hd.voxel_size, hd.feature_map_stride = [-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8.0], 4
        hd.class_id_mapping_each_head = [torch.tensor([0]), torch.tensor([1, 2])]
        h = w = 32
        yy, xx = np.mgrid[0:h, 0:w]
        pred_dicts = []
        for hi, c in enumerate((1, 2)):
            hm = rng.normal(-6, 1, (B, c, h, w))
            for b in range(B):
                for _ in range(14):
                    k, py, px = rng.integers(0, c), rng.uniform(0, h), rng.uniform(0, w)
                    hm[b, k] += 8.0 * np.exp(-((yy - py) ** 2 + (xx - px) ** 2) / (2 * 1.5 ** 2))
            hm = spread_scores(np.clip(hm, -8.9, 6.0), rng, thresh=0.02, fresh=lambda r, n: r.normal(-6, 1, n))
            ang = rng.uniform(-np.pi, np.pi, (B, h, w))
            pd = {"hm": hm, "center": rng.uniform(0, 1, (B, 2, h, w)), "center_z": rng.normal(-1, 0.5, (B, 1, h, w)),
                  "dim": rng.normal(0.7, 0.3, (B, 3, h, w)), "rot": np.stack([np.cos(ang), np.sin(ang)], 1), "vel": rng.normal(0, 2, (B, 2, h, w))}
            pd = {k: torch.from_numpy(np.ascontiguousarray(v, np.float32)) for k, v in pd.items()}
            pred_dicts.append(pd)
            for k, v in pd.items():
                out[f"cp{hi}_{k}"] = v.numpy()
        ret = hd.generate_predicted_boxes(B, pred_dicts)
        for b in range(B):
            assert ret[b]["pred_boxes"].shape[0] > 8
            out[f"cp_boxes{b}"], out[f"cp_scores{b}"], out[f"cp_labels{b}"] = (ret[b][k].numpy() for k in ("pred_boxes", "pred_scores", "pred_labels"))
        # ---- (nms)
        boxes = synth.random_boxes7(1500, 909, extent=25.0)
        scores = torch.sigmoid(torch.from_numpy(spread_scores(rng.normal(-1, 2, 1500), rng, thresh=0.0, fresh=lambda r, n: r.normal(-1, 2, n)))).numpy()
        out["nms_boxes"], out["nms_scores"] = boxes, scores
        for tag, typ, thr, st in (("rot", "nms_gpu", 0.1, 0.2), ("rot_all", "nms_gpu", 0.3, None), ("normal", "nms_normal_gpu", 0.1, 0.2)):
            cfgn = C(NMS_TYPE=typ, NMS_THRESH=thr, NMS_PRE_MAXSIZE=1000, NMS_POST_MAXSIZE=200)
            sel, ssc = nmsu.class_agnostic_nms(torch.from_numpy(scores), torch.from_numpy(boxes), cfgn, score_thresh=st)
            assert 30 < sel.numel() <= 200
            out[f"nms_{tag}_cfg"] = np.array([thr, -1.0 if st is None else st, 1000, 200], np.float64)
            out[f"nms_{tag}_sel"], out[f"nms_{tag}_scores"] = sel.numpy(), ssc.numpy()
        # ---- (multi) multi_classes_nms (model_nms_utils.py:28-66): per-class score cut / top-k / NMS / post cut, three classes
        mc_scores = torch.sigmoid(torch.from_numpy(np.stack([spread_scores(rng.normal(-1.5, 2, 1500), rng, thresh=0.0, fresh=lambda r, n: r.normal(-1.5, 2, n))
                                                             for _ in range(3)], 1))).numpy().astype(np.float32)
        mc_boxes = np.concatenate([boxes, rng.normal(0, 1, (1500, 2)).astype(np.float32)], 1)          # 7 + 2 extra columns (velocities) ride along
        out["mc_scores"], out["mc_boxes"] = mc_scores, mc_boxes
        for tag, typ, thr, st in (("rot", "nms_gpu", 0.2, 0.3), ("normal", "nms_normal_gpu", 0.1, None)):
            cfgn = C(NMS_TYPE=typ, NMS_THRESH=thr, NMS_PRE_MAXSIZE=400, NMS_POST_MAXSIZE=60)
            ps, pl, pb = nmsu.multi_classes_nms(torch.from_numpy(mc_scores), torch.from_numpy(mc_boxes), cfgn, score_thresh=st)
            assert 40 < ps.numel() <= 180 and pl.unique().numel() == 3
            out[f"mc_{tag}_cfg"] = np.array([thr, -1.0 if st is None else st, 400, 60], np.float64)
            out[f"mc_{tag}_scores"], out[f"mc_{tag}_labels"], out[f"mc_{tag}_boxes"] = ps.numpy(), pl.numpy(), pb.numpy()
    save("pcdet_post", **out)


def golden_low_score_nu():
    """PointPillarPostprocess.remove_low_score_nu / get_annotations_indices (clients/postprocess/detector_3d_postprocess.py
    :55-64, :98-133) run unmodified (jsk_recognition_msgs is a placeholder: its classes are only named in an import)."""
    c = ref_harness.clients()
    post = c.PointPillarPostprocess()
    rng = np.random.default_rng(31)
    n = 400
    labels = rng.integers(0, 10, n).astype(np.int64)
    scores = (np.round(rng.uniform(0, 1, n) * 20) / 20).astype(np.float32)      # many values exactly on a threshold (>=)
    boxes = rng.normal(0, 10, (n, 9)).astype(np.float32)
    res = post.remove_low_score_nu({"box3d_lidar": boxes, "scores": scores, "labels": labels, "metadata": {"token": "x"}}, 0.45)
    assert "metadata" not in res and 50 < len(res["scores"]) < n
    idx = post.get_annotations_indices(6, 0.15, labels, scores)
    save("low_score_nu", boxes=boxes, scores=scores, labels=labels, out_boxes=res["box3d_lidar"], out_scores=res["scores"], out_labels=res["labels"],
         motorcycle_idx=np.array(idx, np.int64))


golden_overlay.__doc__ = golden_overlay.__doc__ % np.__version__


def main():
    assert ref_harness.available(), "reference not mounted"
    os.makedirs(GOLD, exist_ok=True)
    which = sys.argv[1:] or [n[len("golden_"):] for n in globals() if n.startswith("golden_")]
    for w in which:
        globals()["golden_" + w]()


if __name__ == "__main__":
    torch.set_num_threads(1)
    main()
