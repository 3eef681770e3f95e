"""Oracle pinning for the 3D postprocess: the torch/C restatements against golden outputs of the reference's own
functions (oracle/make_golden.py: ResidualCoder.decode_torch, limit_period, AnchorGenerator, decode_bbox_from_heatmap)
and against the reference's own compiled iou3d_cpu.cpp (oracle/_ref) for the rotated IoU."""
import ctypes
import os

import numpy as np
import pytest
import torch

from conftest import ROOT, golden
from oracle import box3d
from triton_client_b200 import synth

KITTI_ANCHORS = dict(anchor_range=[0, -39.68, -3, 69.12, 39.68, 1],
                     anchor_sizes=[[3.9, 1.6, 1.56], [0.8, 0.6, 1.73], [1.76, 0.6, 1.73]],
                     anchor_bottoms=[-1.78, -0.6, -0.6], anchor_rotations=[[0, 1.57]] * 3)
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libiou3d_ref.so")


def test_golden_decode_and_limit_period():
    g = golden("box3d")
    dec = box3d.decode_torch(torch.from_numpy(g["dec_enc"]), torch.from_numpy(g["dec_anchors"])).numpy()
    assert np.array_equal(dec, g["dec_out"])
    lp = box3d.limit_period(torch.from_numpy(g["lp_val"]) - 0.78539, 0.0, 2 * np.pi / 2).numpy()
    assert np.array_equal(lp, g["lp_out"])


def test_golden_anchors():
    g = golden("box3d")
    a = box3d.generate_anchors((216, 248), KITTI_ANCHORS["anchor_range"], KITTI_ANCHORS["anchor_sizes"],
                               KITTI_ANCHORS["anchor_rotations"], KITTI_ANCHORS["anchor_bottoms"]).numpy()
    assert a.shape == (321408, 7)
    np.testing.assert_allclose(a[g["anchor_idx"]][:, :2], g["anchor_val"][:, :2], rtol=1e-6, atol=1e-6)  # SIMD-width dependent ulp
    assert np.array_equal(a[g["anchor_idx"]][:, 2:], g["anchor_val"][:, 2:])
    np.testing.assert_allclose(np.stack([a[0], a[-1]]), g["anchor_first_last"], rtol=1e-6)


def test_golden_centerpoint_decode():
    g = golden("box3d")
    t = {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("cp_")}
    res = box3d.decode_bbox_from_heatmap(heatmap=t["hm"], rot_cos=t["rot_cos"], rot_sin=t["rot_sin"], center=t["center"], center_z=t["center_z"],
                                         dim=t["dim"], vel=t["vel"], point_cloud_range=[-51.2, -51.2, -5.0, 51.2, 51.2, 3.0],
                                         voxel_size=[0.2, 0.2, 8], feature_map_stride=4, K=40, score_thresh=0.1,
                                         post_center_limit_range=torch.tensor([-61.2, -61.2, -10.0, 61.2, 61.2, 10.0]))[0]
    assert np.array_equal(res["pred_boxes"].numpy(), g["cp_boxes"])
    assert np.array_equal(res["pred_scores"].numpy(), g["cp_scores"])
    assert np.array_equal(res["pred_labels"].numpy(), g["cp_labels"])


@pytest.mark.skipif(not os.path.exists(REF_SO), reason="oracle/_ref not built (reference was not mounted at build time)")
def test_rotated_iou_restatement_equals_reference_binary():
    ref = ctypes.CDLL(REF_SO)
    a, b = synth.random_boxes7(300, 5, extent=20), synth.random_boxes7(400, 6, extent=20)
    b[:40] = a[:40]
    b[40:80, :2] += 0.01
    a[80:100, 6] = 0
    b[80:100, 6] = np.pi / 2
    want = np.zeros((300, 400), np.float32)
    P = ctypes.c_void_p
    ref.ref_boxes_iou_bev_cpu(P(a.ctypes.data), 300, P(b.ctypes.data), 400, P(want.ctypes.data))
    got = box3d.boxes_iou_bev(a, b)
    assert (want > 0).sum() > 5000
    assert np.array_equal(got, want)


def test_rotated_nms_basics():
    boxes = np.array([[0, 0, 0, 4, 2, 1, 0], [0.2, 0.1, 0, 4, 2, 1, 0.1], [10, 10, 0, 4, 2, 1, 1.0], [0, 0, 0, 4, 2, 1, 0]], np.float32)
    scores = np.array([0.5, 0.9, 0.3, 0.5], np.float32)
    keep = box3d.nms_gpu(boxes, scores, 0.1).tolist()
    assert keep == [1, 2]
    iou = box3d.boxes_iou_bev(boxes[:1], boxes[3:])
    assert abs(iou[0, 0] - 1.0) < 1e-5
    sel, sc = box3d.class_agnostic_nms(torch.from_numpy(scores), torch.from_numpy(boxes), 0.1, 4096, 500, score_thresh=0.4)
    assert sel.tolist() == [1] and sc.tolist() == pytest.approx([0.9])


def test_remove_low_score_nu_order_and_omission():
    labels = np.array([0, 6, 8, 1, 0, 8, 3, 6], np.int64)
    scores = np.array([0.5, 0.9, 0.1, 0.45, 0.3, 0.05, 0.3, 0.2], np.float32)
    out = box3d.remove_low_score_nu({"labels": labels, "scores": scores, "box3d_lidar": np.arange(8)})
    assert out["box3d_lidar"].tolist() == [0, 2, 6, 3]   # car, pedestrian, bus, truck; motorcycles (6) never appended


def test_full_chain_runs_and_is_deterministic():
    cls, box, dirp = synth.pp_head_outputs(9, h=40, w=36, planted=10)
    r1 = box3d.pp_anchor_post(cls, box, dirp, **KITTI_ANCHORS)
    r2 = box3d.pp_anchor_post(cls, box, dirp, **KITTI_ANCHORS)
    assert r1["pred_boxes"].shape[0] > 0 and torch.equal(r1["pred_boxes"], r2["pred_boxes"])
    assert r1["pred_labels"].min() >= 1 and r1["pred_labels"].max() <= 3


PP_SMALL = dict(anchor_sizes=KITTI_ANCHORS["anchor_sizes"], anchor_bottoms=KITTI_ANCHORS["anchor_bottoms"],
                anchor_rotations=KITTI_ANCHORS["anchor_rotations"])


def _same_dets(got, g, tag, b, exact_cols=None):
    """labels and row order exact; float columns exact where the restatement repeats torch's CPU arithmetic."""
    assert np.array_equal(got["pred_labels"].numpy(), g[f"{tag}_labels{b}"])
    assert np.array_equal(got["pred_scores"].numpy(), g[f"{tag}_scores{b}"])
    bx, want = got["pred_boxes"].numpy(), g[f"{tag}_boxes{b}"]
    cols = [c for c in range(want.shape[1]) if c != 6]
    assert bx.shape == want.shape and np.array_equal(bx[:, cols], want[:, cols])
    np.testing.assert_allclose(bx[:, 6], want[:, 6], rtol=2e-7, atol=1e-7)     # torch's vectorised atan2 depends on the tensor's position in the batch (1 ulp)


def test_golden_pointpillars_glue_batch():
    """generate_predicted_boxes -> post_processing -> class_agnostic_nms -> nms_gpu run unmodified on a batch of two
    (oracle/make_golden.py: golden_pcdet_post) vs the restated chain, sample by sample."""
    g = golden("pcdet_post")
    res = box3d.pp_anchor_post_batch(g["pp_cls"], g["pp_box"], g["pp_dir"], g["pp_range"], **PP_SMALL)
    for b in range(2):
        _same_dets(res[b], g, "pp", b)


def test_golden_centerhead_glue_batch():
    g = golden("pcdet_post")
    heads = [{k: g[f"cp{h}_{k}"] for k in ("hm", "center", "center_z", "dim", "rot", "vel")} for h in range(2)]
    res = box3d.centerhead_generate_predicted_boxes_batch(heads, [[0], [1, 2]], [-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8.0], 4,
                                                          [-61.2, -61.2, -10.0, 61.2, 61.2, 10.0], K=60, score_thresh=0.1, nms_thresh=0.2,
                                                          pre_max=1000, post_max=30)
    for b in range(2):
        _same_dets(res[b], g, "cp", b)


@pytest.mark.parametrize("tag,nms_type", [("rot", "nms_gpu"), ("rot_all", "nms_gpu"), ("normal", "nms_normal_gpu")])
def test_golden_class_agnostic_nms(tag, nms_type):
    g = golden("pcdet_post")
    thr, st, pre, post = g[f"nms_{tag}_cfg"]
    sel, sc = box3d.class_agnostic_nms(torch.from_numpy(g["nms_scores"]), torch.from_numpy(g["nms_boxes"]), float(thr), int(pre), int(post),
                                       None if st < 0 else float(st), nms_type)
    assert np.array_equal(sel.numpy(), g[f"nms_{tag}_sel"]) and np.array_equal(sc.numpy(), g[f"nms_{tag}_scores"])


@pytest.mark.parametrize("tag,nms_type", [("rot", "nms_gpu"), ("normal", "nms_normal_gpu")])
def test_golden_multi_classes_nms(tag, nms_type):
    g = golden("pcdet_post")
    thr, st, pre, post = g[f"mc_{tag}_cfg"]
    ps, pl, pb = box3d.multi_classes_nms(g["mc_scores"], g["mc_boxes"], float(thr), int(pre), int(post), None if st < 0 else float(st), nms_type)
    assert np.array_equal(ps.numpy(), g[f"mc_{tag}_scores"]) and np.array_equal(pl.numpy(), g[f"mc_{tag}_labels"])
    assert np.array_equal(pb.numpy(), g[f"mc_{tag}_boxes"])


def test_golden_remove_low_score_nu():
    g = golden("low_score_nu")
    res = box3d.remove_low_score_nu({"box3d_lidar": g["boxes"], "scores": g["scores"], "labels": g["labels"], "metadata": {}})
    assert np.array_equal(res["box3d_lidar"], g["out_boxes"]) and np.array_equal(res["scores"], g["out_scores"])
    assert np.array_equal(res["labels"], g["out_labels"])
    assert box3d.get_annotations_indices(6, 0.15, g["labels"], g["scores"]) == g["motorcycle_idx"].tolist()


def test_fma32_is_the_correctly_rounded_fused_multiply_add():
    """box3d.fma32 (what iou_normal's `Sa + Sb` contraction needs) against exact rational arithmetic, including sums that cancel."""
    from fractions import Fraction
    rng = np.random.default_rng(0)
    a = rng.normal(0, 3, 3000).astype(np.float32)
    b = rng.normal(0, 3, 3000).astype(np.float32)
    c = (-(a * b) + rng.normal(0, 1e-6, 3000)).astype(np.float32)
    c[::3] = rng.normal(0, 1e4, len(c[::3])).astype(np.float32)
    got = box3d.fma32(a, b, c)
    for i in range(0, 3000, 3):
        ex = Fraction(float(a[i])) * Fraction(float(b[i])) + Fraction(float(c[i]))
        v = np.float32(float(ex))
        cands = [np.nextafter(v, np.float32(-np.inf)), v, np.nextafter(v, np.float32(np.inf))]
        best = min(cands, key=lambda t: (abs(Fraction(float(t)) - ex), int(np.float32(t).view(np.uint32)) & 1))
        assert best == got[i], (i, a[i], b[i], c[i])


def test_nms_type_names_resolve_like_getattr_on_iou3d_nms_utils():
    from triton_client_b200.gpu import _nms_type
    assert _nms_type("nms_gpu") == 0 and _nms_type("nms_normal_gpu") == 1
    with pytest.raises(AttributeError):
        _nms_type("nms_cpu")
