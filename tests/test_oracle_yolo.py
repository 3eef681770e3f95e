"""Oracle pinning for the YOLOv5 postprocess: golden outputs of the unmodified reference and the numpy
restatement of torchvision's CPU NMS against torchvision itself."""
import numpy as np
import pytest
import torch
import torchvision

from conftest import golden
from oracle import ref_harness, yolo_post
from triton_client_b200 import synth

CASES = 4


def _cfg(g, k):
    conf, iou, agn, max_det = g[f"cfg{k}"]
    classes = g[f"classes{k}"].tolist() or None
    return dict(conf_thres=float(conf), iou_thres=float(iou), classes=classes, agnostic=bool(agn), max_det=int(max_det))


@pytest.mark.parametrize("as_written", [False, True])
def test_golden_extract_boxes(as_written):
    g = golden("yolo_post")
    for k in range(CASES):
        pred = g[f"pred{k}"]
        res = yolo_post.extract_boxes(yolo_post.FakeResponse([pred]), as_written=as_written, **_cfg(g, k))
        assert isinstance(res, list) and len(res) == pred.shape[0]
        for bi, t in enumerate(res):
            want = g[f"out{k}_{bi}"]
            assert t.dtype == torch.float64 or t.shape[0] == 0
            assert np.array_equal(t.numpy().astype(np.float64).reshape(-1, 6), want.reshape(-1, 6)), (k, bi)


def test_golden_empty_returns_assertion():
    g = golden("yolo_post")
    res = yolo_post.extract_boxes(np.zeros((1, 64, 7), np.float32), conf_thres=0.3)
    assert isinstance(res, AssertionError) and str(res) == str(g["empty_msg"])


def test_restated_nms_equals_torchvision():
    rng = np.random.default_rng(5)
    for n in [0, 1, 2, 50, 400]:
        xy = rng.uniform(0, 200, (n, 2))
        wh = rng.uniform(5, 80, (n, 2))
        boxes = np.concatenate([xy, xy + wh], 1)
        scores = np.round(rng.uniform(0, 1, n) * 64) / 64  # many ties
        for thr in [0.0, 0.3, 0.45, 1.0]:
            a = yolo_post.nms_torchvision_restated(boxes, scores, thr)
            b = torchvision.ops.nms(torch.from_numpy(boxes), torch.from_numpy(scores), thr).numpy()
            assert np.array_equal(a, b), (n, thr)


def test_extract_boxes_with_restated_nms_is_identical():
    pred = synth.yolo_output(2, 800, 4, 99, objects=10, rows_per_object=15, tie_frac=0.3)
    a = yolo_post.extract_boxes(pred, conf_thres=0.3, as_written=False)
    b = yolo_post.extract_boxes(pred, conf_thres=0.3, as_written=False, nms="restated")
    for x, y in zip(a, b):
        assert torch.equal(x, y)


def test_nms_cpu_matches_stable_variant_on_distinct_scores():
    rng = np.random.default_rng(2)
    xy = rng.uniform(0, 100, (60, 2))
    boxes = np.concatenate([xy, xy + rng.uniform(5, 40, (60, 2))], 1).astype(np.float32)
    conf = rng.permutation(60).astype(np.float32) / 60
    assert np.array_equal(yolo_post.nms_cpu(boxes, conf, 0.5), yolo_post.nms_cpu(boxes, conf, 0.5, stable=True))


@pytest.mark.skipif(not ref_harness.available(), reason="reference not mounted (GPU box)")
def test_against_live_reference():
    c = ref_harness.clients()
    post = c.Yolov5client().get_postprocess()
    pred = synth.yolo_output(1, 400, 3, 123, objects=6, rows_per_object=10, tie_frac=0.3)
    ref = post.extract_boxes(yolo_post.FakeResponse([pred]), conf_thres=0.3)
    got = yolo_post.extract_boxes(pred, conf_thres=0.3, as_written=False)
    assert len(ref) == len(got) and all(torch.equal(a, b) for a, b in zip(ref, got))
