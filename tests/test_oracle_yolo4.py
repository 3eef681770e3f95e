"""Oracle pinning for the YOLOv4 head decode / two-tensor postprocess against golden outputs of the reference's own
tools/yolo_layer.py and utils/postprocess.py (oracle/make_golden.py: golden_yolo4)."""
import numpy as np
import torch

from conftest import golden
from oracle import yolo_head

NC = 4
MASKS = [([0, 1, 2], 8), ([3, 4, 5], 16), ([6, 7, 8], 32)]


def _anchors(g, mask, stride):
    a = g["anchors"].tolist()
    m = []
    for k in mask:
        m += a[k * 2:(k + 1) * 2]
    return [v / stride for v in m]


def test_golden_yolo_forward_dynamic():
    g = golden("yolo4")
    for hi, (mask, stride) in enumerate(MASKS):
        b, c = yolo_head.yolo_forward_dynamic(g[f"head{hi}"], NC, _anchors(g, mask, stride), 3)
        assert np.array_equal(b.numpy(), g[f"boxes{hi}"]) and np.array_equal(c.numpy(), g[f"confs{hi}"])


def test_golden_extract_boxes_onnx():
    g = golden("yolo4")
    boxes = np.concatenate([g[f"boxes{i}"] for i in range(3)], 1)
    confs = np.concatenate([g[f"confs{i}"] for i in range(3)], 1)
    res = yolo_head.extract_boxes_onnx(0.2, 0.4, [boxes, confs])
    for i, rows in enumerate(res):
        assert np.array_equal(np.array(rows, np.float32).reshape(-1, 7), g[f"post{i}"])
        assert g[f"post{i}"].shape[0] > 5


def test_yolov5_detect_shapes():
    rng = np.random.default_rng(0)
    heads = [rng.normal(0, 1, (1, 3 * 7, s, s)).astype(np.float32) for s in (8, 4, 2)]
    out = yolo_head.yolov5_detect(heads, [[10, 13, 16, 30, 33, 23], [30, 61, 62, 45, 59, 119], [116, 90, 156, 198, 373, 326]], [8, 16, 32], 2)
    assert out.shape == (1, 3 * (64 + 16 + 4), 7) and float(out[..., 4:].min()) > 0
