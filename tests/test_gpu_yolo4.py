"""GPU parity: YOLO raw-head decode (v4 vs the reference's golden vectors, v5 vs the restatement), YOLOv4 two-tensor
postprocess, numpy-style NMS and the IoU matrix, through the C ABI."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import yolo_head, yolo_post

pytestmark = pytest.mark.gpu

NC = 4
MASKS = [([0, 1, 2], 8), ([3, 4, 5], 16), ([6, 7, 8], 32)]


def _anchors(g, mask, stride):
    a = g["anchors"].tolist()
    m = []
    for k in mask:
        m += a[k * 2:(k + 1) * 2]
    return [v / stride for v in m]


def test_v4_decode_golden(engine):
    g = golden("yolo4")
    heads = [torch.from_numpy(g[f"head{i}"]).cuda() for i in range(3)]
    boxes, confs = engine.yolo_head_decode_v4(heads, [_anchors(g, m, s) for m, s in MASKS], NC)
    want_b = np.concatenate([g[f"boxes{i}"] for i in range(3)], 1)
    want_c = np.concatenate([g[f"confs{i}"] for i in range(3)], 1)
    np.testing.assert_allclose(boxes.cpu().numpy(), want_b, rtol=1e-5, atol=1e-6)   # exp/sigmoid differ by ulps from torch CPU
    np.testing.assert_allclose(confs.cpu().numpy(), want_c, rtol=1e-5, atol=1e-7)


def test_v4_decode_full_size_and_scale(engine):
    rng = np.random.default_rng(1)
    anchors = [12, 16, 19, 36, 40, 28, 36, 75, 76, 55, 72, 146, 142, 110, 192, 243, 459, 401]
    heads, per = [], []
    for mask, stride, hw in [([0, 1, 2], 8, 64), ([3, 4, 5], 16, 32), ([6, 7, 8], 32, 16)]:   # 512x512 -> 16128 rows
        heads.append(rng.normal(0, 1.5, (2, 255, hw, hw)).astype(np.float32))
        per.append([anchors[2 * k + j] / stride for k in mask for j in (0, 1)])
    for scale in (1.0, 1.05):
        want = yolo_head.get_region_boxes([yolo_head.yolo_forward_dynamic(h, 80, a, 3, scale) for h, a in zip(heads, per)])
        b, c = engine.yolo_head_decode_v4([torch.from_numpy(h).cuda() for h in heads], per, 80, scale_x_y=scale)
        assert b.shape == (2, 16128, 1, 4) and c.shape == (2, 16128, 80)
        np.testing.assert_allclose(b.cpu().numpy(), want[0].numpy(), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(c.cpu().numpy(), want[1].numpy(), rtol=1e-5, atol=1e-7)


def test_v5_decode_feeds_yolo_nms(engine):
    rng = np.random.default_rng(2)
    anchors = [[10, 13, 16, 30, 33, 23], [30, 61, 62, 45, 59, 119], [116, 90, 156, 198, 373, 326]]
    heads = [rng.normal(-1, 2, (2, 3 * 7, s, s)).astype(np.float32) for s in (80, 40, 20)]
    want = yolo_head.yolov5_detect(heads, anchors, [8, 16, 32], 2).numpy()
    pred = engine.yolo_head_decode_v5([torch.from_numpy(h).cuda() for h in heads], anchors, [8, 16, 32], 2)
    assert pred.shape == (2, 25200, 7)
    np.testing.assert_allclose(pred.cpu().numpy(), want, rtol=1e-5, atol=1e-5)
    out, cnt, _ = engine.yolo_nms(pred, conf_thres=0.3)
    ref = yolo_post.extract_boxes(pred.cpu().numpy(), conf_thres=0.3, as_written=False)
    for b in range(2):
        assert torch.equal(out[b, :cnt[b]].cpu(), ref[b].double().reshape(-1, 6))


def test_yolo4_post_golden(engine):
    g = golden("yolo4")
    boxes = np.concatenate([g[f"boxes{i}"] for i in range(3)], 1)
    confs = np.concatenate([g[f"confs{i}"] for i in range(3)], 1)
    out, cnt = engine.yolo4_post(torch.from_numpy(boxes).cuda(), torch.from_numpy(confs).cuda(), 0.2, 0.4)
    for i in range(boxes.shape[0]):
        assert np.array_equal(out[i, :int(cnt[0, i])].cpu().numpy(), g[f"post{i}"])
    assert int(cnt[1].sum()) == 0


def test_yolo4_post_full_size_vs_oracle(engine):
    rng = np.random.default_rng(3)
    n, nc = 16128, 80
    xy = rng.uniform(0, 0.9, (2, n, 2))
    wh = rng.uniform(0.02, 0.2, (2, n, 2))
    boxes = np.concatenate([xy, xy + wh], -1).astype(np.float32).reshape(2, n, 1, 4)
    confs = rng.uniform(0, 0.05, (2, n, nc))
    for b in range(2):   # ~900 rows with one confident class each, clustered so that NMS has work to do
        rows = rng.choice(n, 900, replace=False)
        confs[b, rows, rng.integers(0, 12, 900)] = rng.uniform(0.4, 1.0, 900)
        boxes[b, rows, 0, :2] = (rng.integers(0, 30, (900, 2)) / 40.0 + rng.normal(0, 0.004, (900, 2))).astype(np.float32)
        boxes[b, rows, 0, 2:] = boxes[b, rows, 0, :2] + rng.uniform(0.05, 0.1, (900, 2)).astype(np.float32)
    confs = np.round(confs * 256) / 256   # ties -> index order
    want = yolo_head.extract_boxes_onnx(0.4, 0.6, [boxes, confs], stable=True)
    out, cnt = engine.yolo4_post(torch.from_numpy(boxes).cuda(), torch.from_numpy(confs.astype(np.float32)).cuda(), 0.4, 0.6)
    for i in range(2):
        w = np.array(want[i], np.float32).reshape(-1, 7)
        assert w.shape[0] > 50 and w.shape[0] <= 1024
        assert np.array_equal(out[i, :int(cnt[0, i])].cpu().numpy(), w)
    assert int(cnt[1].sum()) == 0


@pytest.mark.parametrize("n", [400, 6000])     # single-CTA path and the wide path
def test_yolo4_post_reports_what_it_cuts(engine, n):
    """The reference is unbounded; the kernel takes 4096 candidates per image into the NMS. When more pass the confidence
    cut it must say so (count[b + i]) and the drop-ins must refuse to return an incomplete result."""
    from triton_client_b200.utils import postprocess as up
    rng = np.random.default_rng(5)
    xy = rng.uniform(0, 0.9, (2, n, 2))
    boxes = np.concatenate([xy, xy + 0.01], -1).astype(np.float32).reshape(2, n, 1, 4)
    confs = rng.uniform(0, 0.05, (2, n, 3)).astype(np.float32)
    confs[1, :, 2] = 0.9                              # image 1: every row passes
    out, cnt = engine.yolo4_post(torch.from_numpy(boxes).cuda(), torch.from_numpy(confs).cuda(), 0.4, 0.6)
    cnt = cnt.cpu().numpy()
    assert cnt[1].tolist() == [0, 1 if n > 4096 else 0] and cnt[0, 0] == 0
    if n > 4096:
        with pytest.raises(OverflowError):
            up.extract_boxes_onnx(0.4, 0.6, [boxes, confs])
    many = np.tile(boxes[0, :, 0], (5000 // n + 1, 1))[:5000]
    with pytest.raises(OverflowError):                # > 4096 boxes into nms_cpu
        up.nms_cpu(many, np.ones(5000, np.float32), 0.5)


@pytest.mark.parametrize("min_mode", [False, True])
def test_nms_axis_aligned_vs_nms_cpu(engine, min_mode):
    rng = np.random.default_rng(4)
    xy = rng.uniform(0, 300, (800, 2))
    boxes = np.concatenate([xy, xy + rng.uniform(5, 80, (800, 2))], 1).astype(np.float32)
    conf = (np.round(rng.uniform(0, 1, 800) * 128) / 128).astype(np.float32)
    want = yolo_post.nms_cpu(boxes, conf, 0.5, min_mode=min_mode, stable=True)
    keep, cnt = engine.nms_axis_aligned(torch.from_numpy(boxes).cuda(), torch.from_numpy(conf).cuda(), 0.5, min_mode)
    assert keep[:int(cnt[0])].cpu().tolist() == want.tolist() and int(cnt[1]) == 0


def test_box_iou_matrix(engine):
    rng = np.random.default_rng(5)
    def mk(n):
        xy = rng.uniform(0, 100, (n, 2))
        return np.concatenate([xy, xy + rng.uniform(1, 50, (n, 2))], 1).astype(np.float32)
    a, b = mk(300), mk(200)
    want = yolo_post.box_iou(torch.from_numpy(a), torch.from_numpy(b)).numpy()
    got = engine.box_iou(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()).cpu().numpy()
    assert np.array_equal(got, want)
