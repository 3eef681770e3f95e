"""GPU parity: YOLOv5 postprocess (filter + decode + ordering + NMS through the C ABI) vs the
reference's golden outputs and the oracle. Bit-exact: keep lists, class ids, float64 rows."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import yolo_post
from triton_client_b200 import synth

pytestmark = pytest.mark.gpu


def _gpu(engine, pred, **kw):
    out, cnt, xc = engine.yolo_nms(torch.from_numpy(pred).cuda(), **kw)
    torch.cuda.synchronize()
    out, cnt = out.cpu().numpy(), cnt.cpu().numpy()
    return [out[b, :cnt[b]] for b in range(pred.shape[0])], xc.cpu().numpy()


def _same(got, want):
    assert len(got) == len(want)
    for b, (g, w) in enumerate(zip(got, want)):
        w = w.numpy().astype(np.float64).reshape(-1, 6) if isinstance(w, torch.Tensor) else np.asarray(w).reshape(-1, 6)
        assert g.shape == w.shape, (b, g.shape, w.shape)
        assert np.array_equal(g, w), (b, np.abs(g - w).max())


def test_golden_reference_vectors(engine):
    g = golden("yolo_post")
    for k in range(4):
        conf, iou, agn, max_det = g[f"cfg{k}"]
        classes = g[f"classes{k}"].tolist() or None
        pred = g[f"pred{k}"]
        got, _ = _gpu(engine, pred, conf_thres=float(conf), iou_thres=float(iou), classes=classes,
                      agnostic=bool(agn), max_det=int(max_det))
        _same(got, [g[f"out{k}_{b}"] for b in range(pred.shape[0])])


# (3, 1003, 2) / (2, 777, 80) / (5, 333, 7): frame starts and last chunks that are not 16-byte aligned (lane copies instead of bulk copies)
@pytest.mark.parametrize("b,n,nc,conf", [(1, 25200, 80, 0.3), (4, 25200, 2, 0.3), (1, 16128, 2, 0.3), (3, 1000, 1, 0.6), (2, 5000, 33, 0.25),
                                         (3, 1003, 2, 0.3), (2, 777, 80, 0.3), (5, 333, 7, 0.25), (1, 31, 2, 0.1)])
def test_matches_oracle_bit_exact(engine, b, n, nc, conf):
    pred = synth.yolo_output(b, n, nc, 500 + n + nc)
    want = yolo_post.extract_boxes(pred, conf_thres=conf, as_written=False)
    got, xc = _gpu(engine, pred, conf_thres=conf)
    _same(got, want)
    assert np.array_equal(xc, (pred[..., 4].astype(np.float64) > conf).sum(1))


@pytest.mark.parametrize("b,n,nc", [(16, 10647, 80), (12, 22743, 1), (40, 1003, 2), (33, 777, 6)])
def test_warps_revisit_buffers_with_mixed_bulk_and_lane_copies(engine, b, n, nc):
    """n*(5+nc) not a multiple of 4 floats: frames alternate between 16-byte aligned (bulk copy) and unaligned (copied
    by the lanes), and b*chunks > 2*warps so every warp comes back to each of its two buffers several times — the
    barrier phase of a buffer must follow its bulk copies only (YOLOv5-416 / -608 COCO heads are such shapes)."""
    pred = synth.yolo_output(b, n, nc, 4242 + n, objects=12, rows_per_object=10)
    want = yolo_post.extract_boxes(pred, conf_thres=0.3, as_written=False)
    for _ in range(2):
        got, xc = _gpu(engine, pred, conf_thres=0.3)
        _same(got, want)
        assert np.array_equal(xc, (pred[..., 4].astype(np.float64) > 0.3).sum(1))


def test_ties_agnostic_filter_and_max_det(engine):
    pred = synth.yolo_output(2, 4000, 6, 77, objects=60, rows_per_object=40, tie_frac=0.6)
    for kw in [dict(conf_thres=0.3, agnostic=True), dict(conf_thres=0.3, classes=[1, 4]), dict(conf_thres=0.3, max_det=7),
               dict(conf_thres=0.3, iou_thres=0.0), dict(conf_thres=0.3, iou_thres=1.0), dict(conf_thres=0.0, iou_thres=0.45)]:
        want = yolo_post.extract_boxes(pred, as_written=False, **kw)
        got, _ = _gpu(engine, pred, **kw)
        _same(got, want)


def test_score_distributions_of_the_counting_order(engine):
    """The in-shared-memory ordering counts inside 1024 score buckets and falls back to the compare-exchange sort when a
    bucket is crowded: equal scores, scores clustered in a sliver of the range, one outlier stretching the range, 1-4
    records per thread — all must give the reference's order."""
    rng = np.random.default_rng(21)
    for n, mode in [(300, "equal"), (900, "equal"), (3000, "sliver"), (1500, "outlier"), (2500, "spread"), (4000, "two")]:
        pred = np.zeros((2, n, 6), np.float32)
        pred[:, :, 0:2] = rng.uniform(0, 640, (2, n, 2))
        pred[:, :, 2:4] = rng.uniform(10, 60, (2, n, 2))
        pred[:, :, 5] = 1.0
        if mode == "equal":
            pred[:, :, 4] = 0.75
        elif mode == "sliver":
            pred[:, :, 4] = 0.5 + rng.uniform(0, 1e-6, (2, n))
        elif mode == "outlier":
            pred[:, :, 4] = 0.5 + rng.uniform(0, 1e-3, (2, n))
            pred[:, 7, 4] = 40.0
        elif mode == "two":
            pred[:, :, 4] = np.where(rng.uniform(0, 1, (2, n)) < 0.5, 0.4, 0.9)
        else:
            pred[:, :, 4] = rng.uniform(0.3, 1.0, (2, n))
        want = yolo_post.extract_boxes(pred, conf_thres=0.25, as_written=False)
        got, _ = _gpu(engine, pred, conf_thres=0.25)
        _same(got, want)


def test_random_shapes_against_the_oracle(engine):
    """Seeded sweep over batch, row count, class count, thresholds, class whitelists and score distributions (the
    counting order, the sort fallbacks, split and unsplit class groups, bulk and lane-copied chunks all get hit)."""
    rng = np.random.default_rng(2024)
    for case in range(24):
        b = int(rng.integers(1, 6))
        n = int(rng.choice([17, 128, 333, 1000, 2049, 4097, 6000]))
        nc = int(rng.choice([1, 2, 3, 7, 20, 80]))
        objects = int(rng.integers(1, 40))
        pred = synth.yolo_output(b, n, nc, 9000 + case, objects=objects, rows_per_object=int(rng.integers(1, 60)),
                                 tie_frac=float(rng.choice([0.0, 0.05, 0.5])))
        if case % 4 == 1:
            pred[..., 4] = np.round(pred[..., 4] * 8) / 8            # few distinct objectness values: crowded score buckets
        if case % 6 == 2:
            pred[..., 0:2] += 3000.0                                   # boxes beyond +-2048: class groups must not split
        kw = dict(conf_thres=float(rng.choice([0.05, 0.25, 0.3, 0.6])), iou_thres=float(rng.choice([0.2, 0.45, 0.7])),
                  agnostic=bool(case % 5 == 3), max_det=int(rng.choice([5, 100, 300, 1000])))
        if case % 7 == 4 and nc > 2:
            kw["classes"] = sorted(set(int(c) for c in rng.integers(0, nc, 3)))
        want = yolo_post.extract_boxes(pred, as_written=False, **kw)
        if isinstance(want, Exception):
            continue
        got, _ = _gpu(engine, pred, **kw)
        _same(got, want)


def test_many_candidates_sorts_in_global_scratch(engine):
    """> 4096 candidates per image: the ordering falls back to global scratch; dense overlapping boxes."""
    rng = np.random.default_rng(8)
    n = 9000
    pred = np.zeros((1, n, 7), np.float32)
    pred[0, :, 0:2] = rng.uniform(100, 400, (n, 2))
    pred[0, :, 2:4] = rng.uniform(20, 90, (n, 2))
    pred[0, :, 4] = rng.uniform(0.5, 1, n)
    pred[0, :, 5:] = rng.uniform(0.6, 1, (n, 2))
    want = yolo_post.extract_boxes(pred, conf_thres=0.25, as_written=False)
    got, _ = _gpu(engine, pred, conf_thres=0.25)
    _same(got, want)


def test_class_groups_and_their_fallbacks(engine):
    """Class-aware NMS runs as independent per-class-group CTAs merged by score. Frames where that split is not exact
    are done whole: boxes reaching beyond +-2048 px (classes 4096 apart can then overlap and suppress each other, as
    they do in the reference), and frames with more candidates than max_nms (the reference keeps the max_nms best over
    all classes first)."""
    pred = synth.yolo_output(3, 3000, 5, 91, objects=30, rows_per_object=25, tie_frac=0.4)
    rng = np.random.default_rng(4)
    huge = rng.choice(3000, 60, replace=False)          # frame 1: giant boxes of several classes around the same centre
    pred[1, huge, 0:2] = 3000 + rng.normal(0, 30, (60, 2))
    pred[1, huge, 2:4] = rng.uniform(5000, 9000, (60, 2))
    pred[1, huge, 4] = rng.uniform(0.5, 1, 60)
    pred[1, huge, 5:] = rng.uniform(0, 0.3, (60, 5))
    pred[1, huge, 5 + rng.integers(0, 5, 60)] = rng.uniform(0.7, 1, 60)
    for kw in [dict(conf_thres=0.3), dict(conf_thres=0.3, max_det=11), dict(conf_thres=0.3, iou_thres=0.2)]:
        want = yolo_post.extract_boxes(pred, as_written=False, **kw)
        got, _ = _gpu(engine, pred, **kw)
        _same(got, want)
    assert len({int(c) for c in want[1][:, 5]}) >= 1
    # more candidates than max_nms
    dev = torch.from_numpy(pred).cuda()
    out, cnt, _ = engine.yolo_nms(dev, conf_thres=0.3, max_nms=200)
    want = yolo_post.extract_boxes(pred, conf_thres=0.3, as_written=False, max_nms=200)
    _same([out[i, :int(cnt[i])].cpu().numpy() for i in range(3)], want)


def test_empty_and_ragged(engine):
    pred = synth.yolo_output(3, 300, 2, 5)
    pred[1, :, 4] = 0.0  # image 1 has no candidate at all
    got, xc = _gpu(engine, pred, conf_thres=0.3)
    want = yolo_post.extract_boxes(pred, conf_thres=0.3, as_written=False)
    _same(got, want)
    assert got[1].shape == (0, 6) and xc[1] == 0
    got, xc = _gpu(engine, np.zeros((2, 64, 7), np.float32), conf_thres=0.3)
    assert all(g.shape == (0, 6) for g in got) and xc.sum() == 0


def test_full_batch_c2_properties(engine):
    """BASELINE config 2 at full size (64x25200x7): all 64 frames equal the oracle's rows bit for bit, every image
    equals its single-image result, kept boxes are score-sorted, and re-running NMS on the kept set is the identity."""
    pred = synth.yolo_output(64, 25200, 2, 2002)
    dev = torch.from_numpy(pred).cuda()
    out, cnt, _ = engine.yolo_nms(dev, conf_thres=0.3)
    one, cnt1, _ = engine.yolo_nms(dev[9:10].contiguous(), conf_thres=0.3)
    torch.cuda.synchronize()
    assert int(cnt1[0]) == int(cnt[9]) and torch.equal(out[9, :cnt[9]], one[0, :cnt1[0]])
    want = yolo_post.extract_boxes(pred, conf_thres=0.3, as_written=False)
    _same([out[b, :cnt[b]].cpu().numpy() for b in range(64)], want)
    for b in (0, 31, 63):
        k = out[b, :cnt[b]].cpu()
        assert torch.all(k[:-1, 4] >= k[1:, 4])
        import torchvision
        keep = torchvision.ops.nms(k[:, :4] + k[:, 5:6] * 4096, k[:, 4], 0.45)
        assert keep.numel() == k.shape[0]
