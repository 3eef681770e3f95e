"""GPU parity: detectron2-style dense-detector inference (RetinaNet 201 600 anchors / FCOS 22 400 locations at
800x1333) vs the oracle restatement (parity unpinned: detectron2 is not vendored by the reference)."""
import numpy as np
import pytest
import torch

from oracle import detectron_post as dp
from triton_client_b200 import synth

pytestmark = pytest.mark.gpu
STRIDES = (8, 16, 32, 64, 128)
HW = (800, 1333)


def _run(engine, logits, deltas, cells, ctr=None, **kw):
    c = lambda ts: [torch.from_numpy(t).cuda() for t in ts]
    ob, osc, oc, cnt = engine.detectron_post(c(logits), c(deltas), STRIDES, cells, HW, ctrness=c(ctr) if ctr is not None else None, **kw)
    k = int(cnt.item())
    return ob[:k].cpu().numpy(), osc[:k].cpu().numpy(), oc[:k].cpu().numpy()


def _check(got, want):
    assert got[0].shape == want[0].shape and got[0].shape[0] > 10
    assert np.array_equal(got[2], want[2])
    np.testing.assert_allclose(got[1], want[1], rtol=1e-5)
    np.testing.assert_allclose(got[0], want[0], rtol=1e-5, atol=1e-3)


@pytest.mark.parametrize("seed", [3001, 3002])
def test_retinanet_c3(engine, seed):
    logits, deltas = synth.detectron_head_outputs(seed)
    cells = dp.retinanet_cell_anchors()
    assert sum(l.shape[1] * l.shape[2] * 9 for l in logits) == 201600
    _check(_run(engine, logits, deltas, cells), dp.dense_inference(logits, deltas, STRIDES, cells, HW))


def test_fcos_c3(engine):
    logits, deltas, ctr = synth.detectron_head_outputs(3003, anchors_per_loc=1, fcos=True, planted=400)
    cells = dp.fcos_cell_anchors(STRIDES)
    assert sum(l.shape[1] * l.shape[2] for l in logits) == 22400
    kw = dict(score_thresh=0.2, nms_thresh=0.6)
    _check(_run(engine, logits, deltas, cells, ctr, **kw), dp.dense_inference(logits, deltas, STRIDES, cells, HW, ctrness=ctr, **kw))


def test_more_kept_rows_wanted_than_the_best_1024_give(engine):
    """One or two classes and a tiny IoU threshold: of ~5000 merged rows a few hundred survive, fewer than max_det, so the NMS
    over the best 1024 rows cannot settle the result and the full width must (kept < max_det is only reachable that way)."""
    cells = dp.retinanet_cell_anchors()
    for nc, kw in ((1, dict(max_det=1000, nms_thresh=0.0)), (1, dict(max_det=1000, nms_thresh=0.01)), (2, dict(max_det=600, nms_thresh=0.0))):
        logits, deltas = synth.detectron_head_outputs(3005, num_classes=nc, planted=3000)
        got, want = _run(engine, logits, deltas, cells, **kw), dp.dense_inference(logits, deltas, STRIDES, cells, HW, **kw)
        _check(got, want)
        assert 100 < got[0].shape[0] < kw["max_det"]


def test_many_candidates_and_nothing(engine):
    logits, deltas = synth.detectron_head_outputs(3004, num_classes=4, planted=3000)
    cells = dp.retinanet_cell_anchors()
    _check(_run(engine, logits, deltas, cells, topk=300, max_det=200), dp.dense_inference(logits, deltas, STRIDES, cells, HW, topk=300, max_det=200))
    empty = [np.full_like(l, -8.0) for l in logits]
    assert _run(engine, empty, deltas, cells)[0].shape[0] == 0


def test_batched_images_equal_their_single_calls(engine):
    """Three images in one call ([B,A*K,H,W] levels): each image's detections are those of its own call and of the oracle."""
    cells = dp.retinanet_cell_anchors()
    imgs = [synth.detectron_head_outputs(3100 + i, num_classes=8) for i in range(3)]
    logits = [torch.from_numpy(np.stack([im[0][l] for im in imgs])).cuda() for l in range(5)]
    deltas = [torch.from_numpy(np.stack([im[1][l] for im in imgs])).cuda() for l in range(5)]
    ob, osc, oc, cnt = engine.detectron_post(logits, deltas, STRIDES, cells, HW)
    assert ob.shape == (3, 100, 4) and cnt.shape == (3,)
    for i, im in enumerate(imgs):
        one = _run(engine, im[0], im[1], cells)
        k = int(cnt[i])
        assert np.array_equal(ob[i, :k].cpu().numpy(), one[0]) and np.array_equal(osc[i, :k].cpu().numpy(), one[1]) and np.array_equal(oc[i, :k].cpu().numpy(), one[2])
    _check((ob[2, :int(cnt[2])].cpu().numpy(), osc[2, :int(cnt[2])].cpu().numpy(), oc[2, :int(cnt[2])].cpu().numpy()),
           dp.dense_inference(imgs[2][0], imgs[2][1], STRIDES, cells, HW))
    fc = [synth.detectron_head_outputs(3110 + i, anchors_per_loc=1, fcos=True, planted=300) for i in range(2)]
    fcells = dp.fcos_cell_anchors(STRIDES)
    stack = lambda j: [torch.from_numpy(np.stack([im[j][l] for im in fc])).cuda() for l in range(5)]
    ob, osc, oc, cnt = engine.detectron_post(stack(0), stack(1), STRIDES, fcells, HW, ctrness=stack(2), score_thresh=0.2, nms_thresh=0.6)
    for i, im in enumerate(fc):
        one = _run(engine, im[0], im[1], fcells, im[2], score_thresh=0.2, nms_thresh=0.6)
        k = int(cnt[i])
        assert k == len(one[1]) and np.array_equal(ob[i, :k].cpu().numpy(), one[0]) and np.array_equal(oc[i, :k].cpu().numpy(), one[2])
