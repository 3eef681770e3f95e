"""Oracle pinning for the box rescale + overlay step: golden frames annotated by the unmodified
RosInference._callback (oracle/make_golden.py::golden_overlay), and — where cv2 is importable — OpenCV itself."""
import zlib

import numpy as np
import pytest

from conftest import golden
from oracle import overlay, ref_harness
from triton_client_b200 import synth

NAMES = ["crop", "weed"]
GREEN = (0, 255, 0)


def _case(g, k):
    h, w, size, seed = (int(v) for v in g[f"cfg{k}"])
    frame = synth.image_noise(h, w, seed)
    assert zlib.crc32(frame.tobytes()) == int(g[f"frame_crc{k}"]), "synthetic frame generator drifted from the fixture"
    painted = np.unpackbits(g[f"painted{k}"])[:h * w].reshape(h, w).astype(bool)
    return frame, size, g[f"det{k}"], painted


@pytest.mark.parametrize("k", [0, 1])
def test_golden_callback_overlay(k):
    g = golden("overlay")
    frame, size, det, painted = _case(g, k)
    assert len(det) >= 5 and painted.sum() > 5000
    # the fixture was made under numpy >= 2 (float32 scalar arithmetic); on this data float64 gives the same pixels
    for f64 in (False, True):
        out = overlay.draw_detections(frame.copy(), det, NAMES, GREEN, [size, size], f64=f64)
        mine = (out != frame).any(-1)
        assert np.array_equal(mine, painted)
        assert (out[painted] == np.array(GREEN, np.uint8)).all() and np.array_equal(out[~painted], frame[~painted])


def test_rect_mask_matches_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    for _ in range(3000):
        h, w = (int(v) for v in rng.integers(8, 60, 2))
        t = int(rng.integers(1, 3))
        x1, x2 = (int(v) for v in rng.integers(-20, w + 20, 2))
        y1, y2 = (int(v) for v in rng.integers(-20, h + 20, 2))
        img = np.zeros((h, w), np.uint8)
        cv2.rectangle(img, (x1, y1), (x2, y2), 255, t)
        assert np.array_equal(img > 0, overlay.rect_mask(h, w, x1, y1, x2, y2, t)), (h, w, x1, y1, x2, y2, t)
    # far outside / huge boxes on a full-size frame
    for (x1, y1, x2, y2) in [(-5000, -3000, 9000, 7000), (-1, -1, 1280, 720), (1279, 719, 4000, 4000), (640, -50, 641, 900)]:
        for t in (1, 2):
            img = np.zeros((720, 1280), np.uint8)
            cv2.rectangle(img, (x1, y1), (x2, y2), 255, t)
            assert np.array_equal(img > 0, overlay.rect_mask(720, 1280, x1, y1, x2, y2, t))


def test_text_mask_matches_cv2_inside_the_frame():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(1)
    names = ["crop", "weed", "person", "traffic light", "Quxz@gp|"]
    for _ in range(400):
        text = overlay.label_text(rng.uniform(0, 1), names[int(rng.integers(0, len(names)))])
        ox, oy = int(rng.integers(0, 60)), int(rng.integers(14, 60))
        img = np.zeros((70, 240), np.uint8)
        cv2.putText(img, text, (ox, oy), cv2.FONT_HERSHEY_SIMPLEX, 0.5, 255, 2)
        assert np.array_equal(img > 0, overlay.text_mask(70, 240, text, ox, oy)), (text, ox, oy)


def test_text_cut_by_the_frame_edge_differs_only_near_the_edge():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(2)
    worst = 0
    for _ in range(300):
        text = overlay.label_text(rng.uniform(0, 1), "weed")
        ox, oy = int(rng.integers(-40, 10)), int(rng.integers(-4, 20))
        img = np.zeros((64, 128), np.uint8)
        cv2.putText(img, text, (ox, oy), cv2.FONT_HERSHEY_SIMPLEX, 0.5, 255, 2)
        diff = (img > 0) != overlay.text_mask(64, 128, text, ox, oy)
        worst = max(worst, int(diff.sum()))
        ys, xs = np.nonzero(diff)
        assert diff.sum() <= 16 and (len(xs) == 0 or min(xs.min(), ys.min()) < 20), (text, ox, oy)   # inside a glyph the edge cuts
    assert worst > 0   # documents that the difference exists (see oracle/overlay.py)


def test_scale_boxes_against_the_reference_methods():
    if not ref_harness.available():
        pytest.skip("reference not mounted")
    comm = ref_harness.communicator()
    node = object.__new__(comm.RosInference)
    ev = object.__new__(comm.EvaluateInference)
    rng = np.random.default_rng(3)
    numpy2 = int(np.__version__.split(".")[0]) >= 2
    for _ in range(200):
        oh, ow = (int(v) for v in rng.integers(100, 2000, 2))
        ins = [int(v) for v in rng.integers(64, 1400, 2)]
        node.orig_size, node.input_size = (oh, ow), ins
        ev.orig_size, ev.input_size = (oh, ow), ins
        det = rng.uniform(-20, 1400, (7, 6))
        for normalized in (False, True):
            box = det[0, :4].astype(np.float32)
            want = node._scale_boxes(box, normalized=normalized)
            got = overlay.scale_box(box, (oh, ow), ins, normalized, f64=not numpy2)
            assert [float(v) for v in want] == [float(v) for v in got] and type(want[0]) is type(got[0])
        import torch
        want = ev._scale_box_array(torch.from_numpy(det), normalized=False)
        assert np.array_equal(np.asarray(want), overlay.scale_box_array(det, (oh, ow), ins))


def test_label_text_rounding():
    # '{:.2f}' rounds the exact binary value, ties to even: 0.125 -> 0.12, 0.375 -> 0.38, 0.005 (just above) -> 0.01
    assert overlay.label_text(0.125, "a") == "0.12 a" and overlay.label_text(0.375, "a") == "0.38 a"
    assert overlay.label_text(0.005, "a") == "0.01 a" and overlay.label_text(0.995, "a") == "0.99 a" and overlay.label_text(1.0, "b") == "1.00 b"
