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


def test_text_is_cv2_puttext_at_any_scale_thickness_and_position():
    """Stroke-level text: identical to cv2.putText also where the frame edge cuts the label, at the reference's two
    settings (0.5 / thickness 2: ros_inference.py:163-169; 1.2 / thickness 1: plot_boxes_cv2) and others."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(2)
    chars = [chr(c) for c in range(32, 127)]
    for k in range(260):
        text = overlay.label_text(rng.uniform(0, 1), "weed") if k % 2 else "".join(rng.choice(chars, int(rng.integers(1, 9))))
        scale, thick = [(0.5, 2), (1.2, 1), (0.8, 3), (2.0, 1), (0.35, 1), (1.0, 2)][k % 6]
        ox, oy = int(rng.integers(-60, 130)), int(rng.integers(-10, 80))
        img = np.zeros((64, 128), np.uint8)
        cv2.putText(img, text, (ox, oy), cv2.FONT_HERSHEY_SIMPLEX, scale, 255, thick)
        got = overlay.text_mask(64, 128, text, ox, oy, scale, thick)
        assert np.array_equal(img > 0, got), (text, ox, oy, scale, thick, int(((img > 0) != got).sum()))


def test_drawing_primitives_match_cv2():
    """The rasterisers under putText, each against its cv2 call: thin and thick lines with sub-pixel endpoints (also far
    outside the frame: clipLine), filled discs, filled convex quads."""
    cv2 = pytest.importorskip("cv2")
    from oracle import cvdraw
    rng = np.random.default_rng(11)
    H, W = 48, 64
    for k in range(500):
        pts = [(int(rng.integers(-20 * 65536, (W + 20) * 65536)), int(rng.integers(-20 * 65536, (H + 20) * 65536))) for _ in range(2)]
        th = int(rng.integers(1, 6))
        img = np.zeros((H, W), np.uint8)
        cv2.line(img, pts[0], pts[1], 255, th, cv2.LINE_8, 16)
        m = np.zeros((H, W), bool)
        cvdraw.thick_line(m, pts[0], pts[1], th, 3)
        assert np.array_equal(img > 0, m), (pts, th)
    for k in range(200):
        cx, cy, r = int(rng.integers(-5, W + 5)), int(rng.integers(-5, H + 5)), int(rng.integers(0, 7))
        img = np.zeros((H, W), np.uint8)
        cv2.circle(img, (cx, cy), r, 255, -1)
        m = np.zeros((H, W), bool)
        cvdraw.circle_filled(m, cx, cy, r)
        assert np.array_equal(img > 0, m), (cx, cy, r)
    for k in range(200):
        c = (rng.uniform(-5, W + 5), rng.uniform(-5, H + 5))
        ang, rad = np.sort(rng.uniform(0, 2 * np.pi, 4)), rng.uniform(2, 25, 4)
        pts = [(int((c[0] + r * np.cos(a)) * 65536), int((c[1] + r * np.sin(a)) * 65536)) for a, r in zip(ang, rad)]
        img = np.zeros((H, W), np.uint8)
        cv2.fillConvexPoly(img, np.array(pts, np.int32), 255, cv2.LINE_8, 16)
        m = np.zeros((H, W), bool)
        cvdraw.fill_convex_poly(m, pts)
        assert np.array_equal(img > 0, m), pts


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


def test_golden_plot_boxes_cv2():
    """Frames annotated by the unmodified Yolov5postprocess.plot_boxes_cv2 (oracle/make_golden.py: golden_plot_boxes):
    per-class colours, class-name labels at scale 1.2 / thickness 1, rectangles thickness 1, boxes over every edge."""
    g = golden("plot_boxes")
    names = [str(v) for v in g["names"]]
    f64 = int(str(g["numpy_version"]).split(".")[0]) < 2
    for k in range(2):
        boxes, cls = g[f"boxes{k}"], g[f"cls{k}"]
        det = np.concatenate([boxes[:, :4], boxes[:, 5:6], cls[:, None].astype(np.float32)], 1).astype(np.float64)
        got = overlay.draw_detections(g[f"img{k}"].copy(), det, names, overlay.plot_colors(len(names)), None, thickness=1, labels="name", f64=f64,
                                      normalized=True, per_class_color=True, font_scale=1.2, text_thickness=1, text_dy=0)
        assert np.array_equal(got, g[f"out{k}"]), int((got != g[f"out{k}"]).any(-1).sum())
