"""Oracle pinning for the 2D preprocess: golden vectors produced by the unmodified reference
(oracle/make_golden.py) and the numpy restatement of OpenCV's u8 bilinear against cv2 itself."""
import numpy as np
import pytest

from conftest import golden
from oracle import pre2d, ref_harness


def test_golden_yolo_and_fcos_preprocess():
    g = golden("pre2d")
    for k in range(5):
        h, w, oh, ow = g[f"shape{k}"]
        img = g[f"img{k}"]
        got = pre2d.image_adjust_yolov5(pre2d.stretch_resize(img, ow, oh))
        assert got.dtype == np.float32 and got.shape == (1, 3, oh, ow)
        assert np.array_equal(got, g[f"yolo{k}"])
        fused = pre2d.preprocess_frame(img, oh, ow, mode="stretch", div=255.0)
        assert np.array_equal(fused, g[f"yolo{k}"])
    assert np.array_equal(pre2d.image_adjust_fcos(pre2d.stretch_resize(g["img0"], 32, 32)), g["fcos0"])


def test_restated_bilinear_matches_cv2_bit_for_bit():
    import cv2
    rng = np.random.default_rng(7)
    shapes = [(720, 1280, 640, 640), (720, 1280, 360, 640), (480, 640, 512, 512), (64, 64, 640, 640),
              (100, 100, 300, 277), (800, 1333, 800, 1333), (2, 2, 5, 7), (7, 5, 1, 1)]
    for _ in range(40):
        h, w = rng.integers(2, 300, 2)
        oh, ow = rng.integers(1, 400, 2)
        shapes.append((int(h), int(w), int(oh), int(ow)))
    for h, w, oh, ow in shapes:
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        assert np.array_equal(pre2d.cv_resize_linear_u8(img, ow, oh), cv2.resize(img, (ow, oh))), (h, w, oh, ow)


def test_letterbox_geometry_c2():
    r, nw, nh, left, top, right, bottom = pre2d.letterbox_geometry(720, 1280, 640, 640)
    assert (r, nw, nh, left, top, right, bottom) == (0.5, 640, 360, 0, 140, 0, 140)
    out, _ = pre2d.letterbox(np.zeros((720, 1280, 3), np.uint8), 640, 640)
    assert out.shape == (640, 640, 3) and out[0, 0, 0] == 114 and out[320, 320, 0] == 0


def test_scaling_modes_and_pad():
    img = np.random.default_rng(3).integers(0, 256, (20, 30, 3), dtype=np.uint8)
    inc = pre2d.image_adjust_utils(img, 16, 12, "INCEPTION")
    assert inc.shape == (1, 3, 12, 16) and inc.min() >= -1 and inc.max() <= 1
    fused = pre2d.preprocess_frame(img, 12, 16, swap=True, div=127.5, sub=(1, 1, 1))
    assert np.array_equal(fused, inc)
    vgg = pre2d.image_adjust_utils(img, 16, 12, "VGG")
    assert np.array_equal(pre2d.preprocess_frame(img, 12, 16, swap=True, div=1.0, sub=(123, 117, 104)), vgg)
    ros = pre2d.image_adjust_ros(img)
    assert np.array_equal(pre2d.preprocess_frame(img, 36, 30, mode="pad", pad_value=0), ros)


@pytest.mark.skipif(not ref_harness.available(), reason="reference not mounted (GPU box)")
def test_against_live_reference():
    import cv2
    c = ref_harness.clients()
    pre = c.Yolov5client().get_preprocess()
    img = np.random.default_rng(11).integers(0, 256, (90, 160, 3), dtype=np.uint8)
    ref = pre.image_adjust(cv2.resize(img, (64, 64)))
    assert np.array_equal(pre2d.preprocess_frame(img, 64, 64), ref)


def test_golden_utils_image_adjust():
    """utils/preprocess.py::image_adjust run unmodified (oracle/make_golden.py: golden_utils_preprocess): the cv2 path
    (which ignores `scaling` and always divides by 255) and the scaling arithmetic of the PIL path."""
    g = golden("utils_preprocess")
    for k in range(3):
        w, h = (int(v) for v in g[f"cv_wh{k}"])
        assert np.array_equal(pre2d.image_adjust_utils(g[f"cv_bgr{k}"], w, h, "COCO"), g[f"cv_out{k}"])
    rgb = g["pil_rgb"]
    for sc in ("NONE", "INCEPTION", "VGG", "COCO"):
        got = pre2d.image_adjust_utils(rgb[:, :, ::-1], rgb.shape[1], rgb.shape[0], sc)[0]      # the oracle flips BGR -> RGB
        assert np.array_equal(got, g[f"pil_{sc}_nchw"]), sc
        assert np.array_equal(np.transpose(got, (1, 2, 0)), g[f"pil_{sc}_nhwc"]), sc
