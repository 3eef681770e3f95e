"""GPU parity: fused preprocess kernel (through the C ABI) vs the oracle and the reference's golden
vectors. Standard: north_star allows <= 1/255 per pixel; the kernel is written to be bit-identical to
cv2.resize + image_adjust, so the tests ask for exact equality."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import pre2d
from triton_client_b200 import synth

pytestmark = pytest.mark.gpu


def _run(engine, imgs, oh, ow, **kw):
    src = torch.from_numpy(np.stack(imgs)).cuda()
    out, aff = engine.preprocess_images(src, oh, ow, **kw)
    torch.cuda.synchronize()
    return out.cpu().numpy(), aff


def test_golden_reference_vectors(engine):
    g = golden("pre2d")
    for k in range(5):
        h, w, oh, ow = (int(v) for v in g[f"shape{k}"])
        got, _ = _run(engine, [g[f"img{k}"]], oh, ow, mode="stretch", div=255.0)
        assert np.array_equal(got, g[f"yolo{k}"]), k
    got, _ = _run(engine, [g["img0"]], 32, 32, mode="stretch", div=1.0)
    assert np.array_equal(got, g["fcos0"])


@pytest.mark.parametrize("h,w,oh,ow", [
    (720, 1280, 640, 640),   # C2 stretch (what the reference does)
    (640, 640, 640, 640),    # C1 (identity resample)
    (480, 640, 512, 512),    # shipped 512x512 model, upscale in y
    (800, 1333, 800, 1333),  # C3, odd width -> scalar store path
    (375, 1242, 640, 640),   # KITTI camera frame
    (1080, 1920, 640, 640),
    (2160, 3840, 320, 320),  # heavy downscale -> unstaged path
    (37, 53, 131, 257),      # upscale, ragged tiles
    (5, 3, 9, 2), (2, 2, 1, 1), (1, 1, 4, 4),
])
def test_stretch_matches_oracle_exactly(engine, h, w, oh, ow):
    imgs = [synth.image_noise(h, w, 10 + i) for i in range(2)]
    got, aff = _run(engine, imgs, oh, ow, mode="stretch", div=255.0)
    for i, img in enumerate(imgs):
        want = pre2d.preprocess_frame(img, oh, ow, mode="stretch", div=255.0)
        assert np.array_equal(got[i:i + 1], want), (h, w, oh, ow, i, np.abs(got[i:i + 1] - want).max())
    assert np.allclose(aff, [ow / w, oh / h, 0, 0])


@pytest.mark.parametrize("h,w,oh,ow", [(720, 1280, 640, 640), (1280, 720, 640, 640), (480, 640, 640, 640),
                                         (375, 1242, 640, 640), (100, 300, 64, 96), (640, 640, 640, 640), (30, 20, 64, 64)])
def test_letterbox_matches_oracle_exactly(engine, h, w, oh, ow):
    imgs = [synth.image_smooth(h, w, 3), synth.image_noise(h, w, 4)]
    got, aff = _run(engine, imgs, oh, ow, mode="letterbox", div=255.0, pad_value=114)
    r, nw, nh, left, top, _, _ = pre2d.letterbox_geometry(h, w, oh, ow)
    for i, img in enumerate(imgs):
        want = pre2d.preprocess_frame(img, oh, ow, mode="letterbox", div=255.0, pad_value=114)
        assert np.array_equal(got[i:i + 1], want), (h, w, i)
    assert np.allclose(aff, [r, r, left, top])


@pytest.mark.parametrize("h,w,oh,ow,mode", [
    (256, 160, 128, 128, "letterbox"),   # 2:1, left pad 24 (a multiple of 4): fused 2x2 pass
    (256, 168, 128, 128, "letterbox"),   # 2:1, left pad 22: falls back to the two-pass kernel
    (64, 96, 32, 48, "stretch"),         # 2:1 stretch, one tile
    (600, 1032, 300, 516, "stretch"),    # 2:1, ragged right tile, several tile rows
    (720, 1280, 360, 640, "stretch"),
    (2, 8, 1, 4, "stretch"),
    (480, 1280, 512, 640, "stretch"),    # width halved, height up-scaled: fused pass with bilinear rows
    (100, 64, 37, 32, "stretch"),
    (375, 1242, 200, 621, "stretch"),    # width halved but 621 is not a multiple of 4: general kernel
    (480, 640, 512, 640, "stretch"),     # width unchanged, height up-scaled: fused identity-width pass
    (100, 64, 37, 64, "stretch"),        # width unchanged, height reduced
    (300, 640, 640, 640, "letterbox"),   # scale 1, rows padded above and below
    (64, 64, 64, 64, "stretch"),
])
def test_exact_half_scale_paths(engine, h, w, oh, ow, mode):
    """cv2's INTER_LINEAR at exactly 2:1 is the 2x2 box mean; aligned cases take the fused pass, the rest the general
    kernel. Also with channel swap + mean subtraction, a batch (first/last tile rows copy chunk-wise) and a row-strided,
    4-byte aligned source that is not 16-byte aligned (fallback)."""
    imgs = [synth.image_noise(h, w, 70 + i) for i in range(3)]
    got, _ = _run(engine, imgs, oh, ow, mode=mode, div=255.0, pad_value=114)
    for i, img in enumerate(imgs):
        assert np.array_equal(got[i:i + 1], pre2d.preprocess_frame(img, oh, ow, mode=mode, div=255.0, pad_value=114)), (h, w, i)
    got, _ = _run(engine, imgs[:1], oh, ow, mode=mode, swap=True, div=127.5, sub=(1, 1, 1))
    want = pre2d.preprocess_frame(np.ascontiguousarray(imgs[0][..., ::-1]), oh, ow, mode=mode, div=127.5, pad_value=114) - np.float32(1)
    if mode == "stretch":
        assert np.array_equal(got, want)
    if w >= 16:
        big = torch.from_numpy(synth.image_noise(h + 3, w + 8, 80)).cuda()
        view = big[1:h + 1, 4:w + 4]
        out, _ = engine.preprocess_images(view, oh, ow, mode=mode, pad_value=114)
        assert np.array_equal(out.cpu().numpy(), pre2d.preprocess_frame(view.cpu().numpy(), oh, ow, mode=mode, pad_value=114))
    if w % 16 == 0:   # strided but 16-byte aligned rows: still the fused pass
        big = torch.from_numpy(synth.image_noise(h, w + 16, 81)).cuda()
        view = big[:, 16:]
        out, _ = engine.preprocess_images(view, oh, ow, mode=mode, pad_value=114)
        assert np.array_equal(out.cpu().numpy(), pre2d.preprocess_frame(view.cpu().numpy(), oh, ow, mode=mode, pad_value=114))


def test_scaling_modes_swap_and_pad(engine):
    img = synth.image_noise(90, 160, 5)
    for scaling, div, sub in [("INCEPTION", 127.5, (1, 1, 1)), ("VGG", 1.0, (123, 117, 104)), ("COCO", 255.0, (0, 0, 0)), ("NONE", 1.0, (0, 0, 0))]:
        want = pre2d.image_adjust_utils(img, 64, 48, scaling)
        got, _ = _run(engine, [img], 48, 64, mode="stretch", swap=True, div=div, sub=sub)
        assert np.array_equal(got, want), scaling
    # a divisor for which v*(1/div) cannot be corrected to the exact quotient -> true-division path
    got, _ = _run(engine, [img], 48, 64, mode="stretch", div=3.1)
    assert np.array_equal(got, pre2d.preprocess_frame(img, 48, 64, div=3.1))
    # image_adjust_ros: 16 zero rows appended (utils/preprocess.py:177-183)
    got, _ = _run(engine, [img], 106, 160, mode="pad", pad_value=0)
    assert np.array_equal(got, pre2d.image_adjust_ros(img))


def test_utils_image_adjust_reference_golden(engine):
    """utils/preprocess.py::image_adjust run unmodified (golden_utils_preprocess): the cv2 path through the drop-in twin
    (file path in, like the reference), and the scaling modes of its PIL path as kernel parameters."""
    import cv2
    import tempfile, os
    from triton_client_b200.utils import preprocess as up
    g = golden("utils_preprocess")
    with tempfile.TemporaryDirectory() as d:
        for k in range(3):
            w, h = (int(v) for v in g[f"cv_wh{k}"])
            path = os.path.join(d, f"f{k}.png")
            cv2.imwrite(path, g[f"cv_bgr{k}"])
            assert np.array_equal(up.image_adjust(path, w, h, "COCO", engine=engine), g[f"cv_out{k}"])
            assert np.array_equal(up.image_adjust(g[f"cv_bgr{k}"], w, h, engine=engine), g[f"cv_out{k}"])
    rgb = g["pil_rgb"]
    for sc in ("NONE", "INCEPTION", "VGG", "COCO"):
        got = up.image_adjust(np.ascontiguousarray(rgb[:, :, ::-1]), rgb.shape[1], rgb.shape[0], sc, engine=engine)
        assert np.array_equal(got[0], g[f"pil_{sc}_nchw"]), sc


def test_strided_source_and_input_not_mutated(engine):
    big = torch.from_numpy(synth.image_noise(200, 300, 6)).cuda()
    view = big[10:150, 20:221]                      # row stride 900 B, offset not 16-B aligned
    before = big.clone()
    out, _ = engine.preprocess_images(view, 96, 96)
    torch.cuda.synchronize()
    want = pre2d.preprocess_frame(view.cpu().numpy(), 96, 96)
    assert np.array_equal(out.cpu().numpy(), want)
    assert torch.equal(big, before)


def test_full_batch_c2_properties(engine):
    """BASELINE config 2 at full size (64 x 1280x720 -> 640x640): all 64 frames equal the oracle bit for bit, batch
    entries equal the single-frame result, and a constant frame maps to the constant."""
    base = synth.image_smooth(720, 1280, 2000)
    frames = np.stack([np.roll(base, 7 * i, axis=1) for i in range(64)])
    frames[5] = 200
    src = torch.from_numpy(frames).cuda()
    out, _ = engine.preprocess_images(src, 640, 640, mode="letterbox")
    one, _ = engine.preprocess_images(src[17], 640, 640, mode="letterbox")
    torch.cuda.synchronize()
    assert torch.equal(out[17:18], one)
    host = out.cpu().numpy()
    for i in range(64):
        assert np.array_equal(host[i:i + 1], pre2d.preprocess_frame(frames[i], 640, 640, mode="letterbox")), i
    content = out[5, :, 140:500, :]
    assert torch.all(content == np.float32(200) / np.float32(255))
    assert torch.all(out[:, :, :140, :] == np.float32(114) / np.float32(255))


def test_errors_are_returned_not_fatal(engine):
    from triton_client_b200._lib import TcbError
    src = torch.zeros((4, 4, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(TcbError):
        engine.preprocess_images(src, 8, 8, div=0.0)
    with pytest.raises(TcbError):
        engine.preprocess_images(src, 2, 2, mode="pad")
    out, _ = engine.preprocess_images(src, 8, 8)  # context still usable afterwards
    assert out.shape == (1, 3, 8, 8)


@pytest.mark.parametrize("h,w", [(800, 1333), (36, 53), (37, 53), (64, 64), (3, 4)])
def test_same_size_frames_flat_path(engine, h, w):
    """Output the size of the source (FCOS / RetinaNet image_adjust on model-sized frames): a flat u8 HWC -> f32 CHW stream
    when the pixel count is a multiple of 4, the tiled kernel otherwise; every normalisation, channel swap, several frames."""
    imgs = [synth.image_noise(h, w, 300 + i) for i in range(3)]
    for div, sub, swap in [(1.0, None, False), (255.0, None, False), (127.5, (1, 1, 1), True), (1.0, (123, 117, 104), True), (3.1, (0.5, 0, 2), False)]:
        got, _ = _run(engine, imgs, h, w, mode="stretch", div=div, sub=sub, swap=swap)
        for i, img in enumerate(imgs):
            want = pre2d.preprocess_frame(img, h, w, mode="stretch", swap=swap, div=div, sub=sub or (0.0, 0.0, 0.0))
            assert np.array_equal(got[i:i + 1], want), (h, w, div, sub, swap, i)
    for mode in ("letterbox", "pad"):     # the other modes at the same size are the identity too
        got, _ = _run(engine, imgs[:1], h, w, mode=mode, div=255.0, pad_value=114)
        assert np.array_equal(got, pre2d.preprocess_frame(imgs[0], h, w, mode=mode, div=255.0, pad_value=114))
