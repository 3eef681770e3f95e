"""Oracle pinning: the JPEG restatement (oracle/c/jpeg.c) against cv2.imdecode itself — what the reference calls
(communicator/ros_inference.py:123-131) — on frames encoded by cv2.imencode: qualities, chroma samplings, restart
intervals, sizes that are not multiples of the MCU, grey streams, optimised Huffman tables. Bit-exact."""
import cv2
import numpy as np
import pytest

from oracle import jpeg as oj
from triton_client_b200 import synth

S = {"444": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444, "422": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422, "420": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420,
     "440": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_440}


def encode(img, quality=95, sampling=None, rst=0, optimize=False):
    p = [cv2.IMWRITE_JPEG_QUALITY, quality]
    if sampling:
        p += [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, S[sampling]]
    if rst:
        p += [cv2.IMWRITE_JPEG_RST_INTERVAL, rst]
    if optimize:
        p += [cv2.IMWRITE_JPEG_OPTIMIZE, 1]
    ok, buf = cv2.imencode(".jpg", img, p)
    assert ok
    return buf.tobytes()


def cases():
    out = []
    for k, (h, w) in enumerate([(64, 64), (67, 93), (120, 200), (17, 33), (240, 321), (8, 8), (1, 1), (9, 300)]):
        img = synth.image_smooth(h, w, 7000 + k) if k % 2 == 0 else synth.image_noise(h, w, 7000 + k)
        for q in (50, 75, 95, 100):
            for smp in (None, "444", "422", "420", "440"):
                out.append((img, dict(quality=q, sampling=smp, rst=(0, 1, 5)[(k + q) % 3], optimize=(q == 75))))
    return out


def test_matches_cv2_imdecode_bit_for_bit():
    n = 0
    for img, kw in cases():
        data = encode(img, **kw)
        want = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_ANYCOLOR)
        got = oj.imdecode(data)
        assert got.shape == want.shape, (img.shape, kw)
        assert np.array_equal(got, want), (img.shape, kw, int(np.abs(got.astype(int) - want).max()), int((got != want).sum()))
        n += 1
    assert n == 160


def test_camera_frame_and_the_reference_channel_flip():
    img = synth.image_smooth(720, 1280, 7100)
    data = encode(img, 90)
    want = cv2.imdecode(np.ndarray(shape=(1, len(data)), dtype=np.uint8, buffer=data), cv2.IMREAD_ANYCOLOR)[..., ::-1]   # ros_inference.py:125-131
    assert np.array_equal(oj.decode_compressed_image(data), want)


def test_grey_and_unsupported():
    g = synth.image_smooth(50, 70, 7200)[:, :, 0]
    data = encode(g, 85)
    want = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_ANYCOLOR)
    assert want.ndim == 2 and np.array_equal(oj.imdecode(data), want)
    ok, prog = cv2.imencode(".jpg", synth.image_noise(32, 32, 1), [cv2.IMWRITE_JPEG_PROGRESSIVE, 1])
    with pytest.raises(ValueError):
        oj.imdecode(prog.tobytes())
    with pytest.raises(ValueError):
        oj.imdecode(b"not a jpeg")
