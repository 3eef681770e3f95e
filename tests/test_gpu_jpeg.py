"""GPU parity: baseline JPEG decode through the C ABI against cv2.imdecode itself (what the reference calls,
communicator/ros_inference.py:123-131) and against the oracle. Bit-exact."""
import cv2
import numpy as np
import pytest
import torch

from oracle import jpeg as oj
from test_oracle_jpeg import cases, encode
from triton_client_b200 import synth
from triton_client_b200._lib import TcbError

pytestmark = pytest.mark.gpu


def _cv(data):
    return cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_ANYCOLOR)


def test_every_sampling_quality_and_restart_interval(engine):
    for img, kw in cases():
        data = encode(img, **kw)
        want = _cv(data)
        got = engine.jpeg_decode([data], bgr=True)[0].cpu().numpy()
        assert got.shape == want.shape and np.array_equal(got, want), (img.shape, kw, int((got != want).sum()))


def test_camera_batch_c2_frames(engine):
    """64 camera frames (1280x720, quality 90, 4:2:0, no restart markers: one serial entropy-coded segment each)."""
    frames = [np.roll(synth.image_smooth(720, 1280, 7300 + i % 3), 31 * i, axis=1) for i in range(64)]
    frames[5] = synth.image_noise(720, 1280, 7400)             # a busy frame: long codes, many coefficients
    streams = [encode(f, 90) for f in frames]
    got = engine.jpeg_decode(streams)                          # RGB, as the reference has it after [..., ::-1]
    torch.cuda.synchronize()
    assert got.shape == (64, 720, 1280, 3)
    host = got.cpu().numpy()
    for i in (0, 5, 17, 63):
        assert np.array_equal(host[i], _cv(streams[i])[..., ::-1]), i
    again = engine.jpeg_decode(streams[::-1])
    assert torch.equal(again.flip(0), got)
    assert np.array_equal(host[5], oj.decode_compressed_image(streams[5]))


def test_restart_intervals_and_grey(engine):
    img = synth.image_noise(243, 517, 7500)
    for rst in (1, 3, 33, 1000):
        for smp in ("420", "444"):
            data = encode(img, 92, sampling=smp, rst=rst)
            assert np.array_equal(engine.jpeg_decode([data], bgr=True)[0].cpu().numpy(), _cv(data)), (rst, smp)
    g = synth.image_smooth(300, 410, 7600)[:, :, 1]
    data = encode(g, 80)
    got = engine.jpeg_decode([data, data])
    assert got.shape == (2, 300, 410) and np.array_equal(got[1].cpu().numpy(), _cv(data))


def test_dropin_branch_and_refusals(engine):
    from triton_client_b200.clients import wire
    img = synth.image_smooth(180, 320, 7700)
    data = encode(img, 85)
    want = cv2.imdecode(np.ndarray(shape=(1, len(data)), dtype=np.uint8, buffer=data), cv2.IMREAD_ANYCOLOR)[..., ::-1]   # the reference's lines
    assert np.array_equal(wire.decode_compressed_image(data, engine=engine), want)
    dev = wire.decode_compressed_image([data, data], engine=engine, device=True)
    assert dev.is_cuda and dev.shape == (2, 180, 320, 3) and np.array_equal(dev[1].cpu().numpy(), want)
    ok, prog = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_PROGRESSIVE, 1])
    for bad in (prog.tobytes(), b"\xff\xd8 nothing here", data[:200]):
        with pytest.raises(TcbError):
            engine.jpeg_decode([bad])
    other = encode(synth.image_smooth(90, 160, 1), 85)
    with pytest.raises(TcbError):
        engine.jpeg_decode([data, other])                       # sizes differ
    assert engine.jpeg_decode([data]).shape == (1, 180, 320, 3)   # still usable
