"""Oracle: baseline JPEG decode as cv2.imdecode does it for the reference's CompressedImage branch
(communicator/ros_inference.py:123-131). Test infrastructure only. The decoder is the C restatement in oracle/c/jpeg.c
of libjpeg-turbo's default decode path (libjpeg-turbo is a third-party dependency of OpenCV, not part of the reference
tree: OpenCV 4.13 bundles 3.1.2); it is pinned bit for bit against cv2.imdecode itself (tests/test_oracle_jpeg.py)."""
import ctypes

import numpy as np

from .voxelize import clib


def info(data):
    buf = np.frombuffer(bytes(data), np.uint8)
    w, h, n = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    rc = clib().oracle_jpeg_info(ctypes.c_void_p(buf.ctypes.data), ctypes.c_size_t(buf.size), ctypes.byref(w), ctypes.byref(h), ctypes.byref(n))
    if rc != 0:
        raise ValueError(f"not a baseline JPEG this decoder handles (code {rc})")
    return h.value, w.value, n.value


def imdecode(data, bgr=True):
    """bytes -> (H, W, 3) uint8 in BGR order like cv2.imdecode (RGB when bgr=False), or (H, W) for a grey stream."""
    buf = np.frombuffer(bytes(data), np.uint8)
    h, w, n = info(data)
    out = np.empty((h, w, 3) if n == 3 else (h, w), np.uint8)
    lib = clib()
    lib.oracle_jpeg_decode.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int]
    rc = lib.oracle_jpeg_decode(ctypes.c_void_p(buf.ctypes.data), buf.size, ctypes.c_void_p(out.ctypes.data), 1 if bgr else 0)
    if rc != 0:
        raise ValueError(f"JPEG outside the decoder's scope or corrupt (code {rc})")
    return out


def decode_compressed_image(data):
    """communicator/ros_inference.py:124-131: imdecode(IMREAD_ANYCOLOR) then [..., ::-1] -> RGB (H, W, 3)."""
    return imdecode(data)[..., ::-1]
