"""Wire/host boundary helpers.

The reference parses Triton's `raw_output_contents` one element at a time with struct.unpack_from
(clients/postprocess/base_postprocess.py:15-37). Here the little-endian bytes ARE the tensor: they are
copied to the device (through a reusable pinned staging buffer when the source is pageable) and
reinterpreted there; nothing is parsed.
"""
import os
import threading

import numpy as np
import torch

_engine = None
_lock = threading.Lock()


def default_engine():
    """Process-wide Engine (one process per GPU: device = LOCAL_RANK, or TCB_DEVICE)."""
    global _engine
    with _lock:
        if _engine is None:
            from ..gpu import Engine
            dev = int(os.environ.get("TCB_DEVICE", os.environ.get("LOCAL_RANK", "0")))
            ws = int(os.environ.get("TCB_WORKSPACE_MB", "512")) << 20
            _engine = Engine(dev, workspace_bytes=ws)
    return _engine


class _Out:
    def __init__(self, shape):
        self.shape = list(shape)


class HostResponse:
    """Minimal stand-in for tritonclient's ModelInferResponse: `.outputs[i].shape` and
    `.raw_output_contents[i]`. Contents may be bytes (as gRPC delivers them) or host arrays/tensors
    (e.g. pinned), which are used without a copy."""

    def __init__(self, arrays, shapes=None):
        self.raw_output_contents = list(arrays)
        if shapes is None:
            shapes = [tuple(a.shape) for a in arrays]
        self.outputs = [_Out(s) for s in shapes]


class Stager:
    """Reusable pinned host buffers: pageable bytes -> pinned -> device, and device -> pinned -> numpy."""

    def __init__(self, device):
        self.device = device
        self._pin = {}

    def pinned(self, key, nbytes):
        buf = self._pin.get(key)
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(max(int(nbytes), 1), dtype=torch.uint8).pin_memory()
            self._pin[key] = buf
        return buf[:nbytes]

    def to_device(self, raw, np_dtype, shape, key="in"):
        """bytes / bytearray / memoryview / ndarray / CPU tensor / CUDA tensor -> CUDA tensor of `shape`."""
        tdtype = torch.from_numpy(np.empty(0, dtype=np_dtype)).dtype
        if isinstance(raw, torch.Tensor):
            t = raw
            if t.is_cuda:
                return t.reshape(shape).contiguous()
            if t.dtype != tdtype:
                t = t.view(torch.uint8).reshape(-1).view(tdtype)
            t = t.reshape(shape)
            if t.is_pinned():
                return t.to(self.device, non_blocking=True)
            host = t.contiguous().view(torch.uint8).reshape(-1)
        else:
            if isinstance(raw, np.ndarray):
                arr = np.ascontiguousarray(raw)
                host = torch.from_numpy(arr.view(np.uint8).reshape(-1))
            else:   # bytes / bytearray / memoryview straight from gRPC: one memcpy into the pinned buffer
                arr = np.frombuffer(raw, dtype=np.uint8)
                stage = self.pinned(key, arr.size)
                if arr.size:
                    stage.numpy()[:] = arr
                return stage.to(self.device, non_blocking=True).view(tdtype).reshape(shape)
        stage = self.pinned(key, host.numel())
        stage.copy_(host)
        return stage.to(self.device, non_blocking=True).view(tdtype).reshape(shape)

    def to_host(self, t, key="out", copy=True):
        """CUDA tensor -> numpy array through a pinned buffer. copy=True returns memory the caller owns
        (the reference's contract); copy=False returns a view of the pinned buffer, valid until the next
        call with the same key."""
        t = t.contiguous()
        stage = self.pinned(key, t.numel() * t.element_size())
        view = stage.view(t.dtype).reshape(t.shape)
        view.copy_(t, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return view.numpy().copy() if copy else view.numpy()


def decode_compressed_image(data, engine=None, device=False):
    """The compressed-image branch of the reference's 2D callback (communicator/ros_inference.py:123-131):
        cv_image = cv2.imdecode(np.ndarray(shape=(1, len(msg.data)), dtype=np.uint8, buffer=msg.data), cv2.IMREAD_ANYCOLOR)[..., ::-1]
    `data` = msg.data (bytes-like, a baseline JPEG). Returns the RGB frame (H, W, 3) uint8 — a host ndarray like the
    reference's, or with device=True the CUDA tensor, ready for Yolov5preprocess.image_adjust_device / resize_adjust
    without the pixels ever visiting the host. Decoded by tcb_jpeg_decode, bit-identical to OpenCV. A list of streams
    of one size gives (n, H, W, 3)."""
    eng = engine or default_engine()
    single = isinstance(data, (bytes, bytearray, memoryview, np.ndarray))
    frames = eng.jpeg_decode([data] if single else list(data), bgr=False)
    frames = frames[0] if single else frames
    return frames if device else frames.cpu().numpy()
