"""ctypes binding of libtcb200.so (the C ABI declared in include/tcb200.h).

There is no CPU fallback: if the shared library is missing or a call fails, this raises.
PyTorch is only the tensor carrier — callers pass `tensor.data_ptr()` and the current stream.
"""
import ctypes
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TCB_LIB_PATH") or os.path.join(_HERE, "libtcb200.so")   # TCB_LIB_PATH: A/B runs of two builds
CSRC = os.path.join(_HERE, "csrc")

_lib = None
_lock = threading.Lock()

c_int, c_ll, c_float, c_double, c_void_p, c_size_t = (
    ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_double, ctypes.c_void_p, ctypes.c_size_t)
P = ctypes.POINTER


class TcbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libtcb200 error {code}: {msg}")
        self.code = code


# name -> (restype, argtypes); every symbol include/tcb200.h declares must be listed here
# (tests/test_abi.py cross-checks the two).
SIGNATURES = {
    "tcb_create": (c_int, [c_int, c_size_t, P(c_void_p)]),
    "tcb_destroy": (c_int, [c_void_p]),
    "tcb_last_error": (ctypes.c_char_p, [c_void_p]),
    "tcb_workspace_bytes": (c_size_t, [c_void_p]),
    "tcb_version": (ctypes.c_char_p, []),
    "tcb_launch_count": (c_ll, [c_void_p]),
    "tcb_preprocess_image": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_ll, c_ll, c_int, c_int, c_int, c_int,
                                     c_float, P(c_float), c_int, c_void_p, P(c_float), c_void_p]),
    "tcb_jpeg_info": (c_int, [c_void_p, c_size_t, P(c_int), P(c_int), P(c_int)]),
    "tcb_jpeg_staging_bytes": (c_size_t, [c_size_t, c_int]),
    "tcb_jpeg_workspace": (c_size_t, [c_size_t, c_int, c_int, c_int]),
    "tcb_jpeg_decode": (c_int, [c_void_p, P(c_void_p), P(c_size_t), c_int, c_void_p, c_size_t, c_int, c_void_p, c_int, c_int, c_int, c_void_p]),
    "tcb_content_rect": (c_int, [c_int, c_int, c_int, c_int, c_int, P(c_int)]),
    "tcb_copy_rect_d2h": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "tcb_yolo_nms_workspace": (c_size_t, [c_int, c_int]),
    "tcb_yolo_nms": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_double, c_double, P(c_int), c_int, c_int,
                             c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "tcb_yolo_head_decode": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, P(c_float), c_int, c_double, c_float,
                                     c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "tcb_nms_axis_aligned": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_float, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "tcb_yolo4_post_workspace": (c_size_t, [c_int, c_int]),
    "tcb_yolo4_post": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_float, c_int, c_void_p, c_void_p, c_void_p]),
    "tcb_box_iou": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "tcb_detectron_post_workspace": (c_size_t, [c_int, c_int, P(c_int), P(c_int), c_int, c_int, c_int]),
    "tcb_detectron_post": (c_int, [c_void_p, c_int, c_int, P(c_void_p), P(c_void_p), P(c_void_p), P(c_int), P(c_int), P(c_int), c_int, c_int,
                                   P(c_float), c_int, c_float, c_int, c_float, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p]),
    "tcb_voxelize_workspace": (c_size_t, [c_int, c_int, c_int, c_ll]),
    "tcb_voxelize": (c_int, [c_void_p, c_void_p, c_int, P(c_int), c_int, c_int, c_int, P(c_double), P(c_double), c_int, c_int, c_int, c_int,
                             c_int, c_float, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "tcb_pillar_decorate": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, P(c_double),
                                    P(c_double), c_void_p, c_void_p]),
    "tcb_dynamic_pillar_workspace": (c_size_t, [c_int, c_int, c_int, c_int]),
    "tcb_dynamic_pillar": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, P(c_double), P(c_double), c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p]),
    "tcb_rotated_iou": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "tcb_rotated_nms": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "tcb_pp_anchor_post_workspace": (c_size_t, [c_int, c_int, c_int, c_int, c_int]),
    "tcb_pp_anchor_post": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, P(c_float), P(c_float),
                                   P(c_float), P(c_float), c_float, c_float, c_int, c_float, c_float, c_int, c_int, c_int,
                                   c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "tcb_centerpoint_post_workspace": (c_size_t, [c_int, c_int, c_int, c_int]),
    "tcb_centerpoint_post": (c_int, [c_void_p, c_int, c_int, P(c_void_p), P(c_void_p), P(c_void_p), P(c_void_p), P(c_void_p), P(c_void_p),
                                     P(c_int), P(c_int), c_int, c_int, c_int, P(c_float), P(c_float), P(c_float), c_int,
                                     c_float, c_float, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "tcb_class_score_filter": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, P(c_int), P(c_float), c_void_p, c_void_p, c_void_p]),
    "tcb_scale_boxes": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "tcb_draw_detections_workspace": (c_size_t, [c_int, c_int]),
    "tcb_draw_detections": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_ll, c_ll, c_void_p, c_void_p, c_int,
                                    c_int, c_int, c_int, c_int, ctypes.c_char_p, c_int, ctypes.c_char_p, c_int, c_int, c_int, c_double, c_int,
                                    c_int, c_void_p]),
    "tcb_pointcloud2_unpack_workspace": (c_size_t, [c_int, c_int]),
    "tcb_pointcloud2_unpack": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_ll, c_int, P(c_int), P(c_int), c_int, c_int,
                                       c_void_p, c_void_p, c_void_p]),
    "tcb_eval_workspace": (c_size_t, [c_int, c_int]),
    "tcb_eval_match": (c_int, [c_void_p, c_void_p, P(c_int), c_void_p, P(c_int), c_int, c_void_p, c_void_p, c_void_p]),
    "tcb_ap_per_class": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_ll, P(c_int), c_void_p, c_ll, P(c_int), c_int, c_int,
                                 c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "tcb_calculate_metrics": (c_int, [c_void_p, c_void_p, P(c_int), c_void_p, P(c_int), c_int, c_int, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "tcb_boxes_to_jsk": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_ll, c_float, c_double, c_void_p, c_void_p,
                                 c_void_p, c_void_p]),
    "tcb_yolo_nms_static_scratch_bytes": (c_size_t, []),
    "tcb_yolo_nms_static": (c_int, [c_void_p, c_int, c_int, c_int, c_double, c_double, P(c_int), c_int, c_int,
                                    c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
}


def build(verbose=False):
    """Compile libtcb200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", CSRC, "-j", str(min(8, os.cpu_count() or 1))]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise RuntimeError("building libtcb200.so failed:\n" + res.stdout[-4000:])
    return LIB_PATH


def lib():
    """The loaded library; raises if it has not been built (no silent fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    f"{LIB_PATH} not found: build it with `make -C {CSRC}` or `python -c 'import __graft_entry__ as g; g.build()'`. "
                    "triton_client_b200 has no CPU fallback.")
            l = ctypes.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(l, name)
                fn.restype = res
                fn.argtypes = args
            _lib = l
    return _lib


class Context:
    """Owns one tcb_ctx (device + workspace). One per pre/post object, i.e. one frame in flight."""

    def __init__(self, device=0, workspace_bytes=256 << 20):
        self._l = lib()
        h = c_void_p()
        rc = self._l.tcb_create(int(device), int(workspace_bytes), ctypes.byref(h))
        if rc != 0:
            raise TcbError(rc, (self._l.tcb_last_error(None) or b"").decode())
        self.handle = h
        self.device = int(device)

    def check(self, rc):
        if rc != 0:
            raise TcbError(rc, (self._l.tcb_last_error(self.handle) or b"").decode())

    def call(self, name, *args):
        self.check(getattr(self._l, name)(self.handle, *args))

    @property
    def launches(self):
        return int(self._l.tcb_launch_count(self.handle))

    def close(self):
        if getattr(self, "handle", None):
            self._l.tcb_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_raw_stream = None


def stream_ptr():
    """cudaStream_t of torch's current stream (current device) as an integer for the C ABI."""
    global _raw_stream
    import torch
    if _raw_stream is None:      # torch's own fast path to the raw handle (what its extensions use); the public objects otherwise
        fast = getattr(torch._C, "_cuda_getCurrentRawStream", None)
        _raw_stream = (lambda: fast(torch.cuda.current_device())) if fast is not None else (lambda: torch.cuda.current_stream().cuda_stream)
    return c_void_p(_raw_stream())
