"""Mirror of the hot-path functions of the reference's utils/preprocess.py (image_adjust :127-175, image_adjust_ros
:177-183, model_dtype_to_np :17-40), on the GPU. requestGenerator/parse_model (gRPC plumbing) are out of scope."""
import numpy as np

from ..clients import wire

_SCALING = {"NONE": (1.0, (0, 0, 0)), "INCEPTION": (127.5, (1, 1, 1)), "VGG": (1.0, (123, 117, 104)), "COCO": (255.0, (0, 0, 0))}


def model_dtype_to_np(model_dtype):
    table = {"BOOL": bool, "INT8": np.int8, "INT16": np.int16, "INT32": np.int32, "INT64": np.int64, "UINT8": np.uint8,
             "UINT16": np.uint16, "FP16": np.float16, "FP32": np.float32, "FP64": np.float64, "BYTES": np.dtype(object)}
    return table.get(model_dtype)


def image_adjust(img, w, h, scaling="COCO", engine=None):
    """utils/preprocess.py:169-175 (the reachable cv2 path): BGR frame (ndarray, or a path read with cv2.imread) ->
    resize (w,h) INTER_LINEAR -> BGR2RGB -> scaling -> CHW -> batch dim, fused in one kernel. Returns (1,3,h,w) float32."""
    if isinstance(img, str):
        import cv2
        img = cv2.imread(img)
    eng = engine or wire.default_engine()
    st = wire.Stager(eng.device)
    div, sub = _SCALING[scaling]
    src = st.to_device(np.ascontiguousarray(img), np.uint8, tuple(img.shape), "pre_in")
    out, _ = eng.preprocess_images(src, int(h), int(w), mode="stretch", swap=True, div=div, sub=sub)
    return st.to_host(out, "pre_out")


def image_adjust_ros(cv_image, pad_rows=16, engine=None):
    """utils/preprocess.py:177-183: append pad_rows zero rows, CHW, /255 (the reference forgets to return it)."""
    eng = engine or wire.default_engine()
    st = wire.Stager(eng.device)
    h, w = cv_image.shape[:2]
    src = st.to_device(np.ascontiguousarray(cv_image), np.uint8, tuple(cv_image.shape), "pre_in")
    out, _ = eng.preprocess_images(src, h + pad_rows, w, mode="pad", div=255.0, pad_value=0)
    return st.to_host(out, "pre_out")
