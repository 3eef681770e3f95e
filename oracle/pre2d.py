"""Oracle: 2D image preprocessing. Test infrastructure only.

cv2 (opencv-python, pinned 4.5.1.48 by the reference's environment.yml:102; 4.13.0 in this image) is
the third-party library the reference itself calls for resampling, so the oracle calls it too;
`cv_resize_linear_u8` is an independent numpy restatement of its 8-bit INTER_LINEAR arithmetic that
tests/test_oracle_pre2d.py checks against cv2 bit for bit.
"""
import numpy as np

try:  # cv2 is present in the build image and on the GPU box; keep the import soft for the ABI-only tests
    import cv2
except Exception:  # pragma: no cover
    cv2 = None


# --- reference: communicator/ros_inference.py:140, bag_inference2d.py:110, utils/preprocess.py:170 ----
def stretch_resize(img, out_w, out_h):
    """cv2.resize(img, (w, h)) with the default INTER_LINEAR — aspect ratio not preserved."""
    return cv2.resize(img, (int(out_w), int(out_h)))


# --- reference: clients/preprocess/yolov5_preprocess.py:12-24 -------------------------------------
def image_adjust_yolov5(cv_image):
    x = np.transpose(cv_image, (2, 0, 1)).astype(np.float32)
    x = np.expand_dims(x, axis=0)
    x /= 255.0
    return x.astype(np.float32)


# --- reference: clients/preprocess/detectron_preprocess.py:12-24 ----------------------------------
def image_adjust_fcos(cv_image):
    x = np.transpose(cv_image, (2, 0, 1)).astype(np.float32)
    return np.expand_dims(x, axis=0).astype(np.float32)


# --- reference: utils/preprocess.py:127-175 (scaling modes :147-159, cv2 path :169-175) -------------
def image_adjust_utils(bgr_image, w, h, scaling="COCO"):
    """resize(w,h) -> BGR2RGB -> scaling -> CHW -> batch dim, for an already-decoded BGR frame."""
    img = cv2.resize(bgr_image, (w, h))
    img = img[:, :, ::-1]  # cv2.cvtColor(img, cv2.COLOR_BGR2RGB)
    typed = img.astype(np.float32)
    if scaling == "INCEPTION":
        scaled = (typed / 127.5) - 1
    elif scaling == "VGG":
        scaled = typed - np.asarray((123, 117, 104), dtype=np.float32)
    elif scaling == "COCO":
        scaled = typed / 255.0
    else:
        scaled = typed
    return np.expand_dims(np.transpose(scaled, (2, 0, 1)), 0).astype(np.float32)


# --- reference: utils/preprocess.py:177-183 (image_adjust_ros; dead code, returns nothing) ----------
def image_adjust_ros(cv_image, pad_rows=16):
    pad = np.zeros((pad_rows, cv_image.shape[1], 3), dtype=np.uint8)
    x = np.concatenate((cv_image, pad), axis=0)
    x = np.transpose(x, (2, 0, 1)).astype(np.float32)
    x = np.expand_dims(x, axis=0)
    x /= 255.0
    return x


# --- NOT in the reference: ultralytics/yolov5 letterbox, parity unpinned ----------
def letterbox_geometry(h, w, out_h, out_w):
    r = min(out_h / h, out_w / w)
    new_w, new_h = int(round(w * r)), int(round(h * r))
    new_w = min(max(new_w, 1), out_w)
    new_h = min(max(new_h, 1), out_h)
    dw, dh = (out_w - new_w) / 2, (out_h - new_h) / 2
    top, bottom = int(round(dh - 0.1)), int(round(dh + 0.1))
    left, right = int(round(dw - 0.1)), int(round(dw + 0.1))
    return r, new_w, new_h, left, top, right, bottom


def letterbox(img, out_h, out_w, pad_value=114):
    h, w = img.shape[:2]
    r, new_w, new_h, left, top, right, bottom = letterbox_geometry(h, w, out_h, out_w)
    if (w, h) != (new_w, new_h):
        img = cv2.resize(img, (new_w, new_h), interpolation=cv2.INTER_LINEAR)
    out = np.full((out_h, out_w, 3), pad_value, dtype=np.uint8)
    out[top:top + new_h, left:left + new_w] = img
    return out, (r, left, top)


# --- independent restatement of OpenCV's u8 INTER_LINEAR -----------------------------
def _taps(dst, src, clamp_weights):
    scale = 1.0 / (dst / src)
    d = np.arange(dst)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int32)
    f = f - s.astype(np.float32)
    if clamp_weights:  # columns: taps outside the image get weight 0
        lo = s < 0
        f[lo] = 0
        s[lo] = 0
        hi = s >= src - 1
        f[hi] = 0
        s[hi] = src - 1
    w1 = np.rint(f * np.float32(2048)).astype(np.int32)
    w0 = np.rint((np.float32(1) - f) * np.float32(2048)).astype(np.int32)
    return s, w0, w1


def cv_resize_linear_u8(img, out_w, out_h):
    """Bit-exact numpy model of cv2.resize(u8 HxWx3, INTER_LINEAR): 11-bit taps, two-pass fixed point,
    2x2 averaging when both scale factors are exactly 2."""
    h, w = img.shape[:2]
    i = img.astype(np.int32)
    if w == 2 * out_w and h == 2 * out_h:
        return ((i[0::2, 0::2] + i[0::2, 1::2] + i[1::2, 0::2] + i[1::2, 1::2] + 2) >> 2).astype(np.uint8)
    sx, a0, a1 = _taps(out_w, w, True)
    sy, b0, b1 = _taps(out_h, h, False)
    sx1 = np.minimum(sx + 1, w - 1)
    hp = i[:, sx, :] * a0[None, :, None] + i[:, sx1, :] * a1[None, :, None]
    s0 = hp[np.clip(sy, 0, h - 1)]
    s1 = hp[np.clip(sy + 1, 0, h - 1)]
    out = (((b0[:, None, None] * (s0 >> 4)) >> 16) + ((b1[:, None, None] * (s1 >> 4)) >> 16) + 2) >> 2
    return out.astype(np.uint8)


def preprocess_frame(img, out_h, out_w, mode="stretch", swap=False, div=255.0, sub=(0.0, 0.0, 0.0), pad_value=114):
    """Whole fused op as the CUDA kernel performs it, composed from the pieces above."""
    if mode == "stretch":
        x = stretch_resize(img, out_w, out_h) if img.shape[:2] != (out_h, out_w) else img
    elif mode == "letterbox":
        x, _ = letterbox(img, out_h, out_w, pad_value)
    elif mode == "pad":
        x = np.full((out_h, out_w, 3), pad_value, dtype=np.uint8)
        x[: img.shape[0], : img.shape[1]] = img
    else:
        raise ValueError(mode)
    if swap:
        x = x[:, :, ::-1]
    t = np.transpose(x, (2, 0, 1)).astype(np.float32)
    t = t / np.float32(div) - np.asarray(sub, dtype=np.float32)[:, None, None]
    return t[None].astype(np.float32)
