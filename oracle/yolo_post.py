"""Oracle: YOLOv5 postprocess. Test infrastructure only.

Restates clients/postprocess/yolov5_postprocess.py:28-125 and clients/postprocess/base_postprocess.py
operation for operation, keeping the reference's precisions: the wire bytes become Python floats (f64
values of the f32s), conf = obj*cls is an f64 product, boxes are rounded to f32 after xywh2xyxy, NMS
runs in f64 on boxes offset by cls*4096 through torchvision.ops.nms (the library the reference calls;
requirements.txt:2 pins 0.9.1, this image has 0.26.0). `nms_torchvision_restated` is an independent
numpy restatement of that CPU kernel.
"""
import struct
import numpy as np
import torch
import torchvision


# --- reference: clients/postprocess/base_postprocess.py:15-25 -------------------------------------
def deserialize_bytes_float(encoded_tensor, as_written=True):
    """as_written=True keeps the reference's per-element struct.unpack_from loop (its dominant cost)."""
    if not as_written:
        return np.frombuffer(encoded_tensor, dtype="<f4").astype(np.float64).astype(np.object_)
    vals = []
    offset = 0
    size = struct.calcsize("f")
    while offset < len(encoded_tensor):
        vals.append(struct.unpack_from("f", encoded_tensor, offset)[0])
        offset += size
    return np.array(vals, dtype=np.object_)


# --- reference: clients/postprocess/base_postprocess.py:27-37 ('l' = 8 bytes; utils/postprocess.py:23-34 uses 'i') ---
def deserialize_bytes_int(encoded_tensor, fmt="l"):
    vals = []
    offset = 0
    size = struct.calcsize(fmt)
    while offset < len(encoded_tensor):
        vals.append(struct.unpack_from(fmt, encoded_tensor, offset)[0])
        offset += size
    return np.array(vals, dtype=np.object_)


# --- reference: clients/postprocess/base_postprocess.py:39-46 -------------------------------------
def xywh2xyxy(x):
    y = x.clone() if isinstance(x, torch.Tensor) else np.copy(x)
    y[:, 0] = x[:, 0] - x[:, 2] / 2
    y[:, 1] = x[:, 1] - x[:, 3] / 2
    y[:, 2] = x[:, 0] + x[:, 2] / 2
    y[:, 3] = x[:, 1] + x[:, 3] / 2
    return y


# --- reference: clients/postprocess/base_postprocess.py:48-70 -------------------------------------
def box_iou(box1, box2):
    def box_area(box):
        return (box[2] - box[0]) * (box[3] - box[1])

    area1 = box_area(box1.T)
    area2 = box_area(box2.T)
    inter = (torch.min(box1[:, None, 2:], box2[:, 2:]) - torch.max(box1[:, None, :2], box2[:, :2])).clamp(0).prod(2)
    return inter / (area1[:, None] + area2 - inter)


# --- reference: clients/postprocess/base_postprocess.py:72-106 (= utils/postprocess.py:69-103, tools/utils.py:61-95) ---
def nms_cpu(boxes, confs, nms_thresh=0.5, min_mode=False, stable=False):
    """numpy greedy NMS. `stable=True` replaces the reference's unstable argsort()[::-1] with the tie rule
    the build fixes everywhere (score descending, lower index first) — identical when scores are distinct."""
    x1, y1, x2, y2 = boxes[:, 0], boxes[:, 1], boxes[:, 2], boxes[:, 3]
    areas = (x2 - x1) * (y2 - y1)
    if stable:
        order = np.lexsort((np.arange(len(confs)), -confs))
    else:
        order = confs.argsort()[::-1]
    keep = []
    while order.size > 0:
        idx_self = order[0]
        idx_other = order[1:]
        keep.append(idx_self)
        xx1 = np.maximum(x1[idx_self], x1[idx_other])
        yy1 = np.maximum(y1[idx_self], y1[idx_other])
        xx2 = np.minimum(x2[idx_self], x2[idx_other])
        yy2 = np.minimum(y2[idx_self], y2[idx_other])
        w = np.maximum(0.0, xx2 - xx1)
        h = np.maximum(0.0, yy2 - yy1)
        inter = w * h
        if min_mode:
            over = inter / np.minimum(areas[order[0]], areas[order[1:]])
        else:
            over = inter / (areas[order[0]] + areas[order[1:]] - inter)
        inds = np.where(over <= nms_thresh)[0]
        order = order[inds + 1]
    return np.array(keep)


# --- restatement of torchvision/csrc/ops/cpu/nms_kernel.cpp ---------------------------
def nms_torchvision_restated(boxes, scores, iou_threshold):
    boxes = np.asarray(boxes)
    scores = np.asarray(scores)
    n = boxes.shape[0]
    if n == 0:
        return np.zeros((0,), dtype=np.int64)
    x1, y1, x2, y2 = boxes[:, 0], boxes[:, 1], boxes[:, 2], boxes[:, 3]
    areas = (x2 - x1) * (y2 - y1)
    order = np.argsort(-scores, kind="stable")
    suppressed = np.zeros(n, dtype=bool)
    keep = []
    for _i in range(n):
        i = order[_i]
        if suppressed[i]:
            continue
        keep.append(i)
        rest = order[_i + 1:]
        xx1 = np.maximum(x1[i], x1[rest])
        yy1 = np.maximum(y1[i], y1[rest])
        xx2 = np.minimum(x2[i], x2[rest])
        yy2 = np.minimum(y2[i], y2[rest])
        w = np.maximum(0, xx2 - xx1)
        h = np.maximum(0, yy2 - yy1)
        inter = w * h
        with np.errstate(invalid="ignore", divide="ignore"):
            ovr = inter / (areas[i] + areas[rest] - inter)
        suppressed[rest[ovr > iou_threshold]] = True
    return np.asarray(keep, dtype=np.int64)


class FakeResponse:
    """Stand-in for tritonclient's ModelInferResponse: .outputs[i].shape and .raw_output_contents[i]."""

    class _Out:
        def __init__(self, shape):
            self.shape = list(shape)

    def __init__(self, arrays):
        self.outputs = [FakeResponse._Out(a.shape) for a in arrays]
        self.raw_output_contents = [np.ascontiguousarray(a).tobytes() for a in arrays]


# --- reference: clients/postprocess/yolov5_postprocess.py:28-125 ----------------------------------
def extract_boxes(prediction, conf_thres=0.6, iou_thres=0.45, classes=None, agnostic=False, max_det=300,
                  as_written=True, nms="torchvision", max_nms=30000):
    """`prediction`: FakeResponse (wire bytes) or an f32 ndarray [B,N,5+nc]. Returns what the reference
    returns: list of (n,6) float64 tensors, or the caught exception object (:124-125)."""
    if isinstance(prediction, np.ndarray):
        shape = prediction.shape
        boxes = deserialize_bytes_float(np.ascontiguousarray(prediction, dtype="<f4").tobytes(), as_written)
    else:
        shape = prediction.outputs[0].shape
        boxes = deserialize_bytes_float(prediction.raw_output_contents[0], as_written)
    prediction = np.reshape(boxes, shape)
    nc = prediction.shape[2] - 5
    xc = prediction[..., 4] > conf_thres
    try:
        assert np.any(np.array(xc)) == True, 'No valid detections found'  # noqa: E712  (as the reference writes it)
        assert 0 <= conf_thres <= 1, f'Invalid Confidence threshold {conf_thres}, valid values are between 0.0 and 1.0'
        assert 0 <= iou_thres <= 1, f'Invalid IoU {iou_thres}, valid values are between 0.0 and 1.0'
        max_wh = 4096
        # max_nms: 30000 in the reference (a local there; a parameter here so the tests can reach the cut)
        output = [torch.zeros((0, 6))] * prediction.shape[0]
        for xi, x in enumerate(prediction):
            x = x[xc[xi]]
            if not x.shape[0]:
                continue
            x[:, 5:] *= x[:, 4:5]
            box = xywh2xyxy(x[:, :4]).astype(np.float32)
            x = torch.tensor(x.astype(np.float64))
            conf, j = x[:, 5:].max(1, keepdim=True)
            x = torch.cat((torch.from_numpy(box), conf, j.float()), 1)[conf.view(-1) > conf_thres]
            if classes is not None:
                x = x[(x[:, 5:6] == torch.tensor(classes, device=x.device)).any(1)]
            n = x.shape[0]
            if not n:
                continue
            elif n > max_nms:
                x = x[x[:, 4].argsort(descending=True, stable=True)[:max_nms]]
            c = x[:, 5:6] * (0 if agnostic else max_wh)
            bxs, scores = x[:, :4] + c, x[:, 4]
            if nms == "torchvision":
                i = torchvision.ops.nms(bxs, scores, iou_thres)
            else:
                i = torch.from_numpy(nms_torchvision_restated(bxs.numpy(), scores.numpy(), iou_thres))
            if i.shape[0] > max_det:
                i = i[:max_det]
            output[xi] = x[i]
        return output
    except Exception as e:  # the reference returns the exception instead of raising
        return e
