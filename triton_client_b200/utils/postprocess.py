"""Mirror of the reference's utils/postprocess.py free functions (the older twins of clients/postprocess): same
names, arguments and return values; these RAISE on errors where the class version returns them."""
import numpy as np
import torch

from ..clients import wire
from ..clients.postprocess.yolov5_postprocess import Yolov5postprocess
from ..clients.postprocess.detectron_postprocess import FCOSpostprocess

_post = None


def _p():
    global _post
    if _post is None:
        _post = Yolov5postprocess()
    return _post


def deserialize_bytes_float(encoded_tensor):      # utils/postprocess.py:12-22
    return np.frombuffer(encoded_tensor, dtype="<f4").astype(np.float64).astype(np.object_)


def deserialize_bytes_int(encoded_tensor):        # utils/postprocess.py:23-34 — 4-byte 'i' here, 8-byte 'l' in clients/
    return np.frombuffer(encoded_tensor, dtype="<i4").astype(np.object_)


def xywh2xyxy(x):                                 # utils/postprocess.py:36-43
    return _p().xywh2xyxy(x)


def box_iou(box1, box2):                          # utils/postprocess.py:45-67
    return _p().box_iou(box1, box2)


def nms_cpu(boxes, confs, nms_thresh=0.5, min_mode=False):   # utils/postprocess.py:69-103
    return _p().nms_cpu(boxes, confs, nms_thresh, min_mode)


def extract_boxes_yolov5(prediction, conf_thres=0.6, iou_thres=0.45, classes=None, agnostic=False, multi_label=False, labels=(), max_det=300):
    """utils/postprocess.py:105-199 — like Yolov5postprocess.extract_boxes but exceptions propagate."""
    res = _p().extract_boxes(prediction, conf_thres, iou_thres, classes, agnostic, multi_label, labels, max_det)
    if isinstance(res, Exception):
        raise res
    return res


def extract_boxes_onnx(conf_thresh, nms_thresh, output):
    """utils/postprocess.py:265-322: output = [boxes [B,N,1,4], confs [B,N,nc]] -> list (per image) of rows
    [x1,y1,x2,y2,conf,conf,cls], grouped by class, score-descending inside a class."""
    eng = wire.default_engine()
    box_array, confs = output[0], output[1]
    b = torch.as_tensor(np.asarray(box_array) if not isinstance(box_array, torch.Tensor) else box_array, dtype=torch.float32).to(eng.device)
    c = torch.as_tensor(np.asarray(confs) if not isinstance(confs, torch.Tensor) else confs, dtype=torch.float32).to(eng.device)
    rows, cnt = eng.yolo4_post(b, c, conf_thresh, nms_thresh)
    rows, cnt = rows.cpu().numpy(), cnt.cpu().numpy()
    if cnt[1].any() or (cnt[0] >= rows.shape[1]).any():
        # the reference has no bound; the kernel keeps 4096 candidates / 1024 rows per image and says when that cut anything
        raise OverflowError("extract_boxes_onnx: more than 4096 rows above conf_thresh or 1024 kept rows in an image "
                            "(images %s): raise conf_thresh" % np.nonzero(cnt[1] | (cnt[0] >= rows.shape[1]))[0].tolist())
    cnt = cnt[0]
    out = []
    for i in range(rows.shape[0]):
        r = rows[i, :cnt[i]]
        out.append([[x[0], x[1], x[2], x[3], x[4], x[5], np.int64(x[6])] for x in r])
    return out


def extract_boxes_triton(response):
    """utils/postprocess.py:201-262 calls an undefined deserialize_bytes_tensor (:208) and cannot run in the reference;
    here it unpacks confs/boxes with frombuffer and applies the same thresholds (0.4 / 0.6)."""
    confs = np.frombuffer(response.raw_output_contents[0], dtype="<f4").reshape(response.outputs[0].shape)
    boxes = np.frombuffer(response.raw_output_contents[1], dtype="<f4").reshape(response.outputs[1].shape)
    return extract_boxes_onnx(0.4, 0.6, [boxes, confs])


def extract_boxes_detectron(prediction):          # utils/postprocess.py:368-468
    return FCOSpostprocess().extract_boxes(prediction)


def plot_boxes_cv2(img, boxes, savename=None, class_names=None, color=None):      # utils/postprocess.py:324-366 (= tools/utils.py:99-141)
    from ..clients.overlay import plot_boxes_cv2 as _plot
    return _plot(img, boxes, savename, class_names, color)
