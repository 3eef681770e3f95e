"""Box rescale + detection overlay on the GPU.

The reference does this inline in its ROS callback, on the host, one detection at a time
(communicator/ros_inference.py:100-115 `_scale_boxes`, :148-169 cv2.rectangle + cv2.putText;
communicator/evaluate_inference.py:237-256 `_scale_box_array`). `Overlay` is that loop as one call:

    ov = Overlay(class_names, color=(0, 255, 0), input_size=[h, w])
    annotated = ov.annotate(orig_image, prediction)          # numpy in -> numpy out, like the callback's result
    ov.annotate_device(frames, det, counts)                   # everything stays on the GPU

`prediction` is what `Yolov5postprocess.extract_boxes` returns (a list of (n,6) tensors; an exception object draws
nothing, like the callback's isinstance check). No CPU path: without the CUDA library these calls raise.
"""
import numpy as np
import torch

from . import wire


def _numpy2():
    return int(np.__version__.split(".")[0]) >= 2


class Overlay:
    def __init__(self, class_names, color=(0, 255, 0), input_size=(640, 640), thickness=2, labels=True, math_mode=None, engine=None):
        self.class_names = list(class_names)
        self.color = tuple(int(c) for c in color)
        self.input_size = [int(input_size[0]), int(input_size[1])]
        self.thickness, self.labels = int(thickness), bool(labels)
        # `np.float32 scalar * Python number` is float64 under numpy < 2 and float32 from numpy 2 on; default to what the
        # reference's own expression does in this interpreter
        self.math_mode = (2 if _numpy2() else 1) if math_mode is None else int(math_mode)
        self._engine = engine

    @property
    def engine(self):
        if self._engine is None:
            self._engine = wire.default_engine()
        return self._engine

    # -- _scale_boxes / _scale_box_array ----------------------------------------------------------------
    def scale_boxes(self, box, orig_size, normalized=False):
        """`_scale_boxes` (ros_inference.py:100-115): one float32 box [x1,y1,x2,y2] -> [xtl, ytl, xbr, ybr]."""
        row = torch.zeros((1, 6), dtype=torch.float64)
        row[0, :4] = torch.as_tensor(np.asarray(box, dtype=np.float32).astype(np.float64))
        out = self.engine.scale_boxes(row.to(self.engine.device), orig_size, self.input_size, normalized, self.math_mode).cpu().numpy()[0]
        cast = np.float32 if self.math_mode == 2 else np.float64
        return [cast(v) for v in out[:4]]

    def scale_box_array(self, box, orig_size, normalized=False):
        """`_scale_box_array` (evaluate_inference.py:237-256): (n,6) -> (n,6) float64 numpy."""
        t = torch.as_tensor(np.asarray(box, dtype=np.float64)) if not isinstance(box, torch.Tensor) else box.to(torch.float64)
        if t.numel() == 0:
            return np.zeros((0, 6), np.float64)
        return self.engine.scale_boxes(t.reshape(-1, 6).to(self.engine.device), orig_size, self.input_size, normalized, 0).cpu().numpy()

    # -- the drawing loop ---------------------------------------------------------------------------------
    def annotate_device(self, frames, det, counts):
        """frames [b,h,w,3] (or [h,w,3]) uint8 CUDA, modified in place; det [b,max_det,6] float64, counts [b] int32."""
        return self.engine.draw_detections(frames, det, counts, self.class_names, self.color, self.input_size, self.thickness,
                                           self.labels, self.math_mode)

    def annotate(self, orig_image, prediction):
        """One frame, host in / host out: the `for object in self.prediction[0]` loop of the callback."""
        img = np.ascontiguousarray(orig_image)
        if img.dtype != np.uint8 or img.ndim != 3 or img.shape[2] != 3:
            raise TypeError("annotate: orig_image must be (H,W,3) uint8")
        if isinstance(prediction, BaseException) or prediction is None:
            return img.copy()
        det = prediction[0] if isinstance(prediction, (list, tuple)) else prediction
        det = torch.as_tensor(np.asarray(det, dtype=np.float64)) if not isinstance(det, torch.Tensor) else det.to(torch.float64)
        det = det.reshape(-1, 6)
        if det.shape[0] == 0:
            return img.copy()
        dev = self.engine.device
        frame = torch.from_numpy(img).to(dev)
        self.annotate_device(frame, det.to(dev).unsqueeze(0), None)
        return frame.cpu().numpy()


def _plot_colors(num_classes):
    """plot_boxes_cv2's get_color table (clients/postprocess/yolov5_postprocess.py:130-161): (red, green, blue) per class id."""
    import math
    colors = np.array([[1, 0, 1], [0, 0, 1], [0, 1, 1], [0, 1, 0], [1, 1, 0], [1, 0, 0]], dtype=np.float32)

    def get_color(c, x, max_val):
        ratio = float(x) / max_val * 5
        i, j = int(math.floor(ratio)), int(math.ceil(ratio))
        ratio = ratio - i
        return int(((1 - ratio) * colors[i][c] + ratio * colors[j][c]) * 255)
    table = []
    for cls_id in range(num_classes):
        offset = cls_id * 123457 % num_classes
        table.append((get_color(2, offset, num_classes), get_color(1, offset, num_classes), get_color(0, offset, num_classes)))
    return table


def plot_boxes_cv2(img, boxes, savename=None, class_names=None, color=None, engine=None):
    """Yolov5postprocess.plot_boxes_cv2 (clients/postprocess/yolov5_postprocess.py:127-169; twins tools/utils.py:99 and
    utils/postprocess.py:324) on the GPU: `boxes` rows hold normalised x1, y1, x2, y2 and, when they have 7 entries and
    class_names is given, cls_conf at [5] and cls_id at [6]; every box is drawn with cv2.rectangle(thickness 1) in its
    class's get_color colour (or `color`, or red), labelled with cv2.putText(class name, (x1, y1), FONT_HERSHEY_SIMPLEX,
    1.2, rgb, 1). Returns the annotated copy (the input is not modified); `savename` writes it with cv2.imwrite like the
    reference. Pixels are those OpenCV paints."""
    eng = engine or wire.default_engine()
    img = np.ascontiguousarray(img)
    out = torch.from_numpy(img.copy()).to(eng.device)
    rows = [list(b) for b in boxes]
    if rows:
        labelled = bool(class_names) and all(len(b) >= 7 for b in rows)
        det = np.zeros((len(rows), 6), np.float64)
        for k, b in enumerate(rows):
            det[k, :4] = [float(np.float32(v)) for v in b[:4]]
            if labelled:
                det[k, 4], det[k, 5] = float(b[5]), int(b[6])
                print('%s: %f' % (class_names[int(b[6])], b[5]))      # the reference prints every labelled box (:155)
        if labelled:
            names = list(class_names)
            table = [tuple(int(v) for v in color)] * len(names) if color else _plot_colors(len(names))
            eng.draw_detections(out, torch.from_numpy(det).to(eng.device).unsqueeze(0), None, names, table, (1, 1), thickness=1, labels="name",
                                math_mode=2 if _numpy2() else 1, normalized=True, per_class_color=True, font_scale=1.2, text_thickness=1, text_dy=0)
        else:
            eng.draw_detections(out, torch.from_numpy(det).to(eng.device).unsqueeze(0), None, [""], tuple(int(v) for v in color) if color else (255, 0, 0),
                                (1, 1), thickness=1, labels=False, math_mode=2 if _numpy2() else 1, normalized=True)
    res = out.cpu().numpy()
    if savename:
        import cv2
        print("save plot results to %s" % savename)
        cv2.imwrite(savename, res)
    return res
