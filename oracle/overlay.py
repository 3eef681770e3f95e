"""CPU restatement of the step right after the path: box rescale + overlay drawing.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py). Follows, per detection of a frame,
communicator/ros_inference.py:148-169:
    box = np.array(object[0:4], dtype=np.float32); box = self._scale_boxes(box)            (:100-115)
    cv2.rectangle(orig_image, (int(box[0]), int(box[1])), (int(box[2]), int(box[3])), color, thickness=2)
    cv2.putText(orig_image, '{:.2f} {}'.format(conf, name), (int(box[0]), int(box[1] - 10)),
                cv2.FONT_HERSHEY_SIMPLEX, fontScale=0.5, thickness=2, color)
and communicator/evaluate_inference.py:237-256 (_scale_box_array, float64 array arithmetic).

and, per box, Yolov5postprocess.plot_boxes_cv2 (clients/postprocess/yolov5_postprocess.py:127-169: normalised boxes,
per-class colours, putText of the class name at scale 1.2 thickness 1, rectangle thickness 1).

Pinned: tests/golden/overlay.npz holds frames annotated by the unmodified RosInference._callback and plot_boxes_cv2 run
in the build container (oracle/make_golden.py::golden_overlay), and tests/test_oracle_overlay.py checks rect_mask /
text_mask against cv2 itself. OpenCV is a dependency of the reference, not part of its tree: the rectangle rule below is
the observed raster of cv2.rectangle (thickness 1: the outline; thickness 2: the outline grown by its 4-neighbourhood),
the text is oracle/cvdraw.py's restatement of putText (Hershey strokes + OpenCV's line / polygon / disc rasterisers),
identical to cv2.putText for every scale, thickness and position, labels cut by the frame edge included.
"""
import os

import numpy as np

from . import cvdraw


def scale_box(box4, orig_size, input_size, normalized=False, f64=True):
    """_scale_boxes (ros_inference.py:100-115) on one float32 box. The reference multiplies numpy float32 scalars by
    Python numbers: numpy < 2 promotes that to float64 (f64=True, the reference's era), numpy >= 2 keeps float32."""
    b = np.asarray(box4, np.float32)
    oh, ow = int(orig_size[0]), int(orig_size[1])
    if f64:
        b = b.astype(np.float64)
        if normalized:
            return [b[0] * ow, b[1] * oh, b[2] * ow, b[3] * oh]
        rx = ow / input_size[0]
        return [b[0] * rx, b[1] * oh / input_size[1], b[2] * rx, b[3] * oh / input_size[1]]
    f = np.float32
    if normalized:
        return [b[0] * f(ow), b[1] * f(oh), b[2] * f(ow), b[3] * f(oh)]
    rx = f(ow / input_size[0])
    return [b[0] * rx, b[1] * f(oh) / f(input_size[1]), b[2] * rx, b[3] * f(oh) / f(input_size[1])]


def scale_box_array(det, orig_size, input_size, normalized=False):
    """_scale_box_array (evaluate_inference.py:237-256): (n,6) float64 in, (n,6) float64 out."""
    d = np.asarray(det, np.float64).reshape(-1, 6)
    oh, ow = int(orig_size[0]), int(orig_size[1])
    out = d.copy()
    if normalized:
        out[:, 0], out[:, 2] = d[:, 0] * ow, d[:, 2] * ow
        out[:, 1], out[:, 3] = d[:, 1] * oh, d[:, 3] * oh
    else:
        rx = ow / input_size[0]
        out[:, 0], out[:, 2] = d[:, 0] * rx, d[:, 2] * rx
        out[:, 1], out[:, 3] = d[:, 1] * oh / input_size[1], d[:, 3] * oh / input_size[1]
    return out


def rect_mask(h, w, x1, y1, x2, y2, thickness):
    """Pixels cv2.rectangle(img, (x1,y1), (x2,y2), color, thickness) writes, thickness 1 or 2."""
    xa, xb, ya, yb = min(x1, x2), max(x1, x2), min(y1, y2), max(y1, y2)
    ys, xs = np.mgrid[0:h, 0:w]

    def outline(x, y):
        return ((x >= xa) & (x <= xb) & ((y == ya) | (y == yb))) | ((y >= ya) & (y <= yb) & ((x == xa) | (x == xb)))
    m = outline(xs, ys)
    if thickness == 2:
        m = m | outline(xs - 1, ys) | outline(xs + 1, ys) | outline(xs, ys - 1) | outline(xs, ys + 1)
    elif thickness != 1:
        raise ValueError("thickness 1 or 2")
    return m


def text_mask(h, w, text, ox, oy, font_scale=0.5, thickness=2):
    """Pixels cv2.putText(img, text, (ox,oy), FONT_HERSHEY_SIMPLEX, font_scale, color, thickness) writes."""
    out = np.zeros((h, w), bool)
    cvdraw.put_text(out, "".join(c if 32 <= ord(c) < 127 else "?" for c in text), (int(ox), int(oy)), font_scale, thickness)
    return out


def label_text(conf, name):
    return "{:.2f} {}".format(float(conf), name)


def draw_detections(frame, det, class_names, color, input_size, thickness=2, labels=True, f64=True, normalized=False,
                    per_class_color=False, font_scale=0.5, text_thickness=2, text_dy=-10):
    """The loop body of RosInference._callback over one frame. frame (H,W,3) u8 is modified in place and returned;
    det (n,6) float64 rows [x1,y1,x2,y2,conf,cls] in model-input pixels. per_class_color: `color` is a table indexed
    by class (plot_boxes_cv2's colouring, yolov5_postprocess.py:127-169); later detections paint over earlier ones.
    labels: False, True ('{:.2f} {name}') or "name" (the class name alone, plot_boxes_cv2)."""
    h, w = frame.shape[:2]
    for row in np.asarray(det, np.float64).reshape(-1, 6):
        col = np.asarray(color[int(row[5])] if per_class_color else color, np.uint8)
        b = scale_box(row[:4].astype(np.float32), (h, w), input_size, normalized, f64)
        x1, y1, x2, y2 = (int(v) for v in b)
        frame[rect_mask(h, w, x1, y1, x2, y2, thickness)] = col
        if labels:
            text = class_names[int(row[5])] if labels == "name" else label_text(row[4], class_names[int(row[5])])
            oy = y1 if text_dy == 0 else int(b[1] + (np.float64(text_dy) if f64 else np.float32(text_dy)))
            frame[text_mask(h, w, text, x1, oy, font_scale, text_thickness)] = col
    return frame


def plot_colors(num_classes):
    """The per-class colour table of plot_boxes_cv2's get_color (yolov5_postprocess.py:130-161): (red, green, blue) per class id."""
    import math
    colors = np.array([[1, 0, 1], [0, 0, 1], [0, 1, 1], [0, 1, 0], [1, 1, 0], [1, 0, 0]], dtype=np.float32)

    def get_color(c, x, max_val):
        ratio = float(x) / max_val * 5
        i, j = int(math.floor(ratio)), int(math.ceil(ratio))
        ratio = ratio - i
        return int(((1 - ratio) * colors[i][c] + ratio * colors[j][c]) * 255)
    out = []
    for cls_id in range(num_classes):
        offset = cls_id * 123457 % num_classes
        out.append((get_color(2, offset, num_classes), get_color(1, offset, num_classes), get_color(0, offset, num_classes)))
    return out
