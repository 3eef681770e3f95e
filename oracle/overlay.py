"""CPU restatement of the step right after the path: box rescale + overlay drawing.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py). Follows, per detection of a frame,
communicator/ros_inference.py:148-169:
    box = np.array(object[0:4], dtype=np.float32); box = self._scale_boxes(box)            (:100-115)
    cv2.rectangle(orig_image, (int(box[0]), int(box[1])), (int(box[2]), int(box[3])), color, thickness=2)
    cv2.putText(orig_image, '{:.2f} {}'.format(conf, name), (int(box[0]), int(box[1] - 10)),
                cv2.FONT_HERSHEY_SIMPLEX, fontScale=0.5, thickness=2, color)
and communicator/evaluate_inference.py:237-256 (_scale_box_array, float64 array arithmetic).

Pinned: tests/golden/overlay.npz holds frames annotated by the unmodified RosInference._callback run in the build
container (oracle/make_golden.py::golden_overlay), and tests/test_oracle_overlay.py checks rect_mask / text_mask
against cv2 itself when cv2 is importable. OpenCV is a dependency of the reference, not part of its tree: the
rectangle rule below is the observed raster of cv2.rectangle (thickness 1: the outline; thickness 2: the outline
grown by its 4-neighbourhood), the glyph rasters are measured from cv2.putText (csrc/gen_font_atlas.py).
Known difference: a label cut by the frame edge can differ from OpenCV in a few pixels (OpenCV clips each stroke
outline before walking it).
"""
import os

import numpy as np

_ATLAS = None


def atlas():
    global _ATLAS
    if _ATLAS is None:
        z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "hershey_s05_t2.npz"))
        _ATLAS = (z["adv2"].astype(np.int64), z["rows"].astype(np.int64), [int(v) for v in z["cell"]])
    return _ATLAS


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


def text_mask(h, w, text, ox, oy):
    """Pixels cv2.putText(img, text, (ox,oy), FONT_HERSHEY_SIMPLEX, 0.5, color, 2) writes (in-frame labels exactly)."""
    adv2, rows, (cw, ch, y0) = atlas()
    out = np.zeros((h, w), bool)
    pen2 = 0
    for c in text:
        g = ord(c) - 32
        if not 0 <= g < 95:
            g = ord("?") - 32
        px = ox + (pen2 >> 1)
        for r in range(ch):
            bits = int(rows[g, pen2 & 1, r])
            y = oy + y0 + r
            if bits and 0 <= y < h:
                for b in range(cw):
                    if (bits >> b) & 1 and 0 <= px + b < w:
                        out[y, px + b] = True
        pen2 += int(adv2[g])
    return out


def label_text(conf, name):
    return "{:.2f} {}".format(float(conf), name)


def draw_detections(frame, det, class_names, color, input_size, thickness=2, labels=True, f64=True, normalized=False,
                    per_class_color=False):
    """The loop body of RosInference._callback over one frame. frame (H,W,3) u8 is modified in place and returned;
    det (n,6) float64 rows [x1,y1,x2,y2,conf,cls] in model-input pixels. per_class_color: `color` is a table indexed
    by class (plot_boxes_cv2's colouring, yolov5_postprocess.py:127-169); later detections paint over earlier ones."""
    h, w = frame.shape[:2]
    for row in np.asarray(det, np.float64).reshape(-1, 6):
        col = np.asarray(color[int(row[5])] if per_class_color else color, np.uint8)
        b = scale_box(row[:4].astype(np.float32), (h, w), input_size, normalized, f64)
        x1, y1, x2, y2 = (int(v) for v in b)
        frame[rect_mask(h, w, x1, y1, x2, y2, thickness)] = col
        if labels:
            ten = np.float64(10) if f64 else np.float32(10)
            frame[text_mask(h, w, label_text(row[4], class_names[int(row[5])]), x1, int(b[1] - ten))] = col
    return frame
