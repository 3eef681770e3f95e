"""Oracle: the OpenCV drawing primitives behind the reference's overlay calls, restated pixel for pixel.

TEST INFRASTRUCTURE ONLY. The reference draws its detections with cv2.rectangle / cv2.putText
(communicator/ros_inference.py:158-169: thickness 2, FONT_HERSHEY_SIMPLEX scale 0.5) and, in plot_boxes_cv2
(clients/postprocess/yolov5_postprocess.py:127-169), with thickness 1 and scale 1.2. OpenCV is the reference's
dependency (opencv-python), not part of its tree; what follows restates the rasterisation rules of its drawing.cpp —
clipLine, Line (LineIterator's Bresenham, what thin lines use after their endpoints are rounded to pixels), Line2 (the
16.16 fixed-point line that outlines filled polygons), FillConvexPoly, the filled Circle, ThickLine, PolyLine and putText's
pen arithmetic — and is pinned against cv2 itself by tests/test_oracle_overlay.py
(random segments, quads, circles and strings at several scales, thicknesses and frame-edge positions).
Glyph strokes come from oracle/hershey_simplex.npz, written by oracle/gen_font_strokes.py.
"""
import math
import os

import numpy as np

XY_SHIFT = 16
XY_ONE = 1 << XY_SHIFT


def _trunc(v):
    return int(v)          # C cast of a double to int64: towards zero


def clip_line(width, height, p1, p2):
    """clipLine(Size2l, Point2l&, Point2l&): integer endpoints, double arithmetic for the cuts. Returns (inside, p1, p2)."""
    x1, y1 = p1
    x2, y2 = p2
    right, bottom = width - 1, height - 1
    if width <= 0 or height <= 0:
        return False, p1, p2
    c1 = (x1 < 0) + (x1 > right) * 2 + (y1 < 0) * 4 + (y1 > bottom) * 8
    c2 = (x2 < 0) + (x2 > right) * 2 + (y2 < 0) * 4 + (y2 > bottom) * 8
    if (c1 & c2) == 0 and (c1 | c2) != 0:
        if c1 & 12:
            a = 0 if c1 < 8 else bottom
            x1 += _trunc(float(a - y1) * float(x2 - x1) / float(y2 - y1))
            y1 = a
            c1 = (x1 < 0) + (x1 > right) * 2
        if c2 & 12:
            a = 0 if c2 < 8 else bottom
            x2 += _trunc(float(a - y2) * float(x2 - x1) / float(y2 - y1))
            y2 = a
            c2 = (x2 < 0) + (x2 > right) * 2
        if (c1 & c2) == 0 and (c1 | c2) != 0:
            if c1:
                a = 0 if c1 == 1 else right
                y1 += _trunc(float(a - x1) * float(y2 - y1) / float(x2 - x1))
                x1 = a
                c1 = 0
            if c2:
                a = 0 if c2 == 1 else right
                y2 += _trunc(float(a - x2) * float(y2 - y1) / float(x2 - x1))
                x2 = a
                c2 = 0
    return (c1 | c2) == 0, (x1, y1), (x2, y2)


def _cdiv(a, b):
    """C integer division (towards zero) of Python ints."""
    q = abs(a) // abs(b)
    return q if (a >= 0) == (b >= 0) else -q


def line2(mask, p1, p2):
    """Line2: endpoints in 16.16 fixed point; sets the pixels it writes in the boolean mask."""
    h, w = mask.shape
    ok, p1, p2 = clip_line(w << XY_SHIFT, h << XY_SHIFT, p1, p2)
    if not ok:
        return
    (x1, y1), (x2, y2) = p1, p2
    dx, dy = x2 - x1, y2 - y1
    ax, ay = abs(dx), abs(dy)

    def put(x, y):
        if 0 <= x < w and 0 <= y < h:
            mask[y, x] = True
    if ax > ay:
        if dx < 0:
            dy = -dy
            x1, x2, y1, y2 = x2, x1, y2, y1
        y_step = _cdiv(dy << XY_SHIFT, ax | 1)
        ecount = (x2 - x1) >> XY_SHIFT
        x1 += XY_ONE >> 1
        y1 += XY_ONE >> 1
        put((x2 + (XY_ONE >> 1)) >> XY_SHIFT, (y2 + (XY_ONE >> 1)) >> XY_SHIFT)
        x = x1 >> XY_SHIFT
        while ecount >= 0:
            put(x, y1 >> XY_SHIFT)
            x += 1
            y1 += y_step
            ecount -= 1
    else:
        if dy < 0:
            dx = -dx
            x1, x2, y1, y2 = x2, x1, y2, y1
        x_step = _cdiv(dx << XY_SHIFT, ay | 1)
        ecount = (y2 - y1) >> XY_SHIFT
        x1 += XY_ONE >> 1
        y1 += XY_ONE >> 1
        put((x2 + (XY_ONE >> 1)) >> XY_SHIFT, (y2 + (XY_ONE >> 1)) >> XY_SHIFT)
        y = y1 >> XY_SHIFT
        while ecount >= 0:
            put(x1 >> XY_SHIFT, y)
            x1 += x_step
            y += 1
            ecount -= 1


def line_int(mask, p1, p2):
    """Line(img, pt1, pt2, color, 8): integer endpoints, clipped to the image, then LineIterator's 8-connected Bresenham
    walked from the left endpoint (leftToRight = true)."""
    h, w = mask.shape
    ok, p1, p2 = clip_line(w, h, p1, p2)
    if not ok:
        return
    (x1, y1), (x2, y2) = p1, p2
    dx, dy = x2 - x1, y2 - y1
    if dx < 0:
        dx, dy = -dx, -dy
        x1, y1 = x2, y2
    sy = -1 if dy < 0 else 1
    dy = abs(dy)
    steep = dy > dx
    if steep:
        dx, dy = dy, dx
    err, plus, minus = dx - (dy + dy), dx + dx, -(dy + dy)
    x, y = x1, y1
    for _ in range(dx + 1):
        mask[y, x] = True
        m = err < 0
        err += minus + (plus if m else 0)
        if steep:                       # the major axis is y
            y += sy
            if m:
                x += 1
        else:
            x += 1
            if m:
                y += sy


def fill_convex_poly(mask, v):
    """FillConvexPoly(img, v, npts, color, LINE_8, shift = XY_SHIFT): vertices in 16.16 fixed point."""
    h, w = mask.shape
    n = len(v)
    delta = XY_ONE >> 1
    delta1 = delta2 = XY_ONE >> 1
    p0 = v[-1]
    xmin = xmax = v[0][0]
    ymin = ymax = v[0][1]
    imin = 0
    for i, p in enumerate(v):
        if p[1] < ymin:
            ymin, imin = p[1], i
        ymax, xmax, xmin = max(ymax, p[1]), max(xmax, p[0]), min(xmin, p[0])
        line2(mask, p0, p)
        p0 = p
    xmin, xmax, ymin, ymax = (xmin + delta) >> XY_SHIFT, (xmax + delta) >> XY_SHIFT, (ymin + delta) >> XY_SHIFT, (ymax + delta) >> XY_SHIFT
    if n < 3 or xmax < 0 or ymax < 0 or xmin >= w or ymin >= h:
        return
    ymax = min(ymax, h - 1)
    edges = n
    e = [dict(idx=imin, di=1, x=-XY_ONE, dx=0, ye=ymin), dict(idx=imin, di=n - 1, x=-XY_ONE, dx=0, ye=ymin)]
    y = ymin
    while True:
        for i in range(2):
            if y >= e[i]["ye"]:
                idx0, di = e[i]["idx"], e[i]["di"]
                idx = idx0 + di
                if idx >= n:
                    idx -= n
                while True:
                    edges -= 1
                    if edges < 0:        # `for (; edges-- > 0; )`: the test sees the old value, the decrement always happens
                        break
                    ty = (v[idx][1] + delta) >> XY_SHIFT
                    if ty > y:
                        xs, xe = v[idx0][0], v[idx][0]
                        e[i]["ye"] = ty
                        e[i]["dx"] = _cdiv((xe - xs) * 2 + (ty - y), 2 * (ty - y))
                        e[i]["x"] = xs
                        e[i]["idx"] = idx
                        break
                    idx0 = idx
                    idx += di
                    if idx >= n:
                        idx -= n
        if edges < 0:
            break
        if y >= 0:
            left, right = (1, 0) if e[0]["x"] > e[1]["x"] else (0, 1)
            xx1 = (e[left]["x"] + delta1) >> XY_SHIFT
            xx2 = (e[right]["x"] + delta2) >> XY_SHIFT
            if xx2 >= 0 and xx1 < w:
                mask[y, max(xx1, 0):min(xx2, w - 1) + 1] = True
        e[0]["x"] += e[0]["dx"]
        e[1]["x"] += e[1]["dx"]
        y += 1
        if y > ymax:
            break


def circle_filled(mask, cx, cy, radius):
    """Circle(img, center, radius, color, fill = 1): integer centre."""
    h, w = mask.shape

    def hline(y, xa, xb):
        if 0 <= y < h:
            xa, xb = max(xa, 0), min(xb, w - 1)
            if xa <= xb:
                mask[y, xa:xb + 1] = True
    err, dx, dy, plus, minus = 0, radius, 0, 1, (radius << 1) - 1
    while dx >= dy:
        hline(cy - dy, cx - dx, cx + dx)
        hline(cy + dy, cx - dx, cx + dx)
        hline(cy - dx, cx - dy, cx + dy)
        hline(cy + dx, cx - dy, cx + dy)
        dy += 1
        err += plus
        plus += 2
        m = -1 if err > 0 else 0          # mask = (err <= 0) - 1
        err -= minus & m
        dx += m
        minus -= m & 2


def _cv_round(v):
    return int(np.rint(v))                # cvRound: round half to even


def thick_line(mask, p0, p1, thickness, flags=3):
    """ThickLine(img, p0, p1, color, thickness, LINE_8, flags, shift = XY_SHIFT)."""
    if thickness <= 1:               # thin lines: endpoints rounded to whole pixels, then the integer line
        h = XY_ONE >> 1
        line_int(mask, ((p0[0] + h) >> XY_SHIFT, (p0[1] + h) >> XY_SHIFT), ((p1[0] + h) >> XY_SHIFT, (p1[1] + h) >> XY_SHIFT))
        return
    inv = 1.0 / XY_ONE
    dx, dy = (p0[0] - p1[0]) * inv, (p1[1] - p0[1]) * inv
    r = dx * dx + dy * dy
    odd = thickness & 1
    t = thickness << (XY_SHIFT - 1)
    if abs(r) > 2.220446049250313e-16:
        r = (t + odd * XY_ONE * 0.5) / math.sqrt(r)
        dpx, dpy = _cv_round(dy * r), _cv_round(dx * r)
        fill_convex_poly(mask, [(p0[0] + dpx, p0[1] + dpy), (p0[0] - dpx, p0[1] - dpy), (p1[0] - dpx, p1[1] - dpy), (p1[0] + dpx, p1[1] + dpy)])
    for i, p in enumerate((p0, p1)):
        if flags & (i + 1):
            circle_filled(mask, (p[0] + (XY_ONE >> 1)) >> XY_SHIFT, (p[1] + (XY_ONE >> 1)) >> XY_SHIFT, (t + (XY_ONE >> 1)) >> XY_SHIFT)


def poly_line(mask, pts, thickness):
    flags = 3
    p0 = pts[0]
    for p in pts[1:]:
        thick_line(mask, p0, p, thickness, flags)
        p0 = p
        flags = 2


_GLYPHS = None


def glyphs():
    """FONT_HERSHEY_SIMPLEX: per printable ASCII character (left, right, [polyline, ...]) in Hershey units, and the base line."""
    global _GLYPHS
    if _GLYPHS is None:
        z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "hershey_simplex.npz"))
        out = {}
        for c in range(32, 127):
            d = z[f"g{c}"]
            left, right, k = int(d[0]), int(d[1]), 2
            polys = []
            while k < len(d):
                n = int(d[k])
                k += 1
                polys.append([(int(d[k + 2 * i]), int(d[k + 2 * i + 1])) for i in range(n)])
                k += 2 * n
            out[chr(c)] = (left, right, polys)
        _GLYPHS = (out, int(z["base_line"]))
    return _GLYPHS


def put_text(mask, text, org, font_scale, thickness):
    """putText(img, text, org, FONT_HERSHEY_SIMPLEX, font_scale, color, thickness) -> pixels written."""
    table, base_line = glyphs()
    hscale = _cv_round(font_scale * XY_ONE)
    vscale = hscale
    view_x = org[0] << XY_SHIFT
    view_y = (org[1] << XY_SHIFT) + base_line * vscale
    for ch in text:
        left, right, polys = table.get(ch, table["?"])
        dx = right * hscale
        view_x -= left * hscale
        for poly in polys:
            pts = [(x * hscale + view_x, y * vscale + view_y) for x, y in poly]
            if len(pts) > 1:
                poly_line(mask, pts, thickness)
        view_x += dx
