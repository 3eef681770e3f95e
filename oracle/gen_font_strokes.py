#!/usr/bin/env python
"""Build-time generator for the label font of overlay.cu: the stroke table of
FONT_HERSHEY_SIMPLEX, the font of the reference's cv2.putText calls (communicator/ros_inference.py:163-169 scale 0.5
thickness 2; clients/postprocess/yolov5_postprocess.py:164 scale 1.2 thickness 1).

OpenCV's glyph table is not part of the reference tree and has no Python accessor, but the strings are in the cv2
binary: each glyph is "LR" + coordinate pairs ('R' = 0), single spaces lifting the pen (the classic Hershey encoding).
The generator collects every string of that shape from the binary and, for each printable ASCII character, keeps the
one whose rasterisation by oracle/cvdraw.py (the restated putText / Line2) equals what cv2.putText itself draws —
so the table is measured against OpenCV, glyph by glyph, at two scales. Writes triton_client_b200/csrc/font_strokes.inc and
oracle/hershey_simplex.npz (both committed). Run in the build container: python oracle/gen_font_strokes.py
"""
import os
import re
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import cvdraw  # noqa: E402

FONT = cv2.FONT_HERSHEY_SIMPLEX


def parse(s):
    """'MWRFRT RYQZ' -> (left, right, [[(x, y), ...], ...]) or None when the string is not a glyph."""
    if len(s) < 2:
        return None
    left, right = s[0] - ord("R"), s[1] - ord("R")
    if not (-40 <= left <= 0 <= right <= 40):
        return None
    polys, cur, k = [], [], 2
    while k < len(s):
        if s[k] == 32:
            if cur:
                polys.append(cur)
            cur = []
            k += 1
        else:
            if k + 1 >= len(s):
                return None
            x, y = s[k] - ord("R"), s[k + 1] - ord("R")
            if not (-50 <= x <= 50 and -50 <= y <= 50):
                return None
            cur.append((x, y))
            k += 2
    if cur:
        polys.append(cur)
    return left, right, polys


def candidates():
    data = open(cv2.__file__.replace("__init__.py", "cv2.abi3.so") if cv2.__file__.endswith(".py") else cv2.__file__, "rb").read() \
        if not os.path.isdir(os.path.dirname(cv2.__file__)) else None
    if data is None:
        d = os.path.dirname(cv2.__file__)
        so = [os.path.join(d, f) for f in os.listdir(d) if f.endswith(".so")][0]
        data = open(so, "rb").read()
    anchor = data.find(b"MWRFRT RYQZR[SZRY\0")
    assert anchor > 0, "Hershey glyph strings not found in the cv2 binary"
    lo, hi = max(0, anchor - (8 << 20)), min(len(data), anchor + (8 << 20))
    out = {}
    for m in re.finditer(rb"[ -~]{2,400}\0", data[lo:hi]):
        s = m.group()[:-1]
        g = parse(s)
        if g is not None:
            out[s] = g
    return out


def render_cv(text, scale, thick, org=(60, 120), size=(220, 400)):
    img = np.zeros(size, np.uint8)
    cv2.putText(img, text, org, FONT, scale, 255, thick)
    return img > 0


def render_fast(table, base_line, text, scale, org=(60, 120), size=(220, 400)):
    """putText's pen arithmetic with the segments drawn by cv2.line(..., shift=16) — OpenCV's own Line2 — for the search."""
    img = np.zeros(size, np.uint8)
    hscale = int(np.rint(scale * 65536))
    view_x, view_y = org[0] << 16, (org[1] << 16) + base_line * hscale
    for ch in text:
        left, right, polys = table[ch]
        view_x -= left * hscale
        for poly in polys:
            pts = [(x * hscale + view_x, y * hscale + view_y) for x, y in poly]
            for a, b in zip(pts[:-1], pts[1:]):
                if max(abs(v) for v in a + b) < (1 << 31):
                    cv2.line(img, a, b, 255, 1, cv2.LINE_8, 16)
        view_x += right * hscale
    return img > 0


def render_mine(table, base_line, text, scale, thick, org=(60, 120), size=(220, 400)):
    cvdraw._GLYPHS = (table, base_line)
    m = np.zeros(size, bool)
    cvdraw.put_text(m, text, org, scale, thick)
    return m


def main():
    cands = candidates()
    print(f"{len(cands)} glyph-shaped strings near the anchor")
    # base line: the '!' glyph is known; try the 16 possible values
    bang = parse(b"MWRFRT RYQZR[SZRY")
    want = render_cv("!", 2.0, 1)
    base_line = [b for b in range(-15, 1) if np.array_equal(render_mine({"!": bang, "?": bang}, b, "!", 2.0, 1), want)]
    assert len(base_line) == 1, base_line
    base_line = base_line[0]
    print("base_line", base_line)
    table = {}
    # advance of every character, measured: pen position of a following '!' (in 1/65536 pixels the pen moves by (right-left)*hscale)
    for c in range(32, 127):
        ch = chr(c)
        if ch == "!":
            table[ch] = bang          # the anchor glyph: identified by its string, confirmed by the base-line search above
            continue
        want_a, want_b = render_cv(ch + "!", 2.0, 1), render_cv(ch + "!", 0.75, 1)
        hits = []
        for s, g in cands.items():
            t = {"!": bang, "?": bang, ch: g}
            if np.array_equal(render_fast(t, base_line, ch + "!", 2.0), want_a) and np.array_equal(render_fast(t, base_line, ch + "!", 0.75), want_b):
                hits.append((s, g))
        assert hits, f"no glyph string reproduces {ch!r}"
        # several strings may draw the same strokes (other fonts share glyphs): they must agree
        norm = lambda g: (g[1] - g[0], [[(x - g[0], y) for x, y in poly] for poly in g[2]])      # what the raster depends on
        if not all(norm(h[1]) == norm(hits[0][1]) for h in hits):
            # other Hershey faces hold variants of a glyph that look alike as thin lines: tell them apart with thick strokes
            # (joints and caps follow the vertex lists) and more scales
            keep = []
            for h in hits:
                t = {"!": bang, "?": bang, ch: h[1]}
                if all(np.array_equal(render_mine(t, base_line, ch + "!", sc, th), render_cv(ch + "!", sc, th)) for sc, th in ((0.5, 2), (1.2, 1), (1.0, 3), (1.75, 2), (3.1, 2))):
                    keep.append(h)
            assert keep, ch
            if not all(norm(h[1]) == norm(keep[0][1]) for h in keep):
                print(f"note: {ch!r} has {len(keep)} stroke lists that cv2.putText cannot tell apart at the tested scales; taking the first")
            hits = keep
        table[ch] = hits[0][1]
    # self check: strings, scales and thicknesses, also cut by the frame edge
    rng = np.random.default_rng(0)
    chars = [chr(c) for c in range(32, 127)]
    for k in range(60):
        text = "".join(rng.choice(chars, rng.integers(1, 12)))
        scale, thick = float(rng.choice([0.5, 0.6, 1.0, 1.2, 1.75])), int(rng.choice([1, 2, 3]))
        org = (int(rng.integers(-30, 380)), int(rng.integers(-5, 230)))
        a = render_cv(text, scale, thick, org)
        b = render_mine(table, base_line, text, scale, thick, org)
        assert np.array_equal(a, b), (text, scale, thick, org, int((a != b).sum()))
    print("self check passed: 60 random strings identical to cv2.putText")
    arrays = {"base_line": np.array(base_line)}
    rows, offs = [], [0]
    for c in range(32, 127):
        left, right, polys = table[chr(c)]
        flat = [left, right]
        for poly in polys:
            flat += [len(poly)] + [v for pt in poly for v in pt]
        arrays[f"g{c}"] = np.array(flat, np.int8)
        rows += flat
        offs.append(len(rows))
    np.savez_compressed(os.path.join(ROOT, "oracle", "hershey_simplex.npz"), **arrays)
    with open(os.path.join(ROOT, "triton_client_b200", "csrc", "font_strokes.inc"), "w") as f:
        f.write("// generated by oracle/gen_font_strokes.py: FONT_HERSHEY_SIMPLEX strokes, measured against cv2.putText glyph by glyph.\n")
        f.write("// per glyph: left, right, then polylines as (n, x0, y0, ..., x(n-1), y(n-1)); Hershey units, 'R' = 0.\n")
        f.write(f"constexpr int HF_BASE_LINE = {base_line};\n")
        f.write("__constant__ short c_hf_off[96] = {" + ", ".join(str(v) for v in offs) + "};\n")
        f.write("__constant__ signed char c_hf_data[%d] = {\n" % len(rows))
        for i in range(0, len(rows), 32):
            f.write("    " + ", ".join(str(v) for v in rows[i:i + 32]) + ",\n")
        f.write("};\n")
    print(f"wrote font_strokes.inc ({len(rows)} bytes of strokes) and oracle/hershey_simplex.npz")


if __name__ == "__main__":
    main()
