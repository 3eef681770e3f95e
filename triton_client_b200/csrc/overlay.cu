// Box rescale + overlay drawing on device (the step right after the path).
//
// Replaces, per frame, the host loop of communicator/ros_inference.py:148-169 —
//   box = np.array(object[0:4], float32); box = _scale_boxes(box)                    (:100-115)
//   cv2.rectangle(orig_image, (int(x1),int(y1)), (int(x2),int(y2)), color, thickness=2)
//   cv2.putText(orig_image, '{:.2f} {}'.format(conf, name), (int(x1), int(y1-10)), FONT_HERSHEY_SIMPLEX, 0.5, color, 2)
// and _scale_box_array (communicator/evaluate_inference.py:237-256) — so the detections never leave the device
// between NMS and the annotated frame.
//
// Two kernels: (1) one thread per detection scales the box, truncates to pixel coordinates like Python's int() and
// formats the label text; (2) one CTA per frame walks the frame's detections in order — like the reference's successive
// cv2 calls, the last writer wins — and for each one all threads draw its rectangle, then its label, stroke by stroke,
// straight into the frame. Frames nothing is detected in are not touched.
//
// Raster rules: OpenCV's own (it is the reference's dependency, not in its tree), restated in oracle/cvdraw.py and pinned
// there against cv2 itself: cv2.rectangle thickness 1 = the outline, thickness 2 = the outline grown by its
// 4-neighbourhood; cv2.putText = FONT_HERSHEY_SIMPLEX strokes (font_strokes.inc, measured against cv2.putText glyph by
// glyph by oracle/gen_font_strokes.py) placed with putText's 16.16 pen arithmetic and drawn by ThickLine: thin strokes round
// their endpoints to pixels, clip them to the frame (clipLine, in double) and walk LineIterator's Bresenham; thick
// strokes fill a quad (FillConvexPoly: Line2 outline + edge-stepped spans) and a disc at the polyline's vertices.
// Any scale, any thickness, labels cut by the frame edge included.
#include "common.cuh"

namespace {

#include "font_strokes.inc"

constexpr int MAXC = 64;              // label characters kept (the rest is dropped)
constexpr int NAME_BYTES = 2560, MAX_CLASSES = 128;
constexpr int DRAW_T = 256;
constexpr int XY_SHIFT = 16;
constexpr long long XY_ONE = 1ll << XY_SHIFT, XY_HALF = XY_ONE >> 1;

struct __align__(16) DrawItem {
    int xa, ya, xb, yb;               // rectangle corners, ordered
    int ox, oy;                       // text origin
    int nchar, pad;
    unsigned char rgb[4];
    unsigned char text[MAXC];         // printable ASCII - 32
};

struct DrawParams {
    int h, w, max_det;
    long long row_stride, frame_stride;
    int orig_h, orig_w, in0, in1;     // _scale_boxes: x *= orig_w / input_size[0], y = y * orig_h / input_size[1]
    int normalized, math_mode, thickness, labels, per_class_color, num_classes;
    int text_thickness, text_dy, hscale;   // putText: thickness, origin = (x1, y1 + text_dy), hscale = cvRound(fontScale * 65536)
    unsigned char color[4];
    unsigned char class_color[MAX_CLASSES * 3];
    unsigned short name_off[MAX_CLASSES + 1];
    char names[NAME_BYTES];
};

// math_mode 0: float64 array arithmetic on the float64 detections (_scale_box_array)
//           1: box rounded to float32, then float64 arithmetic (_scale_boxes under numpy < 2)
//           2: box rounded to float32, float32 arithmetic (_scale_boxes under numpy >= 2)
__device__ __forceinline__ void scale4(const double* d, int orig_h, int orig_w, int in0, int in1, int normalized, int mode, double* o) {
    if (mode == 2) {
        const float b0 = (float)d[0], b1 = (float)d[1], b2 = (float)d[2], b3 = (float)d[3];
        if (normalized) {
            o[0] = b0 * (float)orig_w; o[1] = b1 * (float)orig_h; o[2] = b2 * (float)orig_w; o[3] = b3 * (float)orig_h;
        } else {
            const float rx = (float)((double)orig_w / (double)in0);
            o[0] = b0 * rx; o[2] = b2 * rx;
            o[1] = (b1 * (float)orig_h) / (float)in1; o[3] = (b3 * (float)orig_h) / (float)in1;
        }
        return;
    }
    double b0 = d[0], b1 = d[1], b2 = d[2], b3 = d[3];
    if (mode == 1) { b0 = (double)(float)b0; b1 = (double)(float)b1; b2 = (double)(float)b2; b3 = (double)(float)b3; }
    if (normalized) {
        o[0] = b0 * (double)orig_w; o[1] = b1 * (double)orig_h; o[2] = b2 * (double)orig_w; o[3] = b3 * (double)orig_h;
    } else {
        const double rx = (double)orig_w / (double)in0;
        o[0] = b0 * rx; o[2] = b2 * rx;
        o[1] = (b1 * (double)orig_h) / (double)in1; o[3] = (b3 * (double)orig_h) / (double)in1;
    }
}

__global__ void __launch_bounds__(256) scale_boxes_kernel(const double* __restrict__ det, int n, int orig_h, int orig_w, int in0, int in1,
                                                          int normalized, int mode, double* __restrict__ out) {
    pdl_wait();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double o[4];
    scale4(det + 6ll * i, orig_h, orig_w, in0, in1, normalized, mode, o);
    out[6ll * i + 0] = o[0]; out[6ll * i + 1] = o[1]; out[6ll * i + 2] = o[2]; out[6ll * i + 3] = o[3];
    out[6ll * i + 4] = det[6ll * i + 4]; out[6ll * i + 5] = det[6ll * i + 5];
}

// Python's int(float): truncate toward zero (clamped; int() of nan/inf raises there, here it becomes 0 / the clamp)
__device__ __forceinline__ int py_int(double v) {
    if (!(v == v)) return 0;
    if (v >= 1.0e9) return 1000000000;
    if (v <= -1.0e9) return -1000000000;
    return (int)v;
}

// '{:.2f}'.format(v): correctly rounded (ties to even on the exact binary value), like CPython's float formatting.
// An fma computes v*100 - c exactly before its single rounding, so its sign is the sign of the exact difference.
__device__ int format_2f(double v, char* s, int cap) {
    int len = 0;
    auto put = [&](char c) { if (len < cap) s[len] = c; len++; };
    if (!(v == v)) { put('n'); put('a'); put('n'); return min(len, cap); }
    if (signbit(v)) { put('-'); v = -v; }
    if (v > 1.0e15) { put('i'); put('n'); put('f'); return min(len, cap); }   // inf (and numbers no confidence reaches)
    unsigned long long n = (unsigned long long)floor(v * 100.0);          // floor of the rounded product: off by one at most
    if (fma(v, 100.0, -(double)n) < 0.0) n--;
    else if (fma(v, 100.0, -(double)(n + 1ull)) >= 0.0) n++;
    const double d = fma(v, 100.0, -((double)n + 0.5));                    // sign of v*100 - (n + 1/2), exact
    if (d > 0.0 || (d == 0.0 && (n & 1ull))) n++;
    char digits[24];
    int nd = 0;
    unsigned long long ip = n / 100ull;
    do { digits[nd++] = (char)('0' + (int)(ip % 10ull)); ip /= 10ull; } while (ip);
    while (nd) put(digits[--nd]);
    put('.');
    put((char)('0' + (int)((n / 10ull) % 10ull)));
    put((char)('0' + (int)(n % 10ull)));
    return min(len, cap);
}

__global__ void __launch_bounds__(128) draw_items_kernel(const DrawParams p, const double* __restrict__ det, const int* __restrict__ count,
                                                         DrawItem* __restrict__ items) {
    pdl_wait();
    const int f = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = min(count ? count[f] : p.max_det, p.max_det);
    if (i >= n) return;
    const double* d = det + ((long long)f * p.max_det + i) * 6;
    double o[4];
    scale4(d, p.orig_h, p.orig_w, p.in0, p.in1, p.normalized, p.math_mode, o);
    DrawItem it;
    const int x1 = py_int(o[0]), y1 = py_int(o[1]), x2 = py_int(o[2]), y2 = py_int(o[3]);
    it.xa = min(x1, x2); it.xb = max(x1, x2); it.ya = min(y1, y2); it.yb = max(y1, y2);
    it.ox = x1;
    // ros_inference.py:165 int(box[1] - 10): the subtraction happens before the truncation, in the box's arithmetic
    it.oy = p.text_dy == 0 ? y1 : (p.math_mode == 2 ? py_int((double)((float)o[1] + (float)p.text_dy)) : py_int(o[1] + (double)p.text_dy));
    const int cls = py_int(d[5]);
    const bool cls_ok = cls >= 0 && cls < p.num_classes;
    const unsigned char* col = (p.per_class_color && cls_ok) ? p.class_color + 3 * cls : p.color;
    it.rgb[0] = col[0]; it.rgb[1] = col[1]; it.rgb[2] = col[2]; it.rgb[3] = 0;
    it.nchar = 0; it.pad = 0;
    if (p.labels) {
        char text[MAXC];
        int len = 0;
        if (p.labels == 1) {                       // '{:.2f} {}'.format(conf, name)  (ros_inference.py:164)
            len = format_2f(d[4], text, MAXC);
            if (len < MAXC) text[len++] = ' ';
        }
        if (cls_ok)                                // labels == 2: the class name alone (plot_boxes_cv2, yolov5_postprocess.py:164)
            for (int k = p.name_off[cls]; k < p.name_off[cls + 1] && len < MAXC; k++) text[len++] = p.names[k];
        for (int k = 0; k < len; k++) {
            int g = (int)(unsigned char)text[k] - 32;
            if (g < 0 || g >= 95) g = '?' - 32;
            it.text[k] = (unsigned char)g;
        }
        it.nchar = len;
    }
    items[(long long)f * p.max_det + i] = it;
}

// ---------------------------------------------------------------------------------------------------------------------
// OpenCV's rasterisation rules (drawing.cpp), as restated and pinned in oracle/cvdraw.py
// ---------------------------------------------------------------------------------------------------------------------
struct Canvas { unsigned char* px; long long row_stride; int w, h; unsigned char r, g, b; };
__device__ __forceinline__ void put_px(const Canvas& c, int x, int y) {
    if (x >= 0 && x < c.w && y >= 0 && y < c.h) { unsigned char* q = c.px + (long long)y * c.row_stride + 3ll * x; q[0] = c.r; q[1] = c.g; q[2] = c.b; }
}
__device__ __forceinline__ void hline(const Canvas& c, int y, int xa, int xb) {
    if (y < 0 || y >= c.h) return;
    xa = max(xa, 0); xb = min(xb, c.w - 1);
    for (int x = xa; x <= xb; x++) { unsigned char* q = c.px + (long long)y * c.row_stride + 3ll * x; q[0] = c.r; q[1] = c.g; q[2] = c.b; }
}

// clipLine(Size2l, Point2l&, Point2l&): integer endpoints, the cuts in double, truncated
__device__ bool clip_line(long long width, long long height, long long& x1, long long& y1, long long& x2, long long& y2) {
    const long long right = width - 1, bottom = height - 1;
    if (width <= 0 || height <= 0) return false;
    int c1 = (x1 < 0) + (x1 > right) * 2 + (y1 < 0) * 4 + (y1 > bottom) * 8;
    int c2 = (x2 < 0) + (x2 > right) * 2 + (y2 < 0) * 4 + (y2 > bottom) * 8;
    if ((c1 & c2) == 0 && (c1 | c2) != 0) {
        long long a;
        if (c1 & 12) {
            a = c1 < 8 ? 0 : bottom;
            x1 += (long long)((double)(a - y1) * (double)(x2 - x1) / (double)(y2 - y1));
            y1 = a;
            c1 = (x1 < 0) + (x1 > right) * 2;
        }
        if (c2 & 12) {
            a = c2 < 8 ? 0 : bottom;
            x2 += (long long)((double)(a - y2) * (double)(x2 - x1) / (double)(y2 - y1));
            y2 = a;
            c2 = (x2 < 0) + (x2 > right) * 2;
        }
        if ((c1 & c2) == 0 && (c1 | c2) != 0) {
            if (c1) {
                a = c1 == 1 ? 0 : right;
                y1 += (long long)((double)(a - x1) * (double)(y2 - y1) / (double)(x2 - x1));
                x1 = a;
                c1 = 0;
            }
            if (c2) {
                a = c2 == 1 ? 0 : right;
                y2 += (long long)((double)(a - x2) * (double)(y2 - y1) / (double)(x2 - x1));
                x2 = a;
                c2 = 0;
            }
        }
    }
    return (c1 | c2) == 0;
}

// Line2: endpoints in 16.16 fixed point (the outline FillConvexPoly draws around its spans)
__device__ void line2(const Canvas& c, long long x1, long long y1, long long x2, long long y2) {
    if (!clip_line((long long)c.w << XY_SHIFT, (long long)c.h << XY_SHIFT, x1, y1, x2, y2)) return;
    long long dx = x2 - x1, dy = y2 - y1;
    const long long ax = dx < 0 ? -dx : dx, ay = dy < 0 ? -dy : dy;
    if (ax > ay) {
        if (dx < 0) { dy = -dy; long long t = x1; x1 = x2; x2 = t; t = y1; y1 = y2; y2 = t; }
        const long long y_step = (dy << XY_SHIFT) / (ax | 1);       // C division: towards zero
        int ecount = (int)((x2 - x1) >> XY_SHIFT);
        x1 += XY_HALF; y1 += XY_HALF;
        put_px(c, (int)((x2 + XY_HALF) >> XY_SHIFT), (int)((y2 + XY_HALF) >> XY_SHIFT));
        int x = (int)(x1 >> XY_SHIFT);
        while (ecount >= 0) { put_px(c, x, (int)(y1 >> XY_SHIFT)); x++; y1 += y_step; ecount--; }
    } else {
        if (dy < 0) { dx = -dx; long long t = x1; x1 = x2; x2 = t; t = y1; y1 = y2; y2 = t; }
        const long long x_step = (dx << XY_SHIFT) / (ay | 1);
        int ecount = (int)((y2 - y1) >> XY_SHIFT);
        x1 += XY_HALF; y1 += XY_HALF;
        put_px(c, (int)((x2 + XY_HALF) >> XY_SHIFT), (int)((y2 + XY_HALF) >> XY_SHIFT));
        int y = (int)(y1 >> XY_SHIFT);
        while (ecount >= 0) { put_px(c, (int)(x1 >> XY_SHIFT), y); x1 += x_step; y++; ecount--; }
    }
}

// Line(img, pt1, pt2, color, 8): integer endpoints, clipped to the frame, LineIterator's 8-connected Bresenham from the left end
__device__ void line_int(const Canvas& c, long long x1, long long y1, long long x2, long long y2) {
    if (!clip_line(c.w, c.h, x1, y1, x2, y2)) return;
    int dx = (int)(x2 - x1), dy = (int)(y2 - y1);
    if (dx < 0) { dx = -dx; dy = -dy; x1 = x2; y1 = y2; }
    const int sy = dy < 0 ? -1 : 1;
    dy = dy < 0 ? -dy : dy;
    const bool steep = dy > dx;
    if (steep) { const int t = dx; dx = dy; dy = t; }
    int err = dx - (dy + dy);
    const int plus = dx + dx, minus = -(dy + dy);
    int x = (int)x1, y = (int)y1;
    for (int k = 0; k <= dx; k++) {
        put_px(c, x, y);
        const bool m = err < 0;
        err += minus + (m ? plus : 0);
        if (steep) { y += sy; if (m) x++; }
        else { x++; if (m) y += sy; }
    }
}

// FillConvexPoly(img, v, 4, color, LINE_8, XY_SHIFT)
__device__ void fill_quad(const Canvas& c, const long long (&vx)[4], const long long (&vy)[4]) {
    const int n = 4;
    long long xmin = vx[0], xmax = vx[0], ymin = vy[0], ymax = vy[0];
    int imin = 0;
    {
        long long px = vx[n - 1], py = vy[n - 1];
        for (int i = 0; i < n; i++) {
            if (vy[i] < ymin) { ymin = vy[i]; imin = i; }
            ymax = max(ymax, vy[i]); xmax = max(xmax, vx[i]); xmin = min(xmin, vx[i]);
            line2(c, px, py, vx[i], vy[i]);
            px = vx[i]; py = vy[i];
        }
    }
    xmin = (xmin + XY_HALF) >> XY_SHIFT; xmax = (xmax + XY_HALF) >> XY_SHIFT;
    ymin = (ymin + XY_HALF) >> XY_SHIFT; ymax = (ymax + XY_HALF) >> XY_SHIFT;
    if ((int)xmax < 0 || (int)ymax < 0 || (int)xmin >= c.w || (int)ymin >= c.h) return;
    ymax = min(ymax, (long long)c.h - 1);
    int edges = n;
    int e_idx[2] = {imin, imin}, e_di[2] = {1, n - 1}, e_ye[2] = {(int)ymin, (int)ymin};
    long long e_x[2] = {-XY_ONE, -XY_ONE}, e_dx[2] = {0, 0};
    int y = (int)ymin;
    do {
        for (int i = 0; i < 2; i++) {
            if (y >= e_ye[i]) {
                int idx0 = e_idx[i];
                const int di = e_di[i];
                int idx = idx0 + di;
                if (idx >= n) idx -= n;
                for (; edges-- > 0;) {
                    const int ty = (int)((vy[idx] + XY_HALF) >> XY_SHIFT);
                    if (ty > y) {
                        const long long xs = vx[idx0], xe = vx[idx];
                        e_ye[i] = ty;
                        e_dx[i] = ((xe - xs) * 2 + (ty - y)) / (2 * (ty - y));
                        e_x[i] = xs;
                        e_idx[i] = idx;
                        break;
                    }
                    idx0 = idx;
                    idx += di;
                    if (idx >= n) idx -= n;
                }
            }
        }
        if (edges < 0) break;
        if (y >= 0) {
            const int left = e_x[0] > e_x[1] ? 1 : 0, right = 1 - left;
            const int xx1 = (int)((e_x[left] + XY_HALF) >> XY_SHIFT), xx2 = (int)((e_x[right] + XY_HALF) >> XY_SHIFT);
            if (xx2 >= 0 && xx1 < c.w) hline(c, y, xx1, xx2);
        }
        e_x[0] += e_dx[0];
        e_x[1] += e_dx[1];
    } while (++y <= (int)ymax);
}

// Circle(img, center, radius, color, fill = 1)
__device__ void disc(const Canvas& c, int cx, int cy, int radius) {
    int err = 0, dx = radius, dy = 0, plus = 1, minus = (radius << 1) - 1;
    while (dx >= dy) {
        hline(c, cy - dy, cx - dx, cx + dx);
        hline(c, cy + dy, cx - dx, cx + dx);
        hline(c, cy - dx, cx - dy, cx + dy);
        hline(c, cy + dx, cx - dy, cx + dy);
        dy++;
        err += plus;
        plus += 2;
        const int mask = (err <= 0) - 1;
        err -= minus & mask;
        dx += mask;
        minus -= mask & 2;
    }
}

// ThickLine(img, p0, p1, color, thickness, LINE_8, flags, XY_SHIFT): flags bit 0 / 1 = a disc at p0 / p1
__device__ void thick_line(const Canvas& c, long long x0, long long y0, long long x1, long long y1, int thickness, int flags) {
    if (thickness <= 1) {
        line_int(c, (x0 + XY_HALF) >> XY_SHIFT, (y0 + XY_HALF) >> XY_SHIFT, (x1 + XY_HALF) >> XY_SHIFT, (y1 + XY_HALF) >> XY_SHIFT);
        return;
    }
    const double inv = 1.0 / (double)XY_ONE;
    const double dx = (double)(x0 - x1) * inv, dy = (double)(y1 - y0) * inv;
    double r = dx * dx + dy * dy;
    const int odd = thickness & 1;
    const long long t = (long long)thickness << (XY_SHIFT - 1);
    if (fabs(r) > 2.220446049250313e-16) {
        r = ((double)t + (double)odd * (double)XY_ONE * 0.5) / sqrt(r);
        const long long dpx = (long long)rint(dy * r), dpy = (long long)rint(dx * r);      // cvRound
        const long long vx[4] = {x0 + dpx, x0 - dpx, x1 - dpx, x1 + dpx}, vy[4] = {y0 + dpy, y0 - dpy, y1 - dpy, y1 + dpy};
        fill_quad(c, vx, vy);
    }
    const int radius = (int)((t + XY_HALF) >> XY_SHIFT);
    if (flags & 1) disc(c, (int)((x0 + XY_HALF) >> XY_SHIFT), (int)((y0 + XY_HALF) >> XY_SHIFT), radius);
    if (flags & 2) disc(c, (int)((x1 + XY_HALF) >> XY_SHIFT), (int)((y1 + XY_HALF) >> XY_SHIFT), radius);
}

// segments of glyph g (ASCII - 32): number, and the k-th one (endpoints in Hershey units, is it the first of its polyline)
__device__ __forceinline__ int glyph_segments(int g) {
    int n = 0;
    for (int k = c_hf_off[g] + 2; k < c_hf_off[g + 1];) { const int m = c_hf_data[k]; n += m > 1 ? m - 1 : 0; k += 1 + 2 * m; }
    return n;
}
__device__ __forceinline__ void glyph_segment(int g, int j, int& xa, int& ya, int& xb, int& yb, bool& first) {
    for (int k = c_hf_off[g] + 2; k < c_hf_off[g + 1];) {
        const int m = c_hf_data[k], ns = m > 1 ? m - 1 : 0;
        if (j < ns) {
            const signed char* q = c_hf_data + k + 1 + 2 * j;
            xa = q[0]; ya = q[1]; xb = q[2]; yb = q[3]; first = j == 0;
            return;
        }
        j -= ns;
        k += 1 + 2 * m;
    }
}

// One detection: the rectangle first, then the label; all its pixels have one colour, so the threads that draw it need
// no order among themselves. pen / seg0: MAXC(+1) entries of shared memory; ends with a __syncthreads().
__device__ void draw_one(const DrawParams& p, Canvas& c, const DrawItem& it, long long* pen, int* seg0) {
    const int tid = threadIdx.x;
    c.r = it.rgb[0]; c.g = it.rgb[1]; c.b = it.rgb[2];
    // ---- cv2.rectangle: the outline (thickness 1), grown by its 4-neighbourhood (thickness 2) ----
    {
        const int xa = it.xa, xb = it.xb, ya = it.ya, yb = it.yb, grow = p.thickness == 2 ? 1 : 0;
        // outline pixels inside the frame grown by one: top / bottom rows, then the left / right columns between them
        const int cx0 = max(xa, -1), cx1 = min(xb, p.w), cy0 = max(ya + 1, -1), cy1 = min(yb - 1, p.h);
        const int nrow = max(cx1 - cx0 + 1, 0), ncol = max(cy1 - cy0 + 1, 0);
        const int total = 2 * nrow + 2 * ncol;
        for (int k = tid; k < total; k += DRAW_T) {
            int x, y;
            if (k < 2 * nrow) { x = cx0 + (k >> 1); y = (k & 1) ? yb : ya; }
            else { const int q = k - 2 * nrow; y = cy0 + (q >> 1); x = (q & 1) ? xb : xa; }
            put_px(c, x, y);
            if (grow) { put_px(c, x - 1, y); put_px(c, x + 1, y); put_px(c, x, y - 1); put_px(c, x, y + 1); }
        }
    }
    __syncthreads();
    // ---- cv2.putText ----
    const int nchar = it.nchar;
    if (nchar > 0) {
        if (tid == 0) {                                    // putText's pen: view_x -= left * hscale; draw; view_x += right * hscale
            long long view_x = (long long)it.ox << XY_SHIFT;
            int s0 = 0;
            for (int k = 0; k < nchar; k++) {
                const int g = it.text[k];
                view_x -= (long long)c_hf_data[c_hf_off[g]] * p.hscale;
                pen[k] = view_x;
                seg0[k] = s0;
                s0 += glyph_segments(g);
                view_x += (long long)c_hf_data[c_hf_off[g] + 1] * p.hscale;
            }
            seg0[nchar] = s0;
        }
        __syncthreads();
        const long long view_y = ((long long)it.oy << XY_SHIFT) + (long long)HF_BASE_LINE * p.hscale;
        const int total = seg0[nchar];
        for (int sidx = tid; sidx < total; sidx += DRAW_T) {
            int lo = 0, hi = nchar - 1;                    // the character that owns segment sidx
            while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (seg0[mid] <= sidx) lo = mid; else hi = mid - 1; }
            int xa, ya, xb, yb;
            bool first;
            glyph_segment(it.text[lo], sidx - seg0[lo], xa, ya, xb, yb, first);
            thick_line(c, (long long)xa * p.hscale + pen[lo], (long long)ya * p.hscale + view_y, (long long)xb * p.hscale + pen[lo],
                       (long long)yb * p.hscale + view_y, p.text_thickness, first ? 3 : 2);
        }
    }
    __syncthreads();
}

// ORDERED: one CTA per frame draws its detections one after the other (later ones paint over earlier ones, as the
// reference's loop does) — needed when detections have different colours. Otherwise every pixel any detection touches
// ends up with the one colour whatever the order, and each detection gets a CTA of its own (grid max_det x frames).
template <bool ORDERED>
__global__ void __launch_bounds__(DRAW_T) draw_frames_kernel(const DrawParams p, const int* __restrict__ count, const DrawItem* __restrict__ items,
                                                             unsigned char* __restrict__ frames) {
    pdl_wait();
    __shared__ long long pen[MAXC];        // view_x of every character of the current label (16.16)
    __shared__ int seg0[MAXC + 1];         // first segment of every character
    const int f = ORDERED ? blockIdx.x : blockIdx.y;
    const int n = min(count ? count[f] : p.max_det, p.max_det);
    Canvas c;
    c.px = frames + (long long)f * p.frame_stride; c.row_stride = p.row_stride; c.w = p.w; c.h = p.h;
    if (ORDERED) {
        for (int i = 0; i < n; i++) draw_one(p, c, items[(long long)f * p.max_det + i], pen, seg0);
    } else if ((int)blockIdx.x < n) {
        draw_one(p, c, items[(long long)f * p.max_det + blockIdx.x], pen, seg0);
    }
}

}  // namespace

// =========================================================================================================
// C ABI
// =========================================================================================================
extern "C" int tcb_scale_boxes(tcb_ctx* ctx, const double* det_dev, int n, int orig_h, int orig_w, int input_size0, int input_size1,
                               int normalized, int math_mode, double* out_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, n >= 0 && orig_h > 0 && orig_w > 0 && input_size0 > 0 && input_size1 > 0, "tcb_scale_boxes: bad sizes");
    TCB_REQUIRE(ctx, math_mode >= 0 && math_mode <= 2, "tcb_scale_boxes: math_mode must be 0, 1 or 2");
    if (n == 0) return TCB_OK;
    TCB_REQUIRE(ctx, det_dev && out_dev, "tcb_scale_boxes: null buffer");
    TCB_KLAUNCH(ctx, scale_boxes_kernel, tcb_div_up(n, 256), 256, 0, (cudaStream_t)stream, det_dev, n, orig_h, orig_w, input_size0, input_size1,
                                                                             normalized ? 1 : 0, math_mode, out_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

extern "C" size_t tcb_draw_detections_workspace(int b, int max_det) {
    return (size_t)(b > 0 ? b : 0) * (size_t)(max_det > 0 ? max_det : 0) * sizeof(DrawItem) + 512;
}

extern "C" int tcb_draw_detections(tcb_ctx* ctx, unsigned char* frames_dev, int b, int h, int w, long long row_stride, long long frame_stride,
                                   const double* det_dev, const int* count_dev, int max_det, int input_size0, int input_size1,
                                   int normalized, int math_mode, const char* class_names, int num_classes,
                                   const unsigned char* colors, int per_class_color, int thickness, int labels, double font_scale,
                                   int text_thickness, int text_dy, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, b >= 0 && h > 0 && w > 0 && max_det >= 0, "tcb_draw_detections: bad sizes");
    TCB_REQUIRE(ctx, row_stride >= 3ll * w && (b <= 1 || frame_stride >= row_stride * h), "tcb_draw_detections: strides smaller than the frame");
    TCB_REQUIRE(ctx, input_size0 > 0 && input_size1 > 0, "tcb_draw_detections: bad model input size");
    TCB_REQUIRE(ctx, math_mode >= 0 && math_mode <= 2, "tcb_draw_detections: math_mode must be 0, 1 or 2");
    TCB_REQUIRE(ctx, thickness == 1 || thickness == 2, "tcb_draw_detections: thickness must be 1 or 2");
    TCB_REQUIRE(ctx, labels >= 0 && labels <= 2, "tcb_draw_detections: labels must be 0 (none), 1 ('conf name') or 2 (name)");
    TCB_REQUIRE(ctx, !labels || (font_scale > 0.0 && font_scale <= 64.0 && text_thickness >= 1 && text_thickness <= 32), "tcb_draw_detections: font scale in (0, 64], text thickness in [1, 32]");
    TCB_REQUIRE(ctx, colors != nullptr, "tcb_draw_detections: colors is null");
    TCB_REQUIRE(ctx, num_classes >= 0 && (num_classes == 0 || class_names || (!labels && !per_class_color)), "tcb_draw_detections: class names missing");
    if (num_classes > MAX_CLASSES) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_draw_detections: %d classes > compiled limit %d", num_classes, MAX_CLASSES);
    if (b == 0 || max_det == 0) return TCB_OK;
    TCB_REQUIRE(ctx, frames_dev && det_dev, "tcb_draw_detections: null buffer");

    DrawParams p;    // 3.3 KB, passed to both kernels by value
    memset(&p, 0, sizeof(p));
    p.h = h; p.w = w; p.max_det = max_det; p.row_stride = row_stride; p.frame_stride = frame_stride;
    p.orig_h = h; p.orig_w = w; p.in0 = input_size0; p.in1 = input_size1;
    p.normalized = normalized ? 1 : 0; p.math_mode = math_mode; p.thickness = thickness; p.labels = labels;
    p.text_thickness = text_thickness; p.text_dy = text_dy; p.hscale = (int)nearbyint(font_scale * 65536.0);   // cvRound(fontScale * XY_ONE)
    p.per_class_color = per_class_color ? 1 : 0; p.num_classes = num_classes;
    if (per_class_color) memcpy(p.class_color, colors, (size_t)num_classes * 3);
    else memcpy(p.color, colors, 3);
    if (class_names) {
        const char* s = class_names;   // num_classes strings, each ending in '\0'
        int used = 0;
        for (int c = 0; c < num_classes; c++) {
            const size_t len = strlen(s);
            if (used + (int)len > NAME_BYTES) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_draw_detections: class names exceed %d bytes", NAME_BYTES);
            p.name_off[c] = (unsigned short)used;
            memcpy(p.names + used, s, len);
            used += (int)len;
            s += len + 1;
        }
        p.name_off[num_classes] = (unsigned short)used;
    }
    TCB_WS(ctx, items, DrawItem, (size_t)b * max_det);
    cudaStream_t st = (cudaStream_t)stream;
    TCB_KLAUNCH(ctx, draw_items_kernel, dim3(tcb_div_up(max_det, 128), b), 128, 0, st, p, det_dev, count_dev, items);
    TCB_LAUNCHED(ctx);
    if (per_class_color) { TCB_KLAUNCH(ctx, draw_frames_kernel<true>, b, DRAW_T, 0, st, p, count_dev, items, frames_dev); }
    else { TCB_KLAUNCH(ctx, draw_frames_kernel<false>, dim3(max_det, b), DRAW_T, 0, st, p, count_dev, items, frames_dev); }
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}
