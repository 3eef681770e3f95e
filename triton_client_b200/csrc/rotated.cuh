// Rotated BEV rectangle overlap / IoU in float32, device side.
// Same algorithm and the same float32 operation order as the reference's only native code on the path,
// OPCD/pcdet/ops/iou3d_nms/src/iou3d_nms_kernel.cu:104-234 (CPU twin iou3d_cpu.cpp:58-234): rotate the four
// corners of both boxes, collect the up-to-16 edge/edge crossings and the corners of one box lying inside the
// other (1 cm margin), order the points by angle around their centroid, shoelace area, IoU = ov/max(sa+sb-ov,1e-8).
// Written from the algorithm, not copied: points live in plain float arrays, headings enter as precomputed
// (cos, sin) pairs so that one sincos per box serves every pair it takes part in.
#pragma once
#include <cuda_runtime.h>

namespace rot {

struct RBox {   // what the BEV test needs of a [x, y, z, dx, dy, dz, heading] box
    float x, y, dx, dy, c, s, r;  // centre, size, cos/sin(heading), half diagonal (pre-test radius)
};

__device__ __forceinline__ RBox make_rbox(const float* b7) {
    RBox r;
    r.x = b7[0]; r.y = b7[1]; r.dx = b7[3]; r.dy = b7[4];
    r.c = cosf(b7[6]); r.s = sinf(b7[6]);
    r.r = 0.5f * sqrtf(r.dx * r.dx + r.dy * r.dy);
    return r;
}

// Conservative: true only when the rectangles (inflated by the reference's 1 cm corner margin) cannot touch,
// in which case the reference's polygon is empty and its IoU is exactly 0.
__device__ __forceinline__ bool far_apart(const RBox& a, const RBox& b) {
    const float ddx = a.x - b.x, ddy = a.y - b.y, lim = a.r + b.r + 0.05f;
    return ddx * ddx + ddy * ddy > lim * lim;
}

__device__ __forceinline__ float cross3(float p1x, float p1y, float p2x, float p2y, float p0x, float p0y) {
    return (p1x - p0x) * (p2y - p0y) - (p2x - p0x) * (p1y - p0y);
}

__device__ __forceinline__ void corners(const RBox& b, float* cx, float* cy) {  // 5 entries, [4] = [0]
    const float hx = b.dx / 2, hy = b.dy / 2;
    const float x1 = b.x - hx, y1 = b.y - hy, x2 = b.x + hx, y2 = b.y + hy;
    const float rx[4] = {x1, x2, x2, x1}, ry[4] = {y1, y1, y2, y2};
#pragma unroll
    for (int k = 0; k < 4; k++) {
        cx[k] = (rx[k] - b.x) * b.c + (ry[k] - b.y) * (-b.s) + b.x;
        cy[k] = (rx[k] - b.x) * b.s + (ry[k] - b.y) * b.c + b.y;
    }
    cx[4] = cx[0]; cy[4] = cy[0];
}

__device__ __forceinline__ bool inside(const RBox& b, float px, float py) {
    // rotate the point by -heading: cos(-h) = c, sin(-h) = -s
    const float ac = b.c, as = -b.s;
    const float rx = (px - b.x) * ac + (py - b.y) * (-as);
    const float ry = (px - b.x) * as + (py - b.y) * ac;
    return fabsf(rx) < b.dx / 2 + 1e-2f && fabsf(ry) < b.dy / 2 + 1e-2f;
}

// segment p0->p1 against q0->q1
__device__ __forceinline__ bool seg_x(float p1x, float p1y, float p0x, float p0y, float q1x, float q1y, float q0x, float q0y,
                                      float& ox, float& oy) {
    const bool touch = fminf(p0x, p1x) <= fmaxf(q0x, q1x) && fminf(q0x, q1x) <= fmaxf(p0x, p1x) &&
                       fminf(p0y, p1y) <= fmaxf(q0y, q1y) && fminf(q0y, q1y) <= fmaxf(p0y, p1y);
    if (!touch) return false;
    const float s1 = cross3(q0x, q0y, p1x, p1y, p0x, p0y), s2 = cross3(p1x, p1y, q1x, q1y, p0x, p0y);
    const float s3 = cross3(p0x, p0y, q1x, q1y, q0x, q0y), s4 = cross3(q1x, q1y, p1x, p1y, q0x, q0y);
    if (!(s1 * s2 > 0 && s3 * s4 > 0)) return false;
    const float s5 = cross3(q1x, q1y, p1x, p1y, p0x, p0y);
    if (fabsf(s5 - s1) > 1e-8f) {
        ox = (s5 * q0x - s1 * q1x) / (s5 - s1);
        oy = (s5 * q0y - s1 * q1y) / (s5 - s1);
    } else {
        const float a0 = p0y - p1y, b0 = p1x - p0x, c0 = p0x * p1y - p1x * p0y;
        const float a1 = q0y - q1y, b1 = q1x - q0x, c1 = q0x * q1y - q1x * q0y;
        const float D = a0 * b1 - a1 * b0;
        ox = (b0 * c1 - b1 * c0) / D;
        oy = (a1 * c0 - a0 * c1) / D;
    }
    return true;
}

__device__ __noinline__ float overlap(const RBox a, const RBox b) {
    float ax[5], ay[5], bx[5], by[5];
    corners(a, ax, ay);
    corners(b, bx, by);
    float px[16], py[16];
    float sx = 0.f, sy = 0.f;
    int cnt = 0;
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float ox, oy;
            if (seg_x(ax[i + 1], ay[i + 1], ax[i], ay[i], bx[j + 1], by[j + 1], bx[j], by[j], ox, oy)) {
                sx = sx + ox; sy = sy + oy;
                px[cnt] = ox; py[cnt] = oy; cnt++;
            }
        }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (inside(a, bx[k], by[k])) { sx = sx + bx[k]; sy = sy + by[k]; px[cnt] = bx[k]; py[cnt] = by[k]; cnt++; }
        if (inside(b, ax[k], ay[k])) { sx = sx + ax[k]; sy = sy + ay[k]; px[cnt] = ax[k]; py[cnt] = ay[k]; cnt++; }
    }
    if (cnt == 0) return 0.f;  // (the reference divides 0/0 here and then sums an empty polygon: also 0)
    sx /= (float)cnt; sy /= (float)cnt;
    float ang[16];
    for (int k = 0; k < cnt; k++) ang[k] = atan2f(py[k] - sy, px[k] - sx);
    for (int j = 0; j < cnt - 1; j++)       // the reference's bubble sort, same swap sequence
        for (int i = 0; i < cnt - j - 1; i++)
            if (ang[i] > ang[i + 1]) {
                float t = ang[i]; ang[i] = ang[i + 1]; ang[i + 1] = t;
                t = px[i]; px[i] = px[i + 1]; px[i + 1] = t;
                t = py[i]; py[i] = py[i + 1]; py[i + 1] = t;
            }
    float area = 0.f;
    for (int k = 0; k < cnt - 1; k++)
        area += (px[k] - px[0]) * (py[k + 1] - py[0]) - (py[k] - py[0]) * (px[k + 1] - px[0]);
    return fabsf(area) / 2.0f;
}

__device__ __forceinline__ float iou_bev(const RBox& a, const RBox& b) {
    const float sa = a.dx * a.dy, sb = b.dx * b.dy;
    const float ov = overlap(a, b);
    return ov / fmaxf(sa + sb - ov, 1e-8f);
}

// iou_normal (iou3d_nms_kernel.cu:314-325), the predicate of nms_normal_gpu: axis-aligned BEV overlap of (x, y, dx, dy),
// heading ignored. a is the box of the mask row (the earlier, higher-scoring one), b the later one. The reference has no
// CPU twin of this function; its CUDA build (nvcc 12.9, sm_100a SASS read back) contracts `Sa + Sb` into one
// fma(b.dx, b.dy, Sa) with Sa = a.dx * a.dy rounded first, and nothing else — restated with explicit intrinsics.
__device__ __forceinline__ float iou_normal(const RBox& a, const RBox& b) {
    const float left = fmaxf(a.x - a.dx / 2, b.x - b.dx / 2), right = fminf(a.x + a.dx / 2, b.x + b.dx / 2);
    const float top = fmaxf(a.y - a.dy / 2, b.y - b.dy / 2), bottom = fminf(a.y + a.dy / 2, b.y + b.dy / 2);
    const float width = fmaxf(right - left, 0.f), height = fmaxf(bottom - top, 0.f);
    const float inter = width * height;
    const float sa = a.dx * a.dy;
    return inter / fmaxf(__fmaf_rn(b.dx, b.dy, sa) - inter, 1e-8f);
}

// Second conservative pre-test, for the pairs the circle test lets through: a separating axis between the two rectangles
// (the four edge directions) with a 3 cm margin. The reference counts a corner as inside a box up to 1 cm outside its edge
// and finds edge crossings only where edges cross, so with a gap above that margin its point list is empty and its IoU
// exactly 0; for nms_normal_gpu the boxes are axis-aligned and a gap along x or y makes the intersection exactly 0.
// Any NaN makes every comparison false: such a pair is left to the exact test.
__device__ __forceinline__ bool separated(const RBox& a, const RBox& b, int normal) {
    const float m = 0.03f;
    const float ox = b.x - a.x, oy = b.y - a.y;
    if (normal) return fabsf(ox) > (a.dx + b.dx) * 0.5f + m || fabsf(oy) > (a.dy + b.dy) * 0.5f + m;
    const float ahx = fabsf(a.dx) * 0.5f, ahy = fabsf(a.dy) * 0.5f, bhx = fabsf(b.dx) * 0.5f, bhy = fabsf(b.dy) * 0.5f;
    const float cd = fabsf(a.c * b.c + a.s * b.s), sd = fabsf(b.s * a.c - b.c * a.s);   // |cos|, |sin| of the heading difference
    float rx = ox * a.c + oy * a.s, ry = oy * a.c - ox * a.s;                             // b's centre in a's frame
    if (fabsf(rx) > ahx + bhx * cd + bhy * sd + m || fabsf(ry) > ahy + bhx * sd + bhy * cd + m) return true;
    rx = ox * b.c + oy * b.s; ry = oy * b.c - ox * b.s;                                   // a's centre in b's frame (sign dropped)
    return fabsf(rx) > bhx + ahx * cd + ahy * sd + m || fabsf(ry) > bhy + ahx * sd + ahy * cd + m;
}

// the suppression test of nms_gpu (normal == 0) / nms_normal_gpu (normal != 0)
__device__ __forceinline__ bool overlaps(const RBox& a, const RBox& b, float thresh, int normal) {
    return (normal ? iou_normal(a, b) : iou_bev(a, b)) > thresh;
}

}  // namespace rot
