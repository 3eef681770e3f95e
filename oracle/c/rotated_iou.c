/* Oracle (test infrastructure only): rotated BEV overlap / IoU and the greedy rotated NMS sweep.
 * Restates, operation for operation in float32, the arithmetic of the reference's
 *   OPCD/pcdet/ops/iou3d_nms/src/iou3d_cpu.cpp:58-234  (cross, check_rect_cross, check_in_box2d, intersection,
 *                                                       rotate_around_center, point_cmp, box_overlap, iou_bev)
 *   OPCD/pcdet/ops/iou3d_nms/src/iou3d_nms_kernel.cu:104-234 (identical device twin)
 * and the host sweep of OPCD/pcdet/ops/iou3d_nms/src/iou3d_nms.cpp:116-132 (keep box i unless an earlier kept
 * box has iou_bev(kept, i) > thresh). Pinned against the reference's own compiled iou3d_cpu.cpp
 * (oracle/_ref/libiou3d_ref.so) in tests/test_oracle_3d.py.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float x, y; } pt2;

static const float EPS_F = 1e-8f;

static float cross_o(pt2 p1, pt2 p2, pt2 p0) { return (p1.x - p0.x) * (p2.y - p0.y) - (p2.x - p0.x) * (p1.y - p0.y); }
static float cross_v(pt2 a, pt2 b) { return a.x * b.y - a.y * b.x; }
static float fmin2(float a, float b) { return a > b ? b : a; }
static float fmax2(float a, float b) { return a > b ? a : b; }

static int rects_touch(pt2 p1, pt2 p2, pt2 q1, pt2 q2) {
    return fmin2(p1.x, p2.x) <= fmax2(q1.x, q2.x) && fmin2(q1.x, q2.x) <= fmax2(p1.x, p2.x) &&
           fmin2(p1.y, p2.y) <= fmax2(q1.y, q2.y) && fmin2(q1.y, q2.y) <= fmax2(p1.y, p2.y);
}

/* point inside the (margin-inflated) rotated box; c,s = cos,sin of the box heading */
static int inside_box(const float* box, float c, float s, pt2 p) {
    const float margin = 1e-2f;
    float ac = c, as = -s; /* cos(-h), sin(-h) */
    float rx = (p.x - box[0]) * ac + (p.y - box[1]) * (-as);
    float ry = (p.x - box[0]) * as + (p.y - box[1]) * ac;
    return fabsf(rx) < box[3] / 2 + margin && fabsf(ry) < box[4] / 2 + margin;
}

static int seg_intersection(pt2 p1, pt2 p0, pt2 q1, pt2 q0, pt2* ans) {
    if (!rects_touch(p0, p1, q0, q1)) return 0;
    float s1 = cross_o(q0, p1, p0), s2 = cross_o(p1, q1, p0), s3 = cross_o(p0, q1, q0), s4 = cross_o(q1, p1, q0);
    if (!(s1 * s2 > 0 && s3 * s4 > 0)) return 0;
    float s5 = cross_o(q1, p1, p0);
    if (fabsf(s5 - s1) > EPS_F) {
        ans->x = (s5 * q0.x - s1 * q1.x) / (s5 - s1);
        ans->y = (s5 * q0.y - s1 * q1.y) / (s5 - s1);
    } else {
        float a0 = p0.y - p1.y, b0 = p1.x - p0.x, c0 = p0.x * p1.y - p1.x * p0.y;
        float a1 = q0.y - q1.y, b1 = q1.x - q0.x, c1 = q0.x * q1.y - q1.x * q0.y;
        float D = a0 * b1 - a1 * b0;
        ans->x = (b0 * c1 - b1 * c0) / D;
        ans->y = (a1 * c0 - a0 * c1) / D;
    }
    return 1;
}

static void corners_of(const float* b, float c, float s, pt2* out /*5*/) {
    float hx = b[3] / 2, hy = b[4] / 2;
    float x1 = b[0] - hx, y1 = b[1] - hy, x2 = b[0] + hx, y2 = b[1] + hy;
    pt2 raw[4] = {{x1, y1}, {x2, y1}, {x2, y2}, {x1, y2}};
    for (int k = 0; k < 4; k++) {
        float nx = (raw[k].x - b[0]) * c + (raw[k].y - b[1]) * (-s) + b[0];
        float ny = (raw[k].x - b[0]) * s + (raw[k].y - b[1]) * c + b[1];
        out[k].x = nx; out[k].y = ny;
    }
    out[4] = out[0];
}

float oracle_box_overlap(const float* a, const float* b) {
    float ca = cosf(a[6]), sa = sinf(a[6]), cb = cosf(b[6]), sb = sinf(b[6]);
    pt2 A[5], B[5], poly[16], ctr = {0.f, 0.f};
    corners_of(a, ca, sa, A);
    corners_of(b, cb, sb, B);
    int cnt = 0;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++)
            if (seg_intersection(A[i + 1], A[i], B[j + 1], B[j], &poly[cnt])) {
                ctr.x = ctr.x + poly[cnt].x; ctr.y = ctr.y + poly[cnt].y;
                cnt++;
            }
    for (int k = 0; k < 4; k++) {
        if (inside_box(a, ca, sa, B[k])) { ctr.x = ctr.x + B[k].x; ctr.y = ctr.y + B[k].y; poly[cnt++] = B[k]; }
        if (inside_box(b, cb, sb, A[k])) { ctr.x = ctr.x + A[k].x; ctr.y = ctr.y + A[k].y; poly[cnt++] = A[k]; }
    }
    ctr.x /= cnt; ctr.y /= cnt;
    for (int j = 0; j < cnt - 1; j++)
        for (int i = 0; i < cnt - j - 1; i++)
            if (atan2f(poly[i].y - ctr.y, poly[i].x - ctr.x) > atan2f(poly[i + 1].y - ctr.y, poly[i + 1].x - ctr.x)) {
                pt2 t = poly[i]; poly[i] = poly[i + 1]; poly[i + 1] = t;
            }
    float area = 0;
    for (int k = 0; k < cnt - 1; k++) {
        pt2 u = {poly[k].x - poly[0].x, poly[k].y - poly[0].y}, v = {poly[k + 1].x - poly[0].x, poly[k + 1].y - poly[0].y};
        area += cross_v(u, v);
    }
    return (float)(fabsf(area) / 2.0);
}

float oracle_iou_bev(const float* a, const float* b) {
    float sa = a[3] * a[4], sb = b[3] * b[4];
    float ov = oracle_box_overlap(a, b);
    return ov / fmaxf(sa + sb - ov, EPS_F);
}

void oracle_boxes_iou_bev(const float* a, int n, const float* b, int m, float* out) {
    for (int i = 0; i < n; i++)
        for (int j = 0; j < m; j++) out[(size_t)i * m + j] = oracle_iou_bev(a + 7 * i, b + 7 * j);
}

/* Greedy NMS over boxes ALREADY in descending-score order (iou3d_nms.cpp:116-132 sweep over the mask that
 * nms_kernel fills with iou_bev(box_i, box_j) > thresh for j > i). Returns the number kept; keep[] holds
 * positions in the sorted order. */
int oracle_rotated_nms_sorted(const float* boxes, int n, float thresh, int* keep) {
    char* removed = (char*)calloc((size_t)n + 1, 1);
    int nk = 0;
    for (int i = 0; i < n; i++) {
        if (removed[i]) continue;
        keep[nk++] = i;
        for (int j = i + 1; j < n; j++)
            if (!removed[j] && oracle_iou_bev(boxes + 7 * i, boxes + 7 * j) > thresh) removed[j] = 1;
    }
    free(removed);
    return nk;
}
