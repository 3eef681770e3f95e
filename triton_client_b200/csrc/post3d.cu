// 3D detection postprocess on device: rotated BEV IoU / NMS, PointPillars anchor
// decode + class-agnostic NMS, CenterPoint heatmap top-K decode + per-head NMS, nuScenes per-class score filter.
//
// In the reference these run inside the Triton Python backend (examples/pointpillar_kitti/1/model.py:163) through
// the vendored pcdet code cited at each kernel; the drop-in runs them client-side behind PointPillarPostprocess.
// Up to 512 candidates one 1024-thread CTA per frame (or per frame x head) does ordering + greedy NMS entirely in
// shared memory; above that the pairwise IoU mask is built by the whole grid and swept by one CTA ("wide path"
// below). Either way nothing leaves the device: no D2H of the mask, no serial host sweep (iou3d_nms.cpp:105-134).
#include <cstddef>
#include "common.cuh"
#include <stdlib.h>
#include "cta_select.cuh"
#include "rotated.cuh"
#include "cta_nms.cuh"
#include "wide_nms.cuh"

namespace {

constexpr int T = sel::THREADS;
constexpr int PRE_CAP = sel::KCAP;  // NMS_PRE_MAXSIZE limit (4096: PointPillars' value, data/pointpillar.yaml:141)
constexpr int POST_CAP = 1024;      // NMS_POST_MAXSIZE limit

// ---------------------------------------------------------------------------------------------------------
// rotated IoU matrix  (boxes_iou_bev_kernel, iou3d_nms_kernel.cu:251-265)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) rotated_iou_kernel(const float* __restrict__ a, int n, const float* __restrict__ b, int m,
                                                          float* __restrict__ out) {
    pdl_wait();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n * m) return;
    const int ia = (int)(i / m), ib = (int)(i - (long long)ia * m);
    const rot::RBox ra = rot::make_rbox(a + 7 * ia), rb = rot::make_rbox(b + 7 * ib);
    out[i] = rot::far_apart(ra, rb) ? 0.f : rot::iou_bev(ra, rb);
}

// ---------------------------------------------------------------------------------------------------------
// greedy rotated NMS over m boxes already in descending score order, resident in shared memory
//   nms_gpu (iou3d_nms_utils.py:84-99) = nms_kernel bitmask (iou3d_nms_kernel.cu:267-311) + host sweep
//   (iou3d_nms.cpp:116-132): box i is kept unless an earlier kept box k has iou_bev(k, i) > thresh.
// ---------------------------------------------------------------------------------------------------------
constexpr int QCAP = 8192;   // pending exact IoU tests per greedy round
struct RotPred {
    static constexpr bool kQueue = true;
    float thresh;
    int normal;    // 0: nms_gpu (rotated BEV IoU), 1: nms_normal_gpu (axis-aligned BEV IoU)
    // far apart: the reference's polygon (or axis-aligned intersection) is empty, IoU 0, never > thresh (>= 0)
    __device__ __forceinline__ bool maybe(const rot::RBox& a, const rot::RBox& b) const { return !rot::far_apart(a, b); }
    __device__ __forceinline__ bool exact(const rot::RBox& a, const rot::RBox& b) const { return rot::overlaps(a, b, thresh, normal); }
    __device__ __forceinline__ bool operator()(const rot::RBox& a, const rot::RBox& b) const { return maybe(a, b) && exact(a, b); }
};
typedef nms::Shared<rot::RBox, POST_CAP, PRE_CAP, QCAP> NmsShared;

__device__ __forceinline__ int cta_rotated_nms(const rot::RBox* __restrict__ R, int m, float thresh, int normal, int post_max, NmsShared& S) {
    return nms::cta_greedy_nms(R, m, RotPred{thresh, normal}, post_max, S);
}

// shared memory of the one-CTA-per-frame kernels
struct FrameShared {
    unsigned long long skeys[sel::KCAP];
    rot::RBox R[PRE_CAP];
    sel::Scratch sel;
    NmsShared nms;
};
static_assert(sizeof(FrameShared) <= 227 * 1024, "FrameShared exceeds the 227 KB of shared memory a CTA can opt in to");

// ---------------------------------------------------------------------------------------------------------
// tcb_rotated_nms: scores + boxes in any order -> top pre_max by score -> greedy NMS -> keep indices
// ---------------------------------------------------------------------------------------------------------
// blockIdx.x = box set of the batch (boxes [b,n,7], scores [b,n], keep [b,post_max], count [b])
__global__ void __launch_bounds__(T, 1) rotated_nms_kernel(const float* __restrict__ boxes, const float* __restrict__ scores, int n,
                                                           float thresh, int normal, int pre_max, int post_max,
                                                           int* __restrict__ keep, int* __restrict__ count) {
    pdl_wait();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FrameShared& S = *reinterpret_cast<FrameShared*>(smem_raw);
    boxes += 7ll * n * blockIdx.x; scores += (long long)n * blockIdx.x; keep += (long long)post_max * blockIdx.x; count += blockIdx.x;
    auto key_of = [&](int i) { return sel::make_key(__ldg(scores + i), (unsigned)i); };
    const int m = sel::cta_topk_sorted(key_of, n, pre_max, S.skeys, S.sel);
    for (int c = threadIdx.x; c < m; c += T) S.R[c] = rot::make_rbox(boxes + 7ll * sel::key_index(S.skeys[c]));
    __syncthreads();
    const int nk = cta_rotated_nms(S.R, m, thresh, normal, post_max, S.nms);
    for (int k = threadIdx.x; k < nk; k += T) keep[k] = (int)sel::key_index(S.skeys[S.nms.kpos[k]]);
    if (threadIdx.x == 0) *count = nk;
}

// ---------------------------------------------------------------------------------------------------------
// Wide path for large candidate sets (NMS_PRE_MAXSIZE > WIDE_MIN): the exact rotated IoU tests are the bulk of
// the work and a single CTA cannot run them faster than ~0.4 ms at 4096 boxes, so they are spread over the
// whole GPU as the reference does (nms_kernel, iou3d_nms_kernel.cu:267-311: one 64-bit word per row and
// 64-column block) — but the mask stays on the device and the sweep over it (iou3d_nms.cpp:116-132, a serial
// host loop there) is done by one warp that streams the rows of still-open candidates through shared memory.
//   order kernel (1 CTA)  : top pre_max by score, sorted -> workspace; a grid-wide kernel then gathers / decodes
//                           their boxes in that order
//   mask kernels (grid)   : 64x64 tiles of the upper triangle run the cheap distance test and list the surviving
//                           pairs; a second kernel evaluates the listed pairs, one per thread
//   sweep kernel (1 warp) : greedy pass over the mask (rows streamed through shared memory), writes the results
// ---------------------------------------------------------------------------------------------------------
constexpr int WIDE_MIN = 512;
constexpr int MW = wide::MW;          // mask words per row
static_assert(PRE_CAP == wide::PRE_CAP && POST_CAP == wide::POST_CAP, "wide_nms.cuh limits");
constexpr int MASK_T = 256;

constexpr int ORDER_CACHE = 16384;   // keys the order kernels keep in (dynamic) shared memory across the selection passes
struct OrderShared {
    unsigned long long skeys[sel::KCAP];
    sel::Scratch sel;
};

// Mask, step 1: the cheap distance test of every pair of the upper triangle. A CTA takes 64 rows x 256 columns: thread =
// (row, 64-column block) keeps its row in registers, walks the block's columns (x, y, radius + margin as one float4
// each, broadcast from shared memory) and collects the pairs it cannot rule out as a 64-bit word — no votes, no atomics
// in the loop. Those pairs go to a global list (one slot reservation per CTA) for rot_pairs_kernel; the mask words are
// written as zeros. Should the list be full, the thread evaluates its pairs itself.
// m_dev[0] = number of candidates, m_dev[1] = pairs listed so far (zeroed by the order kernel).
// blockIdx.z = candidate set of the batch; per-set strides: PRE_CAP boxes, WIDE_M ints, mask_stride words, pair_cap pairs
constexpr int WIDE_M = 64;   // ints per set in the m array ([0] candidates, [1] pairs listed)
constexpr int MASK_COLS = MASK_T;   // columns per CTA: MASK_T / 64 blocks of 64
__global__ void __launch_bounds__(MASK_T) rot_mask_kernel(const rot::RBox* __restrict__ R, int* __restrict__ m_dev, float thresh, int normal,
                                                          unsigned long long* __restrict__ mask, size_t mask_stride, unsigned* __restrict__ pairs, int pair_cap) {
    pdl_wait();
    R += (size_t)blockIdx.z * PRE_CAP; m_dev += blockIdx.z * WIDE_M; mask += blockIdx.z * mask_stride; pairs += (size_t)blockIdx.z * pair_cap;
    const int bi = blockIdx.y, c0 = blockIdx.x * MASK_COLS;   // row block, first column
    const int m = m_dev[0];
    if (c0 + MASK_COLS <= bi * 64 || c0 >= m || bi * 64 >= m) return;   // wholly below the diagonal / past the last candidate: never read
    __shared__ float4 cols[MASK_COLS];
    __shared__ int warp_tot[MASK_T / 32];
    __shared__ int qbase;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        const int g = c0 + tid;
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g < m) { const rot::RBox b = R[g]; c = make_float4(b.x, b.y, b.r + 0.05f, 0.f); }
        cols[tid] = c;
    }
    __syncthreads();
    const int r = tid & 63, q = tid >> 6;
    const int gi = bi * 64 + r, bj = (c0 >> 6) + q;           // this thread's row and column block
    unsigned lo = 0u, hi = 0u;
    const bool live = gi < m && bj >= bi && bj * 64 < m;
    if (live) {
        const float ax = R[gi].x, ay = R[gi].y, ar = R[gi].r;
        const float4* cb = cols + q * 64;
#pragma unroll
        for (int j = 0; j < 32; j++) {                        // far apart (rot::far_apart): squared distance > (ra + rb + 0.05)^2
            const float4 c = cb[j], d = cb[j + 32];
            const float x0 = ax - c.x, y0 = ay - c.y, l0 = ar + c.z, x1 = ax - d.x, y1 = ay - d.y, l1 = ar + d.z;
            if (!(x0 * x0 + y0 * y0 > l0 * l0)) lo |= 1u << j;
            if (!(x1 * x1 + y1 * y1 > l1 * l1)) hi |= 1u << j;
        }
        const int left = m - bj * 64;                         // columns of this block that exist
        if (left < 64) { if (left <= 32) { hi = 0u; lo &= (left == 32 ? 0xffffffffu : ((1u << left) - 1u)); } else hi &= (1u << (left - 32)) - 1u; }
        if (bj == bi) {                                       // diagonal block: columns after the row only
            if (r < 32) lo &= (r == 31 ? 0u : ~((2u << r) - 1u));
            else { lo = 0u; hi &= (r == 63 ? 0u : ~((2u << (r - 32)) - 1u)); }
        }
    }
    // slots for the CTA's pairs: exclusive scan of the per-thread counts, one reservation on the set's counter
    const int nq = __popc(lo) + __popc(hi);
    int inc = nq;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, s); if (lane >= s) inc += t; }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    if (tid == 0) {
        int tot = 0;
#pragma unroll
        for (int w = 0; w < MASK_T / 32; w++) { const int t = warp_tot[w]; warp_tot[w] = tot; tot += t; }
        qbase = tot ? atomicAdd(&m_dev[1], tot) : 0;
    }
    __syncthreads();
    if (!live) return;
    int at = qbase + warp_tot[warp] + inc - nq;
    unsigned long long word = 0ull;
    unsigned long long near = ((unsigned long long)hi << 32) | lo;
    while (near) {
        const int j = __ffsll((long long)near) - 1;
        near &= near - 1ull;
        const int gj = bj * 64 + j;
        if (at < pair_cap) pairs[at] = ((unsigned)gi << 16) | (unsigned)gj;
        else if (rot::overlaps(R[gi], R[gj], thresh, normal)) word |= 1ull << j;   // list full: evaluated here
        at++;
    }
    mask[(size_t)gi * MW + bj] = word;
}

// Mask, step 2: the exact IoU of the listed pairs. A warp takes 64 pairs at a time, drops those a separating axis rules
// out (rot::separated: about half of what the circle test lets through) and packs the rest into its slice of shared
// memory, so that the expensive exact test (polygon clip, ~2000 instructions) always runs with the lanes full.
constexpr int PAIR_WARPS = 8;
__global__ void __launch_bounds__(32 * PAIR_WARPS) rot_pairs_kernel(const rot::RBox* __restrict__ R, const int* __restrict__ m_dev, float thresh, int normal,
                                                                    const unsigned* __restrict__ pairs, int pair_cap, unsigned long long* __restrict__ mask, size_t mask_stride) {
    pdl_wait();
    __shared__ unsigned list[PAIR_WARPS][96];
    R += (size_t)blockIdx.y * PRE_CAP; m_dev += blockIdx.y * WIDE_M; mask += blockIdx.y * mask_stride; pairs += (size_t)blockIdx.y * pair_cap;
    const int total = min(m_dev[1], pair_cap);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    unsigned* mine = list[warp];
    auto exact = [&](unsigned e) {
        const int gi = (int)(e >> 16), gj = (int)(e & 0xffffu);
        const bool hit = normal ? rot::iou_normal(R[gi], R[gj]) > thresh : rot::iou_bev(R[gi], R[gj]) > thresh;
        if (hit) atomicOr(&mask[(size_t)gi * MW + (gj >> 6)], 1ull << (gj & 63));
    };
    int have = 0;
    for (int q0 = (blockIdx.x * PAIR_WARPS + warp) * 64; q0 < total; q0 += gridDim.x * PAIR_WARPS * 64) {
#pragma unroll
        for (int half = 0; half < 2; half++) {
            const int q = q0 + half * 32 + lane;
            const unsigned e = q < total ? pairs[q] : 0xffffffffu;
            const bool ok = e != 0xffffffffu && !rot::separated(R[e >> 16], R[e & 0xffffu], normal);
            const unsigned bal = __ballot_sync(0xffffffffu, ok);
            if (ok) mine[have + __popc(bal & lt)] = e;
            have += __popc(bal);
        }
        __syncwarp();
        while (have >= 32) {
            have -= 32;
            exact(mine[have + lane]);
        }
        __syncwarp();
    }
    if (lane < have) exact(mine[lane]);
}

using wide::SweepShared;
using wide::warp_mask_sweep;
using wide::SWEEP_T;

__global__ void __launch_bounds__(T, 1) rot_order_kernel(const float* __restrict__ scores, int n, int pre_max,
                                                         unsigned long long* __restrict__ skeys_out, int* __restrict__ m_out) {
    pdl_wait();
    __shared__ OrderShared S;
    extern __shared__ __align__(16) unsigned long long key_cache[];
    scores += (long long)n * blockIdx.x; skeys_out += (size_t)blockIdx.x * PRE_CAP; m_out += blockIdx.x * WIDE_M;
    auto key_of = [&](int i) { return sel::make_key(__ldg(scores + i), (unsigned)i); };
    const int m = sel::cta_topk_sorted(key_of, n, pre_max, S.skeys, S.sel, key_cache, ORDER_CACHE);
    for (int c = threadIdx.x; c < m; c += T) skeys_out[c] = S.skeys[c];
    if (threadIdx.x == 0) { m_out[0] = m; m_out[1] = 0; }
}

// the ordered candidates' boxes, one thread each over the whole grid (inside the one-CTA order kernel the four
// dependent rounds of scattered loads per thread cost more than the selection itself)
__global__ void __launch_bounds__(256) rot_rbox_kernel(const float* __restrict__ boxes, int n, const unsigned long long* __restrict__ skeys,
                                                       const int* __restrict__ m_dev, rot::RBox* __restrict__ R_out) {
    pdl_wait();
    boxes += 7ll * n * blockIdx.y; skeys += (size_t)blockIdx.y * PRE_CAP; m_dev += blockIdx.y * WIDE_M; R_out += (size_t)blockIdx.y * PRE_CAP;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < m_dev[0]) R_out[c] = rot::make_rbox(boxes + 7ll * sel::key_index(skeys[c]));
}

__global__ void __launch_bounds__(SWEEP_T, 1) rot_sweep_kernel(const unsigned long long* __restrict__ mask, size_t mask_stride, const unsigned long long* __restrict__ skeys,
                                                          const int* __restrict__ m_dev, int post_max, int* __restrict__ keep,
                                                          int* __restrict__ count) {
    pdl_wait();
    mask += blockIdx.x * mask_stride; skeys += (size_t)blockIdx.x * PRE_CAP; m_dev += blockIdx.x * WIDE_M; keep += (long long)post_max * blockIdx.x; count += blockIdx.x;
    extern __shared__ __align__(128) unsigned char sweep_raw[];
    SweepShared& S = *reinterpret_cast<SweepShared*>(sweep_raw);
    __shared__ int nk_s;
    if (threadIdx.x < 32) { const int v = warp_mask_sweep(mask, *m_dev, post_max, S); if (threadIdx.x == 0) nk_s = v; }
    __syncthreads();
    const int nk = nk_s;
    for (int k = threadIdx.x; k < nk; k += SWEEP_T) keep[k] = (int)sel::key_index(skeys[S.kpos[k]]);
    if (threadIdx.x == 0) *count = nk;
}

// ---------------------------------------------------------------------------------------------------------
// PointPillars anchor head post (d2-d4)
// ---------------------------------------------------------------------------------------------------------
constexpr int PP_MAX_CLASSES = 8, PP_MAX_ROT = 4;
struct PPParams {
    int H, W, num_class, num_rot, num_dir_bins;
    float x0, y0;
    double x_stride, y_stride;                   // anchor grid (align_center = False: stride = range/(n-1), float64 in pcdet)
    float size[PP_MAX_CLASSES][3];               // dx, dy, dz per class
    float zc[PP_MAX_CLASSES];                    // anchor centre height = bottom + dz/2 (anchor_generator.py:58)
    float rot[PP_MAX_CLASSES][PP_MAX_ROT];
    float dir_offset, dir_limit_offset, period;
    float score_thresh, nms_thresh;
    int nms_normal;
    int pre_max, post_max;
    int n_anchors;                               // H * W * num_class * num_rot: rows per frame of the batch
};

__device__ __forceinline__ float sigmoidf_ref(float x) { return 1.f / (1.f + expf(-x)); }  // torch.sigmoid

// per anchor: sigmoid of every class logit, max (detector3d_template.py:217,251), keep score >= thresh.
// The logits of all frames are cut into spans of `pa` anchors (pa * num_class floats, a whole number of
// 16-byte groups per lane) that the warps of a persistent grid walk, each on its own: the 16-byte asynchronous copies of
// a warp's NEXT span are in flight while it works on the current one, every lane reading back only what it copied itself.
// A logit below logit_floor cannot reach the threshold (the floor sits 0.01 below the threshold's logit, far more than
// expf's error), so the stream is only compared against it: a bit per logit in a register word, the positions of the few
// that pass listed in shared memory, and the sigmoids / exact test / key computed over the list with the lanes packed (an
// anchor listed through several classes is taken at its first). Keys collect in a per-warp pending list that is flushed
// with one reservation on the frame's counter.
#ifndef TCB_PC_NBUF
#define TCB_PC_NBUF 1
#endif
constexpr int PC_WARPS = 8, PC_PEND = 128, PC_NBUF = TCB_PC_NBUF;
// anchors per span: a multiple of 128 with at most 16 16-byte groups per lane (the lane's bit word has 64 bits)
static inline int pc_anchors(int num_class) { return num_class <= 2 ? 1024 : (num_class <= 4 ? 512 : 256); }
__device__ __forceinline__ void pc_copy16(float* dst_smem, const float* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
// shared memory of one warp: two span buffers, the position list, the pending keys
__host__ __device__ inline size_t pc_warp_bytes(int num_class, int pa) { return (size_t)pa * num_class * (PC_NBUF * 4 + 2) + PC_PEND * 8; }

__device__ __noinline__ float pc_best_of_all(const float* row, int num_class) {
    float best = sigmoidf_ref(row[0]);
    for (int k = 1; k < num_class; k++) best = fmaxf(best, sigmoidf_ref(row[k]));
    return best;
}

// NC: the class count as a compile-time constant (1..4: the bit-word loops unroll with constant shifts), 0 = any
template <int NC>
__global__ void __launch_bounds__(32 * PC_WARPS) pp_candidates_kernel(const float* cls, int b, int n_anchors, int num_class_rt, int pa_rt, float inv_nc, float thresh,
                                                                      float logit_floor, unsigned long long* __restrict__ cand, int* __restrict__ cand_count) {
    extern __shared__ __align__(16) unsigned char pc_raw[];
    pdl_wait();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int num_class = NC ? NC : num_class_rt;
    const int pa = NC ? (NC <= 2 ? 1024 : 512) : pa_rt;                 // = pc_anchors(num_class)
    const int span = pa * num_class, Q = span >> 7;                    // floats per span; 16-byte groups per lane
    unsigned char* mine_raw = pc_raw + (size_t)warp * pc_warp_bytes(num_class, pa);
    unsigned long long* pend = reinterpret_cast<unsigned long long*>(mine_raw);
    float* buf0 = reinterpret_cast<float*>(mine_raw + PC_PEND * 8);
    unsigned short* l_pos = reinterpret_cast<unsigned short*>(buf0 + PC_NBUF * span);
    const int per_frame = (n_anchors + pa - 1) / pa;
    const int total_items = per_frame * b, nworkers = gridDim.x * PC_WARPS;
    const unsigned lt = (1u << lane) - 1u;
    const long long frame_floats = (long long)n_anchors * num_class;

    struct Item { const float* src; int n, f, a0; };
    auto item_of = [&](int it) {
        Item q;
        int sp;
        q.f = it / per_frame; sp = it - q.f * per_frame;
        q.a0 = sp * pa;
        q.src = cls + q.f * frame_floats + (long long)q.a0 * num_class;
        q.n = min(pa, n_anchors - q.a0) * num_class;
        return q;
    };
    auto issue = [&](const Item& q, float* dst) {
        if (q.n == span && (((uintptr_t)q.src) & 15) == 0) {
            if (NC) {
#pragma unroll
                for (int g = 0; g < (NC ? (NC <= 2 ? 1024 : 512) * NC / 128 : 1); g++) pc_copy16(dst + (g * 32 + lane) * 4, q.src + (g * 32 + lane) * 4);
            } else {
                for (int g = 0; g < Q; g++) pc_copy16(dst + (g * 32 + lane) * 4, q.src + (g * 32 + lane) * 4);
            }
        } else {
            for (int g = 0; g < Q; g++) {
                const int e = (g * 32 + lane) * 4;
                if (e + 3 < q.n && ((((uintptr_t)(q.src + e)) & 15) == 0)) pc_copy16(dst + e, q.src + e);
                else {
#pragma unroll
                    for (int j = 0; j < 4; j++) dst[e + j] = e + j < q.n ? ld_coherent(q.src + e + j) : -INFINITY;
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int pend_n = 0, pend_f = -1;
    auto flush = [&]() {                                   // whole warp
        if (pend_n == 0) return;
        int base = 0;
        if (lane == 0) base = atomicAdd(cand_count + pend_f, pend_n);
        base = __shfl_sync(0xffffffffu, base, 0);
        unsigned long long* dst = cand + (long long)pend_f * n_anchors + base;
        for (int i = lane; i < pend_n; i += 32) dst[i] = pend[i];
        __syncwarp();
        pend_n = 0;
    };

    int it = blockIdx.x * PC_WARPS + warp, stage = 0;
    if (it >= total_items) return;
    Item cur = item_of(it);
    issue(cur, buf0);
    for (; it < total_items; it += nworkers, stage ^= (PC_NBUF - 1)) {
        const int nxt = it + nworkers;
        Item next = cur;
        if (PC_NBUF == 2 && nxt < total_items) {
            next = item_of(nxt);
            issue(next, buf0 + (stage ^ 1) * span);
            asm volatile("cp.async.wait_group 1;" ::: "memory");     // this span's copies: all but the group just committed
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        const float* x = buf0 + stage * span;
        unsigned long long above = 0ull;                   // bit g * 4 + j: the lane's element j of copy group g lies above the floor
        if (NC) {
            unsigned w[2] = {0u, 0u};
#pragma unroll
            for (int g = 0; g < (NC ? (NC <= 2 ? 1024 : 512) * NC / 128 : 1); g++) {
                const float4 v = *reinterpret_cast<const float4*>(x + (g * 32 + lane) * 4);   // what this lane copied itself
                if (v.x > logit_floor) w[g >> 3] |= 1u << ((g & 7) * 4);
                if (v.y > logit_floor) w[g >> 3] |= 2u << ((g & 7) * 4);
                if (v.z > logit_floor) w[g >> 3] |= 4u << ((g & 7) * 4);
                if (v.w > logit_floor) w[g >> 3] |= 8u << ((g & 7) * 4);
            }
            above = ((unsigned long long)w[1] << 32) | w[0];
        } else {
            for (int g = 0; g < Q; g++) {
                const float4 v = *reinterpret_cast<const float4*>(x + (g * 32 + lane) * 4);
                const unsigned m = (v.x > logit_floor ? 1u : 0u) | (v.y > logit_floor ? 2u : 0u) | (v.z > logit_floor ? 4u : 0u) | (v.w > logit_floor ? 8u : 0u);
                above |= (unsigned long long)m << (4 * g);
            }
        }
        const int mine = __popcll(above);
        int inc = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
        const int total = __shfl_sync(0xffffffffu, inc, 31);
        if (total > 0) {                                   // uniform over the warp
            if (pend_n && pend_f != cur.f) flush();
            pend_f = cur.f;
            int o = inc - mine;
            while (above) {
                const int bit = __ffsll((long long)above) - 1;
                above &= above - 1ull;
                l_pos[o++] = (unsigned short)((((bit >> 2) * 32 + lane) << 2) + (bit & 3));
            }
            __syncwarp();
            for (int r0 = 0; r0 < total; r0 += 32) {       // the list, lanes packed
                const int e = r0 + lane;
                bool ok = false;
                unsigned long long key = 0ull;
                if (e < total) {
                    const int pos = l_pos[e];
                    int a = (int)((float)pos * inv_nc), c = pos - a * num_class;
                    if (c < 0) { a--; c += num_class; } else if (c >= num_class) { a++; c -= num_class; }
                    const float* row = x + a * num_class;
                    bool first = true;                     // an anchor is evaluated through the first of its classes on the list
                    for (int k = 0; k < c; k++) first = first && !(row[k] > logit_floor);
                    if (first) {
                        // max over the classes of sigmoid(logit): the sigmoid of the largest logit — and of any other within
                        // 1e-5 of it (expf's 2 ulp could reorder only logits that close)
                        float top = row[0];
                        for (int k = 1; k < num_class; k++) top = fmaxf(top, row[k]);
                        float best = sigmoidf_ref(top);
                        bool close = false;
                        for (int k = 0; k < num_class; k++) close = close || (row[k] != top && row[k] > top - 1e-5f);
                        if (close) best = pc_best_of_all(row, num_class);          // (out of line: keeps the sigmoids out of the common path)
                        ok = best >= thresh;
                        key = sel::make_key(best, (unsigned)(cur.a0 + a));
                    }
                }
                const unsigned em = __ballot_sync(0xffffffffu, ok);
                if (em) {
                    if (pend_n + __popc(em) > PC_PEND) flush();
                    if (ok) pend[pend_n + __popc(em & lt)] = key;
                    pend_n += __popc(em);
                }
            }
            __syncwarp();                                  // the span buffer is rewritten by the copy issued in the next iteration
        }
        if (PC_NBUF == 1 && nxt < total_items) {           // one buffer per warp: more warps per SM hold the bytes in flight instead
            next = item_of(nxt);
            issue(next, buf0);
        }
        cur = next;
    }
    flush();
}

// anchor a -> decoded box (anchor_generator.py:17-60 analytically, box_coder_utils.py:45-77, anchor_head_template.py:253-263)
__device__ __forceinline__ void pp_decode(const PPParams& p, int a, const float* __restrict__ box_preds, const float* __restrict__ dir_preds,
                                          float* out7) {
    const int per_loc = p.num_class * p.num_rot;
    const int loc = a / per_loc, k = a - loc * per_loc;
    const int yy = loc / p.W, xx = loc - yy * p.W;
    const int cls = k / p.num_rot, ri = k - cls * p.num_rot;
    const float xa = (float)((double)p.x0 + p.x_stride * (double)xx);   // torch.arange(start, end, step): start + step*i in double
    const float ya = (float)((double)p.y0 + p.y_stride * (double)yy);
    const float za = p.zc[cls];
    const float dxa = p.size[cls][0], dya = p.size[cls][1], dza = p.size[cls][2], ra = p.rot[cls][ri];
    const float* t = box_preds + (long long)a * 7;
    const float xt = __ldg(t), yt = __ldg(t + 1), zt = __ldg(t + 2), dxt = __ldg(t + 3), dyt = __ldg(t + 4), dzt = __ldg(t + 5), rt = __ldg(t + 6);
    const float diagonal = sqrtf(dxa * dxa + dya * dya);
    out7[0] = xt * diagonal + xa;
    out7[1] = yt * diagonal + ya;
    out7[2] = zt * dza + za;
    out7[3] = expf(dxt) * dxa;
    out7[4] = expf(dyt) * dya;
    out7[5] = expf(dzt) * dza;
    float rg = rt + ra;
    if (dir_preds) {
        const float* d = dir_preds + (long long)a * p.num_dir_bins;
        int label = 0;
        float bestd = __ldg(d);
        for (int q = 1; q < p.num_dir_bins; q++) { const float v = __ldg(d + q); if (v > bestd) { bestd = v; label = q; } }
        const float val = rg - p.dir_offset;
        const float dir_rot = val - floorf(val / p.period + p.dir_limit_offset) * p.period;   // common_utils.limit_period
        rg = dir_rot + p.dir_offset + p.period * (float)label;
    }
    out7[6] = rg;
}

__global__ void __launch_bounds__(T, 1) pp_post_kernel(const PPParams p, const float* __restrict__ cls, const float* __restrict__ box_preds,
                                                       const float* __restrict__ dir_preds, const unsigned long long* __restrict__ cand,
                                                       const int* __restrict__ cand_count, int cand_cap, float* __restrict__ sorted_boxes,
                                                       float* __restrict__ out_boxes, float* __restrict__ out_scores,
                                                       long long* __restrict__ out_labels, int* __restrict__ out_count) {
    pdl_wait();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FrameShared& S = *reinterpret_cast<FrameShared*>(smem_raw);
    const int tid = threadIdx.x, f = blockIdx.x;    // one CTA per frame of the batch
    cls += (long long)f * p.n_anchors * p.num_class; box_preds += (long long)f * p.n_anchors * 7;
    if (dir_preds) dir_preds += (long long)f * p.n_anchors * p.num_dir_bins;
    cand += (long long)f * cand_cap; cand_count += f; sorted_boxes += (size_t)f * PRE_CAP * 7;
    out_boxes += (long long)f * p.post_max * 7; out_scores += (long long)f * p.post_max; out_labels += (long long)f * p.post_max; out_count += f;
    const int nc = min(*cand_count, cand_cap);
    auto key_of = [&](int i) { return cand[i]; };
    const int m = sel::cta_topk_sorted(key_of, nc, p.pre_max, S.skeys, S.sel);   // class_agnostic_nms topk (model_nms_utils.py:15)
    for (int c = tid; c < m; c += T) {
        float b7[7];
        pp_decode(p, (int)sel::key_index(S.skeys[c]), box_preds, dir_preds, b7);
#pragma unroll
        for (int j = 0; j < 7; j++) sorted_boxes[c * 7 + j] = b7[j];
        S.R[c] = rot::make_rbox(b7);
    }
    __syncthreads();
    const int nk = cta_rotated_nms(S.R, m, p.nms_thresh, p.nms_normal, p.post_max, S.nms);
    for (int k = tid; k < nk; k += T) {
        const int c = S.nms.kpos[k];
        const unsigned long long key = S.skeys[c];
        const int a = (int)sel::key_index(key);
#pragma unroll
        for (int j = 0; j < 7; j++) out_boxes[k * 7 + j] = sorted_boxes[c * 7 + j];
        out_scores[k] = sel::key_score(key);
        const float* q = cls + (long long)a * p.num_class;   // label = first arg-max of the sigmoids, + 1 (detector3d_template.py:251,256)
        int label = 0;
        float best = sigmoidf_ref(__ldg(q));
        for (int cc = 1; cc < p.num_class; cc++) { const float v = sigmoidf_ref(__ldg(q + cc)); if (v > best) { best = v; label = cc; } }
        out_labels[k] = (long long)(label + 1);
    }
    if (tid == 0) *out_count = nk;
}

// wide path (see rot_mask_kernel): the same steps as pp_post_kernel with the IoU work between them on the whole GPU
__global__ void __launch_bounds__(T, 1) pp_order_kernel(const PPParams p, const unsigned long long* __restrict__ cand, const int* __restrict__ cand_count,
                                                        int cand_cap, unsigned long long* __restrict__ skeys_out, int* __restrict__ m_out) {
    pdl_wait();
    __shared__ OrderShared S;
    extern __shared__ __align__(16) unsigned long long key_cache[];
    const int tid = threadIdx.x, f = blockIdx.x;
    cand += (long long)f * cand_cap; cand_count += f; skeys_out += (size_t)f * PRE_CAP; m_out += f * WIDE_M;
    const int nc = min(*cand_count, cand_cap);
    auto key_of = [&](int i) { return cand[i]; };
    const int m = sel::cta_topk_sorted(key_of, nc, p.pre_max, S.skeys, S.sel, key_cache, ORDER_CACHE);
    for (int c = tid; c < m; c += T) skeys_out[c] = S.skeys[c];
    if (tid == 0) { m_out[0] = m; m_out[1] = 0; }
}

// lazy decode of the ordered candidates only (<= pre_max of the 321 408 anchors), one thread each over the grid
__global__ void __launch_bounds__(256) pp_decode_kernel(const PPParams p, const float* __restrict__ box_preds, const float* __restrict__ dir_preds,
                                                        const unsigned long long* __restrict__ skeys, const int* __restrict__ m_dev,
                                                        float* __restrict__ sorted_boxes, rot::RBox* __restrict__ R_out) {
    pdl_wait();
    const int f = blockIdx.y;
    box_preds += (long long)f * p.n_anchors * 7;
    if (dir_preds) dir_preds += (long long)f * p.n_anchors * p.num_dir_bins;
    skeys += (size_t)f * PRE_CAP; m_dev += f * WIDE_M; sorted_boxes += (size_t)f * PRE_CAP * 7; R_out += (size_t)f * PRE_CAP;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= m_dev[0]) return;
    float b7[7];
    pp_decode(p, (int)sel::key_index(skeys[c]), box_preds, dir_preds, b7);
#pragma unroll
    for (int j = 0; j < 7; j++) sorted_boxes[c * 7 + j] = b7[j];
    R_out[c] = rot::make_rbox(b7);
}

__global__ void __launch_bounds__(SWEEP_T, 1) pp_finish_kernel(const PPParams p, const float* __restrict__ cls, const unsigned long long* __restrict__ mask,
                                                          size_t mask_stride, const unsigned long long* __restrict__ skeys, const int* __restrict__ m_dev,
                                                          const float* __restrict__ sorted_boxes, float* __restrict__ out_boxes,
                                                          float* __restrict__ out_scores, long long* __restrict__ out_labels,
                                                          int* __restrict__ out_count) {
    pdl_wait();
    extern __shared__ __align__(128) unsigned char sweep_raw[];
    SweepShared& S = *reinterpret_cast<SweepShared*>(sweep_raw);
    const int tid = threadIdx.x, f = blockIdx.x;
    cls += (long long)f * p.n_anchors * p.num_class; mask += f * mask_stride; skeys += (size_t)f * PRE_CAP; m_dev += f * WIDE_M;
    sorted_boxes += (size_t)f * PRE_CAP * 7;
    out_boxes += (long long)f * p.post_max * 7; out_scores += (long long)f * p.post_max; out_labels += (long long)f * p.post_max; out_count += f;
    __shared__ int nk_s;
    if (tid < 32) { const int v = warp_mask_sweep(mask, *m_dev, p.post_max, S); if (tid == 0) nk_s = v; }
    __syncthreads();
    const int nk = nk_s;
    for (int k = tid; k < nk; k += SWEEP_T) {
        const int c = S.kpos[k];
        const unsigned long long key = skeys[c];
        const int a = (int)sel::key_index(key);
#pragma unroll
        for (int j = 0; j < 7; j++) out_boxes[k * 7 + j] = sorted_boxes[c * 7 + j];
        out_scores[k] = sel::key_score(key);
        const float* q = cls + (long long)a * p.num_class;
        int label = 0;
        float best = sigmoidf_ref(__ldg(q));
        for (int cc = 1; cc < p.num_class; cc++) { const float v = sigmoidf_ref(__ldg(q + cc)); if (v > best) { best = v; label = cc; } }
        out_labels[k] = (long long)(label + 1);
    }
    if (tid == 0) *out_count = nk;
}

// ---------------------------------------------------------------------------------------------------------
// CenterPoint head post (d6): one CTA per head
// ---------------------------------------------------------------------------------------------------------
constexpr int CP_MAX_HEADS = 8, CP_MAX_HEAD_CLASSES = 4;
struct CPHead {
    const float *hm, *center, *center_z, *dim, *rot, *vel;  // NCHW planes of one frame
    int num_class;
    int class_map[CP_MAX_HEAD_CLASSES];
};
struct CPParams {
    CPHead head[CP_MAX_HEADS];
    int num_heads, H, W, K;
    float stride, vx, vy, rx, ry;        // feature_map_stride, voxel size, range min
    float limit[6];                      // POST_CENTER_LIMIT_RANGE
    float score_thresh, nms_thresh;
    int nms_normal;
    int pre_max, post_max;
};

__global__ void __launch_bounds__(T, 1) cp_head_kernel(const CPParams p, float* __restrict__ head_boxes, float* __restrict__ head_scores,
                                                       int* __restrict__ head_labels, int* __restrict__ head_count) {
    pdl_wait();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FrameShared& S = *reinterpret_cast<FrameShared*>(smem_raw);
    const int tid = threadIdx.x, h = blockIdx.x, f = blockIdx.y;   // head, frame of the batch ([b,c,H,W] tensors)
    CPHead hd = p.head[h];
    const int HW = p.H * p.W, n = hd.num_class * HW;
    hd.hm += (long long)f * n; hd.center += (long long)f * 2 * HW; hd.center_z += (long long)f * HW; hd.dim += (long long)f * 3 * HW;
    hd.rot += (long long)f * 2 * HW;
    if (hd.vel) hd.vel += (long long)f * 2 * HW;
    {
        const long long rows = (long long)p.num_heads * (p.pre_max + p.post_max);
        head_boxes += f * rows * 9; head_scores += f * rows; head_labels += f * rows; head_count += f * CP_MAX_HEADS;
    }
    // _topk (centernet_utils.py:136-151): per-class top-K then top-K across classes == global top-K of the sigmoids
    // The selection makes seven passes over the scores: compute each sigmoid once into shared memory (as ordered
    // 32-bit patterns) — with K <= 1024 only the first 1024 slots of skeys are used, so the rest of skeys and all of
    // R (not needed before the decode below) hold 34 816 scores, every cell of a 2-class 128x128 head.
    unsigned* sc = reinterpret_cast<unsigned*>(&S.skeys[1024]);
    constexpr int SC_CAP = (int)(((sel::KCAP - 1024) * sizeof(unsigned long long) + sizeof(S.R)) / sizeof(unsigned));
    static_assert(offsetof(FrameShared, R) == sizeof(unsigned long long) * sel::KCAP, "score cache assumes R follows skeys");
    const int cached = p.K <= 1024 ? min(n, SC_CAP) : 0;
#pragma unroll 8
    for (int i = tid; i < cached; i += T) sc[i] = sel::f32_ordered(sigmoidf_ref(__ldg(hd.hm + i)));
    __syncthreads();
    auto key_of = [&](int i) {
        return i < cached ? (((unsigned long long)sc[i] << 32) | (unsigned long long)(0xffffffffu - (unsigned)i))
                          : sel::make_key(sigmoidf_ref(__ldg(hd.hm + i)), (unsigned)i);
    };
    const int m = sel::cta_topk_sorted(key_of, n, p.K, S.skeys, S.sel, nullptr, 0, cached ? 1024 : sel::KCAP);   // the score cache starts at skeys[1024]
    // decode_bbox_from_heatmap (centernet_utils.py:154-216) + range/score mask, order preserved
    // scratch rows: [0, heads*post_max) final rows per head, then per head pre_max rows of masked, still sorted boxes
    float* boxes = head_boxes + (long long)h * p.post_max * 9;
    // masked compaction must preserve score order: do it in chunks of T with an ordered scan
    int base = 0;
    for (int c0 = 0; c0 < m; c0 += T) {
        const int c = c0 + tid;
        bool ok = false;
        float b9[9];
        float score = 0.f;
        int cls = 0;
        if (c < m) {
            const unsigned long long key = S.skeys[c];
            const int i = (int)sel::key_index(key);
            score = sel::key_score(key);
            cls = i / HW;
            const int ind = i - cls * HW;
            const float ys = (float)(ind / p.W), xs = (float)(ind % p.W);
            const float cx = __ldg(hd.center + ind), cy = __ldg(hd.center + HW + ind);
            b9[0] = (xs + cx) * p.stride * p.vx + p.rx;
            b9[1] = (ys + cy) * p.stride * p.vy + p.ry;
            b9[2] = __ldg(hd.center_z + ind);
            b9[3] = expf(__ldg(hd.dim + ind)); b9[4] = expf(__ldg(hd.dim + HW + ind)); b9[5] = expf(__ldg(hd.dim + 2 * HW + ind));
            b9[6] = atan2f(__ldg(hd.rot + HW + ind), __ldg(hd.rot + ind));   // rot[:,1] = sin, rot[:,0] = cos (center_head.py:266-267)
            b9[7] = hd.vel ? __ldg(hd.vel + ind) : 0.f;
            b9[8] = hd.vel ? __ldg(hd.vel + HW + ind) : 0.f;
            ok = b9[0] >= p.limit[0] && b9[1] >= p.limit[1] && b9[2] >= p.limit[2] &&
                 b9[0] <= p.limit[3] && b9[1] <= p.limit[4] && b9[2] <= p.limit[5] && score > p.score_thresh;
        }
        int total;
        const int r = sel::cta_exclusive_scan(ok ? 1 : 0, S.sel.warp_tot, total);
        if (ok && base + r < p.pre_max) {   // class_agnostic_nms topk(min(NMS_PRE_MAXSIZE, n)) on an already sorted list
            const int d = base + r;
            S.R[d] = rot::make_rbox(b9);
            float* row = head_boxes + ((long long)p.num_heads * p.post_max + (long long)h * p.pre_max + d) * 9;
#pragma unroll
            for (int j = 0; j < 9; j++) row[j] = b9[j];
            head_scores[p.num_heads * p.post_max + h * p.pre_max + d] = score;
            head_labels[p.num_heads * p.post_max + h * p.pre_max + d] = hd.class_map[cls];
        }
        base += total;
        __syncthreads();
    }
    const int mv = min(base, p.pre_max);
    __syncthreads();
    const int nk = cta_rotated_nms(S.R, mv, p.nms_thresh, p.nms_normal, p.post_max, S.nms);
    for (int k = tid; k < nk; k += T) {
        const int c = S.nms.kpos[k];
        const long long src = (long long)p.num_heads * p.post_max + (long long)h * p.pre_max + c;
#pragma unroll
        for (int j = 0; j < 9; j++) boxes[k * 9 + j] = head_boxes[src * 9 + j];
        head_scores[h * p.post_max + k] = head_scores[src];
        head_labels[h * p.post_max + k] = head_labels[src];
    }
    if (tid == 0) head_count[h] = nk;
}

// concatenate the heads in order, labels + 1 (center_head.py:299-302)
__global__ void __launch_bounds__(256) cp_concat_kernel(int num_heads, int post_max, const float* __restrict__ head_boxes,
                                                        const float* __restrict__ head_scores, const int* __restrict__ head_labels,
                                                        const int* __restrict__ head_count, float* __restrict__ out_boxes,
                                                        float* __restrict__ out_scores, long long* __restrict__ out_labels, int* __restrict__ out_count,
                                                        int pre_max) {
    pdl_wait();
    {
        const long long f = blockIdx.x, rows = (long long)num_heads * (pre_max + post_max), orows = (long long)num_heads * post_max;
        head_boxes += f * rows * 9; head_scores += f * rows; head_labels += f * rows; head_count += f * CP_MAX_HEADS;
        out_boxes += f * orows * 9; out_scores += f * orows; out_labels += f * orows; out_count += f;
    }
    int off = 0;
    for (int h = 0; h < num_heads; h++) {
        const int n = head_count[h];
        for (int k = threadIdx.x; k < n; k += blockDim.x) {
            for (int j = 0; j < 9; j++) out_boxes[(off + k) * 9 + j] = head_boxes[((long long)h * post_max + k) * 9 + j];
            out_scores[off + k] = head_scores[h * post_max + k];
            out_labels[off + k] = (long long)head_labels[h * post_max + k] + 1;
        }
        off += n;
    }
    if (threadIdx.x == 0) *out_count = off;
}

// ---------------------------------------------------------------------------------------------------------
// nuScenes per-class score cut (clients/postprocess/detector_3d_postprocess.py:55-64,98-133): indices of
// label == cls && score >= thr[cls], concatenated in the reference's fixed class order.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) class_filter_kernel(const long long* __restrict__ labels, const float* __restrict__ scores, int n,
                                                            int n_groups, const int* __restrict__ group_cls, const float* __restrict__ group_thr,
                                                            int* __restrict__ out_idx, int* __restrict__ out_count) {
    pdl_wait();
    __shared__ int warp_tot[32];
    int base = 0;
    for (int g = 0; g < n_groups; g++) {
        const int cls = group_cls[g];
        const float thr = group_thr[g];
        for (int c0 = 0; c0 < n; c0 += 1024) {
            const int i = c0 + threadIdx.x;
            const bool ok = i < n && labels[i] == (long long)cls && scores[i] >= thr;
            int total;
            const int r = sel::cta_exclusive_scan(ok ? 1 : 0, warp_tot, total);
            if (ok) out_idx[base + r] = i;
            base += total;
        }
    }
    if (threadIdx.x == 0) *out_count = base;
}

}  // namespace

// =========================================================================================================
// C ABI
// =========================================================================================================
extern "C" int tcb_rotated_iou(tcb_ctx* ctx, const float* boxes_a_dev, int n, const float* boxes_b_dev, int m, float* out_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, n >= 0 && m >= 0, "tcb_rotated_iou: bad sizes");
    if (n == 0 || m == 0) return TCB_OK;
    TCB_REQUIRE(ctx, boxes_a_dev && boxes_b_dev && out_dev, "tcb_rotated_iou: null buffer");
    TCB_KLAUNCH(ctx, rotated_iou_kernel, tcb_div_up((long long)n * m, 256), 256, 0, (cudaStream_t)stream, boxes_a_dev, n, boxes_b_dev, m, out_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

static int check_nms_limits(tcb_ctx* ctx, const char* who, float thresh, int nms_type, int pre_max, int post_max) {
    if (!(thresh >= 0.f)) return tcb_fail(ctx, TCB_ERR_ARG, "%s: nms threshold must be >= 0", who);
    if (nms_type != TCB_NMS_GPU && nms_type != TCB_NMS_NORMAL) return tcb_fail(ctx, TCB_ERR_ARG, "%s: nms_type must be TCB_NMS_GPU or TCB_NMS_NORMAL", who);
    if (pre_max < 1 || post_max < 1) return tcb_fail(ctx, TCB_ERR_ARG, "%s: pre/post max size must be >= 1", who);
    if (pre_max > PRE_CAP) return tcb_fail(ctx, TCB_ERR_LIMIT, "%s: NMS_PRE_MAXSIZE %d > compiled limit %d", who, pre_max, PRE_CAP);
    if (post_max > POST_CAP) return tcb_fail(ctx, TCB_ERR_LIMIT, "%s: NMS_POST_MAXSIZE %d > compiled limit %d", who, post_max, POST_CAP);
    return TCB_OK;
}

static int init_post3d(tcb_ctx* ctx) {   // run by tcb_create
    TCB_SMEM_OPT_IN(ctx, pp_candidates_kernel<0>, 200 * 1024);
    TCB_SMEM_OPT_IN(ctx, pp_candidates_kernel<1>, 200 * 1024);
    TCB_SMEM_OPT_IN(ctx, pp_candidates_kernel<2>, 200 * 1024);
    TCB_SMEM_OPT_IN(ctx, pp_candidates_kernel<3>, 200 * 1024);
    TCB_SMEM_OPT_IN(ctx, pp_candidates_kernel<4>, 200 * 1024);
    TCB_SMEM_OPT_IN(ctx, rotated_nms_kernel, sizeof(FrameShared));
    TCB_SMEM_OPT_IN(ctx, pp_post_kernel, sizeof(FrameShared));
    TCB_SMEM_OPT_IN(ctx, cp_head_kernel, sizeof(FrameShared));
    TCB_SMEM_OPT_IN(ctx, rot_order_kernel, ORDER_CACHE * 8);
    TCB_SMEM_OPT_IN(ctx, pp_order_kernel, ORDER_CACHE * 8);
    TCB_SMEM_OPT_IN(ctx, rot_sweep_kernel, sizeof(SweepShared));
    TCB_SMEM_OPT_IN(ctx, pp_finish_kernel, sizeof(SweepShared));
    return TCB_OK;
}
static TcbInitReg reg_post3d(&init_post3d);

// workspace of the wide path for `sets` candidate sets; false (nothing taken) when the context's workspace has no room for it
struct WideBufs {
    unsigned long long* skeys;     // [sets][PRE_CAP]
    rot::RBox* R;                  // [sets][PRE_CAP]
    unsigned long long* mask;      // [sets][mask_stride]
    size_t mask_stride;
    int* m;                        // [sets][WIDE_M]
    unsigned* pairs;               // [sets][pair_cap]
    int pair_cap;
};
static int wide_pair_cap(int m_max, int sets) {   // every pair of a small set; 2 M pairs (8 MB) beyond that (512 K per set in a batch), the tiles take the overflow
    const long long all = (long long)m_max * (m_max - 1) / 2, cap = sets > 1 ? (512ll << 10) : (2ll << 20);
    return (int)(all < cap ? (all > 0 ? all : 1) : cap);
}
static size_t wide_bytes(int m_max, int sets) {
    return (size_t)sets * ((size_t)PRE_CAP * (8 + sizeof(rot::RBox)) + (size_t)m_max * MW * 8 + (size_t)wide_pair_cap(m_max, sets) * 4 + WIDE_M * 4) + 6 * 256;
}
static bool wide_bufs(tcb_ctx* ctx, int m_max, int sets, WideBufs& wb) {
    if (((ctx->ws_used + 255) & ~(size_t)255) + wide_bytes(m_max, sets) > ctx->ws_bytes) return false;
    wb.skeys = (unsigned long long*)tcb_ws_alloc(ctx, (size_t)sets * PRE_CAP * 8);
    wb.R = (rot::RBox*)tcb_ws_alloc(ctx, (size_t)sets * PRE_CAP * sizeof(rot::RBox));
    wb.mask_stride = (size_t)m_max * MW;
    wb.mask = (unsigned long long*)tcb_ws_alloc(ctx, (size_t)sets * wb.mask_stride * 8);
    wb.m = (int*)tcb_ws_alloc(ctx, (size_t)sets * WIDE_M * 4);
    wb.pair_cap = wide_pair_cap(m_max, sets);
    wb.pairs = (unsigned*)tcb_ws_alloc(ctx, (size_t)sets * wb.pair_cap * 4);
    return wb.skeys && wb.R && wb.mask && wb.m && wb.pairs;
}
// mask (tiles + listed pairs) of `sets` ordered candidate sets already in wb.R / wb.m
static int wide_mask(tcb_ctx* ctx, const WideBufs& wb, int m_max, int sets, float thresh, int normal, cudaStream_t st) {
    const int nb = tcb_div_up(m_max, 64);
    TCB_KLAUNCH(ctx, rot_mask_kernel, dim3(tcb_div_up(nb * 64, MASK_COLS), nb, sets), MASK_T, 0, st, wb.R, wb.m, thresh, normal, wb.mask, wb.mask_stride, wb.pairs, wb.pair_cap);
    TCB_LAUNCHED(ctx);
    const int per = sets > 1 ? tcb_div_up(ctx->num_sms * 8, sets) : ctx->num_sms * 8;
    TCB_KLAUNCH(ctx, rot_pairs_kernel, dim3(per, sets), 256, 0, st, wb.R, wb.m, thresh, normal, wb.pairs, wb.pair_cap, wb.mask, wb.mask_stride);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

extern "C" int tcb_rotated_nms(tcb_ctx* ctx, const float* boxes_dev, const float* scores_dev, int b, int n, float thresh, int nms_type, int pre_max,
                               int post_max, int* keep_dev, int* count_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, b >= 1 && n >= 0 && keep_dev && count_dev && (n == 0 || (boxes_dev && scores_dev)), "tcb_rotated_nms: bad arguments");
    if (b > 65535) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_rotated_nms: more than 65535 box sets");
    int rc = check_nms_limits(ctx, "tcb_rotated_nms", thresh, nms_type, pre_max, post_max);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    WideBufs wb;
    const int m_max = n < pre_max ? n : pre_max;
    if (m_max > WIDE_MIN && wide_bufs(ctx, m_max, b, wb)) {
        TCB_KLAUNCH(ctx, rot_order_kernel, b, T, ORDER_CACHE * 8, st, scores_dev, n, pre_max, wb.skeys, wb.m);
        TCB_LAUNCHED(ctx);
        TCB_KLAUNCH(ctx, rot_rbox_kernel, dim3(tcb_div_up(m_max, 256), b), 256, 0, st, boxes_dev, n, wb.skeys, wb.m, wb.R);
        TCB_LAUNCHED(ctx);
        if ((rc = wide_mask(ctx, wb, m_max, b, thresh, nms_type, st))) return rc;
        TCB_KLAUNCH(ctx, rot_sweep_kernel, b, SWEEP_T, sizeof(SweepShared), st, wb.mask, wb.mask_stride, wb.skeys, wb.m, post_max, keep_dev, count_dev);
        TCB_LAUNCHED(ctx);
        return TCB_OK;
    }
    TCB_KLAUNCH(ctx, rotated_nms_kernel, b, T, sizeof(FrameShared), st, boxes_dev, scores_dev, n, thresh, nms_type, pre_max, post_max, keep_dev, count_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

extern "C" size_t tcb_pp_anchor_post_workspace(int b, int h, int w, int num_class, int num_rot) {
    const size_t n = (size_t)h * w * num_class * num_rot, bb = (size_t)(b > 0 ? b : 1);
    return bb * (n * 8 + (size_t)PRE_CAP * 7 * 4 + 512) + 4096 + wide_bytes(PRE_CAP, (int)bb);
}

extern "C" int tcb_pp_anchor_post(tcb_ctx* ctx, const float* cls_dev, const float* box_dev, const float* dir_dev, int b, int h, int w,
                                  int num_class, int num_rot, const float anchor_range[6], const float* anchor_sizes,
                                  const float* anchor_bottoms, const float* anchor_rotations, float dir_offset,
                                  float dir_limit_offset, int num_dir_bins, float score_thresh, float nms_thresh, int nms_type,
                                  int pre_max, int post_max, float* out_boxes_dev, float* out_scores_dev, long long* out_labels_dev,
                                  int* out_count_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, cls_dev && box_dev && anchor_range && anchor_sizes && anchor_bottoms && anchor_rotations &&
                     out_boxes_dev && out_scores_dev && out_labels_dev && out_count_dev, "tcb_pp_anchor_post: null argument");
    TCB_REQUIRE(ctx, b >= 1 && h >= 2 && w >= 2, "tcb_pp_anchor_post: need b >= 1 and a feature map of at least 2x2");
    if (b > 65535) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_pp_anchor_post: more than 65535 frames");
    if (num_class < 1 || num_class > PP_MAX_CLASSES || num_rot < 1 || num_rot > PP_MAX_ROT)
        return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_pp_anchor_post: num_class %d (<= %d) / num_rot %d (<= %d)", num_class, PP_MAX_CLASSES, num_rot, PP_MAX_ROT);
    TCB_REQUIRE(ctx, dir_dev == nullptr || num_dir_bins >= 1, "tcb_pp_anchor_post: num_dir_bins must be >= 1");
    int rc = check_nms_limits(ctx, "tcb_pp_anchor_post", nms_thresh, nms_type, pre_max, post_max);
    if (rc) return rc;
    const long long n_anchors = (long long)h * w * num_class * num_rot;
    if (n_anchors > (1ll << 30)) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_pp_anchor_post: too many anchors");
    cudaStream_t st = (cudaStream_t)stream;

    PPParams p;
    memset(&p, 0, sizeof(p));
    p.H = h; p.W = w; p.num_class = num_class; p.num_rot = num_rot; p.num_dir_bins = num_dir_bins; p.n_anchors = (int)n_anchors;
    // anchor_generator.py:28-31: float32 range difference divided by a numpy int64 grid size -> float64 stride
    p.x0 = anchor_range[0]; p.y0 = anchor_range[1];
    p.x_stride = (double)(anchor_range[3] - anchor_range[0]) / (double)(w - 1);
    p.y_stride = (double)(anchor_range[4] - anchor_range[1]) / (double)(h - 1);
    for (int c = 0; c < num_class; c++) {
        for (int j = 0; j < 3; j++) p.size[c][j] = anchor_sizes[c * 3 + j];
        p.zc[c] = anchor_bottoms[c] + p.size[c][2] / 2;
        for (int r = 0; r < num_rot; r++) p.rot[c][r] = anchor_rotations[c * num_rot + r];
    }
    p.dir_offset = dir_offset; p.dir_limit_offset = dir_limit_offset;
    p.period = (float)(2.0 * 3.141592653589793 / (double)(num_dir_bins > 0 ? num_dir_bins : 1));
    p.score_thresh = score_thresh; p.nms_thresh = nms_thresh; p.nms_normal = nms_type; p.pre_max = pre_max; p.post_max = post_max;

    TCB_WS(ctx, cand_count, int, b + 64);
    TCB_WS(ctx, cand, unsigned long long, (size_t)b * n_anchors);
    TCB_WS(ctx, sorted_boxes, float, (size_t)b * PRE_CAP * 7);
    TCB_CUDA(ctx, cudaMemsetAsync(cand_count, 0, sizeof(int) * b, st));
    const double st_d = (double)score_thresh;
    const float logit_floor = st_d <= 0.0 ? -INFINITY : (st_d >= 1.0 ? INFINITY : (float)(log(st_d / (1.0 - st_d)) - 0.01));
    {
        const size_t pc_smem = pc_warp_bytes(num_class, pc_anchors(num_class)) * PC_WARPS;
        int per_sm = (int)((220 * 1024) / (pc_smem + 1024));
        per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
        const int pa = pc_anchors(num_class);
        const long long items = (long long)tcb_div_up(n_anchors, pa) * b;
        if (items >= (1ll << 31)) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_pp_anchor_post: batch too large");
        const long long want = tcb_div_up(items, PC_WARPS);
        const int ctas = (int)(want < (long long)ctx->num_sms * per_sm ? want : (long long)ctx->num_sms * per_sm);
        auto launch = [&](auto kernel) -> int {
            TCB_KLAUNCH(ctx, kernel, ctas, 32 * PC_WARPS, pc_smem, st, cls_dev, b, (int)n_anchors, num_class, pa, 1.0f / (float)num_class, score_thresh,
                        logit_floor, cand, cand_count);
            TCB_LAUNCHED(ctx);
            return TCB_OK;
        };
        if ((rc = num_class == 1 ? launch(pp_candidates_kernel<1>) : num_class == 2 ? launch(pp_candidates_kernel<2>) : num_class == 3 ? launch(pp_candidates_kernel<3>)
                : num_class == 4 ? launch(pp_candidates_kernel<4>) : launch(pp_candidates_kernel<0>))) return rc;
    }
    WideBufs wb;
    const int m_max = (int)(n_anchors < pre_max ? n_anchors : pre_max);
    if (m_max > WIDE_MIN && wide_bufs(ctx, m_max, b, wb)) {
        TCB_KLAUNCH(ctx, pp_order_kernel, b, T, ORDER_CACHE * 8, st, p, cand, cand_count, (int)n_anchors, wb.skeys, wb.m);
        TCB_LAUNCHED(ctx);
        TCB_KLAUNCH(ctx, pp_decode_kernel, dim3(tcb_div_up(m_max, 256), b), 256, 0, st, p, box_dev, dir_dev, wb.skeys, wb.m, sorted_boxes, wb.R);
        TCB_LAUNCHED(ctx);
        if ((rc = wide_mask(ctx, wb, m_max, b, nms_thresh, nms_type, st))) return rc;
        TCB_KLAUNCH(ctx, pp_finish_kernel, b, SWEEP_T, sizeof(SweepShared), st, p, cls_dev, wb.mask, wb.mask_stride, wb.skeys, wb.m, sorted_boxes, out_boxes_dev, out_scores_dev,
                                                             out_labels_dev, out_count_dev);
        TCB_LAUNCHED(ctx);
        return TCB_OK;
    }
    TCB_KLAUNCH(ctx, pp_post_kernel, b, T, sizeof(FrameShared), st, p, cls_dev, box_dev, dir_dev, cand, cand_count, (int)n_anchors, sorted_boxes,
                                                       out_boxes_dev, out_scores_dev, out_labels_dev, out_count_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

extern "C" size_t tcb_centerpoint_post_workspace(int b, int num_heads, int pre_max, int post_max) {
    return (size_t)(b > 0 ? b : 1) * ((size_t)num_heads * (size_t)(pre_max + post_max) * (9 * 4 + 4 + 4) + 256) + 4096;
}

extern "C" int tcb_centerpoint_post(tcb_ctx* ctx, int b, int num_heads, const float* const* hm_dev, const float* const* center_dev,
                                    const float* const* center_z_dev, const float* const* dim_dev, const float* const* rot_dev,
                                    const float* const* vel_dev, const int* head_num_class, const int* class_id_map, int h, int w,
                                    int feature_map_stride, const float voxel_size_xy[2], const float range_min_xy[2],
                                    const float limit_range[6], int k, float score_thresh, float nms_thresh, int nms_type, int pre_max, int post_max,
                                    float* out_boxes_dev, float* out_scores_dev, long long* out_labels_dev, int* out_count_dev,
                                    void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, hm_dev && center_dev && center_z_dev && dim_dev && rot_dev && head_num_class && class_id_map && voxel_size_xy &&
                     range_min_xy && limit_range && out_boxes_dev && out_scores_dev && out_labels_dev && out_count_dev,
                "tcb_centerpoint_post: null argument");
    if (num_heads < 1 || num_heads > CP_MAX_HEADS) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_centerpoint_post: %d heads (limit %d)", num_heads, CP_MAX_HEADS);
    TCB_REQUIRE(ctx, b >= 1 && b <= 65535 && h >= 1 && w >= 1 && k >= 1, "tcb_centerpoint_post: bad shape");
    if (k > PRE_CAP) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_centerpoint_post: K %d > %d", k, PRE_CAP);
    int rc = check_nms_limits(ctx, "tcb_centerpoint_post", nms_thresh, nms_type, pre_max, post_max);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;

    CPParams p;
    memset(&p, 0, sizeof(p));
    p.num_heads = num_heads; p.H = h; p.W = w; p.K = k;
    int cm = 0;
    for (int i = 0; i < num_heads; i++) {
        CPHead& hd = p.head[i];
        hd.hm = hm_dev[i]; hd.center = center_dev[i]; hd.center_z = center_z_dev[i]; hd.dim = dim_dev[i]; hd.rot = rot_dev[i];
        hd.vel = vel_dev ? vel_dev[i] : nullptr;
        hd.num_class = head_num_class[i];
        if (hd.num_class < 1 || hd.num_class > CP_MAX_HEAD_CLASSES)
            return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_centerpoint_post: head %d has %d classes (limit %d)", i, hd.num_class, CP_MAX_HEAD_CLASSES);
        TCB_REQUIRE(ctx, hd.hm && hd.center && hd.center_z && hd.dim && hd.rot, "tcb_centerpoint_post: head %d has a null tensor", i);
        for (int c = 0; c < hd.num_class; c++) hd.class_map[c] = class_id_map[cm++];
    }
    p.stride = (float)feature_map_stride; p.vx = voxel_size_xy[0]; p.vy = voxel_size_xy[1]; p.rx = range_min_xy[0]; p.ry = range_min_xy[1];
    for (int j = 0; j < 6; j++) p.limit[j] = limit_range[j];
    p.score_thresh = score_thresh; p.nms_thresh = nms_thresh; p.nms_normal = nms_type; p.pre_max = pre_max; p.post_max = post_max;

    const size_t rows = (size_t)b * num_heads * (size_t)(pre_max + post_max);
    TCB_WS(ctx, head_count, int, (size_t)b * CP_MAX_HEADS + 64);
    TCB_WS(ctx, head_boxes, float, rows * 9);
    TCB_WS(ctx, head_scores, float, rows);
    TCB_WS(ctx, head_labels, int, rows);
    TCB_KLAUNCH(ctx, cp_head_kernel, dim3(num_heads, b), T, sizeof(FrameShared), st, p, head_boxes, head_scores, head_labels, head_count);
    TCB_LAUNCHED(ctx);
    TCB_KLAUNCH(ctx, cp_concat_kernel, b, 256, 0, st, num_heads, post_max, head_boxes, head_scores, head_labels, head_count, out_boxes_dev,
                                        out_scores_dev, out_labels_dev, out_count_dev, pre_max);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

extern "C" int tcb_class_score_filter(tcb_ctx* ctx, const long long* labels_dev, const float* scores_dev, int n, int n_groups,
                                      const int* group_class_host, const float* group_thresh_host, int* out_idx_dev,
                                      int* out_count_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, n >= 0 && n_groups >= 0 && n_groups <= 64 && out_idx_dev && out_count_dev, "tcb_class_score_filter: bad arguments");
    TCB_REQUIRE(ctx, n == 0 || (labels_dev && scores_dev), "tcb_class_score_filter: null input");
    TCB_REQUIRE(ctx, n_groups == 0 || (group_class_host && group_thresh_host), "tcb_class_score_filter: null group tables");
    cudaStream_t st = (cudaStream_t)stream;
    TCB_WS(ctx, gcls, int, 64);
    TCB_WS(ctx, gthr, float, 64);
    if (n_groups) {
        int rc;
        if ((rc = tcb_stage_h2d(ctx, gcls, group_class_host, sizeof(int) * n_groups, st))) return rc;
        if ((rc = tcb_stage_h2d(ctx, gthr, group_thresh_host, sizeof(float) * n_groups, st))) return rc;
    }
    TCB_KLAUNCH(ctx, class_filter_kernel, 1, 1024, 0, st, labels_dev, scores_dev, n, n_groups, gcls, gthr, out_idx_dev, out_count_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}
