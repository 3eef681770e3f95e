// Fused 2D preprocess: u8 HWC -> OpenCV-exact bilinear resample -> channel order -> x/div - sub -> f32 NCHW.
// One pass over each frame. HBM-bound: 3hw bytes in, 12 out_h out_w bytes out.
//
// Tiling: one CTA of 256 threads owns a 128-column x TH-row output tile. The source footprint of the
// tile (a few KB..tens of KB of u8) is staged into shared memory with 16-byte global loads, then each
// warp walks RPW consecutive output rows, each lane producing 4 adjacent pixels x 3 channels and
// issuing three 16-byte streaming stores (one per colour plane, 512 B contiguous per warp).
// The horizontal pass of a source row is kept in registers and reused by the next output row when
// both need it (vertical scale near 1).
#include "common.cuh"
#include <math.h>

namespace {

constexpr int TW = 128;       // output tile width (32 lanes x 4 px)
constexpr int NTHREADS = 256; // 8 warps

struct PreParams {
    const uint8_t* src;
    const uint8_t* batch_end;  // one past the last valid source byte of the whole batch
    long long img_stride, row_stride;
    int h, w;
    float* dst;
    int out_h, out_w;
    int left, top, new_w, new_h;  // content rectangle inside the output
    int swap;
    int exact_recip;  // 1: v/div == fma-corrected v*rdiv for every v in 0..255 (checked on the host)
    float div, rdiv;
    float sub[3];
    float padv[3];  // normalised pad value per output channel
    double scale_x, scale_y;
    int th;     // output rows per tile (8, 16 or 32)
    int pitch;  // shared-memory row pitch in bytes (multiple of 16)
    int tpr_shift;  // log2(threads cooperating on one staged row)
    int stage_bytes, hs_bytes, nr_cap;  // two-pass kernel: staged-byte area, 16-bit row buffer (3 planes x nr_cap rows x TW)
    float rlo;  // float32 residual of 1/div (two-instruction exact division)
    int exact2;  // fma(v, rdiv, v*rlo) == v/div for all 256 inputs
    int vec_store;
    int fast_area;  // 1: width halved exactly, 2: width unchanged — one fused pass straight from the staged bytes (aligned column groups)
};

// u8 -> normalised f32, identical to numpy's `x.astype(float32) / div - sub` (true IEEE division).
// EXACT: v*(1/div) corrected by two FMAs equals v/div for every v in 0..255 (verified on the host).
template <bool EXACT>
__device__ __forceinline__ float norm_px(int v, float div, float rdiv, float sub) {
    const float f = __int_as_float(0x4B000000 | v) - 8388608.0f;  // exact int->float for 0..255 without I2F
    float q;
    if (EXACT) {
        const float q0 = f * rdiv;
        q = __fmaf_rn(__fmaf_rn(-q0, div, f), rdiv, q0);
    } else {
        q = f / div;
    }
    return q - sub;
}

// OpenCV resizeLinear column taps for destination column dx (modules/imgproc resize.cpp semantics,
// restated from its documented behaviour: half-pixel centres, clamp to edge with zero weight, 11-bit taps).
__host__ __device__ __forceinline__ int round_i(float v) {
#ifdef __CUDA_ARCH__
    return __float2int_rn(v);
#else
    return (int)nearbyintf(v);   // default rounding mode: ties to even, like cvRound
#endif
}
__host__ __device__ __forceinline__ void x_tap(int dx, double scale, int w, int& sx, int& a0, int& a1) {
    float f = (float)(((double)dx + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    if (s < 0) { f = 0.f; s = 0; }
    if (s >= w - 1) { f = 0.f; s = w - 1; }
    a0 = round_i((1.f - f) * 2048.f);
    a1 = round_i(f * 2048.f);
    sx = s;
}
// Row taps: weights are NOT clamped, only the two row indices are (so border rows blend a row with itself).
__host__ __device__ __forceinline__ void y_tap(int dy, double scale, int h, int& r0, int& r1, int& b0, int& b1) {
    float f = (float)(((double)dy + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    b0 = round_i((1.f - f) * 2048.f);
    b1 = round_i(f * 2048.f);
    r0 = s < 0 ? 0 : (s > h - 1 ? h - 1 : s);
    r1 = s + 1 < 0 ? 0 : (s + 1 > h - 1 ? h - 1 : s + 1);
}

// Rows of one warp. INTERIOR: every pixel of the lane's 4 columns and every row lies inside the content
// rectangle, so no per-pixel predicate is needed (the common case: all tiles but the borders).
template <bool AREA, bool STAGED, bool EXACT, bool INTERIOR>
__device__ __forceinline__ void tile_rows(const PreParams& p, const uint8_t* __restrict__ smem, const uint8_t* __restrict__ img,
                                          int z, int ox0, int ty0, int warp, int xmin, int rmin, bool has_content,
                                          const int (&off0)[4], const int (&d1)[4], const int (&a0)[4], const int (&a1)[4],
                                          const bool (&inx)[4]) {
    const int rpw = p.th >> 3;  // consecutive rows per warp
    // channel c of the source lands in plane pl[c]; normalisation constants follow the plane
    int pl[3];
    float subc[3], padc[3];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        pl[c] = p.swap ? 2 - c : c;
        subc[c] = p.swap ? p.sub[2 - c] : p.sub[c];
        padc[c] = p.swap ? p.padv[2 - c] : p.padv[c];
    }
    int hprev[12];
    int rprev = -1;
#pragma unroll
    for (int i = 0; i < 12; i++) hprev[i] = 0;

    for (int rr = 0; rr < rpw; rr++) {
        const int oy = ty0 + warp * rpw + rr;
        if (oy >= p.out_h) break;
        const int dy = oy - p.top;
        const bool iny = INTERIOR || (has_content && dy >= 0 && dy < p.new_h);
        int v[12];
        if (iny) {
            int r0, r1, b0 = 0, b1 = 0;
            if (AREA) { r0 = 2 * dy; r1 = 2 * dy + 1; }
            else y_tap(dy, p.scale_y, p.h, r0, r1, b0, b1);
            const uint8_t *s0, *s1;
            if (STAGED) {
                const long long g0 = (long long)r0 * p.row_stride + (long long)xmin * 3;
                const long long g1 = (long long)r1 * p.row_stride + (long long)xmin * 3;
                s0 = smem + (r0 - rmin) * p.pitch + (int)((uintptr_t)(img + g0) & 15);
                s1 = smem + (r1 - rmin) * p.pitch + (int)((uintptr_t)(img + g1) & 15);
            } else {
                s0 = img + (long long)r0 * p.row_stride + (long long)xmin * 3;
                s1 = img + (long long)r1 * p.row_stride + (long long)xmin * 3;
            }
            if (AREA) {
                int t[48];
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int c = 0; c < 3; c++) {
                        const int k = off0[i] + c;
                        const bool ok = INTERIOR || inx[i];
                        t[(i * 3 + c) * 4 + 0] = ok ? (int)s0[k] : 0;
                        t[(i * 3 + c) * 4 + 1] = ok ? (int)s0[k + 3] : 0;
                        t[(i * 3 + c) * 4 + 2] = ok ? (int)s1[k] : 0;
                        t[(i * 3 + c) * 4 + 3] = ok ? (int)s1[k + 3] : 0;
                    }
#pragma unroll
                for (int j = 0; j < 12; j++) v[j] = (t[j * 4] + t[j * 4 + 1] + t[j * 4 + 2] + t[j * 4 + 3] + 2) >> 2;
            } else {
                int h0[12], h1[12];
                const bool reuse = (r0 == rprev);
                if (reuse) {
#pragma unroll
                    for (int j = 0; j < 12; j++) h0[j] = hprev[j];
                } else {
#pragma unroll
                    for (int i = 0; i < 4; i++)
#pragma unroll
                        for (int c = 0; c < 3; c++) {
                            const int k = off0[i] + c;
                            const bool ok = INTERIOR || inx[i];
                            h0[i * 3 + c] = ok ? (int)s0[k] * a0[i] + (int)s0[k + d1[i]] * a1[i] : 0;
                        }
                }
                if (r1 == r0) {
#pragma unroll
                    for (int j = 0; j < 12; j++) h1[j] = h0[j];
                } else {
#pragma unroll
                    for (int i = 0; i < 4; i++)
#pragma unroll
                        for (int c = 0; c < 3; c++) {
                            const int k = off0[i] + c;
                            const bool ok = INTERIOR || inx[i];
                            h1[i * 3 + c] = ok ? (int)s1[k] * a0[i] + (int)s1[k + d1[i]] * a1[i] : 0;
                        }
                }
#pragma unroll
                for (int j = 0; j < 12; j++) {
                    v[j] = (((b0 * (h0[j] >> 4)) >> 16) + ((b1 * (h1[j] >> 4)) >> 16) + 2) >> 2;
                    hprev[j] = h1[j];
                }
                rprev = r1;
            }
        }
#pragma unroll
        for (int c = 0; c < 3; c++) {
            float o[4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const bool ok = INTERIOR || (iny && inx[i]);
                o[i] = ok ? norm_px<EXACT>(v[i * 3 + c], p.div, p.rdiv, subc[c]) : padc[c];
            }
            float* d = p.dst + (((long long)z * 3 + pl[c]) * p.out_h + oy) * (long long)p.out_w + ox0;
            if (p.vec_store) {
                st_stream_f4(d, make_float4(o[0], o[1], o[2], o[3]));
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if (ox0 + i < p.out_w) d[i] = o[i];
            }
        }
    }
}

template <bool AREA, bool STAGED, bool EXACT>
__global__ void __launch_bounds__(NTHREADS) preprocess_kernel(const PreParams p) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int z = blockIdx.z;
    const int tx0 = blockIdx.x * TW, ty0 = blockIdx.y * p.th;
    const int tx1 = min(tx0 + TW, p.out_w) - 1, ty1 = min(ty0 + p.th, p.out_h) - 1;
    const uint8_t* img = p.src + (long long)z * p.img_stride;

    // ---- content part of this tile, in content coordinates --------------------------------------
    const int cx0 = max(tx0 - p.left, 0), cx1 = min(tx1 - p.left, p.new_w - 1);
    const int cy0 = max(ty0 - p.top, 0), cy1 = min(ty1 - p.top, p.new_h - 1);
    const bool has_content = (cx0 <= cx1) && (cy0 <= cy1);
    // whole 128 x th tile inside the content rectangle?
    const bool interior = has_content && (tx0 - p.left >= 0) && (tx0 + TW - 1 - p.left < p.new_w) && (tx0 + TW <= p.out_w) &&
                          (ty0 - p.top >= 0) && (ty0 + p.th - 1 - p.top < p.new_h) && (ty0 + p.th <= p.out_h);

    int xmin = 0, xmax = 0, rmin = 0, rmax = 0;
    if (has_content) {
        if (AREA) {
            xmin = 2 * cx0; xmax = 2 * cx1 + 1; rmin = 2 * cy0; rmax = 2 * cy1 + 1;
        } else {
            int s, a, b, r0, r1;
            x_tap(cx0, p.scale_x, p.w, s, a, b); xmin = s;
            x_tap(cx1, p.scale_x, p.w, s, a, b); xmax = min(s + 1, p.w - 1);
            y_tap(cy0, p.scale_y, p.h, r0, r1, a, b); rmin = r0;
            y_tap(cy1, p.scale_y, p.h, r0, r1, a, b); rmax = r1;
        }
    }

    // ---- stage the source footprint: rows rmin..rmax, bytes xmin*3 .. xmax*3+2 -------------------
    // 2^tpr_shift threads walk the 16-byte chunks of one row; the rest of the CTA takes other rows.
    if (STAGED && has_content) {
        const int nrows = rmax - rmin + 1;
        const int seg = (xmax - xmin + 1) * 3;
        const int cpr = p.pitch >> 4;
        const int k = tid & ((1 << p.tpr_shift) - 1);
        const int rstep = NTHREADS >> p.tpr_shift;
        if (k < cpr) {
            for (int r = tid >> p.tpr_shift; r < nrows; r += rstep) {
                const uint8_t* a = img + (long long)(rmin + r) * p.row_stride + (long long)xmin * 3;
                const uint8_t* g = (const uint8_t*)((uintptr_t)a & ~(uintptr_t)15) + 16 * k;
                if (g >= a + seg) continue;
                uint4 v;
                if (g >= p.src && g + 16 <= p.batch_end) {
                    v = ld_stream_u4(g);
                } else {  // first/last few bytes of the caller's buffer: never read outside it
                    uint8_t t[16];
#pragma unroll
                    for (int i = 0; i < 16; i++) t[i] = (g + i >= p.src && g + i < p.batch_end) ? g[i] : 0;
                    v = *reinterpret_cast<uint4*>(t);
                }
                *reinterpret_cast<uint4*>(smem + r * p.pitch + 16 * k) = v;
            }
        }
    }
    __syncthreads();

    // ---- per-lane column taps (fixed for the whole tile) -----------------------------------------
    const int ox0 = tx0 + lane * 4;
    int off0[4], d1[4], a0[4], a1[4];
    bool inx[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int dx = ox0 + i - p.left;
        inx[i] = has_content && dx >= 0 && dx < p.new_w;
        off0[i] = 0; d1[i] = 0; a0[i] = 0; a1[i] = 0;
        if (inx[i]) {
            if (AREA) {
                off0[i] = (2 * dx - xmin) * 3;
            } else {
                int s;
                x_tap(dx, p.scale_x, p.w, s, a0[i], a1[i]);
                off0[i] = (s - xmin) * 3;
                d1[i] = (min(s + 1, p.w - 1) - s) * 3;
            }
        }
    }
    if (ox0 >= p.out_w) return;
    if (interior) tile_rows<AREA, STAGED, EXACT, true>(p, smem, img, z, ox0, ty0, warp, xmin, rmin, has_content, off0, d1, a0, a1, inx);
    else tile_rows<AREA, STAGED, EXACT, false>(p, smem, img, z, ox0, ty0, warp, xmin, rmin, has_content, off0, d1, a0, a1, inx);
}

// -------------------------------------------------------------------------------------------------------------
// Two-pass tile kernel (the staged path): like OpenCV itself, first the horizontal pass of every staged source row
// into a 16-bit row buffer, then the vertical pass. Compared with recomputing the horizontal taps per output row this
// halves the byte loads, and the vertical pass reads 12 consecutive 16-bit values per lane with 8-byte loads.
//   stage : cp.async 16-byte copies of the footprint (all in flight at once), tap tables built meanwhile
//   pass H: thread = (column, row parity): hs[r][x*3+c] = (p0*a0 + p1*a1) >> 4        (AREA: p0 + p1)
//   pass V: warp = output rows, lane = 4 columns: v = (((b0*h0) >> 16) + ((b1*h1) >> 16) + 2) >> 2 (32-bit multiply + shift)
//           (AREA: (h0 + h1 + 2) >> 2), normalise, three 16-byte streaming stores per row
// -------------------------------------------------------------------------------------------------------------
struct XTap { int off; short a0, a1; };          // byte offset of the left tap inside the staged row; weights
struct YTap { int r0, r1; unsigned b0s, b1s; };  // staged-row indices; weights << 16; r0 < 0: row outside the content


// Source footprint of every tile column / tile row, computed once on the host with the same tap arithmetic
// (count 0 = no content in that tile column/row). Passed by value with the kernel parameters.
constexpr int GEO_MAX_TX = 64, GEO_MAX_TY = 1024;
struct TileGeo {
    int xmin[GEO_MAX_TX]; short xcnt[GEO_MAX_TX];
    int rmin[GEO_MAX_TY]; short rcnt[GEO_MAX_TY];
    int first_ty, last_ty;   // first / last tile row with content (the only ones that can touch the ends of the caller's buffer)
};

struct TileHdr { int xmin, xmax, rmin, rmax, has_content, interior, safe; };

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned phase) {
    unsigned done = 0;
    while (!done) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(phase) : "memory");
    }
}

// u8 -> normalised f32 with two FP instructions when the host has verified, for all 256 inputs, that
// fma(v, r, fl(v*rlo)) == v/div (r = fl(1/div), rlo = its float32 residual); otherwise the IEEE division.
template <bool EXACT, bool HAS_SUB>
__device__ __forceinline__ float norm2(int v, float div, float r, float rlo, float sub) {
    const float f = (float)v;
    float q = EXACT ? __fmaf_rn(f, r, f * rlo) : f / div;
    if (HAS_SUB) q = q - sub;
    return q;
}

template <bool AREA, bool EXACT, bool HAS_SUB, bool INTERIOR>
__device__ __forceinline__ void pass_v(const PreParams& p, const unsigned short* __restrict__ hs, const XTap* __restrict__ xt,
                                       const YTap* __restrict__ yt, int nr_cap, int z, int tx0, int ty0) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ox0 = tx0 + lane * 4;
    if (ox0 >= p.out_w) return;
    bool inx[4];
#pragma unroll
    for (int i = 0; i < 4; i++) inx[i] = INTERIOR || xt[lane * 4 + i].off >= 0;
    const int rpw = p.th >> 3;
    const long long plane = (long long)p.out_h * p.out_w;
    float* d0 = p.dst + ((long long)z * 3) * plane + (long long)(ty0 + warp * rpw) * p.out_w + ox0;
    for (int rr = 0; rr < rpw; rr++, d0 += p.out_w) {
        const int j = warp * rpw + rr;
        if (ty0 + j >= p.out_h) break;
        const YTap y = yt[j];
        const bool iny = INTERIOR || y.r0 >= 0;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const int cc = p.swap ? 2 - c : c;   // destination plane of source channel c
            float o[4];
            if (iny) {
                const uint2 a = *reinterpret_cast<const uint2*>(hs + (c * nr_cap + y.r0) * TW + lane * 4);
                const uint2 b = *reinterpret_cast<const uint2*>(hs + (c * nr_cap + y.r1) * TW + lane * 4);
                const unsigned av[4] = {a.x & 0xffffu, a.x >> 16, a.y & 0xffffu, a.y >> 16};
                const unsigned bv[4] = {b.x & 0xffffu, b.x >> 16, b.y & 0xffffu, b.y >> 16};
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const int v = AREA ? (int)((av[i] + bv[i] + 2u) >> 2) : (int)(((((y.b0s >> 16) * av[i]) >> 16) + (((y.b1s >> 16) * bv[i]) >> 16) + 2u) >> 2);
                    o[i] = (INTERIOR || inx[i]) ? norm2<EXACT, HAS_SUB>(v, p.div, p.rdiv, p.rlo, p.sub[cc]) : p.padv[cc];
                }
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++) o[i] = p.padv[cc];
            }
            float* d = d0 + cc * plane;
            if (p.vec_store) {
                st_stream_f4(d, make_float4(o[0], o[1], o[2], o[3]));
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if (ox0 + i < p.out_w) d[i] = o[i];
            }
        }
    }
}

// Width halved exactly (the reference's own 1280 -> 640, and the 2:1 letterbox), fused into one pass: warp = output
// rows, lane = 4 output columns = 24 contiguous source bytes on each of two staged rows, read as three 8-byte words
// per row (stride 24 B across lanes: conflict-free). Channel c of output pixel k is bytes 6k+c and 6k+c+3; each
// horizontal sum is a short dp4a chain with constant byte weights.
//   AREA  (height halved too): cv2's INTER_LINEAR is the 2x2 box mean (a+b+c+d+2)>>2; weights 1, the 2 rides in as
//         the accumulator
//   !AREA (any height): the x taps are (1024, 1024), so cv2's horizontal pass >> 4 is 64*(l + r): weights 64; the
//         vertical pass is the general one
// The host only selects this path when every 4-column group is entirely content or entirely padding and the
// column offsets are 8-byte aligned (tcb_preprocess_image: fast_area).
template <bool AREA, bool EXACT, bool HAS_SUB, bool INTERIOR>
__device__ __forceinline__ void pass_area(const PreParams& p, const uint8_t* __restrict__ stg, const XTap* __restrict__ xt,
                                          const YTap* __restrict__ yt, unsigned shift, int z, int tx0, int ty0) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ox0 = tx0 + lane * 4;
    if (ox0 >= p.out_w) return;
    const int xoff = xt[lane * 4].off;
    const bool inx = INTERIOR || xoff >= 0;
    const uint8_t* colbase = stg + shift + (inx ? xoff : 0);
    const int rpw = p.th >> 3;
    const long long plane = (long long)p.out_h * p.out_w;
    float* d0 = p.dst + ((long long)z * 3) * plane + (long long)(ty0 + warp * rpw) * p.out_w + ox0;
    for (int rr = 0; rr < rpw; rr++, d0 += p.out_w) {
        const int j = warp * rpw + rr;
        if (ty0 + j >= p.out_h) break;
        const YTap y = yt[j];
        float o[3][4];
        if (inx && (INTERIOR || y.r0 >= 0)) {
            const uint2* r0 = reinterpret_cast<const uint2*>(colbase + y.r0 * p.pitch);
            const uint2* r1 = reinterpret_cast<const uint2*>(colbase + y.r1 * p.pitch);
            const uint2 a0 = r0[0], a1 = r0[1], a2 = r0[2], b0 = r1[0], b1 = r1[1], b2 = r1[2];
            const unsigned A[6] = {a0.x, a0.y, a1.x, a1.y, a2.x, a2.y}, B[6] = {b0.x, b0.y, b1.x, b1.y, b2.x, b2.y};
#pragma unroll
            for (int h = 0; h < 2; h++) {   // pixel pair h: bytes 12h .. 12h+11 = words 3h .. 3h+2
                const unsigned u0 = A[3 * h], u1 = A[3 * h + 1], u2 = A[3 * h + 2], v0 = B[3 * h], v1 = B[3 * h + 1], v2 = B[3 * h + 2];
                unsigned s[6];
                if (AREA) {
                    s[0] = __dp4a(v0, 0x01000001u, __dp4a(u0, 0x01000001u, 2u));                                                        // B0 + B3
                    s[1] = __dp4a(v1, 0x00000001u, __dp4a(v0, 0x00000100u, __dp4a(u1, 0x00000001u, __dp4a(u0, 0x00000100u, 2u))));    // B1 + B4
                    s[2] = __dp4a(v1, 0x00000100u, __dp4a(v0, 0x00010000u, __dp4a(u1, 0x00000100u, __dp4a(u0, 0x00010000u, 2u))));    // B2 + B5
                    s[3] = __dp4a(v2, 0x00000100u, __dp4a(v1, 0x00010000u, __dp4a(u2, 0x00000100u, __dp4a(u1, 0x00010000u, 2u))));    // B6 + B9
                    s[4] = __dp4a(v2, 0x00010000u, __dp4a(v1, 0x01000000u, __dp4a(u2, 0x00010000u, __dp4a(u1, 0x01000000u, 2u))));    // B7 + B10
                    s[5] = __dp4a(v2, 0x01000001u, __dp4a(u2, 0x01000001u, 2u));                                                        // B8 + B11
                } else {
                    // x taps (1024, 1024): cv2's horizontal pass gives 1024*(l + r), its >> 4 leaves 64*(l + r) (weights 0x40);
                    // then the vertical pass as everywhere else: ((b0*S0) >> 16) + ((b1*S1) >> 16) + 2
                    const unsigned t0[6] = {__dp4a(u0, 0x40000040u, 0u), __dp4a(u1, 0x00000040u, __dp4a(u0, 0x00004000u, 0u)),
                                            __dp4a(u1, 0x00004000u, __dp4a(u0, 0x00400000u, 0u)), __dp4a(u2, 0x00004000u, __dp4a(u1, 0x00400000u, 0u)),
                                            __dp4a(u2, 0x00400000u, __dp4a(u1, 0x40000000u, 0u)), __dp4a(u2, 0x40000040u, 0u)};
                    const unsigned t1[6] = {__dp4a(v0, 0x40000040u, 0u), __dp4a(v1, 0x00000040u, __dp4a(v0, 0x00004000u, 0u)),
                                            __dp4a(v1, 0x00004000u, __dp4a(v0, 0x00400000u, 0u)), __dp4a(v2, 0x00004000u, __dp4a(v1, 0x00400000u, 0u)),
                                            __dp4a(v2, 0x00400000u, __dp4a(v1, 0x40000000u, 0u)), __dp4a(v2, 0x40000040u, 0u)};
                    // ((b*S) >> 16 as a 32-bit multiply and a shift: b <= 2048, S <= 32640, the product stays below 2^27. The
                    //  high-half multiply that does it in one instruction runs on a throttled pipe: 0.0945 -> 0.0874 ms on C2)
                    const unsigned wb0 = y.b0s >> 16, wb1 = y.b1s >> 16;
#pragma unroll
                    for (int q = 0; q < 6; q++) s[q] = ((wb0 * t0[q]) >> 16) + ((wb1 * t1[q]) >> 16) + 2u;
                }
#pragma unroll
                for (int k = 0; k < 2; k++)
#pragma unroll
                    for (int c = 0; c < 3; c++) {
                        const int cc = p.swap ? 2 - c : c;
                        o[c][2 * h + k] = norm2<EXACT, HAS_SUB>((int)(s[3 * k + c] >> 2), p.div, p.rdiv, p.rlo, p.sub[cc]);
                    }
            }
        } else {
#pragma unroll
            for (int c = 0; c < 3; c++) {
                const float pv = p.padv[p.swap ? 2 - c : c];
#pragma unroll
                for (int i = 0; i < 4; i++) o[c][i] = pv;
            }
        }
#pragma unroll
        for (int c = 0; c < 3; c++) st_stream_f4(d0 + (p.swap ? 2 - c : c) * plane, make_float4(o[c][0], o[c][1], o[c][2], o[c][3]));
    }
}

// Width unchanged (image_adjust on an already model-sized frame, padding modes, vertical-only resizes): the x taps are
// (2048, 0), so cv2's horizontal pass >> 4 is 128*p and the vertical pass ((b0*128p0)>>16) + ((b1*128p1)>>16) equals
// ((b0*p0)>>9) + ((b1*p1)>>9). Lane = 4 output columns = 12 contiguous staged bytes per row (three 4-byte words,
// stride 12 B across lanes: conflict-free); no row buffer. Same host conditions as pass_area (4-byte aligned groups).
template <bool EXACT, bool HAS_SUB, bool INTERIOR>
__device__ __forceinline__ void pass_ident(const PreParams& p, const uint8_t* __restrict__ stg, const XTap* __restrict__ xt,
                                           const YTap* __restrict__ yt, unsigned shift, int z, int tx0, int ty0) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ox0 = tx0 + lane * 4;
    if (ox0 >= p.out_w) return;
    const int xoff = xt[lane * 4].off;
    const bool inx = INTERIOR || xoff >= 0;
    const uint8_t* colbase = stg + shift + (inx ? xoff : 0);
    const int rpw = p.th >> 3;
    const long long plane = (long long)p.out_h * p.out_w;
    float* d0 = p.dst + ((long long)z * 3) * plane + (long long)(ty0 + warp * rpw) * p.out_w + ox0;
    for (int rr = 0; rr < rpw; rr++, d0 += p.out_w) {
        const int j = warp * rpw + rr;
        if (ty0 + j >= p.out_h) break;
        const YTap y = yt[j];
        float o[3][4];
        if (inx && (INTERIOR || y.r0 >= 0)) {
            const unsigned* r0 = reinterpret_cast<const unsigned*>(colbase + y.r0 * p.pitch);
            const unsigned* r1 = reinterpret_cast<const unsigned*>(colbase + y.r1 * p.pitch);
            const unsigned b0 = y.b0s >> 16, b1 = y.b1s >> 16;
            const unsigned A[3] = {r0[0], r0[1], r0[2]};
            unsigned B[3] = {0u, 0u, 0u};
            if (b1) { B[0] = r1[0]; B[1] = r1[1]; B[2] = r1[2]; }     // (uniform per output row)
#pragma unroll
            for (int e = 0; e < 12; e++) {                            // byte e = pixel e/3, channel e%3
                const unsigned p0 = __byte_perm(A[e >> 2], 0u, 0x4440u | (unsigned)(e & 3));
                const unsigned p1 = __byte_perm(B[e >> 2], 0u, 0x4440u | (unsigned)(e & 3));
                const unsigned v = (((b0 * p0) >> 9) + ((b1 * p1) >> 9) + 2u) >> 2;
                const int c = e % 3, k = e / 3;
                o[c][k] = norm2<EXACT, HAS_SUB>((int)v, p.div, p.rdiv, p.rlo, p.sub[p.swap ? 2 - c : c]);
            }
        } else {
#pragma unroll
            for (int c = 0; c < 3; c++) {
                const float pv = p.padv[p.swap ? 2 - c : c];
#pragma unroll
                for (int i = 0; i < 4; i++) o[c][i] = pv;
            }
        }
#pragma unroll
        for (int c = 0; c < 3; c++) st_stream_f4(d0 + (p.swap ? 2 - c : c) * plane, make_float4(o[c][0], o[c][1], o[c][2], o[c][3]));
    }
}

// Output the size of the source, nothing padded, packed rows: cv2.resize is then the identity (x and y taps (2048, 0)) and
// image_adjust is u8 HWC -> normalised f32 CHW of one flat pixel array per frame, whatever the row length (FCOS /
// RetinaNet frames are 1333 wide: their rows are not 16-byte aligned, the planes are). Thread = 4 consecutive pixels =
// three aligned 4-byte loads (12 B stride across the lanes), one 16-byte store per plane (512 B contiguous per warp).
#ifndef TCB_FLAT_UNROLL
#define TCB_FLAT_UNROLL 4
#endif
constexpr int FLAT_UNROLL = TCB_FLAT_UNROLL;
template <bool EXACT, bool HAS_SUB>
__global__ void __launch_bounds__(256) preprocess_flat_kernel(const PreParams p) {
    pdl_enter();
    const long long plane = (long long)p.out_h * p.out_w, groups = plane >> 2;     // groups of 4 pixels (plane % 4 == 0)
    const unsigned* src = reinterpret_cast<const unsigned*>(p.src + (long long)blockIdx.y * p.img_stride);
    float* dst = p.dst + (long long)blockIdx.y * 3 * plane;
    const long long g0 = ((long long)blockIdx.x * FLAT_UNROLL) * 256 + threadIdx.x;
    unsigned w[FLAT_UNROLL][3];
#pragma unroll
    for (int u = 0; u < FLAT_UNROLL; u++) {
        const long long g = g0 + u * 256;
        if (g < groups) { w[u][0] = __ldg(src + 3 * g); w[u][1] = __ldg(src + 3 * g + 1); w[u][2] = __ldg(src + 3 * g + 2); }   // the three loads of a warp share their lines: through L1
    }
#pragma unroll
    for (int u = 0; u < FLAT_UNROLL; u++) {
        const long long g = g0 + u * 256;
        if (g >= groups) break;
        float o[3][4];
#pragma unroll
        for (int e = 0; e < 12; e++) {                                // byte e = pixel e/3, channel e%3
            const unsigned v = __byte_perm(w[u][e >> 2], 0u, 0x4440u | (unsigned)(e & 3));
            const int c = e % 3, k = e / 3;
            o[c][k] = norm2<EXACT, HAS_SUB>((int)v, p.div, p.rdiv, p.rlo, p.sub[p.swap ? 2 - c : c]);
        }
#pragma unroll
        for (int c = 0; c < 3; c++) st_stream_f4(dst + (p.swap ? 2 - c : c) * plane + 4 * g, make_float4(o[c][0], o[c][1], o[c][2], o[c][3]));
    }
}

template <bool AREA, bool EXACT, bool HAS_SUB>
__global__ void __launch_bounds__(NTHREADS) preprocess2_kernel(const PreParams p, const TileGeo geo) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int tid = threadIdx.x;
    pdl_enter();
    const int z = blockIdx.z;
    const int tx0 = blockIdx.x * TW, ty0 = blockIdx.y * p.th;
    const uint8_t* img = p.src + (long long)z * p.img_stride;
    // shared layout: staged bytes | hs (u16, 3 planes x nr_cap rows x TW) | x taps | y taps | mbarrier
    uint8_t* stg = smem;
    unsigned short* hs = reinterpret_cast<unsigned short*>(smem + p.stage_bytes);
    XTap* xt = reinterpret_cast<XTap*>(smem + p.stage_bytes + p.hs_bytes);
    YTap* yt = reinterpret_cast<YTap*>(xt + TW);
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(yt + 32);
    const int nr_cap = p.nr_cap;

    // ---- tile geometry from the host table (no arithmetic here) ----------------------------------------------------
    TileHdr H;
    H.xmin = geo.xmin[blockIdx.x]; H.rmin = geo.rmin[blockIdx.y];
    const int xcnt = geo.xcnt[blockIdx.x], rcnt = geo.rcnt[blockIdx.y];
    H.has_content = xcnt > 0 && rcnt > 0;
    H.xmax = H.xmin + xcnt - 1; H.rmax = H.rmin + rcnt - 1;
    H.interior = H.has_content && (tx0 - p.left >= 0) && (tx0 + TW - 1 - p.left < p.new_w) && (tx0 + TW <= p.out_w) &&
                 (ty0 - p.top >= 0) && (ty0 + p.th - 1 - p.top < p.new_h) && (ty0 + p.th <= p.out_h);
    // 16-byte-granular copies can only leave the caller's buffer before its first / after its last content row
    H.safe = !((z == 0 && (int)blockIdx.y == geo.first_ty) || (z == (int)gridDim.z - 1 && (int)blockIdx.y == geo.last_ty));
    const int nrows = H.has_content ? rcnt : 0;
    const int seg = xcnt * 3;

    // ---- stage the footprint: one bulk async copy (TMA unit) per source row, completion on an mbarrier ----------------
    if (tid == 0 && H.has_content && H.safe) mbar_init(bar, nrows);
    __syncthreads();
    if (H.has_content) {
        if (H.safe) {
            for (int r = tid; r < nrows; r += NTHREADS) {   // one arrival + one bulk copy per row (barrier count == nrows)
                const uint8_t* a = img + (long long)(H.rmin + r) * p.row_stride + (long long)H.xmin * 3;
                const uint8_t* g = (const uint8_t*)((uintptr_t)a & ~(uintptr_t)15);
                const unsigned bytes = (unsigned)(((a + seg) - g + 15) & ~(long long)15);
                mbar_arrive_expect_tx(bar, bytes);
                bulk_g2s(stg + r * p.pitch, g, bytes, bar);
            }
        } else {   // tiles touching the ends of the caller's buffer: chunk by chunk, never reading outside it
            const int k = tid & ((1 << p.tpr_shift) - 1);
            const int rstep = NTHREADS >> p.tpr_shift;
            if (k < (p.pitch >> 4)) {
                for (int r = tid >> p.tpr_shift; r < nrows; r += rstep) {
                    const uint8_t* a = img + (long long)(H.rmin + r) * p.row_stride + (long long)H.xmin * 3;
                    const uint8_t* g = (const uint8_t*)((uintptr_t)a & ~(uintptr_t)15) + 16 * k;
                    if (g >= a + seg) continue;
                    uint8_t t[16];
#pragma unroll
                    for (int i = 0; i < 16; i++) t[i] = (g + i >= p.src && g + i < p.batch_end) ? g[i] : 0;
                    *reinterpret_cast<uint4*>(stg + r * p.pitch + 16 * k) = *reinterpret_cast<uint4*>(t);
                }
            }
        }
    }
    // ---- tap tables (overlap the copies) ---------------------------------------------------------------------------
    if (tid >= NTHREADS - TW) {
        const int c = tid - (NTHREADS - TW);
        const int dx = tx0 + c - p.left;
        XTap e; e.off = -1; e.a0 = 0; e.a1 = 0;
        if (H.has_content && dx >= 0 && dx < p.new_w && tx0 + c < p.out_w) {
            if (AREA) { e.off = (2 * dx - H.xmin) * 3; e.a0 = 1; e.a1 = 1; }
            else {
                int s, a0, a1;
                x_tap(dx, p.scale_x, p.w, s, a0, a1);
                e.off = (s - H.xmin) * 3; e.a0 = (short)a0; e.a1 = (short)a1;   // s == w-1 has a1 == 0: the right tap may read padding
            }
        }
        xt[c] = e;
    } else if (tid >= NTHREADS - TW - p.th) {
        const int j = tid - (NTHREADS - TW - p.th);
        const int dy = ty0 + j - p.top;
        YTap e; e.r0 = -1; e.r1 = -1; e.b0s = 0; e.b1s = 0;
        if (H.has_content && dy >= 0 && dy < p.new_h && ty0 + j < p.out_h) {
            if (AREA) { e.r0 = 2 * dy - H.rmin; e.r1 = 2 * dy + 1 - H.rmin; }
            else {
                int r0, r1, b0, b1;
                y_tap(dy, p.scale_y, p.h, r0, r1, b0, b1);
                e.r0 = r0 - H.rmin; e.r1 = r1 - H.rmin; e.b0s = (unsigned)b0 << 16; e.b1s = (unsigned)b1 << 16;
            }
        }
        yt[j] = e;
    }
    if (H.has_content && H.safe && tid < 32) mbar_wait(bar, 0);   // one warp polls; the rest sleep in the barrier below
    __syncthreads();

    if (p.fast_area == 2) {   // width unchanged
        const unsigned shift = (unsigned)((uintptr_t)(img + (long long)H.rmin * p.row_stride + (long long)H.xmin * 3) & 15u);
        if (H.interior) pass_ident<EXACT, HAS_SUB, true>(p, stg, xt, yt, shift, z, tx0, ty0);
        else pass_ident<EXACT, HAS_SUB, false>(p, stg, xt, yt, shift, z, tx0, ty0);
        return;
    }
    if (p.fast_area) {   // AREA: 2x2 box mean; otherwise x exactly 2:1 with bilinear y
        const unsigned shift = (unsigned)((uintptr_t)(img + (long long)H.rmin * p.row_stride + (long long)H.xmin * 3) & 15u);
        if (H.interior) pass_area<AREA, EXACT, HAS_SUB, true>(p, stg, xt, yt, shift, z, tx0, ty0);
        else pass_area<AREA, EXACT, HAS_SUB, false>(p, stg, xt, yt, shift, z, tx0, ty0);
        return;
    }
    // ---- pass H: thread = (column, row parity); 6 source bytes per row come in as three aligned words -----------------
    {
        const int col = tid & (TW - 1), par = tid >> 7;  // NTHREADS == 2*TW
        const XTap e = xt[col];
        if (e.off >= 0) {
            const unsigned wts = (unsigned)(unsigned short)e.a0 | ((unsigned)(unsigned short)e.a1 << 16);
            unsigned addr0 = (unsigned)((uintptr_t)(img + (long long)(H.rmin + par) * p.row_stride + (long long)H.xmin * 3) & 15u);
            const unsigned astep = (unsigned)((2 * p.row_stride) & 15);
            unsigned short* d = hs + par * TW + col;
            const int plane = nr_cap * TW;
            if (astep == 0 && (p.row_stride & 15) == 0) {
                // every staged row has the same alignment shift: constant word offset and funnel distance
                const unsigned b = addr0 + (unsigned)e.off;
                const unsigned* w = reinterpret_cast<const unsigned*>(stg + par * p.pitch + (b & ~3u));
                const unsigned sh = (b & 3u) * 8u;
                const int wstep = (2 * p.pitch) >> 2;
                for (int r = par; r < nrows; r += 2, w += wstep, d += 2 * TW) {
                    const unsigned x0 = __funnelshift_r(w[0], w[1], sh), x1 = __funnelshift_r(w[1], w[2], sh);
                    const unsigned p0 = __byte_perm(x0, x1, 0x4430), p1 = __byte_perm(x0, x1, 0x4441), p2 = __byte_perm(x0, x1, 0x4452);
                    if (AREA) {
                        d[0] = (unsigned short)((p0 & 0xff) + ((p0 >> 8) & 0xff));
                        d[plane] = (unsigned short)((p1 & 0xff) + ((p1 >> 8) & 0xff));
                        d[2 * plane] = (unsigned short)((p2 & 0xff) + ((p2 >> 8) & 0xff));
                    } else {
                        d[0] = (unsigned short)(__dp2a_lo(wts, p0, 0u) >> 4);
                        d[plane] = (unsigned short)(__dp2a_lo(wts, p1, 0u) >> 4);
                        d[2 * plane] = (unsigned short)(__dp2a_lo(wts, p2, 0u) >> 4);
                    }
                }
            } else {
                for (int r = par; r < nrows; r += 2, addr0 = (addr0 + astep) & 15u, d += 2 * TW) {
                    const unsigned b = (unsigned)(r * p.pitch) + addr0 + (unsigned)e.off;   // byte offset of the left tap in the staged area
                    const unsigned* w = reinterpret_cast<const unsigned*>(stg + (b & ~3u));
                    const unsigned sh = (b & 3u) * 8u;
                    const unsigned x0 = __funnelshift_r(w[0], w[1], sh), x1 = __funnelshift_r(w[1], w[2], sh);   // bytes b..b+3, b+4..b+7
                    // (left, right) taps of channel c are bytes c and c+3
                    const unsigned p0 = __byte_perm(x0, x1, 0x4430), p1 = __byte_perm(x0, x1, 0x4441), p2 = __byte_perm(x0, x1, 0x4452);
                    if (AREA) {
                        d[0] = (unsigned short)((p0 & 0xff) + ((p0 >> 8) & 0xff));
                        d[plane] = (unsigned short)((p1 & 0xff) + ((p1 >> 8) & 0xff));
                        d[2 * plane] = (unsigned short)((p2 & 0xff) + ((p2 >> 8) & 0xff));
                    } else {
                        d[0] = (unsigned short)(__dp2a_lo(wts, p0, 0u) >> 4);
                        d[plane] = (unsigned short)(__dp2a_lo(wts, p1, 0u) >> 4);
                        d[2 * plane] = (unsigned short)(__dp2a_lo(wts, p2, 0u) >> 4);
                    }
                }
            }
        }
    }
    __syncthreads();

    // ---- pass V: warp = output rows, lane = 4 columns ---------------------------------------------------------------
    if (H.interior) pass_v<AREA, EXACT, HAS_SUB, true>(p, hs, xt, yt, nr_cap, z, tx0, ty0);
    else pass_v<AREA, EXACT, HAS_SUB, false>(p, hs, xt, yt, nr_cap, z, tx0, ty0);
}

int iround_half_even(double v) { return (int)nearbyint(v); }  // python round() == ties-to-even

}  // namespace

// where the resampled frame lands in the (out_h, out_w) output and the gains of the mapping, per mode
static void content_geometry(int mode, int h, int w, int out_h, int out_w, int& left, int& top, int& new_w, int& new_h, double& gain_x, double& gain_y) {
    if (mode == TCB_RESIZE_STRETCH) {
        left = 0; top = 0; new_w = out_w; new_h = out_h;
        gain_x = (double)out_w / w; gain_y = (double)out_h / h;
    } else if (mode == TCB_RESIZE_LETTERBOX) {
        // ultralytics letterbox: r=min(S/h,S/w); new=round(w r),round(h r); pads split with round(+-0.1)
        double r = fmin((double)out_h / h, (double)out_w / w);
        new_w = iround_half_even(w * r); new_h = iround_half_even(h * r);
        if (new_w < 1) new_w = 1;
        if (new_h < 1) new_h = 1;
        if (new_w > out_w) new_w = out_w;
        if (new_h > out_h) new_h = out_h;
        double dw = (out_w - new_w) / 2.0, dh = (out_h - new_h) / 2.0;
        left = iround_half_even(dw - 0.1); top = iround_half_even(dh - 0.1);
        gain_x = r; gain_y = r;
    } else {
        left = 0; top = 0; new_w = w; new_h = h;
        gain_x = 1.0; gain_y = 1.0;
    }
}

extern "C" int tcb_content_rect(int mode, int h, int w, int out_h, int out_w, int rect[4]) {
    if (!rect || h < 1 || w < 1 || out_h < 1 || out_w < 1) return tcb_fail(nullptr, TCB_ERR_ARG, "tcb_content_rect: bad arguments");
    if (mode != TCB_RESIZE_STRETCH && mode != TCB_RESIZE_LETTERBOX && mode != TCB_RESIZE_PAD) return tcb_fail(nullptr, TCB_ERR_ARG, "tcb_content_rect: unknown mode %d", mode);
    if (mode == TCB_RESIZE_PAD && (h > out_h || w > out_w)) return tcb_fail(nullptr, TCB_ERR_ARG, "tcb_content_rect: PAD mode needs src <= out");
    double gx, gy;
    content_geometry(mode, h, w, out_h, out_w, rect[0], rect[1], rect[2], rect[3], gx, gy);
    return TCB_OK;
}

extern "C" int tcb_copy_rect_d2h(tcb_ctx* ctx, const void* src_dev, void* dst_host, int planes, int plane_h, int plane_w, int elem_bytes,
                                 int left, int top, int rect_w, int rect_h, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, planes >= 0 && plane_h >= 1 && plane_w >= 1 && elem_bytes >= 1, "tcb_copy_rect_d2h: bad plane shape");
    TCB_REQUIRE(ctx, left >= 0 && top >= 0 && rect_w >= 0 && rect_h >= 0 && left + rect_w <= plane_w && top + rect_h <= plane_h, "tcb_copy_rect_d2h: rectangle outside the plane");
    if (planes == 0 || rect_w == 0 || rect_h == 0) return TCB_OK;
    TCB_REQUIRE(ctx, src_dev && dst_host, "tcb_copy_rect_d2h: null buffer");
    const size_t pitch = (size_t)plane_w * elem_bytes;
    if (rect_w == plane_w) {
        // whole rows: one run of rect_h rows per plane -> a pitched 2-D copy (long contiguous runs for the copy engine)
        const size_t off = (size_t)top * pitch, run = (size_t)rect_h * pitch, ppitch = (size_t)plane_h * pitch;
        TCB_CUDA(ctx, cudaMemcpy2DAsync((char*)dst_host + off, ppitch, (const char*)src_dev + off, ppitch, run, (size_t)planes, cudaMemcpyDeviceToHost,
                                        (cudaStream_t)stream));
        return TCB_OK;
    }
    cudaMemcpy3DParms q;
    memset(&q, 0, sizeof(q));
    q.srcPtr = make_cudaPitchedPtr(const_cast<void*>(src_dev), pitch, (size_t)plane_w * elem_bytes, (size_t)plane_h);
    q.dstPtr = make_cudaPitchedPtr(dst_host, pitch, (size_t)plane_w * elem_bytes, (size_t)plane_h);
    q.srcPos = make_cudaPos((size_t)left * elem_bytes, (size_t)top, 0);
    q.dstPos = q.srcPos;
    q.extent = make_cudaExtent((size_t)rect_w * elem_bytes, (size_t)rect_h, (size_t)planes);
    q.kind = cudaMemcpyDeviceToHost;
    TCB_CUDA(ctx, cudaMemcpy3DAsync(&q, (cudaStream_t)stream));
    return TCB_OK;
}

static const size_t SMEM_MAX = 100 * 1024, SMEM_GOOD = 36 * 1024;   // tile budgets of tcb_preprocess_image
typedef void (*kern1_t)(const PreParams);
typedef void (*kern2_t)(const PreParams, const TileGeo);
static const kern1_t one_pass[2][2] = {{preprocess_kernel<false, false, false>, preprocess_kernel<false, false, true>},
                                       {preprocess_kernel<true, false, false>, preprocess_kernel<true, false, true>}};
static const kern2_t two_pass[2][2][2] = {
    {{preprocess2_kernel<false, false, false>, preprocess2_kernel<false, false, true>},
     {preprocess2_kernel<false, true, false>, preprocess2_kernel<false, true, true>}},
    {{preprocess2_kernel<true, false, false>, preprocess2_kernel<true, false, true>},
     {preprocess2_kernel<true, true, false>, preprocess2_kernel<true, true, true>}}};

static int init_preprocess(tcb_ctx* ctx) {   // run by tcb_create
    for (int i = 0; i < 8; i++) TCB_SMEM_OPT_IN(ctx, two_pass[i >> 2][(i >> 1) & 1][i & 1], SMEM_MAX);
    return TCB_OK;
}
static TcbInitReg reg_preprocess(&init_preprocess);

extern "C" int tcb_preprocess_image(tcb_ctx* ctx, const uint8_t* src_dev, int n, int h, int w,
                                    long long src_row_stride, long long src_img_stride, int channel_swap,
                                    int mode, int out_h, int out_w, float div, const float sub[3],
                                    int pad_value, float* dst_dev, float* inv_affine_host, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, src_dev && dst_dev, "tcb_preprocess_image: null buffer");
    TCB_REQUIRE(ctx, n >= 0 && h > 0 && w > 0 && out_h > 0 && out_w > 0, "tcb_preprocess_image: bad shape n=%d h=%d w=%d out=%dx%d", n, h, w, out_h, out_w);
    TCB_REQUIRE(ctx, n <= 65535, "tcb_preprocess_image: at most 65535 frames per call (got %d)", n);
    TCB_REQUIRE(ctx, src_row_stride >= (long long)w * 3, "tcb_preprocess_image: row stride %lld < w*3", src_row_stride);
    TCB_REQUIRE(ctx, src_img_stride >= src_row_stride * (h - 1) + (long long)w * 3 || n <= 1, "tcb_preprocess_image: frame stride %lld too small", src_img_stride);
    TCB_REQUIRE(ctx, mode == TCB_RESIZE_STRETCH || mode == TCB_RESIZE_LETTERBOX || mode == TCB_RESIZE_PAD, "tcb_preprocess_image: unknown mode %d", mode);
    TCB_REQUIRE(ctx, div != 0.f && div == div, "tcb_preprocess_image: div must be a non-zero number");
    TCB_REQUIRE(ctx, pad_value >= 0 && pad_value <= 255, "tcb_preprocess_image: pad_value %d not a byte", pad_value);

    PreParams p;
    memset(&p, 0, sizeof(p));
    p.src = src_dev;
    p.batch_end = src_dev + (long long)(n > 0 ? n - 1 : 0) * src_img_stride + (long long)(h - 1) * src_row_stride + (long long)w * 3;
    p.img_stride = src_img_stride; p.row_stride = src_row_stride;
    p.h = h; p.w = w; p.dst = dst_dev; p.out_h = out_h; p.out_w = out_w;
    p.swap = channel_swap ? 1 : 0;

    double gain_x, gain_y;
    if (mode == TCB_RESIZE_PAD) TCB_REQUIRE(ctx, h <= out_h && w <= out_w, "tcb_preprocess_image: PAD mode needs src (%dx%d) <= out (%dx%d)", h, w, out_h, out_w);
    content_geometry(mode, h, w, out_h, out_w, p.left, p.top, p.new_w, p.new_h, gain_x, gain_y);
    if (inv_affine_host) {
        inv_affine_host[0] = (float)gain_x; inv_affine_host[1] = (float)gain_y;
        inv_affine_host[2] = (float)p.left; inv_affine_host[3] = (float)p.top;
    }
    if (n == 0) return TCB_OK;

    // cv2: inv_scale = dsize/ssize (double); scale = 1/inv_scale
    p.scale_x = 1.0 / ((double)p.new_w / (double)w);
    p.scale_y = 1.0 / ((double)p.new_h / (double)h);
    const bool area = (w == 2 * p.new_w) && (h == 2 * p.new_h);

    p.div = div; p.rdiv = 1.0f / div;
    p.rlo = (float)(1.0 / (double)div - (double)p.rdiv);
    if (!ctx->pre_div_known || ctx->pre_div != div) {      // 2 x 256 checks: once per divisor, not once per frame
        int e1 = 1, e2 = 1;
        for (int v = 0; v < 256; v++) {
            float f = (float)v, q0 = f * p.rdiv;
            float q = fmaf(fmaf(-q0, div, f), p.rdiv, q0);
            if (q != f / div) { e1 = 0; break; }
        }
        for (int v = 0; v < 256; v++) {
            const float f = (float)v;
            if (fmaf(f, p.rdiv, f * p.rlo) != f / div) { e2 = 0; break; }
        }
        ctx->pre_div = div; ctx->pre_exact_recip = e1; ctx->pre_exact2 = e2; ctx->pre_div_known = 1;
    }
    p.exact_recip = ctx->pre_exact_recip; p.exact2 = ctx->pre_exact2;
    for (int c = 0; c < 3; c++) {
        p.sub[c] = sub ? sub[c] : 0.f;
        p.padv[c] = (float)pad_value / div - p.sub[c];
    }
    p.vec_store = (out_w % 4 == 0) && (((uintptr_t)dst_dev & 15) == 0);

    const bool has_sub_any = p.sub[0] != 0.f || p.sub[1] != 0.f || p.sub[2] != 0.f;
    if (h == out_h && w == out_w && p.new_h == h && p.new_w == w && p.left == 0 && p.top == 0 && src_row_stride == (long long)w * 3 &&
        ((long long)h * w) % 4 == 0 && (((uintptr_t)src_dev) & 3) == 0 && (n == 1 || src_img_stride % 4 == 0) && (((uintptr_t)dst_dev) & 15) == 0) {
        // same size, packed rows: a flat stream per frame (preprocess_flat_kernel)
        typedef void (*kernf_t)(const PreParams);
        static const kernf_t flat[2][2] = {{preprocess_flat_kernel<false, false>, preprocess_flat_kernel<false, true>},
                                           {preprocess_flat_kernel<true, false>, preprocess_flat_kernel<true, true>}};
        const long long groups = ((long long)h * w) >> 2;
        TCB_CUDA(ctx, tcb_launch(ctx->pdl != 0, flat[p.exact2 ? 1 : 0][has_sub_any ? 1 : 0], dim3(tcb_div_up(groups, 256 * FLAT_UNROLL), n), dim3(256), 0,
                                 (cudaStream_t)stream, p));
        TCB_LAUNCHED(ctx);
        return TCB_OK;
    }

    // Tile height: the largest of 32/16/8 output rows whose shared-memory need (staged source bytes + 16-bit row
    // buffer + tap tables) stays <= 36 KB, so that six CTAs fit an SM; failing that the largest that fits 100 KB;
    // failing that (extreme down-scaling) the one-pass kernel that reads the source straight from global memory.
    const bool half_x = (w == 2 * p.new_w);   // x taps are then (1024, 1024) on columns 2dx, 2dx+1 for every dx
    const bool same_x = (w == p.new_w);       // x taps (2048, 0) on column dx
    p.fast_area = (area || half_x || same_x) && p.vec_store && (((uintptr_t)src_dev & 15) == 0) && (src_row_stride % 16 == 0) &&
                  (n == 1 || src_img_stride % 16 == 0) && (p.left % 4 == 0) && (p.new_w % 4 == 0);
    if (p.fast_area && same_x && !area && !half_x) p.fast_area = 2;
    bool staged = false;
    size_t smem = 0;
    const double sx = area ? 2.0 : p.scale_x, sy = area ? 2.0 : p.scale_y;
    for (int pass = 0; pass < 2 && !staged; pass++) {
        for (int th = 32; th >= 8; th >>= 1) {
            long long cols = (long long)floor((TW - 1) * sx) + 4;
            if (cols > w) cols = w;
            long long rows = (long long)floor((th - 1) * sy) + 4;
            if (rows > h) rows = h;
            const long long pitch = ((cols * 3 + 15) / 16) * 16 + 32;   // +16 alignment shift, +16 so a clamped right tap stays inside
            const size_t stage = (size_t)(rows * pitch), hsb = p.fast_area ? 0 : (size_t)rows * TW * 3 * 2;
            const size_t need = stage + hsb + TW * sizeof(XTap) + 32 * sizeof(YTap) + 16;   // + the mbarrier
            if (need <= (pass == 0 ? SMEM_GOOD : SMEM_MAX)) {
                staged = true; smem = need; p.th = th; p.pitch = (int)pitch; p.stage_bytes = (int)stage; p.hs_bytes = (int)hsb; p.nr_cap = (int)rows;
                break;
            }
        }
    }
    if (!staged) { p.th = 8; p.pitch = 16; smem = 0; }
    p.tpr_shift = 0;
    while ((1 << p.tpr_shift) < (p.pitch >> 4) && p.tpr_shift < 8) p.tpr_shift++;

    dim3 grid(tcb_div_up(out_w, TW), tcb_div_up(out_h, p.th), n);
    TCB_REQUIRE(ctx, grid.y <= 65535, "tcb_preprocess_image: output too tall (%d rows)", out_h);
    cudaStream_t st = (cudaStream_t)stream;
    if (staged && ((int)grid.x > GEO_MAX_TX || (int)grid.y > GEO_MAX_TY)) {   // geometry table too small: one-pass kernel
        staged = false; p.th = 8; smem = 0;
        grid = dim3(tcb_div_up(out_w, TW), tcb_div_up(out_h, p.th), n);
        TCB_REQUIRE(ctx, grid.y <= 65535, "tcb_preprocess_image: output too tall (%d rows)", out_h);
    }
    const bool has_sub = p.sub[0] != 0.f || p.sub[1] != 0.f || p.sub[2] != 0.f;
    if (staged) {
        // source footprint of every tile column / row, with the kernel's own tap arithmetic
        static thread_local TileGeo geo;
        geo.first_ty = -1; geo.last_ty = -1;
        for (int tx = 0; tx < (int)grid.x; tx++) {
            const int t0 = tx * TW, t1 = (t0 + TW < out_w ? t0 + TW : out_w) - 1;
            const int c0 = t0 - p.left > 0 ? t0 - p.left : 0, c1 = t1 - p.left < p.new_w - 1 ? t1 - p.left : p.new_w - 1;
            geo.xmin[tx] = 0; geo.xcnt[tx] = 0;
            if (c0 <= c1) {
                int lo, hi;
                if (area) { lo = 2 * c0; hi = 2 * c1 + 1; }
                else {
                    int s0, s1, a, b;
                    x_tap(c0, p.scale_x, w, s0, a, b);
                    x_tap(c1, p.scale_x, w, s1, a, b);
                    lo = s0; hi = s1 + 1 < w - 1 ? s1 + 1 : w - 1;
                }
                geo.xmin[tx] = lo; geo.xcnt[tx] = (short)(hi - lo + 1);
            }
        }
        for (int ty = 0; ty < (int)grid.y; ty++) {
            const int t0 = ty * p.th, t1 = (t0 + p.th < out_h ? t0 + p.th : out_h) - 1;
            const int c0 = t0 - p.top > 0 ? t0 - p.top : 0, c1 = t1 - p.top < p.new_h - 1 ? t1 - p.top : p.new_h - 1;
            geo.rmin[ty] = 0; geo.rcnt[ty] = 0;
            if (c0 <= c1) {
                int lo, hi;
                if (area) { lo = 2 * c0; hi = 2 * c1 + 1; }
                else {
                    int r0, r1, a, b;
                    y_tap(c0, p.scale_y, h, r0, r1, a, b); lo = r0;
                    y_tap(c1, p.scale_y, h, r0, r1, a, b); hi = r1;
                }
                geo.rmin[ty] = lo; geo.rcnt[ty] = (short)(hi - lo + 1);
                if (geo.first_ty < 0) geo.first_ty = ty;
                geo.last_ty = ty;
            }
        }
        kern2_t kern = two_pass[area ? 1 : 0][p.exact2 ? 1 : 0][has_sub ? 1 : 0];
        TCB_CUDA(ctx, tcb_launch(ctx->pdl != 0, kern, grid, dim3(NTHREADS), smem, st, p, geo));
    } else {
        kern1_t kern = one_pass[area ? 1 : 0][p.exact_recip ? 1 : 0];
        kern<<<grid, NTHREADS, smem, st>>>(p);
    }
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}
