// YOLOv5 postprocess on device: candidate filter + conf = obj*cls (f64) +
// best class + xywh->xyxy (f64 -> f32) in one pass over the [B,N,5+nc] tensor, then one CTA per image:
// exact (score desc, row asc) ordering and greedy class-aware NMS in float64 with the arithmetic of
// torchvision.ops.nms, stopping at max_det. Nothing goes back to the host but the <= max_det rows.
//
// Follows clients/postprocess/yolov5_postprocess.py:28-125 (twin utils/postprocess.py:105-199):
//   :41 xc = obj > conf_thres           :76 cls *= obj (Python floats = f64)
//   :79 xywh2xyxy(...).astype(float32)   :87 conf, j = max(1) (first index on ties)
//   :88 keep conf > conf_thres           :91-92 class whitelist
//   :102-103 top max_nms by score        :106-110 boxes + cls*4096 -> torchvision nms -> [:max_det]
#include "common.cuh"

namespace {

constexpr int NMS_THREADS = 1024;
constexpr int CHUNK = 256;          // candidates resolved per greedy round
constexpr int SORT_SMEM_CAP = 4096; // entries sorted in shared memory; larger sets sort in global scratch
constexpr int MAX_DET_CAP = 1024;

struct SortEnt {
    unsigned long long key;  // f64 score bits (scores are > conf_thres >= 0, so bit order == numeric order)
    unsigned row;            // row index inside the image: the deterministic tie-break (lower first)
    unsigned slot;           // position in the unordered candidate arrays
};

struct ClassMask { unsigned long long w[4]; };  // whitelist for classes 0..255

// ---------------------------------------------------------------------------------------------
// Pass 1: filter + decode. One warp scans 32 rows at a time (one objectness load per lane), then the
// whole warp cooperates on each surviving row (coalesced class-score loads, shuffle arg-max).
// ---------------------------------------------------------------------------------------------
constexpr int FILTER_UNROLL = 4;   // 32-row groups in flight per warp (independent objectness loads)
constexpr int FILTER_SMALL_NC = 8;  // up to this many classes a surviving lane decodes its own row

__device__ __forceinline__ void emit_candidate(const float* __restrict__ rp, int r, double best, int bidx, long long k,
                                               double* __restrict__ c_score, float4* __restrict__ c_box, int2* __restrict__ c_meta) {
    const double x = (double)__ldg(rp), y = (double)__ldg(rp + 1);
    const double w = (double)__ldg(rp + 2), h = (double)__ldg(rp + 3);
    float4 bx;
    bx.x = (float)(x - w / 2); bx.y = (float)(y - h / 2);
    bx.z = (float)(x + w / 2); bx.w = (float)(y + h / 2);
    c_score[k] = best; c_box[k] = bx; c_meta[k] = make_int2(r, bidx);
}

__global__ void __launch_bounds__(256) yolo_filter_kernel(
    const float* __restrict__ pred, int B, int N, int nc, double conf_thres, int use_filter, ClassMask cmask,
    double* __restrict__ c_score, float4* __restrict__ c_box, int2* __restrict__ c_meta,
    int* __restrict__ c_count, int* __restrict__ xc_count) {
    const int lane = threadIdx.x & 31;
    const long long warp_global = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int stride = 5 + nc;
    const long long groups_per_img = (N + 31) / 32;
    const long long total_groups = groups_per_img * B;

    for (long long g0 = warp_global * FILTER_UNROLL; g0 < total_groups; g0 += nwarps * FILTER_UNROLL) {
        float obj[FILTER_UNROLL];
#pragma unroll
        for (int u = 0; u < FILTER_UNROLL; u++) {  // all objectness loads first
            const long long g = g0 + u;
            obj[u] = -1.f;
            if (g < total_groups) {
                const int b = (int)(g / groups_per_img);
                const int row = (int)(g - (long long)b * groups_per_img) * 32 + lane;
                if (row < N) obj[u] = __ldg(pred + ((long long)b * N + row) * stride + 4);
            }
        }
#pragma unroll
        for (int u = 0; u < FILTER_UNROLL; u++) {
            const long long g = g0 + u;
            if (g >= total_groups) break;
            const int b = (int)(g / groups_per_img);
            const int row0 = (int)(g - (long long)b * groups_per_img) * 32;
            const float* base = pred + ((long long)b * N) * stride;
            const bool pass = (double)obj[u] > conf_thres;  // padding lanes hold -1 and conf_thres >= 0
            unsigned m = __ballot_sync(0xffffffffu, pass);
            if (m == 0) continue;
            if (lane == 0) atomicAdd(&xc_count[b], __popc(m));
            if (nc <= FILTER_SMALL_NC) {
                // few classes: each surviving lane decodes its own row (the lines are already in L1)
                double best = -1.0;
                int bidx = 0;
                bool ok = false;
                const float* rp = base + (long long)(row0 + lane) * stride;
                if (pass) {
                    const double o = (double)obj[u];
                    bool nan = false;
                    for (int c = 0; c < nc; c++) {
                        const double v = (double)__ldg(rp + 5 + c) * o;
                        nan |= (v != v);
                        if (v > best) { best = v; bidx = c; }  // strict: first index wins
                    }
                    ok = !nan && nc > 0 && best > conf_thres;
                    if (ok && use_filter) ok = (cmask.w[bidx >> 6] >> (bidx & 63)) & 1ull;
                }
                const unsigned em = __ballot_sync(0xffffffffu, ok);
                if (em) {
                    int basei = 0;
                    if (lane == 0) basei = atomicAdd(&c_count[b], __popc(em));
                    basei = __shfl_sync(0xffffffffu, basei, 0);
                    if (ok) emit_candidate(rp, row0 + lane, best, bidx, (long long)b * N + basei + __popc(em & ((1u << lane) - 1)),
                                           c_score, c_box, c_meta);
                }
            } else {
                // many classes: the warp cooperates on each surviving row (coalesced loads, shuffle arg-max)
                while (m) {
                    const int src_lane = __ffs(m) - 1;
                    m &= m - 1;
                    const int r = row0 + src_lane;
                    const double o = (double)__shfl_sync(0xffffffffu, obj[u], src_lane);
                    const float* rp = base + (long long)r * stride;
                    double best = -1.0;
                    int bidx = 0x7fffffff;
                    bool nan = false;
                    for (int c = lane; c < nc; c += 32) {
                        const double v = (double)__ldg(rp + 5 + c) * o;
                        nan |= (v != v);
                        if (v > best) { best = v; bidx = c; }  // strict: first index wins inside a lane
                    }
#pragma unroll
                    for (int s = 16; s > 0; s >>= 1) {
                        const double ob = __shfl_xor_sync(0xffffffffu, best, s);
                        const int oi = __shfl_xor_sync(0xffffffffu, bidx, s);
                        if (ob > best || (ob == best && oi < bidx)) { best = ob; bidx = oi; }
                    }
                    nan = __any_sync(0xffffffffu, nan);
                    if (lane == 0 && !nan && best > conf_thres) {
                        bool ok = true;
                        if (use_filter) ok = (bidx < 256) && ((cmask.w[bidx >> 6] >> (bidx & 63)) & 1ull);
                        if (ok) emit_candidate(rp, r, best, bidx, (long long)b * N + atomicAdd(&c_count[b], 1), c_score, c_box, c_meta);
                    }
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Pass 2: per-image ordering + greedy NMS.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool ent_before(const SortEnt& a, const SortEnt& b) {
    return a.key > b.key || (a.key == b.key && a.row < b.row);
}

// Bitonic sort of npad (power of two) entries by the whole CTA; buf may be shared or global memory.
__device__ void bitonic_sort(SortEnt* buf, int npad) {
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < npad; i += blockDim.x) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    SortEnt a = buf[i], b = buf[ixj];
                    const bool up = ((i & k) == 0);
                    if (up ? ent_before(b, a) : ent_before(a, b)) { buf[i] = b; buf[ixj] = a; }
                }
            }
            __syncthreads();
        }
    }
}

// torchvision/csrc/ops/cpu/nms_kernel.cpp arithmetic in double: suppress iff inter/(ai+aj-inter) > thr.
__device__ __forceinline__ bool iou_gt(double ax1, double ay1, double ax2, double ay2, double aarea,
                                       double bx1, double by1, double bx2, double by2, double barea, double thr) {
    const double xx1 = fmax(ax1, bx1), xx2 = fmin(ax2, bx2);
    const double w = fmax(0.0, xx2 - xx1);
    if (w == 0.0) return false;  // inter == 0 -> ovr is 0 (or NaN): never > thr for thr >= 0
    const double yy1 = fmax(ay1, by1), yy2 = fmin(ay2, by2);
    const double h = fmax(0.0, yy2 - yy1);
    if (h == 0.0) return false;
    const double inter = w * h;
    const double ovr = inter / (aarea + barea - inter);
    return ovr > thr;
}

struct NmsShared {
    double kx1[MAX_DET_CAP], ky1[MAX_DET_CAP], kx2[MAX_DET_CAP], ky2[MAX_DET_CAP], karea[MAX_DET_CAP];
    double cx1[CHUNK], cy1[CHUNK], cx2[CHUNK], cy2[CHUNK], carea[CHUNK];
    unsigned long long mask[CHUNK][CHUNK / 64];
    int alive[CHUNK];
    unsigned kslot[MAX_DET_CAP];
    int nkept;
};

__global__ void __launch_bounds__(NMS_THREADS, 1) yolo_nms_kernel(
    int N, const double* __restrict__ c_score, const float4* __restrict__ c_box, const int2* __restrict__ c_meta,
    const int* __restrict__ c_count, SortEnt* __restrict__ gsort, int gsort_stride,
    double iou_thres, double class_offset, int max_det, int max_nms,
    double* __restrict__ out, int* __restrict__ out_count) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    NmsShared& S = *reinterpret_cast<NmsShared*>(smem_raw);
    SortEnt* ssort = reinterpret_cast<SortEnt*>(smem_raw + sizeof(NmsShared));

    const int b = blockIdx.x, tid = threadIdx.x;
    int n = min(c_count[b], N);
    const long long cbase = (long long)b * N;
    if (tid == 0) S.nkept = 0;

    // ---- order candidates: score descending, row ascending ------------------------------------
    int npad = 1;
    while (npad < n) npad <<= 1;
    SortEnt* buf = (npad <= SORT_SMEM_CAP) ? ssort : (gsort + (long long)b * gsort_stride);
    for (int i = tid; i < npad; i += NMS_THREADS) {
        SortEnt e;
        if (i < n) {
            e.key = (unsigned long long)__double_as_longlong(c_score[cbase + i]);
            e.row = (unsigned)c_meta[cbase + i].x;
            e.slot = (unsigned)i;
        } else {
            e.key = 0ull; e.row = 0xffffffffu; e.slot = 0u;
        }
        buf[i] = e;
    }
    __syncthreads();
    bitonic_sort(buf, npad);
    if (n > max_nms) n = max_nms;

    // ---- greedy NMS, CHUNK candidates per round -------------------------------------------------
    double* orow = out + (long long)b * max_det * 6;
    for (int c0 = 0; c0 < n; c0 += CHUNK) {
        const int cn = min(CHUNK, n - c0);
        if (tid < CHUNK) {
            int alive = 0;
            if (tid < cn) {
                const unsigned slot = buf[c0 + tid].slot;
                const float4 bx = c_box[cbase + slot];
                const double off = (double)c_meta[cbase + slot].y * class_offset;
                const double x1 = (double)bx.x + off, y1 = (double)bx.y + off;
                const double x2 = (double)bx.z + off, y2 = (double)bx.w + off;
                S.cx1[tid] = x1; S.cy1[tid] = y1; S.cx2[tid] = x2; S.cy2[tid] = y2;
                S.carea[tid] = (x2 - x1) * (y2 - y1);
                alive = 1;
            }
            S.alive[tid] = alive;
        }
        __syncthreads();
        const int nk = S.nkept;
        const int ci = tid & (CHUNK - 1), part = tid >> 8;  // 4 threads share one candidate
        // (1) against everything kept in earlier rounds
        if (ci < cn) {
            const double x1 = S.cx1[ci], y1 = S.cy1[ci], x2 = S.cx2[ci], y2 = S.cy2[ci], ar = S.carea[ci];
            bool dead = false;
            for (int k = part; k < nk && !dead; k += NMS_THREADS / CHUNK)
                dead = iou_gt(S.kx1[k], S.ky1[k], S.kx2[k], S.ky2[k], S.karea[k], x1, y1, x2, y2, ar, iou_thres);
            if (dead) S.alive[ci] = 0;
        }
        __syncthreads();
        // (2) suppression bitmask inside the chunk: bit j of row i set iff i precedes j and iou(i,j) > thr
        {
            unsigned long long bits = 0ull;
            if (ci < cn && S.alive[ci]) {
                const double x1 = S.cx1[ci], y1 = S.cy1[ci], x2 = S.cx2[ci], y2 = S.cy2[ci], ar = S.carea[ci];
                const int j0 = part * 64;
                for (int jj = 0; jj < 64; jj++) {
                    const int j = j0 + jj;
                    if (j > ci && j < cn && S.alive[j] &&
                        iou_gt(x1, y1, x2, y2, ar, S.cx1[j], S.cy1[j], S.cx2[j], S.cy2[j], S.carea[j], iou_thres))
                        bits |= 1ull << jj;
                }
            }
            S.mask[ci][part] = bits;
        }
        __syncthreads();
        // (3) serial sweep by one warp, hopping from kept box to kept box (shared memory only: the
        //     output rows are written by the whole CTA after the last round)
        if (tid < 32) {
            unsigned long long avail[CHUNK / 64];
#pragma unroll
            for (int w = 0; w < CHUNK / 64; w++) {
                const unsigned lo = __ballot_sync(0xffffffffu, S.alive[w * 64 + tid] != 0);
                const unsigned hi = __ballot_sync(0xffffffffu, S.alive[w * 64 + 32 + tid] != 0);
                avail[w] = ((unsigned long long)hi << 32) | lo;
            }
            int kept = nk;
            while (kept < max_det) {
                int i = -1;
#pragma unroll
                for (int w = CHUNK / 64 - 1; w >= 0; w--)
                    if (avail[w]) i = w * 64 + __ffsll((long long)avail[w]) - 1;
                if (i < 0) break;
#pragma unroll
                for (int w = 0; w < CHUNK / 64; w++) {
                    unsigned long long mrow = S.mask[i][w];
                    if (w == (i >> 6)) mrow |= 1ull << (i & 63);
                    avail[w] &= ~mrow;
                }
                if (tid == 0) {
                    S.kx1[kept] = S.cx1[i]; S.ky1[kept] = S.cy1[i]; S.kx2[kept] = S.cx2[i]; S.ky2[kept] = S.cy2[i];
                    S.karea[kept] = S.carea[i];
                    S.kslot[kept] = buf[c0 + i].slot;
                }
                kept++;
            }
            if (tid == 0) S.nkept = kept;
        }
        __syncthreads();
        if (S.nkept >= max_det) break;
    }
    __syncthreads();
    // ---- write the kept rows: [x1,y1,x2,y2 (f32 box widened), conf (f64), cls] -------------------------
    const int nk_final = S.nkept;
    for (int k = tid; k < nk_final; k += NMS_THREADS) {
        const unsigned slot = S.kslot[k];
        const float4 bx = c_box[cbase + slot];
        double* o = orow + (long long)k * 6;
        o[0] = (double)bx.x; o[1] = (double)bx.y; o[2] = (double)bx.z; o[3] = (double)bx.w;
        o[4] = c_score[cbase + slot];
        o[5] = (double)c_meta[cbase + slot].y;
    }
    if (tid == 0) out_count[b] = S.nkept;
}

size_t yolo_ws_bytes(int b, int n) {
    size_t per = (size_t)b * n;
    size_t bytes = 256 * 8;
    bytes += per * sizeof(double) + 256;
    bytes += per * sizeof(float4) + 256;
    bytes += per * sizeof(int2) + 256;
    bytes += (size_t)b * 2 * sizeof(int) + 256;
    int npad = 1;
    while (npad < n) npad <<= 1;
    if (npad > SORT_SMEM_CAP) bytes += (size_t)b * npad * sizeof(SortEnt) + 256;
    return bytes;
}

}  // namespace

extern "C" size_t tcb_yolo_nms_workspace(int b, int n) { return (b > 0 && n > 0) ? yolo_ws_bytes(b, n) : 0; }

extern "C" int tcb_yolo_nms(tcb_ctx* ctx, const float* pred_dev, int b, int n, int nc, double conf_thres,
                            double iou_thres, const int* classes_host, int n_classes, int agnostic, int max_det,
                            int max_nms, double* out_dev, int* count_dev, int* xc_count_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, pred_dev && out_dev && count_dev, "tcb_yolo_nms: null buffer");
    TCB_REQUIRE(ctx, b >= 0 && n >= 0 && nc >= 0, "tcb_yolo_nms: bad shape b=%d n=%d nc=%d", b, n, nc);
    TCB_REQUIRE(ctx, conf_thres >= 0.0 && conf_thres <= 1.0, "tcb_yolo_nms: Invalid Confidence threshold %g, valid values are between 0.0 and 1.0", conf_thres);
    TCB_REQUIRE(ctx, iou_thres >= 0.0 && iou_thres <= 1.0, "tcb_yolo_nms: Invalid IoU %g, valid values are between 0.0 and 1.0", iou_thres);
    TCB_REQUIRE(ctx, max_det >= 1 && max_nms >= 1, "tcb_yolo_nms: max_det and max_nms must be >= 1");
    if (max_det > MAX_DET_CAP) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_yolo_nms: max_det %d > compiled limit %d", max_det, MAX_DET_CAP);
    TCB_REQUIRE(ctx, n_classes == 0 || classes_host, "tcb_yolo_nms: classes_host is null");
    cudaStream_t st = (cudaStream_t)stream;
    if (b == 0) return TCB_OK;

    ClassMask cm;
    memset(&cm, 0, sizeof(cm));
    for (int i = 0; i < n_classes; i++) {
        int c = classes_host[i];
        if (c >= 256) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_yolo_nms: class filter id %d >= 256", c);
        if (c >= 0) cm.w[c >> 6] |= 1ull << (c & 63);
    }

    const size_t per = (size_t)b * (n > 0 ? n : 1);
    TCB_WS(ctx, counts, int, (size_t)b * 2);
    TCB_WS(ctx, c_score, double, per);
    TCB_WS(ctx, c_box, float4, per);
    TCB_WS(ctx, c_meta, int2, per);
    int npad = 1;
    while (npad < n) npad <<= 1;
    SortEnt* gsort = nullptr;
    if (npad > SORT_SMEM_CAP) {
        TCB_WS(ctx, gs, SortEnt, (size_t)b * npad);
        gsort = gs;
    }
    int* c_count = counts;
    int* xc_count = counts + b;
    TCB_CUDA(ctx, cudaMemsetAsync(counts, 0, sizeof(int) * 2 * (size_t)b, st));

    if (n > 0) {
        const long long groups = (long long)b * ((n + 31) / 32);
        long long blocks = (groups + 7) / 8;
        const long long cap = (long long)ctx->num_sms * 16;
        if (blocks > cap) blocks = cap;
        yolo_filter_kernel<<<(int)blocks, 256, 0, st>>>(pred_dev, b, n, nc, conf_thres, n_classes > 0, cm,
                                                        c_score, c_box, c_meta, c_count, xc_count);
        TCB_LAUNCHED(ctx);
    }
    const size_t smem = sizeof(NmsShared) + sizeof(SortEnt) * SORT_SMEM_CAP;
    TCB_CUDA(ctx, cudaFuncSetAttribute(yolo_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    yolo_nms_kernel<<<b, NMS_THREADS, smem, st>>>(n > 0 ? n : 1, c_score, c_box, c_meta, c_count, gsort, npad,
                                                  iou_thres, agnostic ? 0.0 : 4096.0, max_det, max_nms, out_dev, count_dev);
    TCB_LAUNCHED(ctx);
    if (xc_count_dev)
        TCB_CUDA(ctx, cudaMemcpyAsync(xc_count_dev, xc_count, sizeof(int) * (size_t)b, cudaMemcpyDeviceToDevice, st));
    return TCB_OK;
}

// =================================================================================================
// tcb_yolo_nms_static — the same two kernels, for callers that must not allocate GPU memory.
//
// Who that is: a plugin living inside somebody else's process (a Triton backend, a ROS nodelet next to
// the inference server) whose host has already handed all of HBM to its own pools, so a cudaMalloc of
// ours would fail or eat into a budget we do not own; and latency-bound frame callbacks, for which
// cudaMalloc/cudaFree are off limits because they synchronise the whole device and take unbounded time.
// tcb_create makes exactly such an allocation (the context workspace), so this entry point takes no
// context at all: its scratch is the fixed array below, which every GPU that loads libtcb200.so keeps
// resident until the process exits (TCB_STATIC_SCRATCH_BYTES, about 8.4 MB).
//
// Consequences of sharing one array, which the caller has to respect:
//   * sizes are capped: b <= 64, b*n <= 131072 rows, b*pow2ceil(n) <= 262144 sort slots when n > 4096;
//     anything larger is refused with TCB_ERR_LIMIT (use tcb_yolo_nms with a context instead);
//   * the array still holds the previous call's candidates; the per-image counters are the only part
//     that has to start at zero and they are cleared on the caller's stream at the start of every call;
//   * two calls that overlap on the GPU (different streams, different threads) race on it — issue them
//     on one stream or serialise them yourself;
//   * the first call on each device looks up the array's device address and opts the NMS kernel in to
//     its shared-memory size; make that first call with the device current and before starting any
//     stream capture.
// =================================================================================================
#include <mutex>

namespace {

constexpr int STATIC_MAX_B = 64;
constexpr long long STATIC_MAX_ROWS = 131072;        // b * n
constexpr long long STATIC_MAX_SORT = 262144;        // b * pow2ceil(n), only when pow2ceil(n) > SORT_SMEM_CAP
constexpr size_t ST_OFF_COUNTS = 0;                                             // 2*b ints
constexpr size_t ST_OFF_SCORE = 1024;                                           // rows doubles
constexpr size_t ST_OFF_BOX = ST_OFF_SCORE + STATIC_MAX_ROWS * sizeof(double);  // rows float4
constexpr size_t ST_OFF_META = ST_OFF_BOX + STATIC_MAX_ROWS * sizeof(float4);   // rows int2
constexpr size_t ST_OFF_SORT = ST_OFF_META + STATIC_MAX_ROWS * sizeof(int2);    // sort slots
constexpr size_t TCB_STATIC_SCRATCH_BYTES = ST_OFF_SORT + STATIC_MAX_SORT * sizeof(SortEnt);

// In host code this name is only a handle for cudaGetSymbolAddress, never a pointer to pass to a kernel.
static __device__ __align__(256) unsigned char g_yolo_static_scratch[TCB_STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
std::mutex g_static_mu;
unsigned char* g_static_ptr[STATIC_MAX_DEVICES] = {nullptr};

// One-time set-up for the current device: resolve the device address of the scratch array.
int static_scratch_for_current_device(unsigned char** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return tcb_fail(nullptr, TCB_ERR_CUDA, "tcb_yolo_nms_static: cudaGetDevice -> %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= STATIC_MAX_DEVICES) return tcb_fail(nullptr, TCB_ERR_LIMIT, "tcb_yolo_nms_static: device index %d >= %d", dev, STATIC_MAX_DEVICES);
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        unsigned char* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_yolo_static_scratch);
        if (e != cudaSuccess || !p)
            return tcb_fail(nullptr, TCB_ERR_CUDA, "tcb_yolo_nms_static: cudaGetSymbolAddress -> %s", cudaGetErrorString(e));
        e = cudaFuncSetAttribute(yolo_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)(sizeof(NmsShared) + sizeof(SortEnt) * SORT_SMEM_CAP));
        if (e != cudaSuccess)
            return tcb_fail(nullptr, TCB_ERR_CUDA, "tcb_yolo_nms_static: cudaFuncSetAttribute -> %s", cudaGetErrorString(e));
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return TCB_OK;
}

}  // namespace

extern "C" size_t tcb_yolo_nms_static_scratch_bytes(void) { return TCB_STATIC_SCRATCH_BYTES; }

extern "C" int tcb_yolo_nms_static(const float* pred_dev, int b, int n, int nc, double conf_thres, double iou_thres,
                                   const int* classes_host, int n_classes, int agnostic, int max_det, int max_nms,
                                   double* out_dev, int* count_dev, int* xc_count_dev, void* stream) {
#define ST_REQUIRE(cond, ...) do { if (!(cond)) return tcb_fail(nullptr, TCB_ERR_ARG, __VA_ARGS__); } while (0)
#define ST_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) \
        return tcb_fail(nullptr, TCB_ERR_CUDA, "tcb_yolo_nms_static: %s -> %s", #call, cudaGetErrorString(e__)); } while (0)
    ST_REQUIRE(pred_dev && out_dev && count_dev, "tcb_yolo_nms_static: null buffer");
    ST_REQUIRE(b >= 0 && n >= 0 && nc >= 0, "tcb_yolo_nms_static: bad shape b=%d n=%d nc=%d", b, n, nc);
    ST_REQUIRE(conf_thres >= 0.0 && conf_thres <= 1.0, "tcb_yolo_nms_static: Invalid Confidence threshold %g, valid values are between 0.0 and 1.0", conf_thres);
    ST_REQUIRE(iou_thres >= 0.0 && iou_thres <= 1.0, "tcb_yolo_nms_static: Invalid IoU %g, valid values are between 0.0 and 1.0", iou_thres);
    ST_REQUIRE(max_det >= 1 && max_nms >= 1, "tcb_yolo_nms_static: max_det and max_nms must be >= 1");
    ST_REQUIRE(n_classes == 0 || classes_host, "tcb_yolo_nms_static: classes_host is null");
    if (max_det > MAX_DET_CAP) return tcb_fail(nullptr, TCB_ERR_LIMIT, "tcb_yolo_nms_static: max_det %d > compiled limit %d", max_det, MAX_DET_CAP);
    int npad = 1;
    while (npad < n) npad <<= 1;
    if (b > STATIC_MAX_B || (long long)b * n > STATIC_MAX_ROWS || (npad > SORT_SMEM_CAP && (long long)b * npad > STATIC_MAX_SORT))
        return tcb_fail(nullptr, TCB_ERR_LIMIT,
                        "tcb_yolo_nms_static: b=%d n=%d exceeds the static scratch (b <= %d, b*n <= %lld, b*pow2ceil(n) <= %lld); "
                        "use tcb_yolo_nms with a context workspace for larger batches", b, n, STATIC_MAX_B, STATIC_MAX_ROWS, STATIC_MAX_SORT);
    if (b == 0) return TCB_OK;

    ClassMask cm;
    memset(&cm, 0, sizeof(cm));
    for (int i = 0; i < n_classes; i++) {
        const int c = classes_host[i];
        if (c >= 256) return tcb_fail(nullptr, TCB_ERR_LIMIT, "tcb_yolo_nms_static: class filter id %d >= 256", c);
        if (c >= 0) cm.w[c >> 6] |= 1ull << (c & 63);
    }

    unsigned char* scratch = nullptr;  // device address of g_yolo_static_scratch on the current device
    const int rc = static_scratch_for_current_device(&scratch);
    if (rc != TCB_OK) return rc;
    int* counts = reinterpret_cast<int*>(scratch + ST_OFF_COUNTS);
    double* c_score = reinterpret_cast<double*>(scratch + ST_OFF_SCORE);
    float4* c_box = reinterpret_cast<float4*>(scratch + ST_OFF_BOX);
    int2* c_meta = reinterpret_cast<int2*>(scratch + ST_OFF_META);
    SortEnt* gsort = reinterpret_cast<SortEnt*>(scratch + ST_OFF_SORT);
    int* c_count = counts;
    int* xc_count = counts + b;

    cudaStream_t st = (cudaStream_t)stream;
    ST_CUDA(cudaMemsetAsync(counts, 0, sizeof(int) * 2 * (size_t)b, st));  // the only state that must start at zero
    if (n > 0) {
        int num_sms = 148;
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
        const long long groups = (long long)b * ((n + 31) / 32);
        long long blocks = (groups + 7) / 8;
        if (blocks > (long long)num_sms * 16) blocks = (long long)num_sms * 16;
        yolo_filter_kernel<<<(int)blocks, 256, 0, st>>>(pred_dev, b, n, nc, conf_thres, n_classes > 0, cm,
                                                        c_score, c_box, c_meta, c_count, xc_count);
        ST_CUDA(cudaGetLastError());
    }
    const size_t smem = sizeof(NmsShared) + sizeof(SortEnt) * SORT_SMEM_CAP;
    yolo_nms_kernel<<<b, NMS_THREADS, smem, st>>>(n > 0 ? n : 1, c_score, c_box, c_meta, c_count, gsort, npad,
                                                  iou_thres, agnostic ? 0.0 : 4096.0, max_det, max_nms, out_dev, count_dev);
    ST_CUDA(cudaGetLastError());
    if (xc_count_dev) ST_CUDA(cudaMemcpyAsync(xc_count_dev, xc_count, sizeof(int) * (size_t)b, cudaMemcpyDeviceToDevice, st));
    return TCB_OK;
#undef ST_REQUIRE
#undef ST_CUDA
}
