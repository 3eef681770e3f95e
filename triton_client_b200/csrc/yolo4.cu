// YOLO raw-head decode (v4 as in the reference, v5 per ultralytics) and the YOLOv4-style two-tensor postprocess
// with numpy-semantics per-class NMS; axis-aligned NMS / IoU helpers.
//
//   c6  tools/yolo_layer.py:148-288 (yolo_forward_dynamic), YoloLayer.forward :313-321, get_region_boxes
//       tools/torch_utils.py:48-64; YOLOv5 Detect per SURVEY Appendix A5 (not in the reference, parity unpinned)
//   c4  utils/postprocess.py:265-322 (extract_boxes_onnx), tools/utils.py:166-233 (post_processing)
//   c3  clients/postprocess/base_postprocess.py:72-106 (nms_cpu; = utils/postprocess.py:69-103, tools/utils.py:61-95)
//   c5  clients/postprocess/base_postprocess.py:48-70 (box_iou)
#include "common.cuh"
#include "cta_select.cuh"
#include "cta_nms.cuh"
#include "wide_nms.cuh"

namespace {

constexpr int T = sel::THREADS;
constexpr int PRE_CAP = sel::KCAP;  // candidates per image that can be ordered/resolved (after the confidence cut)
constexpr int POST_CAP = 1024;      // boxes kept per image

__device__ __forceinline__ float sigmoidf_ref(float x) { return 1.f / (1.f + expf(-x)); }  // torch.sigmoid

// ---------------------------------------------------------------------------------------------------------
// c6: one thread per (batch, anchor, y, x) cell of an NCHW head [B, A*(5+nc), H, W]
// ---------------------------------------------------------------------------------------------------------
struct DecodeParams {
    const float* head; int B, H, W, A, nc, variant;
    float aw[8], ah[8];         // v4: anchors / stride ; v5: anchors in pixels
    float scale_x_y, xy_bias, stride;   // v4: sigmoid*scale - bias, bias = 0.5*(scale-1) evaluated in double like the Python scalar
    int row_offset, total_rows;
    float *boxes, *confs, *pred;
};

// One CTA per 32 consecutive cells of one (image, anchor): the head is channel-major (channel c of a cell at
// t[c*HW]), the outputs are row-major (all values of a cell together). Warps read whole channels of the 32 cells
// (128-byte coalesced loads), apply the activation and park the values in a shared-memory tile; then the CTA writes the
// 32 rows, which are contiguous in the output, linearly. (One thread per cell walking its 85 channels was a chain of
// 85 dependent round trips: 50 us even for a 16x16 head.)
constexpr int HD_CELLS = 32, HD_THREADS = 256;
__global__ void __launch_bounds__(HD_THREADS) yolo_head_decode_kernel(const DecodeParams p) {
    pdl_wait();
    extern __shared__ float hd_tile[];              // [HD_CELLS][no + 1]
    const int HW = p.H * p.W;
    const int no = 5 + p.nc, pitch = no + 1;
    const int tiles = (HW + HD_CELLS - 1) / HD_CELLS;
    const int tile = blockIdx.x % tiles, ba = blockIdx.x / tiles;   // ba = b * A + a
    const int a = ba % p.A, b = ba / p.A;
    const int sp0 = tile * HD_CELLS, ncell = min(HD_CELLS, HW - sp0);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float* t = p.head + (long long)ba * no * HW + sp0;
    // activations, channel by channel (warp = channel, lane = cell)
    for (int c = warp; c < no; c += HD_THREADS / 32) {
        if (lane < ncell) {
            const float v = __ldg(t + (long long)c * HW + lane);
            float r;
            if (p.variant == 0) r = (c == 2 || c == 3) ? expf(v) : sigmoidf_ref(v);
            else r = sigmoidf_ref(v);
            hd_tile[lane * pitch + c] = r;
        }
    }
    __syncthreads();
    const long long row0 = (long long)b * p.total_rows + p.row_offset + (long long)a * HW + sp0;   // rows ordered (anchor, y, x)
    if (p.variant == 0) {
        // YOLOv4 (yolo_layer.py:195-276): bxy = sigmoid*s - 0.5(s-1) + grid; bwh = exp*anchor; /W,/H; xyxy; conf = cls*obj
        if (threadIdx.x < ncell) {
            const int i = threadIdx.x, sp = sp0 + i, y = sp / p.W, x = sp - y * p.W;
            const float* r = hd_tile + i * pitch;
            const float bx = (r[0] * p.scale_x_y - p.xy_bias + (float)x) / (float)p.W;
            const float by = (r[1] * p.scale_x_y - p.xy_bias + (float)y) / (float)p.H;
            const float bw = r[2] * p.aw[a] / (float)p.W;
            const float bh = r[3] * p.ah[a] / (float)p.H;
            const float x1 = bx - bw * 0.5f, y1 = by - bh * 0.5f;
            reinterpret_cast<float4*>(p.boxes)[row0 + i] = make_float4(x1, y1, x1 + bw, y1 + bh);
        }
        float* c = p.confs + row0 * p.nc;
        for (int e = threadIdx.x; e < ncell * p.nc; e += HD_THREADS) {
            const int i = e / p.nc, k = e - i * p.nc;
            c[e] = hd_tile[i * pitch + 5 + k] * hd_tile[i * pitch + 4];
        }
    } else {
        // YOLOv5 Detect: y = sigmoid(t); xy = (y*2 - 0.5 + grid)*stride; wh = (y*2)^2 * anchor; obj, cls = y
        float* o = p.pred + row0 * no;
        for (int e = threadIdx.x; e < ncell * no; e += HD_THREADS) {
            const int i = e / no, c = e - i * no;
            const float s = hd_tile[i * pitch + c];
            float v = s;
            if (c < 2) {
                const int sp = sp0 + i, y = sp / p.W, x = sp - y * p.W;
                v = (s * 2.f - 0.5f + (float)(c == 0 ? x : y)) * p.stride;
            } else if (c < 4) {
                v = (s * 2.f) * (s * 2.f) * (c == 2 ? p.aw[a] : p.ah[a]);
            }
            o[e] = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// axis-aligned boxes, numpy nms_cpu semantics in float32
// ---------------------------------------------------------------------------------------------------------
struct ABox { float x1, y1, x2, y2, area; int cls; };

struct NumpyNmsPred {
    static constexpr bool kQueue = false;   // cheap predicate: tested in place
    __device__ __forceinline__ bool maybe(const ABox&, const ABox&) const { return true; }
    __device__ __forceinline__ bool exact(const ABox& a, const ABox& b) const { return (*this)(a, b); }
    float thresh; int min_mode;
    __device__ __forceinline__ bool operator()(const ABox& a, const ABox& b) const {
        if (a.cls != b.cls) return false;  // per-class NMS (extract_boxes_onnx :296-303)
        const float xx1 = fmaxf(a.x1, b.x1), yy1 = fmaxf(a.y1, b.y1), xx2 = fminf(a.x2, b.x2), yy2 = fminf(a.y2, b.y2);
        const float w = fmaxf(0.f, xx2 - xx1), h = fmaxf(0.f, yy2 - yy1);
        const float inter = w * h;
        const float over = min_mode ? inter / fminf(a.area, b.area) : inter / (a.area + b.area - inter);
        return !(over <= thresh);  // the reference keeps `over <= nms_thresh`
    }
};

typedef nms::Shared<ABox, POST_CAP, PRE_CAP> ANmsShared;
struct AFrameShared {
    unsigned long long skeys[sel::KCAP];
    ABox R[PRE_CAP];
    sel::Scratch sel;
    ANmsShared nms;
};

__device__ __forceinline__ ABox make_abox(const float* b4, int cls) {
    ABox r;
    r.x1 = b4[0]; r.y1 = b4[1]; r.x2 = b4[2]; r.y2 = b4[3];
    r.area = (r.x2 - r.x1) * (r.y2 - r.y1);
    r.cls = cls;
    return r;
}

// c3: nms_cpu(boxes, confs, thresh, min_mode) -> keep indices (score descending, ties lower index first)
__global__ void __launch_bounds__(T, 1) aa_nms_kernel(const float* __restrict__ boxes, const float* __restrict__ scores, int n, float thresh,
                                                      int min_mode, int post_max, int* __restrict__ keep, int* __restrict__ count) {
    pdl_wait();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    AFrameShared& S = *reinterpret_cast<AFrameShared*>(smem_raw);
    auto key_of = [&](int i) { return sel::make_key(__ldg(scores + i), (unsigned)i); };
    const int m = sel::cta_topk_sorted(key_of, n, PRE_CAP, S.skeys, S.sel);
    for (int c = threadIdx.x; c < m; c += T) S.R[c] = make_abox(boxes + 4ll * sel::key_index(S.skeys[c]), 0);
    __syncthreads();
    const int nk = nms::cta_greedy_nms(S.R, m, NumpyNmsPred{thresh, min_mode}, post_max, S.nms);
    for (int k = threadIdx.x; k < nk; k += T) keep[k] = (int)sel::key_index(S.skeys[S.nms.kpos[k]]);
    if (threadIdx.x == 0) { count[0] = nk; count[1] = n > PRE_CAP ? 1 : 0; }
}

// c4 pass 1: per row max / first arg-max over the class confidences, keep max > conf_thresh.
// key = [4095 - class : 12][score : 32][0xFFFFF - row : 20]: descending key order == class ascending, score
// descending, row ascending — the order in which extract_boxes_onnx emits its rows.
__global__ void __launch_bounds__(256) yolo4_candidates_kernel(const float* __restrict__ confs, int B, int N, int nc, float conf_thresh,
                                                               unsigned long long* __restrict__ cand, int* __restrict__ cand_count) {
    pdl_wait();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    bool ok = false;
    unsigned long long key = 0ull;
    if (i < N) {
        const float* c = confs + ((long long)b * N + i) * nc;
        float best = __ldg(c);
        int cls = 0;
        for (int k = 1; k < nc; k++) { const float v = __ldg(c + k); if (v > best) { best = v; cls = k; } }
        ok = best > conf_thresh;
        key = ((unsigned long long)(4095 - cls) << 52) | ((unsigned long long)sel::f32_ordered(best) << 20) | (unsigned long long)(0xFFFFF - (unsigned)i);
    }
    const unsigned em = __ballot_sync(0xffffffffu, ok);
    if (em) {
        const int lane = threadIdx.x & 31;
        int base = 0;
        if (lane == 0) base = atomicAdd(&cand_count[b], __popc(em));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (ok) cand[(long long)b * N + base + __popc(em & ((1u << lane) - 1))] = key;
    }
}

__global__ void __launch_bounds__(T, 1) yolo4_nms_kernel(const float* __restrict__ boxes, int N, const unsigned long long* __restrict__ cand,
                                                         const int* __restrict__ cand_count, float nms_thresh, int post_max,
                                                         float* __restrict__ out, int* __restrict__ out_count) {
    pdl_wait();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    AFrameShared& S = *reinterpret_cast<AFrameShared*>(smem_raw);
    const int b = blockIdx.x, tid = threadIdx.x;
    const int nc = min(cand_count[b], N);
    const unsigned long long* ck = cand + (long long)b * N;
    auto key_of = [&](int i) { return ck[i]; };
    const int m = sel::cta_topk_sorted(key_of, nc, PRE_CAP, S.skeys, S.sel);
    for (int c = tid; c < m; c += T) {
        const unsigned long long k = S.skeys[c];
        const int row = (int)(0xFFFFF - (unsigned)(k & 0xFFFFF)), cls = 4095 - (int)(k >> 52);
        S.R[c] = make_abox(boxes + ((long long)b * N + row) * 4, cls);
    }
    __syncthreads();
    const int nk = nms::cta_greedy_nms(S.R, m, NumpyNmsPred{nms_thresh, 0}, post_max, S.nms);
    float* o = out + (long long)b * post_max * 7;
    for (int k = tid; k < nk; k += T) {
        const int c = S.nms.kpos[k];
        const unsigned long long key = S.skeys[c];
        const ABox bx = S.R[c];
        const float conf = sel::f32_from_ordered((unsigned)((key >> 20) & 0xFFFFFFFFull));
        float* r = o + k * 7;  // [x1, y1, x2, y2, conf, conf, cls] (utils/postprocess.py:314-318)
        r[0] = bx.x1; r[1] = bx.y1; r[2] = bx.x2; r[3] = bx.y2; r[4] = conf; r[5] = conf; r[6] = (float)bx.cls;
    }
    if (tid == 0) { out_count[b] = nk; out_count[gridDim.x + b] = cand_count[b] > PRE_CAP ? 1 : 0; }
}

// ---------------------------------------------------------------------------------------------------------
// Wide path (more than WIDE_MIN rows): order kernel (one CTA per image) -> wide::tile_mask_kernel over the grid ->
// sweep kernel (wide::warp_mask_sweep). The single CTA above walks the candidates in rounds of 64 survivors, which with
// thousands of kept boxes (80 classes) is hundreds of rounds.
// ---------------------------------------------------------------------------------------------------------
constexpr int WIDE_MIN = 512;
constexpr int ORDER_CACHE = 16384;   // keys held in dynamic shared memory across the selection passes
struct OrderShared {
    unsigned long long skeys[sel::KCAP];
    sel::Scratch sel;
};

// mode 0: nms_cpu (scores + boxes, one set); mode 1: YOLOv4 candidates of image blockIdx.x
__global__ void __launch_bounds__(T, 1) aa_order_kernel(int mode, const float* __restrict__ boxes, const float* __restrict__ scores, int N,
                                                        const unsigned long long* __restrict__ cand, const int* __restrict__ cand_count,
                                                        unsigned long long* __restrict__ skeys_out, ABox* __restrict__ R_out, int* __restrict__ m_out,
                                                        int* __restrict__ cut_out) {
    pdl_wait();
    __shared__ OrderShared S;
    extern __shared__ __align__(16) unsigned long long key_cache[];
    const int b = blockIdx.x, tid = threadIdx.x;
    int m;
    if (mode == 0) {
        auto key_of = [&](int i) { return sel::make_key(__ldg(scores + i), (unsigned)i); };
        m = sel::cta_topk_sorted(key_of, N, PRE_CAP, S.skeys, S.sel, key_cache, ORDER_CACHE);
    } else {
        const unsigned long long* ck = cand + (long long)b * N;
        auto key_of = [&](int i) { return ck[i]; };
        m = sel::cta_topk_sorted(key_of, min(cand_count[b], N), PRE_CAP, S.skeys, S.sel, key_cache, ORDER_CACHE);
    }
    for (int c = tid; c < m; c += T) {
        const unsigned long long k = S.skeys[c];
        skeys_out[(long long)b * PRE_CAP + c] = k;
        if (mode == 0) R_out[c] = make_abox(boxes + 4ll * sel::key_index(k), 0);
        else {
            const int row = (int)(0xFFFFF - (unsigned)(k & 0xFFFFF)), cls = 4095 - (int)(k >> 52);
            R_out[(long long)b * PRE_CAP + c] = make_abox(boxes + ((long long)b * N + row) * 4, cls);
        }
    }
    if (tid == 0) {
        m_out[b] = m;
        if (mode == 1) cut_out[gridDim.x + b] = cand_count[b] > PRE_CAP ? 1 : 0;   // candidates beyond the 4096 best were dropped
    }
}

__global__ void __launch_bounds__(wide::SWEEP_T, 1) aa_sweep_kernel(int mode, const unsigned long long* __restrict__ mask_all,
                                                                   const unsigned long long* __restrict__ skeys_all, const ABox* __restrict__ R_all,
                                                                   const int* __restrict__ m_all, int n_rows, int post_max,
                                                                   int* __restrict__ keep, float* __restrict__ out, int* __restrict__ count) {
    pdl_wait();
    extern __shared__ __align__(128) unsigned char sweep_raw[];
    wide::SweepShared& S = *reinterpret_cast<wide::SweepShared*>(sweep_raw);
    __shared__ int nk_s;
    const int b = blockIdx.x, tid = threadIdx.x;
    const unsigned long long* skeys = skeys_all + (long long)b * PRE_CAP;
    if (tid < 32) {
        const int v = wide::warp_mask_sweep(mask_all + (size_t)b * PRE_CAP * wide::MW, m_all[b], post_max, S);
        if (tid == 0) nk_s = v;
    }
    __syncthreads();
    const int nk = nk_s;
    if (mode == 0) {
        for (int k = tid; k < nk; k += wide::SWEEP_T) keep[k] = (int)sel::key_index(skeys[S.kpos[k]]);
        if (tid == 0) { count[0] = nk; count[1] = n_rows > PRE_CAP ? 1 : 0; }
    } else {
        const ABox* R = R_all + (long long)b * PRE_CAP;
        float* o = out + (long long)b * post_max * 7;
        for (int k = tid; k < nk; k += wide::SWEEP_T) {
            const int c = S.kpos[k];
            const unsigned long long key = skeys[c];
            const ABox bx = R[c];
            const float conf = sel::f32_from_ordered((unsigned)((key >> 20) & 0xFFFFFFFFull));
            float* r = o + k * 7;  // [x1, y1, x2, y2, conf, conf, cls] (utils/postprocess.py:314-318)
            r[0] = bx.x1; r[1] = bx.y1; r[2] = bx.x2; r[3] = bx.y2; r[4] = conf; r[5] = conf; r[6] = (float)bx.cls;
        }
        if (tid == 0) count[b] = nk;
    }
}

// c5: pairwise IoU (base_postprocess.py:48-70), float32
__global__ void __launch_bounds__(256) box_iou_kernel(const float* __restrict__ a, int n, const float* __restrict__ b, int m, float* __restrict__ out) {
    pdl_wait();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n * m) return;
    const int ia = (int)(i / m), ib = (int)(i - (long long)ia * m);
    const float4 p = __ldg(reinterpret_cast<const float4*>(a) + ia), q = __ldg(reinterpret_cast<const float4*>(b) + ib);
    const float area1 = (p.z - p.x) * (p.w - p.y), area2 = (q.z - q.x) * (q.w - q.y);
    const float w = fmaxf(fminf(p.z, q.z) - fmaxf(p.x, q.x), 0.f), h = fmaxf(fminf(p.w, q.w) - fmaxf(p.y, q.y), 0.f);
    const float inter = w * h;
    out[i] = inter / (area1 + area2 - inter);
}

}  // namespace

static int init_yolo4(tcb_ctx* ctx) {   // run by tcb_create
    TCB_SMEM_OPT_IN(ctx, yolo_head_decode_kernel, 200 * 1024);
    TCB_SMEM_OPT_IN(ctx, aa_order_kernel, ORDER_CACHE * 8);
    TCB_SMEM_OPT_IN(ctx, aa_sweep_kernel, sizeof(wide::SweepShared));
    TCB_SMEM_OPT_IN(ctx, aa_nms_kernel, sizeof(AFrameShared));
    TCB_SMEM_OPT_IN(ctx, yolo4_nms_kernel, sizeof(AFrameShared));
    return TCB_OK;
}
static TcbInitReg reg_yolo4(&init_yolo4);

extern "C" int tcb_yolo_head_decode(tcb_ctx* ctx, const float* head_dev, int b, int h, int w, int num_anchors, int num_classes,
                                    const float* anchors_host, int variant, double scale_x_y, float stride, int row_offset,
                                    int total_rows, float* boxes_dev, float* confs_dev, float* pred_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, head_dev && anchors_host, "tcb_yolo_head_decode: null argument");
    TCB_REQUIRE(ctx, b >= 0 && h >= 1 && w >= 1 && num_classes >= 0, "tcb_yolo_head_decode: bad shape");
    if (num_anchors < 1 || num_anchors > 8) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_yolo_head_decode: %d anchors per head (limit 8)", num_anchors);
    TCB_REQUIRE(ctx, variant == 0 || variant == 1, "tcb_yolo_head_decode: variant must be 0 (YOLOv4) or 1 (YOLOv5)");
    TCB_REQUIRE(ctx, variant == 0 ? (boxes_dev && confs_dev) : (pred_dev != nullptr), "tcb_yolo_head_decode: missing output buffer");
    TCB_REQUIRE(ctx, row_offset >= 0 && row_offset + (long long)num_anchors * h * w <= total_rows, "tcb_yolo_head_decode: rows do not fit (offset %d + %lld > %d)",
                row_offset, (long long)num_anchors * h * w, total_rows);
    TCB_REQUIRE(ctx, variant != 0 || (((uintptr_t)boxes_dev & 15) == 0), "tcb_yolo_head_decode: boxes buffer must be 16-byte aligned");
    if (b == 0) return TCB_OK;
    DecodeParams p;
    memset(&p, 0, sizeof(p));
    p.head = head_dev; p.B = b; p.H = h; p.W = w; p.A = num_anchors; p.nc = num_classes; p.variant = variant;
    for (int i = 0; i < num_anchors; i++) { p.aw[i] = anchors_host[2 * i]; p.ah[i] = anchors_host[2 * i + 1]; }
    p.scale_x_y = (float)scale_x_y; p.xy_bias = (float)(0.5 * (scale_x_y - 1.0)); p.stride = stride; p.row_offset = row_offset; p.total_rows = total_rows;
    p.boxes = boxes_dev; p.confs = confs_dev; p.pred = pred_dev;
    const long long ctas = (long long)b * num_anchors * tcb_div_up((long long)h * w, HD_CELLS);
    TCB_REQUIRE(ctx, ctas < (1ll << 31), "tcb_yolo_head_decode: head too large");
    const size_t tile_bytes = (size_t)HD_CELLS * (5 + num_classes + 1) * sizeof(float);
    if (tile_bytes > 200 * 1024) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_yolo_head_decode: %d classes exceed the shared-memory tile", num_classes);
    TCB_KLAUNCH(ctx, yolo_head_decode_kernel, (unsigned)ctas, HD_THREADS, tile_bytes, (cudaStream_t)stream, p);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

// Workspace + launches of the wide path for `sets` independent candidate sets; false when the workspace has no room
// (the caller then stays on the single-CTA kernels).
struct AAWide { unsigned long long* skeys; ABox* R; unsigned long long* mask; int* m; };
static bool aa_wide_bufs(tcb_ctx* ctx, int sets, AAWide& wb) {
    const size_t per = (size_t)PRE_CAP * (8 + sizeof(ABox) + (size_t)wide::MW * 8) + 256;
    if (((ctx->ws_used + 255) & ~(size_t)255) + per * sets + 4096 > ctx->ws_bytes) return false;
    wb.skeys = (unsigned long long*)tcb_ws_alloc(ctx, (size_t)sets * PRE_CAP * 8);
    wb.R = (ABox*)tcb_ws_alloc(ctx, (size_t)sets * PRE_CAP * sizeof(ABox));
    wb.mask = (unsigned long long*)tcb_ws_alloc(ctx, (size_t)sets * PRE_CAP * wide::MW * 8);
    wb.m = (int*)tcb_ws_alloc(ctx, (size_t)sets * sizeof(int) + 256);
    return wb.skeys && wb.R && wb.mask && wb.m;
}
static int aa_wide_run(tcb_ctx* ctx, int mode, int sets, int n_rows, const float* boxes, const float* scores, const unsigned long long* cand,
                       const int* cand_count, const AAWide& wb, NumpyNmsPred pred, int post_max, int* keep, float* out, int* count,
                       cudaStream_t st) {
    TCB_KLAUNCH(ctx, aa_order_kernel, sets, T, ORDER_CACHE * 8, st, mode, boxes, scores, n_rows, cand, cand_count, wb.skeys, wb.R, wb.m, count);
    TCB_LAUNCHED(ctx);
    const int nb = tcb_div_up(n_rows < PRE_CAP ? n_rows : PRE_CAP, 64);
    TCB_KLAUNCH(ctx, (wide::tile_mask_kernel<ABox, NumpyNmsPred>), dim3(nb, nb, sets), 64, 0, st, wb.R, wb.m, 1, pred, wb.mask);
    TCB_LAUNCHED(ctx);
    TCB_KLAUNCH(ctx, aa_sweep_kernel, sets, wide::SWEEP_T, sizeof(wide::SweepShared), st, mode, wb.mask, wb.skeys, wb.R, wb.m, n_rows, post_max, keep, out, count);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

extern "C" int tcb_nms_axis_aligned(tcb_ctx* ctx, const float* boxes_dev, const float* scores_dev, int n, float thresh, int min_mode,
                                    int post_max, int* keep_dev, int* count_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, n >= 0 && keep_dev && count_dev && (n == 0 || (boxes_dev && scores_dev)), "tcb_nms_axis_aligned: bad arguments");
    TCB_REQUIRE(ctx, post_max >= 1, "tcb_nms_axis_aligned: post_max must be >= 1");
    if (post_max > POST_CAP) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_nms_axis_aligned: post_max %d > compiled limit %d", post_max, POST_CAP);
    AAWide wb;
    if (n > WIDE_MIN && aa_wide_bufs(ctx, 1, wb))
        return aa_wide_run(ctx, 0, 1, n, boxes_dev, scores_dev, nullptr, nullptr, wb, NumpyNmsPred{thresh, min_mode ? 1 : 0}, post_max, keep_dev, nullptr,
                           count_dev, (cudaStream_t)stream);
    TCB_KLAUNCH(ctx, aa_nms_kernel, 1, T, sizeof(AFrameShared), (cudaStream_t)stream, boxes_dev, scores_dev, n, thresh, min_mode ? 1 : 0, post_max, keep_dev, count_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

extern "C" size_t tcb_yolo4_post_workspace(int b, int n) {
    const size_t bb = (size_t)(b > 0 ? b : 0), nn = (size_t)(n > 0 ? n : 0);
    size_t bytes = bb * nn * 8 + bb * 4 + 4096;
    if (n > WIDE_MIN) bytes += bb * ((size_t)PRE_CAP * (8 + sizeof(ABox) + (size_t)wide::MW * 8) + 256) + 8192;   // wide path (optional: without it one CTA per image does the NMS)
    return bytes;
}

extern "C" int tcb_yolo4_post(tcb_ctx* ctx, const float* boxes_dev, const float* confs_dev, int b, int n, int nc, float conf_thresh,
                              float nms_thresh, int post_max, float* out_dev, int* count_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, b >= 0 && n >= 0 && nc >= 1 && out_dev && count_dev, "tcb_yolo4_post: bad arguments");
    TCB_REQUIRE(ctx, (b == 0 || n == 0) || (boxes_dev && confs_dev), "tcb_yolo4_post: null input");
    TCB_REQUIRE(ctx, post_max >= 1, "tcb_yolo4_post: post_max must be >= 1");
    if (post_max > POST_CAP) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_yolo4_post: post_max %d > compiled limit %d", post_max, POST_CAP);
    if (nc > 4096 || n > (1 << 20) || b > 65535) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_yolo4_post: nc <= 4096, n <= 2^20, b <= 65535");
    if (b == 0) return TCB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    TCB_WS(ctx, cand_count, int, b);
    TCB_WS(ctx, cand, unsigned long long, (size_t)b * (n > 0 ? n : 1));
    TCB_CUDA(ctx, cudaMemsetAsync(cand_count, 0, sizeof(int) * b, st));
    if (n > 0) {
        TCB_KLAUNCH(ctx, yolo4_candidates_kernel, dim3(tcb_div_up(n, 256), b), 256, 0, st, confs_dev, b, n, nc, conf_thresh, cand, cand_count);
        TCB_LAUNCHED(ctx);
    }
    AAWide wb;
    if (n > WIDE_MIN && aa_wide_bufs(ctx, b, wb))
        return aa_wide_run(ctx, 1, b, n, boxes_dev, nullptr, cand, cand_count, wb, NumpyNmsPred{nms_thresh, 0}, post_max, nullptr, out_dev, count_dev, st);
    TCB_KLAUNCH(ctx, yolo4_nms_kernel, b, T, sizeof(AFrameShared), st, boxes_dev, n > 0 ? n : 1, cand, cand_count, nms_thresh, post_max, out_dev, count_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

extern "C" int tcb_box_iou(tcb_ctx* ctx, const float* boxes_a_dev, int n, const float* boxes_b_dev, int m, float* out_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, n >= 0 && m >= 0, "tcb_box_iou: bad sizes");
    if (n == 0 || m == 0) return TCB_OK;
    TCB_REQUIRE(ctx, boxes_a_dev && boxes_b_dev && out_dev, "tcb_box_iou: null buffer");
    TCB_REQUIRE(ctx, (((uintptr_t)boxes_a_dev | (uintptr_t)boxes_b_dev) & 15) == 0, "tcb_box_iou: box arrays must be 16-byte aligned");
    TCB_KLAUNCH(ctx, box_iou_kernel, tcb_div_up((long long)n * m, 256), 256, 0, (cudaStream_t)stream, boxes_a_dev, n, boxes_b_dev, m, out_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}
