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
constexpr int SORT_SMEM_CAP = 4096; // entries sorted in shared memory; larger sets sort in global scratch
constexpr int MAX_DET_CAP = 1024;

// Sort record, stored as two arrays so a compare touches 8 bytes and only a swap (or a tie) touches the rest:
//   key  : f64 score bits (scores are > conf_thres >= 0, so bit order == numeric order)
//   meta : x = row index inside the image (the deterministic tie-break, lower first), y = slot in the candidate arrays
struct SortEnt { unsigned long long key; uint2 meta; };  // 16 bytes per slot of sort scratch

struct ClassMask { unsigned long long w[4]; };  // whitelist for classes 0..255

// ---------------------------------------------------------------------------------------------
// Pass 1: filter + decode. One warp scans 32 rows at a time (one objectness load per lane), then the
// whole warp cooperates on each surviving row (coalesced class-score loads, shuffle arg-max).
// ---------------------------------------------------------------------------------------------
constexpr int FILTER_UNROLL = 4;   // 32-row groups in flight per warp (independent objectness loads)
constexpr int FILTER_SMALL_NC = 8;  // up to this many classes a surviving lane decodes its own row

// A frame whose candidates all have every coordinate inside (-2048, 2048) cannot have boxes of different classes
// overlap after the reference's class * 4096 offset, so its NMS splits exactly into independent per-class problems
// (yolo_nms_kernel, class groups). The filters raise big[b] for the frames where that is not guaranteed.
__device__ __forceinline__ void note_big(const float4 bx, int* __restrict__ big, int b) {
    if (big && !(fabsf(bx.x) < 2048.f && fabsf(bx.y) < 2048.f && fabsf(bx.z) < 2048.f && fabsf(bx.w) < 2048.f)) atomicOr(&big[b], 1);
}

__device__ __forceinline__ void emit_candidate(const float* __restrict__ rp, int r, double best, int bidx, long long k,
                                               double* __restrict__ c_score, float4* __restrict__ c_box, int2* __restrict__ c_meta,
                                               int* __restrict__ big, int b) {
    const double x = (double)__ldg(rp), y = (double)__ldg(rp + 1);
    const double w = (double)__ldg(rp + 2), h = (double)__ldg(rp + 3);
    float4 bx;
    bx.x = (float)(x - w / 2); bx.y = (float)(y - h / 2);
    bx.z = (float)(x + w / 2); bx.w = (float)(y + h / 2);
    c_score[k] = best; c_box[k] = bx; c_meta[k] = make_int2(r, bidx);
    note_big(bx, big, b);
}

__global__ void __launch_bounds__(256) yolo_filter_kernel(
    const float* __restrict__ pred, int B, int N, int nc, double conf_thres, int use_filter, ClassMask cmask,
    double* __restrict__ c_score, float4* __restrict__ c_box, int2* __restrict__ c_meta,
    int* __restrict__ c_count, int* __restrict__ xc_count, int* __restrict__ big) {
    // grid: x = warps striding over the 32-row groups of one image, y = image (no index division needed)
    const int lane = threadIdx.x & 31;
    const int warp_in_img = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    const int stride = 5 + nc;
    const int groups = (N + 31) / 32;
    const int b = blockIdx.y;
    const float* base = pred + ((long long)b * N) * stride;

    for (int g0 = warp_in_img * FILTER_UNROLL; g0 < groups; g0 += nwarps * FILTER_UNROLL) {
        float obj[FILTER_UNROLL];
#pragma unroll
        for (int u = 0; u < FILTER_UNROLL; u++) {  // all objectness loads first
            const int row = (g0 + u) * 32 + lane;
            obj[u] = (row < N) ? __ldg(base + (long long)row * stride + 4) : -1.f;
        }
#pragma unroll
        for (int u = 0; u < FILTER_UNROLL; u++) {
            const int row0 = (g0 + u) * 32;
            if (row0 >= N) break;
            const bool pass = (double)obj[u] > conf_thres;  // padding lanes hold -1 and conf_thres >= 0
            unsigned m = __ballot_sync(0xffffffffu, pass);
            if (m == 0) continue;
            if (lane == 0) atomicAdd(&xc_count[b], __popc(m));
            if (nc <= FILTER_SMALL_NC) {
                // few classes: each surviving lane decodes its own row
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
                                           c_score, c_box, c_meta, big, b);
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
                        if (ok) emit_candidate(rp, r, best, bidx, (long long)b * N + atomicAdd(&c_count[b], 1), c_score, c_box, c_meta, big, b);
                    }
                }
            }
        }
    }
}

// Small class counts (rows of <= 13 floats): every 32-byte sector of the tensor is touched anyway, so the rows are
// pulled into shared memory whole (one DRAM round trip instead of the objectness -> class scores -> box chain of
// dependent sector fetches) and filtered / decoded from there.
constexpr int FS_NBUF = 2;     // chunk buffers per warp: one bulk copy in flight behind the chunk being filtered (3 and 4 measured no faster)
// As a stream: every warp owns FS_NBUF buffers of FS_ROWS rows and keeps bulk copies (TMA, completion on one
// mbarrier per buffer) of its next chunks in flight while it filters the current one. The candidates of a chunk go
// to the chunk's own slots (FS_ROWS*c + 0, 1, ...) and its counts to chunk_info, so nothing is reserved atomically
// and nothing has to be zeroed. Chunks that are not 16-byte aligned/sized (the last one of a frame, odd strides,
// unaligned tensors) are copied by the lanes instead.
__device__ __forceinline__ unsigned yf_smem(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void yf_bar_init(unsigned long long* bar) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(yf_smem(bar)) : "memory");
}
__device__ __forceinline__ void yf_bulk_load(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(yf_smem(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(yf_smem(dst)), "l"(src), "r"(bytes), "r"(yf_smem(bar)) : "memory");
}
__device__ __forceinline__ void yf_bar_wait(unsigned long long* bar, unsigned phase) {
    unsigned done = 0;
    while (!done) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(yf_smem(bar)), "r"(phase) : "memory");
    }
}

constexpr int FS_WARPS = 8;
template <int FS_ROWS>   // rows per chunk: 128 for rows of <= 13 floats, 32 for wide rows (COCO: 85 floats) on small batches
__global__ void __launch_bounds__(32 * FS_WARPS) yolo_filter_stream_kernel(
    const float* __restrict__ pred, int B, int N, int nc, double conf_thres, float conf_floor32, int use_filter, ClassMask cmask,
    double* __restrict__ c_score, float4* __restrict__ c_box, int2* __restrict__ c_meta, int2* __restrict__ chunk_info) {
    extern __shared__ __align__(128) unsigned char ys_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int stride = 5 + nc;
    pdl_enter();
    const int buf_floats = FS_ROWS * stride;                        // FS_ROWS rows: a multiple of 4 floats, 16-byte sized
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(ys_raw);            // [FS_WARPS][FS_NBUF]
    unsigned char* plist = ys_raw + 8 * FS_WARPS * FS_NBUF + warp * FS_ROWS;            // rows of the current chunk that pass objectness
    float* bufs = reinterpret_cast<float*>(ys_raw + 8 * FS_WARPS * FS_NBUF + FS_WARPS * FS_ROWS) + (size_t)warp * FS_NBUF * buf_floats;
    unsigned long long* bar = bars + FS_NBUF * warp;
    const unsigned lt_mask = (1u << lane) - 1u;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < FS_NBUF; k++) yf_bar_init(bar + k);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const int chunks = (N + FS_ROWS - 1) / FS_ROWS;
    const int total = B * chunks;                 // the host keeps b * chunks below 2^30
    const int nwarps = (int)gridDim.x * FS_WARPS;
    int ch = (int)blockIdx.x * FS_WARPS + warp;

    // Chunk c of the stream is chunk (c % chunks) of frame (c / chunks). A warp visits c, c + nwarps, ...: the frame and
    // the chunk inside it advance by a fixed step with one carry, so the loop has no division (two of them per chunk
    // were a quarter of this kernel's instructions).
    const int step_b = nwarps / chunks, step_c = nwarps - step_b * chunks;
    struct Pos { int b, c; };
    auto advance = [&](Pos& p) { p.b += step_b; p.c += step_c; if (p.c >= chunks) { p.c -= chunks; p.b++; } };
    auto chunk_src = [&](const Pos& p, int& row0, int& nrow) -> const float* {
        row0 = p.c * FS_ROWS;
        nrow = min(FS_ROWS, N - row0);
        return pred + ((long long)p.b * N + row0) * stride;
    };
    auto bulk_ok = [&](const float* src, int nrow) { return ((((uintptr_t)src) | (uintptr_t)(nrow * stride * 4)) & 15) == 0; };
    auto issue = [&](const Pos& p, int slot) {
        int row0, nrow;
        const float* src = chunk_src(p, row0, nrow);
        if (lane == 0 && bulk_ok(src, nrow)) yf_bulk_load(bufs + slot * buf_floats, src, (unsigned)(nrow * stride * 4), bar + slot);
    };

    Pos cur, ahead;                               // the chunk being filtered and the one FS_NBUF - 1 visits further on
    cur.b = ch / chunks; cur.c = ch - cur.b * chunks;
    ahead = cur;
#pragma unroll
    for (int k = 0; k < FS_NBUF - 1; k++) {
        if (ch + k * nwarps < total) issue(ahead, k);
        advance(ahead);
    }
    const int lane_off = lane * stride + 4, group_off = 32 * stride;
    unsigned phases = 0;   // bit k = parity of buffer k's next completion; flipped only by a real bulk wait, because a
                           // chunk the lanes copy themselves (unaligned frame start, ragged tail) never arrives on the barrier
    for (int it = 0; ch < total; ch += nwarps, it++, advance(cur), advance(ahead)) {
        const int slot = it % FS_NBUF;
        // refill the buffer filtered in the previous iteration (released by the __syncwarp at the end of it)
        if (ch + (FS_NBUF - 1) * nwarps < total) issue(ahead, (it + FS_NBUF - 1) % FS_NBUF);
        const int b = cur.b;
        int row0, nrow;
        const float* src = chunk_src(cur, row0, nrow);
        float* rows = bufs + slot * buf_floats;
        if (bulk_ok(src, nrow)) {
            yf_bar_wait(bar + slot, (phases >> slot) & 1u);
            phases ^= 1u << slot;
        } else {
            for (int i = lane; i < nrow * stride; i += 32) rows[i] = __ldg(src + i);
            __syncwarp();
        }
        // (1) objectness of every row, one load and one compare per row; the rows that pass (a few per cent) are listed
        //     in row order. (2) the expensive part — float64 products over the classes, box conversion, stores — then
        //     runs once per 32 listed rows with the lanes packed, instead of once per 32-row group with one or two
        //     lanes active.
        int n_pass = 0;
#pragma unroll
        for (int u = 0; u < FS_ROWS / 32; u++) {
            const int r = u * 32 + lane;
            const float obj = (r < nrow) ? rows[lane_off + u * group_off] : -1.f;
            const bool pass = obj > conf_floor32;   // == ((double)obj > conf_thres): conf_floor32 is the largest float <= conf_thres
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (pass) plist[n_pass + __popc(m & lt_mask)] = (unsigned char)r;
            n_pass += __popc(m);
        }
        __syncwarp();
        // The chunk's candidates go to slots row0, row0+1, ... of the frame (its own region: no slot reservation, no
        // atomics); the counts and the "a box reaches beyond +-2048" flag go to chunk_info for the NMS kernel.
        int n_ok = 0;
        bool bigc = false;
        for (int base = 0; base < n_pass; base += 32) {
            const int k = base + lane;
            bool ok = false;
            double best = -1.0;
            int bidx = 0, r = 0;
            const float* rp = rows;
            if (k < n_pass) {
                r = plist[k];
                rp = rows + r * stride;
                const double o = (double)rp[4];
                bool nan = false;
                for (int c = 0; c < nc; c++) {
                    const double v = (double)rp[5 + c] * o;
                    nan |= (v != v);
                    if (v > best) { best = v; bidx = c; }  // strict: first index wins
                }
                ok = !nan && nc > 0 && best > conf_thres;
                if (ok && use_filter) ok = bidx < 256 && ((cmask.w[bidx >> 6] >> (bidx & 63)) & 1ull);
            }
            const unsigned okm = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                const double x = (double)rp[0], y = (double)rp[1], w = (double)rp[2], h = (double)rp[3];
                float4 bx;
                bx.x = (float)(x - w / 2); bx.y = (float)(y - h / 2);
                bx.z = (float)(x + w / 2); bx.w = (float)(y + h / 2);
                const long long kk = (long long)b * N + row0 + n_ok + __popc(okm & lt_mask);
                c_score[kk] = best; c_box[kk] = bx; c_meta[kk] = make_int2(row0 + r, bidx);
                bigc |= !(fabsf(bx.x) < 2048.f && fabsf(bx.y) < 2048.f && fabsf(bx.z) < 2048.f && fabsf(bx.w) < 2048.f);
            }
            n_ok += __popc(okm);
        }
        bigc = __any_sync(0xffffffffu, bigc);
        if (lane == 0) chunk_info[ch] = make_int2(n_ok | (bigc ? (int)0x80000000 : 0), n_pass);
        __syncwarp();   // every lane is done with this buffer before lane 0 lets the next copy land in it
    }
}

// ---------------------------------------------------------------------------------------------
// Pass 2: per-image ordering + greedy NMS.
// ---------------------------------------------------------------------------------------------
// Bitonic sort of npad (power of two) records by the whole CTA into (key descending, row ascending) order;
// one compare-exchange per thread and stage (thread p owns the pair i < i|j). Inlined per call site so that
// shared/global loads are typed.
__device__ __forceinline__ void bitonic_sort(unsigned long long* keys, uint2* meta, int npad) {
    const int npairs = npad >> 1;
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int p = threadIdx.x; p < npairs; p += blockDim.x) {
                const int i = ((p & ~(j - 1)) << 1) | (p & (j - 1));
                const int q = i | j;
                const unsigned long long ka = keys[i], kb = keys[q];
                const bool up = ((i & k) == 0);
                bool swap;
                if (ka != kb) swap = up ? (kb > ka) : (ka > kb);   // "before" = larger key
                else { const unsigned ra = meta[i].x, rb = meta[q].x; swap = up ? (rb < ra) : (ra < rb); }
                if (swap) {
                    const uint2 ma = meta[i], mb = meta[q];
                    keys[i] = kb; keys[q] = ka; meta[i] = mb; meta[q] = ma;
                }
            }
            __syncthreads();
        }
    }
}

// Register-resident bitonic sort for up to 1024*E records (E per thread, record i = tid + u*1024):
// strides below 32 are exchanged with warp shuffles, strides >= 1024 inside the thread, and only the
// strides in between go through shared memory (as a mailbox: every thread rewrites only its own record).
__device__ __forceinline__ bool rec_before(unsigned long long ka, unsigned ra, unsigned long long kb, unsigned rb) {
    return ka > kb || (ka == kb && ra < rb);
}

// size: number of slots sorted (a power of two); 1024 * E, or with E == 1 also 512 / 256 / 128 — the threads beyond
// size then only keep the barriers company, and a quarter-size sort has 28 instead of 55 stages.
template <int E, int size = 1024 * E>
__device__ __forceinline__ void bitonic_sort_regs(unsigned long long* skeys, uint2* smeta, int n) {
    const int tid = threadIdx.x;
    const bool act = size >= 1024 || tid < size;  // whole warps (size is a multiple of 32); compile-time true for the full sizes
    unsigned long long k[E];
    uint2 m[E];
#pragma unroll
    for (int u = 0; u < E; u++) {
        const int i = tid + u * 1024;
        if (act) { k[u] = skeys[i]; m[u] = smeta[i]; }
    }
    for (int kk = 2; kk <= size; kk <<= 1) {
        for (int j = kk >> 1; j > 0; j >>= 1) {
            if (j >= 1024) {
                const int du = j >> 10;
#pragma unroll
                for (int dc = 1; dc < E; dc <<= 1) {  // compile-time partner distance keeps k[], m[] in registers
                    if (du != dc) continue;
#pragma unroll
                    for (int u = 0; u < E; u++) {
                        if ((u & dc) == 0) {
                            const int v = u | dc;
                            const bool up = (((tid + u * 1024) & kk) == 0);
                            const bool ub = rec_before(k[u], m[u].x, k[v], m[v].x);
                            if (ub != up) {  // ascending block wants the "before" record at the lower index
                                const unsigned long long tk = k[u]; k[u] = k[v]; k[v] = tk;
                                const uint2 tm = m[u]; m[u] = m[v]; m[v] = tm;
                            }
                        }
                    }
                }
            } else {
                if (j >= 32) {
#pragma unroll
                    for (int u = 0; u < E; u++) { const int i = tid + u * 1024; if (act) { skeys[i] = k[u]; smeta[i] = m[u]; } }
                    __syncthreads();
                }
#pragma unroll
                for (int u = 0; u < E; u++) {
                    if (!act) continue;
                    const int i = tid + u * 1024;
                    unsigned long long pk;
                    uint2 pm;
                    if (j >= 32) {
                        pk = skeys[i ^ j]; pm = smeta[i ^ j];
                    } else {
                        pk = __shfl_xor_sync(0xffffffffu, k[u], j);
                        pm.x = __shfl_xor_sync(0xffffffffu, m[u].x, j);
                        pm.y = __shfl_xor_sync(0xffffffffu, m[u].y, j);
                    }
                    const bool lower = ((i & j) == 0), up = ((i & kk) == 0);
                    const bool want_before = (lower == up);
                    if (rec_before(k[u], m[u].x, pk, pm.x) != want_before) { k[u] = pk; m[u] = pm; }
                }
                if (j >= 32) __syncthreads();
            }
        }
    }
#pragma unroll
    for (int u = 0; u < E; u++) { const int i = tid + u * 1024; if (act) { skeys[i] = k[u]; smeta[i] = m[u]; } }
    __syncthreads();
}

// Ordering by counting instead of comparing-and-swapping (what the candidate sets of a detector look like: a few
// hundred to a few thousand scores spread over (conf_thres, 1]): the scores' range is cut into 1024 buckets, a shared
// histogram + one scan gives every bucket its span of the output, and a record's final position is its bucket's start
// plus the number of records of the same bucket that come before it in the full (score desc, row asc) order — so the
// result is the bitonic sort's, bit for bit, for about a tenth of the instructions. Crowded buckets (equal or tightly
// clustered scores) would make the in-bucket counting quadratic: then nothing is written and the caller sorts instead.
constexpr int BK_N = 1024;       // buckets == threads
constexpr int BK_CROWD = 64;     // give up above this many records in one bucket
struct BucketScratch { int start[BK_N + 1]; int crowded; unsigned long long kmin, kmax; unsigned short perm[SORT_SMEM_CAP]; };

template <int E>
__device__ __forceinline__ bool bucket_order(unsigned long long* skeys, uint2* smeta, int n, BucketScratch& Q, int* wsum) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long k[E];
    uint2 m[E];
    unsigned long long lo = ~0ull, hi = 0ull;
#pragma unroll
    for (int u = 0; u < E; u++) {
        const int i = tid + u * 1024;
        if (i < n) { k[u] = skeys[i]; m[u] = smeta[i]; lo = min(lo, k[u]); hi = max(hi, k[u]); }
    }
    Q.start[tid] = 0;
    if (tid == 0) { Q.crowded = 0; Q.kmin = ~0ull; Q.kmax = 0ull; }
#pragma unroll
    for (int sft = 16; sft > 0; sft >>= 1) {     // scores are positive doubles: their bit patterns order like the values
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, sft));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, sft));
    }
    __syncthreads();
    if (lane == 0 && hi >= lo) { atomicMin(&Q.kmin, lo); atomicMax(&Q.kmax, hi); }
    __syncthreads();
    const double smin = __longlong_as_double((long long)Q.kmin), smax = __longlong_as_double((long long)Q.kmax);
    const double scale = (double)BK_N / (smax - smin);
    int bk[E], ps[E];
#pragma unroll
    for (int u = 0; u < E; u++) {
        const int i = tid + u * 1024;
        if (i < n) {
            const double x = (__longlong_as_double((long long)k[u]) - smin) * scale;      // monotone in the score
            bk[u] = BK_N - 1 - (int)fmin(fmax(x, 0.0), (double)(BK_N - 1));              // higher scores first
            ps[u] = atomicAdd(&Q.start[bk[u]], 1);
        }
    }
    __syncthreads();
    const int cnt = Q.start[tid];
    int inc = cnt;
#pragma unroll
    for (int sft = 1; sft < 32; sft <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, sft); if (lane >= sft) inc += t; }
    if (lane == 31) wsum[warp] = inc;
    if (cnt > BK_CROWD) Q.crowded = 1;
    __syncthreads();
    if (warp == 0) {
        const int t = wsum[lane];
        int ti = t;
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) { const int u = __shfl_up_sync(0xffffffffu, ti, sft); if (lane >= sft) ti += u; }
        wsum[lane] = ti - t;
    }
    __syncthreads();
    Q.start[tid] = wsum[warp] + inc - cnt;
    if (tid == 0) Q.start[BK_N] = n;
    const bool crowded = Q.crowded != 0;
    __syncthreads();
    if (crowded) return false;
#pragma unroll
    for (int u = 0; u < E; u++) {
        const int i = tid + u * 1024;
        if (i < n) Q.perm[Q.start[bk[u]] + ps[u]] = (unsigned short)i;
    }
    __syncthreads();
    int rank[E];
#pragma unroll
    for (int u = 0; u < E; u++) {
        const int i = tid + u * 1024;
        if (i < n) {
            const int s0 = Q.start[bk[u]], e0 = Q.start[bk[u] + 1];
            int r = s0;
            for (int x = s0; x < e0; x++) {
                const int j = Q.perm[x];
                if (j != i && rec_before(skeys[j], smeta[j].x, k[u], m[u].x)) r++;
            }
            rank[u] = r;
        }
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < E; u++) {
        const int i = tid + u * 1024;
        if (i < n) { skeys[rank[u]] = k[u]; smeta[rank[u]] = m[u]; }
    }
    __syncthreads();
    return true;
}

constexpr int ROUND = 1024;      // candidates examined per greedy round (one per thread)
constexpr int ACAP = 64;         // survivors of a round resolved against each other at once (one mask word per row)
constexpr int RESIDENT = 4096;   // sorted candidates kept in shared memory (== SORT_SMEM_CAP: the arrays reuse the sort buffer)
constexpr int MAX_NMS_CAP = 65536;  // size of the resolved-candidate bitmap
constexpr int SMALL_BIT = 1 << 30;

// Offset box in double precision as the reference's float64 tensors hold it: x1,y1,x2,y2 (+cls*off) and area.
struct DBox { double x1, y1, x2, y2, area; };

__device__ __forceinline__ DBox make_dbox(const float4 b, const int cls, const double off) {
    const double o = (double)(cls & (SMALL_BIT - 1)) * off;
    DBox d;
    d.x1 = (double)b.x + o; d.y1 = (double)b.y + o; d.x2 = (double)b.z + o; d.y2 = (double)b.w + o;
    d.area = (d.x2 - d.x1) * (d.y2 - d.y1);
    return d;
}

// fl(inter/denom) > thr, without the division unless the quotient is within a few ulp of thr:
// r = fma(-thr, denom, inter) has the exact sign of inter - thr*denom; |r| > 4e-16*denom puts the real
// quotient more than 2 ulp away from thr (quotients are <= 1), so rounding cannot change the outcome.
__device__ __forceinline__ bool iou_exceeds(const DBox& a, const DBox& b, const double thr) {
    const double w = fmax(0.0, fmin(a.x2, b.x2) - fmax(a.x1, b.x1));
    if (w == 0.0) return false;
    const double h = fmax(0.0, fmin(a.y2, b.y2) - fmax(a.y1, b.y1));
    if (h == 0.0) return false;
    const double inter = w * h;
    const double denom = a.area + b.area - inter;
    if (denom > 0.0) {
        const double r = __fma_rn(-thr, denom, inter), tol = 4e-16 * denom;
        if (r > tol) return true;
        if (r < -tol) return false;
    }
    return inter / denom > thr;
}

// Can the earlier (higher-score) box `a` be ruled out as a suppressor of `b` without float64 work?
// torchvision's NMS (torchvision/csrc/ops/cpu/nms_kernel.cpp) runs on boxes offset by class*off
// (yolov5_postprocess.py:106-108); inter == 0 gives ovr == 0 (or NaN), never > thr for thr >= 0. Exact shortcuts:
//   same class  : offsets are equal and x -> fl(x+off) is monotonic, so min(x2) <= max(x1) in f32 implies w == 0;
//   other class : if every coordinate of both boxes is inside (-2048, 2048) the 4096-spaced offsets cannot overlap.
__device__ __forceinline__ bool cannot_overlap(const float4 a, const int ca, const float4 b, const int cb, const double off) {
    const int cla = ca & (SMALL_BIT - 1), clb = cb & (SMALL_BIT - 1);
    if (off == 0.0 || cla == clb) return fminf(a.z, b.z) <= fmaxf(a.x, b.x) || fminf(a.w, b.w) <= fmaxf(a.y, b.y);
    return (ca & cb & SMALL_BIT) != 0;
}

// Same-class (or class-agnostic) pairs: decide fl64(inter/denom) > thr from a float32 evaluation when it is not
// close. Both evaluations start from the same f32 coordinates; the f32 one carries a relative error < 2e-6
// (each of its ~10 operations rounds to 2^-24) and the reference's f64 one < 1e-9 (offset rounding), so a
// float32 margin of 2e-5*denom on inter - thr*denom is decisive with a 10x reserve. Returns 1 = suppress,
// 0 = keep, -1 = too close: evaluate in float64 exactly like the reference.
__device__ __forceinline__ int iou32_decide(const float4 a, const float4 b, const float thr) {
    const float w = fminf(a.z, b.z) - fmaxf(a.x, b.x), h = fminf(a.w, b.w) - fmaxf(a.y, b.y);  // > 0: cannot_overlap was false
    const float inter = w * h;
    const float denom = (a.z - a.x) * (a.w - a.y) + (b.z - b.x) * (b.w - b.y) - inter;
    if (!(denom > 0.f) || !(inter > 0.f)) return -1;
    const float r = inter - thr * denom, tol = 2e-5f * denom;
    return r > tol ? 1 : (r < -tol ? 0 : -1);
}

#ifdef TCB_NMS_PROFILE
__device__ long long g_nms_prof[16];
__device__ long long g_nms_cta[4096];   // per CTA: cycles, candidates, rounds, kept
#define PROF_T(k) do { __syncthreads(); if (blockIdx.x == 0 && threadIdx.x == 0) { long long t_ = clock64(); atomicAdd((unsigned long long*)&g_nms_prof[k], (unsigned long long)(t_ - prof_t0)); prof_t0 = t_; } } while (0)
#else
#define PROF_T(k) do {} while (0)
#endif

struct NmsShared {
    float4 kbox[MAX_DET_CAP]; int kcls[MAX_DET_CAP]; unsigned kslot[MAX_DET_CAP];  // kept so far
    DBox kd[MAX_DET_CAP];                  // kept boxes, class-offset, float64 (what the reference's NMS sees)
    unsigned resolved[MAX_NMS_CAP / 32];   // bit c: sorted candidate c is kept or suppressed
    unsigned long long mask[ACAP];         // bit j of row i: survivor i suppresses survivor j (i < j)
    int aidx[ACAP];                        // sorted positions of this round's survivors, in score order
    int knew[ACAP];                        // survivors the sweep kept
    int warp_cnt[32];
    int nkept, next_pos, a_total;
};
// After the sort the 64 KB sort buffer is reused for the sorted candidates themselves.
struct CandShared { float4 box[RESIDENT]; int cls[RESIDENT]; unsigned slot[RESIDENT]; };
static_assert(sizeof(BucketScratch) <= sizeof(DBox) * MAX_DET_CAP, "bucket tables overflow the kept-box array they borrow");
static_assert(sizeof(CandShared) <= sizeof(SortEnt) * SORT_SMEM_CAP + 32 * 1024, "candidate arrays overflow their region");

// Class groups (gridDim.y = G > 1): class-aware NMS of a frame whose boxes all lie inside (-2048, 2048) is G
// independent problems (classes c with c % G == g), one CTA each; every CTA leaves its kept rows, in order, in a
// staging list and the last one of the frame to finish merges the G lists by (score desc, row asc) into the
// output. Frames that cannot be split (big[b], or more candidates than max_nms: the reference cuts to the max_nms
// best over all classes first) are done whole by the g == 0 CTA; the merge is then a copy.
struct GroupOut {
    double* rows;                // [b*G][max_det][6]
    unsigned long long* key;     // [b*G][max_det] score bits of each kept row
    unsigned* krow;              // [b*G][max_det] its row in the prediction tensor
    int* count;                  // [b*G]
    int* done;                   // [b] CTAs of the frame that have finished (zeroed per call)
    const int* big;              // [b]
};

__global__ void __launch_bounds__(NMS_THREADS, 1) yolo_nms_kernel(
    // (no __restrict__ on the filter's outputs: this kernel can be resident before the filter has finished, and the
    //  non-coherent loads the qualifier allows are only valid for data that never changes during a kernel's lifetime)
    int N, const double* c_score, const float4* c_box, const int2* c_meta,
    const int* c_count, SortEnt* __restrict__ gsort, int gsort_stride,
    double iou_thres, double class_offset, int max_det, int max_nms,
    double* __restrict__ out, int* __restrict__ out_count, const GroupOut go,
    const int* xc_count, int* __restrict__ xc_out, const int2* chunk_info, int chunks, int chunk_rows) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    NmsShared& S = *reinterpret_cast<NmsShared*>(smem_raw);
    unsigned long long* skeys = reinterpret_cast<unsigned long long*>(smem_raw + sizeof(NmsShared));
    uint2* smeta = reinterpret_cast<uint2*>(skeys + SORT_SMEM_CAP);
    CandShared& C = *reinterpret_cast<CandShared*>(smem_raw + sizeof(NmsShared));

    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_enter();
#ifdef TCB_NMS_PROFILE
    long long prof_t0 = clock64();
    const long long prof_start = prof_t0;
    int prof_rounds = 0;
#endif
    const int G = gridDim.y, g = blockIdx.y;
    const long long cbase = (long long)b * N;
    if (tid == 0) { S.nkept = 0; S.a_total = 0; }

    // ---- this frame's candidates. Two layouts: a dense list of c_count[b] entries (general filter), or one region
    //      per chunk of chunk_rows rows, chunk_info[b][c] = (count | big << 31, rows passing the objectness test), candidates
    //      of chunk c at slots chunk_rows*c + 0, 1, ... (streaming filter: nothing to zero, nothing reserved atomically) ----
    int n_all, xc_total = 0;
    bool big;
    const int2* cinfo = chunk_info ? chunk_info + (long long)b * chunks : nullptr;
    int* cprefix = reinterpret_cast<int*>(S.resolved);   // chunks + 1 exclusive prefix sums of the chunk counts (chunks < 2048);
                                                         // the bitmap proper is cleared after the gather
    if (cinfo) {
        // thread t owns chunks 2t, 2t+1: one coalesced read of the table, then a CTA-wide scan
        const int c0 = 2 * tid, c1 = 2 * tid + 1;
        const int2 v0 = c0 < chunks ? cinfo[c0] : make_int2(0, 0), v1 = c1 < chunks ? cinfo[c1] : make_int2(0, 0);
        const int a0 = v0.x & 0x7fffffff, a1 = v1.x & 0x7fffffff;
        int inc = a0 + a1, x = v0.y + v1.y, bg = (int)(((unsigned)v0.x | (unsigned)v1.x) >> 31);
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, sft); if (lane >= sft) inc += t; }
#pragma unroll
        for (int sft = 16; sft > 0; sft >>= 1) { x += __shfl_xor_sync(0xffffffffu, x, sft); bg |= __shfl_xor_sync(0xffffffffu, bg, sft); }
        if (lane == 31) S.warp_cnt[warp] = inc;
        if (lane == 0) { S.aidx[warp] = x; S.knew[warp] = bg; }
        __syncthreads();
        if (warp == 0) {   // one warp combines the 32 warp totals (every thread walking them costs 3 K shared-memory reads per CTA)
            const int t = S.warp_cnt[lane];
            int ti = t, xw = S.aidx[lane], bw = S.knew[lane];
#pragma unroll
            for (int sft = 1; sft < 32; sft <<= 1) { const int u = __shfl_up_sync(0xffffffffu, ti, sft); if (lane >= sft) ti += u; }
#pragma unroll
            for (int sft = 16; sft > 0; sft >>= 1) { xw += __shfl_xor_sync(0xffffffffu, xw, sft); bw |= __shfl_xor_sync(0xffffffffu, bw, sft); }
            __syncwarp();
            S.warp_cnt[lane] = ti - t;                                  // candidates in the warps before this one
            if (lane == 31) { S.knew[0] = ti; S.knew[1] = xw; S.knew[2] = bw; }
        }
        __syncthreads();
        const int before = S.warp_cnt[warp], tot = S.knew[0], xs = S.knew[1], bgs = S.knew[2];
        const int excl = before + inc - (a0 + a1);
        if (c0 <= chunks) cprefix[c0] = excl;            // entry `chunks` (the total) is written by whoever owns it
        if (c1 <= chunks) cprefix[c1] = excl + a0;
        n_all = tot; xc_total = xs; big = bgs != 0;
    } else {
        n_all = min(c_count[b], N);
        xc_total = xc_count ? xc_count[b] : 0;
        big = G > 1 && go.big[b] != 0;
    }
    const bool split = G > 1 && !big && n_all <= max_nms;
    const int n_bound = (G > 1 && !split && g > 0) ? 0 : n_all;   // an unsplit frame is done whole by its g == 0 CTA

    // ---- order candidates: score descending, row ascending ------------------------------------
    int npad = 1;
    while (npad < n_bound) npad <<= 1;
    const bool sort_in_smem = npad <= SORT_SMEM_CAP;
    // global sort scratch of this CTA: gsort_stride keys followed by gsort_stride metas
    unsigned long long* gkeys = reinterpret_cast<unsigned long long*>(gsort + ((long long)b * G + g) * gsort_stride);
    uint2* gmeta = reinterpret_cast<uint2*>(gkeys + gsort_stride);
    unsigned long long* keys = sort_in_smem ? skeys : gkeys;
    uint2* meta = sort_in_smem ? smeta : gmeta;
    __syncthreads();
    PROF_T(8);
    // gather (any order: the sort fixes it), keeping this CTA's classes when the frame is split; both loads of a
    // candidate are issued together and a warp reserves its output slots with one shared-memory atomic
    if (n_bound > 0) {
        for (int i0 = 0; i0 < n_all; i0 += NMS_THREADS) {
            const int i = i0 + tid;
            int slot = 0;
            bool take = i < n_all;
            if (take && cinfo) {                               // candidate i of the frame -> its chunk by bisection of the prefix sums
                int lo = 0, hi = chunks;                       // invariant: cprefix[lo] <= i < cprefix[hi]
                while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (cprefix[mid] <= i) lo = mid; else hi = mid; }
                slot = lo * chunk_rows + (i - cprefix[lo]);
            } else if (take) {
                slot = i;
            }
            int2 m = make_int2(0, 0);
            double sc = 0.0;
            if (take) { m = c_meta[cbase + slot]; sc = c_score[cbase + slot]; }
            take = take && (!split || (m.y % G) == g);
            const unsigned bal = __ballot_sync(0xffffffffu, take);
            int base = 0;
            if (lane == 0 && bal) base = atomicAdd(&S.a_total, __popc(bal));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (take) {
                const int q = base + __popc(bal & ((1u << lane) - 1u));
                keys[q] = (unsigned long long)__double_as_longlong(sc);
                meta[q] = make_uint2((unsigned)m.x, (unsigned)slot);
            }
        }
    }
    __syncthreads();
    PROF_T(9);
    for (int i = tid; i < MAX_NMS_CAP / 32; i += NMS_THREADS) S.resolved[i] = 0u;   // (the prefix sums lived here)
    int n = S.a_total;
    npad = 1;
    while (npad < n) npad <<= 1;
    const int nfill = sort_in_smem ? (npad <= 128 ? 128 : (npad <= 1024 ? npad : (npad <= 2048 ? 2048 : 4096))) : npad;
    for (int i = n + tid; i < nfill; i += NMS_THREADS) { keys[i] = 0ull; meta[i] = make_uint2(0xffffffffu, 0u); }
    __syncthreads();
    PROF_T(0);
    if (n < 2) {
        // nothing to order
    } else if (sort_in_smem) {
        // the kept-box array is idle until the greedy rounds: its first 12 KB hold the bucket tables
        BucketScratch& Q = *reinterpret_cast<BucketScratch*>(S.kd);
        bool done;
        if (nfill <= 1024) done = bucket_order<1>(skeys, smeta, n, Q, S.warp_cnt);
        else if (nfill == 2048) done = bucket_order<2>(skeys, smeta, n, Q, S.warp_cnt);
        else done = bucket_order<4>(skeys, smeta, n, Q, S.warp_cnt);
        if (done) {
        } else if (nfill == 128) bitonic_sort_regs<1, 128>(skeys, smeta, n);
        else if (nfill == 256) bitonic_sort_regs<1, 256>(skeys, smeta, n);
        else if (nfill == 512) bitonic_sort_regs<1, 512>(skeys, smeta, n);
        else if (nfill == 1024) bitonic_sort_regs<1>(skeys, smeta, n);
        else if (nfill == 2048) bitonic_sort_regs<2>(skeys, smeta, n);
        else bitonic_sort_regs<4>(skeys, smeta, n);
    } else {
        bitonic_sort(gkeys, gmeta, npad);
    }
    PROF_T(1);
    if (n > max_nms) n = max_nms;

    // ---- make the sorted candidates resident: box, class (+SMALL flag), slot ----------------------
    // resident mode (n <= RESIDENT): C.*[c] is sorted candidate c, loaded once (over the dead sort buffer);
    // streaming mode: C.*[t] is candidate pos+t of the current round, gathered every round.
    const bool resident = n <= RESIDENT && sort_in_smem;
    if (resident) {
        unsigned myslot[RESIDENT / NMS_THREADS];
#pragma unroll
        for (int u = 0; u < RESIDENT / NMS_THREADS; u++) {
            const int c = tid + u * NMS_THREADS;
            myslot[u] = c < n ? smeta[c].y : 0u;
        }
        __syncthreads();  // every slot is in registers; the sort buffer may now be overwritten
#pragma unroll
        for (int u = 0; u < RESIDENT / NMS_THREADS; u++) {
            const int c = tid + u * NMS_THREADS;
            if (c < n) {
                const float4 bx = c_box[cbase + myslot[u]];
                int cl = c_meta[cbase + myslot[u]].y;
                if (fabsf(bx.x) < 2048.f && fabsf(bx.y) < 2048.f && fabsf(bx.z) < 2048.f && fabsf(bx.w) < 2048.f) cl |= SMALL_BIT;
                C.box[c] = bx; C.cls[c] = cl; C.slot[c] = myslot[u];
            }
        }
    }
    __syncthreads();

    // ---- greedy NMS in rounds ---------------------------------------------------------------------
    const float thr32 = (float)iou_thres;
    int pos = 0, prev_end = 0, nk_prev = 0;
    while (pos < n) {
        const int cn = min(ROUND, n - pos);
        const int c = pos + tid;               // sorted position of this thread's candidate
        const int ci = resident ? c : tid;     // where it lives in C
        if (!resident) {
            if (tid < cn) {
                const unsigned slot = meta[c].y;
                const float4 bx = c_box[cbase + slot];
                int cl = c_meta[cbase + slot].y;
                if (fabsf(bx.x) < 2048.f && fabsf(bx.y) < 2048.f && fabsf(bx.z) < 2048.f && fabsf(bx.w) < 2048.f) cl |= SMALL_BIT;
                C.box[tid] = bx; C.cls[tid] = cl; C.slot[tid] = slot;
            }
            __syncthreads();
        }
        PROF_T(2);
        // (1) against the boxes kept since this candidate was last examined
        const int nk = S.nkept;
        bool alive = tid < cn && !((S.resolved[c >> 5] >> (c & 31)) & 1u);
        if (alive) {
            const float4 bx = C.box[ci];
            const int cl = C.cls[ci];
            DBox dbx;
            bool have_d = false;
            for (int k = (c < prev_end) ? nk_prev : 0; k < nk; k++) {
                const float4 kb = S.kbox[k];
                const int kc = S.kcls[k];
                if (cannot_overlap(kb, kc, bx, cl, class_offset)) continue;
                int dec = -1;
                if (class_offset == 0.0 || ((kc ^ cl) & (SMALL_BIT - 1)) == 0) dec = iou32_decide(kb, bx, thr32);
                if (dec < 0) {
                    if (!have_d) { dbx = make_dbox(bx, cl, class_offset); have_d = true; }
                    dec = iou_exceeds(S.kd[k], dbx, iou_thres) ? 1 : 0;
                }
                if (dec) { alive = false; break; }
            }
            if (!alive) atomicOr(&S.resolved[c >> 5], 1u << (c & 31));
        }
        // ordered compaction of the survivors
        const unsigned bal = __ballot_sync(0xffffffffu, alive);
        if (lane == 0) S.warp_cnt[warp] = __popc(bal);
        __syncthreads();
        if (warp == 0) {
            int w = S.warp_cnt[lane];
#pragma unroll
            for (int k = 1; k < 32; k <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, k); if (lane >= k) w += t; }
            S.warp_cnt[lane] = w;
            if (lane == 31) { S.a_total = w; S.next_pos = pos + cn; }
        }
        __syncthreads();
        const int rank = (warp ? S.warp_cnt[warp - 1] : 0) + __popc(bal & ((1u << lane) - 1));
        if (alive) {
            if (rank < ACAP) S.aidx[rank] = c;
            else if (rank == ACAP) S.next_pos = c;  // survivors beyond ACAP wait for the next round
        }
        __syncthreads();
        const int A = min(S.a_total, ACAP);
        PROF_T(3);
        // (2) suppression mask among the A survivors: 16 threads per row, 4 columns each
        {
            const int ri = tid >> 4, l16 = tid & 15;
            unsigned long long bits = 0ull;
            if (ri < A) {
                const int pi = resident ? S.aidx[ri] : S.aidx[ri] - pos;
                const float4 bi = C.box[pi];
                const int cli = C.cls[pi];
                DBox di;
                bool have_d = false;
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const int j = l16 * 4 + q;
                    if (j > ri && j < A) {
                        const int pj = resident ? S.aidx[j] : S.aidx[j] - pos;
                        const float4 bj = C.box[pj];
                        const int clj = C.cls[pj];
                        if (!cannot_overlap(bi, cli, bj, clj, class_offset)) {
                            int dec = -1;
                            if (class_offset == 0.0 || ((cli ^ clj) & (SMALL_BIT - 1)) == 0) dec = iou32_decide(bi, bj, thr32);
                            if (dec < 0) {
                                if (!have_d) { di = make_dbox(bi, cli, class_offset); have_d = true; }
                                dec = iou_exceeds(di, make_dbox(bj, clj, class_offset), iou_thres) ? 1 : 0;
                            }
                            if (dec) bits |= 1ull << j;
                        }
                    }
                }
            }
#pragma unroll
            for (int sft = 1; sft < 16; sft <<= 1) bits |= __shfl_xor_sync(0xffffffffu, bits, sft);
            if (l16 == 0 && ri < ACAP) S.mask[ri] = bits;
        }
        __syncthreads();
        PROF_T(4);
        // (3) serial sweep: hop from kept survivor to kept survivor
        if (tid == 0) {
            unsigned long long avail = A >= 64 ? ~0ull : ((1ull << A) - 1);
            int kept = nk;
            while (kept < max_det && avail) {
                const int i = __ffsll((long long)avail) - 1;
                avail &= ~(S.mask[i] | (1ull << i));
                S.knew[kept - nk] = i;
                kept++;
            }
            S.nkept = kept;
        }
        __syncthreads();
        const int nk_new = S.nkept;
        if (tid < nk_new - nk) {  // append the kept survivors
            const int cs = S.aidx[S.knew[tid]];
            const int pi = resident ? cs : cs - pos;
            const float4 bx = C.box[pi];
            const int cl = C.cls[pi];
            S.kbox[nk + tid] = bx; S.kcls[nk + tid] = cl; S.kslot[nk + tid] = C.slot[pi];
            S.kd[nk + tid] = make_dbox(bx, cl, class_offset);
        }
        if (tid < A) { const int cs = S.aidx[tid]; atomicOr(&S.resolved[cs >> 5], 1u << (cs & 31)); }  // all A are decided now
        prev_end = pos + cn; nk_prev = nk;
        const int np = S.next_pos;
        __syncthreads();
        PROF_T(5);
#ifdef TCB_NMS_PROFILE
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd((unsigned long long*)&g_nms_prof[7], 1000000000ull);
        prof_rounds++;
#endif
        if (nk_new >= max_det) break;
        pos = np;
    }
    __syncthreads();
    // ---- write the kept rows: [x1,y1,x2,y2 (f32 box widened), conf (f64), cls] -------------------------
    const long long list = (long long)b * G + g;
    double* orow = G > 1 ? go.rows + list * max_det * 6 : out + (long long)b * max_det * 6;
    const int nk_final = S.nkept;
    for (int k = tid; k < nk_final; k += NMS_THREADS) {
        const unsigned slot = S.kslot[k];
        const float4 bx = S.kbox[k];
        double* o = orow + (long long)k * 6;
        const double sc = c_score[cbase + slot];
        o[0] = (double)bx.x; o[1] = (double)bx.y; o[2] = (double)bx.z; o[3] = (double)bx.w;
        o[4] = sc;
        o[5] = (double)(S.kcls[k] & (SMALL_BIT - 1));
        if (G > 1) {
            go.key[list * max_det + k] = (unsigned long long)__double_as_longlong(sc);
            go.krow[list * max_det + k] = (unsigned)c_meta[cbase + slot].x;
        }
    }
    if (G == 1) {
        if (tid == 0) { out_count[b] = nk_final; if (xc_out) xc_out[b] = xc_total; }
    } else {
        if (tid == 0) go.count[list] = nk_final;
        PROF_T(10);
        __threadfence();
        __syncthreads();
        if (tid == 0) S.next_pos = atomicAdd(&go.done[b], 1);
        __syncthreads();
        PROF_T(11);
        if (S.next_pos == G - 1) {   // last CTA of the frame: merge the G ordered lists
            __threadfence();
            // the G lists' sort keys come into shared memory with one round trip (the candidate arrays are dead by now);
            // the bisections then run there instead of as chains of dependent L2 reads
            unsigned long long* mk = reinterpret_cast<unsigned long long*>(smem_raw + sizeof(NmsShared));
            unsigned* mr = reinterpret_cast<unsigned*>(mk + G * max_det);
            if (tid < G) S.warp_cnt[tid] = __ldcg(&go.count[(long long)b * G + tid]);
            __syncthreads();
            int total = 0;
            for (int q = 0; q < G; q++) total += S.warp_cnt[q];
            for (int e = tid; e < G * max_det; e += NMS_THREADS) {
                const int q = e / max_det, i = e - q * max_det;
                if (i < S.warp_cnt[q]) {
                    const long long lq = (long long)b * G + q;
                    mk[e] = __ldcg(&go.key[lq * max_det + i]);
                    mr[e] = __ldcg(&go.krow[lq * max_det + i]);
                }
            }
            __syncthreads();
            for (int e = tid; e < G * max_det; e += NMS_THREADS) {
                const int q = e / max_det, i = e - q * max_det;
                if (i >= S.warp_cnt[q]) continue;
                const unsigned long long ke = mk[e];
                const unsigned re = mr[e];
                int rank = i;
                for (int o2 = 0; o2 < G; o2++) {
                    if (o2 == q) continue;
                    int lo = 0, hi = S.warp_cnt[o2];   // entries of list o2 that come before e
                    while (lo < hi) {
                        const int mid = (lo + hi) >> 1;
                        if (rec_before(mk[o2 * max_det + mid], mr[o2 * max_det + mid], ke, re)) lo = mid + 1;
                        else hi = mid;
                    }
                    rank += lo;
                }
                if (rank < max_det) {
                    const double* src = go.rows + (((long long)b * G + q) * max_det + i) * 6;
                    double* dst = out + ((long long)b * max_det + rank) * 6;
#pragma unroll
                    for (int c = 0; c < 6; c++) dst[c] = __ldcg(src + c);
                }
            }
            if (tid == 0) { out_count[b] = min(total, max_det); if (xc_out) xc_out[b] = xc_total; go.done[b] = 0; }   // the ticket resets itself
        }
    }
    PROF_T(6);
#ifdef TCB_NMS_PROFILE
    if (blockIdx.x == 0 && threadIdx.x == 0) { atomicAdd((unsigned long long*)&g_nms_prof[7], (unsigned long long)nk_final * 1000000ull + (unsigned long long)n); }
    if (threadIdx.x == 0) {
        const int cta = (blockIdx.x * gridDim.y + blockIdx.y) & 1023;
        g_nms_cta[cta * 4 + 0] = clock64() - prof_start; g_nms_cta[cta * 4 + 1] = n; g_nms_cta[cta * 4 + 2] = prof_rounds; g_nms_cta[cta * 4 + 3] = nk_final;
    }
#endif
}
#ifdef TCB_NMS_PROFILE
}  // namespace
extern "C" void tcb_debug_nms_cta(long long* out4096) { cudaMemcpyFromSymbol(out4096, g_nms_cta, sizeof(long long) * 4096); }
extern "C" void tcb_debug_nms_prof(long long* out8, int reset) {
    cudaMemcpyFromSymbol(out8, g_nms_prof, sizeof(long long) * 16);
    if (reset) { long long z[16] = {0}; cudaMemcpyToSymbol(g_nms_prof, z, sizeof(z)); }
}
namespace {
#endif

constexpr int MAX_GROUPS = 8;   // class groups per frame (CTAs sharing one frame's NMS)
constexpr size_t NMS_SMEM_BYTES = sizeof(NmsShared) + (sizeof(CandShared) > sizeof(SortEnt) * SORT_SMEM_CAP ? sizeof(CandShared) : sizeof(SortEnt) * SORT_SMEM_CAP);

size_t yolo_ws_bytes(int b, int n) {
    size_t per = (size_t)b * n;
    size_t bytes = 256 * 8;
    bytes += per * sizeof(double) + 256;
    bytes += per * sizeof(float4) + 256;
    bytes += per * sizeof(int2) + 256;
    bytes += (size_t)b * 4 * sizeof(int) + 256;
    bytes += (size_t)b * ((n + 31) / 32) * sizeof(int2) + 256;
    int npad = 1;
    while (npad < n) npad <<= 1;
    // class groups: b * G lists with G <= SMs / b, so never more than max(b, SM count) lists (256 bounds today's parts)
    size_t lists = (size_t)b * MAX_GROUPS;
    if (lists > 256) lists = b > 256 ? (size_t)b : 256;
    if (npad > SORT_SMEM_CAP) bytes += lists * npad * sizeof(SortEnt) + 256;
    bytes += lists * ((size_t)MAX_DET_CAP * (6 * 8 + 8 + 4) + 4) + 4 * 256;   // staging lists
    return bytes;
}

}  // namespace

extern "C" size_t tcb_yolo_nms_workspace(int b, int n) { return (b > 0 && n > 0) ? yolo_ws_bytes(b, n) : 0; }

static int init_yolo_post(tcb_ctx* ctx) {   // run by tcb_create
    TCB_SMEM_OPT_IN(ctx, yolo_filter_stream_kernel<128>, TCB_SMEM_LIMIT);
    TCB_SMEM_OPT_IN(ctx, yolo_filter_stream_kernel<32>, TCB_SMEM_LIMIT);
    TCB_SMEM_OPT_IN(ctx, yolo_nms_kernel, NMS_SMEM_BYTES);
    return TCB_OK;
}
static TcbInitReg reg_yolo_post(&init_yolo_post);

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
    if (max_nms > MAX_NMS_CAP) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_yolo_nms: max_nms %d > compiled limit %d", max_nms, MAX_NMS_CAP);
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
    // class groups: as many CTAs per frame as the SM count allows, never more than classes; class-agnostic NMS is one problem
    int G = 1;
    if (!agnostic && nc > 1) {
        G = ctx->num_sms / b;
        if (G > nc) G = nc;
        if (G > MAX_GROUPS) G = MAX_GROUPS;
        if (G < 1) G = 1;
    }
    TCB_REQUIRE(ctx, b <= 65535 && b <= TCB_TICKETS, "tcb_yolo_nms: batch %d > 65535", b);
    // Streaming filter with per-chunk candidate regions: every byte of the tensor is read, which costs nothing extra for
    // small rows (<= 13 floats: every sector is touched anyway, 128-row chunks) and is the faster way for wide rows as
    // long as the tensor is small (a COCO head of one frame is 8.6 MB: 2 us of HBM time against a chain of sparse
    // sector fetches; 32-row chunks). Large batches of wide rows keep the sparse filter. The NMS kernel holds
    // chunks + 1 prefix sums in a 2048-int table.
    int chunk_rows = 0;
    if (nc <= FILTER_SMALL_NC) chunk_rows = 128;
    else if (nc <= 123 && (long long)b * n * (5 + nc) * 4 <= (64ll << 20)) chunk_rows = 32;
    const int chunks = chunk_rows ? (n + chunk_rows - 1) / chunk_rows : 0;
    const bool stream_filter = chunk_rows != 0 && n > 0 && chunks < MAX_NMS_CAP / 32;
    TCB_WS(ctx, counts, int, (size_t)b * 4);
    int2* chunk_info = nullptr;
    if (stream_filter) {
        TCB_WS(ctx, ci, int2, (size_t)b * chunks);
        chunk_info = ci;
    }
    TCB_WS(ctx, c_score, double, per);
    TCB_WS(ctx, c_box, float4, per);
    TCB_WS(ctx, c_meta, int2, per);
    int npad = 1;
    while (npad < n) npad <<= 1;
    SortEnt* gsort = nullptr;
    if (npad > SORT_SMEM_CAP) {
        TCB_WS(ctx, gs, SortEnt, (size_t)b * G * npad);
        gsort = gs;
    }
    int* c_count = counts;
    int* xc_count = counts + b;
    int* big = counts + 2 * (size_t)b;
    GroupOut go;
    memset(&go, 0, sizeof(go));
    if (G > 1) {
        TCB_WS(ctx, g_rows, double, (size_t)b * G * max_det * 6);
        TCB_WS(ctx, g_key, unsigned long long, (size_t)b * G * max_det);
        TCB_WS(ctx, g_krow, unsigned, (size_t)b * G * max_det);
        TCB_WS(ctx, g_count, int, (size_t)b * G);
        go.rows = g_rows; go.key = g_key; go.krow = g_krow; go.count = g_count; go.done = ctx->tickets; go.big = big;
    }
    // dense-list layout only: its three per-frame counters start at zero (the chunk layout has nothing to clear, and
    // the class-group tickets reset themselves)
    if (!stream_filter) TCB_CUDA(ctx, cudaMemsetAsync(counts, 0, sizeof(int) * 3 * (size_t)b, st));

    if (n > 0) {
        // 8 warps per CTA, FILTER_UNROLL 32-row groups per warp and pass; y = image
        int bx = tcb_div_up((n + 31) / 32, 8 * FILTER_UNROLL);
        const int cap = tcb_div_up(ctx->num_sms * 16, b);
        if (bx > cap) bx = cap;
        if (bx < 1) bx = 1;
        if (stream_filter) {
            // FS_NBUF row buffers per warp; as many CTAs as fit, each warp streaming its share of the b * chunks chunks
            const size_t fsm = 8 * FS_WARPS * FS_NBUF + (size_t)FS_WARPS * chunk_rows + (size_t)FS_WARPS * FS_NBUF * chunk_rows * (5 + nc) * sizeof(float);
            auto kernel = chunk_rows == 128 ? yolo_filter_stream_kernel<128> : yolo_filter_stream_kernel<32>;
            if (fsm > (size_t)TCB_SMEM_LIMIT) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_yolo_nms: %d classes exceed the filter's shared-memory rows", nc);
            const int per_sm = (int)((220 * 1024) / (fsm + 1024)) < 4 ? (int)((220 * 1024) / (fsm + 1024)) : 4;
            const long long chunks_total = (long long)b * chunks;
            if (chunks_total >= (1ll << 30)) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_yolo_nms: b*n too large (%d x %d)", b, n);
            long long grid = tcb_div_up(chunks_total, FS_WARPS);
            if (grid > (long long)ctx->num_sms * per_sm) grid = (long long)ctx->num_sms * per_sm;
            float conf_floor32 = (float)conf_thres;   // objectness test in float32, exactly equivalent to the float64 one
            if ((double)conf_floor32 > conf_thres) conf_floor32 = nextafterf(conf_floor32, -INFINITY);
            TCB_CUDA(ctx, tcb_launch(ctx->pdl != 0, kernel, dim3((unsigned)grid), dim3(32 * FS_WARPS), fsm, st,
                                     pred_dev, b, n, nc, conf_thres, conf_floor32, (int)(n_classes > 0), cm, c_score, c_box, c_meta, chunk_info));
        } else {
            yolo_filter_kernel<<<dim3(bx, b), 256, 0, st>>>(pred_dev, b, n, nc, conf_thres, n_classes > 0, cm,
                                                            c_score, c_box, c_meta, c_count, xc_count, big);
        }
        TCB_LAUNCHED(ctx);
    }
    const size_t smem = NMS_SMEM_BYTES;
    TCB_CUDA(ctx, tcb_launch(ctx->pdl != 0, yolo_nms_kernel, dim3(b, G), dim3(NMS_THREADS), smem, st, n > 0 ? n : 1, c_score, c_box, c_meta,
                             c_count, gsort, npad, iou_thres, agnostic ? 0.0 : 4096.0, max_det, max_nms, out_dev, count_dev, go, xc_count,
                             xc_count_dev, chunk_info, chunks, chunk_rows));
    TCB_LAUNCHED(ctx);
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
                                 (int)NMS_SMEM_BYTES);
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
    if (max_nms > MAX_NMS_CAP) return tcb_fail(nullptr, TCB_ERR_LIMIT, "tcb_yolo_nms_static: max_nms %d > compiled limit %d", max_nms, MAX_NMS_CAP);
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
        int bx = tcb_div_up((n + 31) / 32, 8 * FILTER_UNROLL);
        const int cap = tcb_div_up(num_sms * 16, b);
        if (bx > cap) bx = cap;
        if (bx < 1) bx = 1;
        yolo_filter_kernel<<<dim3(bx, b), 256, 0, st>>>(pred_dev, b, n, nc, conf_thres, n_classes > 0, cm,
                                                        c_score, c_box, c_meta, c_count, xc_count, nullptr);
        ST_CUDA(cudaGetLastError());
    }
    const size_t smem = NMS_SMEM_BYTES;
    GroupOut one_group;   // one CTA per frame: the fixed scratch has no room for class-group staging
    memset(&one_group, 0, sizeof(one_group));
    yolo_nms_kernel<<<b, NMS_THREADS, smem, st>>>(n > 0 ? n : 1, c_score, c_box, c_meta, c_count, gsort, npad,
                                                  iou_thres, agnostic ? 0.0 : 4096.0, max_det, max_nms, out_dev, count_dev, one_group,
                                                  xc_count, xc_count_dev, nullptr, 0, 0);
    ST_CUDA(cudaGetLastError());
    return TCB_OK;
#undef ST_REQUIRE
#undef ST_CUDA
}
