// CTA-wide building blocks for the score-ordering steps of the 3D post path (one 1024-thread CTA per frame):
//   * radix select of the K largest 64-bit keys out of n (keys unique: score bits in the high word,
//     ~index in the low word, so "largest" means highest score, then lowest index),
//   * register/shuffle bitonic sort of up to 4096 keys, descending.
// They replace torch.topk / Tensor.sort in OPCD/pcdet/models/model_utils/model_nms_utils.py:15,
// OPCD/pcdet/ops/iou3d_nms/iou3d_nms_utils.py:92 and OPCD/pcdet/models/model_utils/centernet_utils.py:139-145,
// with the tie rule the build fixes everywhere: equal scores -> lower index first.
#pragma once
#include <cuda_runtime.h>

namespace sel {

constexpr int THREADS = 1024;
constexpr int KCAP = 4096;   // most keys kept/sorted in shared memory
constexpr int BINS = 2048;   // 11-bit digits

// monotonic map float -> uint (works for negative scores too)
__device__ __forceinline__ unsigned f32_ordered(float f) {
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float f32_from_ordered(unsigned o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}
__device__ __forceinline__ unsigned long long make_key(float score, unsigned idx) {
    return ((unsigned long long)f32_ordered(score) << 32) | (unsigned long long)(0xffffffffu - idx);
}
__device__ __forceinline__ unsigned key_index(unsigned long long k) { return 0xffffffffu - (unsigned)(k & 0xffffffffull); }
__device__ __forceinline__ float key_score(unsigned long long k) { return f32_from_ordered((unsigned)(k >> 32)); }

struct Scratch {
    int hist[BINS];
    int warp_tot[32];
    int digit, above, count;
};

// exclusive prefix sum of one int per thread over the CTA (thread order); total returned to everyone
__device__ __forceinline__ int cta_exclusive_scan(int v, int* warp_tot, int& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, s); if (lane >= s) inc += t; }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = warp_tot[lane];
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, s); if (lane >= s) w += t; }
        warp_tot[lane] = w;
    }
    __syncthreads();
    total = warp_tot[31];
    const int r = (warp ? warp_tot[warp - 1] : 0) + inc - v;
    __syncthreads();
    return r;
}

// Descending bitonic sort of 1024*E keys held E per thread (key i = tid + u*1024); strides < 32 by shuffle,
// >= 1024 inside the thread, the rest through shared memory used as a mailbox.
template <int E>
__device__ __forceinline__ void sort_desc(unsigned long long* skeys) {
    const int tid = threadIdx.x;
    unsigned long long k[E];
#pragma unroll
    for (int u = 0; u < E; u++) k[u] = skeys[tid + u * 1024];
    for (int kk = 2; kk <= 1024 * E; kk <<= 1) {
        for (int j = kk >> 1; j > 0; j >>= 1) {
            if (j >= 1024) {
                const int du = j >> 10;
#pragma unroll
                for (int dc = 1; dc < E; dc <<= 1) {
                    if (du != dc) continue;
#pragma unroll
                    for (int u = 0; u < E; u++)
                        if ((u & dc) == 0) {
                            const int v = u | dc;
                            const bool up = (((tid + u * 1024) & kk) == 0);
                            if ((k[u] > k[v]) != up) { const unsigned long long t = k[u]; k[u] = k[v]; k[v] = t; }
                        }
                }
            } else {
                if (j >= 32) {
#pragma unroll
                    for (int u = 0; u < E; u++) skeys[tid + u * 1024] = k[u];
                    __syncthreads();
                }
#pragma unroll
                for (int u = 0; u < E; u++) {
                    const int i = tid + u * 1024;
                    const unsigned long long pk = (j >= 32) ? skeys[i ^ j] : __shfl_xor_sync(0xffffffffu, k[u], j);
                    const bool lower = ((i & j) == 0), up = ((i & kk) == 0);
                    if ((k[u] > pk) != (lower == up)) k[u] = pk;  // keys are unique or both zero padding
                }
                if (j >= 32) __syncthreads();
            }
        }
    }
#pragma unroll
    for (int u = 0; u < E; u++) skeys[tid + u * 1024] = k[u];
    __syncthreads();
}

__device__ __forceinline__ void sort_desc_n(unsigned long long* skeys, int m) {  // m valid keys, rest must be 0
    if (m <= 1024) sort_desc<1>(skeys);
    else if (m <= 2048) sort_desc<2>(skeys);
    else sort_desc<4>(skeys);
}

// The K (<= KCAP) largest of the n keys key_of(i), sorted descending into skeys[0..m), m = min(n, K).
// skeys must hold KCAP entries. Every thread of the 1024-thread CTA must call this.
// cache (optional, shared memory, cache_cap entries): the first cache_cap keys are fetched once, with all loads in
// flight together, and the seven passes of the selection read them from there instead of paying the global-memory
// latency of key_of on every pass.
template <typename KeyFn>
__device__ int cta_topk_sorted(KeyFn key_of_src, int n, int K, unsigned long long* skeys, Scratch& S,
                               unsigned long long* cache = nullptr, int cache_cap = 0, int list_cap = KCAP) {
    const int tid = threadIdx.x;
    const int cached = (cache != nullptr && n > K) ? (n < cache_cap ? n : cache_cap) : 0;
    if (cached) {
#pragma unroll 8
        for (int i = tid; i < cached; i += THREADS) cache[i] = key_of_src(i);
        __syncthreads();
    }
    auto key_of = [&](int i) { return i < cached ? cache[i] : key_of_src(i); };
    const int m = n < K ? n : K;
    const int fill = m <= 1024 ? 1024 : (m <= 2048 ? 2048 : 4096);
    if (n <= K) {
        for (int i = tid; i < fill; i += THREADS) skeys[i] = i < n ? key_of(i) : 0ull;
        __syncthreads();
        sort_desc_n(skeys, m);
        return m;
    }
    // radix select of the K-th largest key, 11 bits at a time from the top (last digit 9 bits)
    // As soon as the bin holding the K-th key has no more than list_cap (<= KCAP) keys, those keys (the only ones still undecided) are
    // copied into skeys and the remaining passes run over that short list instead of all n keys.
    unsigned long long prefix = 0ull, pmask = 0ull;
    int remaining = K;
    int listed = -1;   // >= 0: number of undecided keys held in skeys
    for (int shift = 53; shift >= -2; shift -= 11) {
        const int sh = shift < 0 ? 0 : shift;
        const int bits = shift < 0 ? 9 : 11;
        const unsigned long long dmask = (1ull << bits) - 1;
        for (int i = tid; i < BINS; i += THREADS) S.hist[i] = 0;
        __syncthreads();
        if (listed >= 0) {
            for (int i = tid; i < listed; i += THREADS) {
                const unsigned long long k = skeys[i];
                if ((k & pmask) == prefix) atomicAdd(&S.hist[(int)((k >> sh) & dmask)], 1);
            }
        } else {
            for (int i = tid; i < n; i += THREADS) {
                const unsigned long long k = key_of(i);
                if ((k & pmask) == prefix) atomicAdd(&S.hist[(int)((k >> sh) & dmask)], 1);
            }
        }
        __syncthreads();
        // thread t owns bins (BINS-1-2t, BINS-2-2t): scan from the top bin downwards
        const int b0 = BINS - 1 - 2 * tid, b1 = b0 - 1;
        const int h0 = S.hist[b0], h1 = S.hist[b1];
        int total;
        const int before = cta_exclusive_scan(h0 + h1, S.warp_tot, total);  // keys in bins above b0
        if (before < remaining && remaining <= before + h0) { S.digit = b0; S.above = before; }
        else if (before + h0 < remaining && remaining <= before + h0 + h1) { S.digit = b1; S.above = before + h0; }
        __syncthreads();
        prefix |= (unsigned long long)S.digit << sh;
        pmask |= dmask << sh;
        remaining -= S.above;
        const int in_bin = S.hist[S.digit];
        __syncthreads();
        if (listed < 0 && in_bin <= list_cap && shift > 0) {
            if (tid == 0) S.count = 0;
            __syncthreads();
            for (int i = tid; i < n; i += THREADS) {
                const unsigned long long k = key_of(i);
                if ((k & pmask) == prefix) skeys[atomicAdd(&S.count, 1)] = k;
            }
            __syncthreads();
            listed = S.count;
            __syncthreads();
        }
    }
    // prefix is now the K-th largest key exactly: gather everything >= it (exactly K keys)
    if (tid == 0) S.count = 0;
    for (int i = tid; i < fill; i += THREADS) skeys[i] = 0ull;
    __syncthreads();
    for (int i = tid; i < n; i += THREADS) {
        const unsigned long long k = key_of(i);
        if (k >= prefix) { const int p = atomicAdd(&S.count, 1); if (p < KCAP) skeys[p] = k; }
    }
    __syncthreads();
    sort_desc_n(skeys, m);
    return m;
}

}  // namespace sel
