// CTA-wide greedy NMS over m boxes that are already in their processing order (earlier = wins) and
// resident in shared memory. Generic in the box type and in the "earlier box a suppresses later box b"
// predicate, so the same code serves the rotated BEV NMS (pcdet nms_gpu), the YOLOv4 per-class numpy NMS
// and torchvision-style batched NMS.
//
// Instead of the reference's N x N/64 bitmask + serial host sweep (iou3d_nms.cpp:105-134) the CTA works in
// rounds: every not-yet-resolved candidate of a 1024-wide window is tested against the boxes kept since it
// was last looked at; the first 64 survivors are resolved among themselves with a 64x64 bitmask and a
// short serial sweep; everything stops as soon as post_max boxes are kept.
#pragma once
#include <cuda_runtime.h>

namespace nms {

constexpr int THREADS = 1024;
constexpr int ACAP = 64;  // survivors of a round resolved against each other at once (one mask word per row)

template <typename Box, int POST_CAP, int PRE_CAP, int QCAP = 1>
struct Shared {
    unsigned queue[QCAP];                // (candidate slot << 16 | kept index) pairs awaiting the expensive test
    int qcount;
    Box kept[POST_CAP];
    int kpos[POST_CAP];                  // position (in processing order) of each kept box
    unsigned resolved[PRE_CAP / 32];     // bit c: box c is kept or suppressed
    unsigned long long mask[ACAP];
    int aidx[ACAP], knew[ACAP];
    int warp_cnt[32];
    int nkept, next_pos, a_total;
};

// Returns the number kept (<= post_max); S.kpos[0..n) are their positions in order. All 1024 threads call it.
// Predicates with an expensive exact test (Pred::kQueue) are evaluated in two steps in phase (1): every candidate
// first lists the kept boxes its cheap pre-test (Pred::maybe) cannot rule out, then the whole CTA works the list
// off one pair per thread — otherwise the rare expensive tests would each run with 1 of 32 lanes active.
template <typename Box, typename Pred, int POST_CAP, int PRE_CAP, int QCAP>
__device__ int cta_greedy_nms(const Box* __restrict__ R, int m, Pred suppresses, int post_max, Shared<Box, POST_CAP, PRE_CAP, QCAP>& S) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) S.nkept = 0;
    for (int i = tid; i < PRE_CAP / 32; i += THREADS) S.resolved[i] = 0u;
    __syncthreads();
    int pos = 0, prev_end = 0, nk_prev = 0;
    while (pos < m) {
        const int cn = min(THREADS, m - pos);
        const int c = pos + tid;
        const int nk = S.nkept;
        bool alive = tid < cn && !((S.resolved[c >> 5] >> (c & 31)) & 1u);
        if (Pred::kQueue) {
            // (1a) list the (candidate, kept) pairs that need the exact test
            if (tid == 0) S.qcount = 0;
            __syncthreads();
            if (alive) {
                const Box bx = R[c];
                for (int k = (c < prev_end) ? nk_prev : 0; k < nk; k++) {
                    if (!suppresses.maybe(S.kept[k], bx)) continue;
                    const int q = atomicAdd(&S.qcount, 1);
                    if (q < QCAP) S.queue[q] = ((unsigned)tid << 16) | (unsigned)k;
                    else if (suppresses.exact(S.kept[k], bx)) { atomicOr(&S.resolved[c >> 5], 1u << (c & 31)); break; }   // list full: test in place
                }
            }
            __syncthreads();
            // (1b) one pair per thread, all lanes busy
            const int nq = min(S.qcount, QCAP);
            for (int q = tid; q < nq; q += THREADS) {
                const unsigned e = S.queue[q];
                const int cc = pos + (int)(e >> 16);
                if (suppresses.exact(S.kept[e & 0xffffu], R[cc])) atomicOr(&S.resolved[cc >> 5], 1u << (cc & 31));
            }
            __syncthreads();
            alive = alive && !((S.resolved[c >> 5] >> (c & 31)) & 1u);
        } else if (alive) {  // (1) against the boxes kept since this candidate was last examined
            const Box bx = R[c];
            for (int k = (c < prev_end) ? nk_prev : 0; k < nk; k++)
                if (suppresses(S.kept[k], bx)) { alive = false; break; }
            if (!alive) atomicOr(&S.resolved[c >> 5], 1u << (c & 31));
        }
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
        if (Pred::kQueue) {
            // (2) suppression mask among the A survivors, expensive tests again listed first and then shared out
            if (tid == 0) S.qcount = 0;
            if (tid < ACAP) S.mask[tid] = 0ull;
            __syncthreads();
            {
                const int ri = tid >> 4, l16 = tid & 15;
                if (ri < A) {
                    const Box bi = R[S.aidx[ri]];
#pragma unroll 1
                    for (int q = 0; q < 4; q++) {
                        const int j = l16 * 4 + q;
                        if (j > ri && j < A && suppresses.maybe(bi, R[S.aidx[j]])) {
                            const int e = atomicAdd(&S.qcount, 1);   // at most ACAP*ACAP/2 = 2048 pairs <= QCAP
                            if (e < QCAP) S.queue[e] = ((unsigned)ri << 16) | (unsigned)j;
                            else if (suppresses.exact(bi, R[S.aidx[j]])) atomicOr(&S.mask[ri], 1ull << j);
                        }
                    }
                }
            }
            __syncthreads();
            const int nq = min(S.qcount, QCAP);
            for (int q = tid; q < nq; q += THREADS) {
                const unsigned e = S.queue[q];
                const int ri = (int)(e >> 16), j = (int)(e & 0xffffu);
                if (suppresses.exact(R[S.aidx[ri]], R[S.aidx[j]])) atomicOr(&S.mask[ri], 1ull << j);
            }
        } else {  // (2) suppression mask among the A survivors: 16 threads per row, 4 columns each
            const int ri = tid >> 4, l16 = tid & 15;
            unsigned long long bits = 0ull;
            if (ri < A) {
                const Box bi = R[S.aidx[ri]];
#pragma unroll 1
                for (int q = 0; q < 4; q++) {
                    const int j = l16 * 4 + q;
                    if (j > ri && j < A && suppresses(bi, R[S.aidx[j]])) bits |= 1ull << j;
                }
            }
#pragma unroll
            for (int sft = 1; sft < 16; sft <<= 1) bits |= __shfl_xor_sync(0xffffffffu, bits, sft);
            if (l16 == 0 && ri < ACAP) S.mask[ri] = bits;
        }
        __syncthreads();
        if (tid == 0) {  // (3) serial sweep, hopping from kept survivor to kept survivor
            unsigned long long avail = A >= 64 ? ~0ull : ((1ull << A) - 1);
            int kept = nk;
            while (kept < post_max && avail) {
                const int i = __ffsll((long long)avail) - 1;
                avail &= ~(S.mask[i] | (1ull << i));
                S.knew[kept - nk] = i;
                kept++;
            }
            S.nkept = kept;
        }
        __syncthreads();
        const int nk_new = S.nkept;
        if (tid < nk_new - nk) { const int cs = S.aidx[S.knew[tid]]; S.kept[nk + tid] = R[cs]; S.kpos[nk + tid] = cs; }
        if (tid < A) { const int cs = S.aidx[tid]; atomicOr(&S.resolved[cs >> 5], 1u << (cs & 31)); }
        prev_end = pos + cn; nk_prev = nk;
        const int np = S.next_pos;
        __syncthreads();
        if (nk_new >= post_max) break;
        pos = np;
    }
    __syncthreads();
    return S.nkept;
}

}  // namespace nms
