// Greedy NMS for large candidate sets as a mask + sweep (shared by the rotated 3D path, the YOLOv4 per-class NMS and the
// detectron batched NMS): the pairwise "i suppresses j" bitmask of m <= 4096 score-ordered candidates is built by the
// whole grid (64x64 tiles of the upper triangle, one 64-bit word per row and tile: the layout of the reference's
// nms_kernel, OPCD/pcdet/ops/iou3d_nms/src/iou3d_nms_kernel.cu:267-311), then swept by one warp.
#pragma once
#include <cuda_runtime.h>

namespace wide {

constexpr int PRE_CAP = 4096;         // most candidates
constexpr int POST_CAP = 1024;        // most kept
constexpr int MW = PRE_CAP / 64;      // mask words per row

// Tile kernel for predicates cheap enough to evaluate in place (axis-aligned IoU): grid = (column tile, row tile, set),
// 64 threads = the rows of the tile. R / m_dev / mask are per set (set stride PRE_CAP boxes, 1 int, PRE_CAP*MW words).
template <typename Box, typename Pred>
__global__ void __launch_bounds__(64) tile_mask_kernel(const Box* __restrict__ R_all, const int* __restrict__ m_all, int m_stride, Pred pred,
                                                       unsigned long long* __restrict__ mask_all) {
    pdl_wait();
    const int bi = blockIdx.y, bj = blockIdx.x, set = blockIdx.z;
    const int m = m_all[(size_t)set * m_stride];
    if (bj < bi || bj * 64 >= m) return;          // lower triangle / past the last candidate: never read
    const Box* R = R_all + (size_t)set * PRE_CAP;
    unsigned long long* mask = mask_all + (size_t)set * PRE_CAP * MW;
    __shared__ Box cols[64];
    const int tid = threadIdx.x;
    const int gj0 = bj * 64;
    if (gj0 + tid < m) cols[tid] = R[gj0 + tid];
    __syncthreads();
    const int gi = bi * 64 + tid;
    if (gi >= m) return;
    const Box mine = R[gi];
    unsigned long long bits = 0ull;
    const int jend = min(64, m - gj0);
    for (int j = 0; j < jend; j++)
        if (gj0 + j > gi && pred(mine, cols[j])) bits |= 1ull << j;
    mask[(size_t)gi * MW + bj] = bits;
}

// The sweep over the mask (iou3d_nms.cpp:116-132 does it on the host) is a chain over the kept boxes: keep the first
// open candidate, OR its mask row into the "removed" bitmap, find the next open one. One warp does it: lane l holds
// bitmap words 2l and 2l+1 in registers; the mask rows arrive in blocks of 64 (the candidates of one bitmap word)
// through SW_STAGES shared-memory buffers, one 32 KB bulk async copy (TMA) per block that the same warp issues three
// blocks ahead, unless every candidate of the block is already removed (a removed candidate stays removed).
constexpr int SW_ROWS = 64, SW_STAGES = 4;
constexpr int SWEEP_T = 512;          // threads of the sweep kernels: warp 0 sweeps, all of them write the results
struct SweepShared {
    unsigned long long rows[SW_STAGES][SW_ROWS][MW];
    unsigned long long bar[SW_STAGES];
    int kpos[POST_CAP];                    // position (in score order) of each kept box
};

static __device__ __forceinline__ unsigned sw_smem(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// Greedy pass over a mask in the layout above (row i, word w = columns 64w..64w+63, upper triangle), called by the
// first warp of the CTA only; the other warps wait at the caller's barrier. Returns the
// number kept; S.kpos holds their positions.
static __device__ int warp_mask_sweep(const unsigned long long* __restrict__ mask, int m, int post_max, SweepShared& S) {
    const int lane = threadIdx.x & 31;
    const unsigned full = 0xffffffffu;
    const int nblk = (m + 63) >> 6;
    // words >= nblk of a row only touch bits >= m, which are set from the start: garbage there is harmless
    auto init_word = [&](int w) { const int lo = w * 64; return lo + 64 <= m ? 0ull : (lo >= m ? ~0ull : ~((1ull << (m - lo)) - 1ull)); };
    unsigned long long rem0 = init_word(2 * lane), rem1 = init_word(2 * lane + 1);
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < SW_STAGES; k++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sw_smem(&S.bar[k])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    unsigned pending = 0, phases = 0;             // per stage: a copy is in flight / parity of its next completion
#ifdef TCB_SWEEP_PROFILE
    long long t_issue = 0, t_wait = 0, t_hop = 0, t0 = clock64(), tt;
    int n_blocks = 0;
#define SWP(acc) do { tt = clock64(); acc += tt - t0; t0 = tt; } while (0)
#else
#define SWP(acc) do {} while (0)
#endif
    auto issue = [&](int k) {                     // fetch block k (64 consecutive rows = one contiguous 32 KB piece) unless all of it is closed
        const int st = k % SW_STAGES;
        const unsigned long long open = ~__shfl_sync(full, (k & 1) ? rem1 : rem0, k >> 1);
        if (open == 0ull) return;
        const unsigned bytes = (unsigned)min(SW_ROWS, m - SW_ROWS * k) * (unsigned)(MW * 8);
        if (lane == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sw_smem(&S.bar[st])), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(sw_smem(&S.rows[st][0][0])), "l"(mask + (size_t)(SW_ROWS * k) * MW), "r"(bytes), "r"(sw_smem(&S.bar[st])) : "memory");
        }
        pending |= 1u << st;
    };
    for (int k = 0; k < SW_STAGES - 1 && k < nblk; k++) issue(k);
    SWP(t_issue);
    int nk = 0;
    for (int k = 0; k < nblk && nk < post_max; k++) {
        __syncwarp();   // every lane is done with the buffer of block k-1 before copies land in it again
        if (k + SW_STAGES - 1 < nblk) issue(k + SW_STAGES - 1);
        SWP(t_issue);
        const int st = k % SW_STAGES;
        unsigned long long avail = ~__shfl_sync(full, (k & 1) ? rem1 : rem0, k >> 1);
        if (avail == 0ull) {                      // nothing open here; a copy issued for it earlier still has to land
            if (pending & (1u << st)) {
                unsigned done = 0;
                while (!done)
                    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                                 : "=r"(done) : "r"(sw_smem(&S.bar[st])), "r"((phases >> st) & 1u) : "memory");
                phases ^= 1u << st; pending &= ~(1u << st);
            }
            continue;
        }
        {   // open candidates were open when the block was issued, so their rows are on the way
            unsigned done = 0;
            while (!done)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(sw_smem(&S.bar[st])), "r"((phases >> st) & 1u) : "memory");
            phases ^= 1u << st; pending &= ~(1u << st);
        }
        SWP(t_wait);
        // (1) greedy choice inside the block without walking it box by box. Lane l looks after candidates l and l+32 of
        //     the block and first collects, for each, the set of earlier candidates of the block that would suppress it
        //     (a column of the block's 64x64 diagonal sub-mask: 64 broadcast loads, one bit each). Then rounds: a
        //     candidate with a kept suppressor is removed; one whose possible suppressors are all decided and none kept
        //     is kept. Every round settles at least the first undecided candidate; with few suppressions it settles
        //     almost all of them at once.
        {
            const uint2* colk = reinterpret_cast<const uint2*>(&S.rows[st][0][k]);   // word k of row i is colk[i * MW]
            unsigned long long col0 = 0ull, col1 = 0ull;
#pragma unroll
            for (int i = 0; i < 64; i++) {
                const uint2 w = colk[i * MW];
                col0 |= (unsigned long long)((w.x >> lane) & 1u) << i;
                col1 |= (unsigned long long)((w.y >> lane) & 1u) << i;
            }
            col0 &= avail; col1 &= avail;                       // only candidates still open can be kept and suppress
            unsigned long long K = 0ull, R = 0ull;
            while (true) {
                const unsigned long long U = avail & ~(K | R);
                if (U == 0ull) break;
                const bool u0 = (U >> lane) & 1ull, u1 = (U >> (lane + 32)) & 1ull;
                const bool r0 = u0 && (col0 & K) != 0ull, r1 = u1 && (col1 & K) != 0ull;
                const bool k0 = u0 && !r0 && (col0 & U) == 0ull, k1 = u1 && !r1 && (col1 & U) == 0ull;
                K |= (unsigned long long)__ballot_sync(full, k0) | ((unsigned long long)__ballot_sync(full, k1) << 32);
                R |= (unsigned long long)__ballot_sync(full, r0) | ((unsigned long long)__ballot_sync(full, r1) << 32);
            }
            int room = post_max - nk;
            while (__popcll(K) > room) K &= ~(1ull << (63 - __clzll((long long)K)));   // beyond post_max: drop from the far end
            if ((K >> lane) & 1ull) S.kpos[nk + __popcll(K & ((1ull << lane) - 1ull))] = 64 * k + lane;
            if ((K >> (lane + 32)) & 1ull) S.kpos[nk + __popcll(K & ((1ull << (lane + 32)) - 1ull))] = 64 * k + 32 + lane;
            nk += __popcll(K);
            const unsigned klo = (unsigned)K, khi = (unsigned)(K >> 32);
            // (2) what the kept boxes of this block suppress further on: independent loads, two bitmap words per lane
            //     (words before k were never written for these rows)
            const bool own0 = 2 * lane >= k, own1 = 2 * lane + 1 >= k;
            const ulonglong2* mine = reinterpret_cast<const ulonglong2*>(&S.rows[st][0][2 * lane]);   // my two words of row j: mine[j * MW / 2]
            unsigned long long acc0 = 0ull, acc1 = 0ull;
            if (klo | khi) {
                // all 64 rows, unrolled: the loads are independent and stream back to back (a loop over the kept rows only
                // would wait for each load in turn)
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    const ulonglong2 v = mine[j * (MW / 2)], u = mine[(32 + j) * (MW / 2)];
                    const unsigned long long sa = ((klo >> j) & 1u) ? ~0ull : 0ull, sb = ((khi >> j) & 1u) ? ~0ull : 0ull;
                    acc0 |= (v.x & sa) | (u.x & sb);
                    acc1 |= (v.y & sa) | (u.y & sb);
                }
            }
            if (own0) rem0 |= acc0;
            if (own1) rem1 |= acc1;
        }
        SWP(t_hop);
#ifdef TCB_SWEEP_PROFILE
        n_blocks++;
#endif
    }
#ifdef TCB_SWEEP_PROFILE
    if (lane == 0) printf("sweep: m=%d kept=%d blocks=%d cycles issue=%lld wait=%lld hop=%lld\n", m, nk, n_blocks, t_issue, t_wait, t_hop);
#endif
    // copies still in flight must land before the CTA may exit
    for (int st = 0; st < SW_STAGES; st++) {
        if (pending & (1u << st)) {
            unsigned done = 0;
            while (!done)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(sw_smem(&S.bar[st])), "r"((phases >> st) & 1u) : "memory");
        }
    }
    __syncwarp();
    return nk;
}

}  // namespace wide
