// Hard voxelisation + pillar feature decoration on device.
//
// The reference voxelises on the CPU with a sequential first-seen scan (spconv VoxelGenerator through
// OPCD/pcdet/datasets/processor/data_processor.py:15-60,115-143, or det3d's VoxelGenerator through
// clients/preprocess/voxelize.py:19-24,38-42). A GPU hash is unordered, so determinism is rebuilt from
// point indices: voxel id = rank of the cell's FIRST point among all first points, slot inside the voxel
// = rank of the point among the cell's points — bit-identical to the sequential scan.
//
//   K1 cells      : per point, x/y range mask (common_utils.py:60-63, inclusive, no z test), cell =
//                   floor((p - min)/size) per axis in f32 (or f64); table slot = the cell id itself (pillar
//                   grids) or an atomicCAS open-addressing insert of it; atomicMin(first point index),
//                   atomicAdd(point count) -> arrival position
//   K2 mark       : per table slot, the cell's first point gets a record and its 1024-point block counts it
//   K3 scan       : exclusive scan of the block counts (one CTA)
//   K4 assign     : voxel id = block offset + in-block rank; coords; CSR segment; overflow break point
//   K5 scatter    : point index -> its arrival position in its voxel's CSR segment (unordered); zero-fills the voxels
//   K6 place      : per point, slot = number of smaller indices in the segment; slots < P store their row
//   (break mode: K7 counts each voxel's points before the sweep's break point)
// A call takes a batch of sweeps (ragged, back to back in one array); tables, numbering and outputs are per sweep.
//   decorate      : PillarVFE.forward decoration (OPCD/pcdet/models/backbones_3d/vfe/pillar_vfe.py:94-118)
#include "common.cuh"
// Kernels here are launched with programmatic stream serialization (TCB_KLAUNCH) and begin with pdl_enter(). Their const
// pointer parameters carry no __restrict__ and nothing is read with __ldg, so every input is read with coherent loads
// (see common.cuh: that matters once a kernel may be resident while its producer still runs; the point rows themselves can
// come from tcb_pointcloud2_unpack on the same stream).
#include <limits.h>

namespace {

constexpr int VB = 1024;            // points per block; a block never straddles two sweeps
constexpr int VOX_MAX_BATCH = 64;   // sweeps per call (the sweep table travels in the kernel parameters)

// Sweeps of a batch lie back to back in one [n, pt_stride] array. Every kernel that walks the points uses blocks of VB
// points cut per sweep, so a block's sweep is one lookup and the ordered passes (first-seen numbering) restart per sweep.
struct VoxBatch {
    int b;
    int pt_off[VOX_MAX_BATCH + 1];    // first point row of sweep s; pt_off[b] = n
    int blk_off[VOX_MAX_BATCH + 1];   // first block of sweep s;     blk_off[b] = number of blocks
};

struct VoxParams {
    const float* pts; int n; int pt_stride; int nfeat;
    double rmin[3], vs[3];
    float rminf[3], vsf[3], mx0, mx1, my0, my1;  // f32 copies (f32 mode and range mask)
    int grid[3];
    int max_pts, max_vox, break_mode, f64_math, range_mask;
    float z_offset; int intensity_col; const float* intensity_max;  // b1 conditioning (optional); one maximum per sweep
    int dense;   // 1: slot = cell id (pillar grids: the whole grid fits a table); 0: open-addressing hash of the cell id
    int TS;      // table slots per sweep (dense: cells; hash: a power of two >= 2 * the largest sweep)
};

__device__ __forceinline__ int vox_sweep_of_block(const VoxBatch& B, int blk) {
    int lo = 0, hi = B.b - 1;   // the last sweep whose first block is <= blk (empty sweeps share their successor's offset)
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (B.blk_off[mid] <= blk) lo = mid; else hi = mid - 1;
    }
    return lo;
}
// sweep and point row handled by this thread (block of VB threads); false for the ragged tail of a sweep's last block
__device__ __forceinline__ bool vox_locate(const VoxBatch& B, int& s, int& i) {
    s = vox_sweep_of_block(B, (int)blockIdx.x);
    i = B.pt_off[s] + ((int)blockIdx.x - B.blk_off[s]) * VB + (int)threadIdx.x;
    return i < B.pt_off[s + 1];
}
// the same for kernels whose threads do not cooperate: CTAs of VB / SUB threads, SUB of them per block of VB points (a CTA
// retires as soon as its own warps are done instead of waiting for the slowest of 32)
template <int SUB>
__device__ __forceinline__ bool vox_locate_sub(const VoxBatch& B, int& s, int& i) {
    const int blk = (int)blockIdx.x / SUB;
    s = vox_sweep_of_block(B, blk);
    i = B.pt_off[s] + (blk - B.blk_off[s]) * VB + ((int)blockIdx.x % SUB) * (VB / SUB) + (int)threadIdx.x;
    return i < B.pt_off[s + 1];
}
constexpr int PLACE_SUB = 4;

__device__ __forceinline__ float cond_value(const VoxParams& p, int s, float v, int j) {
    if (j == 2) v = v + p.z_offset;
    if (j == p.intensity_col) v = v / ld_coherent(p.intensity_max + s);
    return v;
}
__device__ __forceinline__ float cond_feature(const VoxParams& p, int s, const float* row, int j) {
    return cond_value(p, s, ld_coherent(row + j), j);
}
// a whole 16-byte row (x, y, z, feature) in one load
__device__ __forceinline__ float4 cond_row4(const VoxParams& p, int s, long long i) {
    float4 r = *reinterpret_cast<const float4*>(p.pts + i * 4);
    r.x = cond_value(p, s, r.x, 0); r.y = cond_value(p, s, r.y, 1); r.z = cond_value(p, s, r.z, 2); r.w = cond_value(p, s, r.w, 3);
    return r;
}

// cell id of a point, or -1 when masked / outside the grid
__device__ __forceinline__ int point_cell(const VoxParams& p, float x, float y, float z, int& cx, int& cy, int& cz) {
    if (p.range_mask && !(x >= p.mx0 && x <= p.mx1 && y >= p.my0 && y <= p.my1)) return -1;
    int c[3];
    const float q[3] = {x, y, z};
#pragma unroll
    for (int j = 0; j < 3; j++) {
        if (p.f64_math) {
            const double f = floor(((double)q[j] - p.rmin[j]) / p.vs[j]);
            if (!(f >= 0.0 && f < (double)p.grid[j])) return -1;
            c[j] = (int)f;
        } else {
            const float f = floorf((q[j] - p.rminf[j]) / p.vsf[j]);
            if (!(f >= 0.f && f < (float)p.grid[j])) return -1;
            c[j] = (int)f;
        }
    }
    cx = c[0]; cy = c[1]; cz = c[2];
    return (cz * p.grid[1] + cy) * p.grid[0] + cx;
}
// How a kernel reads point rows: ROW_ANY scalar loads at any stride (the lanes of a warp read neighbouring rows, so the
// loads of one row's columns hit the same lines); ROW_VEC4 one 16-byte load per row (stride 4, aligned: KITTI x, y, z, i).
enum { ROW_ANY = 0, ROW_VEC4 = 1 };

__device__ __forceinline__ unsigned hash_cell(int cell, int H) {
    return ((unsigned)cell * 2654435761u) >> 7 & (unsigned)(H - 1);
}

// Table record of a cell: x = smallest point index seen (atomicMin), y = points seen (atomicAdd).
constexpr long long VOX_EMPTY = 0x000000007fffffffLL;   // {x = INT_MAX, y = 0}

// one pass that prepares everything the call needs: cell records, hash keys (hash mode), first-point records + block
// counts (`zero16` 16-byte units), per-sweep maxima
__global__ void __launch_bounds__(256) vox_init_kernel(long long* __restrict__ recs, long long n_recs, int* __restrict__ hkey, long long n_keys,
                                                       void* __restrict__ flags, long long n_flags16, float* __restrict__ imax, int b) {
    pdl_enter();
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x, step = (long long)gridDim.x * blockDim.x;
    longlong2* r2 = reinterpret_cast<longlong2*>(recs);   // workspace blocks are 256-byte aligned
    for (long long k = t; k < (n_recs >> 1); k += step) r2[k] = make_longlong2(VOX_EMPTY, VOX_EMPTY);
    if (t == 0 && (n_recs & 1)) recs[n_recs - 1] = VOX_EMPTY;
    int4* k4 = reinterpret_cast<int4*>(hkey);
    for (long long k = t; k < (n_keys >> 2); k += step) k4[k] = make_int4(-1, -1, -1, -1);
    if (t < (n_keys & 3)) hkey[n_keys - 1 - t] = -1;
    int4* f4 = reinterpret_cast<int4*>(flags);
    for (long long k = t; k < n_flags16; k += step) f4[k] = make_int4(0, 0, 0, 0);
    if (imax && t < b) imax[t] = -INFINITY;
}

template <int ROWS>
__global__ void __launch_bounds__(VB / PLACE_SUB) vox_cell_kernel(const VoxParams p, const VoxBatch B, int2* __restrict__ pslot,
                                                                  int* __restrict__ hkey, int2* __restrict__ recs) {
    pdl_enter();
    int s, i;
    const bool valid = vox_locate_sub<PLACE_SUB>(B, s, i);
    float x, y, z;
    if (ROWS == ROW_VEC4) {
        if (!valid) return;
        const float4 r = *reinterpret_cast<const float4*>(p.pts + (long long)i * 4);
        x = cond_value(p, s, r.x, 0); y = cond_value(p, s, r.y, 1); z = cond_value(p, s, r.z, 2);
    } else {
        if (!valid) return;
        const float* row = p.pts + (long long)i * p.pt_stride;
        x = cond_feature(p, s, row, 0); y = cond_feature(p, s, row, 1); z = cond_feature(p, s, row, 2);
    }
    int cx, cy, cz;
    const int cell = point_cell(p, x, y, z, cx, cy, cz);
    if (cell < 0) { pslot[i] = make_int2(-1, 0); return; }
    int slot;
    if (p.dense) {
        slot = s * p.TS + cell;
    } else {
        int* keys = hkey + (long long)s * p.TS;
        unsigned h = hash_cell(cell, p.TS);
        while (true) {
            const int prev = atomicCAS(&keys[h], -1, cell);
            if (prev == -1 || prev == cell) break;
            h = (h + 1) & (unsigned)(p.TS - 1);
        }
        slot = s * p.TS + (int)h;
    }
    // Neighbouring points of a scan line mostly share their pillar: the lanes of a warp that hit the same slot send one
    // atomicMin (their lowest index: the group's first lane) and one atomicAdd (their number) between them. The value the
    // add returns numbers the cell's points in arrival order — the point's place in the cell's CSR segment later on.
    const unsigned act = __activemask();
    const unsigned peers = __match_any_sync(act, slot);
    const int leader = __ffs(peers) - 1, lane = threadIdx.x & 31;
    int base = 0;
    if (lane == leader) {
        atomicMin(&recs[slot].x, i);
        base = atomicAdd(&recs[slot].y, __popc(peers));
    }
    base = __shfl_sync(peers, base, leader);
    pslot[i] = make_int2(slot, base + __popc(peers & ((1u << lane) - 1)));
}

// Per table slot: the cell's first point gets a record {slot, points in the cell} (the ordered passes then read these in
// point order instead of chasing every point's table slot) and its bit in a bitmap over the points, from which the scan
// kernel counts the first points of every 1024-point block (32 words each: blocks start at multiples of 1024 bits
// because every sweep's bits start at its first block).
__global__ void __launch_bounds__(256) vox_mark_kernel(const VoxBatch B, const int2* recs, long long n_recs, int TS,
                                                       int2* __restrict__ finfo, unsigned* __restrict__ bitmap) {
    pdl_enter();
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n_recs; k += step) {
        const int2 r = recs[k];
        if (r.x == INT_MAX) continue;
        const int s = (int)(k / TS);
        finfo[r.x] = make_int2((int)k, r.y);
        const int bit = B.blk_off[s] * VB + (r.x - B.pt_off[s]);
        atomicOr(&bitmap[bit >> 5], 1u << (bit & 31));
    }
}

// scalars: [0] CSR cursor, [2 + s] index of the point at which the scan of sweep s breaks
constexpr int SC_CURSOR = 0, SC_BREAK = 2;

// exclusive scan of the block sums of the whole batch by one CTA (blocksum[nblk] = total), then per sweep: occupied
// cells = difference of the prefix at its block range, M = min(cells, max_vox); also resets the per-call scalars.
__global__ void __launch_bounds__(1024) vox_scan_kernel(const VoxBatch B, int nblk, const unsigned* bitmap, int* __restrict__ blocksum, int max_vox,
                                                        int* __restrict__ scalars, int* __restrict__ m_out) {
    pdl_enter();
    __shared__ int sh[32];
    __shared__ int carry;
    __shared__ int cnt[1024];
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nblk; base += 1024) {
        // first points of block i = set bits of its 32 bitmap words = 8 16-byte units: the CTA reads the units of 1024
        // blocks with coalesced loads (thread t takes unit j * 1024 + t), eight neighbouring lanes add up one block
        const uint4* w4 = reinterpret_cast<const uint4*>(bitmap + 32ll * base);
        const long long units = 8ll * min(1024, nblk - base);
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int u = j * 1024 + threadIdx.x;
            int c = 0;
            if (u < units) { const uint4 w = w4[u]; c = __popc(w.x) + __popc(w.y) + __popc(w.z) + __popc(w.w); }
            c += __shfl_xor_sync(0xffffffffu, c, 1);
            c += __shfl_xor_sync(0xffffffffu, c, 2);
            c += __shfl_xor_sync(0xffffffffu, c, 4);
            if ((threadIdx.x & 7) == 0) cnt[u >> 3] = c;
        }
        __syncthreads();
        const int i = base + threadIdx.x;
        const int v = i < nblk ? cnt[threadIdx.x] : 0;
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        int inc = v;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, s); if (lane >= s) inc += t; }
        if (lane == 31) sh[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int w = sh[lane];
#pragma unroll
            for (int s = 1; s < 32; s <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, s); if (lane >= s) w += t; }
            sh[lane] = w;
        }
        __syncthreads();
        const int prefix = carry + (warp ? sh[warp - 1] : 0) + inc - v;
        if (i < nblk) blocksum[i] = prefix;
        __syncthreads();
        if (threadIdx.x == 1023) carry = prefix + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) { blocksum[nblk] = carry; scalars[SC_CURSOR] = 0; }
    __syncthreads();
    if ((int)threadIdx.x < B.b) {
        const int s = threadIdx.x;
        const int cells = blocksum[B.blk_off[s + 1]] - blocksum[B.blk_off[s]];
        m_out[s] = cells < max_vox ? cells : max_vox;
        scalars[SC_BREAK + s] = INT_MAX;
    }
}

// vrec[slot] = {voxel id, CSR segment start (-1: cell beyond max_voxels), points in the cell, -}
__global__ void __launch_bounds__(VB) vox_assign_kernel(const VoxParams p, const VoxBatch B, const int2* finfo,
                                                        const int* hkey, const int* blockoff, int4* __restrict__ vrec,
                                                        int2* __restrict__ vinfo, int* __restrict__ scalars,
                                                        int* __restrict__ coords, int* __restrict__ num_pts) {
    pdl_enter();
    __shared__ int sh[32], sh2[32];
    __shared__ int seg_base;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int s, i, sl = -1, cnt = 0;
    if (vox_locate(B, s, i)) { const int2 f = finfo[i]; sl = f.x; cnt = f.y; }
    const int first = cnt > 0;
    // in-block exclusive rank of `first`
    const unsigned bal = __ballot_sync(0xffffffffu, first);
    if (lane == 0) sh[warp] = __popc(bal);
    __syncthreads();
    if (warp == 0) {
        int w = sh[lane];
#pragma unroll
        for (int k = 1; k < 32; k <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, k); if (lane >= k) w += t; }
        sh[lane] = w;
    }
    __syncthreads();
    const int vid = first ? blockoff[blockIdx.x] - blockoff[B.blk_off[s]] + (warp ? sh[warp - 1] : 0) + __popc(bal & ((1u << lane) - 1)) : 0;
    const bool kept = first && vid < p.max_vox;
    // CSR segments: the block's kept cells take one stretch of the cursor (one atomic per block)
    const int mine = kept ? cnt : 0;
    int inc = mine;
#pragma unroll
    for (int k = 1; k < 32; k <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, k); if (lane >= k) inc += t; }
    if (lane == 31) sh2[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = sh2[lane];
#pragma unroll
        for (int k = 1; k < 32; k <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, k); if (lane >= k) w += t; }
        sh2[lane] = w;
        if (lane == 31) seg_base = w > 0 ? atomicAdd(&scalars[SC_CURSOR], w) : 0;
    }
    __syncthreads();
    if (!first) return;
    if (kept) {
        const int cell = p.dense ? sl - s * p.TS : hkey[sl];
        const int cx = cell % p.grid[0], t = cell / p.grid[0], cy = t % p.grid[1], cz = t / p.grid[1];
        const long long v = (long long)s * p.max_vox + vid;
        const int seg = seg_base + (warp ? sh2[warp - 1] : 0) + inc - mine;
        reinterpret_cast<int4*>(coords)[v] = make_int4(s, cz, cy, cx);  // [batch, z, y, x] (preprocess_3d.py:50)
        vinfo[v] = make_int2(seg, cnt);
        vrec[sl] = make_int4(vid, seg, cnt, 0);
        if (!p.break_mode) num_pts[v] = cnt < p.max_pts ? cnt : p.max_pts;   // "break": counted once the break point is known
    } else {
        vrec[sl] = make_int4(vid, -1, cnt, 0);
        if (vid == p.max_vox) scalars[SC_BREAK + s] = i;   // the point at which the SECOND/det3d scan breaks
    }
}

// Every point of a kept cell goes to its arrival position in the cell's segment (unordered; ranked afterwards). The
// blocks of a sweep also clear the sweep's M voxels between them (the zero padding of the voxel generators), so that
// the placement that follows only writes the rows that exist.
__global__ void __launch_bounds__(VB / PLACE_SUB) vox_scatter_kernel(const VoxParams p, const VoxBatch B, const int* m_dev, const int2* pslot,
                                                                     const int4* vrec, int* __restrict__ csr, float* __restrict__ voxels) {
    pdl_enter();
    int s, i;
    const bool valid = vox_locate_sub<PLACE_SUB>(B, s, i);
    if (valid) {
        const int2 ps = pslot[i];
        if (ps.x >= 0) {
            const int4 vr = vrec[ps.x];
            if (vr.y >= 0) csr[vr.y + ps.y] = i;          // vr.y < 0, "continue": the voxel was never created
        }
    }
    const int nb = (B.blk_off[s + 1] - B.blk_off[s]) * PLACE_SUB, r = (int)blockIdx.x - B.blk_off[s] * PLACE_SUB;   // CTAs of this sweep, this one's rank
    const long long per_vox = (long long)p.max_pts * p.nfeat;
    const long long total = (long long)m_dev[s] * per_vox;                       // floats of this sweep to clear
    float* base = voxels + (long long)s * p.max_vox * per_vox;
    if ((per_vox & 3) == 0 && (((uintptr_t)voxels) & 15) == 0) {
        const long long q4 = total >> 2, chunk = (q4 + nb - 1) / nb;
        const long long lo = (long long)r * chunk, hi = min(q4, lo + chunk);
        for (long long q = lo + threadIdx.x; q < hi; q += VB / PLACE_SUB) st_stream_f4(base + 4 * q, make_float4(0.f, 0.f, 0.f, 0.f));
    } else {
        const long long chunk = (total + nb - 1) / nb;
        const long long lo = (long long)r * chunk, hi = min(total, lo + chunk);
        for (long long q = lo + threadIdx.x; q < hi; q += VB / PLACE_SUB) base[q] = 0.f;
    }
}

// Point-centric placement: a point's slot in its voxel is the number of points of the same cell with a smaller index
// (what the sequential scan's per-voxel counter gives it). The cell's segment is a handful of indices that the lanes
// of a warp mostly share, so every point counts for itself and, when its rank is below max_points, stores its own row
// — read in order — straight into the voxel. The count stops (checked every 8 entries) once max_points smaller indices
// are seen, so a crowded cell costs its late points little; only its first max_points points walk the whole segment.
template <int ROWS>
__global__ void __launch_bounds__(VB / PLACE_SUB) vox_place_kernel(const VoxParams p, const VoxBatch B, const int2* pslot, const int4* vrec,
                                                                   const int* scalars, const int* csr, float* __restrict__ voxels) {
    pdl_enter();
    int s, i;
    if (!vox_locate_sub<PLACE_SUB>(B, s, i)) return;
    const int sl = pslot[i].x;
    if (sl < 0) return;
    const int4 vr = vrec[sl];
    if (vr.y < 0) return;
    if (p.break_mode && i >= scalars[SC_BREAK + s]) return;   // "break": the scan stopped before this point
    const int* seg = csr + vr.y;
    const int P = p.max_pts, F = p.nfeat, L = vr.z;
    int rank = 0, t = 0;
    for (; t + 8 <= L && rank < P; t += 8) {
        int c = 0;
#pragma unroll
        for (int u = 0; u < 8; u++) c += (seg[t + u] < i);
        rank += c;
    }
    if (rank < P) for (; t < L; t++) rank += (seg[t] < i);
    if (rank >= P) return;
    float* vout = voxels + ((long long)s * p.max_vox + vr.x) * P * F;
    if (ROWS == ROW_VEC4) {
        reinterpret_cast<float4*>(vout)[rank] = cond_row4(p, s, i);
    } else {
        const float* row = p.pts + (long long)i * p.pt_stride;
        for (int j = 0; j < F; j++) vout[rank * F + j] = cond_feature(p, s, row, j);
    }
}

// "break" mode: a voxel holds the points of its cell that come before the sweep's break point, at most max_points
__global__ void __launch_bounds__(256) vox_breakcount_kernel(const VoxParams p, int b, const int* m_dev, const int2* vinfo, const int* scalars,
                                                             const int* csr, int* __restrict__ num_pts) {
    pdl_enter();
    const long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int s = (int)(v / p.max_vox);
    if (s >= b || (int)(v - (long long)s * p.max_vox) >= m_dev[s]) return;
    const int2 vi = vinfo[v];
    const int brk = scalars[SC_BREAK + s];
    int lb = vi.y;
    if (brk != INT_MAX) {
        lb = 0;
        for (int t = 0; t < vi.y; t++) lb += (csr[vi.x + t] < brk);
    }
    num_pts[v] = lb < p.max_pts ? lb : p.max_pts;
}

// per-sweep max of one feature column (b1: intensity /= max(intensity)); values assumed finite. grid (x, sweep)
__global__ void __launch_bounds__(256) vox_colmax_kernel(const float* pts, const VoxBatch B, int stride, int col, float* __restrict__ out) {
    pdl_enter();
    const int s = blockIdx.y;
    float m = -INFINITY;
    for (int i = B.pt_off[s] + blockIdx.x * blockDim.x + threadIdx.x; i < B.pt_off[s + 1]; i += gridDim.x * blockDim.x)
        m = fmaxf(m, ld_coherent(pts + (long long)i * stride + col));
#pragma unroll
    for (int k = 16; k > 0; k >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, k));
    if ((threadIdx.x & 31) == 0) {
        // float atomic max through the ordered-int trick
        if (m >= 0.f) atomicMax(reinterpret_cast<int*>(out + s), __float_as_int(m));
        else atomicMin(reinterpret_cast<unsigned*>(out + s), __float_as_uint(m));
    }
}

// ---- pillar decoration ---------------------------------------------------------------------------
// out[v, s, :] = mask * [feat(F), xyz - mean(3), xyz - pillar centre(3)]      (pillar_vfe.py:94-118)
// One warp per voxel. The voxel's P x F block comes in with coalesced 16-byte loads into the warp's slice of shared
// memory, the P x (F+6) result is assembled there and leaves with coalesced 16-byte stores (an output row is 40-44
// bytes: written from registers each lane would store 10 scattered words).
constexpr int DEC_WARPS = 8;
// TP, TF: compile-time max_points / features of the two served configurations (32 x 4 KITTI, 20 x 5 nuScenes), 0 = run time
template <bool VEC4, int TP, int TF>
__global__ void __launch_bounds__(32 * DEC_WARPS, 6) pillar_decorate_kernel(const float* voxels, const int* coords, const int* num_pts,
                                                                            const int* m_dev, int b, int vox_stride,
                                                                            int P_rt, int F_rt, float vx, float vy, float vz, float xo, float yo, float zo,
                                                                            float* __restrict__ out) {
    extern __shared__ __align__(16) float dec_smem[];
    const int P = TP ? TP : P_rt, F = TF ? TF : F_rt;
    __shared__ int pre[VOX_MAX_BATCH + 1];
    pdl_enter();
    if ((int)threadIdx.x < b) pre[threadIdx.x + 1] = m_dev ? m_dev[threadIdx.x] : vox_stride;
    __syncthreads();
    if (threadIdx.x == 0) { int a = 0; for (int s = 0; s < b; s++) { const int m = pre[s + 1]; pre[s] = a; a += m; } pre[b] = a; }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int FO = F + 6, nin = P * F, nout = P * FO;
    float* sin_ = dec_smem + (size_t)warp * (((nin + 3) & ~3) + ((nout + 3) & ~3));
    float* sout = sin_ + ((nin + 3) & ~3);
    const int total = pre[b], nw = gridDim.x * DEC_WARPS;
    for (int k = blockIdx.x * DEC_WARPS + warp; k < total; k += nw) {     // warps stride over the voxels of the whole batch
        int lo = 0, hi = b - 1;
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (pre[mid] <= k) lo = mid; else hi = mid - 1; }
        const long long v = (long long)lo * vox_stride + (k - pre[lo]);
        const float* vin = voxels + v * nin;
        const int n = num_pts[v];
        const int4 c = reinterpret_cast<const int4*>(coords)[v];  // b, z, y, x
        const float px = (float)c.w * vx + xo, py = (float)c.z * vy + yo, pz = (float)c.y * vz + zo;
        if (TF == 4 && VEC4) {
            // a 16-byte row per slot: lane = slot, the slot's row stays in registers from the load to the outputs
            constexpr int RF = TF > 0 ? TF : 1;
            auto load_row = [&](int q, float (&r)[RF]) {
                if (RF == 4 && VEC4) { const float4 t = reinterpret_cast<const float4*>(vin)[q]; r[0] = t.x; r[1] = t.y; r[2] = t.z; r[RF - 1] = t.w; }
                else {
#pragma unroll
                    for (int j = 0; j < RF; j++) r[j] = vin[q * RF + j];
                }
            };
            float sx = 0.f, sy = 0.f, sz = 0.f;
            for (int q = lane; q < P; q += 32) { float r[RF]; load_row(q, r); sx += r[0]; sy += r[1]; sz += r[2]; }
#pragma unroll
            for (int q = 16; q > 0; q >>= 1) {
                sx += __shfl_xor_sync(0xffffffffu, sx, q); sy += __shfl_xor_sync(0xffffffffu, sy, q); sz += __shfl_xor_sync(0xffffffffu, sz, q);
            }
            const float fn = (float)n;
            const float mx = sx / fn, my = sy / fn, mz = sz / fn;
            for (int q = lane; q < P; q += 32) {
                float r[RF];
                load_row(q, r);                                               // P > 32: the second read hits L1
                const float mask = (n > q) ? 1.f : 0.f;
                float* o = sout + q * (RF + 6);
#pragma unroll
                for (int j = 0; j < RF; j++) o[j] = r[j] * mask;
                o[RF + 0] = (r[0] - mx) * mask; o[RF + 1] = (r[1] - my) * mask; o[RF + 2] = (r[2] - mz) * mask;
                o[RF + 3] = (r[0] - px) * mask; o[RF + 4] = (r[1] - py) * mask; o[RF + 5] = (r[2] - pz) * mask;
            }
        } else {
            if (VEC4) { for (int e = lane; e < (nin >> 2); e += 32) reinterpret_cast<float4*>(sin_)[e] = reinterpret_cast<const float4*>(vin)[e]; }
            else      { for (int e = lane; e < nin; e += 32) sin_[e] = vin[e]; }
            __syncwarp();
            float sx = 0.f, sy = 0.f, sz = 0.f;
            for (int q = lane; q < P; q += 32) { sx += sin_[q * F]; sy += sin_[q * F + 1]; sz += sin_[q * F + 2]; }
#pragma unroll
            for (int q = 16; q > 0; q >>= 1) {
                sx += __shfl_xor_sync(0xffffffffu, sx, q); sy += __shfl_xor_sync(0xffffffffu, sy, q); sz += __shfl_xor_sync(0xffffffffu, sz, q);
            }
            const float fn = (float)n;
            const float mx = sx / fn, my = sy / fn, mz = sz / fn;
            for (int q = lane; q < P; q += 32) {
                const float mask = (n > q) ? 1.f : 0.f;
                const float x = sin_[q * F], y = sin_[q * F + 1], z = sin_[q * F + 2];
                float* o = sout + q * FO;
                for (int j = 0; j < F; j++) o[j] = sin_[q * F + j] * mask;
                o[F + 0] = (x - mx) * mask; o[F + 1] = (y - my) * mask; o[F + 2] = (z - mz) * mask;
                o[F + 3] = (x - px) * mask; o[F + 4] = (y - py) * mask; o[F + 5] = (z - pz) * mask;
            }
        }
        __syncwarp();
        float* vo = out + v * nout;
        if (VEC4) { for (int e = lane; e < (nout >> 2); e += 32) st_stream_f4(vo + 4 * e, reinterpret_cast<const float4*>(sout)[e]); }
        else      { for (int e = lane; e < nout; e += 32) vo[e] = sout[e]; }
        __syncwarp();   // the next voxel overwrites both slices
    }
}

// ---- dynamic pillarisation (b7) --------------------------------------------------------------------
// DynamicPillarVFE.forward up to the PFN layers (OPCD/pcdet/models/backbones_3d/vfe/dynamic_pillar_vfe.py:90-142):
// no point cap per pillar; pillars are numbered in sorted key order (torch.unique), which on a dense BEV grid is
// the exclusive scan of the occupancy flags — no sort needed.
struct DynParams {
    const float* pts; int n, ncol;   // rows (batch_idx, x, y, z, features...), ncol = 1 + F
    float r0x, r0y, vx, vy;          // range min / voxel size as float32 tensors (:83-85)
    float xo, yo, zo;                // voxel/2 + range_min, Python doubles used as f32 scalars (:76-78)
    int gx, gy, batch;
};

__global__ void __launch_bounds__(256) dyn_mark_kernel(const DynParams p, int* __restrict__ valid, int* __restrict__ pkey,
                                                       int* __restrict__ mark, double* __restrict__ sums, int* __restrict__ cnt) {
    pdl_enter();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    const float* row = p.pts + (long long)i * p.ncol;
    const float x = ld_coherent(row + 1), y = ld_coherent(row + 2), z = ld_coherent(row + 3);
    const float fx = floorf((x - p.r0x) / p.vx), fy = floorf((y - p.r0y) / p.vy);
    const int b = (int)ld_coherent(row);
    const bool ok = fx >= 0.f && fx < (float)p.gx && fy >= 0.f && fy < (float)p.gy && b >= 0 && b < p.batch;
    valid[i] = ok ? 1 : 0;
    if (!ok) { pkey[i] = -1; return; }
    const int key = b * (p.gx * p.gy) + (int)fx * p.gy + (int)fy;   // merge_coords (:99-101)
    pkey[i] = key;
    mark[key] = 1;
    atomicAdd(&sums[3ll * key], (double)x); atomicAdd(&sums[3ll * key + 1], (double)y); atomicAdd(&sums[3ll * key + 2], (double)z);
    atomicAdd(&cnt[key], 1);
}

// block-local exclusive scan of 0/1 flags + per-block totals; offsets are added by the consumers
__global__ void __launch_bounds__(VB) scan_local_kernel(const int* flags, int n, int* __restrict__ pos, int* __restrict__ blocksum) {
    pdl_enter();
    __shared__ int sh[32];
    const int i = blockIdx.x * VB + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int f = (i < n) ? flags[i] : 0;
    const unsigned bal = __ballot_sync(0xffffffffu, f != 0);
    if (lane == 0) sh[warp] = __popc(bal);
    __syncthreads();
    if (warp == 0) {
        int w = sh[lane];
#pragma unroll
        for (int k = 1; k < 32; k <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, k); if (lane >= k) w += t; }
        sh[lane] = w;
    }
    __syncthreads();
    if (i < n) pos[i] = (warp ? sh[warp - 1] : 0) + __popc(bal & ((1u << lane) - 1));
    if (threadIdx.x == 0) blocksum[blockIdx.x] = sh[31];
}

__global__ void __launch_bounds__(1024) scan_blocks_kernel(int nblk, int* __restrict__ blocksum, int* __restrict__ total_out) {
    pdl_enter();
    __shared__ int sh[32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nblk; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = (i < nblk) ? blocksum[i] : 0;
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        int inc = v;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, s); if (lane >= s) inc += t; }
        if (lane == 31) sh[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int w = sh[lane];
#pragma unroll
            for (int s = 1; s < 32; s <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, s); if (lane >= s) w += t; }
            sh[lane] = w;
        }
        __syncthreads();
        const int prefix = carry + (warp ? sh[warp - 1] : 0) + inc - v;
        if (i < nblk) blocksum[i] = prefix;
        __syncthreads();
        if (threadIdx.x == 1023) carry = prefix + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_out = carry;
}

__global__ void __launch_bounds__(256) dyn_features_kernel(const DynParams p, const int* valid, const int* pkey,
                                                           const int* ppos, const int* pblk,
                                                           const int* crank, const int* cblk,
                                                           const double* sums, const int* cnt,
                                                           float* __restrict__ feats, long long* __restrict__ unq_inv) {
    pdl_enter();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n || !valid[i]) return;
    const int key = pkey[i];
    const long long o = (long long)ppos[i] + pblk[i / VB];      // row among the in-grid points, input order kept
    const int u = crank[key] + cblk[key / VB];                   // pillar id = rank of the key among occupied cells
    const float* row = p.pts + (long long)i * p.ncol;
    const int F = p.ncol - 1;
    float* out = feats + o * (F + 6);
    const float x = ld_coherent(row + 1), y = ld_coherent(row + 2), z = ld_coherent(row + 3);
    const float fc = (float)cnt[key];
    const float mx = (float)sums[3ll * key] / fc, my = (float)sums[3ll * key + 1] / fc, mz = (float)sums[3ll * key + 2] / fc;
    const int within = key % (p.gx * p.gy);
    const int cx = within / p.gy, cy = within - cx * p.gy;
    for (int j = 0; j < F; j++) out[j] = ld_coherent(row + 1 + j);
    out[F] = x - mx; out[F + 1] = y - my; out[F + 2] = z - mz;
    out[F + 3] = x - ((float)cx * p.vx + p.xo);
    out[F + 4] = y - ((float)cy * p.vy + p.yo);
    out[F + 5] = z - p.zo;
    unq_inv[o] = (long long)u;
}

__global__ void __launch_bounds__(256) dyn_coords_kernel(int cells, int gx, int gy, const int* mark, const int* crank,
                                                         const int* cblk, int* __restrict__ coords) {
    pdl_enter();
    const int key = blockIdx.x * blockDim.x + threadIdx.x;
    if (key >= cells || !mark[key]) return;
    const int u = crank[key] + cblk[key / VB];
    const int b = key / (gx * gy), within = key - b * (gx * gy), cx = within / gy, cy = within - cx * gy;
    reinterpret_cast<int4*>(coords)[u] = make_int4(b, 0, cy, cx);   // [:, [0,3,2,1]] of (b, cx, cy, 0)  (:132-138)
}

int next_pow2(long long v) { int r = 1024; while (r < v) r <<= 1; return r; }

// ---------------------------------------------------------------------------------------------------------
// PointCloud2 -> point rows: what
//   np.array(list(point_cloud2.read_points(msg, field_names=(...), skip_nans=True)))   (communicator/ros_inference3d.py:125)
// produces, without the per-point Python generator. sensor_msgs.point_cloud2 is a ROS library that is not part of
// the reference tree; restated from its published algorithm: the selected fields are unpacked in order of their
// byte offset (not of field_names), a point is dropped when any selected value is NaN, points keep their order.
// Rows at and after the returned count are filled with NaN in the first three columns and -inf in the others:
// tcb_voxelize drops points with a NaN coordinate and its column maximum ignores -inf, so the whole width*height
// buffer can be handed on without reading the count back to the host.
// ---------------------------------------------------------------------------------------------------------
constexpr int PC2_MAX_FIELDS = 8;
struct Pc2Params {
    const unsigned char* data;
    int width, height, point_step, nfields, bigendian, skip_nans;
    long long row_step;
    int offset[PC2_MAX_FIELDS], datatype[PC2_MAX_FIELDS];   // sensor_msgs/PointField: 1 i8, 2 u8, 3 i16, 4 u16, 7 f32
};

__device__ __forceinline__ bool pc2_decode(const Pc2Params& p, int i, float* v) {
    const int row = i / p.width, col = i - row * p.width;
    const unsigned char* src = p.data + (long long)row * p.row_step + (long long)col * p.point_step;
    bool ok = true;
    for (int f = 0; f < p.nfields; f++) {
        const unsigned char* q = src + p.offset[f];
        float val;
        switch (p.datatype[f]) {
            case 1: val = (float)(signed char)q[0]; break;
            case 2: val = (float)q[0]; break;
            case 3: { const unsigned u = p.bigendian ? ((unsigned)q[0] << 8 | q[1]) : ((unsigned)q[1] << 8 | q[0]); val = (float)(short)u; break; }
            case 4: { const unsigned u = p.bigendian ? ((unsigned)q[0] << 8 | q[1]) : ((unsigned)q[1] << 8 | q[0]); val = (float)u; break; }
            default: {
                const unsigned u = p.bigendian ? ((unsigned)q[0] << 24 | (unsigned)q[1] << 16 | (unsigned)q[2] << 8 | q[3])
                                               : ((unsigned)q[3] << 24 | (unsigned)q[2] << 16 | (unsigned)q[1] << 8 | q[0]);
                val = __uint_as_float(u);
            }
        }
        if (p.skip_nans && val != val) ok = false;
        v[f] = val;
    }
    return ok;
}

__global__ void __launch_bounds__(256) pc2_flags_kernel(const Pc2Params p, int n, int* __restrict__ flags) {
    pdl_enter();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float v[PC2_MAX_FIELDS];
    flags[i] = pc2_decode(p, i, v) ? 1 : 0;
}

__global__ void __launch_bounds__(256) pc2_scatter_kernel(const Pc2Params p, int n, const int* flags, const int* pos,
                                                          const int* blockoff, const int* count,
                                                          float* __restrict__ out) {
    pdl_enter();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (flags[i]) {
        float v[PC2_MAX_FIELDS];
        pc2_decode(p, i, v);
        float* d = out + (long long)(pos[i] + blockoff[i / VB]) * p.nfields;
        for (int f = 0; f < p.nfields; f++) d[f] = v[f];
    }
    if (i >= *count) {   // tail: rows no kept point lands in
        float* d = out + (long long)i * p.nfields;
        for (int f = 0; f < p.nfields; f++) d[f] = f < 3 ? __int_as_float(0x7fc00000) : -INFINITY;
    }
}

// BoundingBoxArray fields of the detections the reference publishes (communicator/ros_inference3d.py:150-177):
// indices with labels == label_sel and scores > score_thresh, ascending; per box position (z - offset), the
// quaternion of a yaw about z as pyquaternion builds it (w = cos(yaw/2), x = 0*s, y = 0*s, z = s = sin(yaw/2), in
// float64 from the float32 yaw), dimensions with x and y swapped, score, label.
__global__ void __launch_bounds__(1024) jsk_boxes_kernel(const float* boxes, const float* scores,
                                                         const long long* __restrict__ labels, int n, long long label_sel, float score_thresh,
                                                         double z_offset, double* __restrict__ out, int* __restrict__ out_idx,
                                                         int* __restrict__ out_count) {
    pdl_enter();
    __shared__ int warp_tot[32];
    __shared__ int carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const bool sel = i < n && labels[i] == label_sel && scores[i] > score_thresh;
        const unsigned bal = __ballot_sync(0xffffffffu, sel);
        if (lane == 0) warp_tot[warp] = __popc(bal);
        __syncthreads();
        int before = carry;
        for (int w = 0; w < warp; w++) before += warp_tot[w];
        if (sel) {
            const int k = before + __popc(bal & ((1u << lane) - 1));
            const float* b = boxes + 7ll * i;
            const double half = (double)b[6] / 2.0, sn = sin(half), cs = cos(half);
            double* o = out + 12ll * k;
            o[0] = (double)b[0]; o[1] = (double)b[1]; o[2] = (double)b[2] - z_offset;
            o[3] = 0.0 * sn; o[4] = 0.0 * sn; o[5] = 1.0 * sn; o[6] = cs;          // orientation x, y, z, w
            o[7] = (double)b[4]; o[8] = (double)b[3]; o[9] = (double)b[5];          // dimensions x, y, z
            o[10] = (double)scores[i]; o[11] = (double)labels[i];
            out_idx[k] = i;
        }
        __syncthreads();
        if (tid == 0) { int t = 0; for (int w = 0; w < 32; w++) t += warp_tot[w]; carry += t; }
        __syncthreads();
    }
    if (tid == 0) *out_count = carry;
}

}  // namespace

// table geometry shared by the workspace query and the entry point
static void vox_table(long long cells, int n_max, int& dense, int& TS) {
    // pillar grids (a few hundred thousand cells) index the table by cell id: no keys, no probing; sparse 3-D grids
    // (C4': 90 M cells) hash the cell id into 2 * n slots
    dense = (cells <= (1ll << 20) || cells <= 4ll * (n_max > 0 ? n_max : 1)) ? 1 : 0;
    TS = dense ? (int)cells : next_pow2(2LL * (n_max > 0 ? n_max : 1));
}

extern "C" size_t tcb_voxelize_workspace(int n, int b, int max_vox, long long cells) {
    if (n < 0) n = 0;
    if (b < 1) b = 1;
    if (max_vox < 0) max_vox = 0;
    int dense, TS;
    vox_table(cells > 0 ? cells : 1, n, dense, TS);   // n as the bound of the largest sweep
    size_t bytes = 8192;
    bytes += (size_t)n * 12 + 512;                            // pslot (slot, arrival position), csr
    bytes += (size_t)b * TS * (8 + 16 + (dense ? 0 : 4)) + 768; // cell records, voxel records, hash keys
    bytes += (size_t)b * max_vox * 8 + 256;                   // vinfo (segment, points)
    bytes += (size_t)n * 8 + ((size_t)n / VB + b + 4) * (4 + 128) + 1024; // first-point records, block sums, first-point bitmap
    return bytes;
}

extern "C" int tcb_voxelize(tcb_ctx* ctx, const float* pts_dev, int n, const int* sweep_offsets_host, int b, int pt_stride, int nfeat,
                            const double range[6], const double vsize[3], int max_pts, int max_vox, int overflow_mode, int f64_math,
                            int range_mask_xy, float z_offset, int intensity_col, float* voxels_dev, int* coords_dev,
                            int* num_pts_dev, int* m_out_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, range && vsize && voxels_dev && coords_dev && num_pts_dev && m_out_dev, "tcb_voxelize: null argument");
    TCB_REQUIRE(ctx, n >= 0 && (n == 0 || pts_dev), "tcb_voxelize: bad points");
    TCB_REQUIRE(ctx, b >= 1 && (b == 1 || sweep_offsets_host), "tcb_voxelize: b sweeps need b+1 row offsets");
    if (b > VOX_MAX_BATCH) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_voxelize: %d sweeps per call (limit %d)", b, VOX_MAX_BATCH);
    TCB_REQUIRE(ctx, nfeat >= 3 && pt_stride >= nfeat, "tcb_voxelize: need nfeat >= 3 and pt_stride >= nfeat (got %d, %d)", nfeat, pt_stride);
    TCB_REQUIRE(ctx, max_pts >= 1 && max_vox >= 1, "tcb_voxelize: max_pts and max_vox must be >= 1");
    TCB_REQUIRE(ctx, overflow_mode == 0 || overflow_mode == 1, "tcb_voxelize: overflow_mode must be 0 (continue) or 1 (break)");
    TCB_REQUIRE(ctx, intensity_col < nfeat, "tcb_voxelize: intensity_col out of range");
    if (n > (1 << 26)) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_voxelize: more than 2^26 points");
    cudaStream_t st = (cudaStream_t)stream;

    VoxBatch B;
    memset(&B, 0, sizeof(B));
    B.b = b;
    int n_max = 0;
    for (int s = 0; s <= b; s++) {
        const int off = sweep_offsets_host ? sweep_offsets_host[s] : (s == 0 ? 0 : n);
        TCB_REQUIRE(ctx, off >= 0 && off <= n && (s == 0 ? off == 0 : off >= B.pt_off[s - 1]), "tcb_voxelize: sweep offsets must ascend from 0 to n");
        B.pt_off[s] = off;
        if (s > 0) {
            const int ns = off - B.pt_off[s - 1];
            n_max = ns > n_max ? ns : n_max;
            B.blk_off[s] = B.blk_off[s - 1] + tcb_div_up(ns, VB);
        }
    }
    TCB_REQUIRE(ctx, B.pt_off[b] == n, "tcb_voxelize: the last sweep offset must be n");
    const int nblk = B.blk_off[b];

    VoxParams p;
    memset(&p, 0, sizeof(p));
    p.pts = pts_dev; p.n = n; p.pt_stride = pt_stride; p.nfeat = nfeat;
    long long cells = 1;
    for (int j = 0; j < 3; j++) {
        TCB_REQUIRE(ctx, vsize[j] > 0 && range[j + 3] > range[j], "tcb_voxelize: bad range/voxel size on axis %d", j);
        p.rmin[j] = range[j]; p.vs[j] = vsize[j];
        p.rminf[j] = (float)range[j]; p.vsf[j] = (float)vsize[j];
        // grid = round((max - min) / size), in the arithmetic of the chosen precision (data_processor.py:117-118)
        const double g = f64_math ? (range[j + 3] - range[j]) / vsize[j]
                                  : (double)(((float)range[j + 3] - (float)range[j]) / (float)vsize[j]);
        p.grid[j] = (int)nearbyint(g);
        TCB_REQUIRE(ctx, p.grid[j] >= 1, "tcb_voxelize: empty grid on axis %d", j);
        cells *= p.grid[j];
    }
    if (cells >= INT_MAX) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_voxelize: grid of %lld cells does not fit a 32-bit key", cells);
    p.mx0 = (float)range[0]; p.mx1 = (float)range[3]; p.my0 = (float)range[1]; p.my1 = (float)range[4];
    p.max_pts = max_pts; p.max_vox = max_vox; p.break_mode = overflow_mode; p.f64_math = f64_math ? 1 : 0;
    p.range_mask = range_mask_xy ? 1 : 0; p.z_offset = z_offset; p.intensity_col = intensity_col;
    vox_table(cells, n_max, p.dense, p.TS);
    const long long slots = (long long)b * p.TS;
    if (slots >= INT_MAX || (long long)b * max_vox >= INT_MAX) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_voxelize: batch too large for 32-bit slot ids");

    TCB_WS(ctx, scalars, int, 2 * VOX_MAX_BATCH + 64);
    TCB_WS(ctx, pslot, int2, n > 0 ? n : 1);
    TCB_WS(ctx, csr, int, n > 0 ? n : 1);
    TCB_WS(ctx, recs, int2, slots);
    TCB_WS(ctx, vrec, int4, slots);
    TCB_WS(ctx, hkey, int, p.dense ? 1 : slots);
    TCB_WS(ctx, vinfo, int2, (size_t)b * max_vox);
    // zeroed per call, in one stretch: first-point records (8 B per point), then the first-point bitmap (1024 bits a block)
    const long long zero16 = ((long long)n * 8 + 15) / 16 + (long long)nblk * 8;
    TCB_WS(ctx, zeroed, char, zero16 * 16 + 16);
    TCB_WS(ctx, blocksum, int, nblk + 2);
    int2* finfo = reinterpret_cast<int2*>(zeroed);
    unsigned* bitmap = reinterpret_cast<unsigned*>(zeroed + (((size_t)n * 8 + 15) / 16) * 16);
    float* imax = reinterpret_cast<float*>(scalars + SC_BREAK + VOX_MAX_BATCH);
    p.intensity_max = imax;
    const bool use_max = intensity_col >= 0 && n > 0;
    // 16-byte rows read and written whole: KITTI-style (x, y, z, intensity) clouds
    const bool vec4 = pt_stride == 4 && nfeat == 4 && (((uintptr_t)pts_dev | (uintptr_t)voxels_dev) & 15) == 0;

    const long long init_units = (slots >> 1) + (p.dense ? 0 : (slots >> 2)) + zero16 + 1;
    const int init_ctas = (int)(init_units < (long long)ctx->num_sms * 8 * 256 ? tcb_div_up(init_units, 256) : ctx->num_sms * 8);
    TCB_KLAUNCH(ctx, vox_init_kernel, init_ctas, 256, 0, st, reinterpret_cast<long long*>(recs), slots, hkey, p.dense ? 0ll : slots,
                (void*)zeroed, zero16, use_max ? imax : (float*)nullptr, b);
    TCB_LAUNCHED(ctx);
    if (use_max) {
        TCB_KLAUNCH(ctx, vox_colmax_kernel, dim3(tcb_div_up(ctx->num_sms * 2, b), b), 256, 0, st, pts_dev, B, pt_stride, intensity_col, imax);
        TCB_LAUNCHED(ctx);
    }
    if (nblk > 0) {
        if (vec4) TCB_KLAUNCH(ctx, vox_cell_kernel<ROW_VEC4>, nblk * PLACE_SUB, VB / PLACE_SUB, 0, st, p, B, pslot, hkey, recs);
        else      TCB_KLAUNCH(ctx, vox_cell_kernel<ROW_ANY>, nblk * PLACE_SUB, VB / PLACE_SUB, 0, st, p, B, pslot, hkey, recs);
        TCB_LAUNCHED(ctx);
        const int mark_ctas = (int)(slots < (long long)ctx->num_sms * 8 * 256 ? tcb_div_up(slots, 256) : ctx->num_sms * 8);
        TCB_KLAUNCH(ctx, vox_mark_kernel, mark_ctas, 256, 0, st, B, recs, slots, p.TS, finfo, bitmap);
        TCB_LAUNCHED(ctx);
    }
    TCB_KLAUNCH(ctx, vox_scan_kernel, 1, 1024, 0, st, B, nblk, bitmap, blocksum, max_vox, scalars, m_out_dev);
    TCB_LAUNCHED(ctx);
    if (nblk > 0) {
        TCB_KLAUNCH(ctx, vox_assign_kernel, nblk, VB, 0, st, p, B, finfo, hkey, blocksum, vrec, vinfo, scalars, coords_dev, num_pts_dev);
        TCB_LAUNCHED(ctx);
        TCB_KLAUNCH(ctx, vox_scatter_kernel, nblk * PLACE_SUB, VB / PLACE_SUB, 0, st, p, B, m_out_dev, pslot, vrec, csr, voxels_dev);
        TCB_LAUNCHED(ctx);
        if (vec4) TCB_KLAUNCH(ctx, vox_place_kernel<ROW_VEC4>, nblk * PLACE_SUB, VB / PLACE_SUB, 0, st, p, B, pslot, vrec, scalars, csr, voxels_dev);
        else      TCB_KLAUNCH(ctx, vox_place_kernel<ROW_ANY>, nblk * PLACE_SUB, VB / PLACE_SUB, 0, st, p, B, pslot, vrec, scalars, csr, voxels_dev);
        TCB_LAUNCHED(ctx);
        if (overflow_mode) {
            TCB_KLAUNCH(ctx, vox_breakcount_kernel, tcb_div_up((long long)b * max_vox, 256), 256, 0, st, p, b, m_out_dev, vinfo, scalars, csr, num_pts_dev);
            TCB_LAUNCHED(ctx);
        }
    }
    return TCB_OK;
}

static int init_voxelize(tcb_ctx* ctx) {   // run by tcb_create
    TCB_SMEM_OPT_IN(ctx, (pillar_decorate_kernel<true, 32, 4>), 200 * 1024);
    TCB_SMEM_OPT_IN(ctx, (pillar_decorate_kernel<true, 20, 5>), 200 * 1024);
    TCB_SMEM_OPT_IN(ctx, (pillar_decorate_kernel<true, 0, 0>), 200 * 1024);
    TCB_SMEM_OPT_IN(ctx, (pillar_decorate_kernel<false, 0, 0>), 200 * 1024);
    return TCB_OK;
}
static TcbInitReg reg_voxelize(&init_voxelize);

extern "C" int tcb_pillar_decorate(tcb_ctx* ctx, const float* voxels_dev, const int* coords_dev, const int* num_pts_dev,
                                   const int* m_dev, int b, int m_max, int max_pts, int nfeat, const double vsize[3],
                                   const double range_min[3], float* out_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, voxels_dev && coords_dev && num_pts_dev && out_dev && vsize && range_min, "tcb_pillar_decorate: null argument");
    TCB_REQUIRE(ctx, b >= 1 && m_max >= 0 && max_pts >= 1 && nfeat >= 3, "tcb_pillar_decorate: bad shape");
    if (m_max == 0) return TCB_OK;
    // offsets exactly as PillarVFE.__init__ computes them (pillar_vfe.py:76-81): Python doubles, used as f32 scalars
    const float vx = (float)vsize[0], vy = (float)vsize[1], vz = (float)vsize[2];
    const float xo = (float)(vsize[0] / 2 + range_min[0]), yo = (float)(vsize[1] / 2 + range_min[1]), zo = (float)(vsize[2] / 2 + range_min[2]);
    const int nin = max_pts * nfeat, nout = max_pts * (nfeat + 6);
    const size_t smem = (size_t)DEC_WARPS * (((nin + 3) & ~3) + ((nout + 3) & ~3)) * sizeof(float);
    if (smem > 200 * 1024) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_pillar_decorate: %d points x %d features per voxel exceed the shared-memory tile", max_pts, nfeat);
    if (b > VOX_MAX_BATCH) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_pillar_decorate: %d sweeps per call (limit %d)", b, VOX_MAX_BATCH);
    const long long vmax = (long long)b * m_max;
    const int per_sm = (int)((size_t)200 * 1024 / (smem + 1024)) < 8 ? (int)((size_t)200 * 1024 / (smem + 1024)) : 8;
    const int ctas = (int)(vmax < (long long)ctx->num_sms * per_sm * DEC_WARPS ? tcb_div_up(vmax, DEC_WARPS) : ctx->num_sms * (per_sm > 0 ? per_sm : 1));
    auto ctas_of = [&]() { return ctas; };
    const bool vec4 = (nin % 4 == 0) && (nout % 4 == 0) && (((uintptr_t)voxels_dev | (uintptr_t)out_dev) & 15) == 0;
    auto launch = [&](auto kernel) -> int {
        TCB_KLAUNCH(ctx, kernel, ctas_of(), 32 * DEC_WARPS, smem, (cudaStream_t)stream, voxels_dev, coords_dev, num_pts_dev, m_dev, b, m_max,
                    max_pts, nfeat, vx, vy, vz, xo, yo, zo, out_dev);
        TCB_LAUNCHED(ctx);
        return TCB_OK;
    };
    if (vec4 && max_pts == 32 && nfeat == 4) return launch(pillar_decorate_kernel<true, 32, 4>);
    if (vec4 && max_pts == 20 && nfeat == 5) return launch(pillar_decorate_kernel<true, 20, 5>);
    if (vec4) return launch(pillar_decorate_kernel<true, 0, 0>);
    return launch(pillar_decorate_kernel<false, 0, 0>);
}

extern "C" size_t tcb_dynamic_pillar_workspace(int n, int batch, int gx, int gy) {
    const size_t cells = (size_t)(batch > 0 ? batch : 1) * (size_t)(gx > 0 ? gx : 1) * (size_t)(gy > 0 ? gy : 1);
    const size_t nn = (size_t)(n > 0 ? n : 1);
    return nn * 4 * 3 + cells * (4 + 4 + 4 + 24) + (nn / VB + cells / VB + 8) * 4 + 8192;
}

extern "C" int tcb_dynamic_pillar(tcb_ctx* ctx, const float* pts_dev, int n, int ncol, int batch, const double range[6],
                                  const double vsize[3], float* feats_dev, long long* unq_inv_dev, int* voxel_coords_dev,
                                  int* counts_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, range && vsize && feats_dev && unq_inv_dev && voxel_coords_dev && counts_dev, "tcb_dynamic_pillar: null argument");
    TCB_REQUIRE(ctx, n >= 0 && (n == 0 || pts_dev) && ncol >= 4 && batch >= 1, "tcb_dynamic_pillar: bad shape (n=%d ncol=%d batch=%d)", n, ncol, batch);
    cudaStream_t st = (cudaStream_t)stream;
    DynParams p;
    memset(&p, 0, sizeof(p));
    p.pts = pts_dev; p.n = n; p.ncol = ncol; p.batch = batch;
    p.r0x = (float)range[0]; p.r0y = (float)range[1]; p.vx = (float)vsize[0]; p.vy = (float)vsize[1];
    p.xo = (float)(vsize[0] / 2 + range[0]); p.yo = (float)(vsize[1] / 2 + range[1]); p.zo = (float)(vsize[2] / 2 + range[2]);
    p.gx = (int)nearbyint((range[3] - range[0]) / vsize[0]);
    p.gy = (int)nearbyint((range[4] - range[1]) / vsize[1]);
    TCB_REQUIRE(ctx, p.gx >= 1 && p.gy >= 1, "tcb_dynamic_pillar: empty grid");
    const long long cells = (long long)batch * p.gx * p.gy;
    if (cells > (1ll << 24)) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_dynamic_pillar: %lld BEV cells > 2^24", cells);
    if (n > (1 << 26)) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_dynamic_pillar: more than 2^26 points");
    const int nn = n > 0 ? n : 1;
    const int pb = tcb_div_up(nn, VB), cb = tcb_div_up(cells, VB);
    TCB_WS(ctx, valid, int, nn);
    TCB_WS(ctx, pkey, int, nn);
    TCB_WS(ctx, ppos, int, nn);
    TCB_WS(ctx, pblk, int, pb + 1);
    TCB_WS(ctx, mark, int, cells);
    TCB_WS(ctx, cnt, int, cells);
    TCB_WS(ctx, crank, int, cells);
    TCB_WS(ctx, cblk, int, cb + 1);
    TCB_WS(ctx, sums, double, cells * 3);
    TCB_CUDA(ctx, cudaMemsetAsync(mark, 0, sizeof(int) * cells, st));
    TCB_CUDA(ctx, cudaMemsetAsync(cnt, 0, sizeof(int) * cells, st));
    TCB_CUDA(ctx, cudaMemsetAsync(sums, 0, sizeof(double) * cells * 3, st));
    TCB_CUDA(ctx, cudaMemsetAsync(valid, 0, sizeof(int) * nn, st));
    if (n > 0) {
        TCB_KLAUNCH(ctx, dyn_mark_kernel, tcb_div_up(n, 256), 256, 0, st, p, valid, pkey, mark, sums, cnt);
        TCB_LAUNCHED(ctx);
    }
    TCB_KLAUNCH(ctx, scan_local_kernel, pb, VB, 0, st, valid, n, ppos, pblk);
    TCB_LAUNCHED(ctx);
    TCB_KLAUNCH(ctx, scan_blocks_kernel, 1, 1024, 0, st, pb, pblk, counts_dev);
    TCB_LAUNCHED(ctx);
    TCB_KLAUNCH(ctx, scan_local_kernel, cb, VB, 0, st, mark, (int)cells, crank, cblk);
    TCB_LAUNCHED(ctx);
    TCB_KLAUNCH(ctx, scan_blocks_kernel, 1, 1024, 0, st, cb, cblk, counts_dev + 1);
    TCB_LAUNCHED(ctx);
    if (n > 0) {
        TCB_KLAUNCH(ctx, dyn_features_kernel, tcb_div_up(n, 256), 256, 0, st, p, valid, pkey, ppos, pblk, crank, cblk, sums, cnt, feats_dev, unq_inv_dev);
        TCB_LAUNCHED(ctx);
    }
    TCB_KLAUNCH(ctx, dyn_coords_kernel, tcb_div_up(cells, 256), 256, 0, st, (int)cells, p.gx, p.gy, mark, crank, cblk, voxel_coords_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

extern "C" size_t tcb_pointcloud2_unpack_workspace(int width, int height) {
    const size_t n = (size_t)(width > 0 ? width : 0) * (size_t)(height > 0 ? height : 0);
    return n * 8 + (n / VB + 2) * 4 + 4 * 256;
}

extern "C" int tcb_pointcloud2_unpack(tcb_ctx* ctx, const unsigned char* data_dev, int width, int height, int point_step,
                                      long long row_step, int nfields, const int* field_offsets, const int* field_datatypes,
                                      int is_bigendian, int skip_nans, float* out_dev, int* count_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, width >= 0 && height >= 0 && point_step > 0, "tcb_pointcloud2_unpack: bad cloud shape");
    TCB_REQUIRE(ctx, count_dev != nullptr, "tcb_pointcloud2_unpack: count_dev is null");
    TCB_REQUIRE(ctx, nfields >= 1 && field_offsets && field_datatypes, "tcb_pointcloud2_unpack: no fields");
    if (nfields > PC2_MAX_FIELDS) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_pointcloud2_unpack: %d fields > compiled limit %d", nfields, PC2_MAX_FIELDS);
    const long long n = (long long)width * height;
    TCB_REQUIRE(ctx, n < (1ll << 30), "tcb_pointcloud2_unpack: too many points");
    TCB_REQUIRE(ctx, row_step >= (long long)width * point_step || height <= 1, "tcb_pointcloud2_unpack: row_step smaller than a row");
    Pc2Params p;
    memset(&p, 0, sizeof(p));
    p.data = data_dev; p.width = width; p.height = height; p.point_step = point_step; p.row_step = row_step;
    p.nfields = nfields; p.bigendian = is_bigendian ? 1 : 0; p.skip_nans = skip_nans ? 1 : 0;
    for (int f = 0; f < nfields; f++) {
        const int dt = field_datatypes[f];
        const int size = (dt == 1 || dt == 2) ? 1 : ((dt == 3 || dt == 4) ? 2 : 4);
        if (!(dt == 1 || dt == 2 || dt == 3 || dt == 4 || dt == 7))
            return tcb_fail(ctx, TCB_ERR_ARG, "tcb_pointcloud2_unpack: field %d has datatype %d; only INT8/UINT8/INT16/UINT16/FLOAT32 are exact in "
                                              "float32 rows (the reference carries INT32/UINT32/FLOAT64 as float64)", f, dt);
        TCB_REQUIRE(ctx, field_offsets[f] >= 0 && field_offsets[f] + size <= point_step, "tcb_pointcloud2_unpack: field %d lies outside the point", f);
        TCB_REQUIRE(ctx, f == 0 || field_offsets[f] > field_offsets[f - 1], "tcb_pointcloud2_unpack: fields must come in ascending byte offset (read_points' order)");
        p.offset[f] = field_offsets[f]; p.datatype[f] = dt;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) { TCB_CUDA(ctx, cudaMemsetAsync(count_dev, 0, sizeof(int), st)); return TCB_OK; }
    TCB_REQUIRE(ctx, data_dev && out_dev, "tcb_pointcloud2_unpack: null buffer");
    const int nb = tcb_div_up(n, VB);
    TCB_WS(ctx, flags, int, n);
    TCB_WS(ctx, pos, int, n);
    TCB_WS(ctx, blk, int, nb + 1);
    TCB_KLAUNCH(ctx, pc2_flags_kernel, tcb_div_up(n, 256), 256, 0, st, p, (int)n, flags);
    TCB_LAUNCHED(ctx);
    TCB_KLAUNCH(ctx, scan_local_kernel, nb, VB, 0, st, flags, (int)n, pos, blk);
    TCB_LAUNCHED(ctx);
    TCB_KLAUNCH(ctx, scan_blocks_kernel, 1, 1024, 0, st, nb, blk, count_dev);
    TCB_LAUNCHED(ctx);
    TCB_KLAUNCH(ctx, pc2_scatter_kernel, tcb_div_up(n, 256), 256, 0, st, p, (int)n, flags, pos, blk, count_dev, out_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

extern "C" int tcb_boxes_to_jsk(tcb_ctx* ctx, const float* boxes_dev, const float* scores_dev, const long long* labels_dev, int n,
                                long long label_sel, float score_thresh, double z_offset, double* out_dev, int* out_idx_dev,
                                int* out_count_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, n >= 0 && out_count_dev, "tcb_boxes_to_jsk: bad arguments");
    TCB_REQUIRE(ctx, n == 0 || (boxes_dev && scores_dev && labels_dev && out_dev && out_idx_dev), "tcb_boxes_to_jsk: null buffer");
    // plain launch: its inputs come from kernels of other files, and the compiler reads them through the non-coherent path
    TCB_CUDA(ctx, tcb_launch(false, jsk_boxes_kernel, dim3(1), dim3(1024), 0, (cudaStream_t)stream, boxes_dev, scores_dev, labels_dev, n, label_sel, score_thresh, z_offset, out_dev,
                                                          out_idx_dev, out_count_dev));
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}
