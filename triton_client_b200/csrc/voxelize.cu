// Hard voxelisation + pillar feature decoration on device.
//
// The reference voxelises on the CPU with a sequential first-seen scan (spconv VoxelGenerator through
// OPCD/pcdet/datasets/processor/data_processor.py:15-60,115-143, or det3d's VoxelGenerator through
// clients/preprocess/voxelize.py:19-24,38-42). A GPU hash is unordered, so determinism is rebuilt from
// point indices: voxel id = rank of the cell's FIRST point among all first points, slot inside the voxel
// = rank of the point among the cell's points — bit-identical to the sequential scan.
//
//   K1 cells+hash : per point, x/y range mask (common_utils.py:60-63, inclusive, no z test), cell =
//                   floor((p - min)/size) per axis in f32 (or f64), atomicCAS open-addressing insert of
//                   the linear cell id, atomicMin(first point index), atomicAdd(point count)
//   K2 count      : per 1024-point block, how many points are the first of their cell
//   K3 scan       : exclusive scan of the block counts (one CTA)
//   K4 assign     : voxel id = block offset + in-block rank; coords; CSR segment; overflow break point
//   K5 scatter    : point index -> its voxel's CSR segment (unordered, atomicAdd cursor)
//   K6 gather     : one warp per voxel orders the segment by point index, keeps the first P, writes
//                   the zero-padded (P, F) block and the count
//   decorate      : PillarVFE.forward decoration (OPCD/pcdet/models/backbones_3d/vfe/pillar_vfe.py:94-118)
#include "common.cuh"
// Kernels here are launched with programmatic stream serialization (TCB_KLAUNCH) and begin with pdl_enter(). Their const
// pointer parameters carry no __restrict__ and nothing is read with __ldg, so every input is read with coherent loads
// (see common.cuh: that matters once a kernel may be resident while its producer still runs; the point rows themselves can
// come from tcb_pointcloud2_unpack on the same stream).
#include <limits.h>

namespace {

constexpr int VB = 1024;  // points per block in the ordered passes

struct VoxParams {
    const float* pts; int n; int pt_stride; int nfeat;
    double rmin[3], rmax_x, rmax_y, rmin_x, rmin_y;  // range
    double vs[3];
    float rminf[3], vsf[3], mx0, mx1, my0, my1;  // f32 copies (f32 mode and range mask)
    int grid[3];
    int max_pts, max_vox, break_mode, f64_math, range_mask;
    float z_offset; int intensity_col; const float* intensity_max;  // b1 conditioning (optional)
    int H;  // hash size (power of two)
};

__device__ __forceinline__ float cond_feature(const VoxParams& p, const float* row, int j) {
    float v = ld_coherent(row + j);
    if (j == 2) v = v + p.z_offset;
    if (j == p.intensity_col) v = v / ld_coherent(p.intensity_max);
    return v;
}

// cell id of point i, or -1 when masked / outside the grid
__device__ __forceinline__ int point_cell(const VoxParams& p, int i, int& cx, int& cy, int& cz) {
    const float* row = p.pts + (long long)i * p.pt_stride;
    const float x = cond_feature(p, row, 0), y = cond_feature(p, row, 1), z = cond_feature(p, row, 2);
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

__device__ __forceinline__ unsigned hash_cell(int cell, int H) {
    return ((unsigned)cell * 2654435761u) >> 7 & (unsigned)(H - 1);
}

__global__ void __launch_bounds__(256) vox_hash_kernel(const VoxParams p, int* __restrict__ pslot,
                                                       int* __restrict__ hkey, int* __restrict__ hmin, int* __restrict__ hcnt) {
    pdl_enter();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    int cx, cy, cz;
    const int cell = point_cell(p, i, cx, cy, cz);
    if (cell < 0) { pslot[i] = -1; return; }
    unsigned h = hash_cell(cell, p.H);
    while (true) {
        const int prev = atomicCAS(&hkey[h], -1, cell);
        if (prev == -1 || prev == cell) break;
        h = (h + 1) & (unsigned)(p.H - 1);
    }
    atomicMin(&hmin[h], i);
    atomicAdd(&hcnt[h], 1);
    pslot[i] = (int)h;
}

__device__ __forceinline__ int block_sum(int v, int* sh) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    int t = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0;
    if (warp == 0) {
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) t += __shfl_xor_sync(0xffffffffu, t, s);
        if (lane == 0) sh[0] = t;
    }
    __syncthreads();
    const int r = sh[0];
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(VB) vox_count_kernel(int n, const int* pslot, const int* hmin,
                                                       int* __restrict__ blocksum) {
    pdl_enter();
    __shared__ int sh[32];
    const int i = blockIdx.x * VB + threadIdx.x;
    int first = 0;
    if (i < n) { const int s = pslot[i]; first = (s >= 0 && hmin[s] == i); }
    const int tot = block_sum(first, sh);
    if (threadIdx.x == 0) blocksum[blockIdx.x] = tot;
}

// exclusive scan of up to 1024*1024 points' block sums by one CTA (nblk <= 4096)
__global__ void __launch_bounds__(1024) vox_scan_kernel(int nblk, int* __restrict__ blocksum, int max_vox,
                                                        int* __restrict__ scalars, int* __restrict__ m_out) {
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
    if (threadIdx.x == 0) {
        const int total = carry;
        scalars[0] = total;                              // occupied cells
        scalars[1] = INT_MAX;                            // break point index (break mode)
        scalars[2] = 0;                                  // CSR cursor
        *m_out = total < max_vox ? total : max_vox;
    }
}

__global__ void __launch_bounds__(VB) vox_assign_kernel(const VoxParams p, const int* pslot, const int* hmin,
                                                        const int* hcnt, const int* blockoff,
                                                        int* __restrict__ hvid, int* __restrict__ hseg, int* __restrict__ vslot,
                                                        int* __restrict__ scalars, int* __restrict__ coords) {
    pdl_enter();
    __shared__ int sh[32];
    const int i = blockIdx.x * VB + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int s = -1, first = 0;
    if (i < p.n) { s = pslot[i]; first = (s >= 0 && hmin[s] == i); }
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
    if (!first) return;
    const int vid = blockoff[blockIdx.x] + (warp ? sh[warp - 1] : 0) + __popc(bal & ((1u << lane) - 1));
    hvid[s] = vid;
    if (vid < p.max_vox) {
        int cx, cy, cz;
        point_cell(p, i, cx, cy, cz);
        reinterpret_cast<int4*>(coords)[vid] = make_int4(0, cz, cy, cx);  // [batch, z, y, x] (preprocess_3d.py:50)
        vslot[vid] = s;
        hseg[s] = atomicAdd(&scalars[2], hcnt[s]);
    } else if (vid == p.max_vox) {
        scalars[1] = i;  // the point at which the SECOND/det3d scan breaks
    }
}

__global__ void __launch_bounds__(256) vox_scatter_kernel(const VoxParams p, const int* pslot, const int* hvid,
                                                          const int* hseg, int* __restrict__ hfill,
                                                          const int* scalars, int* __restrict__ csr) {
    pdl_enter();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    const int s = pslot[i];
    if (s < 0) return;
    if (hvid[s] >= p.max_vox) return;                 // "continue": the voxel was never created
    if (p.break_mode && i >= scalars[1]) return;      // "break": the scan stopped before this point
    csr[hseg[s] + atomicAdd(&hfill[s], 1)] = i;
}

// One warp per voxel: keep the max_pts smallest point indices of the segment, in ascending order.
__global__ void __launch_bounds__(256) vox_gather_kernel(const VoxParams p, const int* m_dev, const int* vslot,
                                                         const int* hseg, const int* hfill,
                                                         const int* csr, float* __restrict__ voxels, int* __restrict__ num_pts) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (v >= *m_dev) return;
    const int s = vslot[v];
    const int L = hfill[s];
    const int* seg = csr + hseg[s];
    const int P = p.max_pts, F = p.nfeat;
    const int keep = L < P ? L : P;
    float* vout = voxels + (long long)v * P * F;
    if (lane == 0) num_pts[v] = keep;
    if (L <= 32) {
        const int mine = lane < L ? seg[lane] : INT_MAX;
        int rank = 0;
        for (int t = 0; t < L; t++) rank += (__shfl_sync(0xffffffffu, mine, t) < mine);
        if (lane < L && rank < P) {
            const float* row = p.pts + (long long)mine * p.pt_stride;
            for (int j = 0; j < F; j++) vout[rank * F + j] = cond_feature(p, row, j);
        }
    } else if (L <= 32 * 8) {
        // up to 256 points: the segment is read once into registers (8 per lane), then P rounds of "smallest index
        // greater than the last one" run on registers and shuffles only
        int q[8];
#pragma unroll
        for (int u = 0; u < 8; u++) q[u] = lane + 32 * u < L ? seg[lane + 32 * u] : INT_MAX;
        int last = -1;
        for (int r = 0; r < keep; r++) {
            int best = INT_MAX;
#pragma unroll
            for (int u = 0; u < 8; u++) if (q[u] > last && q[u] < best) best = q[u];
#pragma unroll
            for (int k = 16; k > 0; k >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, k));
            last = best;
            const float* row = p.pts + (long long)best * p.pt_stride;
            for (int j = lane; j < F; j += 32) vout[r * F + j] = cond_feature(p, row, j);
        }
    } else {
        int last = -1;
        for (int r = 0; r < keep; r++) {  // selection: next smallest index greater than `last`
            int best = INT_MAX;
            for (int t = lane; t < L; t += 32) { const int q = seg[t]; if (q > last && q < best) best = q; }
#pragma unroll
            for (int k = 16; k > 0; k >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, k));
            last = best;
            const float* row = p.pts + (long long)best * p.pt_stride;
            for (int j = lane; j < F; j += 32) vout[r * F + j] = cond_feature(p, row, j);
        }
    }
    for (int e = keep * F + lane; e < P * F; e += 32) vout[e] = 0.f;  // zero padding
}

// per-frame max of one feature column (b1: intensity /= max(intensity)); values assumed finite
__global__ void __launch_bounds__(256) vox_colmax_kernel(const float* pts, int n, int stride, int col, float* __restrict__ out) {
    pdl_enter();
    float m = -INFINITY;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) m = fmaxf(m, ld_coherent(pts + (long long)i * stride + col));
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, s));
    if ((threadIdx.x & 31) == 0) {
        // float atomic max through the ordered-int trick
        if (m >= 0.f) atomicMax(reinterpret_cast<int*>(out), __float_as_int(m));
        else atomicMin(reinterpret_cast<unsigned*>(out), __float_as_uint(m));
    }
}

__global__ void vox_init_max_kernel(float* out) {
    pdl_enter();
    *out = -INFINITY;
}

// ---- pillar decoration ---------------------------------------------------------------------------
// out[v, s, :] = mask * [feat(F), xyz - mean(3), xyz - pillar centre(3)]      (pillar_vfe.py:94-118)
__global__ void __launch_bounds__(256) pillar_decorate_kernel(const float* voxels, const int* coords,
                                                              const int* num_pts, const int* m_dev, int m_host,
                                                              int P, int F, float vx, float vy, float vz, float xo, float yo, float zo,
                                                              float* __restrict__ out) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int M = m_dev ? *m_dev : m_host;
    if (v >= M) return;
    const float* vin = voxels + (long long)v * P * F;
    const int n = num_pts[v];
    const int4 c = reinterpret_cast<const int4*>(coords)[v];  // b, z, y, x
    float sx = 0.f, sy = 0.f, sz = 0.f;
    for (int s = lane; s < P; s += 32) { sx += vin[s * F]; sy += vin[s * F + 1]; sz += vin[s * F + 2]; }
#pragma unroll
    for (int k = 16; k > 0; k >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, k); sy += __shfl_xor_sync(0xffffffffu, sy, k); sz += __shfl_xor_sync(0xffffffffu, sz, k);
    }
    const float fn = (float)n;
    const float mx = sx / fn, my = sy / fn, mz = sz / fn;
    const float px = (float)c.w * vx + xo, py = (float)c.z * vy + yo, pz = (float)c.y * vz + zo;
    const int FO = F + 6;
    float* vo = out + (long long)v * P * FO;
    for (int s = lane; s < P; s += 32) {
        const float mask = (n > s) ? 1.f : 0.f;
        const float x = vin[s * F], y = vin[s * F + 1], z = vin[s * F + 2];
        float* o = vo + s * FO;
        for (int j = 0; j < F; j++) o[j] = vin[s * F + j] * mask;
        o[F + 0] = (x - mx) * mask; o[F + 1] = (y - my) * mask; o[F + 2] = (z - mz) * mask;
        o[F + 3] = (x - px) * mask; o[F + 4] = (y - py) * mask; o[F + 5] = (z - pz) * mask;
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

extern "C" size_t tcb_voxelize_workspace(int n, int max_vox) {
    if (n < 0) n = 0;
    if (max_vox < 0) max_vox = 0;
    const size_t H = (size_t)next_pow2(2LL * (n > 0 ? n : 1));
    size_t bytes = 4096;
    bytes += 2 * ((size_t)n * 4 + 256);          // pslot, csr
    bytes += 6 * (H * 4 + 256);                  // hkey, hmin, hcnt, hvid, hseg, hfill
    bytes += ((size_t)n / VB + 2) * 4 + 256;     // block sums
    bytes += (size_t)max_vox * 4 + 256;          // vslot
    return bytes;
}

extern "C" int tcb_voxelize(tcb_ctx* ctx, const float* pts_dev, int n, int pt_stride, int nfeat, const double range[6],
                            const double vsize[3], int max_pts, int max_vox, int overflow_mode, int f64_math,
                            int range_mask_xy, float z_offset, int intensity_col, float* voxels_dev, int* coords_dev,
                            int* num_pts_dev, int* m_out_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, range && vsize && voxels_dev && coords_dev && num_pts_dev && m_out_dev, "tcb_voxelize: null argument");
    TCB_REQUIRE(ctx, n >= 0 && (n == 0 || pts_dev), "tcb_voxelize: bad points");
    TCB_REQUIRE(ctx, nfeat >= 3 && pt_stride >= nfeat, "tcb_voxelize: need nfeat >= 3 and pt_stride >= nfeat (got %d, %d)", nfeat, pt_stride);
    TCB_REQUIRE(ctx, max_pts >= 1 && max_vox >= 1, "tcb_voxelize: max_pts and max_vox must be >= 1");
    TCB_REQUIRE(ctx, overflow_mode == 0 || overflow_mode == 1, "tcb_voxelize: overflow_mode must be 0 (continue) or 1 (break)");
    TCB_REQUIRE(ctx, intensity_col < nfeat, "tcb_voxelize: intensity_col out of range");
    if (n > (1 << 26)) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_voxelize: more than 2^26 points");
    cudaStream_t st = (cudaStream_t)stream;

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
    p.H = next_pow2(2LL * (n > 0 ? n : 1));

    const int nblk = tcb_div_up(n > 0 ? n : 1, VB);
    if (nblk > 65536) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_voxelize: too many points");
    TCB_WS(ctx, scalars, int, 64);
    TCB_WS(ctx, pslot, int, n > 0 ? n : 1);
    TCB_WS(ctx, csr, int, n > 0 ? n : 1);
    TCB_WS(ctx, htab, int, (size_t)p.H * 6);
    TCB_WS(ctx, blocksum, int, nblk + 1);
    TCB_WS(ctx, vslot, int, max_vox);
    int *hkey = htab, *hmin = htab + p.H, *hcnt = htab + 2 * (size_t)p.H, *hfill = htab + 3 * (size_t)p.H,
        *hvid = htab + 4 * (size_t)p.H, *hseg = htab + 5 * (size_t)p.H;
    float* imax = reinterpret_cast<float*>(scalars + 8);
    p.intensity_max = imax;

    // hkey = -1, hmin = INT_MAX (0x7f7f7f7f is large enough: indices < 2^26), hcnt = hfill = 0
    TCB_CUDA(ctx, cudaMemsetAsync(hkey, 0xff, (size_t)p.H * 4, st));
    TCB_CUDA(ctx, cudaMemsetAsync(hmin, 0x7f, (size_t)p.H * 4, st));
    TCB_CUDA(ctx, cudaMemsetAsync(hcnt, 0, (size_t)p.H * 8, st));
    if (intensity_col >= 0 && n > 0) {
        TCB_KLAUNCH(ctx, vox_init_max_kernel, 1, 1, 0, st, imax);
        TCB_LAUNCHED(ctx);
        TCB_KLAUNCH(ctx, vox_colmax_kernel, ctx->num_sms * 2, 256, 0, st, pts_dev, n, pt_stride, intensity_col, imax);
        TCB_LAUNCHED(ctx);
    }
    if (n > 0) {
        TCB_KLAUNCH(ctx, vox_hash_kernel, tcb_div_up(n, 256), 256, 0, st, p, pslot, hkey, hmin, hcnt);
        TCB_LAUNCHED(ctx);
        TCB_KLAUNCH(ctx, vox_count_kernel, nblk, VB, 0, st, n, pslot, hmin, blocksum);
        TCB_LAUNCHED(ctx);
    } else {
        TCB_CUDA(ctx, cudaMemsetAsync(blocksum, 0, sizeof(int) * (nblk + 1), st));
    }
    TCB_KLAUNCH(ctx, vox_scan_kernel, 1, 1024, 0, st, nblk, blocksum, max_vox, scalars, m_out_dev);
    TCB_LAUNCHED(ctx);
    if (n > 0) {
        TCB_KLAUNCH(ctx, vox_assign_kernel, nblk, VB, 0, st, p, pslot, hmin, hcnt, blocksum, hvid, hseg, vslot, scalars, coords_dev);
        TCB_LAUNCHED(ctx);
        TCB_KLAUNCH(ctx, vox_scatter_kernel, tcb_div_up(n, 256), 256, 0, st, p, pslot, hvid, hseg, hfill, scalars, csr);
        TCB_LAUNCHED(ctx);
        TCB_KLAUNCH(ctx, vox_gather_kernel, tcb_div_up((long long)max_vox * 32, 256), 256, 0, st, p, m_out_dev, vslot, hseg, hfill, csr, voxels_dev, num_pts_dev);
        TCB_LAUNCHED(ctx);
    }
    return TCB_OK;
}

extern "C" int tcb_pillar_decorate(tcb_ctx* ctx, const float* voxels_dev, const int* coords_dev, const int* num_pts_dev,
                                   const int* m_dev, int m_max, int max_pts, int nfeat, const double vsize[3],
                                   const double range_min[3], float* out_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, voxels_dev && coords_dev && num_pts_dev && out_dev && vsize && range_min, "tcb_pillar_decorate: null argument");
    TCB_REQUIRE(ctx, m_max >= 0 && max_pts >= 1 && nfeat >= 3, "tcb_pillar_decorate: bad shape");
    if (m_max == 0) return TCB_OK;
    // offsets exactly as PillarVFE.__init__ computes them (pillar_vfe.py:76-81): Python doubles, used as f32 scalars
    const float vx = (float)vsize[0], vy = (float)vsize[1], vz = (float)vsize[2];
    const float xo = (float)(vsize[0] / 2 + range_min[0]), yo = (float)(vsize[1] / 2 + range_min[1]), zo = (float)(vsize[2] / 2 + range_min[2]);
    TCB_KLAUNCH(ctx, pillar_decorate_kernel, tcb_div_up((long long)m_max * 32, 256), 256, 0, (cudaStream_t)stream, 
        voxels_dev, coords_dev, num_pts_dev, m_dev, m_max, max_pts, nfeat, vx, vy, vz, xo, yo, zo, out_dev);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
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
