// Evaluation math: what EvaluateInference.calculate_metrics / ap_per_class / compute_ap
// (communicator/evaluate_inference.py:400-446, 158-218, 131-156) do per frame in torch + numpy, for a whole batch of
// frames in two launches, one CTA per frame. All arithmetic is float64 in the reference's operation order (the library
// is built with -fmad=false), so recall/precision/AP/F1 come out identical, not merely close.
#include "common.cuh"
#include <limits.h>
#include <math.h>

#define EV_MAX 2048        // ground truths / predictions per frame
#define EV_T 512
#define EV_IPT (EV_MAX / EV_T)
#define EV_NPX 1000
#define EV_NT 10
#define EV_MAX_CLASSES 256

typedef unsigned long long u64;

// torch.tensor([0.5, 0.55, ...]) is float32: the float64 literals rounded once more.
__constant__ float c_iouv[EV_NT] = {(float)0.5, (float)0.55, (float)0.6, (float)0.65, (float)0.70,
                                    (float)0.75, (float)0.80, (float)0.85, (float)0.9, (float)0.95};

// ---- matching: calculate_metrics :408-425 ----------------------------------------------------------------------------
// The reference lists every (label, detection) pair with IoU >= 0.5 and equal class, orders the list by IoU
// (descending), keeps each detection's first row, then — the list now being in detection order — each label's first row.
// That is: a detection's partner is its highest-IoU label (equal IoUs: the later label, as a stable ascending sort read
// backwards leaves it), and a label keeps the lowest-numbered detection that chose it.
__global__ void __launch_bounds__(256) eval_match_kernel(const double* __restrict__ gts, const int* __restrict__ goff,
                                                         const double* __restrict__ preds, const int* __restrict__ poff,
                                                         unsigned char* __restrict__ correct, int* __restrict__ matched) {
    pdl_wait();
    __shared__ int s_best[EV_MAX];
    __shared__ double s_iou[EV_MAX];
    __shared__ int s_first[EV_MAX];
    __shared__ int s_any;
    const int f = blockIdx.x, tid = threadIdx.x;
    const int g0 = goff[f], G = goff[f + 1] - g0, p0 = poff[f], D = poff[f + 1] - p0;
    for (int i = tid; i < G; i += 256) s_first[i] = INT_MAX;
    if (tid == 0) s_any = 0;
    __syncthreads();
    for (int j = tid; j < D; j += 256) {
        const double* b = preds + (size_t)(p0 + j) * 6;
        const double b0 = b[0], b1 = b[1], b2 = b[2], b3 = b[3], bc = b[5];
        const double area2 = (b2 - b0) * (b3 - b1);
        int best = -1;
        double biou = 0.0;
        for (int i = 0; i < G; i++) {
            const double* a = gts + (size_t)(g0 + i) * 6;
            const double a0 = a[0], a1 = a[1], a2 = a[2], a3 = a[3];
            const double area1 = (a2 - a0) * (a3 - a1);
            const double iw = fmax(fmin(a2, b2) - fmax(a0, b0), 0.0);
            const double ih = fmax(fmin(a3, b3) - fmax(a1, b1), 0.0);
            const double inter = iw * ih;
            const double iou = inter / (area1 + area2 - inter);
            if (iou >= 0.5 && a[5] == bc && (best < 0 || iou >= biou)) { best = i; biou = iou; }
        }
        s_best[j] = best;
        s_iou[j] = biou;
        if (best >= 0) { atomicMin(&s_first[best], j); s_any = 1; }
    }
    __syncthreads();
    for (int j = tid; j < D; j += 256) {
        const int best = s_best[j];
        const bool ok = best >= 0 && s_first[best] == j;
        const float fi = (float)s_iou[j];                 // torch.Tensor(matches): float32 before `>= iouv`
        unsigned char* row = correct + (size_t)(p0 + j) * EV_NT;
#pragma unroll
        for (int t = 0; t < EV_NT; t++) row[t] = (ok && fi >= c_iouv[t]) ? 1 : 0;
    }
    if (tid == 0) matched[f] = s_any;
}

// ---- CTA helpers ------------------------------------------------------------------------------------------------------
// Bitonic sort of n2 (power of two) (key, value) pairs in shared memory. desc: larger key first, equal keys by smaller
// value (= original position: the stable order); asc: smaller key first.
template <bool DESC>
__device__ void cta_bitonic(double* key, int* val, int n2) {
    for (int k = 2; k <= n2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n2; i += EV_T) {
                const int l = i ^ j;
                if (l > i) {
                    const double ka = key[i], kb = key[l];
                    const int va = val[i], vb = val[l];
                    const bool a_first = DESC ? (ka > kb || (ka == kb && va < vb)) : (ka < kb || (ka == kb && va < vb));
                    const bool up = (i & k) == 0;
                    if (up != a_first) { key[i] = kb; key[l] = ka; val[i] = vb; val[l] = va; }
                }
            }
            __syncthreads();
        }
    }
}

// Exclusive scan over the CTA of three packed counters per thread; totals returned to every thread.
__device__ void cta_scan3(u64& a, u64& b, u64& c, u64* wtot /* [3][17] shared */, u64& ta, u64& tb, u64& tc) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    u64 ia = a, ib = b, ic = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u64 xa = __shfl_up_sync(0xffffffffu, ia, o), xb = __shfl_up_sync(0xffffffffu, ib, o), xc = __shfl_up_sync(0xffffffffu, ic, o);
        if (lane >= o) { ia += xa; ib += xb; ic += xc; }
    }
    if (lane == 31) { wtot[w] = ia; wtot[17 + w] = ib; wtot[34 + w] = ic; }
    __syncthreads();
    if (threadIdx.x == 0) {
        u64 sa = 0, sb = 0, sc = 0;
        for (int i = 0; i < EV_T / 32; i++) {
            const u64 va = wtot[i], vb = wtot[17 + i], vc = wtot[34 + i];
            wtot[i] = sa; wtot[17 + i] = sb; wtot[34 + i] = sc;
            sa += va; sb += vb; sc += vc;
        }
        wtot[16] = sa; wtot[33] = sb; wtot[50] = sc;
    }
    __syncthreads();
    a = wtot[w] + ia - a; b = wtot[17 + w] + ib - b; c = wtot[34 + w] + ic - c;
    ta = wtot[16]; tb = wtot[33]; tc = wtot[50];
    __syncthreads();
}

// numpy.interp (arr_interp in numpy/_core/src/multiarray/compiled_base.c) for one abscissa: j = the last index with
// xp[j] <= x; fp[j] on a node or at the last node; otherwise slope * (x - xp[j]) + fp[j].
template <class XP, class FP>
__device__ double interp_np(double x, int n, XP xp, FP fp, double left, double right) {
    if (n == 1) {
        const double xv = xp(0);
        return x < xv ? left : (x > xv ? right : fp(0));
    }
    if (x < xp(0)) return left;
    if (x > xp(n - 1)) return right;
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (xp(mid) <= x) lo = mid; else hi = mid;
    }
    if (lo == n - 1) return fp(lo);
    const double xj = xp(lo);
    if (xj == x) return fp(lo);
    const double fj = fp(lo);
    const double slope = (fp(lo + 1) - fj) / (xp(lo + 1) - xj);
    return slope * (x - xj) + fj;
}

struct ArrayAt { const double* p; __device__ double operator()(int i) const { return p[i]; } };
// recall / precision of the class's k-th prediction (0-based): tpc/(n_l + 1e-16), tpc/(tpc + fpc) with tpc + fpc = k + 1.
struct RecallAt { const unsigned short* tpc; double denom; __device__ double operator()(int k) const { return (double)tpc[k] / denom; } };
struct PrecisionAt { const unsigned short* tpc; __device__ double operator()(int k) const { return (double)tpc[k] / (double)(k + 1); } };
// compute_ap's curves with the sentinels: mrec = [0, recall.., 1]; the envelope lives in shared memory.
struct MrecAt {
    const unsigned short* tpc; double denom; int last;
    __device__ double operator()(int k) const { return k == 0 ? 0.0 : (k == last ? 1.0 : (double)tpc[k - 1] / denom); }
};

struct ApArgs {
    const unsigned char* tp; int nt;
    const double* conf; const double* pcls; long long pstride; const int* poff;
    const double* tcls; long long tstride; const int* toff;
    const int* frame_valid;
    int max_classes;
    double* rows;                 // scratch [frames][max_classes][2][EV_NPX]: precision and recall sampled on px
    int* num_classes; int* classes; double* p; double* r; double* f1; double* ap;
};

struct ApShared {
    double skey[EV_MAX];                  // confidences, descending
    double scls[EV_MAX];                  // predicted class in that order
    double cconf[EV_MAX];                 // the current class's -conf (ascending); first used to sort the target classes
    double env[EV_MAX + 8];               // precision envelope incl. sentinels
    double f1sum[EV_NPX];
    double ucls[EV_MAX_CLASSES];
    double y[104];
    double yw[EV_NT][104];                // interpolated envelope per threshold (one warp per threshold)
    u64 wtot[51];
    int sidx[EV_MAX];
    int ucount[EV_MAX_CLASSES];
    int ustart[EV_MAX_CLASSES + 1];
    unsigned short ctp[EV_NT][EV_MAX];    // inclusive TP counts of the current class per threshold; first an int scratch
    unsigned short stp[EV_MAX];           // TP bits per prediction, sorted order
    unsigned char skip[EV_MAX_CLASSES];
    int best_i;
};

__device__ __forceinline__ double px_at(int t) { return t == EV_NPX - 1 ? 1.0 : (double)t * (1.0 / (EV_NPX - 1)); }   // np.linspace(0, 1, 1000)
__device__ __forceinline__ double x101_at(int t) { return t == 100 ? 1.0 : (double)t * (1.0 / 100); }                 // np.linspace(0, 1, 101)

__global__ void __launch_bounds__(EV_T) eval_ap_kernel(ApArgs A) {
    pdl_wait();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ApShared& S = *reinterpret_cast<ApShared*>(smem_raw);
    const int f = blockIdx.x, tid = threadIdx.x;
    if (A.frame_valid && !A.frame_valid[f]) {
        if (tid == 0) A.num_classes[f] = 0;
        return;
    }
    const int p0 = A.poff[f], n = A.poff[f + 1] - p0, t0 = A.toff[f], m = A.toff[f + 1] - t0;
    const int nt = A.nt;

    // (1) predictions in descending confidence, original order among equals (np.argsort(-conf), :174-175)
    int n2 = 1;
    while (n2 < n) n2 <<= 1;
    for (int k = tid; k < n2; k += EV_T) {
        S.skey[k] = k < n ? A.conf[(size_t)(p0 + k) * A.pstride] : -INFINITY;
        S.sidx[k] = k < n ? k : INT_MAX;
    }
    __syncthreads();
    cta_bitonic<true>(S.skey, S.sidx, n2);
    for (int k = tid; k < n; k += EV_T) {
        const size_t row = (size_t)(p0 + S.sidx[k]);
        S.scls[k] = A.pcls[row * A.pstride];
        unsigned bits = 0;
        for (int t = 0; t < nt; t++) bits |= (A.tp[row * nt + t] ? 1u : 0u) << t;
        S.stp[k] = (unsigned short)bits;
    }
    for (int t = tid; t < EV_NPX; t += EV_T) S.f1sum[t] = 0.0;

    // (2) np.unique(target_cls) (:178) with each class's label count
    int m2 = 1;
    while (m2 < m) m2 <<= 1;
    int* tmpi = reinterpret_cast<int*>(&S.ctp[0][0]);
    for (int k = tid; k < m2; k += EV_T) {
        S.cconf[k] = k < m ? A.tcls[(size_t)(t0 + k) * A.tstride] : INFINITY;
        tmpi[k] = k;
    }
    __syncthreads();
    cta_bitonic<false>(S.cconf, tmpi, m2);
    int nc;
    {
        u64 a = 0, b = 0, c = 0, ta, tb, tc;
        bool flag[EV_IPT];
#pragma unroll
        for (int q = 0; q < EV_IPT; q++) {
            const int k = tid * EV_IPT + q;
            flag[q] = k < m && (k == 0 || S.cconf[k] != S.cconf[k - 1]);
            a += flag[q];
        }
        cta_scan3(a, b, c, S.wtot, ta, tb, tc);
        nc = (int)ta;
        int pos = (int)a;
#pragma unroll
        for (int q = 0; q < EV_IPT; q++) {
            const int k = tid * EV_IPT + q;
            if (flag[q]) {
                if (pos < EV_MAX_CLASSES) { S.ucls[pos] = S.cconf[k]; S.ustart[pos] = k; }
                pos++;
            }
        }
        __syncthreads();
        const int ns = min(nc, EV_MAX_CLASSES);
        for (int c2 = tid; c2 < ns; c2 += EV_T) S.ucount[c2] = (c2 + 1 < nc && c2 + 1 < EV_MAX_CLASSES ? S.ustart[c2 + 1] : m) - S.ustart[c2];
        __syncthreads();
    }
    const int ncp = min(nc, A.max_classes);      // classes with room in the outputs (the host checks num_classes)
    double* rows = A.rows + (size_t)f * A.max_classes * 2 * EV_NPX;

    // (3) per class: cumulative TP/FP, sampled curves, AP per threshold (:185-207)
    for (int ci = 0; ci < ncp; ci++) {
        const double cls = S.ucls[ci];
        const int n_l = S.ucount[ci];
        u64 a = 0, b = 0, c = 0, ta, tb, tc;
        bool mem[EV_IPT];
        unsigned bits[EV_IPT];
#pragma unroll
        for (int q = 0; q < EV_IPT; q++) {
            const int k = tid * EV_IPT + q;
            mem[q] = k < n && S.scls[k] == cls;
            bits[q] = mem[q] ? S.stp[k] : 0u;
            // 12-bit fields: a = members, tp0..tp3; b = tp4..tp8; c = tp9
            a += (u64)mem[q] | ((u64)(bits[q] & 1) << 12) | ((u64)((bits[q] >> 1) & 1) << 24) | ((u64)((bits[q] >> 2) & 1) << 36) |
                 ((u64)((bits[q] >> 3) & 1) << 48);
            b += (u64)((bits[q] >> 4) & 1) | ((u64)((bits[q] >> 5) & 1) << 12) | ((u64)((bits[q] >> 6) & 1) << 24) |
                 ((u64)((bits[q] >> 7) & 1) << 36) | ((u64)((bits[q] >> 8) & 1) << 48);
            c += (u64)((bits[q] >> 9) & 1);
        }
        cta_scan3(a, b, c, S.wtot, ta, tb, tc);
        const int n_p = (int)(ta & 0xfff);
#pragma unroll
        for (int q = 0; q < EV_IPT; q++) {
            const int k = tid * EV_IPT + q;
            if (mem[q]) {
                a += 1ull | ((u64)(bits[q] & 1) << 12) | ((u64)((bits[q] >> 1) & 1) << 24) | ((u64)((bits[q] >> 2) & 1) << 36) |
                     ((u64)((bits[q] >> 3) & 1) << 48);
                b += (u64)((bits[q] >> 4) & 1) | ((u64)((bits[q] >> 5) & 1) << 12) | ((u64)((bits[q] >> 6) & 1) << 24) |
                     ((u64)((bits[q] >> 7) & 1) << 36) | ((u64)((bits[q] >> 8) & 1) << 48);
                c += (u64)((bits[q] >> 9) & 1);
                const int pos = (int)(a & 0xfff) - 1;      // a now holds inclusive counts
                S.cconf[pos] = -S.skey[k];
                S.ctp[0][pos] = (unsigned short)((a >> 12) & 0xfff);
                S.ctp[1][pos] = (unsigned short)((a >> 24) & 0xfff);
                S.ctp[2][pos] = (unsigned short)((a >> 36) & 0xfff);
                S.ctp[3][pos] = (unsigned short)((a >> 48) & 0xfff);
                S.ctp[4][pos] = (unsigned short)(b & 0xfff);
                S.ctp[5][pos] = (unsigned short)((b >> 12) & 0xfff);
                S.ctp[6][pos] = (unsigned short)((b >> 24) & 0xfff);
                S.ctp[7][pos] = (unsigned short)((b >> 36) & 0xfff);
                S.ctp[8][pos] = (unsigned short)((b >> 48) & 0xfff);
                S.ctp[9][pos] = (unsigned short)(c & 0xfff);
            }
        }
        if (tid == 0) S.skip[ci] = (n_p == 0 || n_l == 0);
        __syncthreads();
        double* prow = rows + (size_t)ci * 2 * EV_NPX;
        double* rrow = prow + EV_NPX;
        if (n_p == 0 || n_l == 0) {
            for (int j = tid; j < nt; j += EV_T) A.ap[((size_t)f * A.max_classes + ci) * nt + j] = 0.0;
            continue;                       // p, r rows stay zero (:189-190): F1 contributes 0
        }
        const double denom = (double)n_l + 1e-16;
        const ArrayAt xp{S.cconf};
        const RecallAt rec0{S.ctp[0], denom};
        const PrecisionAt pre0{S.ctp[0]};
        const double rlast = rec0(n_p - 1), plast = pre0(n_p - 1);
        for (int t = tid; t < EV_NPX; t += EV_T) {
            const double x = -px_at(t);
            const double rv = interp_np(x, n_p, xp, rec0, 0.0, rlast);
            const double pv = interp_np(x, n_p, xp, pre0, 1.0, plast);
            prow[t] = pv;
            rrow[t] = rv;
            S.f1sum[t] = S.f1sum[t] + 2 * pv * rv / (pv + rv + 1e-16);       // f1.mean(0) adds the class rows in order
        }
        const int L = n_p + 2;
        // Short curves (the usual case: a few dozen predictions per class and frame): the envelopes of all thresholds fit
        // side by side, so warp j does threshold j on its own — suffix maximum by lane chunks + one shuffle scan (max is
        // exact in any order), the 101 interpolated points, the pairwise trapezoid sum — with no CTA barrier in between.
        const bool side_by_side = nt * L <= EV_MAX + 8;
        if (side_by_side) {
            const int wj = tid >> 5, lane = tid & 31;
            if (wj < nt) {
                const unsigned short* tpc = S.ctp[wj];
                double* env = S.env + wj * L;
                double* y = S.yw[wj];
                for (int k = lane; k < L; k += 32) env[k] = k == 0 ? 1.0 : (k == L - 1 ? 0.0 : (double)tpc[k - 1] / (double)k);
                __syncwarp();
                const int C = (L + 31) / 32, a = lane * C, b = min(L, a + C);
                double m = -INFINITY;
                for (int k = b - 1; k >= a; k--) { m = fmax(m, env[k]); env[k] = m; }
                double sfx = m;                                        // maximum over this lane's chunk and all later ones
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const double t = __shfl_down_sync(0xffffffffu, sfx, o);
                    if (lane + o < 32) sfx = fmax(sfx, t);
                }
                double later = __shfl_down_sync(0xffffffffu, sfx, 1);
                if (lane == 31) later = -INFINITY;
                for (int k = a; k < b; k++) env[k] = fmax(env[k], later);
                __syncwarp();
                const MrecAt mrec{tpc, denom, L - 1};
                const ArrayAt mpre{env};
                for (int t = lane; t < 101; t += 32) y[t] = interp_np(x101_at(t), L, mrec, mpre, env[0], env[L - 1]);
                __syncwarp();
                auto term = [&](int k) { return (x101_at(k + 1) - x101_at(k)) * (y[k + 1] + y[k]) / 2.0; };
                double acc = 0.0;
                if (lane < 8) {
                    acc = term(lane);
                    for (int k = lane + 8; k < 96; k += 8) acc = acc + term(k);
                }
                acc = acc + __shfl_down_sync(0xffffffffu, acc, 1);
                acc = acc + __shfl_down_sync(0xffffffffu, acc, 2);
                acc = acc + __shfl_down_sync(0xffffffffu, acc, 4);
                if (lane == 0) {
                    for (int k = 96; k < 100; k++) acc = acc + term(k);
                    A.ap[((size_t)f * A.max_classes + ci) * nt + wj] = acc;
                }
            }
            __syncthreads();
        }
        for (int j = 0; j < (side_by_side ? 0 : nt); j++) {
            const unsigned short* tpc = S.ctp[j];
            // mpre = [1, precision.., 0], then the running maximum from the right (:143-148)
            for (int k = tid; k < L; k += EV_T) S.env[k] = k == 0 ? 1.0 : (k == L - 1 ? 0.0 : (double)tpc[k - 1] / (double)k);
            __syncthreads();
            for (int off = 1; off < L; off <<= 1) {
                double v[EV_IPT + 1];
#pragma unroll
                for (int q = 0; q < EV_IPT + 1; q++) {
                    const int k = tid + q * EV_T;
                    v[q] = (k + off < L) ? fmax(S.env[k], S.env[k + off]) : (k < L ? S.env[k] : 0.0);
                }
                __syncthreads();
#pragma unroll
                for (int q = 0; q < EV_IPT + 1; q++) {
                    const int k = tid + q * EV_T;
                    if (k < L) S.env[k] = v[q];
                }
                __syncthreads();
            }
            if (tid < 101) {
                const MrecAt mrec{tpc, denom, L - 1};
                const ArrayAt mpre{S.env};
                S.y[tid] = interp_np(x101_at(tid), L, mrec, mpre, S.env[0], S.env[L - 1]);
            }
            __syncthreads();
            if (tid < 32) {
                // np.trapz: (d * (y[1:] + y[:-1]) / 2.0).sum() — numpy adds a contiguous float64 vector of 100 with eight
                // running sums over the first 96 terms (sum q takes terms q, q+8, ...), combines them pairwise, then adds
                // the last four in order. Lanes 0-7 are the eight sums.
                auto term = [&](int k) { return (x101_at(k + 1) - x101_at(k)) * (S.y[k + 1] + S.y[k]) / 2.0; };
                double acc = 0.0;
                if (tid < 8) {
                    acc = term(tid);
                    for (int k = tid + 8; k < 96; k += 8) acc = acc + term(k);
                }
                acc = acc + __shfl_down_sync(0xffffffffu, acc, 1);      // lanes 0,2,4,6: r0+r1, r2+r3, r4+r5, r6+r7
                acc = acc + __shfl_down_sync(0xffffffffu, acc, 2);      // lanes 0,4
                acc = acc + __shfl_down_sync(0xffffffffu, acc, 4);      // lane 0
                if (tid == 0) {
                    for (int k = 96; k < 100; k++) acc = acc + term(k);
                    A.ap[((size_t)f * A.max_classes + ci) * nt + j] = acc;
                }
            }
            __syncthreads();
        }
    }
    __syncthreads();

    // (4) the px index with the largest mean F1, first one on ties (:215), and the per-class values there
    if (ncp > 0) {
        double bv = -INFINITY;
        int bi = INT_MAX;
        for (int t = tid; t < EV_NPX; t += EV_T) {
            const double v = S.f1sum[t] / (double)nc;
            if (v > bv) { bv = v; bi = t; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if ((tid & 31) == 0) { S.y[tid >> 5] = bv; S.sidx[tid >> 5] = bi; }
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < EV_T / 32; w++)
                if (S.y[w] > bv || (S.y[w] == bv && S.sidx[w] < bi)) { bv = S.y[w]; bi = S.sidx[w]; }
            S.best_i = bi == INT_MAX ? 0 : bi;
        }
        __syncthreads();
        const int bi2 = S.best_i;
        for (int ci = tid; ci < ncp; ci += EV_T) {
            double pv = 0.0, rv = 0.0;
            if (!S.skip[ci]) { pv = rows[(size_t)ci * 2 * EV_NPX + bi2]; rv = rows[(size_t)ci * 2 * EV_NPX + EV_NPX + bi2]; }
            const size_t o = (size_t)f * A.max_classes + ci;
            A.p[o] = pv;
            A.r[o] = rv;
            A.f1[o] = 2 * pv * rv / (pv + rv + 1e-16);
            A.classes[o] = (int)S.ucls[ci];
        }
    }
    if (tid == 0) A.num_classes[f] = nc;
}

// ---- host side --------------------------------------------------------------------------------------------------------
static int check_offsets(tcb_ctx* ctx, const char* who, const int* off, int frames, const char* what) {
    if (!off || off[0] != 0) return tcb_fail(ctx, TCB_ERR_ARG, "%s: %s offsets must start at 0", who, what);
    for (int f = 0; f < frames; f++) {
        const int c = off[f + 1] - off[f];
        if (c < 0) return tcb_fail(ctx, TCB_ERR_ARG, "%s: %s offsets decrease at frame %d", who, what, f);
        if (c > EV_MAX) return tcb_fail(ctx, TCB_ERR_LIMIT, "%s: frame %d has %d %s (limit %d)", who, f, c, what, EV_MAX);
    }
    return TCB_OK;
}

static int launch_match(tcb_ctx* ctx, const double* gts, const int* goff_dev, const double* preds, const int* poff_dev, int frames,
                        unsigned char* correct, int* matched, cudaStream_t st) {
    TCB_KLAUNCH(ctx, eval_match_kernel, frames, 256, 0, st, gts, goff_dev, preds, poff_dev, correct, matched);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

static int init_eval(tcb_ctx* ctx) {
    TCB_SMEM_OPT_IN(ctx, eval_ap_kernel, sizeof(ApShared));
    return TCB_OK;
}
static TcbInitReg reg_eval(&init_eval);

static int launch_ap(tcb_ctx* ctx, ApArgs& a, int frames, cudaStream_t st) {
    TCB_KLAUNCH(ctx, eval_ap_kernel, frames, EV_T, sizeof(ApShared), st, a);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}

static int upload_offsets(tcb_ctx* ctx, const int* host, int frames, int* dev, cudaStream_t st) {
    return tcb_stage_h2d(ctx, dev, host, sizeof(int) * (size_t)(frames + 1), st);
}

extern "C" size_t tcb_eval_workspace(int frames, int max_classes) {
    if (frames < 0 || max_classes < 0) return 0;
    return (size_t)frames * (size_t)max_classes * 2 * EV_NPX * sizeof(double) + 2 * (size_t)(frames + 1) * sizeof(int) + 4096;
}

extern "C" int tcb_eval_match(tcb_ctx* ctx, const double* gts_dev, const int* gt_offsets_host, const double* preds_dev,
                              const int* pred_offsets_host, int frames, unsigned char* correct_dev, int* matched_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, frames >= 0 && correct_dev && matched_dev, "tcb_eval_match: bad arguments");
    if (frames == 0) return TCB_OK;
    int rc;
    if ((rc = check_offsets(ctx, "tcb_eval_match", gt_offsets_host, frames, "ground truths"))) return rc;
    if ((rc = check_offsets(ctx, "tcb_eval_match", pred_offsets_host, frames, "predictions"))) return rc;
    TCB_REQUIRE(ctx, (gts_dev || gt_offsets_host[frames] == 0) && (preds_dev || pred_offsets_host[frames] == 0), "tcb_eval_match: null input");
    cudaStream_t st = (cudaStream_t)stream;
    TCB_WS(ctx, goff, int, frames + 1);
    TCB_WS(ctx, poff, int, frames + 1);
    if ((rc = upload_offsets(ctx, gt_offsets_host, frames, goff, st))) return rc;
    if ((rc = upload_offsets(ctx, pred_offsets_host, frames, poff, st))) return rc;
    return launch_match(ctx, gts_dev, goff, preds_dev, poff, frames, correct_dev, matched_dev, st);
}

static int fill_ap_args(tcb_ctx* ctx, const char* who, ApArgs& a, int nt, int frames, int max_classes, int* num_classes_dev,
                        int* classes_dev, double* p_dev, double* r_dev, double* f1_dev, double* ap_dev) {
    TCB_REQUIRE(ctx, nt >= 1 && nt <= EV_NT, "%s: %d IoU thresholds (1..%d supported)", who, nt, EV_NT);
    TCB_REQUIRE(ctx, max_classes >= 1 && max_classes <= EV_MAX_CLASSES, "%s: max_classes %d (1..%d supported)", who, max_classes, EV_MAX_CLASSES);
    TCB_REQUIRE(ctx, num_classes_dev && classes_dev && p_dev && r_dev && f1_dev && ap_dev, "%s: null output", who);
    a.nt = nt; a.max_classes = max_classes;
    a.num_classes = num_classes_dev; a.classes = classes_dev; a.p = p_dev; a.r = r_dev; a.f1 = f1_dev; a.ap = ap_dev;
    return TCB_OK;
}

extern "C" int tcb_ap_per_class(tcb_ctx* ctx, const unsigned char* tp_dev, int nt, const double* conf_dev, const double* pred_cls_dev,
                                long long pred_stride, const int* pred_offsets_host, const double* target_cls_dev, long long target_stride,
                                const int* target_offsets_host, int frames, int max_classes, int* num_classes_dev, int* classes_dev,
                                double* p_dev, double* r_dev, double* f1_dev, double* ap_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, frames >= 0 && pred_stride >= 1 && target_stride >= 1, "tcb_ap_per_class: bad arguments");
    if (frames == 0) return TCB_OK;
    int rc;
    ApArgs a;
    if ((rc = fill_ap_args(ctx, "tcb_ap_per_class", a, nt, frames, max_classes, num_classes_dev, classes_dev, p_dev, r_dev, f1_dev, ap_dev))) return rc;
    if ((rc = check_offsets(ctx, "tcb_ap_per_class", pred_offsets_host, frames, "predictions"))) return rc;
    if ((rc = check_offsets(ctx, "tcb_ap_per_class", target_offsets_host, frames, "targets"))) return rc;
    TCB_REQUIRE(ctx, (pred_offsets_host[frames] == 0 || (tp_dev && conf_dev && pred_cls_dev)) && (target_offsets_host[frames] == 0 || target_cls_dev),
                "tcb_ap_per_class: null input");
    cudaStream_t st = (cudaStream_t)stream;
    TCB_WS(ctx, poff, int, frames + 1);
    TCB_WS(ctx, toff, int, frames + 1);
    TCB_WS(ctx, rows, double, (size_t)frames * max_classes * 2 * EV_NPX);
    if ((rc = upload_offsets(ctx, pred_offsets_host, frames, poff, st))) return rc;
    if ((rc = upload_offsets(ctx, target_offsets_host, frames, toff, st))) return rc;
    a.tp = tp_dev; a.conf = conf_dev; a.pcls = pred_cls_dev; a.pstride = pred_stride; a.poff = poff;
    a.tcls = target_cls_dev; a.tstride = target_stride; a.toff = toff; a.frame_valid = nullptr; a.rows = rows;
    return launch_ap(ctx, a, frames, st);
}

extern "C" int tcb_calculate_metrics(tcb_ctx* ctx, const double* gts_dev, const int* gt_offsets_host, const double* preds_dev,
                                     const int* pred_offsets_host, int frames, int max_classes, unsigned char* correct_dev,
                                     int* matched_dev, int* num_classes_dev, int* classes_dev, double* p_dev, double* r_dev,
                                     double* f1_dev, double* ap_dev, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, frames >= 0 && correct_dev && matched_dev, "tcb_calculate_metrics: bad arguments");
    if (frames == 0) return TCB_OK;
    int rc;
    ApArgs a;
    if ((rc = fill_ap_args(ctx, "tcb_calculate_metrics", a, EV_NT, frames, max_classes, num_classes_dev, classes_dev, p_dev, r_dev, f1_dev, ap_dev))) return rc;
    if ((rc = check_offsets(ctx, "tcb_calculate_metrics", gt_offsets_host, frames, "ground truths"))) return rc;
    if ((rc = check_offsets(ctx, "tcb_calculate_metrics", pred_offsets_host, frames, "predictions"))) return rc;
    TCB_REQUIRE(ctx, (gts_dev || gt_offsets_host[frames] == 0) && (preds_dev || pred_offsets_host[frames] == 0), "tcb_calculate_metrics: null input");
    cudaStream_t st = (cudaStream_t)stream;
    TCB_WS(ctx, goff, int, frames + 1);
    TCB_WS(ctx, poff, int, frames + 1);
    TCB_WS(ctx, rows, double, (size_t)frames * max_classes * 2 * EV_NPX);
    if ((rc = upload_offsets(ctx, gt_offsets_host, frames, goff, st))) return rc;
    if ((rc = upload_offsets(ctx, pred_offsets_host, frames, poff, st))) return rc;
    if ((rc = launch_match(ctx, gts_dev, goff, preds_dev, poff, frames, correct_dev, matched_dev, st))) return rc;
    // ap_per_class on each frame's own rows (:428-436): tp = correct, conf/class = prediction columns 4/5, targets = column 5
    a.tp = correct_dev; a.conf = preds_dev + 4; a.pcls = preds_dev + 5; a.pstride = 6; a.poff = poff;
    a.tcls = gts_dev + 5; a.tstride = 6; a.toff = goff; a.frame_valid = matched_dev; a.rows = rows;
    return launch_ap(ctx, a, frames, st);
}
