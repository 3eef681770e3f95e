// Baseline JPEG decode on device: the compressed-image branch of the 2D callback,
//   buf = np.ndarray(shape=(1, len(msg.data)), dtype=np.uint8, buffer=msg.data)
//   cv_image = cv2.imdecode(buf, cv2.IMREAD_ANYCOLOR)[..., ::-1]            (communicator/ros_inference.py:123-131)
// so that CompressedImage frames reach tcb_preprocess_image without a host decode. cv2.imdecode hands the bytes to
// libjpeg-turbo with default settings; the result here is bit-identical to it: Huffman entropy decoding, dequantisation
// + the "islow" integer inverse DCT (13-bit constants, two passes), "fancy" triangle-filter chroma upsampling and the
// 16-bit fixed-point YCbCr -> RGB conversion.
//
//   host    : marker parse (tables, frame, restart interval), entropy-coded bytes unstuffed (FF00 -> FF) and cut at the
//             restart markers into the caller's pinned staging buffer, one H2D copy
//   huffman : one CTA per image. The entropy-coded segment is serial by nature; it is cut into 128-byte subsequences that
//             are decoded in parallel from guessed states and re-decoded from their predecessor's exit state (bit
//             position, block inside the MCU, zig-zag index) until nothing changes — Huffman streams re-synchronise within
//             a few symbols, so a handful of rounds settle thousands of subsequences (Weissenberger & Schmidt's
//             self-synchronising scheme). A scan of what every subsequence advances — coefficient slots and, per
//             component, the sum of its DC differences — gives the write positions and DC predictors; a last pass
//             writes the coefficients.
//   idct    : one thread per 8x8 block: dequantise, two 8-point passes, clamp, store into the component's plane
//   colour  : four pixels per thread: upsample the chroma on the fly, convert, store RGB or BGR bytes
#include "common.cuh"
#include <stdlib.h>
#include <new>
#include <thread>
#include <vector>

namespace {

constexpr int SUB_BYTES = 128, SUB_BITS = SUB_BYTES * 8;   // subsequence of the entropy-coded stream
constexpr int HUFF_T = 1024;
constexpr int MAX_BPM = 6;                                  // blocks per MCU (4:2:0: 4 + 1 + 1)

struct JHuff {
    unsigned short lut[512];       // 9-bit prefix -> (length << 8) | symbol, 0 = longer code
    int maxcode[18], valptr[17], mincode[17];
    unsigned char huffval[256];
};
struct JSeg { int byte_off, byte_len, mcu_first, sub_first; };   // a restart interval of an image (offsets inside the image)
struct JpegHdr {
    int width, height, ncomp, hmax, vmax, mcux, mcuy, ri, bpm;
    int comp_h[3], comp_v[3], comp_tq[3], comp_td[3], comp_ta[3], comp_bw[3], comp_bh[3], comp_first_j[3];
    int up_hf[3], up_vf[3], up_dw[3], up_dh[3];   // upsampling factors and downsampled size of every component (colour kernel)
    int mcu_comp[MAX_BPM], mcu_bx[MAX_BPM], mcu_by[MAX_BPM];
    long long scan_off;            // bytes, inside the batch's scan buffer (16-byte aligned)
    int scan_len, seg_first, nseg, sub_first, nsub;
    long long coef_off, blocks;    // coefficients (int16) of the image: blocks * 64
    long long plane_off[3];
    unsigned short qt[4][64];      // natural order
    JHuff dc[4], ac[4];
};

__constant__ unsigned char c_zigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                                           35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
static const unsigned char h_zigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                                           35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// ---------------------------------------------------------------------------------------------------------------------
// entropy decoding
// ---------------------------------------------------------------------------------------------------------------------
struct HuffShared {
    unsigned short lut[2 * 3][512];          // [dc/ac][component]
    int maxcode[6][18], valptr[6][17], mincode[6][17];
    unsigned char huffval[6][256];
    int mcu_comp[MAX_BPM];
    int changed;
    int warp_tot[32];
};

// the 33+ bits of the stream that start at bit `bp` of the segment (left-aligned in a 64-bit word)
__device__ __forceinline__ unsigned long long peek64(const unsigned char* seg, long long bp) {
    const unsigned char* p = seg + (bp >> 3);
    const unsigned* w = reinterpret_cast<const unsigned*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
    const unsigned long long hi = __byte_perm(w[0], 0, 0x0123), lo = __byte_perm(w[1], 0, 0x0123);
    const int sh = (int)((reinterpret_cast<uintptr_t>(p) & 3) * 8 + (bp & 7));
    return ((hi << 32) | lo) << sh;          // >= 33 valid bits on top
}

struct DecState { int p, c, z; };   // bit position in the segment, block inside the MCU, zig-zag index (0: a DC symbol comes next)

// Decodes from `st` until the bit position reaches end_bit. dc[comp] accumulates the DC differences met: with WRITE it
// comes in as the predictor of each component at `st` and the coefficients are stored (DC as a value) for blocks
// [.., block_end) at coef + block*64, `block` being the absolute block index of st; without WRITE it comes in as zero
// and leaves as the span's DC sums. Returns the number of coefficient slots advanced.
template <bool WRITE>
__device__ __forceinline__ int decode_span(const HuffShared& S, const unsigned char* seg, DecState& st, int end_bit, int bpm,
                                           short* coef, long long block, long long block_end, int (&dc)[3]) {
    int p = st.p, c = st.c, z = st.z, blocks_done = 0;
    const int z0 = z;
    while (p < end_bit) {
        const int comp = S.mcu_comp[c];
        const int t = (z == 0 ? 0 : 3) + comp;
        const unsigned long long win = peek64(seg, p);
        const unsigned short e = S.lut[t][(unsigned)(win >> 55)];
        int len = e >> 8, sym = e & 255;
        if (len == 0) {
            sym = 0; len = 16;                               // no code matches: libjpeg-turbo yields symbol 0
            for (int l = 10; l <= 16; l++) {
                const int code = (int)(win >> (64 - l));
                if (code <= S.maxcode[t][l] && code >= S.mincode[t][l]) { sym = S.huffval[t][S.valptr[t][l] + code - S.mincode[t][l]]; len = l; break; }
            }
        }
        const int sz = z == 0 ? sym : (sym & 15), run = z == 0 ? 0 : (sym >> 4);
        int val = 0;
        if (sz) {
            const int raw = (int)((win << len) >> (64 - sz));
            val = raw < (1 << (sz - 1)) ? raw - (1 << sz) + 1 : raw;      // HUFF_EXTEND
        }
        p += len + sz;
        bool block_end_now = false;
        if (z == 0) {
            if (comp == 0) dc[0] += val; else if (comp == 1) dc[1] += val; else dc[2] += val;      // (kept in registers)
            if (WRITE && block < block_end) coef[block * 64] = (short)(comp == 0 ? dc[0] : (comp == 1 ? dc[1] : dc[2]));
            z = 1;
        } else if (sz == 0) {
            if (run == 15) { z += 16; if (z > 63) block_end_now = true; }   // ZRL
            else block_end_now = true;                                       // EOB
        } else {
            z += run;
            if (z > 63) block_end_now = true;                                // corrupt run: the block ends (only ever seen while still unsynchronised)
            else {
                if (WRITE && block < block_end) coef[block * 64 + c_zigzag[z]] = (short)val;
                z++;
                if (z > 63) block_end_now = true;
            }
        }
        if (block_end_now) { z = 0; c = c + 1 == bpm ? 0 : c + 1; blocks_done++; block++; }
    }
    st.p = p; st.c = c; st.z = z;
    return blocks_done * 64 + z - z0;
}

__device__ __forceinline__ unsigned long long pack_state(const DecState& s) { return ((unsigned long long)(unsigned)s.p << 16) | ((unsigned)s.c << 8) | (unsigned)s.z; }
__device__ __forceinline__ DecState unpack_state(unsigned long long v) { DecState s; s.p = (int)(v >> 16); s.c = (int)((v >> 8) & 255); s.z = (int)(v & 255); return s; }

__global__ void __launch_bounds__(HUFF_T, 1) jpeg_huffman_kernel(const JpegHdr* __restrict__ hdrs, const unsigned char* __restrict__ scan,
                                                                 const JSeg* __restrict__ segs, unsigned long long* __restrict__ exit_st,
                                                                 unsigned long long* __restrict__ used_in, int4* __restrict__ adv, int4* __restrict__ gpos,
                                                                 short* __restrict__ coef_all) {
    extern __shared__ __align__(16) unsigned char hs_raw[];
    HuffShared& S = *reinterpret_cast<HuffShared*>(hs_raw);
    pdl_wait();
    const JpegHdr& H = hdrs[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // the image's tables: [0..2] DC of component 0..2, [3..5] AC
    for (int t = 0; t < 2 * H.ncomp; t++) {
        const int comp = t % H.ncomp, ac = t / H.ncomp;
        const JHuff& src = ac ? H.ac[H.comp_ta[comp]] : H.dc[H.comp_td[comp]];
        const int d = ac * 3 + comp;
        for (int i = tid; i < 512; i += HUFF_T) S.lut[d][i] = src.lut[i];
        for (int i = tid; i < 256; i += HUFF_T) S.huffval[d][i] = src.huffval[i];
        if (tid < 18) S.maxcode[d][tid] = src.maxcode[tid];
        if (tid < 17) { S.valptr[d][tid] = src.valptr[tid]; S.mincode[d][tid] = src.mincode[tid]; }
    }
    if (tid < MAX_BPM) S.mcu_comp[tid] = H.mcu_comp[tid];
    __syncthreads();
    const unsigned char* img_scan = scan + H.scan_off;
    const JSeg* sg = segs + H.seg_first;
    exit_st += H.sub_first; used_in += H.sub_first; adv += H.sub_first; gpos += H.sub_first;
    short* coef = coef_all + H.coef_off;
    const int nsub = H.nsub, bpm = H.bpm;
    const unsigned long long UNUSED = ~0ull;

    auto seg_of = [&](int j) {                       // the segment that holds subsequence j (segments are in sub order)
        int lo = 0, hi = H.nseg - 1;
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (sg[mid].sub_first <= j) lo = mid; else hi = mid - 1; }
        return lo;
    };

    // ---- rounds: every subsequence decodes from its predecessor's current exit state until no input changes ----
    for (int j = tid; j < nsub; j += HUFF_T) used_in[j] = UNUSED;
    __syncthreads();
    for (int round = 0; round <= nsub; round++) {
        if (tid == 0) S.changed = 0;
        __syncthreads();
        bool any = false;
        for (int j = tid; j < nsub; j += HUFF_T) {
            const JSeg g = sg[seg_of(j)];
            const int k = j - g.sub_first;
            DecState st;
            if (k == 0) { st.p = 0; st.c = 0; st.z = 0; }
            else if (round == 0) { st.p = k * SUB_BITS; st.c = 0; st.z = 0; }                   // a guess: the stream re-synchronises
            else st = unpack_state(*reinterpret_cast<volatile unsigned long long*>(&exit_st[j - 1]));
            const unsigned long long in = pack_state(st);
            if (in == used_in[j]) continue;
            used_in[j] = in;
            const int end_bit = min((k + 1) * SUB_BITS, g.byte_len * 8);
            int dcs[3] = {0, 0, 0};
            const int n = decode_span<false>(S, img_scan + g.byte_off, st, end_bit, bpm, nullptr, 0, 0, dcs);
            adv[j] = make_int4(n, dcs[0], dcs[1], dcs[2]);      // slots advanced, DC differences summed per component
            *reinterpret_cast<volatile unsigned long long*>(&exit_st[j]) = pack_state(st);
            any = true;
        }
        if (any) S.changed = 1;
        __syncthreads();
        if (!S.changed) break;
        __syncthreads();
    }

    // ---- exclusive scan of (slots advanced, DC sums) over the image's subsequences (segments are told apart afterwards) ----
    __shared__ int4 wtot[32];
    __shared__ int4 carry4;
    if (tid == 0) carry4 = make_int4(0, 0, 0, 0);
    __syncthreads();
    for (int base = 0; base < nsub; base += HUFF_T) {
        const int j = base + tid;
        const int4 v = j < nsub ? adv[j] : make_int4(0, 0, 0, 0);
        int4 inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int tx = __shfl_up_sync(0xffffffffu, inc.x, d), ty = __shfl_up_sync(0xffffffffu, inc.y, d);
            const int tz = __shfl_up_sync(0xffffffffu, inc.z, d), tw = __shfl_up_sync(0xffffffffu, inc.w, d);
            if (lane >= d) { inc.x += tx; inc.y += ty; inc.z += tz; inc.w += tw; }
        }
        if (lane == 31) wtot[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int4 w = wtot[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int tx = __shfl_up_sync(0xffffffffu, w.x, d), ty = __shfl_up_sync(0xffffffffu, w.y, d);
                const int tz = __shfl_up_sync(0xffffffffu, w.z, d), tw = __shfl_up_sync(0xffffffffu, w.w, d);
                if (lane >= d) { w.x += tx; w.y += ty; w.z += tz; w.w += tw; }
            }
            wtot[lane] = w;
        }
        __syncthreads();
        const int4 wb = warp ? wtot[warp - 1] : make_int4(0, 0, 0, 0), cr = carry4;
        const int4 excl = make_int4(cr.x + wb.x + inc.x - v.x, cr.y + wb.y + inc.y - v.y, cr.z + wb.z + inc.z - v.z, cr.w + wb.w + inc.w - v.w);
        if (j < nsub) gpos[j] = excl;
        __syncthreads();
        if (tid == HUFF_T - 1) carry4 = make_int4(excl.x + v.x, excl.y + v.y, excl.z + v.z, excl.w + v.w);
        __syncthreads();
    }

    // ---- write pass: every subsequence again, from its settled start state, at its settled position, with the DC
    //      predictors its predecessors of the same restart interval leave ----
    for (int j = tid; j < nsub; j += HUFF_T) {
        const JSeg g = sg[seg_of(j)];
        const int k = j - g.sub_first;
        DecState st = unpack_state(used_in[j]);
        const long long seg_block0 = (long long)g.mcu_first * bpm;
        const int mcus = H.ri ? min(H.ri, H.mcux * H.mcuy - g.mcu_first) : H.mcux * H.mcuy;
        const int4 here = gpos[j], first = gpos[g.sub_first];
        const long long slot = seg_block0 * 64 + (here.x - first.x);
        int pred[3] = {here.y - first.y, here.z - first.z, here.w - first.w};
        const int end_bit = min((k + 1) * SUB_BITS, g.byte_len * 8);
        decode_span<true>(S, img_scan + g.byte_off, st, end_bit, bpm, coef, slot >> 6, seg_block0 + (long long)mcus * bpm, pred);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// inverse DCT (jidctint.c jpeg_idct_islow) — one thread per block
// ---------------------------------------------------------------------------------------------------------------------
#define JD_DESCALE(x, n) (((x) + (1 << ((n) - 1))) >> (n))
__device__ __forceinline__ void idct8(const int in0, const int in1, const int in2, const int in3, const int in4, const int in5, const int in6, const int in7,
                                      int* o, const int shift) {
    int z1, z2, z3, z4, z5, tmp0, tmp1, tmp2, tmp3, tmp10, tmp11, tmp12, tmp13;
    z2 = in2; z3 = in6;
    z1 = (z2 + z3) * 4433;
    tmp2 = z1 + z3 * (-15137);
    tmp3 = z1 + z2 * 6270;
    z2 = in0; z3 = in4;
    tmp0 = (z2 + z3) * 8192;
    tmp1 = (z2 - z3) * 8192;
    tmp10 = tmp0 + tmp3; tmp13 = tmp0 - tmp3; tmp11 = tmp1 + tmp2; tmp12 = tmp1 - tmp2;
    tmp0 = in7; tmp1 = in5; tmp2 = in3; tmp3 = in1;
    z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2; z4 = tmp1 + tmp3;
    z5 = (z3 + z4) * 9633;
    tmp0 *= 2446; tmp1 *= 16819; tmp2 *= 25172; tmp3 *= 12299;
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
    z3 += z5; z4 += z5;
    tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
    o[0] = JD_DESCALE(tmp10 + tmp3, shift); o[7] = JD_DESCALE(tmp10 - tmp3, shift);
    o[1] = JD_DESCALE(tmp11 + tmp2, shift); o[6] = JD_DESCALE(tmp11 - tmp2, shift);
    o[2] = JD_DESCALE(tmp12 + tmp1, shift); o[5] = JD_DESCALE(tmp12 - tmp1, shift);
    o[3] = JD_DESCALE(tmp13 + tmp0, shift); o[4] = JD_DESCALE(tmp13 - tmp0, shift);
}

__global__ void __launch_bounds__(128) jpeg_idct_kernel(const JpegHdr* __restrict__ hdrs, const short* __restrict__ coef_all, unsigned char* __restrict__ planes) {
    pdl_wait();
    const JpegHdr& H = hdrs[blockIdx.y];
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= H.blocks) return;
    const long long mcu = b / H.bpm;
    const int j = (int)(b - mcu * H.bpm), comp = H.mcu_comp[j];
    const int my = (int)(mcu / H.mcux), mx = (int)(mcu - (long long)my * H.mcux);
    const int bx = mx * H.comp_h[comp] + H.mcu_bx[j], by = my * H.comp_v[comp] + H.mcu_by[j];
    const unsigned short* q = H.qt[H.comp_tq[comp]];
    const int4* src = reinterpret_cast<const int4*>(coef_all + H.coef_off + b * 64);
    int ws[64];
#pragma unroll
    for (int r = 0; r < 8; r++) {            // dequantise a row of coefficients (8 shorts = one 16-byte load)
        const int4 v = src[r];
        const int w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            ws[r * 8 + 2 * k] = (int)(short)(w4[k] & 0xffff) * (int)q[r * 8 + 2 * k];
            ws[r * 8 + 2 * k + 1] = (int)(short)(w4[k] >> 16) * (int)q[r * 8 + 2 * k + 1];
        }
    }
#pragma unroll
    for (int c = 0; c < 8; c++) {            // pass 1: columns
        int o[8];
        idct8(ws[c], ws[8 + c], ws[16 + c], ws[24 + c], ws[32 + c], ws[40 + c], ws[48 + c], ws[56 + c], o, 11);
#pragma unroll
        for (int r = 0; r < 8; r++) ws[r * 8 + c] = o[r];
    }
    const int pitch = H.comp_bw[comp] * 8;
    unsigned char* dst = planes + H.plane_off[comp] + (long long)by * 8 * pitch + bx * 8;
#pragma unroll
    for (int r = 0; r < 8; r++) {            // pass 2: rows, range limit, store 8 bytes
        int o[8];
        idct8(ws[r * 8], ws[r * 8 + 1], ws[r * 8 + 2], ws[r * 8 + 3], ws[r * 8 + 4], ws[r * 8 + 5], ws[r * 8 + 6], ws[r * 8 + 7], o, 18);
        unsigned lo = 0, hi = 0;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            lo |= (unsigned)min(max(o[c] + 128, 0), 255) << (8 * c);
            hi |= (unsigned)min(max(o[c + 4] + 128, 0), 255) << (8 * c);
        }
        *reinterpret_cast<uint2*>(dst + (long long)r * pitch) = make_uint2(lo, hi);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// upsampling (jdsample.c, fancy) + colour conversion (jdcolor.c)
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int chroma_at(const unsigned char* pl, int pitch, int dw, int dh, int hf, int vf, int x, int y) {
    if (hf == 1 && vf == 1) return pl[(long long)y * pitch + x];
    const int cy = vf == 2 ? y >> 1 : y, cx = hf == 2 ? x >> 1 : x;
    const unsigned char* r0 = pl + (long long)cy * pitch;
    if (vf == 2) {
        if (hf == 2 && dw <= 2) return r0[cx];                            // tiny widths: plain replication
        const int yn = (y & 1) ? min(cy + 1, dh - 1) : max(cy - 1, 0);
        const unsigned char* r1 = pl + (long long)yn * pitch;
        if (hf == 1) return (r0[cx] * 3 + r1[cx] + ((y & 1) ? 2 : 1)) >> 2;      // h1v2
        const int cur = r0[cx] * 3 + r1[cx];
        if (x & 1) {
            if (cx == dw - 1) return (cur * 4 + 7) >> 4;
            return (cur * 3 + r0[cx + 1] * 3 + r1[cx + 1] + 7) >> 4;
        }
        if (cx == 0) return (cur * 4 + 8) >> 4;
        return (cur * 3 + r0[cx - 1] * 3 + r1[cx - 1] + 8) >> 4;
    }
    if (dw <= 2) return r0[cx];                                            // h2v1
    if (x & 1) return cx == dw - 1 ? r0[cx] : (r0[cx] * 3 + r0[cx + 1] + 2) >> 2;
    return cx == 0 ? r0[0] : (r0[cx] * 3 + r0[cx - 1] + 1) >> 2;
}

// four pixels of a row per thread: one word of luma in, twelve bytes (three words when aligned) out
__global__ void __launch_bounds__(256) jpeg_colour_kernel(const JpegHdr* __restrict__ hdrs, const unsigned char* __restrict__ planes, int bgr,
                                                          unsigned char* __restrict__ out, long long out_img_stride) {
    pdl_wait();
    const JpegHdr& H = hdrs[blockIdx.z];
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x0 >= H.width || y >= H.height) return;
    const int npx = min(4, H.width - x0);
    const unsigned yw = *reinterpret_cast<const unsigned*>(planes + H.plane_off[0] + (long long)y * (H.comp_bw[0] * 8) + x0);   // planes are padded to whole blocks
    unsigned char* o = out + blockIdx.z * out_img_stride;
    if (H.ncomp == 1) {
        unsigned char* px = o + (long long)y * H.width + x0;
        for (int i = 0; i < npx; i++) px[i] = (unsigned char)(yw >> (8 * i));
        return;
    }
    const int hf[2] = {H.up_hf[1], H.up_hf[2]}, vf[2] = {H.up_vf[1], H.up_vf[2]}, dw[2] = {H.up_dw[1], H.up_dw[2]}, dh[2] = {H.up_dh[1], H.up_dh[2]};
    // 4:2:0 away from the left / right edge (nearly every thread of a camera frame): the four pixels share two chroma columns
    // and their two neighbours — eight bytes per component instead of a look-up per pixel. v(c) = 3 * near row + far row;
    // even x: (3 v(c) + v(c - 1) + 8) >> 4, odd x: (3 v(c) + v(c + 1) + 7) >> 4 (chroma_at, h2v2).
    int cbv[4], crv[4];
    const int c0 = x0 >> 1;
    const bool fast = npx == 4 && hf[0] == 2 && vf[0] == 2 && hf[1] == 2 && vf[1] == 2 && c0 >= 1 && c0 + 2 < min(dw[0], dw[1]);
    if (fast) {
#pragma unroll
        for (int c = 0; c < 2; c++) {
            const unsigned char* pl = planes + H.plane_off[1 + c];
            const int pitch = H.comp_bw[1 + c] * 8, cy = y >> 1, yn = (y & 1) ? min(cy + 1, dh[c] - 1) : max(cy - 1, 0);
            const unsigned char* r0 = pl + (long long)cy * pitch + c0 - 1;
            const unsigned char* r1 = pl + (long long)yn * pitch + c0 - 1;
            int v[4];
#pragma unroll
            for (int k = 0; k < 4; k++) v[k] = r0[k] * 3 + r1[k];
            int* o4 = c == 0 ? cbv : crv;
            o4[0] = (3 * v[1] + v[0] + 8) >> 4; o4[1] = (3 * v[1] + v[2] + 7) >> 4;
            o4[2] = (3 * v[2] + v[1] + 8) >> 4; o4[3] = (3 * v[2] + v[3] + 7) >> 4;
        }
    }
    unsigned char rgb[12];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int x = min(x0 + i, H.width - 1);
        const int Y = (int)((yw >> (8 * i)) & 255u);
        const int cb = (fast ? cbv[i] : chroma_at(planes + H.plane_off[1], H.comp_bw[1] * 8, dw[0], dh[0], hf[0], vf[0], x, y)) - 128;
        const int cr = (fast ? crv[i] : chroma_at(planes + H.plane_off[2], H.comp_bw[2] * 8, dw[1], dh[1], hf[1], vf[1], x, y)) - 128;
        const int r = min(max(Y + ((91881 * cr + 32768) >> 16), 0), 255);
        const int g = min(max(Y + ((-22554 * cb + 32768 - 46802 * cr) >> 16), 0), 255);
        const int b = min(max(Y + ((116130 * cb + 32768) >> 16), 0), 255);
        rgb[3 * i] = (unsigned char)(bgr ? b : r); rgb[3 * i + 1] = (unsigned char)g; rgb[3 * i + 2] = (unsigned char)(bgr ? r : b);
    }
    unsigned char* px = o + ((long long)y * H.width + x0) * 3;
    if (npx == 4 && (reinterpret_cast<uintptr_t>(px) & 3) == 0) {
        unsigned* w = reinterpret_cast<unsigned*>(px);
#pragma unroll
        for (int k = 0; k < 3; k++) w[k] = rgb[4 * k] | (rgb[4 * k + 1] << 8) | (rgb[4 * k + 2] << 16) | ((unsigned)rgb[4 * k + 3] << 24);
    } else {
        for (int i = 0; i < 3 * npx; i++) px[i] = rgb[i];
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// host: marker parse
// ---------------------------------------------------------------------------------------------------------------------
void build_huff(JHuff& t, const unsigned char bits[17], const unsigned char* vals, int nvals) {
    memset(&t, 0, sizeof(t));
    memcpy(t.huffval, vals, nvals);
    int code = 0, k = 0;
    for (int l = 1; l <= 16; l++) {
        t.valptr[l] = k; t.mincode[l] = code;
        for (int i = 0; i < bits[l]; i++, k++, code++)
            if (l <= 9)
                for (int f = 0; f < (1 << (9 - l)); f++) t.lut[(code << (9 - l)) + f] = (unsigned short)((l << 8) | vals[k]);
        t.maxcode[l] = bits[l] ? code - 1 : -1;
        if (!bits[l]) t.mincode[l] = 0x7fffffff;          // no code of this length: never matches
        code <<= 1;
    }
    t.maxcode[17] = 0x7fffffff;
}

// Parses one stream. Fills `h` (everything but the batch offsets), appends the unstuffed entropy-coded bytes to `scan`
// (capacity scan_cap) and the restart intervals to `segs` (capacity seg_cap). Returns 0 or an error message.
const char* parse_jpeg(const unsigned char* d, size_t len, JpegHdr& h, unsigned char* scan, size_t scan_cap, size_t& scan_len, JSeg* segs, int seg_cap, int& nseg) {
    memset(&h, 0, sizeof(h));
    if (len < 4 || d[0] != 0xFF || d[1] != 0xD8) return "not a JPEG stream (no SOI)";
    size_t i = 2;
    bool have_sof = false, have_dc[4] = {false, false, false, false}, have_ac[4] = {false, false, false, false};
    int comp_id[3] = {0, 0, 0}, adobe = -1;
    for (;;) {
        if (i + 4 > len || d[i] != 0xFF) return "truncated or corrupt marker segment";
        const int m = d[i + 1];
        if (m == 0xFF) { i++; continue; }
        const size_t L = ((size_t)d[i + 2] << 8) | d[i + 3];
        if (L < 2 || i + 2 + L > len) return "marker segment runs past the end";
        const unsigned char* s = d + i + 4;
        if (m == 0xDB) {
            size_t k = 0;
            while (k < L - 2) {
                const int pq = s[k] >> 4, tq = s[k] & 15;
                if (tq > 3 || pq > 1 || k + 1 + (pq ? 128 : 64) > L - 2) return "bad quantisation table";
                k++;
                for (int z = 0; z < 64; z++) { h.qt[tq][h_zigzag[z]] = (unsigned short)(pq ? ((s[k] << 8) | s[k + 1]) : s[k]); k += pq ? 2 : 1; }
            }
        } else if (m == 0xC4) {
            size_t k = 0;
            while (k + 17 <= L - 2) {
                const int tc = s[k] >> 4, th = s[k] & 15;
                unsigned char bits[17];
                int n = 0;
                bits[0] = 0;
                for (int l = 1; l <= 16; l++) { bits[l] = s[k + l]; n += bits[l]; }
                if (tc > 1 || th > 3 || n > 256 || k + 17 + n > L - 2) return "bad Huffman table";
                build_huff(tc ? h.ac[th] : h.dc[th], bits, s + k + 17, n);
                (tc ? have_ac : have_dc)[th] = true;
                k += 17 + n;
            }
        } else if (m == 0xC0 || m == 0xC1) {
            if (L < 8 || s[0] != 8) return "only 8-bit samples are supported";
            h.height = (s[1] << 8) | s[2]; h.width = (s[3] << 8) | s[4]; h.ncomp = s[5];
            if (h.ncomp != 1 && h.ncomp != 3) return "only 1- and 3-component streams are supported";
            if (L < 8 + 3 * (size_t)h.ncomp || h.width < 1 || h.height < 1) return "bad frame header";
            for (int c = 0; c < h.ncomp; c++) { comp_id[c] = s[6 + 3 * c]; h.comp_h[c] = s[7 + 3 * c] >> 4; h.comp_v[c] = s[7 + 3 * c] & 15; h.comp_tq[c] = s[8 + 3 * c] & 3; }
            have_sof = true;
        } else if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC8 && m != 0xCC)) {
            return "progressive, lossless and arithmetic-coded streams are not supported (baseline Huffman only)";
        } else if (m == 0xDD) {
            if (L < 4) return "bad restart interval";
            h.ri = (s[0] << 8) | s[1];
        } else if (m == 0xEE && L >= 14 && memcmp(s, "Adobe", 5) == 0) {
            adobe = s[11];
        } else if (m == 0xDA) {
            if (!have_sof) return "scan before frame header";
            if (L < 6 + 2 * (size_t)h.ncomp || s[0] != h.ncomp) return "only one interleaved scan with all components is supported";
            for (int k = 0; k < h.ncomp; k++) {
                if (s[1 + 2 * k] != comp_id[k]) return "scan components out of frame order";
                h.comp_td[k] = s[2 + 2 * k] >> 4; h.comp_ta[k] = s[2 + 2 * k] & 15;
                if (h.comp_td[k] > 3 || h.comp_ta[k] > 3 || !have_dc[h.comp_td[k]] || !have_ac[h.comp_ta[k]]) return "scan refers to a missing Huffman table";
            }
            i += 2 + L;
            break;
        }
        i += 2 + L;
    }
    if (h.ncomp == 3 && adobe == 0) return "RGB-coded (Adobe transform 0) streams are not supported";
    if (h.ncomp == 1) { h.comp_h[0] = h.comp_v[0] = 1; }
    h.hmax = h.vmax = 1;
    for (int c = 0; c < h.ncomp; c++) { h.hmax = h.comp_h[c] > h.hmax ? h.comp_h[c] : h.hmax; h.vmax = h.comp_v[c] > h.vmax ? h.comp_v[c] : h.vmax; }
    for (int c = 0; c < h.ncomp; c++)
        if (h.comp_h[c] < 1 || h.comp_v[c] < 1 || h.comp_h[c] > 2 || h.comp_v[c] > 2 || (c == 0 && (h.comp_h[c] != h.hmax || h.comp_v[c] != h.vmax)) ||
            (c > 0 && (h.comp_h[c] != 1 || h.comp_v[c] != 1)))
            return "unsupported sampling factors (4:4:4, 4:2:2, 4:4:0, 4:2:0 and grey are)";
    h.mcux = (h.width + 8 * h.hmax - 1) / (8 * h.hmax); h.mcuy = (h.height + 8 * h.vmax - 1) / (8 * h.vmax);
    h.bpm = 0;
    for (int c = 0; c < h.ncomp; c++) {
        h.comp_first_j[c] = h.bpm;
        for (int v = 0; v < h.comp_v[c]; v++)
            for (int hh = 0; hh < h.comp_h[c]; hh++) { h.mcu_comp[h.bpm] = c; h.mcu_bx[h.bpm] = hh; h.mcu_by[h.bpm] = v; h.bpm++; }
        h.comp_bw[c] = h.mcux * h.comp_h[c]; h.comp_bh[c] = h.mcuy * h.comp_v[c];
        h.up_hf[c] = h.hmax / h.comp_h[c]; h.up_vf[c] = h.vmax / h.comp_v[c];
        h.up_dw[c] = (h.width * h.comp_h[c] + h.hmax - 1) / h.hmax; h.up_dh[c] = (h.height * h.comp_v[c] + h.vmax - 1) / h.vmax;
    }
    h.blocks = (long long)h.mcux * h.mcuy * h.bpm;
    // entropy-coded data: unstuff, cut at RSTn, stop at any other marker (EOI)
    const long long total_mcus = (long long)h.mcux * h.mcuy;
    nseg = 0;
    scan_len = 0;
    size_t seg_start = 0;
    int mcu_first = 0;
    auto close_seg = [&]() -> bool {
        if (nseg >= seg_cap) return false;
        segs[nseg].byte_off = (int)seg_start; segs[nseg].byte_len = (int)(scan_len - seg_start); segs[nseg].mcu_first = mcu_first; segs[nseg].sub_first = 0;
        nseg++;
        // segments start on a 4-byte boundary of the scan buffer (aligned window loads); pad with ones
        while (scan_len & 3) { if (scan_len >= scan_cap) return false; scan[scan_len++] = 0xFF; }
        seg_start = scan_len;
        return true;
    };
    while (i < len) {
        const unsigned char* q = (const unsigned char*)memchr(d + i, 0xFF, len - i);
        const size_t run = q ? (size_t)(q - (d + i)) : len - i;
        if (scan_len + run + 8 > scan_cap) return "staging buffer too small";
        memcpy(scan + scan_len, d + i, run);
        scan_len += run; i += run;
        if (!q) break;
        if (i + 1 >= len) break;
        const int nx = d[i + 1];
        if (nx == 0x00) { scan[scan_len++] = 0xFF; i += 2; }
        else if (nx >= 0xD0 && nx <= 0xD7) {
            if (!h.ri) return "restart marker without a restart interval";
            if (!close_seg()) return "too many restart intervals for the staging buffer";
            mcu_first += h.ri;
            i += 2;
            if (mcu_first >= total_mcus) break;
        } else if (nx == 0xFF) { i++; }                   // fill byte
        else break;                                        // EOI or another marker ends the scan
    }
    if (!close_seg()) return "too many restart intervals for the staging buffer";
    if (h.ri ? ((long long)(nseg - 1) * h.ri >= total_mcus || (long long)nseg * h.ri < total_mcus) : nseg != 1) return "entropy-coded data does not cover the frame";
    if (scan_len > (1u << 28)) return "entropy-coded segment larger than 256 MB";
    h.scan_len = (int)scan_len;
    h.nseg = nseg;
    return nullptr;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

static int init_jpeg(tcb_ctx* ctx) {   // run by tcb_create
    TCB_SMEM_OPT_IN(ctx, jpeg_huffman_kernel, sizeof(HuffShared));
    return TCB_OK;
}
static TcbInitReg reg_jpeg(&init_jpeg);

extern "C" int tcb_jpeg_info(const unsigned char* jpeg_host, size_t len, int* height, int* width, int* channels) {
    if (!jpeg_host || !height || !width || !channels) return tcb_fail(nullptr, TCB_ERR_ARG, "tcb_jpeg_info: null argument");
    size_t i = 2;
    if (len < 4 || jpeg_host[0] != 0xFF || jpeg_host[1] != 0xD8) return tcb_fail(nullptr, TCB_ERR_ARG, "tcb_jpeg_info: not a JPEG stream");
    while (i + 4 <= len) {
        if (jpeg_host[i] != 0xFF) break;
        const int m = jpeg_host[i + 1];
        if (m == 0xFF) { i++; continue; }
        const size_t L = ((size_t)jpeg_host[i + 2] << 8) | jpeg_host[i + 3];
        if ((m == 0xC0 || m == 0xC1) && i + 10 <= len) {
            *height = (jpeg_host[i + 5] << 8) | jpeg_host[i + 6]; *width = (jpeg_host[i + 7] << 8) | jpeg_host[i + 8]; *channels = jpeg_host[i + 9];
            return TCB_OK;
        }
        if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC))
            return tcb_fail(nullptr, TCB_ERR_ARG, "tcb_jpeg_info: progressive, lossless and arithmetic-coded streams are not supported (baseline Huffman only)");
        if (m == 0xDA) break;
        i += 2 + L;
    }
    return tcb_fail(nullptr, TCB_ERR_ARG, "tcb_jpeg_info: no baseline frame header found");
}

extern "C" size_t tcb_jpeg_staging_bytes(size_t total_jpeg_bytes, int n) {
    const size_t nn = (size_t)(n > 0 ? n : 0);
    return align_up(nn * sizeof(JpegHdr), 256) + align_up(total_jpeg_bytes + total_jpeg_bytes / 8 + nn * 80 + 256, 256) + align_up((total_jpeg_bytes / 16 + nn * 16 + 64) * sizeof(JSeg), 256) + 1024;
}

extern "C" size_t tcb_jpeg_workspace(size_t total_jpeg_bytes, int n, int height, int width) {
    const size_t nn = (size_t)(n > 0 ? n : 0);
    const size_t mcu_px = ((size_t)(width > 0 ? width : 0) + 15) / 16 * 16 * (((size_t)(height > 0 ? height : 0) + 15) / 16 * 16);
    const size_t blocks = mcu_px / 64 * 3;                    // 4:4:4 bound
    const size_t subs = total_jpeg_bytes / SUB_BYTES + total_jpeg_bytes / 16 + nn * 16 + 64;
    return tcb_jpeg_staging_bytes(total_jpeg_bytes, n) + subs * (8 + 8 + 16 + 16) + nn * blocks * (128 + 64) + 16 * 256 + 4096;
}

static int tcb_jpeg_decode_(tcb_ctx* ctx, const unsigned char* const* jpeg_host, const size_t* jpeg_len, int n, void* staging_host,
                            size_t staging_bytes, int bgr, unsigned char* out_dev, int out_h, int out_w, int out_channels, void* stream);

extern "C" int tcb_jpeg_decode(tcb_ctx* ctx, const unsigned char* const* jpeg_host, const size_t* jpeg_len, int n, void* staging_host,
                               size_t staging_bytes, int bgr, unsigned char* out_dev, int out_h, int out_w, int out_channels, void* stream) {
    try {       // the host parse uses std::vector / std::thread: nothing may leave a C entry point as an exception
        return tcb_jpeg_decode_(ctx, jpeg_host, jpeg_len, n, staging_host, staging_bytes, bgr, out_dev, out_h, out_w, out_channels, stream);
    } catch (const std::bad_alloc&) {
        return tcb_fail(ctx, TCB_ERR_ALLOC, "tcb_jpeg_decode: host allocation failed");
    } catch (...) {
        return tcb_fail(ctx, TCB_ERR_ARG, "tcb_jpeg_decode: unexpected host error");
    }
}

static int tcb_jpeg_decode_(tcb_ctx* ctx, const unsigned char* const* jpeg_host, const size_t* jpeg_len, int n, void* staging_host,
                            size_t staging_bytes, int bgr, unsigned char* out_dev, int out_h, int out_w, int out_channels, void* stream) {
    TCB_CHECK_CTX(ctx);
    TCB_REQUIRE(ctx, n >= 0 && (n == 0 || (jpeg_host && jpeg_len && staging_host && out_dev)), "tcb_jpeg_decode: null argument");
    TCB_REQUIRE(ctx, out_h >= 1 && out_w >= 1 && (out_channels == 1 || out_channels == 3), "tcb_jpeg_decode: bad output shape");
    if (n == 0) return TCB_OK;
    if (n > 65535) return tcb_fail(ctx, TCB_ERR_LIMIT, "tcb_jpeg_decode: more than 65535 images per call");
    cudaStream_t st = (cudaStream_t)stream;
    size_t total = 0;
    for (int k = 0; k < n; k++) { TCB_REQUIRE(ctx, jpeg_host[k] != nullptr, "tcb_jpeg_decode: stream %d is null", k); total += jpeg_len[k]; }
    if (staging_bytes < tcb_jpeg_staging_bytes(total, n))
        return tcb_fail(ctx, TCB_ERR_WORKSPACE, "tcb_jpeg_decode: staging buffer too small (need %zu B, have %zu B)", tcb_jpeg_staging_bytes(total, n), staging_bytes);
    // staging layout: headers | scan bytes | segments
    char* sbase = (char*)staging_host;
    JpegHdr* hdrs = (JpegHdr*)sbase;
    const size_t scan_at = align_up((size_t)n * sizeof(JpegHdr), 256), scan_cap = align_up(total + total / 8 + (size_t)n * 80 + 256, 256);
    unsigned char* scan = (unsigned char*)(sbase + scan_at);
    const size_t seg_at = scan_at + scan_cap;
    JSeg* segs = (JSeg*)(sbase + seg_at);
    const int seg_cap = (int)((staging_bytes - seg_at) / sizeof(JSeg));
    // Every stream is parsed into its own stretch of the scan area (its raw length bounds the unstuffed bytes), so the
    // streams of a batch are parsed by a few host threads side by side; the segment lists are joined afterwards.
    std::vector<size_t> scan_off(n + 1, 0);
    for (int k = 0; k < n; k++) scan_off[k + 1] = align_up(scan_off[k] + jpeg_len[k] + jpeg_len[k] / 8 + 48, 16);   // unstuffing shrinks a stream, the 4-byte alignment of
                                                                                                      // restart segments can grow it a little; + readable slack behind it
    if (scan_off[n] > scan_cap) return tcb_fail(ctx, TCB_ERR_WORKSPACE, "tcb_jpeg_decode: staging buffer too small for the scan bytes");
    std::vector<std::vector<JSeg>> seg_of(n);
    std::vector<const char*> err_of(n, nullptr);
    auto parse_range = [&](int k0, int k1) {
        for (int k = k0; k < k1; k++) {
            size_t sl = 0;
            int ns = 0;
            std::vector<JSeg>& sv = seg_of[k];
            sv.resize(64);
            const char* err = parse_jpeg(jpeg_host[k], jpeg_len[k], hdrs[k], scan + scan_off[k], scan_off[k + 1] - scan_off[k], sl, sv.data(), (int)sv.size(), ns);
            if (err && !strcmp(err, "too many restart intervals for the staging buffer")) {   // a stream with many restart intervals: room for one per MCU row pair at least
                sv.resize(jpeg_len[k] / 2 + 64);
                err = parse_jpeg(jpeg_host[k], jpeg_len[k], hdrs[k], scan + scan_off[k], scan_off[k + 1] - scan_off[k], sl, sv.data(), (int)sv.size(), ns);
            }
            err_of[k] = err;
            sv.resize(err ? 0 : ns);
        }
    };
    {
        int nthreads = (int)(total >> 24);                    // one thread per 16 MB of streams: a 64-frame batch of 720p frames is parsed in ~3 ms by
                                                              // one thread, less than its kernels take, and extra threads only disturbed the launching thread
        const int hw = (int)std::thread::hardware_concurrency();
        nthreads = nthreads > 8 ? 8 : nthreads;
        if (hw > 0 && nthreads > hw) nthreads = hw;
        { const char* e = getenv("TCB_JPEG_THREADS"); if (e && atoi(e) > 0) nthreads = atoi(e); }   // override (measurement)
        if (nthreads > n) nthreads = n;
        if (nthreads <= 1) parse_range(0, n);
        else {
            std::vector<std::thread> pool;
            int started = 1;                                  // ranges [n*t/nthreads, n*(t+1)/nthreads); the calling thread takes what no thread got
            try {
                for (; started < nthreads; started++)
                    pool.emplace_back(parse_range, (int)((long long)n * started / nthreads), (int)((long long)n * (started + 1) / nthreads));
            } catch (...) {}                                  // no more threads to be had: nothing may escape a C entry point
            parse_range(0, n / nthreads);
            if (started < nthreads) parse_range((int)((long long)n * started / nthreads), n);
            for (auto& th : pool) th.join();
        }
    }
    int seg_used = 0, sub_used = 0;
    long long coef_used = 0, plane_used = 0;
    for (int k = 0; k < n; k++) {
        if (err_of[k]) return tcb_fail(ctx, TCB_ERR_ARG, "tcb_jpeg_decode: stream %d: %s", k, err_of[k]);
        JpegHdr& h = hdrs[k];
        if (h.width != out_w || h.height != out_h || h.ncomp != out_channels)
            return tcb_fail(ctx, TCB_ERR_ARG, "tcb_jpeg_decode: stream %d is %dx%dx%d, the output is %dx%dx%d", k, h.height, h.width, h.ncomp, out_h, out_w, out_channels);
        const int ns = (int)seg_of[k].size();
        if (seg_used + ns > seg_cap) return tcb_fail(ctx, TCB_ERR_WORKSPACE, "tcb_jpeg_decode: staging buffer too small for %d restart intervals", seg_used + ns);
        memcpy(segs + seg_used, seg_of[k].data(), sizeof(JSeg) * (size_t)ns);
        h.scan_off = (long long)scan_off[k]; h.seg_first = seg_used; h.sub_first = sub_used;
        h.nsub = 0;
        for (int s2 = 0; s2 < ns; s2++) {
            segs[seg_used + s2].sub_first = h.nsub;
            h.nsub += segs[seg_used + s2].byte_len > 0 ? tcb_div_up(segs[seg_used + s2].byte_len, SUB_BYTES) : 1;
        }
        h.coef_off = coef_used;
        coef_used += h.blocks * 64;
        for (int c = 0; c < h.ncomp; c++) { h.plane_off[c] = plane_used; plane_used += (long long)h.comp_bw[c] * h.comp_bh[c] * 64; }
        seg_used += ns; sub_used += h.nsub;
    }
    const size_t copy_bytes = seg_at + (size_t)seg_used * sizeof(JSeg);
    TCB_WS(ctx, dev_stage, char, copy_bytes + 256);
    TCB_WS(ctx, exit_st, unsigned long long, sub_used + 1);
    TCB_WS(ctx, used_in, unsigned long long, sub_used + 1);
    TCB_WS(ctx, adv, int4, sub_used + 1);
    TCB_WS(ctx, gpos, int4, sub_used + 1);
    TCB_WS(ctx, coef, short, coef_used + 64);
    TCB_WS(ctx, planes, unsigned char, plane_used + 64);
    TCB_CUDA(ctx, cudaMemcpyAsync(dev_stage, sbase, copy_bytes, cudaMemcpyHostToDevice, st));
    TCB_CUDA(ctx, cudaMemsetAsync(coef, 0, (size_t)coef_used * sizeof(short), st));
    const JpegHdr* d_hdrs = (const JpegHdr*)dev_stage;
    const unsigned char* d_scan = (const unsigned char*)(dev_stage + scan_at);
    const JSeg* d_segs = (const JSeg*)(dev_stage + seg_at);
    TCB_KLAUNCH(ctx, jpeg_huffman_kernel, n, HUFF_T, sizeof(HuffShared), st, d_hdrs, d_scan, d_segs, exit_st, used_in, adv, gpos, coef);
    TCB_LAUNCHED(ctx);
    long long max_blocks = 0;
    for (int k = 0; k < n; k++) max_blocks = hdrs[k].blocks > max_blocks ? hdrs[k].blocks : max_blocks;
    TCB_KLAUNCH(ctx, jpeg_idct_kernel, dim3(tcb_div_up(max_blocks, 128), n), 128, 0, st, d_hdrs, coef, planes);
    TCB_LAUNCHED(ctx);
    TCB_KLAUNCH(ctx, jpeg_colour_kernel, dim3(tcb_div_up(out_w, 1024), out_h, n), 256, 0, st, d_hdrs, planes, bgr ? 1 : 0, out_dev,
                (long long)out_h * out_w * out_channels);
    TCB_LAUNCHED(ctx);
    return TCB_OK;
}
