/* Oracle (test infrastructure, not product code): baseline JPEG decode as cv2.imdecode performs it — what the
 * reference calls at communicator/ros_inference.py:123-131 for sensor_msgs/CompressedImage frames. OpenCV hands the
 * bytes to libjpeg-turbo (not part of the reference tree; OpenCV 4.13 bundles 3.1.2) with default settings:
 * Huffman entropy decoding, dequantisation + the "islow" integer inverse DCT (jidctint.c: 13-bit constants, two
 * passes, PASS1_BITS 2), "fancy" triangle-filter chroma upsampling (jdsample.c h2v1 / h2v2 fancy), and the
 * fixed-point YCbCr -> RGB conversion of jdcolor.c (16-bit scaled constants), output in BGR order.
 * Restated from the published algorithms (ITU T.81 and the files named above); pinned by tests/test_oracle_jpeg.py,
 * which compares it bit for bit with cv2.imdecode on frames encoded by cv2.imencode.
 * Scope: 8-bit baseline / extended-sequential Huffman, one interleaved scan, 1 or 3 components, sampling factors
 * h, v in {1, 2} with full-resolution luma (4:4:4, 4:2:2, 4:4:0, 4:2:0), restart intervals.
 * Serial and simple on purpose. Returns 0, or a negative code for streams outside that scope. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int maxcode[18], valptr[17], mincode[17];
    uint8_t huffval[256];
    int present;
} HuffTab;

typedef struct {
    const uint8_t* p; const uint8_t* end;
    uint32_t acc; int nbits;
    int hit_marker;
} BitReader;

static int br_bit(BitReader* b) {
    if (b->nbits == 0) {
        int c = 0xFF;                                   /* past the data: feed ones (harmless padding) */
        if (!b->hit_marker && b->p < b->end) {
            c = *b->p++;
            if (c == 0xFF) {
                int d = b->p < b->end ? *b->p : 0xD9;
                if (d == 0x00) b->p++;                  /* stuffed zero */
                else { b->hit_marker = 1; b->p--; c = 0xFF; }
            }
        }
        b->acc = (uint32_t)c; b->nbits = 8;
    }
    b->nbits--;
    return (b->acc >> b->nbits) & 1;
}
static int br_bits(BitReader* b, int n) { int v = 0; while (n--) v = (v << 1) | br_bit(b); return v; }

static void build_huff(HuffTab* t, const uint8_t bits[17], const uint8_t* vals, int nvals) {
    int code = 0, k = 0;
    memcpy(t->huffval, vals, nvals);
    for (int l = 1; l <= 16; l++) {
        t->valptr[l] = k; t->mincode[l] = code;
        k += bits[l]; code += bits[l];
        t->maxcode[l] = bits[l] ? code - 1 : -1;
        code <<= 1;
    }
    t->maxcode[17] = 0x7fffffff;
    t->present = 1;
}
static int huff_decode(BitReader* b, const HuffTab* t) {
    int code = 0;
    for (int l = 1; l <= 16; l++) {
        code = (code << 1) | br_bit(b);
        if (t->maxcode[l] >= 0 && code <= t->maxcode[l] && code >= t->mincode[l]) return t->huffval[t->valptr[l] + code - t->mincode[l]];
    }
    return 0;   /* invalid code: libjpeg warns and returns 0 */
}
static int extend(int v, int s) { return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v; }

static const uint8_t ZIGZAG[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                                   35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

/* jidctint.c jpeg_idct_islow */
#define CONST_BITS 13
#define PASS1_BITS 2
#define DESCALE(x, n) (((x) + (1 << ((n) - 1))) >> (n))
static uint8_t clamp_u8(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

static void idct_pass(const int* in, int stride, int* out, int ostride, int shift, int lshift_dc) {
    /* one 8-point pass over in[0], in[stride], ... -> out[0], out[ostride], ... descaled by `shift` */
    int z1, z2, z3, z4, z5, tmp0, tmp1, tmp2, tmp3, tmp10, tmp11, tmp12, tmp13;
    (void)lshift_dc;
    z2 = in[2 * stride]; z3 = in[6 * stride];
    z1 = (z2 + z3) * 4433;
    tmp2 = z1 + z3 * (-15137);
    tmp3 = z1 + z2 * 6270;
    z2 = in[0]; z3 = in[4 * stride];
    tmp0 = (z2 + z3) * (1 << CONST_BITS);
    tmp1 = (z2 - z3) * (1 << CONST_BITS);
    tmp10 = tmp0 + tmp3; tmp13 = tmp0 - tmp3; tmp11 = tmp1 + tmp2; tmp12 = tmp1 - tmp2;
    tmp0 = in[7 * stride]; tmp1 = in[5 * stride]; tmp2 = in[3 * stride]; tmp3 = in[1 * stride];
    z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2; z4 = tmp1 + tmp3;
    z5 = (z3 + z4) * 9633;
    tmp0 *= 2446; tmp1 *= 16819; tmp2 *= 25172; tmp3 *= 12299;
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
    z3 += z5; z4 += z5;
    tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
    out[0 * ostride] = DESCALE(tmp10 + tmp3, shift);
    out[7 * ostride] = DESCALE(tmp10 - tmp3, shift);
    out[1 * ostride] = DESCALE(tmp11 + tmp2, shift);
    out[6 * ostride] = DESCALE(tmp11 - tmp2, shift);
    out[2 * ostride] = DESCALE(tmp12 + tmp1, shift);
    out[5 * ostride] = DESCALE(tmp12 - tmp1, shift);
    out[3 * ostride] = DESCALE(tmp13 + tmp0, shift);
    out[4 * ostride] = DESCALE(tmp13 - tmp0, shift);
}
static void idct_islow(const int16_t* coef, const uint16_t* q, uint8_t* dst, int dst_stride) {
    int in[64], ws[64], row[8];
    for (int i = 0; i < 64; i++) in[i] = (int)coef[i] * (int)q[i];
    for (int c = 0; c < 8; c++) idct_pass(in + c, 8, ws + c, 8, CONST_BITS - PASS1_BITS, 0);     /* columns */
    for (int r = 0; r < 8; r++) {                                                                 /* rows */
        idct_pass(ws + 8 * r, 1, row, 1, CONST_BITS + PASS1_BITS + 3, 0);
        for (int c = 0; c < 8; c++) dst[r * dst_stride + c] = clamp_u8(row[c] + 128);
    }
}

typedef struct { int id, h, v, tq, td, ta; int bw, bh; /* blocks per row / rows (MCU padded) */ int pw, ph; uint8_t* plane; } Comp;

int oracle_jpeg_info(const uint8_t* data, size_t len, int* width, int* height, int* ncomp) {
    size_t i = 2;
    if (len < 4 || data[0] != 0xFF || data[1] != 0xD8) return -1;
    while (i + 4 <= len) {
        if (data[i] != 0xFF) return -1;
        int m = data[i + 1];
        if (m == 0xFF) { i++; continue; }
        size_t L = ((size_t)data[i + 2] << 8) | data[i + 3];
        if (m == 0xC0 || m == 0xC1) {
            *height = (data[i + 5] << 8) | data[i + 6]; *width = (data[i + 7] << 8) | data[i + 8]; *ncomp = data[i + 9];
            return 0;
        }
        if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)) return -2;
        i += 2 + L;
    }
    return -1;
}

/* out: height x width x 3 (BGR when bgr != 0, else RGB), or height x width for a 1-component stream */
int oracle_jpeg_decode(const uint8_t* data, size_t len, uint8_t* out, int bgr) {
    uint16_t qt[4][64];
    HuffTab dc[4], ac[4];
    Comp comp[3];
    int ncomp = 0, width = 0, height = 0, ri = 0, have_sof = 0, adobe_transform = -1;
    memset(dc, 0, sizeof(dc)); memset(ac, 0, sizeof(ac)); memset(qt, 0, sizeof(qt));
    size_t i = 2;
    if (len < 4 || data[0] != 0xFF || data[1] != 0xD8) return -1;
    for (;;) {
        if (i + 4 > len || data[i] != 0xFF) return -1;
        int m = data[i + 1];
        if (m == 0xFF) { i++; continue; }
        size_t L = ((size_t)data[i + 2] << 8) | data[i + 3];
        const uint8_t* s = data + i + 4;
        if (i + 2 + L > len) return -1;
        if (m == 0xDB) {                                       /* DQT */
            size_t k = 0;
            while (k + 1 <= L - 2) {
                int pq = s[k] >> 4, tq = s[k] & 15;
                if (tq > 3) return -1;
                k++;
                for (int z = 0; z < 64; z++) {
                    int v = pq ? ((s[k] << 8) | s[k + 1]) : s[k];
                    k += pq ? 2 : 1;
                    qt[tq][ZIGZAG[z]] = (uint16_t)v;
                }
            }
        } else if (m == 0xC4) {                                /* DHT */
            size_t k = 0;
            while (k + 17 <= L - 2) {
                int tc = s[k] >> 4, th = s[k] & 15;
                uint8_t bits[17];
                int n = 0;
                bits[0] = 0;
                for (int l = 1; l <= 16; l++) { bits[l] = s[k + l]; n += bits[l]; }
                if (th > 3 || n > 256) return -1;
                build_huff(tc ? &ac[th] : &dc[th], bits, s + k + 17, n);
                k += 17 + n;
            }
        } else if (m == 0xC0 || m == 0xC1) {                   /* SOF0 / SOF1 */
            if (s[0] != 8) return -2;
            height = (s[1] << 8) | s[2]; width = (s[3] << 8) | s[4]; ncomp = s[5];
            if (ncomp != 1 && ncomp != 3) return -2;
            for (int c = 0; c < ncomp; c++) { comp[c].id = s[6 + 3 * c]; comp[c].h = s[7 + 3 * c] >> 4; comp[c].v = s[7 + 3 * c] & 15; comp[c].tq = s[8 + 3 * c]; }
            have_sof = 1;
        } else if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC8 && m != 0xCC)) {
            return -2;                                         /* progressive / lossless / arithmetic */
        } else if (m == 0xDD) {
            ri = (s[0] << 8) | s[1];
        } else if (m == 0xEE && L >= 14 && memcmp(s, "Adobe", 5) == 0) {
            adobe_transform = s[11];
        } else if (m == 0xDA) {                                /* SOS */
            if (!have_sof || s[0] != ncomp) return -2;
            for (int k = 0; k < ncomp; k++) {
                int cid = s[1 + 2 * k], found = 0;
                for (int c = 0; c < ncomp; c++) if (comp[c].id == cid) { comp[c].td = s[2 + 2 * k] >> 4; comp[c].ta = s[2 + 2 * k] & 15; found = (c == k); }
                if (!found) return -2;                         /* components must come in frame order */
            }
            i += 2 + L;
            break;
        }
        i += 2 + L;
    }
    if (ncomp == 3 && adobe_transform == 0) return -2;          /* RGB-coded JPEG */
    int hmax = 1, vmax = 1;
    for (int c = 0; c < ncomp; c++) { if (comp[c].h > hmax) hmax = comp[c].h; if (comp[c].v > vmax) vmax = comp[c].v; }
    if (ncomp == 1) { comp[0].h = comp[0].v = 1; hmax = vmax = 1; }
    for (int c = 0; c < ncomp; c++)
        if (comp[c].h < 1 || comp[c].v < 1 || comp[c].h > 2 || comp[c].v > 2 || (c == 0 && (comp[c].h != hmax || comp[c].v != vmax)) || (c > 0 && (comp[c].h != 1 || comp[c].v != 1)))
            return -2;
    const int mcuw = 8 * hmax, mcuh = 8 * vmax, mcux = (width + mcuw - 1) / mcuw, mcuy = (height + mcuh - 1) / mcuh;
    for (int c = 0; c < ncomp; c++) {
        comp[c].bw = mcux * comp[c].h; comp[c].bh = mcuy * comp[c].v;
        comp[c].pw = comp[c].bw * 8; comp[c].ph = comp[c].bh * 8;
        comp[c].plane = (uint8_t*)malloc((size_t)comp[c].pw * comp[c].ph);
        if (!dc[comp[c].td].present || !ac[comp[c].ta].present) return -1;
    }
    /* entropy-coded segment */
    BitReader br = {data + i, data + len, 0, 0, 0};
    int pred[3] = {0, 0, 0}, mcus_left = ri;
    for (int my = 0; my < mcuy; my++) {
        for (int mx = 0; mx < mcux; mx++) {
            if (ri && mcus_left == 0) {                        /* restart marker: byte-align, skip FF Dn, reset predictors */
                br.nbits = 0; br.hit_marker = 0;
                while (br.p + 1 < br.end && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)) br.p++;
                br.p += 2;
                pred[0] = pred[1] = pred[2] = 0;
                mcus_left = ri;
            }
            mcus_left--;
            for (int c = 0; c < ncomp; c++) {
                for (int v = 0; v < comp[c].v; v++) {
                    for (int h = 0; h < comp[c].h; h++) {
                        int16_t blk[64];
                        memset(blk, 0, sizeof(blk));
                        int s = huff_decode(&br, &dc[comp[c].td]);
                        int diff = s ? extend(br_bits(&br, s), s) : 0;
                        pred[c] += diff;
                        blk[0] = (int16_t)pred[c];
                        for (int k = 1; k < 64;) {
                            int rs = huff_decode(&br, &ac[comp[c].ta]);
                            int r = rs >> 4, sz = rs & 15;
                            if (sz == 0) { if (r == 15) { k += 16; continue; } break; }
                            k += r;
                            if (k > 63) break;
                            blk[ZIGZAG[k]] = (int16_t)extend(br_bits(&br, sz), sz);
                            k++;
                        }
                        const int bx = mx * comp[c].h + h, by = my * comp[c].v + v;
                        idct_islow(blk, qt[comp[c].tq], comp[c].plane + (size_t)by * 8 * comp[c].pw + bx * 8, comp[c].pw);
                    }
                }
            }
        }
    }
    if (ncomp == 1) {
        for (int y = 0; y < height; y++) memcpy(out + (size_t)y * width, comp[0].plane + (size_t)y * comp[0].pw, width);
        free(comp[0].plane);
        return 0;
    }
    /* upsampling (jdsample.c, fancy) to full resolution rows, then jdcolor.c ycc_rgb_convert */
    uint8_t* up[3];
    up[0] = comp[0].plane;
    for (int c = 1; c < 3; c++) {
        const int hf = hmax / comp[c].h, vf = vmax / comp[c].v;
        const int dw = (width * comp[c].h + hmax - 1) / hmax, dh = (height * comp[c].v + vmax - 1) / vmax;   /* downsampled_width / height */
        const uint8_t* src = comp[c].plane;
        const int sp = comp[c].pw;
        if (hf == 1 && vf == 1) { up[c] = comp[c].plane; continue; }
        uint8_t* dst = (uint8_t*)malloc((size_t)(dw * hf + 2) * (dh * vf + 2));
        const int dp = dw * hf;
        for (int y = 0; y < dh; y++) {
            for (int v = 0; v < vf; v++) {
                uint8_t* o = dst + (size_t)(y * vf + v) * dp;
                const uint8_t* r0 = src + (size_t)y * sp;
                if (vf == 2) {
                    int yn = v == 0 ? y - 1 : y + 1;
                    if (yn < 0) yn = 0;
                    if (yn > dh - 1) yn = dh - 1;
                    const uint8_t* r1 = src + (size_t)yn * sp;
                    if (hf == 2) {                                            /* h2v2_fancy_upsample */
                        if (dw <= 2) { for (int x = 0; x < dw; x++) { o[2 * x] = o[2 * x + 1] = r0[x]; } continue; }   /* (plain replication for tiny widths) */
                        int this_ = r0[0] * 3 + r1[0], next = r0[1] * 3 + r1[1], last;
                        o[0] = (uint8_t)((this_ * 4 + 8) >> 4);
                        o[1] = (uint8_t)((this_ * 3 + next + 7) >> 4);
                        last = this_; this_ = next;
                        for (int x = 1; x < dw - 1; x++) {
                            next = r0[x + 1] * 3 + r1[x + 1];
                            o[2 * x] = (uint8_t)((this_ * 3 + last + 8) >> 4);
                            o[2 * x + 1] = (uint8_t)((this_ * 3 + next + 7) >> 4);
                            last = this_; this_ = next;
                        }
                        o[2 * (dw - 1)] = (uint8_t)((this_ * 3 + last + 8) >> 4);
                        o[2 * (dw - 1) + 1] = (uint8_t)((this_ * 4 + 7) >> 4);
                    } else {                                                  /* h1v2_fancy_upsample (4:4:0) */
                        const int bias = v == 0 ? 1 : 2;
                        for (int x = 0; x < dw; x++) o[x] = (uint8_t)((r0[x] * 3 + r1[x] + bias) >> 2);
                    }
                } else {                                                      /* h2v1_fancy_upsample (4:2:2) */
                    if (dw <= 2) { for (int x = 0; x < dw; x++) { o[2 * x] = o[2 * x + 1] = r0[x]; } continue; }
                    o[0] = r0[0];
                    o[1] = (uint8_t)((r0[0] * 3 + r0[1] + 2) >> 2);
                    for (int x = 1; x < dw - 1; x++) {
                        o[2 * x] = (uint8_t)((r0[x] * 3 + r0[x - 1] + 1) >> 2);
                        o[2 * x + 1] = (uint8_t)((r0[x] * 3 + r0[x + 1] + 2) >> 2);
                    }
                    o[2 * (dw - 1)] = (uint8_t)((r0[dw - 1] * 3 + r0[dw - 2] + 1) >> 2);
                    o[2 * (dw - 1) + 1] = r0[dw - 1];
                }
            }
        }
        up[c] = dst;
        comp[c].pw = dp;   /* row pitch of the upsampled plane */
    }
    for (int y = 0; y < height; y++) {
        for (int x = 0; x < width; x++) {
            const int Y = up[0][(size_t)y * comp[0].pw + x], cb = up[1][(size_t)y * comp[1].pw + x] - 128, cr = up[2][(size_t)y * comp[2].pw + x] - 128;
            const int r = Y + ((91881 * cr + 32768) >> 16);
            const int g = Y + ((-22554 * cb + 32768 - 46802 * cr) >> 16);
            const int b = Y + ((116130 * cb + 32768) >> 16);
            uint8_t* o = out + ((size_t)y * width + x) * 3;
            o[bgr ? 2 : 0] = clamp_u8(r); o[1] = clamp_u8(g); o[bgr ? 0 : 2] = clamp_u8(b);
        }
    }
    for (int c = 0; c < 3; c++) { if (up[c] != comp[c].plane) free(up[c]); free(comp[c].plane); }
    return 0;
}
