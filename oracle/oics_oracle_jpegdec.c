/* TEST INFRASTRUCTURE -- CPU oracle, never on the product path.
 *
 * N2 (input side): imgcodecs::imread(path, IMREAD_COLOR) (omr.rs:349, transfer.rs:43,62) of a
 * baseline JPEG = libjpeg(-turbo)'s decompressor with its defaults, restated: jdmarker (DQT / DHT /
 * SOF0 / DRI / SOS), jdhuff (sequential Huffman decode, DC prediction, restart intervals),
 * jidctint (islow IDCT), jdsample (h2v1 / h2v2 fancy upsampling with jdmainct's edge rows),
 * jdcolor (YCC -> RGB tables), output BGR.  The codec is a third-party dependency of the reference
 * (libjpeg-turbo inside OpenCV, not vendored); pinned bit for bit against cv2.imdecode
 * (tests/test_oracle_vs_cv2.py::test_jpeg_decode_equals_cv2).
 *
 * Supported: 8-bit baseline / extended sequential, one interleaved scan, 1 component or Y Cb Cr with
 * luma sampling 1x1, 2x1 or 2x2 and chroma 1x1.  Anything else returns a negative code. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))
#define ERR_FORMAT (-1)
#define ERR_UNSUPPORTED (-2)

typedef struct {
    int w, h, ncomp;
    int hs[3], vs[3], tq[3], td[3], ta[3];
    uint16_t q[4][64];          /* natural order */
    uint8_t bits[8][17];        /* [tc * 4 + th] */
    uint8_t vals[8][256];
    int have_q[4], have_h[8];
    int restart;
    const uint8_t *scan;        /* first byte of entropy-coded data */
    long scan_len;
} Hdr;

static const uint8_t ZZ[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                               41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                               30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};


/* cv::imread applies the Exif Orientation tag (IMREAD_COLOR without IMREAD_IGNORE_ORIENTATION): returns the
   tag's value (1 = upright / absent) from an APP1 "Exif" segment */
static int exif_orientation(const uint8_t *s, long n)
{
    if (n < 14 || memcmp(s, "Exif\0\0", 6)) return 1;
    const uint8_t *t = s + 6;
    const long tn = n - 6;
    const int le = t[0] == 'I' && t[1] == 'I';
    if (!le && !(t[0] == 'M' && t[1] == 'M')) return 1;
#define RD16(o) (le ? (t[o] | (t[(o) + 1] << 8)) : ((t[o] << 8) | t[(o) + 1]))
#define RD32(o) (le ? ((long)t[o] | ((long)t[(o) + 1] << 8) | ((long)t[(o) + 2] << 16) | ((long)t[(o) + 3] << 24)) \
                    : (((long)t[o] << 24) | ((long)t[(o) + 1] << 16) | ((long)t[(o) + 2] << 8) | (long)t[(o) + 3]))
    if (tn < 8) return 1;
    const long ifd = RD32(4);
    if (ifd < 8 || ifd + 2 > tn) return 1;
    const int cnt = RD16(ifd);
    for (int i = 0; i < cnt; i++) {
        const long e = ifd + 2 + 12L * i;
        if (e + 12 > tn) break;
        if (RD16(e) == 0x0112) return RD16(e + 8);
    }
#undef RD16
#undef RD32
    return 1;
}

static int parse(const uint8_t *f, long n, Hdr *H)
{
    memset(H, 0, sizeof(*H));
    if (n < 4 || f[0] != 0xFF || f[1] != 0xD8) return ERR_FORMAT;
    long p = 2;
    int have_sof = 0;
    while (p + 4 <= n) {
        if (f[p] != 0xFF) return ERR_FORMAT;
        while (p < n && f[p] == 0xFF) p++;          /* fill bytes */
        if (p >= n) return ERR_FORMAT;
        const int m = f[p++];
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
        if (m == 0xD9) return ERR_FORMAT;
        if (p + 2 > n) return ERR_FORMAT;
        const long L = (f[p] << 8) | f[p + 1];
        if (L < 2 || p + L > n) return ERR_FORMAT;
        const uint8_t *s = f + p + 2;
        const long sl = L - 2;
        if (m == 0xDB) {
            long i = 0;
            while (i < sl) {
                const int pq = s[i] >> 4, tq = s[i] & 15;
                i++;
                if (tq > 3 || i + (pq ? 128 : 64) > sl) return ERR_FORMAT;
                for (int k = 0; k < 64; k++) {
                    H->q[tq][ZZ[k]] = pq ? (uint16_t)((s[i] << 8) | s[i + 1]) : s[i];
                    i += pq ? 2 : 1;
                }
                H->have_q[tq] = 1;
            }
        } else if (m == 0xC4) {
            long i = 0;
            while (i < sl) {
                const int tc = s[i] >> 4, th = s[i] & 15;
                i++;
                if (tc > 1 || th > 3 || i + 16 > sl) return ERR_FORMAT;
                int cnt = 0;
                H->bits[tc * 4 + th][0] = 0;
                for (int k = 1; k <= 16; k++) cnt += (H->bits[tc * 4 + th][k] = s[i + k - 1]);
                i += 16;
                if (cnt > 256 || i + cnt > sl) return ERR_FORMAT;
                memcpy(H->vals[tc * 4 + th], s + i, cnt);
                i += cnt;
                H->have_h[tc * 4 + th] = 1;
            }
        } else if (m == 0xC0 || m == 0xC1) {
            if (sl < 6 || s[0] != 8) return ERR_UNSUPPORTED;
            H->h = (s[1] << 8) | s[2];
            H->w = (s[3] << 8) | s[4];
            H->ncomp = s[5];
            if (H->w == 0 || H->h == 0) return ERR_UNSUPPORTED;
            if ((H->ncomp != 1 && H->ncomp != 3) || sl < 6 + 3 * H->ncomp) return ERR_UNSUPPORTED;
            for (int c = 0; c < H->ncomp; c++) {
                if (s[6 + 3 * c] != c + 1) return ERR_UNSUPPORTED;   /* JFIF ids 1 2 3 = Y Cb Cr */
                H->hs[c] = s[7 + 3 * c] >> 4;
                H->vs[c] = s[7 + 3 * c] & 15;
                H->tq[c] = s[8 + 3 * c];
                if (H->tq[c] > 3) return ERR_FORMAT;
            }
            if (H->ncomp == 1) {
                H->hs[0] = H->vs[0] = 1;                             /* a single component is never subsampled */
            } else {
                if (H->hs[1] != 1 || H->vs[1] != 1 || H->hs[2] != 1 || H->vs[2] != 1) return ERR_UNSUPPORTED;
                const int k = H->hs[0] * 10 + H->vs[0];
                if (k != 11 && k != 21 && k != 22) return ERR_UNSUPPORTED;
            }
            have_sof = 1;
        } else if ((m >= 0xC2 && m <= 0xCF) && m != 0xC4 && m != 0xC8 && m != 0xCC) {
            return ERR_UNSUPPORTED;                                  /* progressive, lossless, arithmetic */
        } else if (m == 0xDD) {
            if (sl < 2) return ERR_FORMAT;
            H->restart = (s[0] << 8) | s[1];
        } else if (m == 0xE1) {
            const int o = exif_orientation(s, sl);
            if (o >= 2 && o <= 8) return ERR_UNSUPPORTED;            /* imread would rotate / mirror the image */
        } else if (m == 0xEE) {
            if (sl >= 12 && !memcmp(s, "Adobe", 5) && H->ncomp == 3 && s[11] != 1) return ERR_UNSUPPORTED;
        } else if (m == 0xDA) {
            if (!have_sof || sl < 1 || s[0] != H->ncomp || sl < 1 + 2 * H->ncomp + 3) return ERR_UNSUPPORTED;
            for (int c = 0; c < H->ncomp; c++) {
                if (s[1 + 2 * c] != c + 1) return ERR_UNSUPPORTED;
                H->td[c] = s[2 + 2 * c] >> 4;
                H->ta[c] = s[2 + 2 * c] & 15;
                if (H->td[c] > 3 || H->ta[c] > 3) return ERR_FORMAT;
                if (!H->have_h[H->td[c]] || !H->have_h[4 + H->ta[c]] || !H->have_q[H->tq[c]]) return ERR_FORMAT;
            }
            H->scan = f + p + L;
            H->scan_len = n - (p + L);
            return 0;
        }
        p += L;
    }
    return ERR_FORMAT;
}

/* ---- jdhuff.c: canonical codes, bit reader over the stuffed stream ---- */
typedef struct {
    int maxcode[18], valptr[17];
    const uint8_t *vals;
} HTab;

static void build(const uint8_t *bits, const uint8_t *vals, HTab *T)
{
    int code = 0, k = 0;
    for (int l = 1; l <= 16; l++) {
        T->valptr[l] = k - code;
        code += bits[l];
        k += bits[l];
        T->maxcode[l] = bits[l] ? code - 1 : -1;
        code <<= 1;
    }
    T->maxcode[17] = 0x7fffffff;
    T->vals = vals;
}

typedef struct {
    const uint8_t *s;
    long n, p;
    uint64_t acc;
    int cnt;
    int marker;     /* a marker was hit: feed zeros (libjpeg's behaviour on truncated data) */
} Bits;

static void fill(Bits *b)
{
    while (b->cnt <= 56) {
        int c = 0;
        if (!b->marker && b->p < b->n) {
            c = b->s[b->p];
            if (c == 0xFF) {
                const int d = b->p + 1 < b->n ? b->s[b->p + 1] : 0xD9;
                if (d == 0) { b->p += 2; }
                else { b->marker = 1; c = 0; }
            } else {
                b->p++;
            }
        }
        b->acc |= (uint64_t)c << (56 - b->cnt);
        b->cnt += 8;
    }
}

static int getbits(Bits *b, int n)
{
    if (!n) return 0;
    if (b->cnt < n) fill(b);
    const int v = (int)(b->acc >> (64 - n));
    b->acc <<= n;
    b->cnt -= n;
    return v;
}

static int decode_sym(Bits *b, const HTab *T)
{
    int code = 0;
    for (int l = 1; l <= 16; l++) {
        code = (code << 1) | getbits(b, 1);
        if (T->maxcode[l] >= 0 && code <= T->maxcode[l] && code + T->valptr[l] >= 0) {
            /* codes of length l occupy [maxcode - bits + 1, maxcode]; valptr maps them to vals */
            return T->vals[code + T->valptr[l]];
        }
    }
    return 0;   /* corrupt code: libjpeg warns and returns 0 */
}

static int extend(int v, int s) { return s && v < (1 << (s - 1)) ? v - (1 << s) + 1 : v; }

static void next_restart(Bits *b)
{
    /* discard the rest of the interval, step over the RSTn marker */
    while (b->p + 1 < b->n && !(b->s[b->p] == 0xFF && b->s[b->p + 1] >= 0xD0 && b->s[b->p + 1] <= 0xD7)) b->p++;
    if (b->p + 1 < b->n) b->p += 2;
    b->acc = 0;
    b->cnt = 0;
    b->marker = 0;
}

/* ---- jidctint.c jpeg_idct_islow: dequantised block -> 8x8 samples ---- */
#define FIX(x) ((int32_t)((x) * 8192.0 + 0.5))
#define DESCALE(x, n) (((x) + (1 << ((n) - 1))) >> (n))

static void idct_1d(const int32_t *in, int stride, int32_t *o, int shift)
{
    int32_t z2 = in[2 * stride], z3 = in[6 * stride];
    int32_t z1 = (z2 + z3) * FIX(0.541196100);
    const int32_t tmp2 = z1 + z3 * (-FIX(1.847759065)), tmp3 = z1 + z2 * FIX(0.765366865);
    z2 = in[0];
    z3 = in[4 * stride];
    const int32_t tmp0 = (z2 + z3) * 8192, tmp1 = (z2 - z3) * 8192;
    const int32_t tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    int32_t t0 = in[7 * stride], t1 = in[5 * stride], t2 = in[3 * stride], t3 = in[1 * stride];
    z1 = t0 + t3;
    z2 = t1 + t2;
    z3 = t0 + t2;
    int32_t z4 = t1 + t3;
    const int32_t z5 = (z3 + z4) * FIX(1.175875602);
    t0 *= FIX(0.298631336);
    t1 *= FIX(2.053119869);
    t2 *= FIX(3.072711026);
    t3 *= FIX(1.501321110);
    z1 *= -FIX(0.899976223);
    z2 *= -FIX(2.562915447);
    z3 = z3 * (-FIX(1.961570560)) + z5;
    z4 = z4 * (-FIX(0.390180644)) + z5;
    t0 += z1 + z3;
    t1 += z2 + z4;
    t2 += z2 + z3;
    t3 += z1 + z4;
    o[0] = DESCALE(tmp10 + t3, shift);
    o[7] = DESCALE(tmp10 - t3, shift);
    o[1] = DESCALE(tmp11 + t2, shift);
    o[6] = DESCALE(tmp11 - t2, shift);
    o[2] = DESCALE(tmp12 + t1, shift);
    o[5] = DESCALE(tmp12 - t1, shift);
    o[3] = DESCALE(tmp13 + t0, shift);
    o[4] = DESCALE(tmp13 - t0, shift);
}

static void idct_block(const int16_t *coef, const uint16_t *q, uint8_t *out, long pitch)
{
    int32_t deq[64], ws[64], o[8];
    for (int i = 0; i < 64; i++) deq[i] = (int32_t)coef[i] * q[i];
    for (int c = 0; c < 8; c++) {                       /* columns, scaled up by 2^PASS1_BITS */
        idct_1d(deq + c, 8, o, 13 - 2);
        for (int r = 0; r < 8; r++) ws[8 * r + c] = o[r];
    }
    for (int r = 0; r < 8; r++) {
        idct_1d(ws + 8 * r, 1, o, 13 + 2 + 3);
        for (int c = 0; c < 8; c++) {
            const int v = o[c] + 128;
            out[r * pitch + c] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    }
}

ORC_API int orc_jpeg_info(const uint8_t *file, long n, int *w, int *h, int *ncomp)
{
    Hdr H;
    const int rc = parse(file, n, &H);
    if (rc) return rc;
    *w = H.w;
    *h = H.h;
    *ncomp = H.ncomp;
    return 0;
}

/* file -> BGR (IMREAD_COLOR), rows `stride` bytes apart */
ORC_API int orc_jpeg_decode(const uint8_t *file, long n, uint8_t *bgr, long stride)
{
    Hdr H;
    int rc = parse(file, n, &H);
    if (rc) return rc;
    const int hmax = H.hs[0], vmax = H.vs[0];
    const int mcux = (H.w + 8 * hmax - 1) / (8 * hmax), mcuy = (H.h + 8 * vmax - 1) / (8 * vmax);
    long pw[3], ph[3];
    uint8_t *plane[3] = {0, 0, 0};
    for (int c = 0; c < H.ncomp; c++) {
        pw[c] = (long)mcux * H.hs[c] * 8;
        ph[c] = (long)mcuy * H.vs[c] * 8;
        plane[c] = (uint8_t *)malloc((size_t)(pw[c] * ph[c]));
        if (!plane[c]) return -3;
    }
    HTab dc[3], ac[3];
    for (int c = 0; c < H.ncomp; c++) {
        build(H.bits[H.td[c]], H.vals[H.td[c]], &dc[c]);
        build(H.bits[4 + H.ta[c]], H.vals[4 + H.ta[c]], &ac[c]);
    }
    Bits b = {H.scan, H.scan_len, 0, 0, 0, 0};
    int pred[3] = {0, 0, 0};
    long mcu_count = 0;
    for (int my = 0; my < mcuy; my++) {
        for (int mx = 0; mx < mcux; mx++, mcu_count++) {
            if (H.restart && mcu_count && mcu_count % H.restart == 0) {
                next_restart(&b);
                pred[0] = pred[1] = pred[2] = 0;
            }
            for (int c = 0; c < H.ncomp; c++) {
                for (int by = 0; by < H.vs[c]; by++) {
                    for (int bx = 0; bx < H.hs[c]; bx++) {
                        int16_t blk[64];
                        memset(blk, 0, sizeof(blk));
                        int s = decode_sym(&b, &dc[c]);
                        int v = s ? extend(getbits(&b, s), s) : 0;
                        pred[c] += v;
                        blk[0] = (int16_t)pred[c];
                        for (int k = 1; k < 64; k++) {
                            const int rs = decode_sym(&b, &ac[c]);
                            const int r = rs >> 4;
                            s = rs & 15;
                            if (s) {
                                k += r;
                                v = extend(getbits(&b, s), s);
                                if (k < 64) blk[ZZ[k]] = (int16_t)v;
                            } else {
                                if (r != 15) break;
                                k += 15;
                            }
                        }
                        idct_block(blk, H.q[H.tq[c]],
                                   plane[c] + ((long)(my * H.vs[c] + by) * 8) * pw[c] + (long)(mx * H.hs[c] + bx) * 8, pw[c]);
                    }
                }
            }
        }
    }
    /* jdsample.c fancy upsampling of the chroma planes (edge rows as jdmainct.c presents them: the
       row above the first and below the last REAL downsampled row are those rows themselves),
       jdcolor.c conversion */
    const long cw = H.ncomp == 3 ? (H.w + hmax - 1) / hmax : 0, ch = H.ncomp == 3 ? (H.h + vmax - 1) / vmax : 0;
    uint8_t *up[2] = {0, 0};
    const long uw = cw * hmax;       /* upsampled row length (>= w) */
    for (int c = 1; c < H.ncomp; c++) {
        up[c - 1] = (uint8_t *)malloc((size_t)(uw * H.h + 16));
        const uint8_t *P = plane[c];
        for (long y = 0; y < H.h; y++) {
            uint8_t *o = up[c - 1] + y * uw;
            if (hmax == 1) {
                memcpy(o, P + y * pw[c], (size_t)cw);
            } else if (cw <= 2) {                /* jinit_upsampler: fancy only if downsampled_width > 2 */
                const uint8_t *in = P + (y / vmax) * pw[c];
                for (long x = 0; x < cw; x++) o[2 * x] = o[2 * x + 1] = in[x];
            } else if (vmax == 1) {              /* h2v1_fancy_upsample */
                const uint8_t *in = P + y * pw[c];
                o[0] = in[0];
                o[1] = (uint8_t)((in[0] * 3 + in[1] + 2) >> 2);
                for (long x = 1; x < cw - 1; x++) {
                    o[2 * x] = (uint8_t)((in[x] * 3 + in[x - 1] + 1) >> 2);
                    o[2 * x + 1] = (uint8_t)((in[x] * 3 + in[x + 1] + 2) >> 2);
                }
                o[2 * (cw - 1)] = (uint8_t)((in[cw - 1] * 3 + in[cw - 2] + 1) >> 2);
                o[2 * (cw - 1) + 1] = in[cw - 1];
            } else {                              /* h2v2_fancy_upsample */
                const long r0 = y >> 1;
                long r1 = (y & 1) ? r0 + 1 : r0 - 1;
                if (r1 < 0) r1 = 0;
                if (r1 > ch - 1) r1 = ch - 1;
                const uint8_t *a = P + r0 * pw[c], *bb = P + r1 * pw[c];
                int last, cur = a[0] * 3 + bb[0], next = a[1] * 3 + bb[1];
                o[0] = (uint8_t)((cur * 4 + 8) >> 4);
                o[1] = (uint8_t)((cur * 3 + next + 7) >> 4);
                for (long x = 1; x < cw - 1; x++) {
                    last = cur;
                    cur = next;
                    next = a[x + 1] * 3 + bb[x + 1];
                    o[2 * x] = (uint8_t)((cur * 3 + last + 8) >> 4);
                    o[2 * x + 1] = (uint8_t)((cur * 3 + next + 7) >> 4);
                }
                last = cur;
                cur = next;
                o[2 * (cw - 1)] = (uint8_t)((cur * 3 + last + 8) >> 4);
                o[2 * (cw - 1) + 1] = (uint8_t)((cur * 4 + 7) >> 4);
            }
        }
    }
    for (long y = 0; y < H.h; y++) {
        uint8_t *o = bgr + y * stride;
        const uint8_t *Y = plane[0] + y * pw[0];
        if (H.ncomp == 1) {
            for (long x = 0; x < H.w; x++) o[3 * x] = o[3 * x + 1] = o[3 * x + 2] = Y[x];
            continue;
        }
        const uint8_t *Cb = up[0] + y * uw, *Cr = up[1] + y * uw;
        for (long x = 0; x < H.w; x++) {
            const int yy = Y[x], cb = Cb[x] - 128, cr = Cr[x] - 128;
            const int r = yy + (int)((91881 * cr + 32768) >> 16);
            const int g = yy + (int)((-22554 * cb + 32768 - 46802 * cr) >> 16);
            const int bl = yy + (int)((116130 * cb + 32768) >> 16);
            o[3 * x] = (uint8_t)(bl < 0 ? 0 : bl > 255 ? 255 : bl);
            o[3 * x + 1] = (uint8_t)(g < 0 ? 0 : g > 255 ? 255 : g);
            o[3 * x + 2] = (uint8_t)(r < 0 ? 0 : r > 255 ? 255 : r);
        }
    }
    for (int c = 0; c < 3; c++) free(plane[c]);
    free(up[0]);
    free(up[1]);
    return 0;
}
