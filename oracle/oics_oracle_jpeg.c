/*
 * oics_oracle_jpeg.c -- CPU restatement of the output side of omr::correct_default:
 * TransformableMatrix::im_write(path, JPEG, 100) (transfer.rs:162-179, omr.rs:447-448) =
 * cv::imwrite(".jpg", bgr, {IMWRITE_JPEG_QUALITY, q}) = libjpeg(-turbo)'s baseline encoder
 * with OpenCV's defaults: JFIF APP0, two DQT, SOF0 with 4:2:0 sampling, the four standard
 * Huffman tables (no optimisation), one scan, no restart markers.
 *
 * TEST INFRASTRUCTURE ONLY (see oics_oracle.c).  The arithmetic lives in a third-party
 * dependency that is not in the reference tree (libjpeg-turbo inside OpenCV); what is restated
 * is its published integer pipeline -- jccolor.c rgb_ycc_convert, jcsample.c h2v2_downsample
 * (alternating bias 1, 2), jcprepct.c edge expansion, jfdctint.c (CONST_BITS 13, PASS1_BITS 2),
 * jcdctmgr.c quantisation (divide by 8 q, round half away from zero), jccoefct.c dummy blocks at
 * the right / bottom edge (AC zero, DC of the previous block), jchuff.c encode_one_block with
 * byte stuffing -- and it is pinned byte for byte against cv2.imencode in
 * tests/test_oracle_vs_cv2.py.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

static const uint8_t ZZ[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                               41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                               30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
static const uint8_t QL[64] = {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57,
                               69, 56, 14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64,
                               81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
static const uint8_t QC[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99,
                               99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                               99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
static const uint8_t DC_L_BITS[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
static const uint8_t DC_C_BITS[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
static const uint8_t DC_VALS[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
static const uint8_t AC_L_BITS[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
static const uint8_t AC_L_VALS[162] = {
    0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71, 0x14, 0x32,
    0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0a, 0x16,
    0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45,
    0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69,
    0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94,
    0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6,
    0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8,
    0xd9, 0xda, 0xe1, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
    0xf9, 0xfa};
static const uint8_t AC_C_BITS[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
static const uint8_t AC_C_VALS[162] = {
    0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22, 0x32, 0x81,
    0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16, 0x24, 0x34,
    0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44,
    0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68,
    0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92,
    0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4,
    0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6,
    0xd7, 0xd8, 0xd9, 0xda, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
    0xf9, 0xfa};

typedef struct { uint16_t code[256]; uint8_t len[256]; } hufftab;

static void build_huff(const uint8_t *bits, const uint8_t *vals, hufftab *t)
{
    memset(t, 0, sizeof(*t));
    unsigned code = 0;
    int k = 0;
    for (int l = 1; l <= 16; l++) {
        for (int i = 0; i < bits[l - 1]; i++, k++) {
            t->code[vals[k]] = (uint16_t)code++;
            t->len[vals[k]] = (uint8_t)l;
        }
        code <<= 1;
    }
}

/* jcparam.c: jpeg_quality_scaling + jpeg_add_quant_table(force_baseline = TRUE), natural order */
static void quant_table(const uint8_t *base, int quality, uint8_t *out)
{
    if (quality <= 0) quality = 1;
    if (quality > 100) quality = 100;
    const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
    for (int i = 0; i < 64; i++) {
        long v = ((long)base[i] * scale + 50L) / 100L;
        if (v <= 0) v = 1;
        if (v > 255) v = 255;
        out[i] = (uint8_t)v;
    }
}

/* jfdctint.c jpeg_fdct_islow on one row / column */
#define FIX13(x) ((int32_t)((x) * 8192.0 + 0.5))
static void fdct_1d(const int32_t *x, int32_t *o, int first)
{
    const int32_t t0 = x[0] + x[7], t7 = x[0] - x[7], t1 = x[1] + x[6], t6 = x[1] - x[6];
    const int32_t t2 = x[2] + x[5], t5 = x[2] - x[5], t3 = x[3] + x[4], t4 = x[3] - x[4];
    const int32_t t10 = t0 + t3, t13 = t0 - t3, t11 = t1 + t2, t12 = t1 - t2;
    const int sh = first ? 13 - 2 : 13 + 2;
    const int32_t rnd = (int32_t)1 << (sh - 1);
    if (first) {
        o[0] = (t10 + t11) * 4;
        o[4] = (t10 - t11) * 4;
    } else {
        o[0] = (t10 + t11 + 2) >> 2;
        o[4] = (t10 - t11 + 2) >> 2;
    }
    int32_t z1 = (t12 + t13) * FIX13(0.541196100);
    o[2] = (z1 + t13 * FIX13(0.765366865) + rnd) >> sh;
    o[6] = (z1 + t12 * (-FIX13(1.847759065)) + rnd) >> sh;
    z1 = t4 + t7;
    int32_t z2 = t5 + t6, z3 = t4 + t6, z4 = t5 + t7;
    const int32_t z5 = (z3 + z4) * FIX13(1.175875602);
    const int32_t u4 = t4 * FIX13(0.298631336), u5 = t5 * FIX13(2.053119869);
    const int32_t u6 = t6 * FIX13(3.072711026), u7 = t7 * FIX13(1.501321110);
    z1 *= -FIX13(0.899976223);
    z2 *= -FIX13(2.562915447);
    z3 = z3 * (-FIX13(1.961570560)) + z5;
    z4 = z4 * (-FIX13(0.390180644)) + z5;
    o[7] = (u4 + z1 + z3 + rnd) >> sh;
    o[5] = (u5 + z2 + z4 + rnd) >> sh;
    o[3] = (u6 + z2 + z3 + rnd) >> sh;
    o[1] = (u7 + z1 + z4 + rnd) >> sh;
}

/* samples [8][8] (already - 128) -> quantised coefficients in zigzag order */
static void fdct_quant(const int32_t *blk, const uint8_t *q, int16_t *zz)
{
    int32_t a[64], b[64], col[8], o[8];
    for (int r = 0; r < 8; r++) fdct_1d(blk + 8 * r, a + 8 * r, 1);
    for (int c = 0; c < 8; c++) {
        for (int r = 0; r < 8; r++) col[r] = a[8 * r + c];
        fdct_1d(col, o, 0);
        for (int r = 0; r < 8; r++) b[8 * r + c] = o[r];
    }
    for (int k = 0; k < 64; k++) {
        const int32_t v = b[ZZ[k]], d = (int32_t)q[ZZ[k]] << 3;
        zz[k] = (int16_t)(v < 0 ? -((-v + (d >> 1)) / d) : (v + (d >> 1)) / d);
    }
}

typedef struct { uint8_t *out; size_t cap, n; uint64_t acc; int nbits; } bitw;

static void put_byte(bitw *w, uint8_t b)
{
    if (w->n < w->cap) w->out[w->n] = b;
    w->n++;
}
static void put_bits(bitw *w, unsigned code, int size)
{
    if (!size) return;
    w->acc = (w->acc << size) | (code & ((1u << size) - 1u));
    w->nbits += size;
    while (w->nbits >= 8) {
        const uint8_t b = (uint8_t)(w->acc >> (w->nbits - 8));
        put_byte(w, b);
        if (b == 0xFF) put_byte(w, 0);
        w->nbits -= 8;
    }
}
static int nbits_of(int v) { int n = 0; while (v) { n++; v >>= 1; } return n; }

/* jchuff.c encode_one_block */
static void emit_block(bitw *w, const int16_t *zz, int *pred, const hufftab *dc, const hufftab *ac)
{
    int t = zz[0] - *pred, t2 = t;
    *pred = zz[0];
    if (t < 0) { t = -t; t2--; }
    int nb = nbits_of(t);
    put_bits(w, dc->code[nb], dc->len[nb]);
    put_bits(w, (unsigned)t2, nb);
    int r = 0;
    for (int k = 1; k < 64; k++) {
        int v = zz[k];
        if (v == 0) { r++; continue; }
        while (r > 15) { put_bits(w, ac->code[0xF0], ac->len[0xF0]); r -= 16; }
        t2 = v;
        if (v < 0) { v = -v; t2--; }
        nb = nbits_of(v);
        put_bits(w, ac->code[(r << 4) + nb], ac->len[(r << 4) + nb]);
        put_bits(w, (unsigned)t2, nb);
        r = 0;
    }
    if (r > 0) put_bits(w, ac->code[0], ac->len[0]);
}

static void put_marker_tables(bitw *w, int wd, int ht, const uint8_t *ql, const uint8_t *qc)
{
    static const uint8_t app0[] = {0xFF, 0xD8, 0xFF, 0xE0, 0, 16, 'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};
    for (size_t i = 0; i < sizeof(app0); i++) put_byte(w, app0[i]);
    for (int t = 0; t < 2; t++) {
        put_byte(w, 0xFF); put_byte(w, 0xDB); put_byte(w, 0); put_byte(w, 67); put_byte(w, (uint8_t)t);
        for (int k = 0; k < 64; k++) put_byte(w, (t ? qc : ql)[ZZ[k]]);
    }
    const uint8_t sof[] = {0xFF, 0xC0, 0, 17, 8, (uint8_t)(ht >> 8), (uint8_t)ht, (uint8_t)(wd >> 8), (uint8_t)wd,
                           3, 1, 0x22, 0, 2, 0x11, 1, 3, 0x11, 1};
    for (size_t i = 0; i < sizeof(sof); i++) put_byte(w, sof[i]);
    const struct { uint8_t id; const uint8_t *bits, *vals; int n; } dht[4] = {
        {0x00, DC_L_BITS, DC_VALS, 12}, {0x10, AC_L_BITS, AC_L_VALS, 162},
        {0x01, DC_C_BITS, DC_VALS, 12}, {0x11, AC_C_BITS, AC_C_VALS, 162}};
    for (int t = 0; t < 4; t++) {
        const int L = 2 + 1 + 16 + dht[t].n;
        put_byte(w, 0xFF); put_byte(w, 0xC4); put_byte(w, (uint8_t)(L >> 8)); put_byte(w, (uint8_t)L); put_byte(w, dht[t].id);
        for (int i = 0; i < 16; i++) put_byte(w, dht[t].bits[i]);
        for (int i = 0; i < dht[t].n; i++) put_byte(w, dht[t].vals[i]);
    }
    static const uint8_t sos[] = {0xFF, 0xDA, 0, 12, 3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 0x3F, 0};
    for (size_t i = 0; i < sizeof(sos); i++) put_byte(w, sos[i]);
}

#define FIX16(x) ((int32_t)((x) * 65536.0 + 0.5))
static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : v > hi ? hi : v; }

/* Returns the encoded size (the buffer is filled up to `cap`; a larger return value = too small). */
ORC_API long orc_jpeg_encode(const uint8_t *bgr, int w, int h, int stride, int quality, uint8_t *out, long cap)
{
    uint8_t ql[64], qc[64];
    quant_table(QL, quality, ql);
    quant_table(QC, quality, qc);
    hufftab dcl, dcc, acl, acc;
    build_huff(DC_L_BITS, DC_VALS, &dcl);
    build_huff(DC_C_BITS, DC_VALS, &dcc);
    build_huff(AC_L_BITS, AC_L_VALS, &acl);
    build_huff(AC_C_BITS, AC_C_VALS, &acc);
    bitw bw = {out, (size_t)(cap < 0 ? 0 : cap), 0, 0, 0};
    put_marker_tables(&bw, w, h, ql, qc);
    const int mcux = (w + 15) / 16, mcuy = (h + 15) / 16;
    const int wibY = (w + 7) / 8, hibY = (h + 7) / 8;
    int pred[3] = {0, 0, 0};
    int32_t blk[64];
    int16_t buf[4][64], cz[64];
    for (int my = 0; my < mcuy; my++)
        for (int mx = 0; mx < mcux; mx++) {
            /* Y: 2 x 2 blocks; samples past the edge replicate the last column / row (jcsample / jcprepct),
             * whole blocks past it are dummies (jccoefct.c) */
            int nb = 0;
            for (int dy = 0; dy < 2; dy++) {
                const int br = my * 2 + dy;
                for (int dx = 0; dx < 2; dx++, nb++) {
                    const int bc = mx * 2 + dx;
                    if (br >= hibY) {                     /* dummy row: DC of the last block of the row above */
                        memset(buf[nb], 0, sizeof(buf[nb]));
                        buf[nb][0] = buf[1][0];
                    } else if (bc >= wibY) {              /* dummy block at the right edge: DC of its left neighbour */
                        memset(buf[nb], 0, sizeof(buf[nb]));
                        buf[nb][0] = buf[nb - 1][0];
                    } else {
                        for (int y = 0; y < 8; y++)
                            for (int x = 0; x < 8; x++) {
                                const uint8_t *p = bgr + (size_t)clampi(br * 8 + y, 0, h - 1) * stride + 3 * clampi(bc * 8 + x, 0, w - 1);
                                const int Y = (FIX16(0.29900) * p[2] + FIX16(0.58700) * p[1] + FIX16(0.11400) * p[0] + 32768) >> 16;
                                blk[8 * y + x] = Y - 128;
                            }
                        fdct_quant(blk, ql, buf[nb]);
                    }
                }
            }
            for (int i = 0; i < 4; i++) emit_block(&bw, buf[i], &pred[0], &dcl, &acl);
            /* Cb, Cr: h2v2 box average with the alternating bias; the chroma plane is padded to an even
             * number of rows and to 16 wib_C columns before it, and its rows are replicated after it */
            for (int c = 1; c <= 2; c++) {
                for (int y = 0; y < 8; y++)
                    for (int x = 0; x < 8; x++) {
                        const int cy = clampi(my * 8 + y, 0, (h + 1) / 2 - 1), cx = mx * 8 + x;
                        int s = 0;
                        for (int j = 0; j < 2; j++)
                            for (int i = 0; i < 2; i++) {
                                const uint8_t *p = bgr + (size_t)clampi(2 * cy + j, 0, h - 1) * stride + 3 * clampi(2 * cx + i, 0, w - 1);
                                const int R = p[2], G = p[1], B = p[0];
                                s += c == 1 ? (-FIX16(0.16874) * R - FIX16(0.33126) * G + FIX16(0.50000) * B + (128 << 16) + 32767) >> 16
                                            : (FIX16(0.50000) * R - FIX16(0.41869) * G - FIX16(0.08131) * B + (128 << 16) + 32767) >> 16;
                            }
                        blk[8 * y + x] = ((s + ((cx & 1) ? 2 : 1)) >> 2) - 128;
                    }
                fdct_quant(blk, qc, cz);
                emit_block(&bw, cz, &pred[c], &dcc, &acc);
            }
        }
    if (bw.nbits) put_bits(&bw, 0x7F, 7);   /* pad the last byte with ones */
    put_byte(&bw, 0xFF);
    put_byte(&bw, 0xD9);
    return (long)bw.n;
}
