// N2 (input side): imgcodecs::imread(path, IMREAD_COLOR) (omr.rs:349, transfer.rs:43,62) of a
// baseline JPEG = libjpeg(-turbo)'s decompressor with its defaults, bit for bit, on the GPU -- so a
// card reaches the device as the few-MB file it is stored as instead of 26 MB of pixels, and the
// host does no codec work at all.
//
//   host              marker parsing (DQT / DHT / SOF0 / SOS), Huffman lookup tables -> one small upload
//   jdec_unstuff_*    0xFF 0x00 -> 0xFF: count per 32-byte chunk, scan, compact into MSB-first words
//   jdec_sync_kernel  Huffman decoding is sequential, so the stream is cut into 1024-bit subsequences
//                     and every thread decodes one from a guessed state (bit position, coefficient
//                     index, block of the MCU); Huffman streams self-synchronise, so feeding every
//                     thread its predecessor's exit state and repeating until nothing changes
//                     gives each subsequence its true entry state (Weissenberger & Schmidt, "Massively
//                     parallel Huffman decoding on GPUs", restated for JPEG's (position, k, block)
//                     state).  Round 0 from the guesses, then the "wide" launch: iterations inside each
//                     128-subsequence CTA until nothing changes there.  A quality-100 stream hardly
//                     self-synchronises (no EOBs), so the truth travels from the last empty MCU one
//                     subsequence per iteration: the walker (jdec_run) is built for the latency of one symbol.
//   jdec_walk_kernel  carries the states across the CTA borders: one lane per 128-subsequence chunk walks
//                     it in order and decodes only where the entry state differs from the one last used;
//                     launched until a launch decodes nothing.
//   scan              blocks completed per subsequence -> first block index of every subsequence
//   jdec_write_kernel the same decode once more, now storing coefficients [block][64] (natural order)
//   jdec_mcu_dc_*     DC prediction: per-MCU sums of the DC differences (three 21-bit fields in one
//                     64-bit word), scan
//   jdec_idct_kernel  8 threads per block: dequantise, jidctint.c islow IDCT (column then row through
//                     a conflict-free transpose), range limit -> component planes
//   jdec_color_kernel jdsample.c fancy h2v1 / h2v2 upsampling (jdmainct.c edge rows, jinit_upsampler's
//                     width rule), jdcolor.c YCC -> RGB, BGR out; jdec_color420_kernel: the 4:2:0 case,
//                     four pixels per thread
// Covered: 8-bit baseline / extended sequential, one interleaved scan, grey or Y Cb Cr with luma
// sampling 1x1, 2x1 or 2x2, any Huffman tables, restart intervals.  Anything else:
// OICS_ERR_UNSUPPORTED (the caller keeps its imread for those).
#include <algorithm>
#include <cstring>
#include <vector>

#include "common.h"

namespace oics {

static const uint8_t d_zz[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                                 41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// What the kernels need from the header; uploaded once per file.
// Huffman lookup in two levels, every entry ready for the walker's two recurrences (cursor, coefficient index):
//   bits 0-5   bits the symbol consumes: code length + value bits     bits 8-14  what it adds to k: 1 for a DC symbol;
//   bits 16-19 value bits (the symbol's low nibble)                    run + 1 for a coefficient, 16 for ZRL,
//   bits 20-24 code length                                             64 for EOB (= the block is complete)
//   JD_BOTH    a DC table entry that stands for "DC difference 0" AND the EOB behind it (the empty block of a
//              white card in one lookup): consumes both codes, adds 64
//   JD_L2      the first JD_LBITS bits are the prefix of longer codes: bits 0-5 = sub-table i, the symbol is
//              lut2[t][i * 64 + next 6 bits]
// JD_L2 | 63 in lut / 0 in lut2 = not in the tables (an invalid code, or more long prefixes than JD_L2SUBS -- the standard tables have 1 to 6, and
// long codes are the rare symbols of any table): jdhuff.c's canonical walk decides.
#define JD_LBITS 10
#define JD_BOTH 0x80000000u
#define JD_L2 0x40000000u
#define JD_L2SUBS 8
static inline uint32_t jdec_entry(int len, int sym, bool dc)
{
    const int s = sym & 15, r = sym >> 4;
    const int kinc = dc ? 1 : s ? r + 1 : r == 15 ? 16 : 64;
    return (uint32_t)(len + s) | ((uint32_t)kinc << 8) | ((uint32_t)s << 16) | ((uint32_t)len << 20);
}
struct JdecTables {
    uint32_t lut[6][1 << JD_LBITS];   // [2 * comp + (0 DC, 1 AC)]
    uint32_t lut2[6][JD_L2SUBS * 64];
    int32_t maxcode[6][18];   // canonical decode (jdhuff.c)
    int32_t valptr[6][18];
    uint8_t vals[6][256];
    uint16_t quant[3][64];    // natural order
    uint8_t zz[64];           // zigzag position -> natural index
    uint8_t comp_of[8];       // block of the MCU -> component
    int32_t nbm;              // blocks per MCU
};

struct JdecHeader {
    int w = 0, h = 0, ncomp = 0, hs = 1, vs = 1;   // luma sampling; chroma is 1x1
    size_t scan_off = 0, scan_len = 0;             // entropy-coded bytes (stuffed), EOI excluded
    int restart = 0;                               // DRI: MCUs per restart interval (0 = none)
};


// cv::imread applies the Exif Orientation tag (IMREAD_COLOR without IMREAD_IGNORE_ORIENTATION): returns the
// tag's value (1 = upright / absent) from an APP1 "Exif" segment
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

static int32_t parse_header(const uint8_t *f, size_t n, JdecHeader *H, JdecTables *T)
{
    if (n < 4 || f[0] != 0xFF || f[1] != 0xD8) { set_error("jpeg: not a JPEG file"); return OICS_ERR_INVALID_ARG; }
    uint8_t bits[8][17], vals[8][256];
    uint16_t q[4][64];
    bool have_h[8] = {false}, have_q[4] = {false};
    int tq[3] = {0, 0, 0}, hsf[3] = {1, 1, 1}, vsf[3] = {1, 1, 1};
    bool have_sof = false;
    size_t p = 2;
    auto bad = [&](const char *why) { set_error("jpeg: %s", why); return OICS_ERR_INVALID_ARG; };
    auto unsupported = [&](const char *why) { set_error("jpeg: %s is not covered by the CUDA decoder", why); return OICS_ERR_UNSUPPORTED; };
    while (p + 4 <= n) {
        if (f[p] != 0xFF) return bad("marker expected");
        while (p < n && f[p] == 0xFF) p++;
        if (p >= n) return bad("truncated");
        const int m = f[p++];
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
        if (m == 0xD9 || p + 2 > n) return bad("no scan");
        const size_t L = ((size_t)f[p] << 8) | f[p + 1];
        if (L < 2 || p + L > n) return bad("segment length");
        const uint8_t *s = f + p + 2;
        const size_t sl = L - 2;
        if (m == 0xDB) {
            for (size_t i = 0; i < sl;) {
                const int pq = s[i] >> 4, t = s[i] & 15;
                i++;
                if (t > 3 || i + (pq ? 128 : 64) > sl) return bad("DQT");
                for (int k = 0; k < 64; k++, i += pq ? 2 : 1) q[t][d_zz[k]] = pq ? (uint16_t)((s[i] << 8) | s[i + 1]) : s[i];
                have_q[t] = true;
            }
        } else if (m == 0xC4) {
            for (size_t i = 0; i < sl;) {
                const int tc = s[i] >> 4, th = s[i] & 15;
                i++;
                if (tc > 1 || th > 3 || i + 16 > sl) return bad("DHT");
                int cnt = 0;
                bits[tc * 4 + th][0] = 0;
                for (int k = 1; k <= 16; k++) cnt += (bits[tc * 4 + th][k] = s[i + k - 1]);
                i += 16;
                if (cnt > 256 || i + cnt > sl) return bad("DHT");
                memcpy(vals[tc * 4 + th], s + i, cnt);
                i += cnt;
                have_h[tc * 4 + th] = true;
            }
        } else if (m == 0xC0 || m == 0xC1) {
            if (sl < 6 || s[0] != 8) return unsupported("sample precision other than 8 bits");
            H->h = (s[1] << 8) | s[2];
            H->w = (s[3] << 8) | s[4];
            H->ncomp = s[5];
            if (!H->w || !H->h) return unsupported("a DNL-defined height");
            if ((H->ncomp != 1 && H->ncomp != 3) || sl < (size_t)(6 + 3 * H->ncomp)) return unsupported("this component count");
            for (int c = 0; c < H->ncomp; c++) {
                if (s[6 + 3 * c] != c + 1) return unsupported("component ids other than 1 2 3");
                hsf[c] = s[7 + 3 * c] >> 4;
                vsf[c] = s[7 + 3 * c] & 15;
                tq[c] = s[8 + 3 * c];
                if (tq[c] > 3) return bad("SOF");
            }
            if (H->ncomp == 1) {
                hsf[0] = vsf[0] = 1;
            } else {
                const int k = hsf[0] * 10 + vsf[0];
                if (hsf[1] != 1 || vsf[1] != 1 || hsf[2] != 1 || vsf[2] != 1 || (k != 11 && k != 21 && k != 22))
                    return unsupported("this chroma subsampling");
            }
            H->hs = hsf[0];
            H->vs = vsf[0];
            have_sof = true;
        } else if (m >= 0xC2 && m <= 0xCF && m != 0xC8 && m != 0xCC) {
            return unsupported("progressive / lossless / arithmetic coding");
        } else if (m == 0xDD) {
            if (sl < 2) return bad("DRI");
            H->restart = (s[0] << 8) | s[1];
        } else if (m == 0xE1) {
            const int o = exif_orientation(s, (long)sl);
            if (o >= 2 && o <= 8) return unsupported("an Exif orientation other than upright (imread would rotate the image)");
        } else if (m == 0xEE) {
            if (sl >= 12 && !memcmp(s, "Adobe", 5) && H->ncomp == 3 && s[11] != 1) return unsupported("an Adobe RGB / CMYK file");
        } else if (m == 0xDA) {
            if (!have_sof || sl < 1 || s[0] != H->ncomp || sl < (size_t)(1 + 2 * H->ncomp + 3)) return unsupported("a non-interleaved scan");
            memset(T, 0, sizeof(*T));
            int nb = 0;
            for (int c = 0; c < H->ncomp; c++) {
                if (s[1 + 2 * c] != c + 1) return unsupported("a reordered scan");
                const int td = s[2 + 2 * c] >> 4, ta = s[2 + 2 * c] & 15;
                if (td > 3 || ta > 3 || !have_h[td] || !have_h[4 + ta] || !have_q[tq[c]]) return bad("missing table");
                for (int k = 0; k < 64; k++) T->quant[c][k] = q[tq[c]][k];
                for (int a = 0; a < 2; a++) {
                    const uint8_t *b = bits[a ? 4 + ta : td], *v = vals[a ? 4 + ta : td];
                    const int t = 2 * c + a;
                    int code = 0, k = 0, nsubs = 0;
                    for (int l = 1; l <= 16; l++) {
                        T->valptr[t][l] = k - code;
                        for (int i = 0; i < b[l]; i++, code++, k++) {
                            if (code >> l) continue;                      // an over-subscribed table: the canonical walk decides
                            const unsigned c16 = (unsigned)code << (16 - l);   // the code, left-aligned in 16 bits
                            const uint32_t ent = jdec_entry(l, v[k], a == 0);
                            if (l <= JD_LBITS) {
                                for (unsigned fill = 0; fill < (1u << (JD_LBITS - l)); fill++)
                                    T->lut[t][(c16 >> (16 - JD_LBITS)) + fill] = ent;
                                continue;
                            }
                            uint32_t &pe = T->lut[t][c16 >> (16 - JD_LBITS)];
                            if (pe == 0 && nsubs < JD_L2SUBS) pe = JD_L2 | (uint32_t)nsubs++;
                            if (!(pe & JD_L2)) continue;                  // no sub-table left (or a shorter code owns the prefix)
                            uint32_t *sub = T->lut2[t] + (pe & 63u) * 64;
                            const unsigned rest = c16 & 63u;
                            for (unsigned fill = 0; fill < (1u << (16 - l)); fill++) sub[rest + fill] = ent;
                        }
                        T->maxcode[t][l] = b[l] ? code - 1 : -1;
                        code <<= 1;
                    }
                    memcpy(T->vals[t], v, 256);
                }
                // the commonest block of a scanned card -- DC difference 0, then end of block -- in ONE lookup:
                // where a DC code of category 0 and the EOB code behind it fit the window together
                for (int idx = 0; idx < (1 << JD_LBITS); idx++) {
                    const uint32_t e = T->lut[2 * c][idx];
                    if (!e || (e & JD_L2) || ((e >> 16) & 15u)) continue;           // a DC code whose low nibble is 0
                    const int l1 = (int)((e >> 20) & 31u);
                    const uint32_t e2 = T->lut[2 * c + 1][(idx << l1) & ((1 << JD_LBITS) - 1)];
                    if (!e2 || (e2 & JD_L2) || ((e2 >> 8) & 127u) != 64u) continue;  // an EOB behind it
                    const int l2 = (int)((e2 >> 20) & 31u);
                    if (l1 + l2 <= JD_LBITS) T->lut[2 * c][idx] = JD_BOTH | (uint32_t)(l1 + l2) | (64u << 8) | ((uint32_t)(l1 + l2) << 20);
                }
                // what no code of <= JD_LBITS bits and no long prefix claims: one flag test on the walker's path
                for (int a = 0; a < 2; a++)
                    for (int idx = 0; idx < (1 << JD_LBITS); idx++)
                        if (T->lut[2 * c + a][idx] == 0) T->lut[2 * c + a][idx] = JD_L2 | 63u;
                for (int i = 0; i < hsf[c] * vsf[c]; i++) T->comp_of[nb++] = (uint8_t)c;
            }
            T->nbm = nb;
            memcpy(T->zz, d_zz, 64);
            H->scan_off = p + L;
            size_t end = n;                                  // EOI (and anything behind it) is not entropy data
            if (end >= 2 && f[end - 2] == 0xFF && f[end - 1] == 0xD9) end -= 2;
            if (end <= H->scan_off) return bad("empty scan");
            H->scan_len = end - H->scan_off;
            return OICS_OK;
        }
        p += L;
    }
    return bad("no scan");
}

// ---------------------------------------------------------------------------
// unstuffing: byte j of the scan is dropped iff it is the 0x00 behind a 0xFF or, in a file with restart
// intervals, one of the two bytes of an RSTn marker.  Any other byte behind a 0xFF is a marker that does
// not belong into the scan (premature EOI, or RSTn without DRI): *flag |= 1.  cnt[i] = dropped bytes of
// chunk i in the low word, RSTn markers (counted at their second byte) in the high word.
__device__ __forceinline__ bool is_rst(uint8_t b) { return b >= 0xD0 && b <= 0xD7; }

__global__ void __launch_bounds__(256) jdec_unstuff_count_kernel(const uint8_t *__restrict__ s, size_t n, int restart,
                                                                 unsigned long long *__restrict__ cnt, int *__restrict__ flag)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i * 32 >= n) return;
    unsigned c = 0, marks = 0;
    bool marker = false;
    uint8_t prev = i ? s[i * 32 - 1] : 0;
    const size_t j1 = min(n, i * 32 + 32);
    for (size_t j = i * 32; j < j1; j++) {
        const uint8_t b = s[j];
        if (prev == 0xFF) {
            if (b == 0) c++;
            else if (restart && is_rst(b)) { c++; marks++; }
            else marker = true;
        } else if (restart && b == 0xFF && j + 1 < n && is_rst(s[j + 1])) {
            c++;
        }
        prev = b;
    }
    if (j1 == n && prev == 0xFF) marker = true;
    cnt[i] = (unsigned long long)c | ((unsigned long long)marks << 32);
    if (marker) atomicOr(flag, 1);
}

// kept bytes -> MSB-first 32-bit words (byte i of the stream lands at byte address i ^ 3); the bit position
// behind the r-th RSTn marker (= where restart interval r + 1 begins) -> rst[r]
__global__ void __launch_bounds__(256) jdec_unstuff_write_kernel(const uint8_t *__restrict__ s, size_t n, int restart,
                                                                 const unsigned long long *__restrict__ off,
                                                                 const unsigned long long *__restrict__ off_base,
                                                                 uint8_t *__restrict__ out, unsigned long long *__restrict__ rst,
                                                                 unsigned nrst)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i * 32 >= n) return;
    const unsigned long long before = off[i] + off_base[i >> 10];   // both fields at once: dropped bytes | markers << 32
    size_t o = i * 32 - (size_t)(before & 0xffffffffull);
    unsigned mark = (unsigned)(before >> 32);
    uint8_t prev = i ? s[i * 32 - 1] : 0;
    const size_t j1 = min(n, i * 32 + 32);
    for (size_t j = i * 32; j < j1; j++) {
        const uint8_t b = s[j];
        bool drop = false;
        if (prev == 0xFF) {
            if (b == 0) drop = true;
            else if (restart && is_rst(b)) {
                drop = true;
                if (mark < nrst) rst[mark] = 8ull * o;
                mark++;
            }
        } else if (restart && b == 0xFF && j + 1 < n && is_rst(s[j + 1])) {
            drop = true;
        }
        if (!drop) out[(o++) ^ 3] = b;
        prev = b;
    }
}

// ---------------------------------------------------------------------------
// The sequential decoder of jdhuff.c as a state machine: state = (bit position p, coefficient index
// k of the block being filled, block bi of the MCU).  One step = one Huffman symbol plus its value
// bits.  SINK receives (block ordinal since the start, zigzag position, value) of every coefficient
// (the DC value is the DIFFERENCE; prediction is a later scan).
#define JD_SUB 1024   // bits per subsequence (r02m: 2048 measured -- every pass twice as long on one card, the same cards/s in the batch)

struct JdecSmall {            // what every walker keeps in shared memory
    int32_t maxcode[6][18];
    int32_t valptr[6][18];
    uint8_t vals[6][256];
    uint32_t comps;           // component of block i of the MCU in bits 2i, 2i + 1
    int32_t nbm;
};
// The lookup tables (36 KB): a shared-memory copy for the short-lived, all-lanes-busy kernels; the CTAs of the
// synchronisation live long and mostly wait, so they read the tables in place (global memory through L1,
// a few cycles more per symbol) and 13 of them fit an SM instead of 4.
struct JdecLuts {
    uint32_t lut[6][1 << JD_LBITS];
    uint32_t lut2[6][JD_L2SUBS * 64];
};
struct JdecShared {
    JdecLuts luts;            // first: filled with 16-byte copies
    JdecSmall small;
};

__device__ __forceinline__ void jdec_load_small(JdecSmall &S, const JdecTables *__restrict__ T)
{
    for (int i = threadIdx.x; i < 6 * 256 / 4; i += blockDim.x)
        reinterpret_cast<uint32_t *>(&S.vals[0][0])[i] = __ldg(reinterpret_cast<const uint32_t *>(&T->vals[0][0]) + i);
    for (int i = threadIdx.x; i < 6 * 18; i += blockDim.x) {
        (&S.maxcode[0][0])[i] = (&T->maxcode[0][0])[i];
        (&S.valptr[0][0])[i] = (&T->valptr[0][0])[i];
    }
    if (threadIdx.x == 0) {
        uint32_t cs = 0;
        for (int i = 0; i < 8; i++) cs |= (uint32_t)(T->comp_of[i] & 3) << (2 * i);
        S.comps = cs;
        S.nbm = T->nbm;
    }
}

__device__ __forceinline__ void jdec_load_tables(JdecShared &S, const JdecTables *__restrict__ T)
{
    static_assert(offsetof(JdecTables, lut2) == offsetof(JdecTables, lut) + sizeof(JdecLuts::lut), "lut, lut2 adjacent");
    for (int i = threadIdx.x; i < (int)(sizeof(JdecLuts) / 16); i += blockDim.x)
        reinterpret_cast<uint4 *>(&S.luts)[i] = __ldg(reinterpret_cast<const uint4 *>(&T->lut[0][0]) + i);
    jdec_load_small(S.small, T);
    __syncthreads();
}

__device__ __forceinline__ unsigned long long jdec_pack(unsigned long long p, int k, int bi)
{
    return (p << 12) | ((unsigned long long)k << 4) | (unsigned long long)bi;
}

struct NoSink {
    __device__ __forceinline__ void operator()(unsigned, int, int) {}
};

// Decode from state (p, k, bi) until p >= end.  Returns the number of blocks completed.
// The CTA's part of the stream (128 subsequences) staged in shared memory, one ROW per subsequence: its 32 words
// and the 2 behind them (a symbol may start up to 31 bits past the subsequence and reach 32 further), so a
// thread addresses its words with the cursor alone -- row[cursor >> 5] -- and rows 35 words apart keep the
// lanes of a warp in different banks.
#define JD_WPS (JD_SUB / 32)   // words per subsequence
#define JD_ROW (JD_WPS + 3)    // 34 staged words + one slot the look-ahead may touch past the end
#define JD_WIN_WORDS (128 * JD_ROW)
#define JD_DYN_SMEM (sizeof(JdecShared) + JD_WIN_WORDS * sizeof(uint32_t))   // above the 48 KB a kernel gets unasked
#define JD_SYNC_SMEM (sizeof(JdecSmall) + 16 + JD_WIN_WORDS * sizeof(uint32_t))

__device__ __forceinline__ void jdec_stage_words(uint32_t *sw, const uint32_t *__restrict__ words)
{
    const size_t base = (size_t)blockIdx.x * 128 * JD_WPS;
    for (int i = threadIdx.x; i < 128 * (JD_WPS + 2); i += blockDim.x) {
        const int row = i / (JD_WPS + 2), l = i - row * (JD_WPS + 2);
        sw[row * JD_ROW + l] = __ldg(words + base + row * JD_WPS + l);      // (the stream is zero-padded behind its end)
    }
    __syncthreads();
}

// Restart intervals: rst[r] = first bit of interval r + 1.  In front of it the stream holds up to 7 padding
// one-bits, which no Huffman code consists of, so the symbol that would reach across the boundary is the
// padding: the decoder drops it and continues at the boundary with the state of a fresh interval
// (k = 0, block 0; the DC predictors are reset by the DC scan).  The interval of a state follows from its
// bit position alone.
//
// The walk is a chain of dependent symbols, so its speed is the latency of ONE trip:
//   cursor -> funnel shift -> table entry -> cursor          and          entry -> k -> table of the next symbol.
// Everything is 32-bit and relative to the subsequence (`base` = its first bit; row = its staged words), the entry
// carries the sums both recurrences need (jdec_entry), the stream is read through three registers (the two
// words under the cursor and the next one, fetched a symbol ahead), and everything that is not per symbol is
// kept out of the trip: restart boundaries are a template variant (HAS_RST), MCU bookkeeping (the pattern of
// empty MCUs) lives in the block-end branch, value bits / stores hang off the chain.  Branches are structured
// (no continue), so the lanes of a warp that walk different subsequences re-join after every symbol.
template <bool GLOBAL_LUTS, bool HAS_RST, typename SINK>
__device__ __forceinline__ unsigned jdec_run_t(const JdecSmall &S, const uint32_t *__restrict__ L, const uint32_t *row,
                                               unsigned long long base, unsigned long long &p_io, int &k_io, int &bi_io,
                                               unsigned long long end, SINK &sink, const unsigned long long *__restrict__ rst,
                                               unsigned nrst)
{
    unsigned ri = 0;                                 // boundaries at or before p
    if (HAS_RST) {
        unsigned lo = 0, hi = nrst;
        while (lo < hi) {
            const unsigned mid = (lo + hi) >> 1;
            if (__ldg(rst + mid) <= p_io) lo = mid + 1;
            else hi = mid;
        }
        ri = lo;
    }
    auto rel = [&](unsigned long long x) { return x - base > 0x3fffffffull ? 0x3fffffff : (int)(x - base); };
    int bound = HAS_RST && ri < nrst ? rel(__ldg(rst + ri)) : 0x3fffffff;
    int q = (int)(p_io - base), k = k_io, bi = bi_io;
    const int endq = rel(end), nbm = S.nbm;
    const uint32_t comps = S.comps;
    // L: lut[6][..], then lut2[6][..] -- in global memory, or a shared-memory copy addressed as such (through the
    // generic pointer the compiler rebuilds the shared window's base, a special-register read, on every trip)
    const uint32_t lbase = GLOBAL_LUTS ? 0u : (uint32_t)__cvta_generic_to_shared(L);
    auto lut = [&](int i) -> uint32_t {
        if (GLOBAL_LUTS) return __ldg(L + i);
        uint32_t r;
        asm("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(lbase + 4u * (uint32_t)i));
        return r;
    };
    // tables as entry offsets into L: the block's DC table, its AC table behind it, and the DC table of the block
    // after it (looked up while this block is walked, so a block end costs a move)
    auto dc_of = [&](int blk) { return (int)(((comps >> (2 * blk)) & 3u) << (JD_LBITS + 1)); };
    int dc = dc_of(bi), dc_next = dc_of(bi + 1 == nbm ? 0 : bi + 1);
    int tb = k ? dc + (1 << JD_LBITS) : dc;
    uint32_t a = row[q >> 5], b = row[(q >> 5) + 1], c = row[(q >> 5) + 2];
    unsigned done = 0;
    // Runs of empty MCUs (the white of a card: every block "DC difference 0, EOB"): once one MCU has been parsed
    // that way from its first bit, every following copy of its bits (at most 32) parses the same way -- a Huffman
    // parse depends on the state and the bits consumed only -- so the run advances one compare per MCU instead
    // of one table walk per block.  A copy is taken only while it ends at or before `end` and the restart
    // boundary: every state inside it lies before `end`, so the exit state is the one the block-by-block walk
    // reaches.  The blocks of such an MCU store nothing (the coefficient buffer starts at zero).
    int mstart = q, patlen = 0;
    uint32_t vstart = __funnelshift_l(b, a, q & 31), pat = 0;
    uint32_t plain = (k | bi) == 0 ? JD_BOTH : 0u;   // bit 31: the MCU under the cursor has been empty blocks only so far
    // The trip is straight-line code (selects, one look-ahead load) with three rarely taken branches: an entry
    // outside the first table, a restart boundary, the end of a block.
    while (q < endq) {
        const int sh = q & 31;
        const uint32_t v = __funnelshift_l(b, a, sh);   // the next 32 bits of the stream
        uint32_t e = lut(tb + (int)(v >> (32 - JD_LBITS)));
        if (e & JD_L2) {                                 // a long code, or nothing in the first table
            const int t = tb >> JD_LBITS;
            const uint32_t sub = e & 63u;
            e = sub < JD_L2SUBS ? lut((6 << JD_LBITS) + t * (JD_L2SUBS * 64) + (int)(sub * 64 + ((v >> (32 - JD_LBITS - 6)) & 63u))) : 0u;
            if (e == 0) {                                // not in the tables: jdhuff.c's canonical walk
                int len = 16, sym = 0;
                const int code16 = (int)(v >> 16);
#pragma unroll 1
                for (int l = 1; l <= 16; l++) {
                    const int cd = code16 >> (16 - l);
                    if (cd <= S.maxcode[t][l]) {
                        len = l;
                        sym = S.vals[t][(cd + S.valptr[t][l]) & 255];
                        break;
                    }
                }
                const int s = sym & 15, r = sym >> 4;
                const int kinc = k == 0 ? 1 : s ? r + 1 : r == 15 ? 16 : 64;
                e = (uint32_t)(len + s) | ((uint32_t)kinc << 8) | ((uint32_t)s << 16) | ((uint32_t)len << 20);
            }
        }
        const int adv = (int)(e & 63u);
        if (HAS_RST && q + adv > bound) {                    // padding in front of a restart marker
            q = bound;
            k = 0;
            bi = 0;
            dc = dc_of(0);
            dc_next = dc_of(nbm > 1 ? 1 : 0);
            tb = dc;
            ri++;
            bound = ri < nrst ? rel(__ldg(rst + ri)) : 0x3fffffff;
            a = row[q >> 5];
            b = row[(q >> 5) + 1];
            c = row[(q >> 5) + 2];
            mstart = q;
            vstart = __funnelshift_l(b, a, q & 31);
            plain = JD_BOTH;
            continue;
        }
        const int s = (int)((e >> 16) & 15u);
        const int pos = k + (int)((e >> 8) & 127u) - 1;      // where the coefficient goes (zigzag order)
        if (s && pos < 64) {
            const int len = (int)((e >> 20) & 31u);
            int val = (int)((v << len) >> (32 - s));
            if (val < (1 << (s - 1))) val += 1 - (1 << s);
            sink(done, pos, val);
        }
        k = pos + 1;
        plain &= e;
        q += adv;
        {                                                    // at most one word per trip (adv <= 32)
            const bool cross = sh + adv >= 32;
            const uint32_t nw = row[(q >> 5) + 2];
            a = cross ? b : a;
            b = cross ? c : b;
            c = cross ? nw : c;
        }
        tb = dc + (1 << JD_LBITS);
        if (k >= 64) {                                       // the block is complete
            k = 0;
            bi = bi + 1 == nbm ? 0 : bi + 1;
            dc = dc_next;
            dc_next = dc_of(bi + 1 == nbm ? 0 : bi + 1);
            tb = dc;
            done++;
            if (bi == 0) {                                   // ... and so is the MCU
                if ((plain & JD_BOTH) && q - mstart <= 32) {
                    patlen = q - mstart;
                    pat = vstart >> (32 - patlen);
                }
                uint32_t v2 = __funnelshift_l(b, a, q & 31);
                if (patlen) {
                    const int lim = min(endq, bound);
                    while ((v2 >> (32 - patlen)) == pat && q + patlen <= lim) {
                        const int sh2 = q & 31;
                        q += patlen;
                        done += nbm;
                        if (sh2 + patlen >= 32) {
                            a = b;
                            b = c;
                            c = row[(q >> 5) + 2];
                        }
                        v2 = __funnelshift_l(b, a, q & 31);
                    }
                }
                mstart = q;
                vstart = v2;
                plain = JD_BOTH;
            }
        }
    }
    p_io = base + (unsigned long long)q;
    k_io = k;
    bi_io = bi;
    return done;
}

template <bool GLOBAL_LUTS, typename SINK>
__device__ __forceinline__ unsigned jdec_run(const JdecSmall &S, const uint32_t *__restrict__ L, const uint32_t *row,
                                             unsigned long long base, unsigned long long &p, int &k, int &bi, unsigned long long end,
                                             SINK &sink, const unsigned long long *__restrict__ rst, unsigned nrst)
{
    return nrst ? jdec_run_t<GLOBAL_LUTS, true>(S, L, row, base, p, k, bi, end, sink, rst, nrst)
                : jdec_run_t<GLOBAL_LUTS, false>(S, L, row, base, p, k, bi, end, sink, rst, nrst);
}

// One round of the synchronisation: thread j decodes subsequence j from its predecessor's exit state
// unless that is the state it used last time.  round 0: from the guess (j * JD_SUB, 0, 0).
// Uniform stretches (the white of a scanned card: every MCU is the same few bits) do not
// self-synchronise in (k, block): decoders that entered with a wrong phase parse them consistently,
// and their states at the subsequence boundaries cycle through a few values round after round until
// the true state arrives from the left, one subsequence per round.  Each thread therefore remembers
// the last JD_M (entry -> exit, blocks) pairs it decoded: a round whose input it has seen before costs
// a lookup, so only the front of the true chain decodes.
#define JD_M 4

__global__ void __launch_bounds__(128, 10) jdec_sync_kernel(const JdecTables *__restrict__ T, const uint32_t *__restrict__ words,
                                                        const unsigned long long *__restrict__ removed, size_t scan_len,
                                                        unsigned long long *used, unsigned long long *out,
                                                        unsigned long long *__restrict__ nblk, unsigned long long *mkey,
                                                        unsigned long long *mout, unsigned *mnb, unsigned *mcnt, int round,
                                                        int max_it, int *__restrict__ changed,
                                                        const unsigned long long *__restrict__ rst, unsigned nrst)
{
    extern __shared__ uint4 jdec_dyn[];                       // JdecSmall, then the staged window (JD_SYNC_SMEM bytes)
    JdecSmall &S = *reinterpret_cast<JdecSmall *>(jdec_dyn);
    uint32_t *sw = reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(jdec_dyn) + ((sizeof(JdecSmall) + 15) & ~(size_t)15));
    const unsigned long long total = 8ull * (scan_len - (*removed & 0xffffffffull));
    const unsigned long long j = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = j * JD_SUB < total;
    bool tables = false;
    unsigned long long mine = live && round > 0 ? used[j] : ~0ull;   // the entry state of my last decode / lookup
    unsigned cnt = live && round > 0 ? mcnt[j] : 0;
    // Inside the CTA the chain advances without leaving the kernel: one iteration = every thread looks
    // at its predecessor's exit again.  Across CTAs it advances with the next launch.
    const uint32_t *row = sw + JD_ROW * threadIdx.x;          // this thread's subsequence
    for (int it = 0; it < (round ? max_it : 1); it++) {
        bool need = false, hit = false;
        unsigned long long x = jdec_pack(j * JD_SUB, 0, 0);
        if (live) {
            if (round == 0) {
                need = true;
            } else {
                if (j > 0) x = __ldcg(out + j - 1);
                if (x != mine) {
                    need = true;
                    for (unsigned b = 0; b < min(cnt, (unsigned)JD_M); b++) {
                        if (mkey[j * JD_M + b] == x) {          // seen before
                            mine = x;
                            __stcg(out + j, mout[j * JD_M + b]);
                            nblk[j] = mnb[j * JD_M + b];
                            need = false;
                            hit = true;
                            break;
                        }
                    }
                }
            }
        }
        const int any_need = __syncthreads_or(need), any_hit = __syncthreads_or(hit);
        if (!any_need && !any_hit) break;
        if (round > 0 && threadIdx.x == 0) changed[round] = 1;
        if (any_need && !tables) {
            jdec_load_small(S, T);
            jdec_stage_words(sw, words);
            tables = true;
        }
        if (need) {
            unsigned long long p = x >> 12;
            int k = (int)((x >> 4) & 255), bi = (int)(x & 15);
            NoSink sink;
            const unsigned n = jdec_run<true>(S, &T->lut[0][0], row, j * JD_SUB, p, k, bi, min(total, (j + 1) * JD_SUB), sink, rst, nrst);
            const unsigned long long e = jdec_pack(p, k, bi);
            mine = x;
            __stcg(out + j, e);
            nblk[j] = n;
            const unsigned slot = cnt % JD_M;               // the oldest entry is replaced
            mkey[j * JD_M + slot] = x;
            mout[j * JD_M + slot] = e;
            mnb[j * JD_M + slot] = n;
            cnt++;
        }
        __syncthreads();                                    // exits written above are read by the neighbours next
    }
    if (live) {
        used[j] = mine;
        mcnt[j] = cnt;
    }
}

// The chunk walk.  Where the stream rarely holds an EOB (quality ~100: nearly every block has all 63 AC
// coefficients) a decoder that entered with a wrong (k, block) never re-aligns, so the true state has to be
// carried through such stretches subsequence by subsequence.  In the kernel above that keeps a 128-thread CTA
// resident per 128 subsequences with one lane decoding.  Here ONE lane per warp
// walks a chunk of 128 subsequences in order, carrying the exit state along and decoding only where the
// entry differs from the one last used; the other lanes fetch: the chunk's stored states once, and the 34 words
// of the subsequence about to be decoded.  The same decodes as the chain's critical path with a quarter of
// the resident warps and no barriers.  Chunks are walked concurrently; a chunk re-walks when its predecessor's
// last exit changes under it, and the next launch catches what the last one missed (`changed[round]` says
// whether a launch decoded anything).  from_guess: the first walk of a stream whose wide rounds were skipped.
#define JD_CHUNK 128
#define JD_WALK_WARPS 4
__global__ void __launch_bounds__(32 * JD_WALK_WARPS) jdec_walk_kernel(const JdecTables *__restrict__ T, const uint32_t *__restrict__ words,
                                                                      const unsigned long long *__restrict__ removed, size_t scan_len,
                                                                      unsigned long long *used, unsigned long long *out,
                                                                      unsigned long long *__restrict__ nblk, int round, int from_guess,
                                                                      int *__restrict__ changed, const unsigned long long *__restrict__ rst,
                                                                      unsigned nrst)
{
    __shared__ JdecShared S;
    __shared__ unsigned long long s_used[JD_WALK_WARPS][JD_CHUNK], s_out[JD_WALK_WARPS][JD_CHUNK];
    __shared__ uint32_t s_words[JD_WALK_WARPS][JD_ROW];
    jdec_load_tables(S, T);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned long long total = 8ull * (scan_len - (*removed & 0xffffffffull));
    const unsigned long long nsub = (total + JD_SUB - 1) / JD_SUB;
    const unsigned long long j0 = ((unsigned long long)blockIdx.x * JD_WALK_WARPS + warp) * JD_CHUNK;
    if (j0 >= nsub) return;                                   // whole warps leave: no barrier below
    const int n = (int)min((unsigned long long)JD_CHUNK, nsub - j0);
    for (int i = lane; i < n; i += 32) {
        s_used[warp][i] = from_guess ? ~0ull : used[j0 + i];
        s_out[warp][i] = from_guess ? 0ull : out[j0 + i];
    }
    __syncwarp();
    bool any = false;
    unsigned long long entry = 0;
    for (int pass = 0; pass < 4; pass++) {
        // the state my predecessor chunk ends in right now (it may still be walking)
        unsigned long long e0 = j0 == 0 || from_guess ? jdec_pack(j0 * JD_SUB, 0, 0) : __ldcg(out + j0 - 1);
        e0 = __shfl_sync(0xffffffffu, e0, 0);
        if (pass && e0 == entry) break;
        entry = e0;
        unsigned long long prev = entry;
        for (int i = 0; i < n; i++) {
            // lane 0 knows whether subsequence i needs a decode; everybody follows its decision
            const bool need = __shfl_sync(0xffffffffu, (int)(prev != s_used[warp][i]), 0) != 0;
            if (!need) {
                prev = s_out[warp][i];
                continue;
            }
            const unsigned long long wbase = (j0 + i) * JD_WPS;
            s_words[warp][lane] = __ldg(words + wbase + lane);
            if (lane < 2) s_words[warp][JD_WPS + lane] = __ldg(words + wbase + JD_WPS + lane);
            __syncwarp();
            if (lane == 0) {
                unsigned long long pp = prev >> 12;
                int k = (int)((prev >> 4) & 255), bi = (int)(prev & 15);
                NoSink sink;
                const unsigned nb = jdec_run<false>(S.small, &S.luts.lut[0][0], s_words[warp], (j0 + i) * JD_SUB, pp, k, bi, min(total, (j0 + i + 1) * JD_SUB), sink,
                                             rst, nrst);
                const unsigned long long e = jdec_pack(pp, k, bi);
                s_used[warp][i] = prev;
                s_out[warp][i] = e;
                used[j0 + i] = prev;
                __stcg(out + j0 + i, e);
                nblk[j0 + i] = nb;
                prev = e;
            }
            __syncwarp();
            any = true;
        }
        if (from_guess) break;
    }
    if (any && lane == 0) changed[round] = 1;
}

struct CoefSink {
    int16_t *coef;             // [block][64], natural order
    unsigned long long base;   // ordinal of the first block this thread touches
    unsigned long long nblocks;
    const uint8_t *zz;
    __device__ __forceinline__ void operator()(unsigned done, int k, int v)
    {
        const unsigned long long g = base + done;
        if (g < nblocks) coef[g * 64 + zz[k]] = (int16_t)v;
    }
};

__global__ void __launch_bounds__(128) jdec_write_kernel(const JdecTables *__restrict__ T, const uint32_t *__restrict__ words,
                                                         const unsigned long long *__restrict__ removed, size_t scan_len,
                                                         const unsigned long long *__restrict__ used,
                                                         const unsigned long long *__restrict__ first_block,
                                                         const unsigned long long *__restrict__ first_block_base,
                                                         unsigned long long nblocks, int16_t *__restrict__ coef,
                                                         const unsigned long long *__restrict__ rst, unsigned nrst)
{
    extern __shared__ uint4 jdec_dyn[];
    JdecShared &S = *reinterpret_cast<JdecShared *>(jdec_dyn);
    uint32_t *sw = reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(jdec_dyn) + sizeof(JdecShared));
    __shared__ uint8_t zz[64];
    if (threadIdx.x < 64) zz[threadIdx.x] = T->zz[threadIdx.x];
    jdec_load_tables(S, T);
    jdec_stage_words(sw, words);
    const unsigned long long total = 8ull * (scan_len - (*removed & 0xffffffffull));
    const unsigned long long j = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = j * JD_SUB < total;
    const unsigned long long x = live ? used[j] : 0ull;
    unsigned long long p = x >> 12;
    int k = (int)((x >> 4) & 255), bi = (int)(x & 15);
    // a block that straddles the boundary belongs to the subsequence it ends in: its earlier
    // coefficients were stored by the predecessor under the same ordinal
    CoefSink sink{coef, live ? first_block[j] + first_block_base[j >> 10] : nblocks, nblocks, zz};
    if (live)
        jdec_run<false>(S.small, &S.luts.lut[0][0], sw + JD_ROW * threadIdx.x, j * JD_SUB, p, k, bi, min(total, (j + 1) * JD_SUB), sink, rst,
                 nrst);
}

// ---------------------------------------------------------------------------
// DC prediction.  The running sums are the DC values themselves (|DC| < 2^15 even for hostile
// streams that stay decodable), so the three components share one 64-bit scan as 21-bit fields.
struct JdecGeom {
    int w, h, ncomp, hs, vs, mcux, mcuy, nbm;
    int pw[3], ph[3];          // component planes (whole MCUs)
    size_t poff[3];            // byte offsets of the planes
};

__device__ __forceinline__ long long sext21(long long v) { return (v << 43) >> 43; }

__global__ void __launch_bounds__(256) jdec_mcu_dc_kernel(const int16_t *__restrict__ coef, JdecGeom G,
                                                          unsigned long long *__restrict__ sums)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= G.mcux * G.mcuy) return;
    const int ny = G.hs * G.vs;
    long long y = 0, cb = 0, cr = 0;
    for (int b = 0; b < ny; b++) y += coef[((size_t)m * G.nbm + b) * 64];
    if (G.ncomp == 3) {
        cb = coef[((size_t)m * G.nbm + ny) * 64];
        cr = coef[((size_t)m * G.nbm + ny + 1) * 64];
    }
    sums[m] = (unsigned long long)(y + cb * (1ll << 21) + cr * (1ll << 42));
}

// jidctint.c jpeg_idct_islow, one dimension
#define JDFIX(x) ((int)((x) * 8192.0 + 0.5))
__device__ __forceinline__ void idct_1d(const int (&in)[8], int (&o)[8], int shift)
{
    int z2 = in[2], z3 = in[6];
    int z1 = (z2 + z3) * JDFIX(0.541196100);
    const int tmp2 = z1 + z3 * (-JDFIX(1.847759065)), tmp3 = z1 + z2 * JDFIX(0.765366865);
    const int tmp0 = (in[0] + in[4]) * 8192, tmp1 = (in[0] - in[4]) * 8192;
    const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    int t0 = in[7], t1 = in[5], t2 = in[3], t3 = in[1];
    z1 = t0 + t3;
    z2 = t1 + t2;
    z3 = t0 + t2;
    int z4 = t1 + t3;
    const int z5 = (z3 + z4) * JDFIX(1.175875602);
    t0 *= JDFIX(0.298631336);
    t1 *= JDFIX(2.053119869);
    t2 *= JDFIX(3.072711026);
    t3 *= JDFIX(1.501321110);
    z1 *= -JDFIX(0.899976223);
    z2 *= -JDFIX(2.562915447);
    z3 = z3 * (-JDFIX(1.961570560)) + z5;
    z4 = z4 * (-JDFIX(0.390180644)) + z5;
    t0 += z1 + z3;
    t1 += z2 + z4;
    t2 += z2 + z3;
    t3 += z1 + z4;
    const int rnd = 1 << (shift - 1);
    o[0] = (tmp10 + t3 + rnd) >> shift;
    o[7] = (tmp10 - t3 + rnd) >> shift;
    o[1] = (tmp11 + t2 + rnd) >> shift;
    o[6] = (tmp11 - t2 + rnd) >> shift;
    o[2] = (tmp12 + t1 + rnd) >> shift;
    o[5] = (tmp12 - t1 + rnd) >> shift;
    o[3] = (tmp13 + t0 + rnd) >> shift;
    o[4] = (tmp13 - t0 + rnd) >> shift;
}

// 256 threads = 32 blocks; thread = row r of its block (load, dequantise), then column r, then row r
__global__ void __launch_bounds__(256) jdec_idct_kernel(const JdecTables *__restrict__ T, const int16_t *__restrict__ coef,
                                                        const unsigned long long *__restrict__ dc_prefix,
                                                        const unsigned long long *__restrict__ dc_base, JdecGeom G,
                                                        int restart, uint8_t *__restrict__ planes)
{
    __shared__ int sm[32][8][9];
    __shared__ uint16_t sq[3][64];
    for (int i = threadIdx.x; i < 192; i += 256) (&sq[0][0])[i] = (&T->quant[0][0])[i];
    __syncthreads();
    const int lb = threadIdx.x >> 3, r = threadIdx.x & 7;
    // 32-bit indices: a baseline JPEG has at most 8192 x 8192 MCUs of at most 10 blocks (< 2^31 blocks x 64 < 2^37: the
    // coefficient index alone is 64-bit)
    const unsigned g = blockIdx.x * 32u + (unsigned)lb, nblocks = (unsigned)G.mcux * (unsigned)G.mcuy * (unsigned)G.nbm;
    const bool valid = g < nblocks;
    const unsigned gg = valid ? g : 0u;
    const int mcu = (int)(gg / (unsigned)G.nbm), bi = (int)(gg - (unsigned)mcu * (unsigned)G.nbm);
    const int ny = G.hs * G.vs, comp = bi < ny ? 0 : bi - ny + 1;
    const int4 raw = *reinterpret_cast<const int4 *>(coef + (size_t)gg * 64 + 8 * r);
    int x[8], o[8];
    x[0] = (int)(int16_t)(raw.x & 0xffff); x[1] = raw.x >> 16;
    x[2] = (int)(int16_t)(raw.y & 0xffff); x[3] = raw.y >> 16;
    x[4] = (int)(int16_t)(raw.z & 0xffff); x[5] = raw.z >> 16;
    x[6] = (int)(int16_t)(raw.w & 0xffff); x[7] = raw.w >> 16;
    {   // DC = the predictor carried into this MCU + the differences up to this block: the eight lanes of the block
        // fetch one difference each (the luma blocks in front of it in the MCU, and its own) and add them up
        const int b0 = comp == 0 ? 0 : bi;
        int part = b0 + r <= bi ? (int)coef[((size_t)mcu * G.nbm + b0 + r) * 64] : 0;
        if (r == 0) {
            // the predictors restart at 0 with every restart interval: sums relative to the interval's first MCU
            const int m0 = restart ? mcu - mcu % restart : 0;
            const long long P = (long long)(dc_prefix[mcu] + dc_base[mcu >> 10] - (restart ? dc_prefix[m0] + dc_base[m0 >> 10] : 0ull));
            const long long fy = sext21(P), P1 = (P - fy) >> 21, fcb = sext21(P1), fcr = (P1 - fcb) >> 21;
            part += (int)(comp == 0 ? fy : comp == 1 ? fcb : fcr);
        }
        part += __shfl_xor_sync(0xffffffffu, part, 1);
        part += __shfl_xor_sync(0xffffffffu, part, 2);
        part += __shfl_xor_sync(0xffffffffu, part, 4);
        if (r == 0) x[0] = (int)(int16_t)part;
    }
#pragma unroll
    for (int i = 0; i < 8; i++) sm[lb][r][i] = x[i] * (int)sq[comp][8 * r + i];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = sm[lb][i][r];
    idct_1d(x, o, 13 - 2);
#pragma unroll
    for (int i = 0; i < 8; i++) sm[lb][i][r] = o[i];        // only this thread touches column r
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = sm[lb][r][i];
    idct_1d(x, o, 13 + 2 + 3);
    if (!valid) return;
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        lo |= (uint32_t)min(255, max(0, o[i] + 128)) << (8 * i);
        hi |= (uint32_t)min(255, max(0, o[4 + i] + 128)) << (8 * i);
    }
    const int my = (int)((unsigned)mcu / (unsigned)G.mcux), mx = mcu - my * G.mcux;
    int by = my, bx = mx;
    if (comp == 0) {
        const int brow = G.hs == 2 ? bi >> 1 : bi;          // hs is 1 or 2 (parse_header)
        by = my * G.vs + brow;
        bx = mx * G.hs + (bi - brow * G.hs);
    }
    *reinterpret_cast<uint2 *>(planes + G.poff[comp] + ((size_t)by * 8 + r) * G.pw[comp] + (size_t)bx * 8) = make_uint2(lo, hi);
}

// jdsample.c + jdcolor.c: one thread = output pixels (y, 2c) and (y, 2c + 1)
__global__ void __launch_bounds__(256) jdec_color_kernel(const uint8_t *__restrict__ planes, JdecGeom G,
                                                         uint8_t *__restrict__ bgr, size_t stride)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (2 * c >= G.w) return;
    const uint8_t *Y = planes + G.poff[0] + (size_t)y * G.pw[0] + 2 * c;
    uint8_t *o = bgr + (size_t)y * stride + 6 * (size_t)c;
    const bool two = 2 * c + 1 < G.w;
    if (G.ncomp == 1) {
        o[0] = o[1] = o[2] = Y[0];
        if (two) o[3] = o[4] = o[5] = Y[1];
        return;
    }
    int cbv[2], crv[2];
    const int cw = (G.w + G.hs - 1) / G.hs, chh = (G.h + G.vs - 1) / G.vs;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const uint8_t *P = planes + G.poff[1 + k];
        const int pw = G.pw[1 + k];
        int v0, v1;
        if (G.hs == 1) {
            v0 = P[(size_t)y * pw + 2 * c];
            v1 = two ? P[(size_t)y * pw + 2 * c + 1] : 0;
        } else if (cw <= 2) {                       // jinit_upsampler: fancy only if downsampled_width > 2
            v0 = v1 = P[(size_t)(y / G.vs) * pw + c];
        } else if (G.vs == 1) {                     // h2v1_fancy_upsample
            const uint8_t *in = P + (size_t)y * pw;
            const int cur = in[c];
            v0 = c == 0 ? cur : (cur * 3 + in[c - 1] + 1) >> 2;
            v1 = c == cw - 1 ? cur : (cur * 3 + in[c + 1] + 2) >> 2;
        } else {                                    // h2v2_fancy_upsample
            const int r0 = y >> 1, r1 = min(chh - 1, max(0, (y & 1) ? r0 + 1 : r0 - 1));
            const uint8_t *a = P + (size_t)r0 * pw, *b = P + (size_t)r1 * pw;
            const int cur = a[c] * 3 + b[c];
            v0 = c == 0 ? (cur * 4 + 8) >> 4 : (cur * 3 + a[c - 1] * 3 + b[c - 1] + 8) >> 4;
            v1 = c == cw - 1 ? (cur * 4 + 7) >> 4 : (cur * 3 + a[c + 1] * 3 + b[c + 1] + 7) >> 4;
        }
        (k ? crv : cbv)[0] = v0 - 128;
        (k ? crv : cbv)[1] = v1 - 128;
    }
#pragma unroll
    for (int i = 0; i < 2; i++) {
        if (i && !two) break;
        const int yy = Y[i], cb = cbv[i], cr = crv[i];
        const int R = yy + ((91881 * cr + 32768) >> 16);
        const int Gv = yy + ((-22554 * cb + 32768 - 46802 * cr) >> 16);
        const int B = yy + ((116130 * cb + 32768) >> 16);
        o[3 * i] = (uint8_t)min(255, max(0, B));
        o[3 * i + 1] = (uint8_t)min(255, max(0, Gv));
        o[3 * i + 2] = (uint8_t)min(255, max(0, R));
    }
}

// The common case -- 4:2:0, h2v2_fancy_upsample, rows of the output a multiple of 4 bytes apart: one thread = four
// output pixels of a row (two chroma samples and their two neighbours), luma as one 32-bit load, the twelve
// output bytes as three 32-bit stores.  Same arithmetic as the general kernel above.
__global__ void __launch_bounds__(256) jdec_color420_kernel(const uint8_t *__restrict__ planes, JdecGeom G,
                                                            uint8_t *__restrict__ bgr, size_t stride)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    const int x0 = 4 * t;
    if (x0 >= G.w) return;
    const int cw = (G.w + 1) >> 1, chh = (G.h + 1) >> 1;
    const uint32_t yy4 = *reinterpret_cast<const uint32_t *>(planes + G.poff[0] + (size_t)y * G.pw[0] + x0);
    const int r0 = y >> 1, r1 = min(chh - 1, max(0, (y & 1) ? r0 + 1 : r0 - 1));
    const int c0 = 2 * t;
    int cv[2][4];                                   // [Cb, Cr][pixel], minus 128
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const int pw = G.pw[1 + k];
        const uint8_t *a = planes + G.poff[1 + k] + (size_t)r0 * pw, *b = planes + G.poff[1 + k] + (size_t)r1 * pw;
        const uint32_t a2 = *reinterpret_cast<const uint16_t *>(a + c0), b2 = *reinterpret_cast<const uint16_t *>(b + c0);
        const int im = max(c0 - 1, 0), ip = min(c0 + 2, pw - 1);
        const int sm = a[im] * 3 + b[im], s0 = (int)(a2 & 255u) * 3 + (int)(b2 & 255u), s1 = (int)(a2 >> 8) * 3 + (int)(b2 >> 8),
                  sp = a[ip] * 3 + b[ip];
        cv[k][0] = (c0 == 0 ? (s0 * 4 + 8) >> 4 : (s0 * 3 + sm + 8) >> 4) - 128;
        cv[k][1] = (c0 == cw - 1 ? (s0 * 4 + 7) >> 4 : (s0 * 3 + s1 + 7) >> 4) - 128;
        cv[k][2] = ((s1 * 3 + s0 + 8) >> 4) - 128;
        cv[k][3] = (c0 + 1 == cw - 1 ? (s1 * 4 + 7) >> 4 : (s1 * 3 + sp + 7) >> 4) - 128;
    }
    uint8_t o[12];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int yv = (int)((yy4 >> (8 * i)) & 255u), cb = cv[0][i], cr = cv[1][i];
        const int R = yv + ((91881 * cr + 32768) >> 16);
        const int Gv = yv + ((-22554 * cb + 32768 - 46802 * cr) >> 16);
        const int B = yv + ((116130 * cb + 32768) >> 16);
        o[3 * i] = (uint8_t)min(255, max(0, B));
        o[3 * i + 1] = (uint8_t)min(255, max(0, Gv));
        o[3 * i + 2] = (uint8_t)min(255, max(0, R));
    }
    uint8_t *dst = bgr + (size_t)y * stride + 3 * (size_t)x0;
    if (x0 + 4 <= G.w) {
        uint32_t *d32 = reinterpret_cast<uint32_t *>(dst);
#pragma unroll
        for (int j = 0; j < 3; j++)
            d32[j] = (uint32_t)o[4 * j] | ((uint32_t)o[4 * j + 1] << 8) | ((uint32_t)o[4 * j + 2] << 16) | ((uint32_t)o[4 * j + 3] << 24);
    } else {
        const int nbytes = 3 * (G.w - x0);
#pragma unroll
        for (int i = 0; i < 9; i++)
            if (i < nbytes) dst[i] = o[i];
    }
}

// ---------------------------------------------------------------------------
int32_t jpeg_info_host(const uint8_t *file, size_t n, int *w, int *h, int *ncomp)
{
    JdecHeader H;
    JdecTables T;
    OICS_TRY(parse_header(file, n, &H, &T));
    *w = H.w;
    *h = H.h;
    *ncomp = H.ncomp;
    return OICS_OK;
}

// One decode in flight: its stream, scratch and the pointers of the current file.  The batch driver
// keeps several (one per card of a sub-batch, each on its own stream): the synchronisation rounds of
// a card are latency-bound chains that leave the GPU almost empty, so cards overlap.
struct JdecSlot {
    cudaStream_t st = nullptr;
    bool own_stream = false;
    DevBuf file, words, tables, state, coef, planes;
    PinnedBuf pin;            // JdecTables staging + the flags read-back
    JdecHeader H;
    JdecGeom G;
    size_t nsub = 0, nblocks = 0, nmcu = 0, nsum_s = 0, nsum_m = 0;
    unsigned long long *used = nullptr, *out = nullptr, *nblk = nullptr, *sums_s = nullptr, *mcu = nullptr, *sums_m = nullptr;
    unsigned long long *mkey = nullptr, *mout = nullptr, *removed = nullptr, *rst = nullptr;
    unsigned nrst = 0;
    unsigned *mnb = nullptr, *mcnt = nullptr;
    int *flags = nullptr;
    int rounds = 0;           // rounds of the batch whose flags are being read back
    bool in_flight = false;   // begun, and the host has not yet seen its chains converge
    cudaEvent_t ev = nullptr;
    cudaEvent_t done_ev()
    {
        if (!ev) cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
        return ev;
    }
    // read-back block: the four counters at `removed`, then the 32 flags
    unsigned long long *hcounters() { return reinterpret_cast<unsigned long long *>(pin.as<uint8_t>() + sizeof(JdecTables) + 64); }
    int *hflags() { return reinterpret_cast<int *>(hcounters() + 4); }
};

// Rounds of the chunk walk (jdec_walk_kernel): `first` = ordinal of the first one (its slot in the flags).
static int32_t jdec_rounds(oics_ctx *ctx, JdecSlot &S, int first, int rounds, bool from_guess)
{
    const unsigned nchunk = (unsigned)((S.nsub + JD_CHUNK - 1) / JD_CHUNK), gw = (nchunk + JD_WALK_WARPS - 1) / JD_WALK_WARPS;
    static const char *const kRoundNames[] = {"jdec_walk_kernel<1>", "jdec_walk_kernel<2>", "jdec_walk_kernel<3>", "jdec_walk_kernel<4+>"};
    for (int r = 0; r < rounds; r++) {
        ProfScope ps(ctx, kRoundNames[std::min(3, first - 2 + r)], 0.0, S.st);
        jdec_walk_kernel<<<gw, 32 * JD_WALK_WARPS, 0, S.st>>>(S.tables.as<JdecTables>(), S.words.as<uint32_t>(), S.removed, S.H.scan_len,
                                                              S.used, S.out, S.nblk, first + r, from_guess && r == 0 ? 1 : 0, S.flags,
                                                              S.rst, S.nrst);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += rounds;
    S.rounds = first + rounds - 1;
    OICS_CUDA_TRY(cudaMemcpyAsync(S.hcounters(), S.removed, 4 * sizeof(unsigned long long) + 32 * sizeof(int), cudaMemcpyDeviceToHost,
                                  S.st));
    return OICS_OK;
}

// parse, upload, unstuff, first rounds of the synchronisation -- everything enqueued, nothing awaited
static int32_t jdec_begin(oics_ctx *ctx, JdecSlot &S, const uint8_t *file, size_t n)
{
    cudaStream_t st = S.st;
    if (!ctx->attr_jdec) {   // per context = per device
        OICS_CUDA_TRY(cudaFuncSetAttribute(jdec_write_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)JD_DYN_SMEM));
        ctx->attr_jdec = true;
    }
    OICS_TRY(S.pin.ensure(sizeof(JdecTables) + 64 + 4 * sizeof(unsigned long long) + 32 * sizeof(int)));
    // the pinned block (tables out, counters back) is reused: a previous file that was begun but never polled
    // to the end (refused or abandoned) may still have copies in flight -- otherwise the last poll has waited
    if (S.in_flight) OICS_CUDA_TRY(cudaStreamSynchronize(st));
    S.in_flight = true;
    JdecTables *T = S.pin.as<JdecTables>();
    S.H = JdecHeader();                         // the slot is reused: nothing of the previous file may survive
    OICS_TRY(parse_header(file, n, &S.H, T));
    const JdecHeader &H = S.H;
    JdecGeom &G = S.G;
    G.w = H.w; G.h = H.h; G.ncomp = H.ncomp; G.hs = H.hs; G.vs = H.vs;
    G.mcux = cdiv(H.w, 8 * H.hs); G.mcuy = cdiv(H.h, 8 * H.vs);
    G.nbm = T->nbm;
    size_t off = 0;
    for (int c = 0; c < 3; c++) {
        G.pw[c] = c < H.ncomp ? G.mcux * 8 * (c ? 1 : H.hs) : 0;
        G.ph[c] = c < H.ncomp ? G.mcuy * 8 * (c ? 1 : H.vs) : 0;
        G.poff[c] = off;
        off += ((size_t)G.pw[c] * G.ph[c] + 255) & ~(size_t)255;
    }
    S.nmcu = (size_t)G.mcux * G.mcuy;
    S.nblocks = S.nmcu * G.nbm;
    const size_t slen = H.scan_len, nchunks = (slen + 31) / 32, nsub = (slen * 8 + JD_SUB - 1) / JD_SUB;
    const size_t nsum_c = (nchunks + 1023) / 1024 + 8;
    S.nsub = nsub;
    S.nsum_s = (nsub + 1023) / 1024 + 8;
    S.nsum_m = (S.nmcu + 1023) / 1024 + 8;
    OICS_TRY(S.file.ensure(slen + 64));
    OICS_TRY(S.words.ensure(slen + 8704));   // the last CTA stages a whole window: zero padding behind the stream
    OICS_TRY(S.tables.ensure(sizeof(JdecTables)));
    S.nrst = H.restart ? (unsigned)((S.nmcu + H.restart - 1) / H.restart - 1) : 0u;
    const size_t n_state = nchunks + nsum_c + 3 * nsub + S.nsum_s + S.nmcu + S.nsum_m + 2 * nsub * JD_M + 64 + S.nrst + 1;
    OICS_TRY(S.state.ensure(n_state * sizeof(unsigned long long) + (nsub * (JD_M + 1) + 64) * sizeof(unsigned)));
    OICS_TRY(S.coef.ensure(S.nblocks * 64 * sizeof(int16_t) + 256));
    OICS_TRY(S.planes.ensure(off + 256));
    unsigned long long *cnt = S.state.as<unsigned long long>(), *sums_c = cnt + nchunks;
    S.used = sums_c + nsum_c;
    S.out = S.used + nsub;
    S.nblk = S.out + nsub;
    S.sums_s = S.nblk + nsub;
    S.mcu = S.sums_s + S.nsum_s;
    S.sums_m = S.mcu + S.nmcu;
    S.mkey = S.sums_m + S.nsum_m;
    S.mout = S.mkey + nsub * JD_M;
    S.removed = S.mout + nsub * JD_M;                       // [0] removed bytes, [1] blocks, [2] DC total, [4..] flags
    S.flags = reinterpret_cast<int *>(S.removed + 4);       // [0]: a marker inside the scan; [r]: round r changed something
    S.rst = S.removed + 64;                                 // restart boundaries (bit positions); all ones = none
    S.mnb = reinterpret_cast<unsigned *>(S.rst + S.nrst + 1);
    S.mcnt = S.mnb + nsub * JD_M;
    uint8_t *d_scan = S.file.as<uint8_t>();
    uint32_t *words = S.words.as<uint32_t>();
    OICS_CUDA_TRY(cudaMemcpyAsync(S.tables.p, T, sizeof(JdecTables), cudaMemcpyHostToDevice, st));
    OICS_CUDA_TRY(cudaMemcpyAsync(d_scan, file + H.scan_off, slen, cudaMemcpyHostToDevice, st));
    OICS_CUDA_TRY(cudaMemsetAsync(words, 0, slen + 8704, st));
    OICS_CUDA_TRY(cudaMemsetAsync(S.nblk, 0, nsub * sizeof(unsigned long long), st));
    OICS_CUDA_TRY(cudaMemsetAsync(S.removed, 0, 64 * sizeof(unsigned long long), st));
    OICS_CUDA_TRY(cudaMemsetAsync(S.rst, 0xFF, ((size_t)S.nrst + 1) * sizeof(unsigned long long), st));
    OICS_CUDA_TRY(cudaMemsetAsync(S.coef.p, 0, S.nblocks * 64 * sizeof(int16_t), st));
    const unsigned gc = (unsigned)((nchunks + 255) / 256), gs = (unsigned)((nsub + 127) / 128);
    jdec_unstuff_count_kernel<<<gc, 256, 0, st>>>(d_scan, slen, H.restart, cnt, S.flags);
    OICS_TRY(exclusive_scan(ctx, st, cnt, nchunks, sums_c, S.removed, nullptr));
    jdec_unstuff_write_kernel<<<gc, 256, 0, st>>>(d_scan, slen, H.restart, cnt, sums_c, reinterpret_cast<uint8_t *>(words), S.rst, S.nrst);
    // The synchronisation: round 0, every subsequence from the guess (its first bit, k = 0, block 0); then the wide
    // kernel, every thread fed its predecessor's exit until nothing changes inside its CTA (several wavefronts per
    // CTA advance side by side); then the chunk walks carry the states across the CTA borders, until one of them
    // decodes nothing -- the host looks at the flags after 3, then every 4.  OICS_JD_WIDE: iterations of the wide
    // kernel (tuning; 0 = walks only, measured slower on quality-100 cards: one wavefront per chunk).
    static const int wide = [] { const char *e = getenv("OICS_JD_WIDE"); return e ? std::max(0, std::min(128, atoi(e))) : -1; }();
    const int n_wide = wide >= 0 ? wide : 128;
    if (n_wide > 0) {
        {
            ProfScope ps(ctx, "jdec_sync_kernel<round0>", (double)slen, st);
            jdec_sync_kernel<<<gs, 128, JD_SYNC_SMEM, st>>>(S.tables.as<JdecTables>(), words, S.removed, slen, S.used, S.out, S.nblk, S.mkey, S.mout,
                                                  S.mnb, S.mcnt, 0, 1, S.flags, S.rst, S.nrst);
        }
        {
            ProfScope ps(ctx, "jdec_sync_kernel<wide>", 0.0, st);
            jdec_sync_kernel<<<gs, 128, JD_SYNC_SMEM, st>>>(S.tables.as<JdecTables>(), words, S.removed, slen, S.used, S.out, S.nblk, S.mkey, S.mout,
                                                  S.mnb, S.mcnt, 1, n_wide, S.flags, S.rst, S.nrst);
        }
        ctx->launches += 2;
    }
    ctx->launches += 2;
    return jdec_rounds(ctx, S, 2, 3, n_wide == 0);
}

// wait for the rounds in flight; *done = converged, else more rounds are enqueued
static int32_t jdec_poll(oics_ctx *ctx, JdecSlot &S, bool *done)
{
    OICS_CUDA_TRY(cudaStreamSynchronize(S.st));
    const int *hf = S.hflags();
    if (hf[0] & 1) {
        set_error("jpeg: a marker inside the scan (truncated file, or restart markers without DRI) is not covered by the CUDA decoder");
        return OICS_ERR_UNSUPPORTED;
    }
    if ((unsigned)(S.hcounters()[0] >> 32) != S.nrst) {
        set_error("jpeg: %u restart markers in the scan, the header announces %u", (unsigned)(S.hcounters()[0] >> 32), S.nrst);
        return OICS_ERR_INVALID_ARG;
    }
    *done = !hf[S.rounds];
    if (*done) {
        S.in_flight = false;
        return OICS_OK;
    }
    OICS_CUDA_TRY(cudaMemsetAsync(S.flags + 1, 0, 31 * sizeof(int), S.st));
    return jdec_rounds(ctx, S, 5, 4, false);
}

// coefficients, DC prediction, IDCT, upsampling + colour -> d_bgr (enqueued on the slot's stream)
static int32_t jdec_finish(oics_ctx *ctx, JdecSlot &S, uint8_t *d_bgr, size_t stride)
{
    cudaStream_t st = S.st;
    const unsigned gs = (unsigned)((S.nsub + 127) / 128);
    int16_t *coef = S.coef.as<int16_t>();
    OICS_TRY(exclusive_scan(ctx, st, S.nblk, S.nsub, S.sums_s, S.removed + 1, nullptr));
    {
        ProfScope ps(ctx, "jdec_write_kernel", (double)S.H.scan_len + 128.0 * (double)S.nblocks, st);
        jdec_write_kernel<<<gs, 128, JD_DYN_SMEM, st>>>(S.tables.as<JdecTables>(), S.words.as<uint32_t>(), S.removed, S.H.scan_len, S.used, S.nblk,
                                               S.sums_s, (unsigned long long)S.nblocks, coef, S.rst, S.nrst);
    }
    jdec_mcu_dc_kernel<<<(unsigned)((S.nmcu + 255) / 256), 256, 0, st>>>(coef, S.G, S.mcu);
    OICS_TRY(exclusive_scan(ctx, st, S.mcu, S.nmcu, S.sums_m, S.removed + 2, nullptr));
    {
        ProfScope ps(ctx, "jdec_idct_kernel", (128.0 + 64.0) * (double)S.nblocks, st);
        jdec_idct_kernel<<<(unsigned)((S.nblocks + 31) / 32), 256, 0, st>>>(S.tables.as<JdecTables>(), coef, S.mcu, S.sums_m, S.G, S.H.restart, S.planes.as<uint8_t>());
    }
    {
        ProfScope ps(ctx, "jdec_color_kernel", 64.0 * (double)S.nblocks + 3.0 * (double)S.H.w * (double)S.H.h, st);
        const bool fast420 = S.G.ncomp == 3 && S.G.hs == 2 && S.G.vs == 2 && (S.H.w + 1) / 2 > 2 && stride % 4 == 0 &&
                             (reinterpret_cast<uintptr_t>(d_bgr) & 3) == 0;
        if (fast420)
            jdec_color420_kernel<<<dim3((unsigned)cdiv(cdiv(S.H.w, 4), 256), (unsigned)S.H.h), 256, 0, st>>>(S.planes.as<uint8_t>(), S.G, d_bgr, stride);
        else
            jdec_color_kernel<<<dim3((unsigned)cdiv(cdiv(S.H.w, 2), 256), (unsigned)S.H.h), 256, 0, st>>>(S.planes.as<uint8_t>(), S.G, d_bgr, stride);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 4;
    return OICS_OK;
}

static JdecSlot *jdec_slot(oics_ctx *ctx, int i)
{
    if (!ctx->jd_slots[i]) {
        JdecSlot *S = new JdecSlot();
        if (i == 0) {
            S->st = ctx->stream;
        } else {
            if (cudaStreamCreateWithFlags(&S->st, cudaStreamNonBlocking) != cudaSuccess) { delete S; return nullptr; }
            S->own_stream = true;
        }
        ctx->jd_slots[i] = S;
    }
    JdecSlot *S = static_cast<JdecSlot *>(ctx->jd_slots[i]);
    if (i == 0) S->st = ctx->stream;
    return S;
}

void jpeg_decode_release(oics_ctx *ctx)
{
    for (void *&p : ctx->jd_slots) {
        JdecSlot *S = static_cast<JdecSlot *>(p);
        if (!S) continue;
        if (S->own_stream) { cudaStreamSynchronize(S->st); cudaStreamDestroy(S->st); }
        if (S->ev) cudaEventDestroy(S->ev);
        for (DevBuf *b : {&S->file, &S->words, &S->tables, &S->state, &S->coef, &S->planes}) b->release();
        S->pin.release();
        delete S;
        p = nullptr;
    }
}

// file bytes on the HOST -> BGR image on the DEVICE (rows `stride` bytes apart), on ctx->stream.  The
// entropy-coded segment and one table block are what crosses PCIe.
int32_t jpeg_decode_device(oics_ctx *ctx, const uint8_t *file, size_t n, uint8_t *d_bgr, size_t stride, int *w_out, int *h_out)
{
    JdecSlot *S = jdec_slot(ctx, 0);
    if (!S) { set_error("jpeg: no decode slot"); return OICS_ERR_CUDA; }
    OICS_TRY(jdec_begin(ctx, *S, file, n));
    *w_out = S->H.w;
    *h_out = S->H.h;
    if (stride < (size_t)S->H.w * 3) { set_error("jpeg: stride %zu is below 3 * width", stride); return OICS_ERR_INVALID_ARG; }
    for (bool done = false; !done;) OICS_TRY(jdec_poll(ctx, *S, &done));
    return jdec_finish(ctx, *S, d_bgr, stride);
}

// A sub-batch of files (host) -> BGR cards on the device, every card on its own stream, in two calls so
// that the caller can put other work between them:
//   begin   parse, upload, unstuff and the first synchronisation rounds of every file -- enqueued only;
//           statuses[i] per file (a file of another size than w x h is refused)
//   finish  wait for the chains (normally long converged by now), then coefficients .. colour into
//           d_bgr + i * card_bytes; the slot streams wait for `free_ev` (may be null) before writing
//           there, and `ready` is recorded on `join` once all cards are done
int32_t jpeg_decode_batch_begin(oics_ctx *ctx, int slot0, const uint8_t *const *files, const size_t *sizes, int n_files, int w, int h,
                                int32_t *statuses)
{
    if (slot0 < 1 || slot0 + n_files > OICS_JD_SLOTS) { set_error("jpeg: decode slots %d..%d of %d", slot0, slot0 + n_files - 1, OICS_JD_SLOTS); return OICS_ERR_INVALID_ARG; }
    for (int i = 0; i < n_files; i++) {
        statuses[i] = OICS_OK;
        JdecSlot *S = jdec_slot(ctx, slot0 + i);
        if (!S) { set_error("jpeg: no decode slot"); return OICS_ERR_CUDA; }
        S->rounds = 0;
        int32_t rc = files[i] ? jdec_begin(ctx, *S, files[i], sizes[i]) : OICS_ERR_INVALID_ARG;
        if (rc == OICS_OK && (S->H.w != w || S->H.h != h)) {
            set_error("jpeg: file %d is %dx%d, the batch is %dx%d", i, S->H.w, S->H.h, w, h);
            rc = OICS_ERR_INVALID_ARG;
        }
        if (rc == OICS_ERR_CUDA) return rc;
        statuses[i] = rc;
    }
    return OICS_OK;
}

int32_t jpeg_decode_batch_finish(oics_ctx *ctx, int slot0, int n_files, int w, uint8_t *d_bgr, size_t card_bytes, cudaEvent_t free_ev,
                                 cudaStream_t join, cudaEvent_t ready, int32_t *statuses)
{
    std::vector<JdecSlot *> slots((size_t)n_files, nullptr);
    for (int i = 0; i < n_files; i++) slots[i] = jdec_slot(ctx, slot0 + i);
    for (bool all = false; !all;) {
        all = true;
        for (int i = 0; i < n_files; i++) {
            if (statuses[i] != OICS_OK || slots[i]->rounds == 0) continue;
            bool done = false;
            const int32_t rc = jdec_poll(ctx, *slots[i], &done);
            if (rc == OICS_ERR_CUDA) return rc;
            if (rc != OICS_OK) statuses[i] = rc;
            else if (done) slots[i]->rounds = 0;
            else all = false;
        }
    }
    for (int i = 0; i < n_files; i++) {
        JdecSlot &S = *slots[i];
        if (statuses[i] == OICS_OK) {
            if (free_ev) OICS_CUDA_TRY(cudaStreamWaitEvent(S.st, free_ev, 0));
            const int32_t rc = jdec_finish(ctx, S, d_bgr + (size_t)i * card_bytes, (size_t)w * 3);
            if (rc == OICS_ERR_CUDA) return rc;
            statuses[i] = rc;
        }
        OICS_CUDA_TRY(cudaEventRecord(S.done_ev(), S.st));
        OICS_CUDA_TRY(cudaStreamWaitEvent(join, S.done_ev(), 0));
    }
    if (ready) OICS_CUDA_TRY(cudaEventRecord(ready, join));
    return OICS_OK;
}

}  // namespace oics
