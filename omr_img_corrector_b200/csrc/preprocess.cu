// P0 / P0': pre-processing in front of the sweep.
//   OMR flavour        (omr.rs:60-139):       gray -> erode(3x3 cross) x3 -> INTER_AREA -> threshold
//   projection flavour (projection.rs:24-32): INTER_AREA on the colour image -> gray -> threshold
// All integer-exact restatements of the OpenCV calls:
//   gray      (c0*9798 + c1*19235 + c2*3735 + 2^14) >> 15   (COLOR_RGB2GRAY on BGR bytes)
//   erode xk  min over the L1 ball of radius k, outside the image ignored
//   area      integer factors: box sum, (float)sum * (1.f/area), round half even
//             (2x2: (sum+2)>>2); fractional: computeResizeAreaTab's f32 taps,
//             accumulated in OpenCV's order without FMA
//   threshold v > 127 ? 255 : 0
// gray + erode (+ threshold when no resize follows) are one kernel: the BGR tile and
// its halo are converted once into shared memory and the diamond minimum is taken
// there, so the card is read from HBM once and written once.
#include "common.h"

namespace oics {

static constexpr int GE_TX = 64, GE_TY = 16, GE_RMAX = 8;

__device__ __forceinline__ uint8_t gray_px(const uint8_t *p)
{
    return (uint8_t)((p[0] * 9798 + p[1] * 19235 + p[2] * 3735 + 16384) >> 15);
}

// src: CH = 3 -> BGR converted to gray on the fly; CH = 1 -> already gray.
template <int CH>
__global__ void __launch_bounds__(256) gray_erode_kernel(const uint8_t *__restrict__ src, int w, int h, size_t stride,
                                                         int radius, int do_threshold, uint8_t *__restrict__ dst,
                                                         size_t dstride)
{
    __shared__ uint8_t s[(GE_TY + 2 * GE_RMAX)][GE_TX + 2 * GE_RMAX + 4];
    const int bx = blockIdx.x * GE_TX, by = blockIdx.y * GE_TY;
    const int sw = GE_TX + 2 * radius, sh = GE_TY + 2 * radius;
    for (int i = threadIdx.x; i < sw * sh; i += blockDim.x) {
        const int ly = i / sw, lx = i - ly * sw;
        const int gx = bx + lx - radius, gy = by + ly - radius;
        uint8_t v = 255;  // morphologyDefaultBorderValue for erode: never the minimum
        if (gx >= 0 && gx < w && gy >= 0 && gy < h) {
            const uint8_t *p = src + (size_t)gy * stride + (size_t)gx * CH;
            v = CH == 3 ? gray_px(p) : p[0];
        }
        s[ly][lx] = v;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < GE_TX * GE_TY; i += blockDim.x) {
        const int ly = i / GE_TX, lx = i - ly * GE_TX;
        const int gx = bx + lx, gy = by + ly;
        if (gx >= w || gy >= h) continue;
        int m = 255;
        for (int dy = -radius; dy <= radius; dy++) {
            const int span = radius - abs(dy);
            for (int dx = -span; dx <= span; dx++) m = min(m, (int)s[ly + radius + dy][lx + radius + dx]);
        }
        if (do_threshold) m = m > 127 ? 255 : 0;
        dst[(size_t)gy * dstride + gx] = (uint8_t)m;
    }
}

// ---------------------------------------------------------------------------
// Fast path of the same operation for radius 0 or 3 (the two the reference uses): a CTA
// produces a 128 x 64 tile of one card of a batch.  The BGR rows of the tile and its halo are
// fetched as aligned 16-byte chunks straight into shared memory (cp.async, every row in
// flight at once, any row alignment), converted to gray once with two dp4a per pixel
// (weights split into 8-bit halves: 9798 = 38*256 + 70, 19235 = 75*256 + 35, 3735 = 14*256 + 151),
// and the L1 diamond is taken four pixels at a time with byte-SIMD minima:
//     H_k[r] = min over |dx| <= k of gray[r][x + dx]            (k = 1, 2, 3; 2 shifts + 2 vmin each)
//     out[r] = min(H3[r], H2[r +- 1], H1[r +- 2], gray[r +- 3])
static constexpr int GF_TW = 128, GF_TH = 64, GF_WW = GF_TW / 4;   // tile in pixels / words
static constexpr int GF_CHUNKS = 27;                               // 16-byte chunks covering 15 + 412 bytes
static constexpr int GF_RAWW = GF_CHUNKS * 4;                      // words per staged BGR row

__device__ __forceinline__ uint32_t gray4(uint32_t a, uint32_t b, uint32_t c)
{
    // a, b, c: 12 BGR bytes of four pixels; the 4th byte of each dp4a operand has weight 0
    const uint32_t HI = 0x000E4B26u, LO = 0x00972346u;
    const uint32_t p1 = __byte_perm(a, b, 0x0543), p2 = __byte_perm(b, c, 0x0432), p3 = c >> 8;
    const uint32_t g0 = (__dp4a(a, LO, 16384u) + (__dp4a(a, HI, 0u) << 8)) >> 15;
    const uint32_t g1 = (__dp4a(p1, LO, 16384u) + (__dp4a(p1, HI, 0u) << 8)) >> 15;
    const uint32_t g2 = (__dp4a(p2, LO, 16384u) + (__dp4a(p2, HI, 0u) << 8)) >> 15;
    const uint32_t g3 = (__dp4a(p3, LO, 16384u) + (__dp4a(p3, HI, 0u) << 8)) >> 15;
    return g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
}

template <int R>   // R = 0 (gray [+ threshold]) or 3
__global__ void __launch_bounds__(256) gray_erode_fast_kernel(const uint8_t *__restrict__ src0, size_t src_card, int w,
                                                              int h, size_t stride, int do_threshold,
                                                              uint8_t *__restrict__ dst0, size_t dst_card,
                                                              size_t dstride)
{
    constexpr int ROWS = GF_TH + 2 * R;
    __shared__ __align__(16) uint32_t raw[ROWS * GF_RAWW];   // BGR chunks; re-used for H1..H3 after the conversion
    __shared__ uint32_t G[ROWS][GF_WW + 2];                   // gray, one halo word per side
    __shared__ int sMis[ROWS];                                // byte offset of each row segment in its first chunk
    const uint8_t *src = src0 + (size_t)blockIdx.z * src_card;
    uint8_t *dst = dst0 + (size_t)blockIdx.z * dst_card;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bx = blockIdx.x * GF_TW, by = blockIdx.y * GF_TH;
    // phase 0: the aligned 16-byte chunks that cover each row segment, a warp per row.  A chunk
    // is read iff it overlaps the valid bytes of its row (same page as a valid byte: never faults).
    const int seg0 = (bx - 4) * 3, row_bytes = w * 3;
    for (int r = warp; r < ROWS; r += 8) {
        const int gy = by - R + r;
        const bool rin = gy >= 0 && gy < h;
        const uint8_t *a0 = src + (size_t)(rin ? gy : 0) * stride + seg0;
        const int mis = (int)(reinterpret_cast<uintptr_t>(a0) & 15);
        if (lane == 0) sMis[r] = mis;
        if (lane < GF_CHUNKS) {
            const int off = seg0 - mis + 16 * lane;    // first byte of the chunk, relative to the row
            const bool in = rin && off + 16 > 0 && off < row_bytes;
            const unsigned d = (unsigned)__cvta_generic_to_shared(&raw[r * GF_RAWW + 4 * lane]);
            const uint8_t *g = a0 - mis + (in ? 16 * lane : 0);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(g), "r"(in ? 16 : 0) : "memory");
        }
    }
    asm volatile("cp.async.wait_all;\n" ::: "memory");
    __syncthreads();
    // phase 1: gray, four pixels per thread; outside the image = 255 (never the minimum)
    const bool interior = bx >= 4 && bx + GF_TW + 4 <= w && by >= R && by + GF_TH + R <= h;
    for (int i = tid; i < ROWS * (GF_WW + 2); i += 256) {
        const int r = i / (GF_WW + 2), wc = i - r * (GF_WW + 2);
        const int mis = sMis[r];
        const uint32_t *rw = raw + r * GF_RAWW + (mis >> 2) + 3 * wc;
        const int sh = 8 * (mis & 3);
        uint32_t g = gray4(__funnelshift_r(rw[0], rw[1], sh), __funnelshift_r(rw[1], rw[2], sh),
                           __funnelshift_r(rw[2], rw[3], sh));
        if (!interior) {
            const int gy = by - R + r, gx = bx - 4 + 4 * wc;
            const bool rin = gy >= 0 && gy < h;
            uint32_t keep = 0;
            if (rin && gx >= 0 && gx < w) keep |= 0x000000FFu;
            if (rin && gx + 1 >= 0 && gx + 1 < w) keep |= 0x0000FF00u;
            if (rin && gx + 2 >= 0 && gx + 2 < w) keep |= 0x00FF0000u;
            if (rin && gx + 3 >= 0 && gx + 3 < w) keep |= 0xFF000000u;
            g |= ~keep;
        }
        G[r][wc] = g;
    }
    __syncthreads();
    uint32_t *H1 = raw, *H2 = raw + ROWS * GF_WW, *H3 = raw + 2 * ROWS * GF_WW;
    if (R == 3) {
        // phase 2a: horizontal minima of every staged row
        for (int i = tid; i < ROWS * GF_WW; i += 256) {
            const int r = i >> 5, wc = (i & 31) + 1;
            const uint32_t p = G[r][wc - 1], c = G[r][wc], n = G[r][wc + 1];
            uint32_t m = __vminu4(__vminu4(c, __funnelshift_l(p, c, 8)), __funnelshift_r(c, n, 8));
            H1[i] = m;
            m = __vminu4(__vminu4(m, __funnelshift_l(p, c, 16)), __funnelshift_r(c, n, 16));
            H2[i] = m;
            m = __vminu4(__vminu4(m, __funnelshift_l(p, c, 24)), __funnelshift_r(c, n, 24));
            H3[i] = m;
        }
        __syncthreads();
    }
    // phase 2b: vertical combination, threshold, store
    const bool word_ok = ((reinterpret_cast<uintptr_t>(dst) | dstride) & 3) == 0;
    for (int i = tid; i < GF_TH * GF_WW; i += 256) {
        const int r = i >> 5, wc = i & 31;
        const int gy = by + r, gx = bx + 4 * wc;
        if (gy >= h || gx >= w) continue;
        uint32_t m;
        if (R == 3) {
            const int q = r + 3;   // staged row of this output row
            m = __vminu4(H3[q * GF_WW + wc], __vminu4(H2[(q - 1) * GF_WW + wc], H2[(q + 1) * GF_WW + wc]));
            m = __vminu4(m, __vminu4(H1[(q - 2) * GF_WW + wc], H1[(q + 2) * GF_WW + wc]));
            m = __vminu4(m, __vminu4(G[q - 3][wc + 1], G[q + 3][wc + 1]));
        } else {
            m = G[r][wc + 1];
        }
        if (do_threshold) m = __vcmpgtu4(m, 0x7f7f7f7fu);
        uint8_t *o = dst + (size_t)gy * dstride + gx;
        if (word_ok && gx + 3 < w) {
            *reinterpret_cast<uint32_t *>(o) = m;
        } else {
            for (int k = 0; k < 4 && gx + k < w; k++) o[k] = (uint8_t)(m >> (8 * k));
        }
    }
}

// ---------------------------------------------------------------------------
// gray -> erode -> threshold when nothing follows the threshold (the BASELINE operating point:
// no resize).  threshold(min(...)) = min(threshold(...)), so the diamond minimum is taken on
// ONE BIT per pixel (gray > 127): a warp converts one 128-pixel row segment (4 pixels per
// lane, two dp4a each), four ballots turn it into four interleaved bit words (bit l of word k
// = pixel 4 l + k), the horizontal partial ANDs A1..A3 of the L1 diamond cost 12 LOP3 per row
// and the vertical combination 3 LOP3 per word.  A CTA produces 120 x 64 output pixels (the
// 128 staged pixels include a 4-pixel halo per side).
static constexpr int GB_TW = 120, GB_TH = 64;
static constexpr int GB_CHUNKS = 25, GB_RAWW = GB_CHUNKS * 4;   // 15 + 384 bytes of a staged BGR row

template <int R>   // R = 0 (gray + threshold) or 3
__global__ void __launch_bounds__(256) gray_erode_bits_kernel(const uint8_t *__restrict__ src0, size_t src_card, int w,
                                                              int h, size_t stride, uint8_t *__restrict__ dst0,
                                                              size_t dst_card, size_t dstride)
{
    constexpr int ROWS = GB_TH + 2 * R;
    __shared__ __align__(16) uint32_t raw[ROWS * GB_RAWW];
    __shared__ __align__(16) uint32_t Bw[ROWS][4], A1[ROWS][4], A2[ROWS][4], A3[ROWS][4];
    __shared__ uint32_t O[GB_TH][4];
    const uint8_t *src = src0 + (size_t)blockIdx.z * src_card;
    uint8_t *dst = dst0 + (size_t)blockIdx.z * dst_card;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bx = blockIdx.x * GB_TW, by = blockIdx.y * GB_TH;
    const int seg0 = (bx - 4) * 3, row_bytes = w * 3;
    // phase 0: aligned 16-byte chunks covering each row segment, a warp per row, all in flight
    for (int r = warp; r < ROWS; r += 8) {
        const int gy = by - R + r;
        const bool rin = gy >= 0 && gy < h;
        const uint8_t *a0 = src + (size_t)(rin ? gy : 0) * stride + seg0;
        const int mis = (int)(reinterpret_cast<uintptr_t>(a0) & 15);
        if (lane < GB_CHUNKS) {
            const int off = seg0 - mis + 16 * lane;
            const bool in = rin && off + 16 > 0 && off < row_bytes;
            const unsigned d = (unsigned)__cvta_generic_to_shared(&raw[r * GB_RAWW + 4 * lane]);
            const uint8_t *g = a0 - mis + (in ? 16 * lane : 0);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(g), "r"(in ? 16 : 0) : "memory");
        }
    }
    asm volatile("cp.async.wait_all;\n" ::: "memory");
    __syncthreads();
    // phase 1: gray > 127 as interleaved bit words + horizontal partial ANDs, a warp per row
    const bool interior = bx >= 4 && bx + GB_TW + 4 <= w && by >= R && by + GB_TH + R <= h;
    for (int r = warp; r < ROWS; r += 8) {
        const int gy = by - R + r;
        const bool rin = gy >= 0 && gy < h;
        const int mis = (int)(reinterpret_cast<uintptr_t>(src + (size_t)(rin ? gy : 0) * stride + seg0) & 15);
        const uint32_t *rw = raw + r * GB_RAWW + (mis >> 2) + 3 * lane;
        const int sh = 8 * (mis & 3);
        // gray > 127  <=>  c0*9798 + c1*19235 + c2*3735 + 16384 >= 128 << 15: compare the dot
        // product, the gray byte itself is never needed
        const uint32_t a = __funnelshift_r(rw[0], rw[1], sh), b = __funnelshift_r(rw[1], rw[2], sh),
                       c = __funnelshift_r(rw[2], rw[3], sh);
        const uint32_t HI = 0x000E4B26u, LO = 0x00972346u, T = (128u << 15) - 16384u;
        const uint32_t q1 = __byte_perm(a, b, 0x0543), q2 = __byte_perm(b, c, 0x0432), q3 = c >> 8;
        bool p0 = __dp4a(a, LO, 0u) + (__dp4a(a, HI, 0u) << 8) >= T;
        bool p1 = __dp4a(q1, LO, 0u) + (__dp4a(q1, HI, 0u) << 8) >= T;
        bool p2 = __dp4a(q2, LO, 0u) + (__dp4a(q2, HI, 0u) << 8) >= T;
        bool p3 = __dp4a(q3, LO, 0u) + (__dp4a(q3, HI, 0u) << 8) >= T;
        if (!interior) {                               // outside the image: never the minimum
            const int gx = bx - 4 + 4 * lane;
            p0 = p0 || !(rin && gx >= 0 && gx < w);
            p1 = p1 || !(rin && gx + 1 >= 0 && gx + 1 < w);
            p2 = p2 || !(rin && gx + 2 >= 0 && gx + 2 < w);
            p3 = p3 || !(rin && gx + 3 >= 0 && gx + 3 < w);
        }
        const uint32_t w0 = __ballot_sync(0xffffffffu, p0), w1 = __ballot_sync(0xffffffffu, p1);
        const uint32_t w2 = __ballot_sync(0xffffffffu, p2), w3 = __ballot_sync(0xffffffffu, p3);
        if (R == 0) {
            if (lane == 0) *reinterpret_cast<uint4 *>(&Bw[r][0]) = make_uint4(w0, w1, w2, w3);
        } else {
            // pixel x + s lives in word (k + s) mod 4, lane shifted by floor((k + s) / 4)
            const uint32_t u0 = w0 >> 1, u1 = w1 >> 1, u2 = w2 >> 1;   // next lane's pixels 0..2
            const uint32_t d1 = w1 << 1, d2 = w2 << 1, d3 = w3 << 1;   // previous lane's pixels 1..3
            const uint32_t a10 = w0 & w1 & d3, a11 = w1 & w2 & w0, a12 = w2 & w3 & w1, a13 = w3 & u0 & w2;
            const uint32_t a20 = a10 & w2 & d2, a21 = a11 & w3 & d3, a22 = a12 & u0 & w0, a23 = a13 & u1 & w1;
            const uint32_t a30 = a20 & w3 & d1, a31 = a21 & u0 & d2, a32 = a22 & u1 & d3, a33 = a23 & u2 & w0;
            if (lane == 0) {
                *reinterpret_cast<uint4 *>(&Bw[r][0]) = make_uint4(w0, w1, w2, w3);
                *reinterpret_cast<uint4 *>(&A1[r][0]) = make_uint4(a10, a11, a12, a13);
                *reinterpret_cast<uint4 *>(&A2[r][0]) = make_uint4(a20, a21, a22, a23);
                *reinterpret_cast<uint4 *>(&A3[r][0]) = make_uint4(a30, a31, a32, a33);
            }
        }
    }
    __syncthreads();
    // phase 2: vertical combination, one thread per (row, word)
    {
        const int r = tid >> 2, k = tid & 3;
        uint32_t o;
        if (R == 3) {
            const int q = r + 3;
            o = A3[q][k] & A2[q - 1][k] & A2[q + 1][k] & A1[q - 2][k] & A1[q + 2][k] & Bw[q - 3][k] & Bw[q + 3][k];
        } else {
            o = Bw[r][k];
        }
        O[r][k] = o;
    }
    __syncthreads();
    // phase 3: bits -> 0/255 bytes, a warp per row, lane l = staged pixels 4l..4l+3 (lanes 1..30 are the tile)
    const bool word_ok = ((reinterpret_cast<uintptr_t>(dst) | dstride) & 3) == 0;
    for (int r = warp; r < GB_TH; r += 8) {
        const int gy = by + r, gx = bx + 4 * (lane - 1);
        if (gy >= h) break;
        const uint4 o = *reinterpret_cast<const uint4 *>(&O[r][0]);
        if (lane < 1 || lane > 30 || gx >= w) continue;
        const uint32_t m = (((o.x >> lane) & 1u) * 0x000000FFu) | (((o.y >> lane) & 1u) * 0x0000FF00u) |
                           (((o.z >> lane) & 1u) * 0x00FF0000u) | (((o.w >> lane) & 1u) * 0xFF000000u);
        uint8_t *p = dst + (size_t)gy * dstride + gx;
        if (word_ok && gx + 3 < w) {
            *reinterpret_cast<uint32_t *>(p) = m;
        } else {
            for (int k = 0; k < 4 && gx + k < w; k++) p[k] = (uint8_t)(m >> (8 * k));
        }
    }
}

// gray (+ erode radius) (+ threshold) of a BGR image: the fast kernel when the radius is one
// the reference uses (0 or 3), the generic one otherwise.
static int32_t launch_gray_erode(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, int radius,
                                 int do_threshold, uint8_t *dst, size_t dstride, int n_cards = 1, size_t src_card = 0,
                                 size_t dst_card = 0)
{
    if ((radius == 0 || radius == 3) && do_threshold) {
        dim3 grid(cdiv(w, GB_TW), cdiv(h, GB_TH), n_cards);
        ProfScope ps(ctx, radius == 3 ? "gray_erode_bits_kernel<3>" : "gray_erode_bits_kernel<0>", 4.0 * w * h * n_cards);
        if (radius == 3) gray_erode_bits_kernel<3><<<grid, 256, 0, ctx->stream>>>(dbgr, src_card, w, h, stride, dst, dst_card, dstride);
        else gray_erode_bits_kernel<0><<<grid, 256, 0, ctx->stream>>>(dbgr, src_card, w, h, stride, dst, dst_card, dstride);
        OICS_CUDA_TRY(cudaGetLastError());
        ctx->launches++;
    } else if (radius == 0 || radius == 3) {
        dim3 grid(cdiv(w, GF_TW), cdiv(h, GF_TH), n_cards);
        ProfScope ps(ctx, radius == 3 ? "gray_erode_fast_kernel<3>" : "gray_erode_fast_kernel<0>", 4.0 * w * h * n_cards);
        if (radius == 3)
            gray_erode_fast_kernel<3><<<grid, 256, 0, ctx->stream>>>(dbgr, src_card, w, h, stride, do_threshold, dst, dst_card, dstride);
        else
            gray_erode_fast_kernel<0><<<grid, 256, 0, ctx->stream>>>(dbgr, src_card, w, h, stride, do_threshold, dst, dst_card, dstride);
        OICS_CUDA_TRY(cudaGetLastError());
        ctx->launches++;
    } else {
        dim3 grid(cdiv(w, GE_TX), cdiv(h, GE_TY));
        for (int c = 0; c < n_cards; c++) {
            gray_erode_kernel<3><<<grid, 256, 0, ctx->stream>>>(dbgr + (size_t)c * src_card, w, h, stride, radius, do_threshold,
                                                               dst + (size_t)c * dst_card, dstride);
            OICS_CUDA_TRY(cudaGetLastError());
            ctx->launches++;
        }
    }
    return OICS_OK;
}

// One axis of computeResizeAreaTab for destination index d: up to `*n` taps.
struct AreaTaps {
    int s0;        // first source index
    int n;         // number of taps
    float first;   // weight of tap 0
    float mid;     // weight of the full-cell taps
    float last;    // weight of the last tap
    int has_first; // tap 0 is a partial leading cell
    int has_last;  // last tap is a partial trailing cell
};

__device__ __forceinline__ AreaTaps area_taps(int d, double scale, int ssize)
{
    const double fsx1 = __dmul_rn((double)d, scale), fsx2 = __dadd_rn(fsx1, scale);
    const double cell = fmin(scale, __dsub_rn((double)ssize, fsx1));
    int sx1 = (int)ceil(fsx1), sx2 = (int)floor(fsx2);
    sx2 = min(sx2, ssize - 1);
    sx1 = min(sx1, sx2);
    AreaTaps t;
    t.has_first = __dsub_rn((double)sx1, fsx1) > 1e-3;
    t.has_last = __dsub_rn(fsx2, (double)sx2) > 1e-3;
    t.first = (float)__ddiv_rn(__dsub_rn((double)sx1, fsx1), cell);
    t.mid = (float)__ddiv_rn(1.0, cell);
    t.last = (float)__ddiv_rn(fmin(fmin(__dsub_rn(fsx2, (double)sx2), 1.0), cell), cell);
    t.s0 = t.has_first ? sx1 - 1 : sx1;
    t.n = (sx2 - sx1) + t.has_first + t.has_last;
    return t;
}

__device__ __forceinline__ float tap_weight(const AreaTaps &t, int k)
{
    if (k == 0 && t.has_first) return t.first;
    if (k == t.n - 1 && t.has_last) return t.last;
    return t.mid;
}

template <int CH>
__global__ void __launch_bounds__(256) resize_area_kernel(const uint8_t *__restrict__ src0, size_t src_card, int w, int h,
                                                          size_t stride, uint8_t *__restrict__ dst0, size_t dst_card, int dw,
                                                          int dh, size_t dstride, double scale_x, double scale_y, int fast,
                                                          int ix, int iy, float inv_area, int do_threshold)
{
    const int dx = blockIdx.x * blockDim.x + threadIdx.x, dy = blockIdx.y;
    if (dx >= dw) return;
    const uint8_t *src = src0 + (size_t)blockIdx.z * src_card;   // one card of a batch per grid.z
    uint8_t *dst = dst0 + (size_t)blockIdx.z * dst_card;
    int out[CH];
    if (fast) {
#pragma unroll
        for (int c = 0; c < CH; c++) {
            int sum = 0;
            for (int ky = 0; ky < iy; ky++) {
                const uint8_t *p = src + (size_t)(dy * iy + ky) * stride + (size_t)(dx * ix) * CH + c;
                for (int kx = 0; kx < ix; kx++) sum += p[kx * CH];
            }
            out[c] = (ix == 2 && iy == 2) ? (sum + 2) >> 2 : __float2int_rn(__fmul_rn((float)sum, inv_area));
        }
    } else {
        const AreaTaps tx = area_taps(dx, scale_x, w), ty = area_taps(dy, scale_y, h);
#pragma unroll
        for (int c = 0; c < CH; c++) {
            float sum = 0.f;
            for (int j = 0; j < ty.n; j++) {
                const uint8_t *p = src + (size_t)(ty.s0 + j) * stride + (size_t)tx.s0 * CH + c;
                float buf = 0.f;
                for (int k = 0; k < tx.n; k++) buf = __fadd_rn(buf, __fmul_rn((float)p[k * CH], tap_weight(tx, k)));
                const float t = __fmul_rn(tap_weight(ty, j), buf);
                sum = j == 0 ? t : __fadd_rn(sum, t);
            }
            out[c] = __float2int_rn(sum);
        }
    }
#pragma unroll
    for (int c = 0; c < CH; c++) {
        int v = min(255, max(0, out[c]));
        if (do_threshold) v = v > 127 ? 255 : 0;
        dst[(size_t)dy * dstride + (size_t)dx * CH + c] = (uint8_t)v;
    }
}

static int32_t launch_resize(oics_ctx *ctx, const uint8_t *src, int ch, int w, int h, size_t stride, uint8_t *dst,
                             int dw, int dh, int do_threshold, int n_cards = 1, size_t src_card = 0, size_t dst_card = 0)
{
    const double sx = (double)w / dw, sy = (double)h / dh;
    if (sx < 1 || sy < 1) {
        set_error("enlarging resize (%dx%d -> %dx%d) is not supported", w, h, dw, dh);
        return OICS_ERR_UNSUPPORTED;
    }
    const int ix = (int)lrint(sx), iy = (int)lrint(sy);
    const int fast = fabs(sx - ix) < 2.220446049250313e-16 && fabs(sy - iy) < 2.220446049250313e-16;
    const float inv_area = 1.f / (float)(ix * iy);
    if (n_cards > 65535) {
        set_error("resize: at most 65535 cards per launch");
        return OICS_ERR_INVALID_ARG;
    }
    dim3 grid(cdiv(dw, 256), dh, n_cards);
    ProfScope ps(ctx, ch == 1 ? "resize_area_kernel<1>" : "resize_area_kernel<3>",
                 (double)n_cards * ch * ((double)w * h + (double)dw * dh));
    if (ch == 1)
        resize_area_kernel<1><<<grid, 256, 0, ctx->stream>>>(src, src_card, w, h, stride, dst, dst_card, dw, dh, (size_t)dw, sx, sy,
                                                             fast, ix, iy, inv_area, do_threshold);
    else
        resize_area_kernel<3><<<grid, 256, 0, ctx->stream>>>(src, src_card, w, h, stride, dst, dst_card, dw, dh, (size_t)dw * 3, sx,
                                                             sy, fast, ix, iy, inv_area, do_threshold);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    return OICS_OK;
}

// cv::resize(INTER_AREA) of a 1- or 3-channel image (TransformableMatrix::scale_self, transfer.rs:66-91)
int32_t resize_area_device(oics_ctx *ctx, const uint8_t *dsrc, int ch, int w, int h, size_t stride, uint8_t *ddst, int dw,
                           int dh)
{
    return launch_resize(ctx, dsrc, ch, w, h, stride, ddst, dw, dh, 0);
}

int32_t gray_device(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, uint8_t *dgray)
{
    return launch_gray_erode(ctx, dbgr, w, h, stride, 0, 0, dgray, (size_t)w);
}

int32_t preprocess_device(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, int mode, int ow, int oh,
                          uint8_t *dbin)
{
    const bool identity = (ow == w && oh == h);
    if (ow <= 0 || oh <= 0) {
        set_error("pre-processing target size %dx%d is empty", ow, oh);
        return OICS_ERR_INVALID_ARG;
    }
    if (mode == OICS_PRE_OMR) {
        if (identity) return launch_gray_erode(ctx, dbgr, w, h, stride, 3, 1, dbin, (size_t)w);
        OICS_TRY(ctx->tmp8a.ensure((size_t)w * h));
        OICS_TRY(launch_gray_erode(ctx, dbgr, w, h, stride, 3, 0, ctx->tmp8a.as<uint8_t>(), (size_t)w));
        return launch_resize(ctx, ctx->tmp8a.as<uint8_t>(), 1, w, h, (size_t)w, dbin, ow, oh, 1);
    }
    if (mode == OICS_PRE_PROJECTION) {
        const uint8_t *col = dbgr;
        size_t cstride = stride;
        if (!identity) {
            OICS_TRY(ctx->tmp8a.ensure((size_t)ow * oh * 3));
            OICS_TRY(launch_resize(ctx, dbgr, 3, w, h, stride, ctx->tmp8a.as<uint8_t>(), ow, oh, 0));
            col = ctx->tmp8a.as<uint8_t>();
            cstride = (size_t)ow * 3;
        }
        return launch_gray_erode(ctx, col, ow, oh, cstride, 0, 1, dbin, (size_t)ow);
    }
    set_error("unknown pre-processing mode %d", mode);
    return OICS_ERR_INVALID_ARG;
}

// P0 for n cards of one geometry (the batch driver): one launch when no resize follows.
int32_t preprocess_batch_device(oics_ctx *ctx, const uint8_t *dbgrs, int n_cards, size_t card_stride, int w, int h,
                                size_t stride, int mode, int ow, int oh, uint8_t *dbins)
{
    if (mode == OICS_PRE_OMR && ow == w && oh == h)
        return launch_gray_erode(ctx, dbgrs, w, h, stride, 3, 1, dbins, (size_t)w, n_cards, card_stride, (size_t)w * h);
    if (mode == OICS_PRE_OMR && n_cards <= 65535 && ow > 0 && oh > 0) {
        // the shrinking operating point (omr.rs:114-127): gray + erode of every card in one launch, then one
        // INTER_AREA + threshold launch over the batch
        OICS_TRY(ctx->tmp8a.ensure((size_t)n_cards * w * h));
        OICS_TRY(launch_gray_erode(ctx, dbgrs, w, h, stride, 3, 0, ctx->tmp8a.as<uint8_t>(), (size_t)w, n_cards, card_stride,
                                   (size_t)w * h));
        return launch_resize(ctx, ctx->tmp8a.as<uint8_t>(), 1, w, h, (size_t)w, dbins, ow, oh, 1, n_cards, (size_t)w * h,
                             (size_t)ow * oh);
    }
    for (int c = 0; c < n_cards; c++)
        OICS_TRY(preprocess_device(ctx, dbgrs + (size_t)c * card_stride, w, h, stride, mode, ow, oh,
                                   dbins + (size_t)c * ow * oh));
    return OICS_OK;
}

}  // namespace oics
