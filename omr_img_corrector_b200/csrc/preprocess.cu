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
// Fast path of the same operation for radius 0 or 3 (the two
// the reference uses): HBM-bound.  A CTA produces a 128 x 64 tile; the BGR rows of the tile
// and its halo are fetched as whole 32-bit words (coalesced), converted to gray once, and the
// L1 diamond is taken four pixels at a time with byte-SIMD minima:
//     H_k[r] = min over |dx| <= k of gray[r][x + dx]            (k = 1, 2, 3; 2 shifts + 2 vmin each)
//     out[r] = min(H3[r], H2[r +- 1], H1[r +- 2], gray[r +- 3])
static constexpr int GF_TW = 128, GF_TH = 64, GF_WW = GF_TW / 4;   // tile in pixels / words
static constexpr int GF_RAWW = (GF_TW + 8) * 3 / 4 + 1;            // aligned words covering a staged BGR row (136 px)

template <int R>   // R = 0 (gray [+ threshold]) or 3
__global__ void __launch_bounds__(256) gray_erode_fast_kernel(const uint8_t *__restrict__ src, int w, int h,
                                                              size_t stride, int do_threshold,
                                                              uint8_t *__restrict__ dst, size_t dstride)
{
    constexpr int ROWS = GF_TH + 2 * R;
    __shared__ uint32_t raw[ROWS * GF_RAWW];          // BGR words; re-used for H1..H3 after the conversion
    __shared__ uint32_t G[ROWS][GF_WW + 2];           // gray, one halo word per side
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bx = blockIdx.x * GF_TW, by = blockIdx.y * GF_TH;
    __shared__ int sMis[ROWS];                        // byte misalignment of each staged row segment
    // phase 0: the aligned 32-bit words that cover each row segment, a warp per row.  A word is
    // read iff it overlaps the valid bytes of its row (same page as a valid byte: never faults).
    const int seg0 = (bx - 4) * 3, row_bytes = w * 3;
    for (int r = warp; r < ROWS; r += 8) {
        const int gy = by - R + r;
        const bool rin = gy >= 0 && gy < h;
        const uint8_t *a0 = src + (size_t)(rin ? gy : 0) * stride + seg0;
        const int mis = (int)(reinterpret_cast<uintptr_t>(a0) & 3);
        const uint32_t *ab = reinterpret_cast<const uint32_t *>(a0 - mis);
        if (lane == 0) sMis[r] = mis;
#pragma unroll
        for (int k = lane; k < GF_RAWW; k += 32) {
            const int off = seg0 - mis + 4 * k;       // first byte of the word, relative to the row
            const bool in = rin && off + 4 > 0 && off < row_bytes;
            // asynchronous 4-byte copy, zero-filled when the word is outside the image: every
            // row of the tile is in flight at once (no register round trip between rows)
            const unsigned d = (unsigned)__cvta_generic_to_shared(&raw[r * GF_RAWW + k]);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(in ? ab + k : ab),
                         "r"(in ? 4 : 0)
                         : "memory");
        }
    }
    asm volatile("cp.async.wait_all;\n" ::: "memory");
    __syncthreads();
    // phase 1: gray, four pixels per thread; outside the image = 255 (never the minimum)
    for (int i = tid; i < ROWS * (GF_WW + 2); i += 256) {
        const int r = i / (GF_WW + 2), wc = i - r * (GF_WW + 2);
        const int gy = by - R + r, gx = bx - 4 + 4 * wc;
        const uint32_t *rw = raw + r * GF_RAWW + 3 * wc;
        const int sh = 8 * sMis[r];
        const uint32_t a = __funnelshift_r(rw[0], rw[1], sh), b = __funnelshift_r(rw[1], rw[2], sh),
                       c = __funnelshift_r(rw[2], rw[3], sh);
        uint32_t g0 = ((a & 255) * 9798 + ((a >> 8) & 255) * 19235 + ((a >> 16) & 255) * 3735 + 16384) >> 15;
        uint32_t g1 = ((a >> 24) * 9798 + (b & 255) * 19235 + ((b >> 8) & 255) * 3735 + 16384) >> 15;
        uint32_t g2 = (((b >> 16) & 255) * 9798 + (b >> 24) * 19235 + (c & 255) * 3735 + 16384) >> 15;
        uint32_t g3 = (((c >> 8) & 255) * 9798 + ((c >> 16) & 255) * 19235 + (c >> 24) * 3735 + 16384) >> 15;
        const bool rin = gy >= 0 && gy < h;
        if (!(rin && gx >= 0 && gx < w)) g0 = 255;
        if (!(rin && gx + 1 >= 0 && gx + 1 < w)) g1 = 255;
        if (!(rin && gx + 2 >= 0 && gx + 2 < w)) g2 = 255;
        if (!(rin && gx + 3 >= 0 && gx + 3 < w)) g3 = 255;
        G[r][wc] = g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
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

// gray (+ erode radius) (+ threshold) of a BGR image: the fast kernel when the radius is one
// the reference uses (0 or 3), the generic one otherwise.
static int32_t launch_gray_erode(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, int radius,
                                 int do_threshold, uint8_t *dst, size_t dstride)
{
    if (radius == 0 || radius == 3) {
        dim3 grid(cdiv(w, GF_TW), cdiv(h, GF_TH));
        if (radius == 3) gray_erode_fast_kernel<3><<<grid, 256, 0, ctx->stream>>>(dbgr, w, h, stride, do_threshold, dst, dstride);
        else gray_erode_fast_kernel<0><<<grid, 256, 0, ctx->stream>>>(dbgr, w, h, stride, do_threshold, dst, dstride);
    } else {
        dim3 grid(cdiv(w, GE_TX), cdiv(h, GE_TY));
        gray_erode_kernel<3><<<grid, 256, 0, ctx->stream>>>(dbgr, w, h, stride, radius, do_threshold, dst, dstride);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
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
__global__ void __launch_bounds__(256) resize_area_kernel(const uint8_t *__restrict__ src, int w, int h, size_t stride,
                                                          uint8_t *__restrict__ dst, int dw, int dh, size_t dstride,
                                                          double scale_x, double scale_y, int fast, int ix, int iy,
                                                          float inv_area, int do_threshold)
{
    const int dx = blockIdx.x * blockDim.x + threadIdx.x, dy = blockIdx.y;
    if (dx >= dw) return;
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
                             int dw, int dh, int do_threshold)
{
    const double sx = (double)w / dw, sy = (double)h / dh;
    if (sx < 1 || sy < 1) {
        set_error("enlarging resize (%dx%d -> %dx%d) is not supported", w, h, dw, dh);
        return OICS_ERR_UNSUPPORTED;
    }
    const int ix = (int)lrint(sx), iy = (int)lrint(sy);
    const int fast = fabs(sx - ix) < 2.220446049250313e-16 && fabs(sy - iy) < 2.220446049250313e-16;
    const float inv_area = 1.f / (float)(ix * iy);
    dim3 grid(cdiv(dw, 256), dh);
    if (ch == 1)
        resize_area_kernel<1><<<grid, 256, 0, ctx->stream>>>(src, w, h, stride, dst, dw, dh, (size_t)dw, sx, sy, fast, ix,
                                                             iy, inv_area, do_threshold);
    else
        resize_area_kernel<3><<<grid, 256, 0, ctx->stream>>>(src, w, h, stride, dst, dw, dh, (size_t)dw * 3, sx, sy, fast,
                                                             ix, iy, inv_area, do_threshold);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    return OICS_OK;
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

}  // namespace oics
