// W1: cv::warpAffine(INTER_NEAREST | INTER_LINEAR, BORDER_CONSTANT) as the reference
// uses it for the final correction (omr.rs:408-445 NN; transfer.rs:459-523 with the
// caller's flags; core/src/main.rs:41-51,72-92 LINEAR).  Exact OpenCV fixed point:
// AB_BITS = 10 coordinate tables, INTER_BITS = 5 bilinear weights (32-fy)(32-fx)*32,
// (sum + 2^14) >> 15.  Destination-tiled: each thread produces four consecutive
// destination pixels and stores them as 32-bit words; the source is gathered through
// the read-only path (A4 BGR is 26 MB: L2 resident after first touch).
#include "common.h"

namespace oics {

__global__ void warp_tables_kernel(const double *__restrict__ Minv, int dw, int dh, int round_delta,
                                   int2 *__restrict__ ab, int2 *__restrict__ xy)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const double *M = Minv;
    if (i < dw) {
        double x = (double)i;
        ab[i] = make_int2(__double2int_rn(__dmul_rn(__dmul_rn(M[0], x), 1024.0)),
                          __double2int_rn(__dmul_rn(__dmul_rn(M[3], x), 1024.0)));
    }
    if (i < dh) {
        double y = (double)i;
        int X0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(M[1], y), M[2]), 1024.0)) + round_delta;
        int Y0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(M[4], y), M[5]), 1024.0)) + round_delta;
        xy[i] = make_int2(X0, Y0);
    }
}

__device__ __forceinline__ int sat16(int v) { return max(-32768, min(32767, v)); }

template <int CH, bool LINEAR>
__global__ void __launch_bounds__(256) warp_kernel(const uint8_t *__restrict__ src, int w, int h, size_t sstride,
                                                   uint8_t *__restrict__ dst, int dw, int dh, size_t dstride,
                                                   const int2 *__restrict__ ab, const int2 *__restrict__ xy,
                                                   uchar4 border, int aligned)
{
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y = blockIdx.y;
    if (x4 >= dw) return;
    const int2 p = xy[y];
    const uint8_t bval[4] = {border.x, border.y, border.z, border.w};
    uint8_t out[4 * CH];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int x = min(x4 + i, dw - 1);
        const int2 t = __ldg(ab + x);
        const int a = p.x + t.x, b = p.y + t.y;
        if (!LINEAR) {
            const int sx = sat16(a >> 10), sy = sat16(b >> 10);
            if ((unsigned)sx < (unsigned)w && (unsigned)sy < (unsigned)h) {
                const uint8_t *s = src + (size_t)sy * sstride + (size_t)sx * CH;
#pragma unroll
                for (int c = 0; c < CH; c++) out[i * CH + c] = __ldg(s + c);
            } else {
#pragma unroll
                for (int c = 0; c < CH; c++) out[i * CH + c] = bval[c];
            }
        } else {
            const int X = a >> 5, Y = b >> 5;
            const int sx = sat16(X >> 5), sy = sat16(Y >> 5), fx = X & 31, fy = Y & 31;
            const int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32;
            const int w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
            const bool x0in = (unsigned)sx < (unsigned)w, x1in = (unsigned)(sx + 1) < (unsigned)w;
            const bool y0in = (unsigned)sy < (unsigned)h, y1in = (unsigned)(sy + 1) < (unsigned)h;
            const uint8_t *s0 = src + (ptrdiff_t)sy * (ptrdiff_t)sstride + (ptrdiff_t)sx * CH;
            const uint8_t *s1 = s0 + sstride;
#pragma unroll
            for (int c = 0; c < CH; c++) {
                const int p00 = (x0in && y0in) ? __ldg(s0 + c) : bval[c];
                const int p01 = (x1in && y0in) ? __ldg(s0 + CH + c) : bval[c];
                const int p10 = (x0in && y1in) ? __ldg(s1 + c) : bval[c];
                const int p11 = (x1in && y1in) ? __ldg(s1 + CH + c) : bval[c];
                const int v = (p00 * w00 + p01 * w01 + p10 * w10 + p11 * w11 + 16384) >> 15;
                out[i * CH + c] = (uint8_t)min(255, max(0, v));
            }
        }
    }
    uint8_t *d = dst + (size_t)y * dstride + (size_t)x4 * CH;
    if (aligned && x4 + 4 <= dw) {
        uint32_t *d32 = reinterpret_cast<uint32_t *>(d);
#pragma unroll
        for (int k = 0; k < CH; k++)
            d32[k] = (uint32_t)out[4 * k] | ((uint32_t)out[4 * k + 1] << 8) | ((uint32_t)out[4 * k + 2] << 16) |
                     ((uint32_t)out[4 * k + 3] << 24);
    } else {
        const int n = min(4, dw - x4) * CH;
        for (int k = 0; k < n; k++) d[k] = out[k];
    }
}

// ---------------------------------------------------------------------------
// Flat variant for tightly packed destinations (dstride == dw * CH, the layout of every image
// this library returns): thread t produces destination pixels 4t .. 4t+3 in flat (row-major)
// order = CH aligned 32-bit words whatever the row alignment.  All index arithmetic is 32-bit
// (the host takes this path only for images below 2^31 bytes and sides <= 32767, where
// saturate_cast<short> of the source coordinates cannot change an in-range sample), the two
// coordinate tables are one int2 load per pixel, a source pixel is two aligned words + a
// funnel shift (CH = 3, 4) and out-of-range samples are selected, not branched.  The CH words
// of a thread are transposed through shared memory so a warp stores CH x 128 contiguous bytes.
template <int CH>
__device__ __forceinline__ uint32_t fetch_px(const uint32_t *__restrict__ src32, uint32_t byte_off, uint32_t last_word)
{
    if (CH == 1) return __ldg(reinterpret_cast<const uint8_t *>(src32) + byte_off);
    const uint32_t i0 = byte_off >> 2, i1 = min(i0 + 1u, last_word);
    const uint32_t v = __funnelshift_r(__ldg(src32 + i0), __ldg(src32 + i1), 8 * (int)(byte_off & 3u));
    return CH == 3 ? (v & 0x00FFFFFFu) : v;
}

template <int CH, bool LINEAR>
__global__ void __launch_bounds__(256) warp_flat_kernel(const uint32_t *__restrict__ src32, int w, int h, uint32_t sstride,
                                                        uint32_t *__restrict__ dst32, uint32_t dw, uint32_t dh,
                                                        uint32_t npix, const int2 *__restrict__ ab,
                                                        const int2 *__restrict__ xy, uint32_t border, uint32_t last_word)
{
    __shared__ uint32_t sT[8][32 * CH + 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t p0 = 4u * (blockIdx.x * 256u + threadIdx.x);
    uint32_t px[4] = {0u, 0u, 0u, 0u};
    if (p0 < npix) {
        const uint32_t y = p0 / dw, x = p0 - y * dw;
        const int2 r0 = __ldg(xy + y);
        int2 r1 = r0;
        if (x + 3u >= dw) r1 = __ldg(xy + min(y + 1u, dh - 1u));   // the four pixels straddle two rows
#pragma unroll
        for (int i = 0; i < 4; i++) {
            uint32_t xi = x + (uint32_t)i;
            int2 r = r0;
            if (xi >= dw) { xi -= dw; r = r1; }    // pixels past npix fall here too: loaded, never stored
            const int2 t = __ldg(ab + xi);
            const int a = r.x + t.x, b = r.y + t.y;
            if (!LINEAR) {
                const int sx = a >> 10, sy = b >> 10;
                const bool in = (unsigned)sx < (unsigned)w && (unsigned)sy < (unsigned)h;
                const uint32_t v = fetch_px<CH>(src32, in ? (uint32_t)sy * sstride + (uint32_t)sx * CH : 0u, last_word);
                px[i] = in ? v : border;
            } else {
                const int X = a >> 5, Y = b >> 5;
                const int sx = X >> 5, sy = Y >> 5, fx = X & 31, fy = Y & 31;
                const int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32;
                const int w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
                const bool x0in = (unsigned)sx < (unsigned)w, x1in = (unsigned)(sx + 1) < (unsigned)w;
                const bool y0in = (unsigned)sy < (unsigned)h, y1in = (unsigned)(sy + 1) < (unsigned)h;
                const uint32_t base = (uint32_t)sy * sstride + (uint32_t)sx * CH;   // wraps when unused
                const uint32_t f00 = fetch_px<CH>(src32, (x0in && y0in) ? base : 0u, last_word);
                const uint32_t f01 = fetch_px<CH>(src32, (x1in && y0in) ? base + CH : 0u, last_word);
                const uint32_t f10 = fetch_px<CH>(src32, (x0in && y1in) ? base + sstride : 0u, last_word);
                const uint32_t f11 = fetch_px<CH>(src32, (x1in && y1in) ? base + sstride + CH : 0u, last_word);
                const uint32_t q00 = (x0in && y0in) ? f00 : border, q01 = (x1in && y0in) ? f01 : border;
                const uint32_t q10 = (x0in && y1in) ? f10 : border, q11 = (x1in && y1in) ? f11 : border;
                uint32_t o = 0;
#pragma unroll
                for (int c = 0; c < CH; c++) {
                    const int v = (int)((q00 >> (8 * c)) & 255) * w00 + (int)((q01 >> (8 * c)) & 255) * w01 +
                                  (int)((q10 >> (8 * c)) & 255) * w10 + (int)((q11 >> (8 * c)) & 255) * w11 + 16384;
                    o |= (uint32_t)min(255, max(0, v >> 15)) << (8 * c);
                }
                px[i] = o;
            }
        }
    }
    // pack 4 pixels -> CH words, transpose through shared memory, store coalesced
    uint32_t wv[CH];
    if (CH == 1) wv[0] = px[0] | (px[1] << 8) | (px[2] << 16) | (px[3] << 24);
    if (CH == 3) {
        wv[0] = px[0] | (px[1] << 24);
        wv[CH > 1 ? 1 : 0] = (px[1] >> 8) | (px[2] << 16);
        wv[CH > 2 ? 2 : 0] = (px[2] >> 16) | (px[3] << 8);
    }
    if (CH == 4) {
#pragma unroll
        for (int k = 0; k < CH; k++) wv[k] = px[k & 3];
    }
    const uint32_t nbytes = npix * CH, nwords = (nbytes + 3u) >> 2;
    const uint32_t wbase = (blockIdx.x * 256u + warp * 32u) * CH;   // first word of this warp
#pragma unroll
    for (int k = 0; k < CH; k++) sT[warp][CH * lane + k] = wv[k];
    __syncwarp();
#pragma unroll
    for (int k = 0; k < CH; k++) {
        const uint32_t wi = wbase + 32u * k + lane;
        if (wi < nwords) {
            const uint32_t v = sT[warp][32 * k + lane];
            if (4u * wi + 4u <= nbytes) dst32[wi] = v;
            else {   // the last word of the image may be partial
                uint8_t *d8 = reinterpret_cast<uint8_t *>(dst32 + wi);
                for (uint32_t q = 0; q < nbytes - 4u * wi; q++) d8[q] = (uint8_t)(v >> (8 * q));
            }
        }
    }
}

int32_t warp_device(oics_ctx *ctx, const uint8_t *dsrc, int ch, int w, int h, size_t sstride, const double Mfwd[6],
                    int interp, const uint8_t border[4], uint8_t *ddst, int dw, int dh, size_t dstride, cudaStream_t on,
                    DevBuf *tables)
{
    // on / tables: another stream of the context and the table scratch that goes with it (the encode lanes of the
    // batch driver); null = ctx->stream and ctx->misc
    cudaStream_t st = on ? on : ctx->stream;
    DevBuf &tab = tables ? *tables : ctx->misc;
    if (ch != 1 && ch != 3 && ch != 4) {
        set_error("warp: %d channels not supported (1, 3 or 4)", ch);
        return OICS_ERR_UNSUPPORTED;
    }
    if (interp != OICS_INTER_NEAREST && interp != OICS_INTER_LINEAR) {
        set_error("warp: interpolation flag %d not supported (0 nearest, 1 linear)", interp);
        return OICS_ERR_UNSUPPORTED;
    }
    if (dw <= 0 || dh <= 0) return OICS_OK;
    double inv[6];
    invert_affine(Mfwd, inv);
    // tables: [6 doubles][{ad, bd} dw][{X0, Y0} dh]
    const size_t bytes = 64 + sizeof(int2) * ((size_t)dw + (size_t)dh + 2);
    OICS_TRY(tab.ensure(bytes));
    double *dM = tab.as<double>();
    int2 *ab = reinterpret_cast<int2 *>(tab.as<uint8_t>() + 64);
    int2 *xy = ab + dw;
    OICS_CUDA_TRY(cudaMemcpyAsync(dM, inv, sizeof(inv), cudaMemcpyHostToDevice, st));
    warp_tables_kernel<<<cdiv(max(dw, dh), 256), 256, 0, st>>>(dM, dw, dh, interp == OICS_INTER_NEAREST ? 512 : 16, ab, xy);
    OICS_CUDA_TRY(cudaGetLastError());
    const uchar4 bv = make_uchar4(border[0], border[1], border[2], border[3]);
    const bool lin = interp == OICS_INTER_LINEAR;
    const size_t src_bytes = (size_t)(h - 1) * sstride + (size_t)w * ch;
    const long long npix = (long long)dw * dh;
    // 32-bit flat path: packed destination, word-aligned buffers, everything below 2^31 bytes, sides
    // that saturate_cast<short> cannot touch
    const bool flat = dstride == (size_t)dw * ch && ((reinterpret_cast<uintptr_t>(ddst) | reinterpret_cast<uintptr_t>(dsrc)) & 3) == 0 &&
                      src_bytes < ((size_t)1 << 31) && (size_t)npix * ch < ((size_t)1 << 31) && sstride < ((size_t)1 << 31) &&
                      w <= 32767 && h <= 32767 && dw >= 4;
    if (flat) {
        const uint32_t last_word = (uint32_t)((src_bytes - 1) >> 2);
        uint32_t bcol = 0;
        for (int c = 0; c < ch; c++) bcol |= (uint32_t)border[c] << (8 * c);
        const unsigned blocks = (unsigned)((npix + 1023) / 1024);
        const uint32_t *s32 = reinterpret_cast<const uint32_t *>(dsrc);
        uint32_t *d32 = reinterpret_cast<uint32_t *>(ddst);
#define OICS_WARPF(CH, LIN)                                                                                          \
    warp_flat_kernel<CH, LIN><<<blocks, 256, 0, st>>>(s32, w, h, (uint32_t)sstride, d32, (uint32_t)dw, (uint32_t)dh, \
                                                      (uint32_t)npix, ab, xy, bcol, last_word)
        ProfScope ps(ctx, lin ? "warp_flat_kernel<LINEAR>" : "warp_flat_kernel<NEAREST>", (double)ch * ((double)w * h + (double)npix), st);
        if (ch == 1) { if (lin) OICS_WARPF(1, true); else OICS_WARPF(1, false); }
        else if (ch == 3) { if (lin) OICS_WARPF(3, true); else OICS_WARPF(3, false); }
        else { if (lin) OICS_WARPF(4, true); else OICS_WARPF(4, false); }
#undef OICS_WARPF
    } else {
        const int aligned = ((reinterpret_cast<uintptr_t>(ddst) | dstride) & 3) == 0;
        dim3 grid(cdiv(cdiv(dw, 4), 256), dh);
#define OICS_WARP(CH, LIN) \
    warp_kernel<CH, LIN><<<grid, 256, 0, st>>>(dsrc, w, h, sstride, ddst, dw, dh, dstride, ab, xy, bv, aligned)
        ProfScope ps(ctx, lin ? "warp_kernel<LINEAR>" : "warp_kernel<NEAREST>", (double)ch * ((double)w * h + (double)npix), st);
        if (ch == 1) { if (lin) OICS_WARP(1, true); else OICS_WARP(1, false); }
        else if (ch == 3) { if (lin) OICS_WARP(3, true); else OICS_WARP(3, false); }
        else { if (lin) OICS_WARP(4, true); else OICS_WARP(4, false); }
#undef OICS_WARP
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 2;
    return OICS_OK;
}

}  // namespace oics
