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
                                   int *__restrict__ ad, int *__restrict__ bd, int2 *__restrict__ xy)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const double *M = Minv;
    if (i < dw) {
        double x = (double)i;
        ad[i] = __double2int_rn(__dmul_rn(__dmul_rn(M[0], x), 1024.0));
        bd[i] = __double2int_rn(__dmul_rn(__dmul_rn(M[3], x), 1024.0));
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
                                                   const int *__restrict__ ad, const int *__restrict__ bd,
                                                   const int2 *__restrict__ xy, uchar4 border, int aligned)
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
        const int a = p.x + ad[x], b = p.y + bd[x];
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

int32_t warp_device(oics_ctx *ctx, const uint8_t *dsrc, int ch, int w, int h, size_t sstride, const double Mfwd[6],
                    int interp, const uint8_t border[4], uint8_t *ddst, int dw, int dh, size_t dstride)
{
    cudaStream_t st = ctx->stream;
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
    // tables: [6 doubles][ad dw][bd dw][xy dh]
    const size_t bytes = 64 + sizeof(int) * (2 * (size_t)dw + 2 * (size_t)dh + 8);
    OICS_TRY(ctx->misc.ensure(bytes));
    double *dM = ctx->misc.as<double>();
    int *ad = reinterpret_cast<int *>(ctx->misc.as<uint8_t>() + 64);
    int *bd = ad + dw;
    int2 *xy = reinterpret_cast<int2 *>(bd + dw + ((2 * dw) & 1));
    OICS_CUDA_TRY(cudaMemcpyAsync(dM, inv, sizeof(inv), cudaMemcpyHostToDevice, st));
    warp_tables_kernel<<<cdiv(max(dw, dh), 256), 256, 0, st>>>(dM, dw, dh, interp == OICS_INTER_NEAREST ? 512 : 16, ad,
                                                               bd, xy);
    OICS_CUDA_TRY(cudaGetLastError());
    const uchar4 bv = make_uchar4(border[0], border[1], border[2], border[3]);
    const int aligned = ((reinterpret_cast<uintptr_t>(ddst) | dstride) & 3) == 0;
    dim3 grid(cdiv(cdiv(dw, 4), 256), dh);
#define OICS_WARP(CH, LIN) \
    warp_kernel<CH, LIN><<<grid, 256, 0, st>>>(dsrc, w, h, sstride, ddst, dw, dh, dstride, ad, bd, xy, bv, aligned)
    const bool lin = interp == OICS_INTER_LINEAR;
    if (ch == 1) { if (lin) OICS_WARP(1, true); else OICS_WARP(1, false); }
    else if (ch == 3) { if (lin) OICS_WARP(3, true); else OICS_WARP(3, false); }
    else { if (lin) OICS_WARP(4, true); else OICS_WARP(4, false); }
#undef OICS_WARP
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 2;
    return OICS_OK;
}

}  // namespace oics
