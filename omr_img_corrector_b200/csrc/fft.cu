// F1: fft::get_fft_image (fft.rs:124-141 and helpers fft.rs:20-122): the centred,
// log-normalised 8-bit magnitude spectrum of a gray card.
//
//   u8 * (1/255) -> f32                                    fft.rs:127-130
//   2-D DFT, scaled by 1/(W*H)                             fft.rs:42-65
//   quadrant swap (cx = W/2, cy = H/2)                     fft.rs:67-86
//   |.| in f32, min-max -> [0,1], *255                     fft.rs:88-114
//   (+1/255, ln, min-max -> [0,1]) * 255 -> u8             fft.rs:116-122, 137-138
//
// cv::dft is a float32 transform that is not bit-stable between its own code paths
// (SURVEY F13), so the transform here is evaluated in float64 -- closer to the exact
// DFT than either cv path -- and every later step follows OpenCV's conversions
// literally: convertTo(CV_32F, a, b) = f32(f64(v)*a + b), convertTo(CV_8U, 255) =
// saturate(rint(f64(v)*255)), magnitude/log in f32.
//
// The card sizes are not powers of two (2480 = 2^4*5*31, 3508 = 2^2*877): every 1-D
// transform is Bluestein's chirp-z over a power-of-two length m >= 2n-1 held entirely
// in shared memory (m = 8192 complex128 = 128 KB for A4 rows and columns), one CTA per
// row / column: forward radix-2 DIF (natural -> bit-reversed), pointwise product with
// the pre-transformed chirp stored bit-reversed, inverse radix-2 DIT (bit-reversed ->
// natural): no reordering pass.  Power-of-two lengths skip the chirp.
//   pass 1  rows:    u8 row -> complex128 row spectrum          (1 + 16 B / pixel)
//   pass 2  columns: complex128 column -> f32 |.| at its swapped position, global
//                    min/max by atomics                          (16 + 4 B / pixel)
//   pass 3  normalise, log, min/max of the log                   (4 + 4 B / pixel)
//   pass 4  normalise -> u8                                      (4 + 1 B / pixel)
#include <algorithm>
#include <cmath>
#include <complex>

#include "common.h"

namespace oics {

static constexpr int FFT_THREADS = 512;

__device__ __forceinline__ double2 cmul(double2 a, double2 b)
{
    return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ double2 cmul_conj(double2 a, double2 b)  // a * conj(b)
{
    return make_double2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}

__device__ __forceinline__ double2 cadd(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 csub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }

// Two consecutive radix-2 DIF stages (spans 2h and h) on the four elements i0 + {0, h, 2h, 3h}:
// the same operations in the same order as the two separate stages, one pass over shared memory.
__device__ __forceinline__ void dif_pair(double2 *a, int g, int h, int tstep1, const double2 *__restrict__ tw)
{
    const int pos = g & (h - 1);
    const int i0 = ((g - pos) << 2) + pos;
    const double2 a0 = a[i0], a1 = a[i0 + h], a2 = a[i0 + 2 * h], a3 = a[i0 + 3 * h];
    const double2 w1a = __ldg(tw + pos * tstep1), w1b = __ldg(tw + (pos + h) * tstep1), w2 = __ldg(tw + pos * 2 * tstep1);
    const double2 b0 = cadd(a0, a2), b2 = cmul(csub(a0, a2), w1a);
    const double2 b1 = cadd(a1, a3), b3 = cmul(csub(a1, a3), w1b);
    a[i0] = cadd(b0, b1);
    a[i0 + h] = cmul(csub(b0, b1), w2);
    a[i0 + 2 * h] = cadd(b2, b3);
    a[i0 + 3 * h] = cmul(csub(b2, b3), w2);
}

// forward DIF, natural in -> bit-reversed out: stages fused in pairs (one barrier per pair)
__device__ void fft_dif(double2 *a, int m, const double2 *__restrict__ tw)
{
    int s = m >> 1;
    int lg = 0;
    while ((1 << lg) < m) lg++;
    if (lg & 1) {   // odd number of stages: the first one alone
        for (int t = threadIdx.x; t < (m >> 1); t += blockDim.x) {
            const int i = t, j = t + s;           // s = m / 2: pos = t
            const double2 u = a[i], v = a[j];
            a[i] = cadd(u, v);
            a[j] = cmul(csub(u, v), __ldg(tw + t));
        }
        __syncthreads();
        s >>= 1;
    }
    for (; s >= 2; s >>= 2) {
        const int h = s >> 1, tstep1 = (m >> 1) / s;
        for (int g = threadIdx.x; g < (m >> 2); g += blockDim.x) dif_pair(a, g, h, tstep1, tw);
        __syncthreads();
    }
}

// Two consecutive radix-2 DIT stages (spans s and 2s) with conjugated twiddles on i0 + {0, s, 2s, 3s}
__device__ __forceinline__ void dit_pair(double2 *a, int g, int s, int tstepA, const double2 *__restrict__ tw)
{
    const int pos = g & (s - 1);
    const int i0 = ((g - pos) << 2) + pos;
    const double2 a0 = a[i0], a1 = a[i0 + s], a2 = a[i0 + 2 * s], a3 = a[i0 + 3 * s];
    const int tstepB = tstepA >> 1;
    const double2 wa = __ldg(tw + pos * tstepA), wb0 = __ldg(tw + pos * tstepB), wb1 = __ldg(tw + (pos + s) * tstepB);
    const double2 v01 = cmul_conj(a1, wa), v23 = cmul_conj(a3, wa);
    const double2 b0 = cadd(a0, v01), b1 = csub(a0, v01), b2 = cadd(a2, v23), b3 = csub(a2, v23);
    const double2 v02 = cmul_conj(b2, wb0), v13 = cmul_conj(b3, wb1);
    a[i0] = cadd(b0, v02);
    a[i0 + 2 * s] = csub(b0, v02);
    a[i0 + s] = cadd(b1, v13);
    a[i0 + 3 * s] = csub(b1, v13);
}

// inverse DIT with conjugated twiddles, bit-reversed in -> natural out (unscaled)
__device__ void ifft_dit(double2 *a, int m, const double2 *__restrict__ tw)
{
    int s = 1;
    for (; (s << 1) <= (m >> 1); s <<= 2) {
        const int tstepA = (m >> 1) / s;
        for (int g = threadIdx.x; g < (m >> 2); g += blockDim.x) dit_pair(a, g, s, tstepA, tw);
        __syncthreads();
    }
    if (s <= (m >> 1)) {   // odd number of stages: the last one (span m / 2) alone
        for (int t = threadIdx.x; t < (m >> 1); t += blockDim.x) {
            const int i = t, j = t + s;
            const double2 u = a[i];
            const double2 v = cmul_conj(a[j], __ldg(tw + t));
            a[i] = cadd(u, v);
            a[j] = csub(u, v);
        }
        __syncthreads();
    }
}

struct PlanDev {
    int n, m, direct, log2m;
    const double2 *chirp;  // n
    const double2 *bbr;    // m, bit-reversed spectrum of the conjugate chirp
    const double2 *tw;     // m/2
};

// In: a[0..n) = the sequence.  Out: a[0..n) = its DFT (unscaled).  a has m entries.
__device__ void dft_block(double2 *a, const PlanDev &P)
{
    const int n = P.n, m = P.m;
    if (P.direct) {
        __syncthreads();
        fft_dif(a, m, P.tw);
        // bit-reversal permutation in place
        for (int i = threadIdx.x; i < m; i += blockDim.x) {
            const int r = P.log2m ? (int)(__brev((unsigned)i) >> (32 - P.log2m)) : 0;
            if (r > i) {
                const double2 t = a[i];
                a[i] = a[r];
                a[r] = t;
            }
        }
        __syncthreads();
        return;
    }
    for (int i = threadIdx.x; i < m; i += blockDim.x)
        a[i] = i < n ? cmul(a[i], __ldg(P.chirp + i)) : make_double2(0.0, 0.0);
    __syncthreads();
    fft_dif(a, m, P.tw);
    for (int i = threadIdx.x; i < m; i += blockDim.x) a[i] = cmul(a[i], __ldg(P.bbr + i));
    __syncthreads();
    ifft_dit(a, m, P.tw);
    const double inv_m = 1.0 / (double)m;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const double2 c = cmul(a[i], __ldg(P.chirp + i));
        a[i] = make_double2(c.x * inv_m, c.y * inv_m);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(FFT_THREADS) fft_rows_kernel(const uint8_t *__restrict__ gray, size_t stride,
                                                               int w, PlanDev P, double2 *__restrict__ out)
{
    extern __shared__ double2 sm[];
    const int y = blockIdx.x;
    const uint8_t *row = gray + (size_t)y * stride;
    for (int i = threadIdx.x; i < w; i += blockDim.x) {
        // convertTo(CV_32F, 1/255): f32(f64(v) * (1/255))
        const float f = __double2float_rn(__dmul_rn((double)row[i], 1.0 / 255.0));
        sm[i] = make_double2((double)f, 0.0);
    }
    if (P.direct)
        for (int i = w + threadIdx.x; i < P.m; i += blockDim.x) sm[i] = make_double2(0.0, 0.0);
    __syncthreads();
    dft_block(sm, P);
    double2 *o = out + (size_t)y * w;
    for (int i = threadIdx.x; i < w; i += blockDim.x) o[i] = sm[i];
}

__device__ __forceinline__ uint32_t float_key(float f)
{
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// block-wide min/max of keys -> two atomics per CTA
__device__ void block_minmax(uint32_t kmin, uint32_t kmax, uint32_t *gmm)
{
    __shared__ uint32_t smin[32], smax[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
        kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
    }
    if (lane == 0) { smin[warp] = kmin; smax[warp] = kmax; }
    __syncthreads();
    if (warp == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        kmin = lane < nw ? smin[lane] : 0xffffffffu;
        kmax = lane < nw ? smax[lane] : 0u;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
            kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
        }
        if (lane == 0) {
            atomicMin(gmm, kmin);
            atomicMax(gmm + 1, kmax);
        }
    }
}

__global__ void __launch_bounds__(FFT_THREADS) fft_cols_kernel(const double2 *__restrict__ in, int w, int h, PlanDev P,
                                                               float *__restrict__ mag, uint32_t *__restrict__ gmm)
{
    extern __shared__ double2 sm[];
    const int x = blockIdx.x;
    for (int i = threadIdx.x; i < h; i += blockDim.x) sm[i] = in[(size_t)i * w + x];
    if (P.direct)
        for (int i = h + threadIdx.x; i < P.m; i += blockDim.x) sm[i] = make_double2(0.0, 0.0);
    __syncthreads();
    dft_block(sm, P);
    const double total = (double)w * (double)h;  // DFT_SCALE
    const int cx = w / 2, cy = h / 2;
    const int dx = x < 2 * cx ? (x + cx) % (2 * cx) : x;
    uint32_t kmin = 0xffffffffu, kmax = 0u;
    for (int i = threadIdx.x; i < h; i += blockDim.x) {
        const double2 c = sm[i];
        const float re = __double2float_rn(__ddiv_rn(c.x, total)), im = __double2float_rn(__ddiv_rn(c.y, total));
        const float m = __fsqrt_rn(__fadd_rn(__fmul_rn(re, re), __fmul_rn(im, im)));  // cv::magnitude, f32
        const int dy = (i < 2 * cy && x < 2 * cx) ? (i + cy) % (2 * cy) : i;
        const int ddx = i < 2 * cy ? dx : x;
        mag[(size_t)dy * w + ddx] = m;
        const uint32_t k = float_key(m);
        kmin = min(kmin, k);
        kmax = max(kmax, k);
    }
    block_minmax(kmin, kmax, gmm);
}

// fft.rs:88-122: min-max normalise, *255 (-> mag8), +1/255, ln; min/max of the log
__global__ void __launch_bounds__(256) fft_log_kernel(float *__restrict__ mag, size_t n, const uint32_t *__restrict__ gmm,
                                                      uint8_t *__restrict__ mag8, uint32_t *__restrict__ gmm2)
{
    const double mn = (double)key_float(gmm[0]), mx = (double)key_float(gmm[1]);
    const double inv = __ddiv_rn(1.0, __dsub_rn(mx, mn));
    uint32_t kmin = 0xffffffffu, kmax = 0u;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float a = mag[i];
        const float b = __double2float_rn(__dadd_rn((double)a, -mn));
        const float c = __double2float_rn(__dmul_rn((double)b, inv));
        const float m = __double2float_rn(__dmul_rn((double)c, 255.0));
        if (mag8) {
            const double v = __dmul_rn((double)m, 255.0);
            mag8[i] = (uint8_t)min(255, max(0, __double2int_rn(fmin(fmax(v, -1.0), 256.0))));
        }
        const float l0 = __double2float_rn(__dadd_rn((double)m, 1.0 / 255.0));
        const float lg = __double2float_rn(log((double)l0));  // cv::log on CV_32F
        mag[i] = lg;
        const uint32_t k = float_key(lg);
        kmin = min(kmin, k);
        kmax = max(kmax, k);
    }
    block_minmax(kmin, kmax, gmm2);
}

__global__ void __launch_bounds__(256) fft_log8_kernel(const float *__restrict__ lg, size_t n,
                                                       const uint32_t *__restrict__ gmm2, uint8_t *__restrict__ log8)
{
    const double mn = (double)key_float(gmm2[0]), mx = (double)key_float(gmm2[1]);
    const double inv = __ddiv_rn(1.0, __dsub_rn(mx, mn));
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float b = __double2float_rn(__dadd_rn((double)lg[i], -mn));
        const float c = __double2float_rn(__dmul_rn((double)b, inv));
        const double v = __dmul_rn((double)c, 255.0);
        log8[i] = (uint8_t)min(255, max(0, __double2int_rn(fmin(fmax(v, -1.0), 256.0))));
    }
}

// ---------------------------------------------------------------------------
// host: plan tables
static void host_fft(std::vector<std::complex<double>> &a)
{
    const size_t m = a.size();
    for (size_t i = 1, j = 0; i < m; i++) {
        size_t bit = m >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) std::swap(a[i], a[j]);
    }
    const double pi = 3.14159265358979323846;
    for (size_t len = 2; len <= m; len <<= 1) {
        for (size_t i = 0; i < m; i += len)
            for (size_t k = 0; k < len / 2; k++) {
                const double ang = -2.0 * pi * (double)k / (double)len;
                const std::complex<double> wv(std::cos(ang), std::sin(ang));
                const std::complex<double> u = a[i + k], v = a[i + k + len / 2] * wv;
                a[i + k] = u + v;
                a[i + k + len / 2] = u - v;
            }
    }
}

static int32_t build_plan(oics_ctx *ctx, FftPlan &pl, int n)
{
    if (pl.n == n) return OICS_OK;
    const double pi = 3.14159265358979323846;
    int m = 1, lg = 0;
    const bool pow2 = (n & (n - 1)) == 0;
    const int need = pow2 ? n : 2 * n - 1;
    while (m < need) { m <<= 1; lg++; }
    if ((size_t)m * sizeof(double2) > (size_t)ctx->smem_optin - 2048) {
        set_error("fft: a transform of length %d needs %zu bytes of shared memory (have %d)", n, (size_t)m * sizeof(double2),
                  ctx->smem_optin);
        return OICS_ERR_UNSUPPORTED;
    }
    std::vector<std::complex<double>> chirp((size_t)n), b((size_t)m, 0.0), tw((size_t)std::max(1, m / 2));
    for (int k = 0; k < m / 2; k++) {
        const double ang = -2.0 * pi * (double)k / (double)m;
        tw[k] = std::complex<double>(std::cos(ang), std::sin(ang));
    }
    if (m < 2) tw[0] = 1.0;
    if (!pow2) {
        for (int j = 0; j < n; j++) {
            const long long q = ((long long)j * j) % (2LL * n);  // exp(-i pi j^2 / n), argument reduced exactly
            const double ang = -pi * (double)q / (double)n;
            chirp[j] = std::complex<double>(std::cos(ang), std::sin(ang));
        }
        for (int j = 0; j < n; j++) {
            b[j] = std::conj(chirp[j]);
            if (j) b[m - j] = std::conj(chirp[j]);
        }
        host_fft(b);
        std::vector<std::complex<double>> bbr((size_t)m);
        for (int i = 0; i < m; i++) {
            unsigned r = 0;
            for (int k = 0; k < lg; k++) r |= ((i >> k) & 1u) << (lg - 1 - k);
            bbr[i] = b[r];
        }
        b.swap(bbr);
    }
    OICS_TRY(pl.chirp.ensure(sizeof(double2) * (size_t)n));
    OICS_TRY(pl.bbr.ensure(sizeof(double2) * (size_t)m));
    OICS_TRY(pl.tw.ensure(sizeof(double2) * tw.size()));
    OICS_CUDA_TRY(cudaMemcpy(pl.chirp.p, chirp.data(), sizeof(double2) * (size_t)n, cudaMemcpyHostToDevice));
    OICS_CUDA_TRY(cudaMemcpy(pl.bbr.p, b.data(), sizeof(double2) * (size_t)m, cudaMemcpyHostToDevice));
    OICS_CUDA_TRY(cudaMemcpy(pl.tw.p, tw.data(), sizeof(double2) * tw.size(), cudaMemcpyHostToDevice));
    pl.n = n;
    pl.m = m;
    pl.direct = pow2 ? 1 : 0;
    return OICS_OK;
}

static PlanDev plan_dev(const FftPlan &pl)
{
    PlanDev P;
    P.n = pl.n;
    P.m = pl.m;
    P.direct = pl.direct;
    P.log2m = 0;
    while ((1 << P.log2m) < pl.m) P.log2m++;
    P.chirp = pl.chirp.as<double2>();
    P.bbr = pl.bbr.as<double2>();
    P.tw = pl.tw.as<double2>();
    return P;
}

int32_t fft_image_device(oics_ctx *ctx, const uint8_t *dgray, int w, int h, size_t stride, uint8_t *dmag8,
                         uint8_t *dlog8)
{
    cudaStream_t st = ctx->stream;
    OICS_TRY(build_plan(ctx, ctx->fft_row, w));
    OICS_TRY(build_plan(ctx, ctx->fft_col, h));
    const size_t n = (size_t)w * h;
    OICS_TRY(ctx->cplx.ensure(n * sizeof(double2)));
    OICS_TRY(ctx->fmag.ensure(n * sizeof(float) + 64));
    double2 *spec = ctx->cplx.as<double2>();
    uint32_t *gmm = ctx->fmag.as<uint32_t>();  // [min, max, min2, max2] keys
    float *mag = reinterpret_cast<float *>(ctx->fmag.as<uint8_t>() + 64);
    const uint32_t init[4] = {0xffffffffu, 0u, 0xffffffffu, 0u};
    OICS_CUDA_TRY(cudaMemcpyAsync(gmm, init, sizeof(init), cudaMemcpyHostToDevice, st));
    const PlanDev Pr = plan_dev(ctx->fft_row), Pc = plan_dev(ctx->fft_col);
    const size_t smr = (size_t)Pr.m * sizeof(double2), smc = (size_t)Pc.m * sizeof(double2);
    OICS_CUDA_TRY(cudaFuncSetAttribute(fft_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max(smr, (size_t)1024)));
    OICS_CUDA_TRY(cudaFuncSetAttribute(fft_cols_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max(smc, (size_t)1024)));
    {
        ProfScope ps(ctx, "fft_rows_kernel", (double)n * (1 + sizeof(double2)));
        fft_rows_kernel<<<h, FFT_THREADS, smr, st>>>(dgray, stride, w, Pr, spec);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    {
        ProfScope ps(ctx, "fft_cols_kernel", (double)n * (sizeof(double2) + 4));
        fft_cols_kernel<<<w, FFT_THREADS, smc, st>>>(spec, w, h, Pc, mag, gmm);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    const int blocks = ctx->sm_count * 8;
    {
        ProfScope ps(ctx, "fft_log_kernel", (double)n * 8);
        fft_log_kernel<<<blocks, 256, 0, st>>>(mag, n, gmm, dmag8, gmm + 2);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    {
        ProfScope ps(ctx, "fft_log8_kernel", (double)n * 5);
        fft_log8_kernel<<<blocks, 256, 0, st>>>(mag, n, gmm + 2, dlog8);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 4;
    // `init` is a stack temporary consumed by the async copy above
    OICS_CUDA_TRY(cudaStreamSynchronize(st));
    return OICS_OK;
}

}  // namespace oics
