// oics_projection_standard_deviations_static: transfer::get_projection_standard_deviations
// (transfer.rs:527-536) of a device-resident image as is, for callers that may not allocate
// (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  Nothing in here calls cudaMalloc /
// cudaFree or any allocating routine: the per-row and per-column zero counts live in ONE file-scope
// device array of a fixed maximum size that every GPU loading the library holds until exit.
//
//   limit   : w <= 16384 and h <= 16384  ((16384 + 16384) * 4 B = 128 KB of device memory per GPU)
//   sharing : every call on a device uses the same array; it is cleared at the start of each call, so
//             calls follow each other freely on ONE stream, but two calls that overlap on the GPU
//             (different streams, different host threads) race on it and return garbage.
//
// The arithmetic is that of the ordinary entry point (sweep.cu: zero counts, then the population
// standard deviation with the squares summed strictly in index order in f64, calculate.rs order), so the
// two agree bit for bit.
#include <mutex>

#include "common.h"

namespace oics {
namespace {

constexpr int NA_MAX_W = 16384;
constexpr int NA_MAX_H = 16384;
constexpr int NA_ROWS = 64;     // rows per CTA of the counting kernel
constexpr int NA_THREADS = 256; // 4 columns per thread -> 1024 columns per CTA
constexpr int NA_CHUNK = 256;   // squares staged per trip of the sequential sum
constexpr int NA_MAX_DEVICES = 64;

// [0, h): zero pixels per row; [h, h + w): zero pixels per column.  In HOST code this name is only a
// stand-in for the symbol -- the address kernels get comes from cudaGetSymbolAddress (scratch_for_device).
static __device__ __align__(256) int32_t g_na_counts[NA_MAX_H + NA_MAX_W];

std::mutex g_na_mutex;
int32_t *g_na_ptr[NA_MAX_DEVICES];   // device address of g_na_counts on each device, filled on first use

// Set-up, once per device, with that device current.
int32_t scratch_for_device(int device, int32_t **out)
{
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_ptr[device]) {
        int32_t *p = nullptr;
        OICS_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_na_counts));
        if (!p) {
            set_error("cudaGetSymbolAddress returned a null address for the static scratch array");
            return OICS_ERR_CUDA;
        }
        g_na_ptr[device] = p;
    }
    *out = g_na_ptr[device];
    return OICS_OK;
}

__global__ void na_clear_kernel(int32_t *__restrict__ counts, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) counts[i] = 0;
}

// One CTA: NA_ROWS rows x 1024 columns.  Each thread owns 4 neighbouring columns (one 32-bit load per
// row where the address allows), keeps their column counts in registers, and the row counts are reduced
// per warp, then per CTA in shared memory: one global atomic per row and per column per CTA.
__global__ void __launch_bounds__(NA_THREADS) na_counts_kernel(const uint8_t *__restrict__ bin, size_t stride, int w,
                                                               int h, int32_t *__restrict__ counts)
{
    __shared__ int row_cnt[NA_ROWS];
    const int lane = threadIdx.x & 31;
    const int x0 = (blockIdx.x * NA_THREADS + threadIdx.x) * 4;
    const int y0 = blockIdx.y * NA_ROWS;
    const int y1 = min(h, y0 + NA_ROWS);
    if (threadIdx.x < NA_ROWS) row_cnt[threadIdx.x] = 0;
    __syncthreads();
    const bool word_ok = x0 + 4 <= w && ((reinterpret_cast<uintptr_t>(bin) | stride) & 3) == 0;
    uint32_t col = 0;   // four byte lanes: NA_ROWS <= 255 keeps them from carrying into each other
    for (int y = y0; y < y1; y++) {
        const uint8_t *row = bin + (size_t)y * stride;
        uint32_t z = 0;   // 0x01 in byte k where pixel (x0 + k, y) is zero
        if (word_ok) {
            z = __vcmpeq4(__ldg(reinterpret_cast<const uint32_t *>(row + x0)), 0u) & 0x01010101u;
        } else {
            for (int k = 0; k < 4; k++)
                if (x0 + k < w && row[x0 + k] == 0) z |= 1u << (8 * k);
        }
        col += z;
        int r = __popc(z);
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
        if (lane == 0 && r) atomicAdd(&row_cnt[y - y0], r);
    }
    static_assert(NA_ROWS <= 255, "packed byte column counts");
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int c = (col >> (8 * k)) & 0xFF;
        if (c && x0 + k < w) atomicAdd(counts + h + x0 + k, c);
    }
    __syncthreads();
    if (threadIdx.x < y1 - y0 && row_cnt[threadIdx.x]) atomicAdd(counts + y0 + threadIdx.x, row_cnt[threadIdx.x]);
}

// Population standard deviation of v[0..n) in the reference's order (calculate.rs): mean = sum / n, the
// squares of the deviations added one after the other in index order, sqrt(sum / n).  One warp.
__device__ double na_seq_std_dev(const int32_t *__restrict__ v, int n, double *stage, int lane)
{
    long long s = 0;
    for (int i = lane; i < n; i += 32) s += v[i];
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const double mean = __ddiv_rn((double)s, (double)n);
    double acc = 0.0;
    for (int base = 0; base < n; base += NA_CHUNK) {
        const int m = min(NA_CHUNK, n - base);
        for (int i = lane; i < m; i += 32) {
            const double d = __dsub_rn((double)v[base + i], mean);
            stage[i] = __dmul_rn(d, d);
        }
        __syncwarp();
        if (lane == 0) {
            int i = 0;
            if (base == 0) { acc = stage[0]; i = 1; }
            for (; i < m; i++) acc = __dadd_rn(acc, stage[i]);
        }
        __syncwarp();
    }
    return __dsqrt_rn(__ddiv_rn(acc, (double)n));
}

// warp 0: rows -> sd[1] (horizontal projection), warp 1: columns -> sd[0] (vertical projection)
__global__ void __launch_bounds__(64) na_finalize_kernel(const int32_t *__restrict__ counts, int w, int h,
                                                         double *__restrict__ sd)
{
    __shared__ double stage[2][NA_CHUNK];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const double r = warp == 0 ? na_seq_std_dev(counts, h, stage[0], lane) : na_seq_std_dev(counts + h, w, stage[1], lane);
    if (lane == 0) sd[warp == 0 ? 1 : 0] = r;
}

}  // namespace
}  // namespace oics

using namespace oics;

OICS_API int32_t oics_projection_standard_deviations_static(const uint8_t *d_bin, int32_t w, int32_t h, int32_t stride,
                                                            double *d_sd, void *cuda_stream)
{
    if (!d_bin || !d_sd || w <= 0 || h <= 0 || stride < w) {
        set_error("invalid argument: device image, 0 < w <= stride, 0 < h and a device double[2] are required");
        return OICS_ERR_INVALID_ARG;
    }
    if (w > NA_MAX_W || h > NA_MAX_H) {
        set_error("image %d x %d is beyond the static scratch array of this entry point (at most %d x %d); "
                  "use oics_projection_standard_deviations",
                  w, h, NA_MAX_W, NA_MAX_H);
        return OICS_ERR_CAPACITY;
    }
    int device = -1;
    OICS_CUDA_TRY(cudaGetDevice(&device));
    if (device < 0 || device >= NA_MAX_DEVICES) {
        set_error("device ordinal %d is outside the %d this entry point keeps a scratch address for", device, NA_MAX_DEVICES);
        return OICS_ERR_UNSUPPORTED;
    }
    cudaPointerAttributes ai{}, ao{};
    OICS_CUDA_TRY(cudaPointerGetAttributes(&ai, d_bin));
    OICS_CUDA_TRY(cudaPointerGetAttributes(&ao, d_sd));
    if (ai.type != cudaMemoryTypeDevice || ao.type != cudaMemoryTypeDevice || ai.device != device || ao.device != device) {
        set_error("the image and the result must be device memory of the CURRENT device (%d): this entry point "
                  "stages nothing", device);
        return OICS_ERR_INVALID_ARG;
    }
    int32_t *counts = nullptr;
    OICS_TRY(scratch_for_device(device, &counts));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const int n = h + w;
    na_clear_kernel<<<cdiv(n, 256), 256, 0, st>>>(counts, n);   // the array keeps the last call's counts
    OICS_CUDA_TRY(cudaGetLastError());
    dim3 grid(cdiv(w, NA_THREADS * 4), cdiv(h, NA_ROWS));
    na_counts_kernel<<<grid, NA_THREADS, 0, st>>>(d_bin, (size_t)stride, w, h, counts);
    OICS_CUDA_TRY(cudaGetLastError());
    na_finalize_kernel<<<1, 64, 0, st>>>(counts, w, h, d_sd);
    OICS_CUDA_TRY(cudaGetLastError());
    return OICS_OK;
}
