// Bandwidth probes behind bench.py's roofline block.  SURVEY 8(d): the projection sweep
// (omr.rs:147-221) re-reads one card 2N times, so its memory ceiling is the L2 / shared-memory
// read rate and not the HBM copy rate; MEASURED_PEAKS.json holds no L2 figure, so the library
// measures one on the box it runs on.  Not on the product path: nothing calls these but bench.py.
#include "common.h"

namespace oics {

__device__ __forceinline__ uint4 ld_cg_v4(const uint4 *p)
{
    uint4 v;
    // volatile: one real L2 transaction per call, never hoisted out of the repetition loop
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// Every CTA streams the whole buffer `reps` times with 16-byte L1-bypassing loads, four in
// flight per thread.  n16 is a multiple of 4 * blockDim.x * gridDim.x.
__global__ void __launch_bounds__(256) l2_read_kernel(const uint4 *__restrict__ buf, size_t n16, int reps, unsigned *sink)
{
    const size_t span = (size_t)gridDim.x * blockDim.x;
    unsigned acc = 0;
    for (int r = 0; r < reps; ++r) {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += 4 * span) {
            uint4 a = ld_cg_v4(buf + i), b = ld_cg_v4(buf + i + span), c = ld_cg_v4(buf + i + 2 * span),
                  d = ld_cg_v4(buf + i + 3 * span);
            acc ^= a.x ^ a.y ^ a.z ^ a.w ^ b.x ^ b.y ^ b.z ^ b.w ^ c.x ^ c.y ^ c.z ^ c.w ^ d.x ^ d.y ^ d.z ^ d.w;
        }
    }
    if (acc == 0x9e3779b9u) *sink = acc;  // never true for the zero-filled buffer; keeps the loads live
}

// Every thread re-reads its CTA's shared-memory tile with conflict-free 16-byte loads.
__global__ void __launch_bounds__(256) smem_read_kernel(int iters, unsigned *sink)
{
    extern __shared__ uint4 tile[];
    const int n16 = 4096;  // 64 KB
    for (int i = threadIdx.x; i < n16; i += blockDim.x) tile[i] = make_uint4(i, 0, 0, 0);
    __syncthreads();
    unsigned acc = 0;
    int at = threadIdx.x;
    const unsigned base = (unsigned)__cvta_generic_to_shared(tile);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            uint4 v;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                         : "r"(base + (unsigned)(((at + it + k * 256) & (n16 - 1)) * 16))
                         : "memory");  // the address moves with `it`: two iterations never share a load
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    }
    if (acc == 0x9e3779b9u) *sink = acc;
}

}  // namespace oics

using namespace oics;

extern "C" {

OICS_API int32_t oics_probe_l2_read(oics_ctx *ctx, size_t bytes, int32_t reps, double *gb_per_s)
{
    OICS_ARG(ctx && gb_per_s && reps > 0 && reps <= 4096);
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    const int threads = 256, grid = ctx->sm_count * 8;
    const size_t quantum = (size_t)4 * threads * grid * sizeof(uint4);
    OICS_ARG(bytes >= quantum && bytes <= ((size_t)1 << 32));
    const size_t n16 = bytes / quantum * quantum / sizeof(uint4);
    OICS_TRY(ctx->misc.ensure(256));
    DevBuf buf;
    OICS_TRY(buf.ensure(n16 * sizeof(uint4)));
    cudaError_t e = cudaMemsetAsync(buf.p, 0, n16 * sizeof(uint4), ctx->stream);
    float ms = 0.f;
    if (e == cudaSuccess) {
        l2_read_kernel<<<grid, threads, 0, ctx->stream>>>(buf.as<uint4>(), n16, 2, ctx->misc.as<unsigned>());  // warm L2
        cudaEventRecord(ctx->ev0, ctx->stream);
        l2_read_kernel<<<grid, threads, 0, ctx->stream>>>(buf.as<uint4>(), n16, reps, ctx->misc.as<unsigned>());
        cudaEventRecord(ctx->ev1, ctx->stream);
        ctx->launches += 2;
        e = cudaStreamSynchronize(ctx->stream);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    }
    buf.release();
    OICS_CUDA_TRY(e);
    *gb_per_s = (double)n16 * sizeof(uint4) * reps / (ms * 1e-3) / 1e9;
    return OICS_OK;
}

OICS_API int32_t oics_probe_smem_read(oics_ctx *ctx, int32_t iters, double *gb_per_s)
{
    OICS_ARG(ctx && gb_per_s && iters > 0 && iters <= (1 << 20));
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    const int threads = 256, grid = ctx->sm_count * 3, smem = 65536;
    OICS_TRY(ctx->misc.ensure(256));
    OICS_CUDA_TRY(cudaFuncSetAttribute(smem_read_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    float ms = 0.f;
    smem_read_kernel<<<grid, threads, smem, ctx->stream>>>(16, ctx->misc.as<unsigned>());
    cudaEventRecord(ctx->ev0, ctx->stream);
    smem_read_kernel<<<grid, threads, smem, ctx->stream>>>(iters, ctx->misc.as<unsigned>());
    cudaEventRecord(ctx->ev1, ctx->stream);
    ctx->launches += 2;
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    OICS_CUDA_TRY(cudaGetLastError());
    OICS_CUDA_TRY(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    *gb_per_s = (double)grid * threads * 16.0 * 16.0 * iters / (ms * 1e-3) / 1e9;
    return OICS_OK;
}

}  // extern "C"
