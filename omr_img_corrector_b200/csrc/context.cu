// Context, error and staging plumbing behind the C ABI (include/oics_cuda.h).
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <map>

#include "common.h"

namespace oics {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int32_t DevBuf::ensure(size_t bytes)
{
    if (bytes <= cap && p) return OICS_OK;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
        p = nullptr;
        set_error("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
        return e == cudaErrorMemoryAllocation ? OICS_ERR_NOMEM : OICS_ERR_CUDA;
    }
    cap = want;
    return OICS_OK;
}

void DevBuf::release()
{
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
}

int32_t PinnedBuf::ensure(size_t bytes)
{
    if (bytes <= cap && p) return OICS_OK;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMallocHost(&p, want);
    if (e != cudaSuccess) {
        p = nullptr;
        set_error("cudaMallocHost(%zu) failed: %s", want, cudaGetErrorString(e));
        return OICS_ERR_NOMEM;
    }
    cap = want;
    return OICS_OK;
}

void PinnedBuf::release()
{
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
}

bool is_device_ptr(const void *p)
{
    if (!p) return false;
    cudaPointerAttributes a;
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

ProfScope::ProfScope(oics_ctx *c, const char *name, double bytes, cudaStream_t stream) : ctx(c), st(stream ? stream : c->stream)
{
    if (!ctx->prof.on) return;
    Profiler &P = ctx->prof;
    cudaEvent_t ev[2] = {nullptr, nullptr};
    for (cudaEvent_t &e : ev) {
        if (!P.pool.empty()) {
            e = P.pool.back();
            P.pool.pop_back();
        } else if (cudaEventCreate(&e) != cudaSuccess) {
            cudaGetLastError();
            return;
        }
    }
    if (cudaEventRecord(ev[0], st) != cudaSuccess) {
        cudaGetLastError();
        return;
    }
    idx = (int)P.entries.size();
    P.entries.push_back({name, ev[0], ev[1], bytes});
}

ProfScope::~ProfScope()
{
    if (idx < 0) return;
    if (cudaEventRecord(ctx->prof.entries[(size_t)idx].e1, st) != cudaSuccess) cudaGetLastError();
}

CtxGuard::~CtxGuard()
{
    if (ranged) nvtxRangePop();
}

CtxGuard::CtxGuard(oics_ctx *ctx, const char *entry)
{
    status = OICS_OK;
    if (entry) {
        nvtxRangePushA(entry);
        ranged = true;
    }
    if (!ctx) {
        set_error("null context");
        status = OICS_ERR_INVALID_ARG;
        return;
    }
    lk = std::unique_lock<std::mutex>(ctx->mu);
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e != cudaSuccess) {
        set_error("cudaSetDevice(%d) failed: %s", ctx->device, cudaGetErrorString(e));
        status = OICS_ERR_CUDA;
    }
}

int32_t stage_in(oics_ctx *ctx, const void *src, size_t row_bytes, int h, size_t stride, DevBuf &scratch,
                 const uint8_t **dptr, size_t *dstride)
{
    if (is_device_ptr(src)) {
        *dptr = static_cast<const uint8_t *>(src);
        *dstride = stride;
        return OICS_OK;
    }
    OICS_TRY(scratch.ensure(row_bytes * (size_t)h));
    OICS_CUDA_TRY(cudaMemcpy2DAsync(scratch.p, row_bytes, src, stride, row_bytes, (size_t)h, cudaMemcpyHostToDevice,
                                    ctx->stream));
    *dptr = scratch.as<uint8_t>();
    *dstride = row_bytes;
    return OICS_OK;
}

int32_t stage_out(oics_ctx *ctx, const void *dsrc, size_t row_bytes, int h, void *dst, size_t dst_stride)
{
    cudaMemcpyKind kind = is_device_ptr(dst) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    OICS_CUDA_TRY(cudaMemcpy2DAsync(dst, dst_stride, dsrc, row_bytes, row_bytes, (size_t)h, kind, ctx->stream));
    return OICS_OK;
}

int32_t copy_out(oics_ctx *ctx, void *dst, const void *dsrc, size_t bytes)
{
    cudaMemcpyKind kind = is_device_ptr(dst) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    OICS_CUDA_TRY(cudaMemcpyAsync(dst, dsrc, bytes, kind, ctx->stream));
    return OICS_OK;
}

int32_t copy_in(oics_ctx *ctx, void *ddst, const void *src, size_t bytes)
{
    cudaMemcpyKind kind = is_device_ptr(src) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    OICS_CUDA_TRY(cudaMemcpyAsync(ddst, src, bytes, kind, ctx->stream));
    return OICS_OK;
}

}  // namespace oics

using namespace oics;

extern "C" {

OICS_API const char *oics_last_error(void) { return g_err; }

OICS_API int32_t oics_version(void) { return 100; }

// The batch driver keeps up to 32 decode streams busy (jpegdec.cu); the CUDA default of 8 hardware work queues
// makes them alias.  The variable is read once, at CUDA initialisation: set here, when the library is loaded,
// unless the host exported its own value (no effect if the process initialised CUDA earlier).
__attribute__((constructor)) static void oics_on_load(void) { setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0); }

OICS_API int32_t oics_create(int32_t device, oics_ctx **out)
{
    if (!out) {
        set_error("out is null");
        return OICS_ERR_INVALID_ARG;
    }
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        set_error("no CUDA device available (%s); this library has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
        return OICS_ERR_CUDA;
    }
    if (device < 0 || device >= n) {
        set_error("device %d out of range (have %d)", device, n);
        return OICS_ERR_INVALID_ARG;
    }
    OICS_CUDA_TRY(cudaSetDevice(device));
    oics_ctx *ctx = new oics_ctx();
    ctx->device = device;
    cudaDeviceProp prop;
    OICS_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = (int)prop.sharedMemPerBlockOptin;
    OICS_CUDA_TRY(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
    ctx->stream = ctx->own_stream;
    OICS_CUDA_TRY(cudaEventCreate(&ctx->ev0));
    OICS_CUDA_TRY(cudaEventCreate(&ctx->ev1));
    *out = ctx;
    return OICS_OK;
}

OICS_API void oics_destroy(oics_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    DevBuf *bufs[] = {&ctx->in_img, &ctx->in_img2, &ctx->out_img, &ctx->out_img2, &ctx->gray,  &ctx->tmp8a, &ctx->tmp8b,
                      &ctx->bin,    &ctx->bits,    &ctx->mats,    &ctx->tabs,  &ctx->counts, &ctx->sd,
                      &ctx->sums,   &ctx->edges,   &ctx->mag,     &ctx->dxy,   &ctx->labels, &ctx->accum,
                      &ctx->lines,  &ctx->misc,    &ctx->cplx,    &ctx->cplx2, &ctx->twid,   &ctx->fmag,
                      &ctx->wplanes, &ctx->wmeta,  &ctx->wlists, &ctx->jpeg_raw, &ctx->jpeg_hdr, &ctx->jpeg_totals,
                      &ctx->fb_in,  &ctx->fb_out};
    for (DevBuf *b : bufs) b->release();
    jpeg_decode_release(ctx);
    jpeg_lanes_release(ctx);
    for (FftPlan *pl : {&ctx->fft_row, &ctx->fft_col}) {
        pl->chirp.release();
        pl->bbr.release();
        pl->tw.release();
    }
    ctx->pin_in.release();
    ctx->pin_out.release();
    ctx->pin_small.release();
    for (ProfEntry &e : ctx->prof.entries) {
        cudaEventDestroy(e.e0);
        cudaEventDestroy(e.e1);
    }
    for (cudaEvent_t e : ctx->prof.pool) cudaEventDestroy(e);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

OICS_API int32_t oics_set_stream(oics_ctx *ctx, void *cuda_stream)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
    return OICS_OK;
}

OICS_API int32_t oics_synchronize(oics_ctx *ctx)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

OICS_API int64_t oics_launch_count(oics_ctx *ctx) { return ctx ? ctx->launches : 0; }

OICS_API int32_t oics_alloc_pinned(size_t bytes, void **out)
{
    if (!out || bytes == 0) {
        set_error("invalid argument: out, bytes");
        return OICS_ERR_INVALID_ARG;
    }
    *out = nullptr;
    cudaError_t e = cudaMallocHost(out, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *out = nullptr;
        set_error("cudaMallocHost(%zu) failed: %s", bytes, cudaGetErrorString(e));
        return e == cudaErrorMemoryAllocation ? OICS_ERR_NOMEM : OICS_ERR_CUDA;
    }
    return OICS_OK;
}

OICS_API void oics_free_pinned(void *p)
{
    if (p) cudaFreeHost(p);
}

OICS_API int32_t oics_profile(oics_ctx *ctx, int32_t enable)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_CUDA_TRY(cudaDeviceSynchronize());
    for (ProfEntry &e : ctx->prof.entries) {
        ctx->prof.pool.push_back(e.e0);
        ctx->prof.pool.push_back(e.e1);
    }
    ctx->prof.entries.clear();
    ctx->prof.on = enable != 0;
    return OICS_OK;
}

OICS_API int32_t oics_profile_report(oics_ctx *ctx, char *buf, size_t cap, size_t *needed)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_ARG(needed && (cap == 0 || buf));
    OICS_CUDA_TRY(cudaDeviceSynchronize());
    struct Agg { double ms = 0, bytes = 0; long long n = 0; };
    std::map<std::string, Agg> agg;
    std::vector<std::string> order;
    for (const ProfEntry &e : ctx->prof.entries) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, e.e0, e.e1) != cudaSuccess) {
            cudaGetLastError();
            continue;
        }
        if (!agg.count(e.name)) order.push_back(e.name);
        Agg &a = agg[e.name];
        a.ms += t;
        a.bytes += e.bytes;
        a.n++;
    }
    // OICS_PROFILE_TIMELINE=<prefix>: every launch of the recording as "kernel,start_us,end_us" (relative to the
    // first launch) into <prefix>.<n>.csv -- the step's timeline across the context's streams
    if (const char *prefix = getenv("OICS_PROFILE_TIMELINE")) {
        static int dumps = 0;
        if (!ctx->prof.entries.empty()) {
            const std::string path = std::string(prefix) + "." + std::to_string(dumps++) + ".csv";
            if (FILE *f = fopen(path.c_str(), "w")) {
                const cudaEvent_t base = ctx->prof.entries.front().e0;
                for (const ProfEntry &e : ctx->prof.entries) {
                    float t0 = 0.f, t1 = 0.f;
                    if (cudaEventElapsedTime(&t0, base, e.e0) != cudaSuccess || cudaEventElapsedTime(&t1, base, e.e1) != cudaSuccess) {
                        cudaGetLastError();
                        continue;
                    }
                    fprintf(f, "\"%s\",%.1f,%.1f\n", e.name, 1000.0 * t0, 1000.0 * t1);
                }
                fclose(f);
            }
        }
    }
    std::string out = "[";
    char line[512];
    for (size_t i = 0; i < order.size(); i++) {
        const Agg &a = agg[order[i]];
        snprintf(line, sizeof(line), "%s{\"kernel\": \"%s\", \"launches\": %lld, \"ms\": %.6f, \"algorithmic_bytes\": %.0f}",
                 i ? ", " : "", order[i].c_str(), a.n, a.ms, a.bytes);
        out += line;
    }
    out += "]";
    *needed = out.size() + 1;
    if (cap < out.size() + 1) {
        set_error("profile report needs %zu bytes, capacity %zu", out.size() + 1, cap);
        return OICS_ERR_CAPACITY;
    }
    memcpy(buf, out.c_str(), out.size() + 1);
    return OICS_OK;
}

OICS_API int32_t oics_last_sweep_kernel_ms(oics_ctx *ctx, double *ms, int32_t *launches)
{
    if (!ctx) {
        set_error("null context");
        return OICS_ERR_INVALID_ARG;
    }
    if (ms) *ms = ctx->last_sweep_ms;
    if (launches) *launches = ctx->last_sweep_launches;
    return OICS_OK;
}

}  // extern "C"
