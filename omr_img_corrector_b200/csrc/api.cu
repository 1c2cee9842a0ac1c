// The compute entry points of include/oics_cuda.h: argument checks, host/device
// staging, and the reference's call sequences expressed over the device kernels.
#include <chrono>
#include <cmath>
#include <cstdlib>

#include "common.h"

namespace oics {
int32_t plain_std_devs_device(oics_ctx *ctx, const uint8_t *dbin, int w, int h, size_t stride, double *d_vsd,
                              double *d_hsd);

static int32_t check_image(const void *p, int w, int h, int stride, int ch)
{
    if (!p) { set_error("image pointer is null"); return OICS_ERR_INVALID_ARG; }
    if (w <= 0 || h <= 0) { set_error("image size %dx%d is empty", w, h); return OICS_ERR_INVALID_ARG; }
    if (stride < w * ch) { set_error("stride %d smaller than a row (%d bytes)", stride, w * ch); return OICS_ERR_INVALID_ARG; }
    return OICS_OK;
}

// (int)((size.width as f64) * scale) of omr.rs:120-121 / transfer.rs:77-78
static int scaled_dim(int v, double scale) { return (int)((double)v * scale); }

// sweep over device-resident binary card(s); results to host or device pointers
static int32_t sweep_to(oics_ctx *ctx, const uint8_t *dbins, int n_cards, size_t card_stride, int w, int h,
                        size_t stride, const double *M, int n_angles, double *v_sd, double *h_sd, int64_t *sums,
                        int32_t *counts_out_host_or_dev)
{
    const size_t n = (size_t)n_cards * n_angles;
    OICS_TRY(ctx->sd.ensure(2 * n * sizeof(double)));
    double *dv = ctx->sd.as<double>(), *dh = dv + n;
    int64_t *dsums = nullptr;
    if (sums) {
        OICS_TRY(ctx->sums.ensure(4 * n * sizeof(int64_t)));
        dsums = ctx->sums.as<int64_t>();
    }
    int32_t *dcounts = nullptr;
    if (counts_out_host_or_dev) {
        OICS_TRY(ctx->misc.ensure((size_t)n_angles * (h + w) * sizeof(int32_t)));
        dcounts = ctx->misc.as<int32_t>();
    }
    OICS_TRY(sweep_device(ctx, dbins, n_cards, card_stride, w, h, stride, M, n_angles, dv, dh, dsums, dcounts));
    OICS_TRY(copy_out(ctx, v_sd, dv, n * sizeof(double)));
    OICS_TRY(copy_out(ctx, h_sd, dh, n * sizeof(double)));
    if (sums) OICS_TRY(copy_out(ctx, sums, dsums, 4 * n * sizeof(int64_t)));
    if (dcounts) OICS_TRY(copy_out(ctx, counts_out_host_or_dev, dcounts, (size_t)n_angles * (h + w) * sizeof(int32_t)));
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

// P0 + sweep for one BGR card already on the device -> host std-dev arrays
static int32_t estimate_projection(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, int mode,
                                   double scale, uint16_t max_angle, double step, double rot_scale,
                                   std::vector<double> *v_sd, std::vector<double> *h_sd, int *N_out)
{
    const int ow = scale == 1.0 ? w : scaled_dim(w, scale), oh = scale == 1.0 ? h : scaled_dim(h, scale);
    if (ow <= 0 || oh <= 0) { set_error("scaled size %dx%d is empty", ow, oh); return OICS_ERR_INVALID_ARG; }
    OICS_TRY(ctx->bin.ensure((size_t)ow * oh));
    OICS_TRY(preprocess_device(ctx, dbgr, w, h, stride, mode, ow, oh, ctx->bin.as<uint8_t>()));
    const int N = oics_angle_half_range(max_angle, step);
    *N_out = N;
    v_sd->assign((size_t)2 * N, 0.0);
    h_sd->assign((size_t)2 * N, 0.0);
    if (N == 0) return OICS_OK;
    std::vector<double> M((size_t)2 * N * 6);
    int n_angles = 0;
    OICS_TRY(oics_sweep_matrices(ow, oh, max_angle, step, rot_scale, M.data(), 2 * N, &n_angles));
    return sweep_to(ctx, ctx->bin.as<uint8_t>(), 1, 0, ow, oh, (size_t)ow, M.data(), n_angles, v_sd->data(),
                    h_sd->data(), nullptr, nullptr);
}

// Canny -> HoughLinesP -> segment-angle rule for n equally sized device images AT ONCE (the reference runs one
// OS thread per card, thread_pool.rs:41-73: here the cards' PPHTs run as one launch, a cluster per card).
// results[i] = {angle, status, n_candidates} (no candidate list), statuses[i] = OICS_OK or that image's error.
static int32_t edges_results_batch(oics_ctx *ctx, const uint8_t *const *dsrcs, int n, int ch, int w, int h, size_t stride,
                                   double lo, double hi, int threshold, double min_len, double max_gap, int variant,
                                   std::vector<oics_omr_result> *results, std::vector<int32_t> *statuses)
{
    results->assign((size_t)n, oics_omr_result{0, 0, 0, nullptr, 0});
    statuses->assign((size_t)n, OICS_OK);
    if (n <= 0) return OICS_OK;
    const size_t npx = (size_t)w * h;
    OICS_TRY(ctx->edges.ensure((size_t)n * npx));
    if (lo > hi) std::swap(lo, hi);
    std::vector<const uint8_t *> edges((size_t)n);
    for (int i = 0; i < n; i++) {
        edges[(size_t)i] = ctx->edges.as<uint8_t>() + (size_t)i * npx;
        OICS_TRY(canny_device(ctx, dsrcs[i], ch, w, h, stride, (int)std::floor(lo), (int)std::floor(hi),
                              ctx->edges.as<uint8_t>() + (size_t)i * npx));
    }
    std::vector<std::vector<int32_t>> lines;
    OICS_TRY(hough_lines_p_batch_device(ctx, edges.data(), n, w, h, (size_t)w, 1.0, 3.14159265358979323846 / 180.0, threshold,
                                        (int)std::lrint(min_len), (int)std::lrint(max_gap), &lines));
    for (int i = 0; i < n; i++)
        (*statuses)[(size_t)i] = oics_lines_angle(lines[(size_t)i].data(), (int)(lines[(size_t)i].size() / 4), variant, &(*results)[(size_t)i]);
    return OICS_OK;
}

static int32_t edges_result(oics_ctx *ctx, const uint8_t *dsrc, int ch, int w, int h, size_t stride, double lo,
                            double hi, int threshold, double min_len, double max_gap, int variant,
                            oics_omr_result *result)
{
    OICS_TRY(ctx->edges.ensure((size_t)w * h));
    if (lo > hi) std::swap(lo, hi);
    OICS_TRY(canny_device(ctx, dsrc, ch, w, h, stride, (int)std::floor(lo), (int)std::floor(hi), ctx->edges.as<uint8_t>()));
    std::vector<std::vector<int32_t>> lines;
    const uint8_t *e = ctx->edges.as<uint8_t>();
    OICS_TRY(hough_lines_p_batch_device(ctx, &e, 1, w, h, (size_t)w, 1.0, 3.14159265358979323846 / 180.0, threshold,
                                        (int)std::lrint(min_len), (int)std::lrint(max_gap), &lines));
    return oics_lines_angle(lines[0].data(), (int)(lines[0].size() / 4), variant, result);
}

}  // namespace oics

using namespace oics;

extern "C" {

OICS_API int32_t oics_preprocess(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h, int32_t stride,
                                 int32_t mode, int32_t max_w, int32_t max_h, double resize_scale, uint8_t *out_bin,
                                 size_t out_cap, int32_t *out_w, int32_t *out_h, double *out_scale)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bgr, w, h, stride, 3));
    OICS_ARG(out_bin && out_w && out_h);
    OICS_ARG(mode == OICS_PRE_OMR || mode == OICS_PRE_PROJECTION);
    const double scale = mode == OICS_PRE_OMR ? oics_projection_scale(w, h, max_w, max_h) : resize_scale;
    // omr.rs always calls resize (identity size = copy); scale_self early-outs on 1.0
    const int ow = scale == 1.0 ? w : scaled_dim(w, scale), oh = scale == 1.0 ? h : scaled_dim(h, scale);
    *out_w = ow;
    *out_h = oh;
    if (out_scale) *out_scale = scale;
    if (ow <= 0 || oh <= 0) { set_error("scaled size %dx%d is empty", ow, oh); return OICS_ERR_INVALID_ARG; }
    if ((size_t)ow * oh > out_cap) {
        set_error("output needs %zu bytes, capacity %zu", (size_t)ow * oh, out_cap);
        return OICS_ERR_CAPACITY;
    }
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, bgr, (size_t)w * 3, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    uint8_t *dout = is_device_ptr(out_bin) ? out_bin : nullptr;
    if (!dout) {
        OICS_TRY(ctx->bin.ensure((size_t)ow * oh));
        dout = ctx->bin.as<uint8_t>();
    }
    OICS_TRY(preprocess_device(ctx, dsrc, w, h, dstride, mode, ow, oh, dout));
    if (dout != out_bin) OICS_TRY(copy_out(ctx, out_bin, dout, (size_t)ow * oh));
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

OICS_API int32_t oics_projection_sweep(oics_ctx *ctx, const uint8_t *bin, int32_t w, int32_t h, int32_t stride,
                                       const double *M, int32_t n_angles, double *v_sd, double *h_sd, int64_t *sums)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bin, w, h, stride, 1));
    OICS_ARG(n_angles >= 0 && (n_angles == 0 || (M && v_sd && h_sd)));
    if (n_angles == 0) return OICS_OK;
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, bin, (size_t)w, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    return sweep_to(ctx, dsrc, 1, 0, w, h, dstride, M, n_angles, v_sd, h_sd, sums, nullptr);
}

OICS_API int32_t oics_projection_sweep_batch(oics_ctx *ctx, const uint8_t *bins, int32_t n_cards, size_t card_stride,
                                             int32_t w, int32_t h, int32_t stride, const double *M,
                                             int32_t n_angles, double *v_sd, double *h_sd)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bins, w, h, stride, 1));
    OICS_ARG(n_cards >= 0 && n_angles >= 0);
    if (n_cards == 0 || n_angles == 0) return OICS_OK;
    OICS_ARG(M && v_sd && h_sd);
    OICS_ARG(card_stride >= (size_t)stride * (size_t)h);
    const uint8_t *dsrc = bins;
    size_t dcard = card_stride, dstride = (size_t)stride;
    if (!is_device_ptr(bins)) {
        // host batch: upload tightly packed
        OICS_TRY(ctx->in_img.ensure((size_t)n_cards * w * h));
        for (int c = 0; c < n_cards; c++)
            OICS_CUDA_TRY(cudaMemcpy2DAsync(ctx->in_img.as<uint8_t>() + (size_t)c * w * h, (size_t)w,
                                            bins + (size_t)c * card_stride, (size_t)stride, (size_t)w, (size_t)h,
                                            cudaMemcpyHostToDevice, ctx->stream));
        dsrc = ctx->in_img.as<uint8_t>();
        dcard = (size_t)w * h;
        dstride = (size_t)w;
    }
    return sweep_to(ctx, dsrc, n_cards, dcard, w, h, dstride, M, n_angles, v_sd, h_sd, nullptr, nullptr);
}

OICS_API int32_t oics_sweep_sharded(oics_ctx *ctx, const uint8_t *bin, int32_t w, int32_t h, int32_t stride,
                                    const double *M, int32_t n_angles, int32_t rank, int32_t world, double *sd,
                                    int32_t *first, int32_t *count)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bin, w, h, stride, 1));
    OICS_ARG(n_angles >= 0 && world > 0 && rank >= 0 && rank < world && (n_angles == 0 || (M && sd)));
    int32_t f = 0, c = 0, per = 0;
    OICS_TRY(oics_shard_range(n_angles, rank, world, &f, &c));
    OICS_TRY(oics_shard_layout(n_angles, world, &per));
    if (first) *first = f;
    if (count) *count = c;
    if (n_angles == 0) return OICS_OK;
    double *blk = sd + (size_t)rank * 2 * per;   // this rank's block: v_sd slice, then h_sd slice
    // the padding of the block travels through the all-gather too: define it
    if (is_device_ptr(sd)) OICS_CUDA_TRY(cudaMemsetAsync(blk, 0, (size_t)2 * per * sizeof(double), ctx->stream));
    else memset(blk, 0, (size_t)2 * per * sizeof(double));
    if (c == 0) {
        OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        return OICS_OK;
    }
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, bin, (size_t)w, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    return sweep_to(ctx, dsrc, 1, 0, w, h, dstride, M + 6 * (size_t)f, c, blk, blk + per, nullptr, nullptr);
}

OICS_API int32_t oics_projection_counts(oics_ctx *ctx, const uint8_t *bin, int32_t w, int32_t h, int32_t stride,
                                        const double M[6], int32_t *row_counts, int32_t *col_counts)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bin, w, h, stride, 1));
    OICS_ARG(M && row_counts && col_counts);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, bin, (size_t)w, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    std::vector<int32_t> counts((size_t)h + w);
    double v, hh;
    OICS_TRY(sweep_to(ctx, dsrc, 1, 0, w, h, dstride, M, 1, &v, &hh, nullptr, counts.data()));
    if (is_device_ptr(row_counts)) OICS_CUDA_TRY(cudaMemcpy(row_counts, counts.data(), sizeof(int32_t) * h, cudaMemcpyHostToDevice));
    else memcpy(row_counts, counts.data(), sizeof(int32_t) * h);
    if (is_device_ptr(col_counts)) OICS_CUDA_TRY(cudaMemcpy(col_counts, counts.data() + h, sizeof(int32_t) * w, cudaMemcpyHostToDevice));
    else memcpy(col_counts, counts.data() + h, sizeof(int32_t) * w);
    return OICS_OK;
}

OICS_API int32_t oics_projection_standard_deviations(oics_ctx *ctx, const uint8_t *bin, int32_t w, int32_t h,
                                                     int32_t stride, double *v_sd, double *h_sd)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bin, w, h, stride, 1));
    OICS_ARG(v_sd && h_sd);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, bin, (size_t)w, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    OICS_TRY(ctx->sd.ensure(2 * sizeof(double)));
    double *d = ctx->sd.as<double>();
    OICS_TRY(plain_std_devs_device(ctx, dsrc, w, h, dstride, d, d + 1));
    double out[2];
    OICS_CUDA_TRY(cudaMemcpyAsync(out, d, sizeof(out), cudaMemcpyDeviceToHost, ctx->stream));
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    *v_sd = out[0];
    *h_sd = out[1];
    return OICS_OK;
}

OICS_API int32_t oics_result_from_projection(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h,
                                             int32_t stride, uint16_t max_angle, double step, int32_t max_w,
                                             int32_t max_h, oics_omr_result *result)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bgr, w, h, stride, 3));
    OICS_ARG(result);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, bgr, (size_t)w * 3, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    const double scale = oics_projection_scale(w, h, max_w, max_h);
    std::vector<double> v, hh;
    int N = 0;
    // the rotation re-applies the resize scale (omr.rs:159-163)
    OICS_TRY(estimate_projection(ctx, dsrc, w, h, dstride, OICS_PRE_OMR, scale, max_angle, step, scale, &v, &hh, &N));
    return oics_rule_a(hh.data(), v.data(), N, step, result);
}

OICS_API int32_t oics_angle_with_projections(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h,
                                             int32_t stride, uint16_t max_angle, double step, double resize_scale,
                                             size_t threads, double *angle)
{
    (void)threads;  // single-thread semantics, see header
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bgr, w, h, stride, 3));
    OICS_ARG(angle);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, bgr, (size_t)w * 3, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    std::vector<double> v, hh;
    int N = 0;
    OICS_TRY(estimate_projection(ctx, dsrc, w, h, dstride, OICS_PRE_PROJECTION, resize_scale, max_angle, step, 1.0, &v,
                                 &hh, &N));
    if (N == 0) { set_error("empty angle range (the reference indexes vertical_vec[0] and panics)"); return OICS_ERR_INVALID_ARG; }
    return oics_rule_b(v.data(), hh.data(), N, step, angle);
}

OICS_API int32_t oics_warp_affine(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w, int32_t h,
                                  int32_t stride, const double M[6], int32_t interp, const uint8_t border[4],
                                  uint8_t *dst, int32_t dst_w, int32_t dst_h, int32_t dst_stride)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_ARG(ch >= 1 && ch <= 4);
    OICS_TRY(check_image(src, w, h, stride, ch));
    OICS_ARG(M && border && dst && dst_w > 0 && dst_h > 0 && dst_stride >= dst_w * ch);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, src, (size_t)w * ch, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    if (is_device_ptr(dst)) {
        OICS_TRY(warp_device(ctx, dsrc, ch, w, h, dstride, M, interp, border, dst, dst_w, dst_h, (size_t)dst_stride));
    } else {
        const size_t row = (size_t)dst_w * ch;
        OICS_TRY(ctx->out_img.ensure(row * dst_h));
        OICS_TRY(warp_device(ctx, dsrc, ch, w, h, dstride, M, interp, border, ctx->out_img.as<uint8_t>(), dst_w, dst_h, row));
        OICS_TRY(stage_out(ctx, ctx->out_img.p, row, dst_h, dst, (size_t)dst_stride));
    }
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

OICS_API int32_t oics_rotate(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w, int32_t h, int32_t stride,
                             double angle, double scale, int32_t interp, const uint8_t border[4], int32_t clip,
                             uint8_t *dst, size_t dst_cap, int32_t *dst_w, int32_t *dst_h)
{
    OICS_ARG(dst_w && dst_h);
    double M[6];
    OICS_TRY(oics_rotate_geometry(w, h, angle, scale, clip, M, dst_w, dst_h));
    if (*dst_w <= 0 || *dst_h <= 0) { set_error("rotated size %dx%d is empty", *dst_w, *dst_h); return OICS_ERR_INVALID_ARG; }
    const size_t need = (size_t)*dst_w * *dst_h * ch;
    if (need > dst_cap) {
        set_error("rotated image needs %zu bytes, capacity %zu", need, dst_cap);
        return OICS_ERR_CAPACITY;
    }
    return oics_warp_affine(ctx, src, ch, w, h, stride, M, interp, border, dst, *dst_w, *dst_h, *dst_w * ch);
}

// T1: TransformableMatrix::scale_self (transfer.rs:66-91)
OICS_API int32_t oics_scale(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w, int32_t h, int32_t stride,
                            double scale, uint8_t *dst, size_t dst_cap, int32_t *dst_w, int32_t *dst_h)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_ARG(ch == 1 || ch == 3);
    OICS_TRY(check_image(src, w, h, stride, ch));
    OICS_ARG(dst && dst_w && dst_h);
    if (scale > 1.0) {
        set_error("scale_self: enlarging (scale %g, INTER_LINEAR) is outside the accelerated path", scale);
        return OICS_ERR_UNSUPPORTED;
    }
    const int ow = scale == 1.0 ? w : scaled_dim(w, scale), oh = scale == 1.0 ? h : scaled_dim(h, scale);
    *dst_w = ow;
    *dst_h = oh;
    if (ow <= 0 || oh <= 0) { set_error("scaled size %dx%d is empty", ow, oh); return OICS_ERR_INVALID_ARG; }
    const size_t row = (size_t)ow * ch, need = row * oh;
    if (need > dst_cap) {
        set_error("scaled image needs %zu bytes, capacity %zu", need, dst_cap);
        return OICS_ERR_CAPACITY;
    }
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, src, (size_t)w * ch, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    if (scale == 1.0) {   // the early-out: the Mat is left as it is
        OICS_TRY(stage_out(ctx, dsrc, row, oh, dst, row));
    } else if (is_device_ptr(dst)) {
        OICS_TRY(resize_area_device(ctx, dsrc, ch, w, h, dstride, dst, ow, oh));
    } else {
        OICS_TRY(ctx->out_img.ensure(need));
        OICS_TRY(resize_area_device(ctx, dsrc, ch, w, h, dstride, ctx->out_img.as<uint8_t>(), ow, oh));
        OICS_TRY(copy_out(ctx, dst, ctx->out_img.p, need));
    }
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

// T1: TransformableMatrix::shrink_to (transfer.rs:93-126)
OICS_API int32_t oics_shrink_to(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w, int32_t h, int32_t stride,
                                int32_t max_w, int32_t max_h, uint8_t *dst, size_t dst_cap, int32_t *dst_w,
                                int32_t *dst_h, double *out_scale)
{
    if (w <= 0 || h <= 0) { set_error("image size %dx%d is empty", w, h); return OICS_ERR_INVALID_ARG; }
    const double target = oics_projection_scale(w, h, max_w, max_h);   // the same min(max_w / w, max_h / h) rule
    const double scale = target >= 1.0 ? 1.0 : target;                 // never upscales (transfer.rs:121-125)
    if (out_scale) *out_scale = scale;
    return oics_scale(ctx, src, ch, w, h, stride, scale, dst, dst_cap, dst_w, dst_h);
}

OICS_API int32_t oics_canny(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w, int32_t h, int32_t stride,
                            double lo, double hi, uint8_t *edges, int32_t edges_stride)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_ARG(ch == 1 || ch == 3);
    OICS_TRY(check_image(src, w, h, stride, ch));
    OICS_ARG(edges && edges_stride >= w);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, src, (size_t)w * ch, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    if (lo > hi) std::swap(lo, hi);
    OICS_TRY(ctx->edges.ensure((size_t)w * h));
    OICS_TRY(canny_device(ctx, dsrc, ch, w, h, dstride, (int)std::floor(lo), (int)std::floor(hi), ctx->edges.as<uint8_t>()));
    OICS_TRY(stage_out(ctx, ctx->edges.p, (size_t)w, h, edges, (size_t)edges_stride));
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

OICS_API int32_t oics_hough_accumulate(oics_ctx *ctx, const uint8_t *edges, int32_t w, int32_t h, int32_t stride,
                                       double rho, double theta, int32_t *accum)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(edges, w, h, stride, 1));
    OICS_ARG(accum && rho > 0 && theta > 0);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, edges, (size_t)w, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    int na, nr;
    hough_geometry(w, h, rho, theta, &na, &nr, nullptr);
    const size_t bytes = (size_t)na * nr * sizeof(int32_t);
    int32_t *dacc = is_device_ptr(accum) ? accum : nullptr;
    if (!dacc) {
        OICS_TRY(ctx->accum.ensure(bytes));
        dacc = ctx->accum.as<int32_t>();
    }
    OICS_TRY(hough_accumulate_device(ctx, dsrc, w, h, dstride, rho, theta, dacc));
    if (dacc != accum) OICS_TRY(copy_out(ctx, accum, dacc, bytes));
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

OICS_API int32_t oics_hough_peaks(oics_ctx *ctx, const int32_t *accum, int32_t numangle, int32_t numrho,
                                  int32_t threshold, int32_t *peaks, int32_t cap, int32_t *n_peaks)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_ARG(accum && numangle > 0 && numrho > 0 && n_peaks && cap >= 0 && (cap == 0 || peaks));
    const size_t bytes = (size_t)numangle * numrho * sizeof(int32_t);
    const int32_t *dacc = accum;
    if (!is_device_ptr(accum)) {
        OICS_TRY(ctx->accum.ensure(bytes));
        OICS_CUDA_TRY(cudaMemcpyAsync(ctx->accum.p, accum, bytes, cudaMemcpyHostToDevice, ctx->stream));
        dacc = ctx->accum.as<int32_t>();
    }
    std::vector<int32_t> found;
    OICS_TRY(hough_peaks_device(ctx, dacc, numangle, numrho, threshold, &found));
    const int n = (int)(found.size() / 3);
    *n_peaks = n;
    const int ncopy = std::min(n, cap);
    if (ncopy > 0) {
        if (is_device_ptr(peaks)) OICS_CUDA_TRY(cudaMemcpy(peaks, found.data(), sizeof(int32_t) * 3 * (size_t)ncopy, cudaMemcpyHostToDevice));
        else memcpy(peaks, found.data(), sizeof(int32_t) * 3 * (size_t)ncopy);
    }
    if (n > cap) {
        set_error("found %d peaks, capacity %d", n, cap);
        return OICS_ERR_CAPACITY;
    }
    return OICS_OK;
}

OICS_API int32_t oics_hough_lines_p(oics_ctx *ctx, const uint8_t *edges, int32_t w, int32_t h, int32_t stride,
                                    double rho, double theta, int32_t threshold, double min_line_length,
                                    double max_line_gap, int32_t *lines, int32_t cap, int32_t *n_lines)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(edges, w, h, stride, 1));
    OICS_ARG(n_lines && cap >= 0 && (cap == 0 || lines) && rho > 0 && theta > 0);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, edges, (size_t)w, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    int n = 0;
    OICS_TRY(hough_lines_p_device(ctx, dsrc, w, h, dstride, rho, theta, threshold, (int)std::lrint(min_line_length),
                                  (int)std::lrint(max_line_gap), lines, cap, &n));
    *n_lines = n;
    if (n > cap) {
        set_error("found %d segments, capacity %d", n, cap);
        return OICS_ERR_CAPACITY;
    }
    return OICS_OK;
}

OICS_API int32_t oics_result_from_edges_detection(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w,
                                                  int32_t h, int32_t stride, double min_line_length,
                                                  double max_line_gap, oics_omr_result *result)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_ARG(ch == 1 || ch == 3);
    OICS_TRY(check_image(src, w, h, stride, ch));
    OICS_ARG(result);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, src, (size_t)w * ch, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    return edges_result(ctx, dsrc, ch, w, h, dstride, 50.0, 150.0, 0, min_line_length, max_line_gap, OICS_LINES_OMR, result);
}

OICS_API int32_t oics_result_from_edges_detection_batch(oics_ctx *ctx, const uint8_t *srcs, int32_t n_cards,
                                                        size_t card_stride, int32_t ch, int32_t w, int32_t h,
                                                        int32_t stride, double min_line_length, double max_line_gap,
                                                        double *angles, int32_t *result_statuses, int32_t *statuses)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_ARG(ch == 1 || ch == 3);
    OICS_ARG(n_cards >= 0 && angles && result_statuses && statuses);
    if (n_cards == 0) return OICS_OK;
    OICS_TRY(check_image(srcs, w, h, stride, ch));
    OICS_ARG(card_stride >= (size_t)stride * (size_t)h);
    const size_t row = (size_t)w * ch, card_bytes = row * h;
    const bool dev_in = is_device_ptr(srcs);
    // cards in flight: bounded by the per-card Hough scratch (7 B per pixel + the accumulator)
    const size_t per_card = (size_t)w * h * 7 + (size_t)180 * ((size_t)(w + h) * 2 + 1) * 4 + card_bytes;
    // and by the device: 37 clusters of 4 CTAs are co-resident on 148 SMs
    const int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)std::max(1, ctx->sm_count / 4), ((size_t)4 << 30) / per_card));
    for (int c0 = 0; c0 < n_cards; c0 += chunk) {
        const int nb = std::min(chunk, n_cards - c0);
        std::vector<const uint8_t *> ptrs((size_t)nb);
        size_t dstride = (size_t)stride;
        if (dev_in) {
            for (int i = 0; i < nb; i++) ptrs[(size_t)i] = srcs + (size_t)(c0 + i) * card_stride;
        } else {
            OICS_TRY(ctx->in_img.ensure((size_t)nb * card_bytes));
            for (int i = 0; i < nb; i++) {
                OICS_CUDA_TRY(cudaMemcpy2DAsync(ctx->in_img.as<uint8_t>() + (size_t)i * card_bytes, row,
                                                srcs + (size_t)(c0 + i) * card_stride, (size_t)stride, row, (size_t)h,
                                                cudaMemcpyHostToDevice, ctx->stream));
                ptrs[(size_t)i] = ctx->in_img.as<uint8_t>() + (size_t)i * card_bytes;
            }
            dstride = row;
        }
        std::vector<oics_omr_result> res;
        std::vector<int32_t> sts;
        OICS_TRY(edges_results_batch(ctx, ptrs.data(), nb, ch, w, h, dstride, 50.0, 150.0, 0, min_line_length, max_line_gap,
                                     OICS_LINES_OMR, &res, &sts));
        for (int i = 0; i < nb; i++) {
            angles[c0 + i] = res[(size_t)i].angle;
            result_statuses[c0 + i] = res[(size_t)i].status;
            statuses[c0 + i] = sts[(size_t)i];
        }
    }
    return OICS_OK;
}

OICS_API int32_t oics_angle_with_hough(oics_ctx *ctx, const uint8_t *gray, int32_t w, int32_t h, int32_t stride,
                                       double min_line_length, double max_line_gap, double *angle)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(gray, w, h, stride, 1));
    OICS_ARG(angle);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, gray, (size_t)w, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    oics_omr_result r = {0, 0, 0, nullptr, 0};
    OICS_TRY(edges_result(ctx, dsrc, 1, w, h, dstride, 50.0, 150.0, 0, min_line_length, max_line_gap, OICS_LINES_HOUGH, &r));
    *angle = r.angle;
    return OICS_OK;
}

OICS_API int32_t oics_fft_image(oics_ctx *ctx, const uint8_t *gray, int32_t w, int32_t h, int32_t stride,
                                uint8_t *mag8, uint8_t *log8)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(gray, w, h, stride, 1));
    OICS_ARG(log8);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, gray, (size_t)w, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    OICS_TRY(ctx->tmp8a.ensure((size_t)w * h));
    OICS_TRY(ctx->tmp8b.ensure((size_t)w * h));
    OICS_TRY(fft_image_device(ctx, dsrc, w, h, dstride, mag8 ? ctx->tmp8a.as<uint8_t>() : nullptr, ctx->tmp8b.as<uint8_t>()));
    if (mag8) OICS_TRY(copy_out(ctx, mag8, ctx->tmp8a.p, (size_t)w * h));
    OICS_TRY(copy_out(ctx, log8, ctx->tmp8b.p, (size_t)w * h));
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

OICS_API int32_t oics_angle_with_fft(oics_ctx *ctx, const uint8_t *gray, int32_t w, int32_t h, int32_t stride,
                                     double canny1, double canny2, double min_line_length, double max_line_gap,
                                     double *angle)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(gray, w, h, stride, 1));
    OICS_ARG(angle);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, gray, (size_t)w, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    OICS_TRY(ctx->tmp8b.ensure((size_t)w * h));
    OICS_TRY(fft_image_device(ctx, dsrc, w, h, dstride, nullptr, ctx->tmp8b.as<uint8_t>()));
    oics_omr_result r = {0, 0, 0, nullptr, 0};
    OICS_TRY(edges_result(ctx, ctx->tmp8b.as<uint8_t>(), 1, w, h, (size_t)w, canny1, canny2, 100, min_line_length,
                          max_line_gap, OICS_LINES_FFT, &r));
    *angle = r.angle;
    return OICS_OK;
}

OICS_API int32_t oics_result_from_fourier_transform(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h,
                                                    int32_t stride, double canny_weak, double canny_strong,
                                                    double min_line_length, double max_line_gap,
                                                    oics_omr_result *result)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bgr, w, h, stride, 3));
    OICS_ARG(result);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, bgr, (size_t)w * 3, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    OICS_TRY(ctx->gray.ensure((size_t)w * h));
    OICS_TRY(gray_device(ctx, dsrc, w, h, dstride, ctx->gray.as<uint8_t>()));
    OICS_TRY(ctx->tmp8b.ensure((size_t)w * h));
    OICS_TRY(fft_image_device(ctx, ctx->gray.as<uint8_t>(), w, h, (size_t)w, nullptr, ctx->tmp8b.as<uint8_t>()));
    // omr.rs:322-336: Canny on the log image, then get_result_from_edges_detection
    // runs Canny(50,150) AGAIN on that edge map, threshold 0
    OICS_TRY(ctx->tmp8a.ensure((size_t)w * h));
    if (canny_weak > canny_strong) std::swap(canny_weak, canny_strong);
    OICS_TRY(canny_device(ctx, ctx->tmp8b.as<uint8_t>(), 1, w, h, (size_t)w, (int)std::floor(canny_weak),
                          (int)std::floor(canny_strong), ctx->tmp8a.as<uint8_t>()));
    return edges_result(ctx, ctx->tmp8a.as<uint8_t>(), 1, w, h, (size_t)w, 50.0, 150.0, 0, min_line_length, max_line_gap,
                        OICS_LINES_OMR, result);
}

OICS_API int32_t oics_jpeg_encode(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h, int32_t stride,
                                  int32_t quality, uint8_t *out, size_t out_cap, size_t *out_size)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bgr, w, h, stride, 3));
    OICS_ARG(out && out_size);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, bgr, (size_t)w * 3, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    if (is_device_ptr(out)) return jpeg_encode_device(ctx, dsrc, w, h, dstride, quality, out, out_cap, out_size);
    // a file is never larger than a few bytes per pixel; the device copy is sized like the caller's buffer
    const size_t dcap = std::min(out_cap, (size_t)4096 + (size_t)w * h * 6);
    OICS_TRY(ctx->out_img2.ensure(dcap));
    int32_t rc = jpeg_encode_device(ctx, dsrc, w, h, dstride, quality, ctx->out_img2.as<uint8_t>(), dcap, out_size);
    if (rc != OICS_OK) return rc;
    OICS_TRY(copy_out(ctx, out, ctx->out_img2.p, *out_size));
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

OICS_API int32_t oics_jpeg_info(const uint8_t *file, size_t file_size, int32_t *w, int32_t *h, int32_t *channels)
{
    if (!file || !w || !h) {
        set_error("invalid argument: file, w, h");
        return OICS_ERR_INVALID_ARG;
    }
    int iw = 0, ih = 0, nc = 0;
    OICS_TRY(jpeg_info_host(file, file_size, &iw, &ih, &nc));
    *w = iw;
    *h = ih;
    if (channels) *channels = nc;
    return OICS_OK;
}

OICS_API int32_t oics_jpeg_decode(oics_ctx *ctx, const uint8_t *file, size_t file_size, uint8_t *bgr, size_t stride,
                                  size_t bgr_cap, int32_t *w, int32_t *h)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_ARG(file && bgr && w && h);
    OICS_ARG(!is_device_ptr(file));
    int iw = 0, ih = 0, nc = 0;
    OICS_TRY(jpeg_info_host(file, file_size, &iw, &ih, &nc));
    *w = iw;
    *h = ih;
    if (stride == 0) stride = (size_t)iw * 3;
    OICS_ARG(stride >= (size_t)iw * 3);
    if (stride * (size_t)(ih - 1) + (size_t)iw * 3 > bgr_cap) {
        set_error("jpeg: the image needs %zu bytes, capacity %zu", stride * (size_t)(ih - 1) + (size_t)iw * 3, bgr_cap);
        return OICS_ERR_CAPACITY;
    }
    if (is_device_ptr(bgr)) {
        OICS_TRY(jpeg_decode_device(ctx, file, file_size, bgr, stride, &iw, &ih));
        OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        return OICS_OK;
    }
    OICS_TRY(ctx->out_img.ensure((size_t)iw * 3 * ih));
    OICS_TRY(jpeg_decode_device(ctx, file, file_size, ctx->out_img.as<uint8_t>(), (size_t)iw * 3, &iw, &ih));
    OICS_CUDA_TRY(cudaMemcpy2DAsync(bgr, stride, ctx->out_img.p, (size_t)iw * 3, (size_t)iw * 3, (size_t)ih, cudaMemcpyDeviceToHost,
                                    ctx->stream));
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

// correct_default for a BGR card already on the device
static int32_t correct_one(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, uint16_t max_angle,
                           double step, int max_w, int max_h, double min_len, double max_gap, uint8_t *dst,
                           size_t dst_cap, int32_t *dst_w, int32_t *dst_h, double *angle, int32_t *need_check)
{
    const double scale = oics_projection_scale(w, h, max_w, max_h);
    std::vector<double> v, hh;
    int N = 0;
    OICS_TRY(estimate_projection(ctx, dbgr, w, h, stride, OICS_PRE_OMR, scale, max_angle, step, scale, &v, &hh, &N));
    std::vector<double> cand((size_t)std::max(1, 2 * N));
    oics_omr_result p = {0, 0, 0, cand.data(), (int32_t)cand.size()};
    OICS_TRY(oics_rule_a(hh.data(), v.data(), N, step, &p));
    double a;
    int32_t need;
    if (p.status == OICS_BELIEVED) {
        a = p.angle;
        need = 0;
    } else {
        // omr.rs:366: Canny runs on the 3-channel source
        oics_omr_result e = {0, 0, 0, nullptr, 0};
        OICS_TRY(edges_result(ctx, dbgr, 3, w, h, stride, 50.0, 150.0, 0, min_len, max_gap, OICS_LINES_OMR, &e));
        OICS_TRY(oics_reconcile(&p, e.angle, &a, &need));
    }
    *angle = a;
    *need_check = need;
    double M[6];
    int dw, dh;
    OICS_TRY(oics_rotate_geometry(w, h, a, 1.0, OICS_CLIP_CONTAIN, M, &dw, &dh));
    if (dst_w) *dst_w = dw;
    if (dst_h) *dst_h = dh;
    if (!dst) return OICS_OK;
    const size_t row = (size_t)dw * 3, need_bytes = row * dh;
    if (need_bytes > dst_cap) {
        set_error("deskewed card needs %zu bytes, capacity %zu", need_bytes, dst_cap);
        return OICS_ERR_CAPACITY;
    }
    const uint8_t white[4] = {255, 255, 255, 0};
    if (is_device_ptr(dst)) {
        OICS_TRY(warp_device(ctx, dbgr, 3, w, h, stride, M, OICS_INTER_NEAREST, white, dst, dw, dh, row));
    } else {
        OICS_TRY(ctx->out_img.ensure(need_bytes));
        OICS_TRY(warp_device(ctx, dbgr, 3, w, h, stride, M, OICS_INTER_NEAREST, white, ctx->out_img.as<uint8_t>(), dw, dh, row));
        OICS_TRY(copy_out(ctx, dst, ctx->out_img.p, need_bytes));
    }
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return OICS_OK;
}

OICS_API int32_t oics_correct_default(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h, int32_t stride,
                                      uint16_t max_angle, double step, int32_t max_w, int32_t max_h,
                                      double hough_min_line_length, double hough_max_line_gap, uint8_t *dst,
                                      size_t dst_cap, int32_t *dst_w, int32_t *dst_h, double *angle,
                                      int32_t *need_check)
{
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    OICS_TRY(check_image(bgr, w, h, stride, 3));
    OICS_ARG(angle && need_check);
    const uint8_t *dsrc;
    size_t dstride;
    OICS_TRY(stage_in(ctx, bgr, (size_t)w * 3, h, (size_t)stride, ctx->in_img, &dsrc, &dstride));
    return correct_one(ctx, dsrc, w, h, dstride, max_angle, step, max_w, max_h, hough_min_line_length,
                       hough_max_line_gap, dst, dst_cap, dst_w, dst_h, angle, need_check);
}

// The batch driver that replaces the reference's thread pool (task.rs:19-68, thread_pool.rs):
// cards move through a three-stage pipeline of sub-batches -- upload (side stream), estimate +
// deskew (ctx stream: pre-processing and ONE fused sweep launch per sub-batch, exact rule A on
// the host, one warp per card), download (side stream) -- so PCIe in both directions overlaps
// the kernels.  Host buffers should be pinned for the copies to be asynchronous.
// jpeg_quality == 0: dsts receive the deskewed BGR pixels; 1..100: the JPEG file imwrite would produce
// (omr.rs:447-448 writes quality 100), its size in dst_sizes.
static int32_t correct_batch_impl(oics_ctx *ctx, const uint8_t *bgrs, int32_t n_cards, size_t card_stride, int32_t w,
                                  int32_t h, int32_t stride, uint16_t max_angle, double step, int32_t max_w,
                                  int32_t max_h, double hough_min_line_length, double hough_max_line_gap,
                                  int32_t jpeg_quality, uint8_t *const *dsts, const size_t *dst_caps, size_t *dst_sizes,
                                  int32_t *dst_ws, int32_t *dst_hs, double *angles, int32_t *need_checks,
                                  int32_t *statuses, const uint8_t *const *files = nullptr, const size_t *file_sizes = nullptr)
{
    // files != nullptr: the cards arrive as JPEG files in host memory (imread, omr.rs:349) and are decoded
    // on the device, each on its own stream, into the slot the upload would have filled
    CtxGuard g(ctx, __func__);
    if (g.status) return g.status;
    if (!files) OICS_TRY(check_image(bgrs, w, h, stride, 3));
    OICS_ARG(!files || (file_sizes && w > 0 && h > 0));
    OICS_ARG(n_cards >= 0 && angles && need_checks && statuses);
    OICS_ARG(jpeg_quality >= 0 && jpeg_quality <= 100 && (!jpeg_quality || !dsts || dst_sizes));
    OICS_ARG(card_stride >= (size_t)stride * (size_t)h);
    OICS_ARG(!dsts || dst_caps);
    if (n_cards == 0) return OICS_OK;
    const bool dev_in = !files && is_device_ptr(bgrs);
    const size_t row = (size_t)w * 3, card_bytes = row * h;
    // cards per sub-batch: 16 A4 cards (418 MB) amortise the per-sub-batch host round trip (rule A needs the
    // std-devs on the host) and share one read of the sweep's plane set; small cards get proportionally more,
    // up to 64
    // files: the cards of a sub-batch are decoded concurrently, one stream each -- the Huffman synchronisation is
    // latency-bound per card, so cards in flight are what fills the device (OICS_FILES_SUBBATCH: tuning)
    const int slot_set = (OICS_JD_SLOTS - 1) / 2;     // sub-batches alternate between two sets of decode slots
    int files_b = (int)std::max<size_t>(8, std::min<size_t>(slot_set, ((size_t)420 << 20) / std::max<size_t>(1, card_bytes)));
    if (const char *fb = getenv("OICS_FILES_SUBBATCH")) files_b = std::max(1, std::min(slot_set, atoi(fb)));
    const int B_max = files ? files_b : (int)std::max<size_t>(8, std::min<size_t>(64, ((size_t)420 << 20) / std::max<size_t>(1, card_bytes)));
    // with host buffers on either side the copies of one sub-batch overlap the kernels of another: keep at least
    // eight stages in the pipeline; device-resident batches have nothing to overlap and take the largest sub-batch
    bool any_host_dst = false;
    if (dsts)
        for (int c = 0; c < n_cards && !any_host_dst; c++) any_host_dst = dsts[c] && !is_device_ptr(dsts[c]);
    const bool pipelined = files || !is_device_ptr(bgrs) || any_host_dst;
    const int B_target = files ? B_max : pipelined ? std::max(8, std::min(B_max, cdiv(n_cards, 8))) : B_max;
    const int B = std::min<int>(n_cards, B_target);
    std::vector<int32_t> file_status((size_t)n_cards, OICS_OK);
    const int n_sub = cdiv(n_cards, B);
    cudaStream_t st = ctx->stream, up = nullptr, down = nullptr;
    cudaEvent_t in_ready[2] = {}, in_free[2] = {}, out_ready[2] = {}, out_free[2] = {}, lane_go = nullptr;
    // encode lanes: the cards of a sub-batch are deskewed + encoded on this many streams side by side (0 = on ctx->stream)
    static const int lanes_env = [] { const char *e = getenv("OICS_JPEG_LANES_N"); return e ? std::max(0, std::min(OICS_JPEG_LANES, atoi(e))) : OICS_JPEG_LANES; }();
    const int n_lanes = jpeg_quality && dsts ? lanes_env : 0;
    DevBuf *in_slot[2] = {&ctx->in_img, &ctx->in_img2}, *out_slot[2] = {&ctx->out_img, &ctx->out_img2};

    // geometry shared by the batch
    const double scale = oics_projection_scale(w, h, max_w, max_h);
    const int ow = scale == 1.0 ? w : scaled_dim(w, scale), oh = scale == 1.0 ? h : scaled_dim(h, scale);
    if (ow <= 0 || oh <= 0) { set_error("scaled size %dx%d is empty", ow, oh); return OICS_ERR_INVALID_ARG; }
    const int N = oics_angle_half_range(max_angle, step);
    std::vector<double> M((size_t)std::max(1, 2 * N) * 6);
    int n_angles = 0;
    if (N > 0) OICS_TRY(oics_sweep_matrices(ow, oh, max_angle, step, scale, M.data(), 2 * N, &n_angles));
    OICS_TRY(ctx->bin.ensure((size_t)B * ow * oh));
    // CONTAIN output bound: any rotation of a w x h card fits a square of side (w + h) / sqrt(2) + 2
    const size_t side = (size_t)std::ceil((double)(w + h) * 0.70710678118654757) + 2;
    const size_t out_card_cap = (side * side * 3 + 255) & ~(size_t)255;
    bool host_out = false;
    if (dsts)
        for (int c = 0; c < n_cards; c++) host_out = host_out || (dsts[c] && !is_device_ptr(dsts[c]));
    // Cards whose projection estimate is not Believed wait in a device-side queue: their Canny + HoughLinesP
    // fallbacks (omr.rs:366-403) run TOGETHER, one thread-block cluster per card in one launch, when the queue
    // is full or the batch ends -- the reference's concurrency (one OS thread per card, thread_pool.rs:41-73)
    // instead of one 20-100 ms sequential transform stalling every sub-batch.
    const size_t hough_per_card = (size_t)w * h * 7 + (size_t)180 * ((size_t)(w + h) * 2 + 1) * 4;
    const int FQ = (int)std::max<size_t>(1, std::min<size_t>((size_t)std::max(1, ctx->sm_count / 4),
                                                            ((size_t)4 << 30) / (hough_per_card + card_bytes + out_card_cap)));
    if (jpeg_quality && dsts) OICS_TRY(ctx->jpeg_totals.ensure((size_t)std::max(B, FQ) * 2 * sizeof(unsigned long long)));
    if (jpeg_quality && dsts) OICS_TRY(ctx->jpeg_raw.ensure(out_card_cap));   // the deskewed pixels of the card being encoded
    struct Deferred {
        int c;
        oics_omr_result p;
        std::vector<double> cand;
    };
    std::vector<Deferred> queue;

    int32_t rc = OICS_OK;
    auto cleanup = [&]() {
        cudaStreamSynchronize(st);
        if (up) { cudaStreamSynchronize(up); cudaStreamDestroy(up); }
        if (down) { cudaStreamSynchronize(down); cudaStreamDestroy(down); }
        for (int li = 0; li < n_lanes; li++)
            if (ctx->jpeg_lanes[li].st) cudaStreamSynchronize(ctx->jpeg_lanes[li].st);
        for (int i = 0; i < 2; i++)
            for (cudaEvent_t e : {in_ready[i], in_free[i], out_ready[i], out_free[i]})
                if (e) cudaEventDestroy(e);
        if (lane_go) cudaEventDestroy(lane_go);
    };
#define OICS_BATCH_TRY(expr)                                   \
    do {                                                       \
        int32_t _s = (expr);                                   \
        if (_s != OICS_OK) { cleanup(); return _s; }           \
    } while (0)
#define OICS_BATCH_CUDA(expr)                                                                     \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess) {                                                                  \
            set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            cleanup();                                                                            \
            return OICS_ERR_CUDA;                                                                 \
        }                                                                                         \
    } while (0)

    OICS_BATCH_CUDA(cudaEventCreateWithFlags(&lane_go, cudaEventDisableTiming));
    for (int i = 0; i < 2; i++) {
        for (cudaEvent_t *e : {&in_ready[i], &in_free[i], &out_ready[i], &out_free[i]})
            OICS_BATCH_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
        if (!dev_in && i < n_sub) OICS_BATCH_TRY(in_slot[i]->ensure((size_t)B * card_bytes));
        if (host_out && i < n_sub) OICS_BATCH_TRY(out_slot[i]->ensure((size_t)B * out_card_cap));
    }
    if (!dev_in) OICS_BATCH_CUDA(cudaStreamCreateWithFlags(&up, cudaStreamNonBlocking));
    if (host_out) OICS_BATCH_CUDA(cudaStreamCreateWithFlags(&down, cudaStreamNonBlocking));

    auto upload = [&](int k) -> cudaError_t {   // sub-batch k -> in_slot[k & 1]
        const int s = k & 1, c0 = k * B, nb = std::min(B, n_cards - c0);
        cudaError_t e = cudaSuccess;
        if (k >= 2) e = cudaStreamWaitEvent(up, in_free[s], 0);
        const uint8_t *src = bgrs + (size_t)c0 * card_stride;
        uint8_t *dst = in_slot[s]->as<uint8_t>();
        if (e == cudaSuccess) {
            if ((size_t)stride == row && card_stride == card_bytes)
                e = cudaMemcpyAsync(dst, src, (size_t)nb * card_bytes, cudaMemcpyHostToDevice, up);
            else
                for (int i = 0; i < nb && e == cudaSuccess; i++)
                    e = cudaMemcpy2DAsync(dst + (size_t)i * card_bytes, row, src + (size_t)i * card_stride, (size_t)stride,
                                          row, (size_t)h, cudaMemcpyHostToDevice, up);
        }
        if (e == cudaSuccess) e = cudaEventRecord(in_ready[s], up);
        return e;
    };
    // files of sub-batch k -> in_slot[k & 1], in two steps (jpegdec.cu): `begin` only enqueues (upload, unstuffing,
    // first synchronisation rounds, each card on its own stream) and touches nothing but decoder scratch, so it is
    // issued a whole sweep ahead; `finish` writes the cards once the slot is free
    auto decode_begin = [&](int k) -> int32_t {
        const int c0 = k * B, nb = std::min(B, n_cards - c0);
        return jpeg_decode_batch_begin(ctx, 1 + (k & 1) * slot_set, files + c0, file_sizes + c0, nb, w, h, file_status.data() + c0);
    };
    auto decode_finish = [&](int k) -> int32_t {
        const int s = k & 1, c0 = k * B, nb = std::min(B, n_cards - c0);
        return jpeg_decode_batch_finish(ctx, 1 + (k & 1) * slot_set, nb, w, in_slot[s]->as<uint8_t>(), card_bytes, k >= 2 ? in_free[s] : nullptr, up,
                                        in_ready[s], file_status.data() + c0);
    };
    if (files) {
        OICS_BATCH_TRY(decode_begin(0));
        if (n_sub > 1) OICS_BATCH_TRY(decode_begin(1));
        OICS_BATCH_TRY(decode_finish(0));
    } else if (!dev_in) {
        OICS_BATCH_CUDA(upload(0));
        if (n_sub > 1) OICS_BATCH_CUDA(upload(1));
    }

    std::vector<double> v((size_t)B * std::max(1, 2 * N)), hh((size_t)B * std::max(1, 2 * N)), cand((size_t)std::max(1, 2 * N));
    const uint8_t white[4] = {255, 255, 255, 0};
    // the queued fallbacks: Canny on every card, ONE PPHT launch (a cluster per card), reconcile (omr.rs:352-403),
    // then the final warp (+ encode) of each and its copy to the caller
    auto flush_fallbacks = [&]() -> int32_t {
        const int nq = (int)queue.size();
        if (nq == 0) return OICS_OK;
        std::vector<const uint8_t *> srcs((size_t)nq);
        for (int i = 0; i < nq; i++) srcs[(size_t)i] = ctx->fb_in.as<uint8_t>() + (size_t)i * card_bytes;
        std::vector<oics_omr_result> er;
        std::vector<int32_t> es;
        OICS_TRY(edges_results_batch(ctx, srcs.data(), nq, 3, w, h, row, 50.0, 150.0, 0, hough_min_line_length, hough_max_line_gap,
                                     OICS_LINES_OMR, &er, &es));
        bool any_host = false;
        if (dsts)
            for (int i = 0; i < nq; i++) any_host = any_host || (dsts[queue[(size_t)i].c] && !is_device_ptr(dsts[queue[(size_t)i].c]));
        if (any_host) OICS_TRY(ctx->fb_out.ensure((size_t)FQ * out_card_cap));
        std::vector<size_t> obytes((size_t)nq, 0), jcap((size_t)nq, 0);
        size_t jhdr = 0;
        for (int i = 0; i < nq; i++) {
            Deferred &d = queue[(size_t)i];
            const int c = d.c;
            int32_t stc = es[(size_t)i];
            double a = 0.0;
            int32_t need = 0;
            if (stc == OICS_OK) {
                d.p.candidates = d.cand.empty() ? nullptr : d.cand.data();
                d.p.cap = (int32_t)d.cand.size();
                stc = oics_reconcile(&d.p, er[(size_t)i].angle, &a, &need);
            }
            double Mw[6];
            int dw = 0, dh = 0;
            if (stc == OICS_OK) stc = oics_rotate_geometry(w, h, a, 1.0, OICS_CLIP_CONTAIN, Mw, &dw, &dh);
            angles[c] = a;
            need_checks[c] = need;
            if (dst_ws) dst_ws[c] = dw;
            if (dst_hs) dst_hs[c] = dh;
            uint8_t *dst = dsts ? dsts[c] : nullptr;
            const size_t need_bytes = (size_t)dw * 3 * dh;
            if (stc == OICS_OK && dst) {
                const bool ddev = is_device_ptr(dst);
                uint8_t *area = ddev ? dst : ctx->fb_out.as<uint8_t>() + (size_t)i * out_card_cap;
                const size_t area_cap = ddev ? dst_caps[c] : std::min(out_card_cap, dst_caps[c]);
                if (jpeg_quality) {
                    if (need_bytes > out_card_cap) {
                        set_error("internal: CONTAIN bound %zu < %zu", out_card_cap, need_bytes);
                        stc = OICS_ERR_CAPACITY;
                    } else {
                        stc = warp_device(ctx, srcs[(size_t)i], 3, w, h, row, Mw, OICS_INTER_NEAREST, white, ctx->jpeg_raw.as<uint8_t>(), dw, dh,
                                          (size_t)dw * 3);
                        if (stc == OICS_OK)
                            stc = jpeg_encode_enqueue(ctx, ctx->jpeg_raw.as<uint8_t>(), dw, dh, (size_t)dw * 3, jpeg_quality, area, area_cap,
                                                      ctx->jpeg_totals.as<unsigned long long>() + 2 * i, &jhdr);
                        if (stc == OICS_OK) jcap[(size_t)i] = area_cap;
                    }
                } else {
                    if (dst_sizes) dst_sizes[c] = need_bytes;
                    if (need_bytes > dst_caps[c] || need_bytes > (ddev ? need_bytes : out_card_cap)) {
                        set_error("deskewed card needs %zu bytes, capacity %zu", need_bytes, dst_caps[c]);
                        stc = OICS_ERR_CAPACITY;
                    } else {
                        stc = warp_device(ctx, srcs[(size_t)i], 3, w, h, row, Mw, OICS_INTER_NEAREST, white, area, dw, dh, (size_t)dw * 3);
                        if (stc == OICS_OK && !ddev) obytes[(size_t)i] = need_bytes;
                    }
                }
            }
            statuses[c] = stc;
            if (stc == OICS_ERR_CUDA) return OICS_ERR_CUDA;
        }
        if (jpeg_quality && dsts) {
            std::vector<unsigned long long> totals((size_t)2 * nq, 0ull);
            OICS_CUDA_TRY(cudaMemcpyAsync(totals.data(), ctx->jpeg_totals.p, totals.size() * sizeof(unsigned long long),
                                          cudaMemcpyDeviceToHost, st));
            OICS_CUDA_TRY(cudaStreamSynchronize(st));
            for (int i = 0; i < nq; i++) {
                if (!jcap[(size_t)i]) continue;
                const int c = queue[(size_t)i].c;
                const size_t fsz = jpeg_file_size(jhdr, &totals[(size_t)2 * i]);
                dst_sizes[c] = fsz;
                if (fsz > jcap[(size_t)i]) {
                    set_error("jpeg: the file needs %zu bytes, capacity %zu", fsz, jcap[(size_t)i]);
                    statuses[c] = OICS_ERR_CAPACITY;
                } else if (!is_device_ptr(dsts[c])) {
                    obytes[(size_t)i] = fsz;
                }
            }
        }
        for (int i = 0; i < nq; i++)
            if (obytes[(size_t)i])
                OICS_CUDA_TRY(cudaMemcpyAsync(dsts[queue[(size_t)i].c], ctx->fb_out.as<uint8_t>() + (size_t)i * out_card_cap, obytes[(size_t)i],
                                              cudaMemcpyDeviceToHost, st));
        // fb_in / fb_out are reused by the next flush: everything above must have drained
        OICS_CUDA_TRY(cudaStreamSynchronize(st));
        queue.clear();
        return OICS_OK;
    };
    // OICS_BATCH_TRACE=1: host-side wall times of the phases of every sub-batch, on stderr (diagnostics)
    static const bool trace = getenv("OICS_BATCH_TRACE") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_start = trace ? now() : 0.0;
    for (int k = 0; k < n_sub && rc == OICS_OK; k++) {
        double tr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (trace) tr[0] = now();
        const int s = k & 1, c0 = k * B, nb = std::min(B, n_cards - c0);
        const uint8_t *dsrc = dev_in ? bgrs + (size_t)c0 * card_stride : in_slot[s]->as<uint8_t>();
        const size_t dcard = dev_in ? card_stride : card_bytes, dstride = dev_in ? (size_t)stride : row;
        if (!dev_in) OICS_BATCH_CUDA(cudaStreamWaitEvent(st, in_ready[s], 0));
        // P0 for every card of the sub-batch (one launch), then one fused sweep over all of them
        OICS_BATCH_TRY(preprocess_batch_device(ctx, dsrc, nb, dcard, w, h, dstride, OICS_PRE_OMR, ow, oh, ctx->bin.as<uint8_t>()));
        if (trace) tr[1] = now();
        if (N > 0)
            OICS_BATCH_TRY(sweep_to(ctx, ctx->bin.as<uint8_t>(), nb, (size_t)ow * oh, ow, oh, (size_t)ow, M.data(), n_angles,
                                    v.data(), hh.data(), nullptr, nullptr));
        if (trace) tr[2] = now();
        // exact rule A (+ Hough fallback and reconcile, omr.rs:352-403) per card on the host
        if (host_out && k >= 2) OICS_BATCH_CUDA(cudaStreamWaitEvent(st, out_free[s], 0));
        std::vector<size_t> out_off((size_t)nb, 0), out_bytes((size_t)nb, 0), jpeg_cap((size_t)nb, 0);
        size_t jpeg_hdr = 0;
        int lanes_used = 0;
        for (int i = 0; i < nb; i++) {
            const int c = c0 + i;
            const uint8_t *card = dsrc + (size_t)i * dcard;
            oics_omr_result p = {0, 0, 0, cand.data(), (int32_t)cand.size()};
            int32_t stc = file_status[c];       // a file that did not decode: its status, nothing else
            if (stc != OICS_OK) {
                statuses[c] = stc;
                angles[c] = 0.0;
                need_checks[c] = 0;
                if (dst_sizes) dst_sizes[c] = 0;
                continue;
            }
            stc = oics_rule_a(hh.data() + (size_t)i * 2 * N, v.data() + (size_t)i * 2 * N, N, step, &p);
            double a = 0.0;
            int32_t need = 0;
            if (stc == OICS_OK) {
                if (p.status == OICS_BELIEVED) {
                    a = p.angle;
                } else {
                    // omr.rs:366: the fallback needs the 3-channel source; keep a copy (the input slot is recycled)
                    const int slot = (int)queue.size();
                    OICS_BATCH_TRY(ctx->fb_in.ensure((size_t)FQ * card_bytes));
                    OICS_BATCH_CUDA(cudaMemcpy2DAsync(ctx->fb_in.as<uint8_t>() + (size_t)slot * card_bytes, row, card, dstride, row,
                                                      (size_t)h, cudaMemcpyDeviceToDevice, st));
                    Deferred d;
                    d.c = c;
                    d.cand.assign(cand.begin(), cand.begin() + std::min<size_t>(cand.size(), (size_t)std::max(0, p.n_candidates)));
                    d.p = p;
                    queue.push_back(std::move(d));
                    statuses[c] = OICS_OK;
                    angles[c] = 0.0;
                    need_checks[c] = 0;
                    if (dst_sizes) dst_sizes[c] = 0;
                    if ((int)queue.size() == FQ) OICS_BATCH_TRY(flush_fallbacks());
                    continue;
                }
            }
            double Mw[6];
            int dw = 0, dh = 0;
            if (stc == OICS_OK) stc = oics_rotate_geometry(w, h, a, 1.0, OICS_CLIP_CONTAIN, Mw, &dw, &dh);
            angles[c] = a;
            need_checks[c] = need;
            if (dst_ws) dst_ws[c] = dw;
            if (dst_hs) dst_hs[c] = dh;
            uint8_t *dst = dsts ? dsts[c] : nullptr;
            if (dst_sizes) dst_sizes[c] = 0;
            if (stc == OICS_OK && dst && jpeg_quality) {
                // deskew into the raw scratch, encode into the caller's device buffer or the download slot
                const size_t need_bytes = (size_t)dw * 3 * dh;
                if (need_bytes > out_card_cap) {
                    set_error("internal: CONTAIN bound %zu < %zu", out_card_cap, need_bytes);
                    stc = OICS_ERR_CAPACITY;
                } else {
                    // the card's deskew + encode chain goes to one of the encode lanes (own stream and scratch, ordered
                    // behind everything ctx->stream has been given so far; joined again below)
                    JpegLane *lane = nullptr;
                    if (n_lanes > 0) {
                        const int li = lanes_used % n_lanes;
                        OICS_BATCH_TRY(jpeg_lane_get(ctx, li, &lane));
                        if (lanes_used < n_lanes) {
                            if (lanes_used == 0) OICS_BATCH_CUDA(cudaEventRecord(lane_go, st));
                            OICS_BATCH_CUDA(cudaStreamWaitEvent(lane->st, lane_go, 0));
                        }
                        lanes_used++;
                        OICS_BATCH_TRY(lane->raw.ensure(out_card_cap));
                    }
                    uint8_t *raw = lane ? lane->raw.as<uint8_t>() : ctx->jpeg_raw.as<uint8_t>();
                    stc = warp_device(ctx, card, 3, w, h, dstride, Mw, OICS_INTER_NEAREST, white, raw, dw, dh, (size_t)dw * 3,
                                      lane ? lane->st : nullptr, lane ? &lane->tabs : nullptr);
                    if (stc == OICS_OK) {
                        // no host round trip here: the sizes of the sub-batch are read back together below
                        uint8_t *file = dst;
                        size_t cap = dst_caps[c];
                        if (!is_device_ptr(dst)) {
                            out_off[i] = (size_t)i * out_card_cap;
                            file = out_slot[s]->as<uint8_t>() + out_off[i];
                            cap = std::min(out_card_cap, dst_caps[c]);
                        }
                        stc = jpeg_encode_enqueue(ctx, raw, dw, dh, (size_t)dw * 3, jpeg_quality, file, cap,
                                                  ctx->jpeg_totals.as<unsigned long long>() + 2 * i, &jpeg_hdr, lane);
                        if (stc == OICS_OK) jpeg_cap[i] = cap;
                    }
                }
            } else if (stc == OICS_OK && dst) {
                const size_t need_bytes = (size_t)dw * 3 * dh;
                if (dst_sizes) dst_sizes[c] = need_bytes;
                if (need_bytes > dst_caps[c]) {
                    set_error("deskewed card needs %zu bytes, capacity %zu", need_bytes, dst_caps[c]);
                    stc = OICS_ERR_CAPACITY;
                } else if (is_device_ptr(dst)) {
                    stc = warp_device(ctx, card, 3, w, h, dstride, Mw, OICS_INTER_NEAREST, white, dst, dw, dh, (size_t)dw * 3);
                } else if (need_bytes > out_card_cap) {
                    set_error("internal: CONTAIN bound %zu < %zu", out_card_cap, need_bytes);
                    stc = OICS_ERR_CAPACITY;
                } else {
                    out_off[i] = (size_t)i * out_card_cap;
                    out_bytes[i] = need_bytes;
                    stc = warp_device(ctx, card, 3, w, h, dstride, Mw, OICS_INTER_NEAREST, white,
                                      out_slot[s]->as<uint8_t>() + out_off[i], dw, dh, (size_t)dw * 3);
                }
            }
            statuses[c] = stc;
            if (stc == OICS_ERR_CUDA) rc = OICS_ERR_CUDA;
        }
        for (int li = 0; li < std::min(lanes_used, n_lanes); li++) {      // ctx->stream continues behind the lanes
            JpegLane *lane = nullptr;
            OICS_BATCH_TRY(jpeg_lane_get(ctx, li, &lane));
            OICS_BATCH_CUDA(cudaEventRecord(lane->done, lane->st));
            OICS_BATCH_CUDA(cudaStreamWaitEvent(st, lane->done, 0));
        }
        if (rc != OICS_OK) break;
        if (trace) tr[3] = now();
        if (!dev_in) {
            OICS_BATCH_CUDA(cudaEventRecord(in_free[s], st));
            if (files) {
                // in the shadow of the deskew + encode kernels just enqueued: the next sub-batch's cards are
                // written (its chains had this sub-batch's sweep to converge), then the one after it is begun
                // (the two sub-batches use different slot sets: sub-batch k + 2 is begun FIRST, so its chains have
                // this whole deskew + encode phase and the next sweep to converge, and k + 1, begun a sub-batch ago,
                // is normally converged by now: `finish` does not wait and its cards are written beside the encoder)
                if (k + 2 < n_sub) OICS_BATCH_TRY(decode_begin(k + 2));
                if (trace) tr[4] = now();
                if (k + 1 < n_sub) OICS_BATCH_TRY(decode_finish(k + 1));
            } else if (k + 2 < n_sub) {
                OICS_BATCH_CUDA(upload(k + 2));
            }
        }
        if (trace) tr[5] = now();
        if (jpeg_quality && dsts) {   // file sizes of the sub-batch: one read-back
            std::vector<unsigned long long> totals((size_t)2 * nb, 0ull);
            OICS_BATCH_CUDA(cudaMemcpyAsync(totals.data(), ctx->jpeg_totals.p, totals.size() * sizeof(unsigned long long),
                                            cudaMemcpyDeviceToHost, st));
            OICS_BATCH_CUDA(cudaStreamSynchronize(st));
            for (int i = 0; i < nb; i++) {
                if (!jpeg_cap[i]) continue;
                const size_t fsz = jpeg_file_size(jpeg_hdr, &totals[(size_t)2 * i]);
                dst_sizes[c0 + i] = fsz;
                if (fsz > jpeg_cap[i]) {
                    set_error("jpeg: the file needs %zu bytes, capacity %zu", fsz, jpeg_cap[i]);
                    statuses[c0 + i] = OICS_ERR_CAPACITY;
                } else if (!is_device_ptr(dsts[c0 + i])) {
                    out_bytes[i] = fsz;
                }
            }
        }
        if (trace) tr[6] = now();
        if (host_out) {
            OICS_BATCH_CUDA(cudaEventRecord(out_ready[s], st));
            OICS_BATCH_CUDA(cudaStreamWaitEvent(down, out_ready[s], 0));
            for (int i = 0; i < nb; i++)
                if (out_bytes[i])
                    OICS_BATCH_CUDA(cudaMemcpyAsync(dsts[c0 + i], out_slot[s]->as<uint8_t>() + out_off[i], out_bytes[i],
                                                    cudaMemcpyDeviceToHost, down));
            OICS_BATCH_CUDA(cudaEventRecord(out_free[s], down));
        }
        if (trace) {
            tr[7] = now();
            fprintf(stderr, "[oics batch] sub %d at %.2f ms: P0 enqueue %.2f | sweep (waits) %.2f | rule A + warp/encode enqueue %.2f | "
                            "decode begin %.2f | decode finish %.2f | sizes read-back (waits) %.2f | copies out enqueue %.2f\n",
                    k, tr[0] - t_start, tr[1] - tr[0], tr[2] - tr[1], tr[3] - tr[2], tr[4] - tr[3], tr[5] - tr[4], tr[6] - tr[5], tr[7] - tr[6]);
        }
    }
    if (rc == OICS_OK) OICS_BATCH_TRY(flush_fallbacks());
    cleanup();
#undef OICS_BATCH_TRY
#undef OICS_BATCH_CUDA
    return rc;
}

OICS_API int32_t oics_correct_batch(oics_ctx *ctx, const uint8_t *bgrs, int32_t n_cards, size_t card_stride,
                                    int32_t w, int32_t h, int32_t stride, uint16_t max_angle, double step,
                                    int32_t max_w, int32_t max_h, double hough_min_line_length,
                                    double hough_max_line_gap, uint8_t *const *dsts, const size_t *dst_caps,
                                    int32_t *dst_ws, int32_t *dst_hs, double *angles, int32_t *need_checks,
                                    int32_t *statuses)
{
    return correct_batch_impl(ctx, bgrs, n_cards, card_stride, w, h, stride, max_angle, step, max_w, max_h,
                              hough_min_line_length, hough_max_line_gap, 0, dsts, dst_caps, nullptr, dst_ws, dst_hs, angles,
                              need_checks, statuses);
}

OICS_API int32_t oics_correct_batch_files(oics_ctx *ctx, const uint8_t *const *files, const size_t *file_sizes, int32_t n_cards,
                                          uint16_t max_angle, double step, int32_t max_w, int32_t max_h,
                                          double hough_min_line_length, double hough_max_line_gap, int32_t jpeg_quality,
                                          uint8_t *const *dsts, const size_t *dst_caps, size_t *dst_sizes, int32_t *dst_ws,
                                          int32_t *dst_hs, double *angles, int32_t *need_checks, int32_t *statuses)
{
    OICS_ARG(files && file_sizes && n_cards >= 0);
    if (n_cards == 0) return OICS_OK;
    int w = 0, h = 0, nc = 0;
    OICS_ARG(files[0]);
    OICS_TRY(jpeg_info_host(files[0], file_sizes[0], &w, &h, &nc));   // the batch's card size; other sizes fail per card
    return correct_batch_impl(ctx, nullptr, n_cards, (size_t)w * 3 * h, w, h, w * 3, max_angle, step, max_w, max_h,
                              hough_min_line_length, hough_max_line_gap, jpeg_quality, dsts, dst_caps, dst_sizes, dst_ws,
                              dst_hs, angles, need_checks, statuses, files, file_sizes);
}

OICS_API int32_t oics_correct_batch_jpeg(oics_ctx *ctx, const uint8_t *bgrs, int32_t n_cards, size_t card_stride,
                                         int32_t w, int32_t h, int32_t stride, uint16_t max_angle, double step,
                                         int32_t max_w, int32_t max_h, double hough_min_line_length,
                                         double hough_max_line_gap, int32_t jpeg_quality, uint8_t *const *dsts,
                                         const size_t *dst_caps, size_t *dst_sizes, int32_t *dst_ws, int32_t *dst_hs,
                                         double *angles, int32_t *need_checks, int32_t *statuses)
{
    OICS_ARG(jpeg_quality >= 1 && jpeg_quality <= 100);
    return correct_batch_impl(ctx, bgrs, n_cards, card_stride, w, h, stride, max_angle, step, max_w, max_h,
                              hough_min_line_length, hough_max_line_gap, jpeg_quality, dsts, dst_caps, dst_sizes, dst_ws,
                              dst_hs, angles, need_checks, statuses);
}

}  // extern "C"
