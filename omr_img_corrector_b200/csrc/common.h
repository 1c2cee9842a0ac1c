// Internal plumbing shared by the kernels' host wrappers: context, scratch
// buffers, error reporting, host/device pointer staging.  Not part of the ABI.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/oics_cuda.h"

namespace oics {

void set_error(const char *fmt, ...);

#define OICS_CUDA_TRY(expr)                                                             \
    do {                                                                                \
        cudaError_t _e = (expr);                                                        \
        if (_e != cudaSuccess) {                                                        \
            ::oics::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),   \
                              __FILE__, __LINE__);                                      \
            return OICS_ERR_CUDA;                                                       \
        }                                                                               \
    } while (0)

#define OICS_TRY(expr)                 \
    do {                               \
        int32_t _s = (expr);           \
        if (_s != OICS_OK) return _s;  \
    } while (0)

#define OICS_ARG(cond)                                              \
    do {                                                            \
        if (!(cond)) {                                              \
            ::oics::set_error("invalid argument: %s", #cond);       \
            return OICS_ERR_INVALID_ARG;                            \
        }                                                           \
    } while (0)

// Grow-only device scratch buffer.
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int32_t ensure(size_t bytes);
    void release();
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct PinnedBuf {
    void *p = nullptr;
    size_t cap = 0;
    int32_t ensure(size_t bytes);
    void release();
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

// One encode in flight (jpeg.cu): its stream and scratch.  The batch driver spreads the cards of a sub-batch over
// OICS_JPEG_LANES of them: one card's encoder is a chain of short kernels with scans between them, several
// chains side by side keep the device busy through each other's tails.
#define OICS_JPEG_LANES 4
struct JpegLane {
    cudaStream_t st = nullptr;
    cudaEvent_t done = nullptr;
    DevBuf tabs, raw, coef, bits, words, cnt;   // warp tables, deskewed pixels, coefficients, bit counts + sums, bit stream, 0xFF counts + sums
};

// Bluestein / radix-2 plan of one 1-D DFT length (fft.cu): device tables built on first use.
struct FftPlan {
    int n = 0, m = 0, direct = 0;
    DevBuf chirp, bbr, tw;
};

// What the sweep last built on the device, so that calls with the same geometry (a batch
// run, the bench loop) skip the table / plan kernels.
struct SweepCache {
    int w = 0, h = 0;
    std::vector<double> inv;        // inverse matrices the tables were built from
    int plan_a0 = -1, plan_a1 = -1; // angle range the word-kernel planes cover
    bool small = false;             // geometry served by the small-card kernel (sweeps.cu): boxes instead of classes
    bool boxes_ready = false;
    int angles_per_chunk = 0;
    // classification (meta kernel verdicts) and the launch lists derived from it
    std::vector<int> general;       // angles the general kernel takes
    std::vector<int> fast;          // angles the word kernel takes, ascending
    // groups of consecutive fast angles sharing a staged tile, sorted by (chunk, class
    // descending, first angle): {index into `fast`, count, class, first angle}
    struct Group { int start, count, kr, first; };
    std::vector<Group> groups;
    std::vector<int> chunk_group_begin;  // per angle chunk: first group (size n_chunks + 1)
    size_t groups_off = 0;               // offset (ints) in ctx->wlists
    size_t general_off = 0, kf_off = 0, fast_off = 0;
};

// Opt-in per-kernel timing (oics_profile / oics_profile_report): a ProfScope around a launch records two
// CUDA events on the launching stream, so the figure is the kernel's own device time inside the real
// pipeline (no profiler attached, no extra synchronisation).  `bytes` = the ALGORITHMIC bytes of that
// launch, from which the bench derives achieved GB/s per kernel.
struct ProfEntry {
    const char *name;
    cudaEvent_t e0, e1;
    double bytes;
};
struct Profiler {
    bool on = false;
    std::vector<ProfEntry> entries;
    std::vector<cudaEvent_t> pool;
};

}  // namespace oics

#define OICS_JD_SLOTS 65   // slot 0 = ctx->stream; two sets of 32 for the batch driver (sub-batches alternate)

struct oics_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;      // stream all work is issued on
    cudaStream_t own_stream = nullptr;  // created by oics_create
    std::mutex mu;
    int sm_count = 0;
    int smem_optin = 0;
    int64_t launches = 0;
    // per-device function attributes / occupancy answers this context has already set up (a context is
    // bound to one device; process-wide statics would leak one device's answer to another)
    bool attr_sweepw = false, attr_sweeps = false, attr_jdec = false, attr_sweep[5] = {false, false, false, false, false};
    int canny_blocks_per_sm = 0;
    size_t counts_clean = 0;            // leading bytes of `counts` known to be zero (the finalize pass clears what it read)
    // sweep timing evidence
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    double last_sweep_ms = 0.0;
    int last_sweep_launches = 0;
    // scratch
    oics::DevBuf in_img, in_img2, out_img, out_img2, gray, tmp8a, tmp8b, bin, bits, mats, tabs, counts, sd,
        sums, edges, mag, dxy, labels, accum, lines, misc, cplx, cplx2, twid, fmag;
    oics::PinnedBuf pin_in, pin_out, pin_small;
    oics::FftPlan fft_row, fft_col;
    oics::DevBuf wplanes, wmeta, wlists, jpeg_raw, jpeg_hdr, jpeg_totals;
    oics::DevBuf fb_in, fb_out;   // batch driver: cards waiting for the batched Hough fallback, and their outputs
    oics::JpegLane jpeg_lanes[OICS_JPEG_LANES];
    void *jd_slots[OICS_JD_SLOTS] = {};   // jpegdec.cu: decodes in flight (slot 0 = ctx->stream, the others own a stream)
    // jpeg.cu: the file header of the last quality used (host copy and its device cache)
    int jpeg_hdr_quality = 0, jpeg_hdr_sof = 0;
    size_t jpeg_hdr_len = 0;
    uint8_t jpeg_hdr_host[1024];
    oics::SweepCache sweep_cache;
    oics::Profiler prof;
};

namespace oics {

bool is_device_ptr(const void *p);

// Makes `src` (host or device, h rows of row_bytes with `stride`) available on the
// device; returns the device pointer and its stride.  Host data is copied tightly
// packed into `scratch`.
int32_t stage_in(oics_ctx *ctx, const void *src, size_t row_bytes, int h, size_t stride, DevBuf &scratch,
                 const uint8_t **dptr, size_t *dstride);
// Copies a tightly packed device image to `dst` (host or device) with dst_stride.
int32_t stage_out(oics_ctx *ctx, const void *dsrc, size_t row_bytes, int h, void *dst, size_t dst_stride);
// Flat copies that pick the direction from the pointer kinds.
int32_t copy_out(oics_ctx *ctx, void *dst, const void *dsrc, size_t bytes);
int32_t copy_in(oics_ctx *ctx, void *ddst, const void *src, size_t bytes);

struct CtxGuard {
    std::unique_lock<std::mutex> lk;
    int32_t status;
    bool ranged = false;
    // `entry` names the C-ABI entry point: an NVTX range (visible in Nsight Systems / ncu --nvtx) spans the call
    explicit CtxGuard(oics_ctx *ctx, const char *entry = nullptr);
    ~CtxGuard();
};

// RAII timing scope around one kernel launch (or a tight group of launches of one kernel).
struct ProfScope {
    oics_ctx *ctx;
    cudaStream_t st;
    int idx = -1;
    ProfScope(oics_ctx *ctx, const char *name, double bytes, cudaStream_t stream = nullptr);
    ~ProfScope();
};

// ---- host logic (hostlogic.cpp) ----
void invert_affine(const double Min[6], double M[6]);

// ---- kernels' host wrappers (all enqueue on ctx->stream, no sync) ----
// sweep.cu
int32_t sweep_device(oics_ctx *ctx, const uint8_t *dbins, int n_cards, size_t card_stride, int w, int h,
                     size_t stride, const double *M_host, int n_angles, double *d_vsd, double *d_hsd,
                     int64_t *d_sums, int32_t *d_counts_out /* nullable: [n_angles][h+w] of card 0 */);
// warp.cu
int32_t warp_device(oics_ctx *ctx, const uint8_t *dsrc, int ch, int w, int h, size_t sstride, const double Mfwd[6],
                    int interp, const uint8_t border[4], uint8_t *ddst, int dw, int dh, size_t dstride, cudaStream_t on = nullptr,
                    DevBuf *tables = nullptr);
// preprocess.cu
int32_t preprocess_device(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, int mode, int ow,
                          int oh, uint8_t *dbin /* tightly packed ow x oh */);
int32_t preprocess_batch_device(oics_ctx *ctx, const uint8_t *dbgrs, int n_cards, size_t card_stride, int w, int h,
                                size_t stride, int mode, int ow, int oh, uint8_t *dbins /* n x ow x oh */);
int32_t gray_device(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, uint8_t *dgray);
int32_t resize_area_device(oics_ctx *ctx, const uint8_t *dsrc, int ch, int w, int h, size_t stride, uint8_t *ddst, int dw,
                           int dh);
// canny.cu
int32_t canny_device(oics_ctx *ctx, const uint8_t *dsrc, int ch, int w, int h, size_t stride, int low, int high,
                     uint8_t *dedges /* tightly packed */);
// hough.cu
int32_t hough_accumulate_device(oics_ctx *ctx, const uint8_t *dedges, int w, int h, size_t stride, double rho,
                                double theta, int32_t *daccum);
int32_t hough_lines_p_device(oics_ctx *ctx, const uint8_t *dedges, int w, int h, size_t stride, double rho,
                             double theta, int threshold, int line_length, int line_gap, int32_t *h_lines,
                             int cap, int *n_lines);
int32_t hough_lines_p_batch_device(oics_ctx *ctx, const uint8_t *const *dedges, int n_img, int w, int h, size_t stride,
                                   double rho, double theta, int threshold, int line_length, int line_gap,
                                   std::vector<std::vector<int32_t>> *lines_out);
void hough_geometry(int w, int h, double rho, double theta, int *numangle, int *numrho, std::vector<float> *ttab);
int32_t hough_peaks_device(oics_ctx *ctx, const int32_t *daccum, int numangle, int numrho, int threshold,
                           std::vector<int32_t> *peaks /* {theta, rho, count} triples, sorted */);
// fft.cu
int32_t fft_image_device(oics_ctx *ctx, const uint8_t *dgray, int w, int h, size_t stride, uint8_t *dmag8,
                         uint8_t *dlog8);
// N2: imwrite(JPEG, quality) of a BGR image on the device -> file bytes on the device (jpeg.cu)
int32_t jpeg_encode_device(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, int quality, uint8_t *d_out,
                           size_t d_cap, size_t *size);

int32_t jpeg_encode_enqueue(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, int quality, uint8_t *d_out,
                            size_t d_cap, unsigned long long *d_totals, size_t *hdr_len, JpegLane *lane = nullptr);
int32_t jpeg_lane_get(oics_ctx *ctx, int i, JpegLane **lane);
void jpeg_lanes_release(oics_ctx *ctx);
size_t jpeg_file_size(size_t hdr_len, const unsigned long long *totals);
int32_t exclusive_scan(oics_ctx *ctx, cudaStream_t st, unsigned long long *v, size_t n, unsigned long long *sums,
                       unsigned long long *total, const unsigned long long *tb);
// N2: imread(JPEG) (jpegdec.cu)
int32_t jpeg_info_host(const uint8_t *file, size_t n, int *w, int *h, int *ncomp);
int32_t jpeg_decode_device(oics_ctx *ctx, const uint8_t *file, size_t n, uint8_t *d_bgr, size_t stride, int *w_out, int *h_out);
int32_t jpeg_decode_batch_begin(oics_ctx *ctx, int slot0, const uint8_t *const *files, const size_t *sizes, int n_files, int w, int h,
                                int32_t *statuses);
int32_t jpeg_decode_batch_finish(oics_ctx *ctx, int slot0, int n_files, int w, uint8_t *d_bgr, size_t card_bytes, cudaEvent_t free_ev,
                                 cudaStream_t join, cudaEvent_t ready, int32_t *statuses);
void jpeg_decode_release(oics_ctx *ctx);

inline int cdiv(int a, int b) { return (a + b - 1) / b; }

}  // namespace oics
