/*
 * oics_cuda.h -- C ABI of the B200-native deskew hot path of ch1ny/omr-img-corrector.
 *
 * This header is the drop-in boundary.  The reference has no FFI of its own on
 * this path: callers (`core`, the Tauri host) link the Rust crate `oics`
 * (packages/lib/Cargo.toml:16-18) whose public functions call OpenCV.  Each
 * export below replaces the body of one of those public functions (cited as
 * reference file:line, relative to the reference repo root); INTEGRATION.md
 * shows the Rust `extern "C"` shim a maintainer adds.
 *
 * Conventions
 *  - Every function returns OICS_OK (0) or a negative oics_status; the message
 *    for the calling thread is available from oics_last_error().  Nothing
 *    aborts, nothing throws across the ABI.
 *  - Images are 8-bit, row-major, `stride` bytes between rows, `ch` interleaved
 *    channels (BGR order for 3-channel, as cv::imread produces).
 *  - Every image / result pointer may be a HOST pointer (pageable or pinned) or
 *    a CUDA DEVICE pointer on the context's device; the library detects which
 *    (cudaPointerGetAttributes) and stages host buffers itself.  Small scalar
 *    outputs (angles, sizes, status) are always host pointers.
 *  - The caller owns every buffer; the library keeps no caller pointer after a
 *    call returns.  All calls on one context serialise on an internal mutex;
 *    use one context per host thread for concurrency (the reference runs one
 *    OS thread per card, packages/app/src-tauri/src/thread_pool.rs:41-73).
 *  - There is NO CPU fallback: without a CUDA device every compute entry
 *    returns OICS_ERR_CUDA.
 */
#ifndef OICS_CUDA_H
#define OICS_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define OICS_API __declspec(dllexport)
#else
#define OICS_API __attribute__((visibility("default")))
#endif

typedef enum oics_status {
    OICS_OK = 0,
    OICS_ERR_INVALID_ARG = -1,
    OICS_ERR_CUDA = -2,
    OICS_ERR_NO_LINES = -3,      /* the reference panics here (omr.rs:272, hough.rs:74) */
    OICS_ERR_UNIMPLEMENTED = -4,
    OICS_ERR_NOMEM = -5,
    OICS_ERR_UNSUPPORTED = -6,   /* e.g. enlarging resize */
    OICS_ERR_CAPACITY = -7       /* an output buffer was too small; needed size is reported */
} oics_status;

/* omr::ResultStatus (omr.rs:41-45) */
enum { OICS_BELIEVED = 0, OICS_NEED_CHECK = 1, OICS_NOT_A_RESULT = 2 };
/* types::RotateClipStrategy (types.rs:7-11) */
enum { OICS_CLIP_DEFAULT = 0, OICS_CLIP_CONTAIN = 1 };
/* interpolation flags the reference passes to warp_affine: 0 = INTER_NEAREST
 * (== imgproc::WARP_POLAR_LINEAR, omr.rs:170), 1 = INTER_LINEAR (core/src/main.rs:76) */
enum { OICS_INTER_NEAREST = 0, OICS_INTER_LINEAR = 1 };
/* pre-processing flavour */
enum { OICS_PRE_OMR = 0 /* omr.rs:60-139: gray, erode x3, area resize, threshold */,
       OICS_PRE_PROJECTION = 1 /* projection.rs:24-32: colour area resize, gray, threshold */ };
/* segment-angle rule */
enum { OICS_LINES_OMR = 0 /* omr.rs:257-301 */, OICS_LINES_HOUGH = 1 /* hough.rs:48-89 */,
       OICS_LINES_FFT = 2 /* fft.rs:195-246 */ };

typedef struct oics_ctx oics_ctx;

/* omr::OmrResult (omr.rs:46-50).  `candidates` is caller storage of `cap`
 * doubles (may be NULL with cap 0); n_candidates is the true count. */
typedef struct oics_omr_result {
    double angle;
    int32_t status;
    int32_t n_candidates;
    double *candidates;
    int32_t cap;
} oics_omr_result;

/* ---- context ----------------------------------------------------------- */
OICS_API int32_t oics_create(int32_t device, oics_ctx **out);
OICS_API void oics_destroy(oics_ctx *ctx);
OICS_API const char *oics_last_error(void);
OICS_API int32_t oics_version(void);
/* Run all work of this context on an existing CUDA stream (cudaStream_t); NULL
 * restores the context's own stream. */
OICS_API int32_t oics_set_stream(oics_ctx *ctx, void *cuda_stream);
OICS_API int32_t oics_synchronize(oics_ctx *ctx);
/* Number of kernels this context has launched since creation (bench evidence). */
OICS_API int64_t oics_launch_count(oics_ctx *ctx);
/* Page-locked host memory for the buffers a caller hands to the batch entries: copies from / to pinned memory
 * are asynchronous and overlap the kernels (pageable buffers are staged by the driver and serialise the
 * pipeline).  No context needed; free with oics_free_pinned.  The reference allocates its Mats on the heap
 * (imread, transfer.rs:43): this replaces that allocation on the caller's side of the boundary. */
OICS_API int32_t oics_alloc_pinned(size_t bytes, void **out);
OICS_API void oics_free_pinned(void *p);
/* Opt-in per-kernel timing of everything this context launches (CUDA events around each launch on its own
 * stream; no profiler, no added synchronisation).  oics_profile(ctx, 1) clears and starts, (ctx, 0) stops.
 * oics_profile_report synchronises the device and writes a JSON array of
 * {"kernel", "launches", "ms", "algorithmic_bytes"} aggregated by kernel name into buf (OICS_ERR_CAPACITY
 * with *needed set when cap is too small).  Measurement plumbing: no reference counterpart. */
OICS_API int32_t oics_profile(oics_ctx *ctx, int32_t enable);
OICS_API int32_t oics_profile_report(oics_ctx *ctx, char *buf, size_t cap, size_t *needed);
/* Device time, in ms, of the sweep kernel launches of the most recent sweep call
 * (CUDA events on the launching stream) and how many launches that was. */
OICS_API int32_t oics_last_sweep_kernel_ms(oics_ctx *ctx, double *ms, int32_t *launches);
/* Bandwidth probes for the roofline block of the bench (the sweep of omr.rs:147-221
 * re-reads one card 2N times, so its memory ceilings are on-chip; MEASURED_PEAKS.json has none).
 * oics_probe_l2_read: every CTA streams an L2-resident buffer of `bytes` `reps` times with
 * L1-bypassing 16-byte loads; oics_probe_smem_read: every thread re-reads its CTA's 64 KB
 * shared-memory tile with 16-byte loads.  Both time themselves with CUDA events on the context's
 * stream and return GB/s of the whole device.  Measurement only; no product call uses them. */
OICS_API int32_t oics_probe_l2_read(oics_ctx *ctx, size_t bytes, int32_t reps, double *gb_per_s);
OICS_API int32_t oics_probe_smem_read(oics_ctx *ctx, int32_t iters, double *gb_per_s);

/* ---- host-side geometry (pure functions, no device) --------------------- */
/* cv::getRotationMatrix2D as called at omr.rs:157-163 / transfer.rs:472-475. */
OICS_API void oics_rotation_matrix_2d(float cx, float cy, double angle_deg, double scale, double M[6]);
/* N = (max_angle / step) as u16 (omr.rs:140-141, projection.rs:36). */
OICS_API int32_t oics_angle_half_range(uint32_t max_angle, double step);
/* The 2N forward matrices of the sweep, angle_i = (i - N) as f64 * step about
 * ((w as f32)/2, (h as f32)/2) with `scale` (omr.rs:153-163).  M: [2N][6]. */
OICS_API int32_t oics_sweep_matrices(int32_t w, int32_t h, uint32_t max_angle, double step,
                                     double scale, double *M, int32_t cap_angles, int32_t *n_angles);
/* transfer::rotate_mat geometry (transfer.rs:459-523; inlined at omr.rs:408-445). */
OICS_API int32_t oics_rotate_geometry(int32_t cols, int32_t rows, double angle, double scale,
                                      int32_t clip, double M[6], int32_t *dst_w, int32_t *dst_h);
/* Single-huge-image mode (BASELINE config 5): the contiguous slice of the n_angles sweep
 * candidates that `rank` of `world` evaluates (ceil(n/world) each, the tail ranks may get
 * fewer or none).  No reference counterpart: the reference sweeps on one CPU. */
OICS_API int32_t oics_shard_range(int32_t n_angles, int32_t rank, int32_t world, int32_t *first,
                                  int32_t *count);
/* Gather layout of the sharded sweep: per = ceil(n_angles / world) slots per rank.  The exchange buffer is
 * sd[world][2][per] doubles -- rank r's slice of v_sd, then of h_sd, each padded to `per` with zeros -- so ONE
 * fixed-size all-gather of 2 * per doubles per rank (ncclAllGather in place over NVLink) completes it. */
OICS_API int32_t oics_shard_layout(int32_t n_angles, int32_t world, int32_t *per_rank);
/* Unpacks a gathered sd[world][2][per] buffer (host) into contiguous v_sd[n_angles], h_sd[n_angles]. */
OICS_API int32_t oics_shard_unpack(const double *sd, int32_t n_angles, int32_t world, double *v_sd, double *h_sd);
/* scale rule of omr.rs:60-82 */
OICS_API double oics_projection_scale(int32_t w, int32_t h, int32_t max_w, int32_t max_h);

/* ---- P0 / P0': pre-processing ------------------------------------------ */
/* mode OICS_PRE_OMR: scale from (max_w, max_h) (omr.rs:60-82), `resize_scale` ignored.
 * mode OICS_PRE_PROJECTION: scale = resize_scale (projection.rs:24-27), max_* ignored.
 * out_bin: 8UC1 {0,255}, tightly packed (stride = *out_w), capacity out_cap bytes. */
OICS_API int32_t oics_preprocess(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h,
                                 int32_t stride, int32_t mode, int32_t max_w, int32_t max_h,
                                 double resize_scale, uint8_t *out_bin, size_t out_cap,
                                 int32_t *out_w, int32_t *out_h, double *out_scale);

/* ---- P2-P4: the fused projection sweep ---------------------------------- */
/* For each forward matrix M[i] (as get_rotation_matrix_2d returns): NN-rotate the
 * binary card with border 255 (omr.rs:165-173), count zero pixels per row and per
 * column (omr.rs:8-39) and return the population std-devs of those counts
 * (calculate.rs:13-23): v_sd[i] = std of column counts ("vertical"),
 * h_sd[i] = std of row counts ("horizontal").  No rotated image is materialised.
 * sums (nullable): [n][4] int64 = {sum rows, sum rows^2, sum cols, sum cols^2}. */
OICS_API int32_t oics_projection_sweep(oics_ctx *ctx, const uint8_t *bin, int32_t w, int32_t h,
                                       int32_t stride, const double *M, int32_t n_angles,
                                       double *v_sd, double *h_sd, int64_t *sums);
/* The same over a batch of equally sized cards laid out `card_stride` bytes
 * apart (config 4).  v_sd/h_sd: [n_cards][n_angles]. */
OICS_API int32_t oics_projection_sweep_batch(oics_ctx *ctx, const uint8_t *bins, int32_t n_cards,
                                             size_t card_stride, int32_t w, int32_t h, int32_t stride,
                                             const double *M, int32_t n_angles, double *v_sd,
                                             double *h_sd);
/* Single-huge-image mode (BASELINE config 5): the angle loop of omr.rs:153-208 / projection.rs:47-65 split
 * over `world` GPUs.  Every rank holds the image and calls this with the SAME M[n_angles][6]; the rank
 * evaluates its contiguous slice (oics_shard_range) and writes it into its block of the exchange buffer
 * sd[world][2][per] (oics_shard_layout; host or device memory, the other ranks' blocks are not touched).
 * The caller all-gathers the buffer (NCCL; 16 bytes per angle) and applies oics_rule_a / oics_rule_b to the
 * unpacked arrays on every rank: the argmax has tie bookkeeping, so it is gather + local rule, not a max
 * all-reduce.  first / count (nullable) return the slice. */
OICS_API int32_t oics_sweep_sharded(oics_ctx *ctx, const uint8_t *bin, int32_t w, int32_t h, int32_t stride,
                                    const double *M, int32_t n_angles, int32_t rank, int32_t world, double *sd,
                                    int32_t *first, int32_t *count);
/* Parity hook for P3: the count vectors of one candidate angle. */
OICS_API int32_t oics_projection_counts(oics_ctx *ctx, const uint8_t *bin, int32_t w, int32_t h,
                                        int32_t stride, const double M[6], int32_t *row_counts,
                                        int32_t *col_counts);
/* transfer::get_projection_standard_deviations (transfer.rs:527-536) of an image as is. */
OICS_API int32_t oics_projection_standard_deviations(oics_ctx *ctx, const uint8_t *bin, int32_t w,
                                                     int32_t h, int32_t stride, double *v_sd, double *h_sd);
/* NO counterpart in the reference: added for callers that may not
 * call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).  The same result as
 * oics_projection_standard_deviations, bit for bit, but it allocates NOTHING and takes no context: the
 * zero counts live in a file-scope static __device__ array (noalloc.cu) of 128 KB that every GPU loading
 * the library holds until exit.  Limit: w <= 16384 and h <= 16384 -- beyond that OICS_ERR_CAPACITY.
 * d_bin (8-bit, 1 channel) and d_sd (double[2]: [0] = v_sd, [1] = h_sd) are DEVICE memory of the CURRENT
 * device; the three kernels are enqueued on cuda_stream (NULL = the default stream) and the call does not
 * synchronize.  All calls on a device share the array (it is cleared at the start of each call): calls on
 * one stream follow each other safely, calls that overlap on the GPU race. */
OICS_API int32_t oics_projection_standard_deviations_static(const uint8_t *d_bin, int32_t w, int32_t h,
                                                            int32_t stride, double *d_sd, void *cuda_stream);

/* ---- P5: argmax rules (host, exact scalar logic) ------------------------ */
/* rule A: omr.rs:147-221 */
OICS_API int32_t oics_rule_a(const double *h_sd, const double *v_sd, int32_t N, double step,
                             oics_omr_result *result);
/* rule A for every card of a batch: h_sd, v_sd are [n_cards][2N]; per card the selected angle,
 * the status (OICS_BELIEVED / NEED_CHECK / NOT_A_RESULT) and the number of candidates. */
OICS_API int32_t oics_rule_a_batch(const double *h_sd, const double *v_sd, int32_t n_cards, int32_t N,
                                   double step, double *angles, int32_t *statuses, int32_t *n_candidates);
/* rule B: projection.rs:125-190 -> angle */
OICS_API int32_t oics_rule_b(const double *v_sd, const double *h_sd, int32_t N, double step, double *angle);

/* ---- estimator entry points -------------------------------------------- */
/* omr::get_result_from_projection (omr.rs:52-58) */
OICS_API int32_t oics_result_from_projection(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h,
                                             int32_t stride, uint16_t max_angle, double step,
                                             int32_t max_w, int32_t max_h, oics_omr_result *result);
/* projection::get_angle_with_projections (projection.rs:17-23).  `threads` is
 * accepted and ignored: single-thread semantics (the reference's threads>1 branch
 * rotates by whole degrees, projection.rs:94 -- SURVEY F6). */
OICS_API int32_t oics_angle_with_projections(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h,
                                             int32_t stride, uint16_t max_angle, double step,
                                             double resize_scale, size_t threads, double *angle);
/* omr::get_result_from_edges_detection (omr.rs:231-235); src has ch = 1 or 3. */
OICS_API int32_t oics_result_from_edges_detection(oics_ctx *ctx, const uint8_t *src, int32_t ch,
                                                  int32_t w, int32_t h, int32_t stride,
                                                  double min_line_length, double max_line_gap,
                                                  oics_omr_result *result);
/* The same for n equally sized cards `card_stride` bytes apart (host or device) IN FLIGHT TOGETHER: the
 * reference runs one OS thread per card (thread_pool.rs:41-73); here the cards' Canny passes are followed by
 * ONE HoughLinesP launch with a thread-block cluster per card.  Per card: the angle, the ResultStatus
 * (OICS_BELIEVED / NEED_CHECK / NOT_A_RESULT) and statuses[i] = OICS_OK or that card's error (e.g.
 * OICS_ERR_NO_LINES where the reference panics).  Results are identical to n single calls. */
OICS_API int32_t oics_result_from_edges_detection_batch(oics_ctx *ctx, const uint8_t *srcs, int32_t n_cards,
                                                        size_t card_stride, int32_t ch, int32_t w, int32_t h,
                                                        int32_t stride, double min_line_length, double max_line_gap,
                                                        double *angles, int32_t *result_statuses, int32_t *statuses);
/* hough::get_angle_with_hough (hough.rs:17-23) without the debug image. */
OICS_API int32_t oics_angle_with_hough(oics_ctx *ctx, const uint8_t *gray, int32_t w, int32_t h,
                                       int32_t stride, double min_line_length, double max_line_gap,
                                       double *angle);
/* fft::get_angle_with_fft (fft.rs:145-154) without the debug image. */
OICS_API int32_t oics_angle_with_fft(oics_ctx *ctx, const uint8_t *gray, int32_t w, int32_t h,
                                     int32_t stride, double canny1, double canny2,
                                     double min_line_length, double max_line_gap, double *angle);
/* omr::get_result_from_fourier_transform (omr.rs:304-310) */
OICS_API int32_t oics_result_from_fourier_transform(oics_ctx *ctx, const uint8_t *bgr, int32_t w,
                                                    int32_t h, int32_t stride, double canny_weak,
                                                    double canny_strong, double min_line_length,
                                                    double max_line_gap, oics_omr_result *result);

/* ---- H1/H2/H3 building blocks ------------------------------------------- */
/* cv::Canny(src, lo, hi, 3, false) (hough.rs:27, omr.rs:239, fft.rs:164-171) */
OICS_API int32_t oics_canny(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w, int32_t h,
                            int32_t stride, double lo, double hi, uint8_t *edges, int32_t edges_stride);
/* accumulator size for (w, h, rho, theta) */
OICS_API int32_t oics_hough_dims(int32_t w, int32_t h, double rho, double theta, int32_t *numangle,
                                 int32_t *numrho);
/* standard accumulator over all edge pixels, HoughLinesP vote formula: accum[numangle][numrho] */
OICS_API int32_t oics_hough_accumulate(oics_ctx *ctx, const uint8_t *edges, int32_t w, int32_t h,
                                       int32_t stride, double rho, double theta, int32_t *accum);
/* Peak extraction over a standard accumulator accum[numangle][numrho] (host or device), the local-maximum
 * rule of cv::HoughLines (findLocalMaximums + sort): a cell is a peak when its count is > threshold, > its
 * left (rho - 1) and upper (theta - 1) neighbours and >= its right and lower ones (outside = 0); peaks are
 * ordered by count descending, then by (theta, rho) ascending, and the first `cap` are stored as
 * peaks[i] = {theta index, rho index, count}; the line is rho = (r - (numrho - 1) / 2) * rho_step,
 * theta = n * theta_step.  *n_peaks = number found (OICS_ERR_CAPACITY if > cap).  The reference calls
 * hough_lines_p, not this (hough.rs:31-42); it completes the north star's "votes ... then extracts peaks". */
OICS_API int32_t oics_hough_peaks(oics_ctx *ctx, const int32_t *accum, int32_t numangle, int32_t numrho,
                                  int32_t threshold, int32_t *peaks, int32_t cap, int32_t *n_peaks);
/* cv::HoughLinesP (hough.rs:31-42, omr.rs:245-253, fft.rs:183-191): lines [cap][4] int32,
 * *n_lines = number found (OICS_ERR_CAPACITY if > cap; the first cap are stored). */
OICS_API int32_t oics_hough_lines_p(oics_ctx *ctx, const uint8_t *edges, int32_t w, int32_t h,
                                    int32_t stride, double rho, double theta, int32_t threshold,
                                    double min_line_length, double max_line_gap, int32_t *lines,
                                    int32_t cap, int32_t *n_lines);
/* segment-angle rules (host): variant OICS_LINES_* ; OICS_ERR_NO_LINES when n == 0
 * for the two variants where the reference panics. */
OICS_API int32_t oics_lines_angle(const int32_t *lines, int32_t n, int32_t variant,
                                  oics_omr_result *result);

/* ---- F1: FFT image ------------------------------------------------------- */
/* fft::get_fft_image (fft.rs:124-141): mag8 (nullable) and log8, 8UC1 tightly packed. */
OICS_API int32_t oics_fft_image(oics_ctx *ctx, const uint8_t *gray, int32_t w, int32_t h,
                                int32_t stride, uint8_t *mag8, uint8_t *log8);

/* ---- W1: final correction ------------------------------------------------ */
/* cv::warpAffine(flags = interp, BORDER_CONSTANT, border) with forward matrix M. */
OICS_API int32_t oics_warp_affine(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w, int32_t h,
                                  int32_t stride, const double M[6], int32_t interp,
                                  const uint8_t border[4], uint8_t *dst, int32_t dst_w, int32_t dst_h,
                                  int32_t dst_stride);
/* transfer::rotate_mat (transfer.rs:459-467).  dst tightly packed (stride = dst_w*ch),
 * capacity dst_cap bytes; size from oics_rotate_geometry. */
OICS_API int32_t oics_rotate(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w, int32_t h,
                             int32_t stride, double angle, double scale, int32_t interp,
                             const uint8_t border[4], int32_t clip, uint8_t *dst, size_t dst_cap,
                             int32_t *dst_w, int32_t *dst_h);

/* ---- T1: the hot members of transfer::TransformableMatrix ------------------------ */
/* TransformableMatrix::scale_self (transfer.rs:66-91): cv::resize to (((w as f64) * scale) as i32,
 * ((h as f64) * scale) as i32) with INTER_AREA for scale < 1; scale == 1 returns the image unchanged
 * (the early-out at :67-69).  scale > 1 (INTER_LINEAR enlargement) is not on the hot path -- every
 * caller shrinks (core/src/main.rs:70, shrink_to) -- and returns OICS_ERR_UNSUPPORTED.  ch = 1 or 3;
 * dst tightly packed (stride = *dst_w * ch), capacity dst_cap bytes. */
OICS_API int32_t oics_scale(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w, int32_t h, int32_t stride,
                            double scale, uint8_t *dst, size_t dst_cap, int32_t *dst_w, int32_t *dst_h);
/* TransformableMatrix::shrink_to (transfer.rs:93-126; called by app/src-tauri/src/test.rs:309-317):
 * scale = min(max_w / w, max_h / h) with a non-positive bound meaning "no bound"; never upscales
 * (scale >= 1 returns the image unchanged), else scale_self(scale).  *out_scale (nullable) = the scale used. */
OICS_API int32_t oics_shrink_to(oics_ctx *ctx, const uint8_t *src, int32_t ch, int32_t w, int32_t h, int32_t stride,
                                int32_t max_w, int32_t max_h, uint8_t *dst, size_t dst_cap, int32_t *dst_w,
                                int32_t *dst_h, double *out_scale);

/* ---- N2: the file written at the end of omr::correct_default ---------------- */
/* TransformableMatrix::im_write(path, ImageFormat::JPEG, quality) (transfer.rs:162-179, called with
 * quality 100 at omr.rs:447-448) = cv::imwrite: the bytes of the JPEG file OpenCV writes for a BGR
 * image (JFIF, 4:2:0, standard Huffman tables), byte for byte.  bgr and out may be host or device
 * pointers.  *out_size = file size; OICS_ERR_CAPACITY (with the needed size) if out_cap is too small. */
OICS_API int32_t oics_jpeg_encode(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h, int32_t stride,
                                  int32_t quality, uint8_t *out, size_t out_cap, size_t *out_size);

/* ---- one-call pipeline ---------------------------------------------------- */
/* reconcile rule of omr::correct_default (omr.rs:352-403), host */
OICS_API int32_t oics_reconcile(const oics_omr_result *projection, double edges_angle, double *angle,
                                int32_t *need_check);
/* omr::correct_default (omr.rs:339-348) on memory buffers: imread/imwrite stay with the
 * caller.  dst (nullable: estimate only) receives the CONTAIN NN rotation of the BGR card. */
OICS_API int32_t oics_correct_default(oics_ctx *ctx, const uint8_t *bgr, int32_t w, int32_t h,
                                      int32_t stride, uint16_t max_angle, double step, int32_t max_w,
                                      int32_t max_h, double hough_min_line_length,
                                      double hough_max_line_gap, uint8_t *dst, size_t dst_cap,
                                      int32_t *dst_w, int32_t *dst_h, double *angle, int32_t *need_check);

/* N2, input side: imgcodecs::imread(path, IMREAD_COLOR) (omr.rs:349, transfer.rs:43,62) for a baseline
 * JPEG file already in host memory.  oics_jpeg_info parses the header only (host, no context): image
 * size and component count.  oics_jpeg_decode decodes on the GPU -- the pixels are libjpeg(-turbo)'s /
 * cv::imread's bit for bit -- into `bgr` (host or device, rows `stride` bytes apart, 0 = packed).
 * OICS_ERR_UNSUPPORTED for files outside the covered subset (progressive, non-interleaved scans, CMYK,
 * subsampling other than 4:4:4 / 4:2:2 / 4:2:0, an Exif orientation other than upright): the caller keeps
 * imread for those. */
OICS_API int32_t oics_jpeg_info(const uint8_t *file, size_t file_size, int32_t *w, int32_t *h, int32_t *channels);
OICS_API int32_t oics_jpeg_decode(oics_ctx *ctx, const uint8_t *file, size_t file_size, uint8_t *bgr, size_t stride,
                                  size_t bgr_cap, int32_t *w, int32_t *h);

/* Batch form replacing the thread-pool fan-out (task.rs:19-68, thread_pool.rs:56-73):
 * n equally sized BGR cards `card_stride` bytes apart (host or device); per card the
 * selected angle, need_check and status (OICS_OK or an error for that card only).
 * dsts (nullable) = array of n pointers (each nullable) with dst_caps bytes each. */
OICS_API int32_t oics_correct_batch(oics_ctx *ctx, const uint8_t *bgrs, int32_t n_cards,
                                    size_t card_stride, int32_t w, int32_t h, int32_t stride,
                                    uint16_t max_angle, double step, int32_t max_w, int32_t max_h,
                                    double hough_min_line_length, double hough_max_line_gap,
                                    uint8_t *const *dsts, const size_t *dst_caps, int32_t *dst_ws,
                                    int32_t *dst_hs, double *angles, int32_t *need_checks,
                                    int32_t *statuses);
/* The same with the output side of omr::correct_default included (omr.rs:447-448): dsts[c] receives the
 * JPEG file imwrite(JPEG, jpeg_quality) would write for card c (the reference uses 100), dst_sizes[c]
 * its size -- a few MB cross PCIe instead of the 30 MB of pixels. */
OICS_API int32_t oics_correct_batch_jpeg(oics_ctx *ctx, const uint8_t *bgrs, int32_t n_cards, size_t card_stride,
                                         int32_t w, int32_t h, int32_t stride, uint16_t max_angle, double step,
                                         int32_t max_w, int32_t max_h, double hough_min_line_length,
                                         double hough_max_line_gap, int32_t jpeg_quality, uint8_t *const *dsts,
                                         const size_t *dst_caps, size_t *dst_sizes, int32_t *dst_ws, int32_t *dst_hs,
                                         double *angles, int32_t *need_checks, int32_t *statuses);

/* The whole of omr::correct_default for a batch (omr.rs:339-452 including imread at :349 and imwrite at
 * :447-448): files[c] / file_sizes[c] are baseline JPEG files in host memory (pinned for asynchronous
 * copies), decoded on the device -- a few MB per card cross PCIe in each direction and the host does no
 * codec work.  All cards must have the size of files[0]; a file that cannot be decoded (or differs in
 * size) gets its error in statuses[c] and the batch goes on.  jpeg_quality 0: dsts receive BGR pixels,
 * 1..100: the JPEG file imwrite would produce, its size in dst_sizes[c]. */
OICS_API int32_t oics_correct_batch_files(oics_ctx *ctx, const uint8_t *const *files, const size_t *file_sizes, int32_t n_cards,
                                          uint16_t max_angle, double step, int32_t max_w, int32_t max_h,
                                          double hough_min_line_length, double hough_max_line_gap, int32_t jpeg_quality,
                                          uint8_t *const *dsts, const size_t *dst_caps, size_t *dst_sizes, int32_t *dst_ws,
                                          int32_t *dst_hs, double *angles, int32_t *need_checks, int32_t *statuses);

#ifdef __cplusplus
}
#endif
#endif /* OICS_CUDA_H */
