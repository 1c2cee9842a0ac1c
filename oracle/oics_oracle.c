/*
 * oics_oracle.c -- CPU restatement of the deskew hot path of ch1ny/omr-img-corrector.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity oracle and the timed CPU
 * baseline.  Nothing in the product (omr_img_corrector_b200/, include/) links,
 * imports or executes it; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may.
 *
 * Where the arithmetic lives: the reference's Rust crate `oics`
 * (packages/lib) delegates every pixel operation to OpenCV 4.6.0 through the
 * `opencv = "0.77.0"` binding (packages/lib/Cargo.toml:9).  OpenCV is a
 * third-party dependency that is NOT in /root/reference, and the Rust crate
 * itself cannot be compiled here (no cargo/rustc).  So each function below
 * restates (a) the reference's own Rust glue, cited by file:line, and (b) the
 * published OpenCV algorithm that glue calls.  Pinning: the reference holds no
 * golden vectors; every OpenCV restatement here is pinned
 * bit-exactly against the real OpenCV build importable in this image
 * (python cv2 4.13.0) by tests/test_oracle_vs_cv2.py and by the golden
 * fixtures in tests/golden/ minted from cv2 by tests/golden/make_golden.py.
 *
 * Build: see oracle/Makefile  (gcc -O2 -ffp-contract=off -pthread).
 * All f64/f32 arithmetic must stay un-contracted (no FMA) to match OpenCV's
 * baseline SSE2 build.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <pthread.h>

#define ORC_API __attribute__((visibility("default")))

static inline int orc_round(double v) { return (int)lrint(v); } /* cvRound: half-even */
static inline int orc_roundf(float v) { return (int)lrintf(v); }

/* ------------------------------------------------------------------------- */
/* P1: cv::getRotationMatrix2D as called at omr.rs:157-163, transfer.rs:472-475.
 * The centre is a Point2f (f32) promoted to f64 by OpenCV. */
ORC_API void orc_rotation_matrix_2d(float cx, float cy, double angle_deg, double scale,
                                    double M[6])
{
    double a = angle_deg * (M_PI / 180.0); /* OpenCV: angle *= CV_PI/180 (constant folded first) */
    double alpha = scale * cos(a);
    double beta = scale * sin(a);
    double x = (double)cx, y = (double)cy;
    M[0] = alpha;
    M[1] = beta;
    M[2] = (1 - alpha) * x - beta * y;
    M[3] = -beta;
    M[4] = alpha;
    M[5] = beta * x + (1 - alpha) * y;
}

/* cv::warpAffine front end (no WARP_INVERSE_MAP): invert the 2x3 in f64 in
 * OpenCV's exact operation order. */
ORC_API void orc_invert_affine(const double Min[6], double M[6])
{
    memcpy(M, Min, 6 * sizeof(double));
    double D = M[0] * M[4] - M[1] * M[3];
    D = D != 0 ? 1. / D : 0;
    double A11 = M[4] * D, A22 = M[0] * D;
    M[0] = A11;
    M[1] *= -D;
    M[3] *= -D;
    M[4] = A22;
    double b1 = -M[0] * M[2] - M[1] * M[5];
    double b2 = -M[3] * M[2] - M[4] * M[5];
    M[2] = b1;
    M[5] = b2;
}

/* W1 geometry: transfer::rotate_mat (transfer.rs:459-523) and the inlined copy
 * in omr::correct_default (omr.rs:408-445).  clip: 0 = DEFAULT, 1 = CONTAIN.
 * Writes the forward matrix and the destination size. */
ORC_API void orc_rotate_geometry(int cols, int rows, double angle, double scale, int clip,
                                 double M[6], int *dw, int *dh)
{
    if (clip == 0) {
        /* transfer.rs:472-475 : centre = (w as f32)/2, (h as f32)/2 */
        orc_rotation_matrix_2d((float)cols / 2.0f, (float)rows / 2.0f, angle, scale, M);
        *dw = cols;
        *dh = rows;
        return;
    }
    /* transfer.rs:487-519 */
    double r = angle * M_PI / 180.0; /* CV_PI */
    double rw = ceil((double)rows * fabs(sin(r)) + (double)cols * fabs(cos(r)));
    double rh = ceil((double)cols * fabs(sin(r)) + (double)rows * fabs(cos(r)));
    float cx = (float)ceil(rw / 2.0), cy = (float)ceil(rh / 2.0);
    orc_rotation_matrix_2d(cx, cy, angle, scale, M);
    M[2] += ceil((rw - (double)cols) / 2.0);
    M[5] += ceil((rh - (double)rows) / 2.0);
    *dw = (int)rw;
    *dh = (int)rh;
}

/* P2 / W1: cv::warpAffine, flags = INTER_NEAREST (0) -- the value of
 * imgproc::WARP_POLAR_LINEAR the reference passes (omr.rs:170, omr.rs:440,
 * projection.rs:52) -- BORDER_CONSTANT.  AB_BITS = 10 fixed point. */
ORC_API void orc_warp_affine_nn(const uint8_t *src, int w, int h, int sstride, int ch,
                                uint8_t *dst, int dw, int dh, int dstride, const double Mfwd[6],
                                const uint8_t border[4])
{
    double M[6];
    orc_invert_affine(Mfwd, M);
    int *adelta = (int *)malloc(sizeof(int) * 2 * (size_t)dw);
    int *bdelta = adelta + dw;
    for (int x = 0; x < dw; x++) {
        adelta[x] = orc_round(M[0] * x * 1024);
        bdelta[x] = orc_round(M[3] * x * 1024);
    }
    for (int y = 0; y < dh; y++) {
        int X0 = orc_round((M[1] * y + M[2]) * 1024) + 512;
        int Y0 = orc_round((M[4] * y + M[5]) * 1024) + 512;
        uint8_t *d = dst + (size_t)y * dstride;
        for (int x = 0; x < dw; x++) {
            int sx = (X0 + adelta[x]) >> 10;
            int sy = (Y0 + bdelta[x]) >> 10;
            /* OpenCV saturates to int16 here; irrelevant below 32767 px but kept */
            if (sx > 32767) sx = 32767;
            if (sx < -32768) sx = -32768;
            if (sy > 32767) sy = 32767;
            if (sy < -32768) sy = -32768;
            if ((unsigned)sx < (unsigned)w && (unsigned)sy < (unsigned)h) {
                const uint8_t *s = src + (size_t)sy * sstride + (size_t)sx * ch;
                for (int c = 0; c < ch; c++) d[x * ch + c] = s[c];
            } else {
                for (int c = 0; c < ch; c++) d[x * ch + c] = border[c];
            }
        }
    }
    free(adelta);
}

/* W1: cv::warpAffine, flags = INTER_LINEAR (1), BORDER_CONSTANT, as used for
 * the verifier's final rotation and input synthesis (core/src/main.rs:41-51,
 * 72-92; lib.rs:156-166).  INTER_BITS = 5, weights (32-fy)(32-fx)*32 which sum
 * to 32768 exactly, result (sum + 16384) >> 15. */
ORC_API void orc_warp_affine_linear(const uint8_t *src, int w, int h, int sstride, int ch,
                                    uint8_t *dst, int dw, int dh, int dstride,
                                    const double Mfwd[6], const uint8_t border[4])
{
    double M[6];
    orc_invert_affine(Mfwd, M);
    int *adelta = (int *)malloc(sizeof(int) * 2 * (size_t)dw);
    int *bdelta = adelta + dw;
    for (int x = 0; x < dw; x++) {
        adelta[x] = orc_round(M[0] * x * 1024);
        bdelta[x] = orc_round(M[3] * x * 1024);
    }
    for (int y = 0; y < dh; y++) {
        int X0 = orc_round((M[1] * y + M[2]) * 1024) + 16;
        int Y0 = orc_round((M[4] * y + M[5]) * 1024) + 16;
        uint8_t *d = dst + (size_t)y * dstride;
        for (int x = 0; x < dw; x++) {
            int X = (X0 + adelta[x]) >> 5, Y = (Y0 + bdelta[x]) >> 5;
            int sx = X >> 5, sy = Y >> 5, fx = X & 31, fy = Y & 31;
            if (sx > 32767) sx = 32767;
            if (sx < -32768) sx = -32768;
            if (sy > 32767) sy = 32767;
            if (sy < -32768) sy = -32768;
            int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32;
            int w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
            int in00 = (unsigned)sx < (unsigned)w && (unsigned)sy < (unsigned)h;
            int in01 = (unsigned)(sx + 1) < (unsigned)w && (unsigned)sy < (unsigned)h;
            int in10 = (unsigned)sx < (unsigned)w && (unsigned)(sy + 1) < (unsigned)h;
            int in11 = (unsigned)(sx + 1) < (unsigned)w && (unsigned)(sy + 1) < (unsigned)h;
            const uint8_t *s = src + (ptrdiff_t)sy * sstride + (ptrdiff_t)sx * ch;
            for (int c = 0; c < ch; c++) {
                int p00 = in00 ? s[c] : border[c];
                int p01 = in01 ? s[ch + c] : border[c];
                int p10 = in10 ? s[sstride + c] : border[c];
                int p11 = in11 ? s[sstride + ch + c] : border[c];
                int v = (p00 * w00 + p01 * w01 + p10 * w10 + p11 * w11 + 16384) >> 15;
                d[x * ch + c] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
            }
        }
    }
    free(adelta);
}

/* ------------------------------------------------------------------------- */
/* P0 pre-processing pieces (omr.rs:88-139, transfer.rs:283-301).            */

/* cvtColor(COLOR_RGB2GRAY) on BGR-ordered bytes (omr.rs:90, transfer.rs:286):
 * c0 is weighted as "R". 15-bit fixed point. */
ORC_API void orc_gray(const uint8_t *src, int w, int h, int sstride, uint8_t *dst, int dstride)
{
    for (int y = 0; y < h; y++) {
        const uint8_t *s = src + (size_t)y * sstride;
        uint8_t *d = dst + (size_t)y * dstride;
        for (int x = 0; x < w; x++)
            d[x] = (uint8_t)((s[3 * x] * 9798 + s[3 * x + 1] * 19235 + s[3 * x + 2] * 3735 + 16384) >> 15);
    }
}

/* erode with the 3x3 MORPH_ELLIPSE element (a plus-shaped cross), `iters`
 * iterations, BORDER_CONSTANT with morphologyDefaultBorderValue (= ignored)
 * (omr.rs:96-112).  Done literally, one iteration at a time. */
ORC_API void orc_erode_cross(const uint8_t *src, int w, int h, int sstride, uint8_t *dst,
                             int dstride, int iters)
{
    uint8_t *a = (uint8_t *)malloc((size_t)w * h), *b = (uint8_t *)malloc((size_t)w * h);
    for (int y = 0; y < h; y++) memcpy(a + (size_t)y * w, src + (size_t)y * sstride, w);
    for (int it = 0; it < iters; it++) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                uint8_t m = a[(size_t)y * w + x];
                if (x > 0 && a[(size_t)y * w + x - 1] < m) m = a[(size_t)y * w + x - 1];
                if (x + 1 < w && a[(size_t)y * w + x + 1] < m) m = a[(size_t)y * w + x + 1];
                if (y > 0 && a[(size_t)(y - 1) * w + x] < m) m = a[(size_t)(y - 1) * w + x];
                if (y + 1 < h && a[(size_t)(y + 1) * w + x] < m) m = a[(size_t)(y + 1) * w + x];
                b[(size_t)y * w + x] = m;
            }
        uint8_t *t = a;
        a = b;
        b = t;
    }
    for (int y = 0; y < h; y++) memcpy(dst + (size_t)y * dstride, a + (size_t)y * w, w);
    free(a);
    free(b);
}

/* cv::resize(..., INTER_AREA) for shrinking (scale_x, scale_y >= 1), 8-bit,
 * `ch` channels (omr.rs:116-126 on gray; transfer.rs:72-87 on colour).
 *   - identical size: copy
 *   - integer factors: ResizeAreaFast: integer box sum, (float)sum * (1.f/area),
 *     round half even; 2x2 has the special (s+2)>>2 path
 *   - otherwise: ResizeArea_ with the f32 tables of computeResizeAreaTab.
 * Returns 0 on success, -1 when the request is an enlargement (not restated). */
typedef struct { int si, di; float alpha; } orc_atab;

static int orc_area_tab(int ssize, int dsize, double scale, orc_atab *tab)
{
    int k = 0;
    for (int dx = 0; dx < dsize; dx++) {
        double fsx1 = dx * scale, fsx2 = fsx1 + scale;
        double cellWidth = scale < ssize - fsx1 ? scale : ssize - fsx1;
        int sx1 = (int)ceil(fsx1), sx2 = (int)floor(fsx2);
        if (sx2 > ssize - 1) sx2 = ssize - 1;
        if (sx1 > sx2) sx1 = sx2;
        if (sx1 - fsx1 > 1e-3) {
            tab[k].di = dx; tab[k].si = sx1 - 1;
            tab[k++].alpha = (float)((sx1 - fsx1) / cellWidth);
        }
        for (int sx = sx1; sx < sx2; sx++) {
            tab[k].di = dx; tab[k].si = sx;
            tab[k++].alpha = (float)(1.0 / cellWidth);
        }
        if (fsx2 - sx2 > 1e-3) {
            double t = fsx2 - sx2;
            if (t > 1.) t = 1.;
            if (t > cellWidth) t = cellWidth;
            tab[k].di = dx; tab[k].si = sx2;
            tab[k++].alpha = (float)(t / cellWidth);
        }
    }
    return k;
}

ORC_API int orc_resize_area(const uint8_t *src, int w, int h, int sstride, int ch, uint8_t *dst,
                            int dw, int dh, int dstride)
{
    if (dw <= 0 || dh <= 0) return -1;
    if (dw == w && dh == h) {
        for (int y = 0; y < h; y++) memcpy(dst + (size_t)y * dstride, src + (size_t)y * sstride, (size_t)w * ch);
        return 0;
    }
    double scale_x = (double)w / dw, scale_y = (double)h / dh;
    if (scale_x < 1 || scale_y < 1) return -1;
    int iscale_x = orc_round(scale_x), iscale_y = orc_round(scale_y);
    int fast = fabs(scale_x - iscale_x) < 2.220446049250313e-16 &&
               fabs(scale_y - iscale_y) < 2.220446049250313e-16;
    if (fast) {
        int area = iscale_x * iscale_y;
        float scale = 1.f / (float)area;
        for (int y = 0; y < dh; y++)
            for (int x = 0; x < dw; x++)
                for (int c = 0; c < ch; c++) {
                    int sum = 0;
                    for (int ky = 0; ky < iscale_y; ky++)
                        for (int kx = 0; kx < iscale_x; kx++)
                            sum += src[(size_t)(y * iscale_y + ky) * sstride + (size_t)(x * iscale_x + kx) * ch + c];
                    int v;
                    if (iscale_x == 2 && iscale_y == 2) v = (sum + 2) >> 2;
                    else v = orc_roundf((float)sum * scale);
                    dst[(size_t)y * dstride + x * ch + c] = (uint8_t)(v > 255 ? 255 : v < 0 ? 0 : v);
                }
        return 0;
    }
    orc_atab *xtab = (orc_atab *)malloc(sizeof(orc_atab) * ((size_t)w * 2 + 2 * dw));
    orc_atab *ytab = (orc_atab *)malloc(sizeof(orc_atab) * ((size_t)h * 2 + 2 * dh));
    int xn = orc_area_tab(w, dw, scale_x, xtab);
    int yn = orc_area_tab(h, dh, scale_y, ytab);
    int n = dw * ch;
    float *buf = (float *)malloc(sizeof(float) * 2 * (size_t)n), *sum = buf + n;
    int prev_dy = ytab[0].di;
    for (int i = 0; i < n; i++) sum[i] = 0;
    for (int j = 0; j < yn; j++) {
        float beta = ytab[j].alpha;
        int dy = ytab[j].di, sy = ytab[j].si;
        const uint8_t *S = src + (size_t)sy * sstride;
        for (int i = 0; i < n; i++) buf[i] = 0;
        for (int k = 0; k < xn; k++) {
            float alpha = xtab[k].alpha;
            for (int c = 0; c < ch; c++)
                buf[xtab[k].di * ch + c] = buf[xtab[k].di * ch + c] + (float)S[xtab[k].si * ch + c] * alpha;
        }
        if (dy != prev_dy) {
            uint8_t *D = dst + (size_t)prev_dy * dstride;
            for (int i = 0; i < n; i++) {
                int v = orc_roundf(sum[i]);
                D[i] = (uint8_t)(v > 255 ? 255 : v < 0 ? 0 : v);
                sum[i] = beta * buf[i];
            }
            prev_dy = dy;
        } else {
            for (int i = 0; i < n; i++) sum[i] = sum[i] + beta * buf[i];
        }
    }
    {
        uint8_t *D = dst + (size_t)prev_dy * dstride;
        for (int i = 0; i < n; i++) {
            int v = orc_roundf(sum[i]);
            D[i] = (uint8_t)(v > 255 ? 255 : v < 0 ? 0 : v);
        }
    }
    free(buf);
    free(xtab);
    free(ytab);
    return 0;
}

/* threshold(127, 255, THRESH_BINARY) (omr.rs:131-137, transfer.rs:297) */
ORC_API void orc_threshold(const uint8_t *src, int w, int h, int sstride, uint8_t *dst, int dstride)
{
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++)
            dst[(size_t)y * dstride + x] = src[(size_t)y * sstride + x] > 127 ? 255 : 0;
}

/* The scale rule of omr.rs:60-82 (no clamp to <= 1). */
ORC_API double orc_projection_scale(int w, int h, int max_w, int max_h)
{
    double ws = max_w <= 0 ? 1.0 : (double)max_w / (double)w;
    double hs = max_h <= 0 ? 1.0 : (double)max_h / (double)h;
    return ws < hs ? ws : hs;
}

/* ------------------------------------------------------------------------- */
/* P3: omr::get_mat_projection_data (omr.rs:8-39); same numbers as
 * transfer::get_horizontal_projection / get_vertical_projection
 * (transfer.rs:305-333, 380-405).  row[y] = #zeros in row y ("horizontal"),
 * col[x] = #zeros in column x ("vertical"). */
ORC_API void orc_projection_counts(const uint8_t *img, int w, int h, int stride, int32_t *row,
                                   int32_t *col)
{
    memset(col, 0, sizeof(int32_t) * (size_t)w);
    for (int y = 0; y < h; y++) {
        const uint8_t *p = img + (size_t)y * stride;
        int s = 0;
        for (int x = 0; x < w; x++)
            if (p[x] == 0) {
                s++;
                col[x]++;
            }
        row[y] = s;
    }
}

/* P4: calculate::get_arithmetic_mean + get_standard_deviation
 * (calculate.rs:2-23): sequential left-to-right f64 sums, population std-dev.
 * powf(2.0) is restated as x*x and powf(0.5) as sqrt (both correctly rounded;
 * libm pow agrees to <= 1 ulp -- tolerance 1e-4 rel in north_star). */
ORC_API double orc_std_dev_i32(const int32_t *v, int n)
{
    double sum = (double)v[0];
    for (int i = 1; i < n; i++) sum = sum + (double)v[i];
    double mean = sum / (double)n;
    double d = (double)v[0] - mean;
    double acc = d * d;
    for (int i = 1; i < n; i++) {
        d = (double)v[i] - mean;
        acc = acc + d * d;
    }
    return sqrt(acc / (double)n);
}

/* P2+P3+P4 for a list of forward matrices: what one iteration of the hot loop
 * does (omr.rs:153-208 / projection.rs:47-65), including writing the rotated
 * image and reading it back.  v_sd = std of column counts ("vertical"),
 * h_sd = std of row counts ("horizontal").  Angle-parallel over `threads`
 * pthreads with a shared work index (the reference is card-parallel, thread_pool.rs; the work per
 * angle is identical).  rows_out/cols_out (nullable) receive the count vectors
 * of every angle: [n][h] and [n][w]. */
typedef struct {
    const uint8_t *bin; int w, h, stride; const double *Mfwd; int n;
    double *v_sd, *h_sd; int32_t *rows_out, *cols_out;
    int *next; pthread_mutex_t *mu;
} orc_sweep_job;

static void *orc_sweep_worker(void *arg)
{
    orc_sweep_job *j = (orc_sweep_job *)arg;
    const uint8_t border[4] = {255, 255, 255, 0};
    int w = j->w, h = j->h;
    uint8_t *rot = (uint8_t *)malloc((size_t)w * h);
    int32_t *row = (int32_t *)malloc(sizeof(int32_t) * ((size_t)w + h));
    int32_t *col = row + h;
    for (;;) {
        pthread_mutex_lock(j->mu);
        int i = (*j->next)++;
        pthread_mutex_unlock(j->mu);
        if (i >= j->n) break;
        orc_warp_affine_nn(j->bin, w, h, j->stride, 1, rot, w, h, w, j->Mfwd + 6 * (size_t)i, border);
        orc_projection_counts(rot, w, h, w, row, col);
        j->h_sd[i] = orc_std_dev_i32(row, h);
        j->v_sd[i] = orc_std_dev_i32(col, w);
        if (j->rows_out) memcpy(j->rows_out + (size_t)i * h, row, sizeof(int32_t) * (size_t)h);
        if (j->cols_out) memcpy(j->cols_out + (size_t)i * w, col, sizeof(int32_t) * (size_t)w);
    }
    free(rot);
    free(row);
    return NULL;
}

ORC_API void orc_projection_sweep(const uint8_t *bin, int w, int h, int stride,
                                  const double *Mfwd, int n, double *v_sd, double *h_sd,
                                  int32_t *rows_out, int32_t *cols_out, int threads)
{
    if (threads < 1) threads = 1;
    if (threads > 1024) threads = 1024;
    int next = 0;
    pthread_mutex_t mu;
    pthread_mutex_init(&mu, NULL);
    orc_sweep_job job = {bin, w, h, stride, Mfwd, n, v_sd, h_sd, rows_out, cols_out, &next, &mu};
    if (threads == 1) {
        orc_sweep_worker(&job);
    } else {
        pthread_t *t = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
        for (int i = 0; i < threads; i++) pthread_create(&t[i], NULL, orc_sweep_worker, &job);
        for (int i = 0; i < threads; i++) pthread_join(t[i], NULL);
        free(t);
    }
    pthread_mutex_destroy(&mu);
}

/* Number of candidate half-range N = (max_angle / step) as u16 (omr.rs:140-141,
 * projection.rs:36): truncating, saturating cast. */
ORC_API int orc_angle_half_range(unsigned max_angle, double step)
{
    double q = (double)max_angle / step;
    if (!(q > 0)) return 0; /* NaN or <= 0 saturate to 0 */
    if (q >= 65535.0) return 65535;
    return (int)q;
}

/* P5a: argmax rule A of omr::get_result_from_projection (omr.rs:147-221).
 * status: 0 Believed, 1 NeedCheck, 2 NotAResult.  candidates receives up to
 * cap angles; *n_cand the true count. */
ORC_API void orc_rule_a(const double *h_sd, const double *v_sd, int N, double step,
                        double *angle, int *status, double *candidates, int cap, int *n_cand)
{
    double max_h = 0.0, max_v = 0.0;
    unsigned ph = 1, pv = 1;
    int nc = 0;
    for (int i = 0; i < 2 * N; i++) {
        int deg = i - N;
        double a = (double)deg * step;
        double hs = h_sd[i];
        if (max_h < hs) {
            max_h = hs;
            max_v = v_sd[i];
            ph = 1;
            pv = 1;
            nc = 0;
            if (nc < cap) candidates[nc] = a;
            nc = 1;
        } else if (max_h == hs) {
            ph++;
            double vs = v_sd[i];
            if (max_v < vs) {
                pv = 1;
                max_v = vs;
                nc = 0;
                if (nc < cap) candidates[nc] = a;
                nc = 1;
            } else if (max_v == vs) {
                pv++;
                if (nc < cap) candidates[nc] = a;
                nc++;
            }
        }
    }
    *n_cand = nc;
    if (ph == 1 && pv == 1 && nc >= 1) {
        *angle = candidates[0];
        *status = 0;
    } else if (nc == 1) {
        *angle = candidates[0];
        *status = 1;
    } else {
        *angle = 0.0;
        *status = 2;
    }
}

/* P5b: argmax rule B of projection::get_angle_with_projections
 * (projection.rs:125-190).  The reference iterates a HashMap (unspecified
 * order) when several candidates exist; ties on v^2+h^2 are resolved here in
 * ascending index order (documented divergence: the reference is
 * nondeterministic on that exact tie). Returns the winning index. */
ORC_API int orc_rule_b_index(const double *v_sd, const double *h_sd, int n)
{
    double vmax = v_sd[0], hmax = h_sd[0];
    int nv = 0, nh = 0, vfirst = 0, hfirst = 0;
    /* the reference seeds the list with index 0 and then visits index 0 again,
     * which appends 0 a second time (== branch); only the *set* matters later,
     * but list length matters for the "unique" test. */
    int vlen = 1, hlen = 1;
    for (int i = 0; i < n; i++) {
        if (v_sd[i] > vmax) { vmax = v_sd[i]; vlen = 1; vfirst = i; }
        else if (v_sd[i] == vmax) { vlen++; }
        if (h_sd[i] > hmax) { hmax = h_sd[i]; hlen = 1; hfirst = i; }
        else if (h_sd[i] == hmax) { hlen++; }
    }
    (void)nv; (void)nh;
    if (vlen == 1 && hlen == 1 && vfirst == hfirst) return vfirst;
    double sdp = 0.0;
    int best = -1;
    for (int i = 0; i < n; i++) {
        if (v_sd[i] == vmax || h_sd[i] == hmax) {
            double cur = v_sd[i] * v_sd[i] + h_sd[i] * h_sd[i];
            if (sdp < cur) { sdp = cur; best = i; }
        }
    }
    return best >= 0 ? best : n / 2;
}
