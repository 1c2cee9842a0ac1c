/*
 * oics_oracle_lines.c -- CPU restatement of the edge/line half of the deskew
 * hot path (Canny, HoughLinesP, segment-angle modes, correct_default's
 * reconcile rule).  TEST INFRASTRUCTURE ONLY -- see the header of
 * oics_oracle.c for the rules and for how these restatements are pinned
 * (bit-exact against the OpenCV build importable here, cv2 4.13.0).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------- */
/* H1: cv::Canny(src, lo, hi, apertureSize=3, L2gradient=false) as called at
 * hough.rs:27 (gray), omr.rs:239 (3-channel BGR) and fft.rs:164-171.
 * Sobel 3x3 with BORDER_REPLICATE in int16, L1 magnitude, for ch>1 the first
 * channel with maximal magnitude per pixel, TG22 fixed-point sector test,
 * zero-padded magnitude border, 8-connected hysteresis. */
ORC_API void orc_canny(const uint8_t *src, int w, int h, int sstride, int ch, double lo,
                       double hi, uint8_t *edges, int estride)
{
    if (lo > hi) { double t = lo; lo = hi; hi = t; }
    int low = (int)floor(lo), high = (int)floor(hi);
    size_t n = (size_t)w * h;
    int16_t *dx = (int16_t *)malloc(sizeof(int16_t) * n * 2), *dy = dx + n;
    int *mag = (int *)calloc((size_t)(w + 2) * (h + 2), sizeof(int));
    uint8_t *map = (uint8_t *)malloc(n); /* 0 candidate, 1 no edge, 2 edge */
#define SRC(yy, xx, c) ((int)src[(size_t)((yy) < 0 ? 0 : (yy) >= h ? h - 1 : (yy)) * sstride + (size_t)((xx) < 0 ? 0 : (xx) >= w ? w - 1 : (xx)) * ch + (c)])
#define MAG(yy, xx) mag[(size_t)((yy) + 1) * (w + 2) + (xx) + 1]
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            int bm = -1, bdx = 0, bdy = 0;
            for (int c = 0; c < ch; c++) {
                int gx = (SRC(y - 1, x + 1, c) + 2 * SRC(y, x + 1, c) + SRC(y + 1, x + 1, c)) -
                         (SRC(y - 1, x - 1, c) + 2 * SRC(y, x - 1, c) + SRC(y + 1, x - 1, c));
                int gy = (SRC(y + 1, x - 1, c) + 2 * SRC(y + 1, x, c) + SRC(y + 1, x + 1, c)) -
                         (SRC(y - 1, x - 1, c) + 2 * SRC(y - 1, x, c) + SRC(y - 1, x + 1, c));
                int m = abs(gx) + abs(gy);
                if (m > bm) { bm = m; bdx = gx; bdy = gy; }
            }
            dx[(size_t)y * w + x] = (int16_t)bdx;
            dy[(size_t)y * w + x] = (int16_t)bdy;
            MAG(y, x) = bm;
        }
    size_t *stack = (size_t *)malloc(sizeof(size_t) * n);
    size_t sp = 0;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            size_t i = (size_t)y * w + x;
            int m = MAG(y, x);
            int keep = 0;
            if (m > low) {
                int xs = dx[i], ys = dy[i];
                int ax = abs(xs), ay = abs(ys) << 15;
                int tg22x = ax * 13573;
                if (ay < tg22x) {
                    keep = m > MAG(y, x - 1) && m >= MAG(y, x + 1);
                } else {
                    int tg67x = tg22x + (ax << 16);
                    if (ay > tg67x) {
                        keep = m > MAG(y - 1, x) && m >= MAG(y + 1, x);
                    } else {
                        int s = (xs ^ ys) < 0 ? -1 : 1;
                        keep = m > MAG(y - 1, x - s) && m > MAG(y + 1, x + s);
                    }
                }
            }
            if (!keep) map[i] = 1;
            else if (m > high) { map[i] = 2; stack[sp++] = i; }
            else map[i] = 0;
        }
    while (sp) {
        size_t i = stack[--sp];
        int y = (int)(i / w), x = (int)(i % w);
        for (int ddy = -1; ddy <= 1; ddy++)
            for (int ddx = -1; ddx <= 1; ddx++) {
                int yy = y + ddy, xx = x + ddx;
                if ((ddx | ddy) == 0 || yy < 0 || yy >= h || xx < 0 || xx >= w) continue;
                size_t j = (size_t)yy * w + xx;
                if (map[j] == 0) { map[j] = 2; stack[sp++] = j; }
            }
    }
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++)
            edges[(size_t)y * estride + x] = map[(size_t)y * w + x] == 2 ? 255 : 0;
#undef SRC
#undef MAG
    free(stack); free(map); free(mag); free(dx);
}

/* ------------------------------------------------------------------------- */
/* Shared Hough geometry: rho, theta are narrowed to f32 inside OpenCV. */
typedef struct { int numangle, numrho; float *ttab; } orc_hgeom;

static void orc_hough_geom(int w, int h, double rho, double theta, orc_hgeom *g)
{
    float rho_f = (float)rho, theta_f = (float)theta;
    float irho = 1 / rho_f;
    /* computeNumangle(0, CV_PI, theta) */
    int numangle = (int)floor((M_PI - 0.0) / (double)theta_f) + 1;
    if (numangle > 1 && fabs(M_PI - (numangle - 1) * (double)theta_f) < (double)theta_f / 2) --numangle;
    g->numangle = numangle;
    g->numrho = (int)lrint(((w + h) * 2 + 1) / (double)rho_f);
    g->ttab = (float *)malloc(sizeof(float) * 2 * (size_t)numangle);
    for (int n = 0; n < numangle; n++) {
        g->ttab[2 * n] = (float)(cos((double)n * (double)theta_f) * (double)irho);
        g->ttab[2 * n + 1] = (float)(sin((double)n * (double)theta_f) * (double)irho);
    }
}

static inline int orc_vote_r(const orc_hgeom *g, int n, int x, int y)
{
    float v = (float)x * g->ttab[2 * n] + (float)y * g->ttab[2 * n + 1];
    return (int)lrintf(v) + (g->numrho - 1) / 2;
}

ORC_API void orc_hough_dims(int w, int h, double rho, double theta, int *numangle, int *numrho)
{
    orc_hgeom g;
    orc_hough_geom(w, h, rho, theta, &g);
    *numangle = g.numangle;
    *numrho = g.numrho;
    free(g.ttab);
}

/* A.8: the standard (rho, theta) accumulator over ALL edge pixels with the
 * HoughLinesP vote formula.  Not something the reference calls; it is the
 * CPU definition the north_star's "accumulator bit-exact" refers to. */
ORC_API void orc_hough_accumulate(const uint8_t *edges, int w, int h, int stride, double rho,
                                  double theta, int32_t *accum)
{
    orc_hgeom g;
    orc_hough_geom(w, h, rho, theta, &g);
    memset(accum, 0, sizeof(int32_t) * (size_t)g.numangle * g.numrho);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++)
            if (edges[(size_t)y * stride + x])
                for (int n = 0; n < g.numangle; n++) accum[(size_t)n * g.numrho + orc_vote_r(&g, n, x, y)]++;
    free(g.ttab);
}

/* Peak extraction of cv::HoughLines (HoughLinesStandard: findLocalMaximums + sort by hough_cmp_gt) over a
 * given accumulator accum[numangle][numrho]: OpenCV keeps a zero frame around the accumulator, so
 * neighbours outside it read 0.  A cell is kept when
 *     accum > threshold && accum > left && accum >= right && accum > up && accum >= down
 * (left/right = rho -/+ 1, up/down = theta -/+ 1); cells are ordered by count descending, ties by cell
 * index ascending.  peaks: up to cap x 3 int32 {theta index, rho index, count}; returns the number found.
 * Not called by the reference (it uses hough_lines_p, hough.rs:31-42); it is the CPU definition of the
 * north star's "extracts peaks".  Pinned against cv2.HoughLines in tests/test_oracle_vs_cv2.py. */
typedef struct { int32_t cnt; int32_t cell; } orc_peak;
static int orc_peak_cmp(const void *a, const void *b)
{
    const orc_peak *p = (const orc_peak *)a, *q = (const orc_peak *)b;
    if (p->cnt != q->cnt) return p->cnt > q->cnt ? -1 : 1;
    return p->cell < q->cell ? -1 : p->cell > q->cell;
}
ORC_API int orc_hough_peaks(const int32_t *accum, int numangle, int numrho, int threshold, int32_t *peaks, int cap)
{
    orc_peak *list = (orc_peak *)malloc(sizeof(orc_peak) * ((size_t)numangle * numrho / 2 + numangle + 1));
    int n = 0;
    for (int a = 0; a < numangle; a++)
        for (int r = 0; r < numrho; r++) {
            const int32_t *row = accum + (size_t)a * numrho;
            const int v = row[r];
            const int left = r > 0 ? row[r - 1] : 0, right = r + 1 < numrho ? row[r + 1] : 0;
            const int up = a > 0 ? row[r - numrho] : 0, down = a + 1 < numangle ? row[r + numrho] : 0;
            if (v > threshold && v > left && v >= right && v > up && v >= down) {
                list[n].cnt = v;
                list[n].cell = a * numrho + r;
                n++;
            }
        }
    qsort(list, (size_t)n, sizeof(orc_peak), orc_peak_cmp);
    for (int i = 0; i < n && i < cap; i++) {
        peaks[3 * i + 0] = list[i].cell / numrho;
        peaks[3 * i + 1] = list[i].cell % numrho;
        peaks[3 * i + 2] = list[i].cnt;
    }
    free(list);
    return n;
}

/* H2: cv::HoughLinesP(edges, rho, theta, threshold, minLineLength, maxLineGap)
 * = HoughLinesProbabilistic, as called at hough.rs:31-42, omr.rs:245-253
 * (threshold 0) and fft.rs:183-191 (threshold 100).  Sequential, RNG-ordered
 * (cv::RNG seeded with (uint64)-1).  lines: up to cap x 4 int32; returns the
 * number of segments found (may exceed cap; only cap are stored). */
ORC_API int orc_hough_lines_p(const uint8_t *edges, int w, int h, int stride, double rho,
                              double theta, int threshold, double min_len, double max_gap,
                              int32_t *lines, int cap)
{
    orc_hgeom g;
    orc_hough_geom(w, h, rho, theta, &g);
    int lineLength = (int)lrint(min_len), lineGap = (int)lrint(max_gap);
    int numangle = g.numangle, numrho = g.numrho;
    int32_t *accum = (int32_t *)calloc((size_t)numangle * numrho, sizeof(int32_t));
    uint8_t *mask = (uint8_t *)malloc((size_t)w * h);
    int32_t *nz = (int32_t *)malloc(sizeof(int32_t) * 2 * (size_t)w * h);
    int count = 0, nlines = 0;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            if (edges[(size_t)y * stride + x]) {
                mask[(size_t)y * w + x] = 1;
                nz[2 * count] = x;
                nz[2 * count + 1] = y;
                count++;
            } else
                mask[(size_t)y * w + x] = 0;
        }
    uint64_t state = (uint64_t)-1;
    for (; count > 0; count--) {
        state = (uint64_t)(uint32_t)state * 4164903690u + (uint32_t)(state >> 32);
        int idx = (int)((uint32_t)state % (uint32_t)count);
        int max_val = threshold - 1, max_n = 0;
        int j = nz[2 * idx], i = nz[2 * idx + 1];
        int line_end[2][2] = {{0, 0}, {0, 0}}; /* [k] = {x, y} */
        nz[2 * idx] = nz[2 * (count - 1)];
        nz[2 * idx + 1] = nz[2 * (count - 1) + 1];
        if (!mask[(size_t)i * w + j]) continue;
        for (int n = 0; n < numangle; n++) {
            int r = orc_vote_r(&g, n, j, i);
            int val = ++accum[(size_t)n * numrho + r];
            if (max_val < val) { max_val = val; max_n = n; }
        }
        if (max_val < threshold) continue;
        float a = -g.ttab[2 * max_n + 1], b = g.ttab[2 * max_n];
        int x0 = j, y0 = i, dx0, dy0, xflag;
        const int shift = 16;
        if (fabsf(a) > fabsf(b)) {
            xflag = 1;
            dx0 = a > 0 ? 1 : -1;
            dy0 = (int)lrint(b * (1 << shift) / fabs(a));
            y0 = (y0 << shift) + (1 << (shift - 1));
        } else {
            xflag = 0;
            dy0 = b > 0 ? 1 : -1;
            dx0 = (int)lrint(a * (1 << shift) / fabs(b));
            x0 = (x0 << shift) + (1 << (shift - 1));
        }
        for (int k = 0; k < 2; k++) {
            int gap = 0, x = x0, y = y0, dx = dx0, dy = dy0;
            if (k > 0) { dx = -dx; dy = -dy; }
            for (;; x += dx, y += dy) {
                int i1, j1;
                if (xflag) { j1 = x; i1 = y >> shift; }
                else { j1 = x >> shift; i1 = y; }
                if (j1 < 0 || j1 >= w || i1 < 0 || i1 >= h) break;
                if (mask[(size_t)i1 * w + j1]) {
                    gap = 0;
                    line_end[k][1] = i1;
                    line_end[k][0] = j1;
                } else if (++gap > lineGap)
                    break;
            }
        }
        int good = abs(line_end[1][0] - line_end[0][0]) >= lineLength ||
                   abs(line_end[1][1] - line_end[0][1]) >= lineLength;
        for (int k = 0; k < 2; k++) {
            int x = x0, y = y0, dx = dx0, dy = dy0;
            if (k > 0) { dx = -dx; dy = -dy; }
            for (;; x += dx, y += dy) {
                int i1, j1;
                if (xflag) { j1 = x; i1 = y >> shift; }
                else { j1 = x >> shift; i1 = y; }
                uint8_t *m = mask + (size_t)i1 * w + j1;
                if (*m) {
                    if (good)
                        for (int n = 0; n < numangle; n++) accum[(size_t)n * numrho + orc_vote_r(&g, n, j1, i1)]--;
                    *m = 0;
                }
                if (i1 == line_end[k][1] && j1 == line_end[k][0]) break;
            }
        }
        if (good) {
            if (nlines < cap) {
                lines[4 * nlines] = line_end[0][0];
                lines[4 * nlines + 1] = line_end[0][1];
                lines[4 * nlines + 2] = line_end[1][0];
                lines[4 * nlines + 3] = line_end[1][1];
            }
            nlines++;
        }
    }
    free(nz); free(mask); free(accum); free(g.ttab);
    return nlines;
}

/* ------------------------------------------------------------------------- */
/* H3a: segment-angle mode of omr::get_result_from_edges_detection
 * (omr.rs:257-301): f32 atan2 -> degrees -> `% 45.0` (fmodf), widened to f64;
 * O(n^2) count within 0.1; strict-max keeps first, equal count appends.
 * status: 0 Believed, 1 NeedCheck, 2 NotAResult.  Returns -1 when n == 0 (the
 * reference panics on angles[0], omr.rs:272). */
static float orc_seg_angle_f32(const int32_t *l)
{
    /* the reference reads the segment as Vec4f (hough.rs:30, omr.rs:244) */
    float x1 = (float)l[0], y1 = (float)l[1], x2 = (float)l[2], y2 = (float)l[3];
    float angle = atan2f(y2 - y1, x2 - x1) * 180.0f / 3.14159265358979323846f;
    return fmodf(angle, 45.0f);
}

ORC_API int orc_lines_angle_omr(const int32_t *lines, int n, double *angle, int *status,
                                double *candidates, int cap, int *n_cand)
{
    if (n <= 0) return -1;
    double *ang = (double *)malloc(sizeof(double) * (size_t)n);
    for (int i = 0; i < n; i++) ang[i] = (double)orc_seg_angle_f32(lines + 4 * i);
    double target = ang[0];
    int best = 0, nc = 0;
    for (int i = 0; i < n; i++) {
        int count = 0;
        for (int j = 0; j < n; j++)
            if (fabs(ang[i] - ang[j]) < 0.1) count++;
        if (count > best) {
            target = ang[i];
            best = count;
            nc = 0;
            if (nc < cap) candidates[nc] = target;
            nc = 1;
        } else if (count == best) {
            if (nc < cap) candidates[nc] = ang[i];
            nc++;
        }
    }
    *angle = target;
    *n_cand = nc;
    *status = nc == 0 ? 2 : nc == 1 ? 0 : 1;
    free(ang);
    return 0;
}

/* H3b: hough::get_angle_with_hough's mode (hough.rs:48-89): all in f32,
 * `range` is the f32 literal 0.1, first strict max wins. */
ORC_API int orc_lines_angle_hough(const int32_t *lines, int n, double *angle)
{
    if (n <= 0) return -1;
    float *ang = (float *)malloc(sizeof(float) * (size_t)n);
    for (int i = 0; i < n; i++) ang[i] = orc_seg_angle_f32(lines + 4 * i);
    float target = ang[0];
    int best = 0;
    for (int i = 0; i < n; i++) {
        int count = 0;
        for (int j = 0; j < n; j++)
            if (fabsf(ang[i] - ang[j]) < 0.1f) count++;
        if (count > best) { target = ang[i]; best = count; }
    }
    *angle = (double)target;
    free(ang);
    return 0;
}

/* F3: the vote loop of fft::get_angle_with_fft (fft.rs:195-246), including
 * the index bug at fft.rs:231 (the inner loop re-reads line i, so angle_j is
 * line i's un-normalised angle).  Default 0.0. */
ORC_API double orc_lines_angle_fft(const int32_t *lines, int n)
{
    double average_angle = 0.0;
    int max_votes = 0;
    for (int i = 0; i < n; i++) {
        const int32_t *l = lines + 4 * i;
        double x1 = l[0], y1 = l[1], x2 = l[2], y2 = l[3];
        double counted = (atan2(y2 - y1, x2 - x1) * 180.0) / M_PI;
        double angle = counted < -45.0 ? counted + 90.0 : counted > 45.0 ? counted - 90.0 : counted;
        int votes = 0;
        for (int j = 0; j < n; j++) {
            if (i == j) continue;
            double angle_j = (atan2(y2 - y1, x2 - x1) * 180.0) / M_PI; /* sic: line i */
            if (fabs(angle_j - angle) < 0.1) votes++;
        }
        if (votes > max_votes) { max_votes = votes; average_angle = angle; }
    }
    return average_angle;
}

/* R1: the reconcile rule of omr::correct_default (omr.rs:352-403).
 * p_status / p_angle / p_candidates from rule A; e_angle from H3a (only read
 * when p_status != Believed). */
ORC_API void orc_reconcile(int p_status, double p_angle, const double *p_cand, int n_cand,
                           double e_angle, double *angle, int *need_check)
{
    if (p_status == 0) { *angle = p_angle; *need_check = 0; return; }
    if (p_status == 1) {
        if (fabs(p_angle - e_angle) >= 0.1) { *angle = e_angle; *need_check = 1; }
        else { *angle = p_angle; *need_check = 0; }
        return;
    }
    if (n_cand <= 0) { *angle = e_angle; *need_check = 1; return; }
    /* Iterator::min_by returns the FIRST minimum */
    int bi = 0;
    double bd = fabs(p_cand[0] - e_angle);
    for (int i = 1; i < n_cand; i++) {
        double d = fabs(p_cand[i] - e_angle);
        if (d < bd) { bd = d; bi = i; }
    }
    if (fabs(p_cand[bi] - e_angle) < 0.05) { *angle = p_cand[bi]; *need_check = 0; }
    else { *angle = e_angle; *need_check = 1; }
}
