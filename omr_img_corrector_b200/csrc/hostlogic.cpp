// Host-side scalar logic of the deskew path: geometry in f64 exactly as OpenCV
// computes it, the two argmax rules, the segment-angle rules and the reconcile
// rule.  Compiled with -ffp-contract=off: none of these products may fuse.
#include <algorithm>
#include <cmath>
#include <vector>

#include "common.h"

namespace oics {

static const double kPi = 3.1415926535897932384626433832795;  // CV_PI

void invert_affine(const double Min[6], double M[6])
{
    // cv::warpAffine (no WARP_INVERSE_MAP) inverts in this exact order; M[0] is
    // overwritten before the translation terms use it.
    for (int i = 0; i < 6; i++) M[i] = Min[i];
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

void hough_geometry(int w, int h, double rho, double theta, int *numangle, int *numrho, std::vector<float> *ttab)
{
    // cv::HoughLinesProbabilistic: rho and theta are narrowed to f32 first.
    float rho_f = (float)rho, theta_f = (float)theta;
    float irho = 1 / rho_f;
    int na = (int)std::floor((kPi - 0.0) / (double)theta_f) + 1;
    if (na > 1 && std::fabs(kPi - (na - 1) * (double)theta_f) < (double)theta_f / 2) --na;
    *numangle = na;
    *numrho = (int)std::lrint(((w + h) * 2 + 1) / (double)rho_f);
    if (ttab) {
        ttab->resize(2 * (size_t)na);
        for (int n = 0; n < na; n++) {
            (*ttab)[2 * n] = (float)(std::cos((double)n * (double)theta_f) * (double)irho);
            (*ttab)[2 * n + 1] = (float)(std::sin((double)n * (double)theta_f) * (double)irho);
        }
    }
}

static void fill_result(oics_omr_result *r, double angle, int status, const std::vector<double> &cand)
{
    r->angle = angle;
    r->status = status;
    r->n_candidates = (int32_t)cand.size();
    if (r->candidates && r->cap > 0) {
        size_t n = std::min<size_t>(cand.size(), (size_t)r->cap);
        for (size_t i = 0; i < n; i++) r->candidates[i] = cand[i];
    }
}

// f32 segment angle folded with `% 45.0` (omr.rs:259-267, hough.rs:65-68).
static float segment_angle_f32(const int32_t *l)
{
    float x1 = (float)l[0], y1 = (float)l[1], x2 = (float)l[2], y2 = (float)l[3];
    float a = std::atan2(y2 - y1, x2 - x1) * 180.0f / 3.14159265358979323846f;
    return std::fmod(a, 45.0f);
}

}  // namespace oics

using namespace oics;

extern "C" {

OICS_API void oics_rotation_matrix_2d(float cx, float cy, double angle_deg, double scale, double M[6])
{
    // cv::getRotationMatrix2D: angle *= CV_PI/180 (the constant is folded first).
    double a = angle_deg * (kPi / 180.0);
    double alpha = scale * std::cos(a), beta = scale * std::sin(a);
    double x = (double)cx, y = (double)cy;
    M[0] = alpha;
    M[1] = beta;
    M[2] = (1 - alpha) * x - beta * y;
    M[3] = -beta;
    M[4] = alpha;
    M[5] = beta * x + (1 - alpha) * y;
}

OICS_API int32_t oics_angle_half_range(uint32_t max_angle, double step)
{
    // `(max_angle as f64 / step) as u16`: truncating, saturating, NaN -> 0
    double q = (double)max_angle / step;
    if (!(q > 0)) return 0;
    if (q >= 65535.0) return 65535;
    return (int32_t)q;
}

OICS_API int32_t oics_sweep_matrices(int32_t w, int32_t h, uint32_t max_angle, double step, double scale,
                                     double *M, int32_t cap_angles, int32_t *n_angles)
{
    OICS_ARG(w > 0 && h > 0 && n_angles);
    int N = oics_angle_half_range(max_angle, step);
    *n_angles = 2 * N;
    if (2 * N > cap_angles) {
        set_error("sweep needs room for %d matrices, have %d", 2 * N, cap_angles);
        return OICS_ERR_CAPACITY;
    }
    OICS_ARG(M || N == 0);
    float cx = (float)w / 2.0f, cy = (float)h / 2.0f;  // Point2f::new((w as f32)/2.0, (h as f32)/2.0)
    for (int i = 0; i < 2 * N; i++) {
        int deg = i - N;
        oics_rotation_matrix_2d(cx, cy, (double)deg * step, scale, M + 6 * (size_t)i);
    }
    return OICS_OK;
}

OICS_API int32_t oics_shard_range(int32_t n_angles, int32_t rank, int32_t world, int32_t *first, int32_t *count)
{
    OICS_ARG(n_angles >= 0 && world > 0 && rank >= 0 && rank < world && first && count);
    const int per = (n_angles + world - 1) / world;
    const int f = std::min(n_angles, rank * per);
    *first = f;
    *count = std::min(per, n_angles - f);
    return OICS_OK;
}

OICS_API int32_t oics_shard_layout(int32_t n_angles, int32_t world, int32_t *per_rank)
{
    OICS_ARG(n_angles >= 0 && world > 0 && per_rank);
    *per_rank = (n_angles + world - 1) / world;
    return OICS_OK;
}

OICS_API int32_t oics_shard_unpack(const double *sd, int32_t n_angles, int32_t world, double *v_sd, double *h_sd)
{
    OICS_ARG(n_angles >= 0 && world > 0 && (n_angles == 0 || (sd && v_sd && h_sd)));
    const int per = (n_angles + world - 1) / world;
    for (int r = 0; r < world; r++) {
        const int f = std::min(n_angles, r * per), c = std::min(per, n_angles - f);
        const double *blk = sd + (size_t)r * 2 * per;
        for (int i = 0; i < c; i++) {
            v_sd[f + i] = blk[i];
            h_sd[f + i] = blk[per + i];
        }
    }
    return OICS_OK;
}

OICS_API int32_t oics_rotate_geometry(int32_t cols, int32_t rows, double angle, double scale, int32_t clip,
                                      double M[6], int32_t *dst_w, int32_t *dst_h)
{
    OICS_ARG(cols > 0 && rows > 0 && M && dst_w && dst_h);
    OICS_ARG(clip == OICS_CLIP_DEFAULT || clip == OICS_CLIP_CONTAIN);
    if (clip == OICS_CLIP_DEFAULT) {
        oics_rotation_matrix_2d((float)cols / 2.0f, (float)rows / 2.0f, angle, scale, M);
        *dst_w = cols;
        *dst_h = rows;
        return OICS_OK;
    }
    // the Rust glue evaluates `angle * CV_PI / 180.0` left to right (transfer.rs:489-494)
    double r = angle * kPi / 180.0;
    double rw = std::ceil((double)rows * std::fabs(std::sin(r)) + (double)cols * std::fabs(std::cos(r)));
    double rh = std::ceil((double)cols * std::fabs(std::sin(r)) + (double)rows * std::fabs(std::cos(r)));
    oics_rotation_matrix_2d((float)std::ceil(rw / 2.0), (float)std::ceil(rh / 2.0), angle, scale, M);
    M[2] += std::ceil((rw - (double)cols) / 2.0);
    M[5] += std::ceil((rh - (double)rows) / 2.0);
    *dst_w = (int32_t)rw;
    *dst_h = (int32_t)rh;
    return OICS_OK;
}

OICS_API double oics_projection_scale(int32_t w, int32_t h, int32_t max_w, int32_t max_h)
{
    double ws = max_w <= 0 ? 1.0 : (double)max_w / (double)w;
    double hs = max_h <= 0 ? 1.0 : (double)max_h / (double)h;
    return ws < hs ? ws : hs;
}

OICS_API int32_t oics_rule_a(const double *h_sd, const double *v_sd, int32_t N, double step,
                             oics_omr_result *result)
{
    OICS_ARG(result && N >= 0 && (N == 0 || (h_sd && v_sd)));
    double best_h = 0.0, best_v = 0.0;
    uint32_t ties_h = 1, ties_v = 1;
    std::vector<double> most;
    for (int i = 0; i < 2 * N; i++) {
        double ang = (double)(i - N) * step;
        double hs = h_sd[i];
        if (best_h < hs) {
            best_h = hs;
            best_v = v_sd[i];
            ties_h = ties_v = 1;
            most.assign(1, ang);
        } else if (best_h == hs) {
            ties_h++;
            double vs = v_sd[i];
            if (best_v < vs) {
                ties_v = 1;
                best_v = vs;
                most.assign(1, ang);
            } else if (best_v == vs) {
                ties_v++;
                most.push_back(ang);
            }
        }
    }
    if (ties_h == 1 && ties_v == 1 && most.empty()) {
        // no candidate ever became the maximum (N == 0, i.e. max_angle / step < 1): the reference indexes
        // most_possible_deg_vec[0] and panics (omr.rs:210-211); an error, not a fabricated result
        set_error("empty angle range: the reference panics on most_possible_deg_vec[0] (omr.rs:211)");
        return OICS_ERR_INVALID_ARG;
    }
    if (ties_h == 1 && ties_v == 1)
        fill_result(result, most[0], OICS_BELIEVED, most);
    else if (most.size() == 1)
        fill_result(result, most[0], OICS_NEED_CHECK, most);
    else
        fill_result(result, 0.0, OICS_NOT_A_RESULT, most);
    return OICS_OK;
}

OICS_API int32_t oics_rule_a_batch(const double *h_sd, const double *v_sd, int32_t n_cards, int32_t N,
                                   double step, double *angles, int32_t *statuses, int32_t *n_candidates)
{
    OICS_ARG(n_cards >= 0 && N >= 0 && angles && statuses);
    std::vector<double> cand((size_t)std::max(1, 2 * N));
    for (int c = 0; c < n_cards; c++) {
        oics_omr_result r = {0, 0, 0, cand.data(), (int32_t)cand.size()};
        OICS_TRY(oics_rule_a(h_sd + (size_t)c * 2 * N, v_sd + (size_t)c * 2 * N, N, step, &r));
        angles[c] = r.angle;
        statuses[c] = r.status;
        if (n_candidates) n_candidates[c] = r.n_candidates;
    }
    return OICS_OK;
}

OICS_API int32_t oics_rule_b(const double *v_sd, const double *h_sd, int32_t N, double step, double *angle)
{
    OICS_ARG(angle && N > 0 && v_sd && h_sd);
    const int n = 2 * N;
    // (value, list) seeded with index 0, then every index visited -- index 0 twice.
    double vmax = v_sd[0], hmax = h_sd[0];
    std::vector<int> vl(1, 0), hl(1, 0);
    for (int i = 0; i < n; i++) {
        if (v_sd[i] > vmax) { vmax = v_sd[i]; vl.assign(1, i); }
        else if (v_sd[i] == vmax) vl.push_back(i);
        if (h_sd[i] > hmax) { hmax = h_sd[i]; hl.assign(1, i); }
        else if (h_sd[i] == hmax) hl.push_back(i);
    }
    int idx;
    if (vl.size() == 1 && hl.size() == 1 && vl[0] == hl[0]) {
        idx = vl[0];
    } else {
        // union of both candidate lists; the reference walks a HashMap (unspecified
        // order) -- ascending index here, which only matters on an exact v^2+h^2 tie.
        std::vector<int> cand(vl);
        cand.insert(cand.end(), hl.begin(), hl.end());
        std::sort(cand.begin(), cand.end());
        cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
        double sdp = 0.0;
        idx = -1;
        for (int c : cand) {
            double cur = v_sd[c] * v_sd[c] + h_sd[c] * h_sd[c];
            if (sdp < cur) { sdp = cur; idx = c; }
        }
        if (idx < 0) idx = n / 2;
    }
    *angle = ((double)idx - (double)N) * step;
    return OICS_OK;
}

OICS_API int32_t oics_lines_angle(const int32_t *lines, int32_t n, int32_t variant, oics_omr_result *result)
{
    OICS_ARG(result && n >= 0 && (n == 0 || lines));
    std::vector<double> cand;
    if (variant == OICS_LINES_FFT) {
        // fft.rs:195-246, including the index bug at fft.rs:231: the inner loop re-reads
        // line i, so votes = n-1 when line i's raw angle is within 0.1 of its folded angle.
        double best_angle = 0.0;
        int max_votes = 0;
        for (int i = 0; i < n; i++) {
            const int32_t *l = lines + 4 * (size_t)i;
            double raw = (std::atan2((double)l[3] - (double)l[1], (double)l[2] - (double)l[0]) * 180.0) / kPi;
            double folded = raw < -45.0 ? raw + 90.0 : raw > 45.0 ? raw - 90.0 : raw;
            int votes = 0;
            for (int j = 0; j < n; j++)
                if (j != i && std::fabs(raw - folded) < 0.1) votes++;
            if (votes > max_votes) { max_votes = votes; best_angle = folded; }
        }
        fill_result(result, best_angle, OICS_BELIEVED, cand);
        return OICS_OK;
    }
    if (n == 0) {
        set_error("no line segments (the reference indexes angles[0] here and panics)");
        return OICS_ERR_NO_LINES;
    }
    if (variant == OICS_LINES_HOUGH) {
        std::vector<float> ang((size_t)n);
        for (int i = 0; i < n; i++) ang[i] = segment_angle_f32(lines + 4 * (size_t)i);
        float target = ang[0];
        int best = 0;
        for (int i = 0; i < n; i++) {
            int count = 0;
            for (int j = 0; j < n; j++) count += std::fabs(ang[i] - ang[j]) < 0.1f;
            if (count > best) { best = count; target = ang[i]; }
        }
        fill_result(result, (double)target, OICS_BELIEVED, cand);
        return OICS_OK;
    }
    if (variant == OICS_LINES_OMR) {
        std::vector<double> ang((size_t)n);
        for (int i = 0; i < n; i++) ang[i] = (double)segment_angle_f32(lines + 4 * (size_t)i);
        double target = ang[0];
        int best = 0;
        for (int i = 0; i < n; i++) {
            int count = 0;
            for (int j = 0; j < n; j++) count += std::fabs(ang[i] - ang[j]) < 0.1;
            if (count > best) {
                best = count;
                target = ang[i];
                cand.assign(1, target);
            } else if (count == best) {
                cand.push_back(ang[i]);
            }
        }
        int status = cand.empty() ? OICS_NOT_A_RESULT : cand.size() == 1 ? OICS_BELIEVED : OICS_NEED_CHECK;
        fill_result(result, target, status, cand);
        return OICS_OK;
    }
    set_error("unknown lines-angle variant %d", variant);
    return OICS_ERR_INVALID_ARG;
}

OICS_API int32_t oics_reconcile(const oics_omr_result *p, double e_angle, double *angle, int32_t *need_check)
{
    OICS_ARG(p && angle && need_check);
    if (p->status == OICS_BELIEVED) {
        *angle = p->angle;
        *need_check = 0;
    } else if (p->status == OICS_NEED_CHECK) {
        bool far = std::fabs(p->angle - e_angle) >= 0.1;
        *angle = far ? e_angle : p->angle;
        *need_check = far ? 1 : 0;
    } else {
        int n = std::min(p->n_candidates, p->cap);
        if (n <= 0 || !p->candidates) {
            *angle = e_angle;
            *need_check = 1;
            return OICS_OK;
        }
        int bi = 0;  // Iterator::min_by keeps the first minimum
        for (int i = 1; i < n; i++)
            if (std::fabs(p->candidates[i] - e_angle) < std::fabs(p->candidates[bi] - e_angle)) bi = i;
        bool close = std::fabs(p->candidates[bi] - e_angle) < 0.05;
        *angle = close ? p->candidates[bi] : e_angle;
        *need_check = close ? 0 : 1;
    }
    return OICS_OK;
}

OICS_API int32_t oics_hough_dims(int32_t w, int32_t h, double rho, double theta, int32_t *numangle,
                                 int32_t *numrho)
{
    OICS_ARG(w > 0 && h > 0 && rho > 0 && theta > 0 && numangle && numrho);
    int na, nr;
    hough_geometry(w, h, rho, theta, &na, &nr, nullptr);
    *numangle = na;
    *numrho = nr;
    return OICS_OK;
}

}  // extern "C"
