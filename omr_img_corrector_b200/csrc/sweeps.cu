// Small-card projection sweep -- the reference's SHIPPED operating point:
// the card is shrunk to <= 248 x 230 before the sweep (getLibParams.ts:29-34, lib.rs:92-97) and
// get_result_from_projection re-applies the resize scale inside the rotation (omr.rs:157-163),
// so every candidate is a general affine map of a 7 KB bit plane.  One CTA holds the whole card:
//
//   * the bit plane of the card (1 bit per pixel, bit 31 - x % 32 of word x / 32, odd pitch) is staged in
//     shared memory once and re-used for every candidate angle of the CTA;
//   * one WARP per candidate angle: it takes the bounding box of the output pixels whose source can fall
//     inside the card (everything else samples the white border: no black pixel, count 0), walks it in
//     32-column strips with OpenCV's exact fixed-point nearest-neighbour coordinates (the same AB_BITS = 10
//     tables as the other two sweep kernels), counts rows with ballots and columns in registers into its
//     own shared-memory count vectors -- no atomics, no count buffer in HBM;
//   * the same warp then reduces the two count vectors to the population std-devs in the reference's own
//     evaluation order (calculate.rs:2-23: sequential f64 sums), so the (card, angle) item leaves the
//     kernel as 16 bytes.  ONE launch covers every card of a batch and every angle of the sweep.
//
// Same results as sweep_kernel / sweepw_kernel bit for bit (tests/test_gpu_parity.py checks all three
// against the CPU restatement on the same inputs).
#include <algorithm>
#include <cmath>

#include "sweep_common.h"

namespace oics {

static constexpr int SWS_THREADS = 256;
static constexpr int SWS_WARPS = SWS_THREADS / 32;
static constexpr int SWS_CHUNK = 256;   // f64 staging chunk of the sequential std-dev sum (per warp)

struct SweepSArgs {
    const uint32_t *bits;        // [card][h][wp]
    size_t bits_card_words;
    int wp, P, w, h;
    SweepTables T;
    const int4 *boxes;           // [angle] {x0, y0, x1, y1} inclusive output box that can sample the card (x1 < x0: empty)
    int n_angles, angles_per_cta;
    double *v_sd, *h_sd;         // [card][n_angles]
    long long *sums;             // nullable [card][n_angles][4]
    int *counts_out;             // nullable: [n_angles][h + w] of card 0
    size_t item0;                // first (card 0, angle 0) item in v_sd / h_sd
    int warp_words;              // 32-bit words of per-warp scratch
};

// calculate.rs:2-23 over a count vector in shared memory: mean from the exact integer sum, then a strictly
// sequential f64 sum of squares (each square rounded once, like powf(2.0)), / n, sqrt.
__device__ __forceinline__ double sws_std_dev(const int *__restrict__ v, int n, double *stage, int lane, long long *sum_out,
                                              long long *sq_out)
{
    long long s = 0, q = 0;
    for (int i = lane; i < n; i += 32) {
        const long long c = v[i];
        s += c;
        q += c * c;
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    *sum_out = s;
    *sq_out = q;
    const double mean = __ddiv_rn((double)s, (double)n);
    double acc = 0.0;
    for (int base = 0; base < n; base += SWS_CHUNK) {
        const int m = min(SWS_CHUNK, n - base);
        for (int i = lane; i < m; i += 32) {
            const double d = __dsub_rn((double)v[base + i], mean);
            stage[i] = __dmul_rn(d, d);
        }
        __syncwarp();
        if (lane == 0) {
            int i = 0;
            if (base == 0) { acc = stage[0]; i = 1; }
            for (; i + 8 <= m; i += 8) {
                const double t0 = stage[i], t1 = stage[i + 1], t2 = stage[i + 2], t3 = stage[i + 3];
                const double t4 = stage[i + 4], t5 = stage[i + 5], t6 = stage[i + 6], t7 = stage[i + 7];
                acc = __dadd_rn(acc, t0); acc = __dadd_rn(acc, t1); acc = __dadd_rn(acc, t2); acc = __dadd_rn(acc, t3);
                acc = __dadd_rn(acc, t4); acc = __dadd_rn(acc, t5); acc = __dadd_rn(acc, t6); acc = __dadd_rn(acc, t7);
            }
            for (; i < m; i++) acc = __dadd_rn(acc, stage[i]);
        }
        __syncwarp();
    }
    return __dsqrt_rn(__ddiv_rn(acc, (double)n));
}

__global__ void __launch_bounds__(SWS_THREADS) sweeps_kernel(const SweepSArgs A)
{
    extern __shared__ uint32_t smem[];
    uint32_t *tile = smem;                                         // [h][P]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // per-warp scratch: counts [h + w] ints, then 8-byte aligned {X0, Y0}[h] aliased with the f64 staging chunk
    uint32_t *wbase = tile + ((A.h * A.P + 1) & ~1) + (size_t)warp * A.warp_words;
    int *rows = reinterpret_cast<int *>(wbase);
    int *cols = rows + A.h;
    int2 *sxy = reinterpret_cast<int2 *>(wbase + ((A.h + A.w + 1) & ~1));
    double *stage = reinterpret_cast<double *>(sxy);

    const int card = blockIdx.y;
    const uint32_t *bits = A.bits + (size_t)card * A.bits_card_words;
    for (int i = tid; i < A.h * A.wp; i += SWS_THREADS) {
        const int r = i / A.wp, c = i - r * A.wp;
        tile[r * A.P + c] = __ldg(bits + i);
    }
    __syncthreads();

    const int a_begin = blockIdx.x * A.angles_per_cta, a_end = min(A.n_angles, a_begin + A.angles_per_cta);
    for (int ang = a_begin + warp; ang < a_end; ang += SWS_WARPS) {
        for (int i = lane; i < A.h + A.w; i += 32) rows[i] = 0;
        const int4 box = A.boxes[ang];
        if (box.z >= box.x && box.w >= box.y) {
            const int2 *xy_g = A.T.xy + (size_t)ang * A.T.hpad;
            const int *ad_g = A.T.ad + (size_t)ang * A.T.wpad, *bd_g = A.T.bd + (size_t)ang * A.T.wpad;
            for (int y = box.y + lane; y <= box.w; y += 32) sxy[y] = __ldg(xy_g + y);
            __syncwarp();
            for (int xs = box.x; xs <= box.z; xs += 32) {         // one 32-column strip
                const int x = xs + lane;
                const bool xin = x <= box.z;
                const int adl = __ldg(ad_g + min(x, A.w - 1)), bdl = __ldg(bd_g + min(x, A.w - 1));
                int colcnt = 0;
                for (int y = box.y; y <= box.w; y++) {
                    const int2 p = sxy[y];                         // broadcast
                    const int sx = (p.x + adl) >> 10, sy = (p.y + bdl) >> 10;
                    const bool in = xin && (unsigned)sx < (unsigned)A.w && (unsigned)sy < (unsigned)A.h;
                    const uint32_t wv = in ? tile[sy * A.P + (sx >> 5)] : 0u;
                    const bool black = ((wv << (sx & 31)) >> 31) != 0u;
                    const uint32_t b = __ballot_sync(0xffffffffu, black);
                    colcnt += black;
                    if (lane == 0 && b) rows[y] += __popc(b);
                }
                if (xin) cols[x] = colcnt;
            }
        }
        __syncwarp();
        if (A.counts_out && card == 0) {
            int *o = A.counts_out + (size_t)ang * (A.h + A.w);
            for (int i = lane; i < A.h + A.w; i += 32) o[i] = rows[i];
        }
        long long s0, q0, s1, q1;
        const double hsd = sws_std_dev(rows, A.h, stage, lane, &s0, &q0);
        const double vsd = sws_std_dev(cols, A.w, stage, lane, &s1, &q1);
        if (lane == 0) {
            const size_t item = A.item0 + (size_t)card * A.n_angles + ang;
            A.h_sd[item] = hsd;
            A.v_sd[item] = vsd;
            if (A.sums) {
                long long *s = A.sums + 4 * item;
                s[0] = s0; s[1] = q0; s[2] = s1; s[3] = q1;
            }
        }
        __syncwarp();
    }
}

// Output pixels whose source may fall inside the card: forward-map the card's corners (with a margin that
// covers the 1/1024-pixel table rounding and the +-0.5 of nearest-neighbour rounding many times over) and
// clip to the image.  Everything outside samples the border.  inv: output -> source (OpenCV's inverse).
static void sample_box(const double *inv, int w, int h, int4 *box)
{
    // invert inv (2x2 + translation) back to source -> output
    const double a = inv[0], b = inv[1], c = inv[2], d = inv[3], e = inv[4], f = inv[5];
    const double det = a * e - b * d;
    if (!(std::fabs(det) > 1e-12)) {      // degenerate map: no restriction
        *box = make_int4(0, 0, w - 1, h - 1);
        return;
    }
    const double ia = e / det, ib = -b / det, id = -d / det, ie = a / det;
    // the footprint of one source pixel in output pixels bounds the rounding slack
    const double slack = 2.0 + 2.0 * (std::fabs(ia) + std::fabs(ib) + std::fabs(id) + std::fabs(ie));
    double x0 = 1e300, x1 = -1e300, y0 = 1e300, y1 = -1e300;
    for (int k = 0; k < 4; k++) {
        const double sx = (k & 1) ? (double)w : -1.0, sy = (k & 2) ? (double)h : -1.0;
        const double ox = ia * (sx - c) + ib * (sy - f), oy = id * (sx - c) + ie * (sy - f);
        x0 = std::min(x0, ox); x1 = std::max(x1, ox);
        y0 = std::min(y0, oy); y1 = std::max(y1, oy);
    }
    const double bx0 = std::floor(x0 - slack), bx1 = std::ceil(x1 + slack), by0 = std::floor(y0 - slack), by1 = std::ceil(y1 + slack);
    const int ix0 = (int)std::max(0.0, std::min((double)w, bx0)), ix1 = (int)std::max(-1.0, std::min((double)(w - 1), bx1));
    const int iy0 = (int)std::max(0.0, std::min((double)h, by0)), iy1 = (int)std::max(-1.0, std::min((double)(h - 1), by1));
    *box = make_int4(ix0, iy0, ix1, iy1);
}

static size_t sweeps_warp_words(int w, int h)
{
    const size_t counts = ((size_t)h + w + 1) & ~(size_t)1;
    const size_t xy = std::max((size_t)2 * h, (size_t)2 * SWS_CHUNK);
    return counts + xy;
}

size_t sweeps_smem_bytes(int w, int h)
{
    const int wp = cdiv(w, 32), P = wp | 1;
    return ((((size_t)h * P + 1) & ~(size_t)1) + SWS_WARPS * sweeps_warp_words(w, h)) * sizeof(uint32_t);
}

// Can (and should) the small-card kernel take this geometry?  The whole card plus eight warps' scratch must
// leave room for three CTAs per SM.
bool sweeps_eligible(int w, int h) { return sweeps_smem_bytes(w, h) <= 72 * 1024; }

int32_t sweeps_launch(oics_ctx *ctx, const uint32_t *bits, size_t bits_card_words, int w, int h, const SweepTables &T,
                      const double *inv_host, int n_angles, int4 *d_boxes, int n_cards, double *d_vsd, double *d_hsd,
                      long long *d_sums, int *d_counts_out, size_t item0, bool boxes_ready)
{
    cudaStream_t st = ctx->stream;
    if (!boxes_ready) {
        std::vector<int4> boxes((size_t)n_angles);
        for (int i = 0; i < n_angles; i++) sample_box(inv_host + 6 * (size_t)i, w, h, &boxes[(size_t)i]);
        OICS_CUDA_TRY(cudaMemcpyAsync(d_boxes, boxes.data(), boxes.size() * sizeof(int4), cudaMemcpyHostToDevice, st));
        OICS_CUDA_TRY(cudaStreamSynchronize(st));   // `boxes` is a temporary
    }
    const size_t smem = sweeps_smem_bytes(w, h);
    if (!ctx->attr_sweeps) {
        OICS_CUDA_TRY(cudaFuncSetAttribute(sweeps_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024));
        ctx->attr_sweeps = true;
    }
    SweepSArgs A;
    A.bits = bits;
    A.bits_card_words = bits_card_words;
    A.wp = cdiv(w, 32);
    A.P = A.wp | 1;
    A.w = w; A.h = h;
    A.T = T;
    A.boxes = d_boxes;
    A.n_angles = n_angles;
    A.v_sd = d_vsd; A.h_sd = d_hsd;
    A.sums = d_sums;
    A.counts_out = d_counts_out;
    A.item0 = item0;
    A.warp_words = (int)sweeps_warp_words(w, h);
    // enough CTAs to fill the device about four times over, at least one angle per warp
    const int want_ctas = std::max(1, (4 * 3 * ctx->sm_count) / std::max(1, n_cards));
    int per = std::max(SWS_WARPS, cdiv(n_angles, want_ctas));
    per = cdiv(per, SWS_WARPS) * SWS_WARPS;
    A.angles_per_cta = per;
    for (int c0 = 0; c0 < n_cards; c0 += 65535) {
        const int nc = std::min(65535, n_cards - c0);
        SweepSArgs B = A;
        B.bits = bits + (size_t)c0 * bits_card_words;
        B.item0 = item0 + (size_t)c0 * n_angles;
        if (c0) B.counts_out = nullptr;
        dim3 grid((unsigned)cdiv(n_angles, per), (unsigned)nc);
        ProfScope ps(ctx, "sweeps_kernel", (double)nc * n_angles * (double)w * h);
        sweeps_kernel<<<grid, SWS_THREADS, smem, st>>>(B);
        OICS_CUDA_TRY(cudaGetLastError());
        ctx->launches++;
        ctx->last_sweep_launches++;
    }
    return OICS_OK;
}

}  // namespace oics
