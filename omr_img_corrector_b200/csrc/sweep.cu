// Fused projection-std-dev angle sweep.
//
// Replaces, per candidate angle, the reference's warp_affine(flags = 0) into a
// fresh Mat (omr.rs:165-173 / projection.rs:48-57), the zero-count pass over that
// Mat (omr.rs:8-39 / transfer.rs:305-405) and the two std-devs (calculate.rs:13-23).
// No rotated image is ever written: every (card, angle, tile) CTA stages the
// bit-packed source footprint of its output tile in shared memory once, samples it
// with OpenCV's exact fixed-point nearest-neighbour coordinates, and reduces row
// and column zero-counts on chip.
//
// Data layout in HBM / L2
//   bits   : [card][h][wp] uint32, 1 bit per pixel, bit = 1 for a BLACK pixel
//            (byte == 0), pixel x of a row in word x/32 at bit 31-(x%32)
//            (bit-reversed so a left shift brings the wanted bit to the sign
//            position).  1.1 MB per A4 card: L2 resident.
//   tables : per angle, int32 adelta[w], bdelta[w] and int2 {X0,Y0}[h]: OpenCV's
//            AB_BITS = 10 fixed-point tables, computed once per sweep call in f64
//            with explicit round-to-nearest ops (no FMA), shared by every card.
//   counts : [item][h + w] int32 (rows, then columns), item = (card, angle);
//            accumulated with one coalesced RED per warp per tile.
#include <algorithm>
#include <cstdlib>

#include "sweep_common.h"

namespace oics {

static constexpr int SW_THREADS = 256;
static constexpr int SW_WARPS = SW_THREADS / 32;
static constexpr int SW_GROUP = 32;  // rows per warp pass

// ---------------------------------------------------------------------------
// OpenCV's warpAffine tables for every angle of the sweep.
__global__ void sweep_tables_kernel(const double *__restrict__ Minv, int w, int h, int wpad, int hpad,
                                    int *__restrict__ ad, int *__restrict__ bd, int2 *__restrict__ xy)
{
    // 1-D grid: angle-major (2N can exceed the 65535 limit of grid.y: N = (max_angle / step) as u16)
    const int nbx = (max(wpad, hpad) + blockDim.x - 1) / blockDim.x;
    const int a = blockIdx.x / nbx;
    const int i = (blockIdx.x - a * nbx) * blockDim.x + threadIdx.x;
    const double *M = Minv + 6 * (size_t)a;
    if (i < wpad) {
        // adelta[x] = saturate_cast<int>(M[0]*x*AB_SCALE): (M0*x) rounded, then *1024 (exact).
        // Filled up to wpad: the word kernel references the last (virtual) pixel of a word.
        double x = (double)i;
        ad[(size_t)a * wpad + i] = __double2int_rn(__dmul_rn(__dmul_rn(M[0], x), 1024.0));
        bd[(size_t)a * wpad + i] = __double2int_rn(__dmul_rn(__dmul_rn(M[3], x), 1024.0));
    }
    if (i < hpad) {
        // rows past the image (hpad is a whole number of word-kernel tiles) repeat the last row
        double y = (double)min(i, h - 1);
        int X0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(M[1], y), M[2]), 1024.0)) + 512;
        int Y0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(M[4], y), M[5]), 1024.0)) + 512;
        xy[(size_t)a * hpad + i] = make_int2(X0, Y0);
    }
}

// ---------------------------------------------------------------------------
// 8-bit card -> reversed bit plane of black pixels.  One thread per output word.
__global__ void pack_bits_kernel(const uint8_t *__restrict__ bins, size_t card_stride, size_t stride, int w,
                                 int h, int wp, uint32_t *__restrict__ bits, int n_cards)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;  // word in row
    const int y = blockIdx.y;
    const int c = blockIdx.z;
    if (j >= wp) return;
    const uint8_t *row = bins + (size_t)c * card_stride + (size_t)y * stride;
    const int x0 = j * 32;
    uint32_t word = 0;
    if (x0 + 32 <= w && ((reinterpret_cast<uintptr_t>(row + x0) & 15) == 0)) {
        const uint4 *p = reinterpret_cast<const uint4 *>(row + x0);
        uint4 v0 = __ldg(p), v1 = __ldg(p + 1);
        uint32_t v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            uint32_t t = __vcmpeq4(v[k], 0u) & 0x01010101u;  // 1 per zero byte
            uint32_t nib = (t * 0x08040201u) >> 24;          // byte 0 -> bit 3 ... byte 3 -> bit 0
            word |= (nib & 0xFu) << (28 - 4 * k);
        }
    } else {
        for (int k = 0; k < 32; k++) {
            int x = x0 + k;
            if (x < w && row[x] == 0) word |= 0x80000000u >> k;
        }
    }
    bits[((size_t)c * h + y) * wp + j] = word;
}

// ---------------------------------------------------------------------------
struct SweepArgs {
    const uint32_t *bits;
    size_t bits_card_words;
    int wp, w, h;
    const int *ad, *bd;
    const int2 *xy;
    int wpad, hpad;
    int *counts;  // [slot][h + w]
    SweepItems I; // which (card, angle) items this launch covers
    int tiles_x;
    int tile_h;   // rows per tile, multiple of SW_WARPS * SW_GROUP
    int tile_cap; // staged-tile capacity in words
};

// One sample: OpenCV's NN coordinates, one LDS, shift the wanted bit into the
// accumulator.  `tile` words are bit-reversed, so `word << (sx & 31)` puts pixel sx
// at bit 31 and the funnel shift appends it to acc.
template <bool CHECK>
__device__ __forceinline__ uint32_t sample_bit(const uint32_t *__restrict__ tile, int P, int X0, int Y0, int ad,
                                               int bd, int nrows, int nbits, uint32_t acc)
{
    const int a = X0 + ad;
    const int b = Y0 + bd;
    const int ry = b >> 10;
    uint32_t wv;
    if (CHECK) {
        const int sx = a >> 10;
        const bool in = (unsigned)ry < (unsigned)nrows && (unsigned)sx < (unsigned)nbits;
        wv = in ? tile[ry * P + (sx >> 5)] : 0u;
    } else {
        wv = tile[ry * P + (a >> 15)];
    }
    // rotate instead of shift: bit 31 of the result is pixel sx either way, and the funnel shift wraps its
    // amount modulo 32, so (a >> 10) needs no mask
    const uint32_t top = __funnelshift_l(wv, wv, a >> 10);
    return __funnelshift_l(top, acc, 1);  // (acc << 1) | (top >> 31)
}

// The sampling loop of one CTA: each warp takes 32-row groups of the tile; per group
// every lane shifts 32 row bits into one register per column, then the warp transposes
// those registers to turn them into row sums.
template <int CPL, bool CHECK>
__device__ __forceinline__ void sweep_rows(const uint32_t *__restrict__ src, int pitch, const int2 *sXY, int th,
                                           const int (&adl)[CPL], const int (&bdl)[CPL], const bool (&valid)[CPL],
                                           int nR, int nbits, int *rows_out, int (&colsum)[CPL], int lane, int warp)
{
    for (int g = warp * SW_GROUP; g < th; g += SW_WARPS * SW_GROUP) {
        const int nr = min(SW_GROUP, th - g);
        uint32_t cb[CPL];
#pragma unroll
        for (int k = 0; k < CPL; k++) cb[k] = 0;
        if (nr == SW_GROUP) {
#pragma unroll
            for (int r = 0; r < SW_GROUP; r++) {
                const int2 p = sXY[g + r];
#pragma unroll
                for (int k = 0; k < CPL; k++) cb[k] = sample_bit<CHECK>(src, pitch, p.x, p.y, adl[k], bdl[k], nR, nbits, cb[k]);
            }
        } else {
            for (int r = 0; r < SW_GROUP; r++) {
                if (r < nr) {
                    const int2 p = sXY[g + r];
#pragma unroll
                    for (int k = 0; k < CPL; k++)
                        cb[k] = sample_bit<CHECK>(src, pitch, p.x, p.y, adl[k], bdl[k], nR, nbits, cb[k]);
                } else {
#pragma unroll
                    for (int k = 0; k < CPL; k++) cb[k] <<= 1;
                }
            }
        }
        // row r of the group sits at bit 31-r of cb[k]
        int rowsum = 0;
#pragma unroll
        for (int k = 0; k < CPL; k++) {
            uint32_t v = valid[k] ? cb[k] : 0u;
            colsum[k] += __popc(v);
            rowsum += __popc(warp_bit_transpose(v, lane));
        }
        const int r = 31 - lane;
        if (r < nr && rowsum) atomicAdd(rows_out + g + r, rowsum);
    }
}

template <int CPL>
__global__ void __launch_bounds__(SW_THREADS) sweep_kernel(const SweepArgs A)
{
    extern __shared__ uint32_t smem[];
    const int TW = 32 * CPL;
    int2 *sXY = reinterpret_cast<int2 *>(smem);              // [tile_h]
    uint32_t *sCol = smem + 2 * A.tile_h;                    // [TW]
    uint32_t *tile = sCol + TW;                              // [tile_cap]
    __shared__ int sBox[6];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int card_local = blockIdx.y / A.I.n_list;
    const int ang = A.I.alist[blockIdx.y - card_local * A.I.n_list];
    const int slot = card_local * A.I.na_chunk + (ang - A.I.a0);
    const int card = A.I.card0 + card_local;
    const int tx = blockIdx.x % A.tiles_x, ty = blockIdx.x / A.tiles_x;
    const int x0 = tx * TW, y0 = ty * A.tile_h;
    const int tw = min(TW, A.w - x0), th = min(A.tile_h, A.h - y0);

    const int *ad_g = A.ad + (size_t)ang * A.wpad;
    const int *bd_g = A.bd + (size_t)ang * A.wpad;
    const int2 *xy_g = A.xy + (size_t)ang * A.hpad;
    const uint32_t *bits = A.bits + (size_t)card * A.bits_card_words;

    // Source footprint of the tile.  X0[y]+adelta[x] is monotone in x and in y
    // separately, so the extremes of (sx, sy) over the tile sit at its corners.
    if (tid == 0) {
        int ax0 = ad_g[x0], ax1 = ad_g[x0 + tw - 1], bx0 = bd_g[x0], bx1 = bd_g[x0 + tw - 1];
        int2 p0 = xy_g[y0], p1 = xy_g[y0 + th - 1];
        int sx00 = (p0.x + ax0) >> 10, sx01 = (p0.x + ax1) >> 10, sx10 = (p1.x + ax0) >> 10, sx11 = (p1.x + ax1) >> 10;
        int sy00 = (p0.y + bx0) >> 10, sy01 = (p0.y + bx1) >> 10, sy10 = (p1.y + bx0) >> 10, sy11 = (p1.y + bx1) >> 10;
        int sxmin = min(min(sx00, sx01), min(sx10, sx11)), sxmax = max(max(sx00, sx01), max(sx10, sx11));
        int symin = min(min(sy00, sy01), min(sy10, sy11)), symax = max(max(sy00, sy01), max(sy10, sy11));
        const bool interior = sxmin >= 0 && sxmax < A.w && symin >= 0 && symax < A.h;
        // clip to the image; what falls outside reads as border (white = bit 0)
        int cx0 = max(sxmin, 0), cx1 = min(sxmax, A.w - 1), cy0 = max(symin, 0), cy1 = min(symax, A.h - 1);
        int wx0 = cx0 >> 5, wx1 = cx1 >> 5;
        int nW = wx1 - wx0 + 1, nR = cy1 - cy0 + 1;
        int empty = (cx1 < cx0) || (cy1 < cy0);
        if (empty) { nW = 0; nR = 0; wx0 = 0; cy0 = 0; }
        sBox[0] = wx0;
        sBox[1] = cy0;
        sBox[2] = nW;
        sBox[3] = nR;
        sBox[4] = empty ? 2 : interior ? 0 : 1;
        sBox[5] = min(cx1, A.w - 1) - wx0 * 32 + 1;  // valid bits per staged row (CHECK path)
    }
    for (int i = tid; i < TW; i += SW_THREADS) sCol[i] = 0;
    __syncthreads();
    const int wx0 = sBox[0], cy0 = sBox[1], nW = sBox[2], nR = sBox[3], kind = sBox[4], nbits = sBox[5];
    if (kind == 2) return;  // the whole tile samples the border: no black pixels
    const int P = nW | 1;   // odd pitch: consecutive source rows land in different banks
    const bool staged = nR * P <= A.tile_cap;

    // per-row table slice, re-based to the staged window
    const int xoff = (wx0 * 32) << 10, yoff = cy0 << 10;
    for (int i = tid; i < th; i += SW_THREADS) {
        int2 p = xy_g[y0 + i];
        sXY[i] = make_int2(p.x - xoff, p.y - yoff);
    }
    if (staged) {
        // the footprint is clipped to the image, so every staged word exists: asynchronous 4-byte copies with
        // incremental addresses (a (half-)warp per source row); the integer divide per word of a flat loop was a
        // fifth of the kernel's instructions on config 5's large angles
        if (nW <= 32) {
            const int lanes = nW <= 16 ? 16 : 32, rows_per_pass = SW_THREADS / lanes;
            const int c = tid & (lanes - 1);
            int r = tid / lanes;
            if (c < nW) {
                const uint32_t *src = bits + (size_t)(cy0 + r) * A.wp + wx0 + c;
                unsigned sdst = (unsigned)__cvta_generic_to_shared(tile + r * P + c);
                const size_t gstep = (size_t)rows_per_pass * A.wp;
                const unsigned sstep = (unsigned)(rows_per_pass * P * 4);
#pragma unroll 4
                for (; r < nR; r += rows_per_pass, src += gstep, sdst += sstep)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(sdst), "l"(src) : "memory");
            }
            asm volatile("cp.async.wait_all;\n" ::: "memory");
        } else {
            for (int i = tid; i < nR * nW; i += SW_THREADS) {
                int r = i / nW, j = i - r * nW;
                tile[r * P + j] = __ldg(bits + (size_t)(cy0 + r) * A.wp + wx0 + j);
            }
        }
    }
    // else: footprint larger than the shared-memory budget (strong down-scaling inside
    // the rotation): the L2-resident bit plane is sampled directly.
    // this lane's columns
    int adl[CPL], bdl[CPL];
    bool valid[CPL];
#pragma unroll
    for (int k = 0; k < CPL; k++) {
        int x = x0 + lane + 32 * k;
        valid[k] = x < A.w;
        int xc = valid[k] ? x : A.w - 1;
        adl[k] = ad_g[xc];
        bdl[k] = bd_g[xc];
    }
    __syncthreads();

    int colsum[CPL];
#pragma unroll
    for (int k = 0; k < CPL; k++) colsum[k] = 0;
    int *rows_out = A.counts + (size_t)slot * (A.h + A.w);
    int *cols_out = rows_out + A.h;

    // Three instantiations so the hot one reads the tile with LDS (not generic LD):
    // staged + interior (no bounds tests), staged + edge, and un-staged (global).
    if (staged && kind == 0)
        sweep_rows<CPL, false>(tile, P, sXY, th, adl, bdl, valid, nR, nbits, rows_out + y0, colsum, lane, warp);
    else if (staged)
        sweep_rows<CPL, true>(tile, P, sXY, th, adl, bdl, valid, nR, nbits, rows_out + y0, colsum, lane, warp);
    else
        sweep_rows<CPL, true>(bits + (size_t)cy0 * A.wp + wx0, A.wp, sXY, th, adl, bdl, valid, nR, nbits, rows_out + y0,
                              colsum, lane, warp);
#pragma unroll
    for (int k = 0; k < CPL; k++)
        if (colsum[k]) atomicAdd(&sCol[lane + 32 * k], (uint32_t)colsum[k]);
    __syncthreads();
    for (int i = tid; i < tw; i += SW_THREADS) {
        uint32_t v = sCol[i];
        if (v) atomicAdd(cols_out + x0 + i, (int)v);
    }
}

// ---------------------------------------------------------------------------
// P4: population std-dev of the count vectors in the reference's own evaluation
// order (calculate.rs:2-23): the mean from a left-to-right f64 sum (exact for
// integers, any order), then a strictly sequential sum of (c - mean)^2 with each
// square rounded before the add (powf(2.0)), divided by n, square root.  One warp
// per vector: lanes stage a chunk, lane 0 runs the dependent f64 chain.
// The pass is bound by the LATENCY of the ordered sum (one dependent DADD per element on lane 0, ~26 cycles
// each with its shared-memory operand), so what matters is how many chains an SM holds: a 256-element staging
// chunk keeps a CTA at 4 KB of shared memory and 32 CTAs (64 chains) resident per SM.  (Measured and rejected,
// r02i/j: 32 chains per warp, every lane on its own -- FP64 issue then bounds it, slower; tabulating the
// squares per distinct count -- no gain, the squares were never the limiter.)
static constexpr int FIN_CHUNK = 256;

__device__ double seq_std_dev(int *__restrict__ v, int n, double *stage, int lane, long long *sum_out,
                              long long *sq_out, bool zero_after)
{
    long long s = 0, q = 0;
    for (int i = lane; i < n; i += 32) {
        long long c = v[i];
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
    for (int base = 0; base < n; base += FIN_CHUNK) {
        const int m = min(FIN_CHUNK, n - base);
        // every lane squares its share (each square rounded once, like powf(2.0)) ...
        for (int i = lane; i < m; i += 32) {
            const double d = __dsub_rn((double)v[base + i], mean);
            stage[i] = __dmul_rn(d, d);
            if (zero_after) v[base + i] = 0;
        }
        __syncwarp();
        // ... lane 0 runs the strictly sequential sum: one dependent DADD per element
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

__global__ void __launch_bounds__(64) sweep_finalize_kernel(int *__restrict__ counts, int w, int h,
                                                            long long item0, double *__restrict__ v_sd,
                                                            double *__restrict__ h_sd, long long *__restrict__ sums,
                                                            int zero_after)
{
    __shared__ double stage[2][FIN_CHUNK];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int *base = counts + (size_t)blockIdx.x * (h + w);
    const size_t item = (size_t)item0 + blockIdx.x;
    long long s, q;
    if (warp == 0) {
        double sd = seq_std_dev(base, h, stage[0], lane, &s, &q, zero_after != 0);
        if (lane == 0) {
            h_sd[item] = sd;
            if (sums) { sums[4 * item + 0] = s; sums[4 * item + 1] = q; }
        }
    } else {
        double sd = seq_std_dev(base + h, w, stage[1], lane, &s, &q, zero_after != 0);
        if (lane == 0) {
            v_sd[item] = sd;
            if (sums) { sums[4 * item + 2] = s; sums[4 * item + 3] = q; }
        }
    }
}

// (Measured and rejected, r02i: a finalize pass with 32 vectors per warp and every lane on its own chain --
// the FP64 lane-operations are the same, and they, not latency, bound the pass.)

// Plain zero counts of an image as is (transfer::get_projection_standard_deviations
// on an un-rotated Mat): identity sampling of the bit plane.
__global__ void plain_counts_kernel(const uint32_t *__restrict__ bits, int wp, int w, int h, int *__restrict__ counts)
{
    // one warp per row: popcounts for the row sum, bit tests for the columns
    const int lane = threadIdx.x & 31;
    const int y = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (y >= h) return;
    int rs = 0;
    for (int j = lane; j < wp; j += 32) {
        uint32_t v = bits[(size_t)y * wp + j];
        rs += __popc(v);
        while (v) {
            int b = __clz(v);
            v &= ~(0x80000000u >> b);
            atomicAdd(counts + h + j * 32 + b, 1);
        }
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) rs += __shfl_xor_sync(0xffffffffu, rs, o);
    if (lane == 0) counts[y] = rs;
}

// ---------------------------------------------------------------------------
static int pick_cpl(int w)
{
    int best = 4, best_waste = 1 << 30;
    for (int cpl = 4; cpl >= 2; cpl--) {
        int tw = 32 * cpl;
        int waste = cdiv(w, tw) * tw - w;
        if (waste < best_waste) { best_waste = waste; best = cpl; }
    }
    return best;
}

template <int CPL>
static cudaError_t launch_sweep(oics_ctx *ctx, const SweepArgs &A, dim3 grid, size_t smem, cudaStream_t st)
{
    if (!ctx->attr_sweep[CPL]) {   // per context = per device
        cudaError_t e = cudaFuncSetAttribute(sweep_kernel<CPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024));
        if (e != cudaSuccess) return e;
        ctx->attr_sweep[CPL] = true;
    }
    sweep_kernel<CPL><<<grid, SW_THREADS, smem, st>>>(A);
    return cudaGetLastError();
}

int32_t sweep_device(oics_ctx *ctx, const uint8_t *dbins, int n_cards, size_t card_stride, int w, int h,
                     size_t stride, const double *M_host, int n_angles, double *d_vsd, double *d_hsd,
                     int64_t *d_sums, int32_t *d_counts_out)
{
    cudaStream_t st = ctx->stream;
    if (n_cards <= 0 || n_angles <= 0) return OICS_OK;
    if (w > 32767 || h > 32767) {
        set_error("image larger than 32767 px per side is not supported (OpenCV saturates coordinates to int16)");
        return OICS_ERR_UNSUPPORTED;
    }
    // inverse matrices, host f64 in OpenCV's order
    std::vector<double> inv((size_t)n_angles * 6);
    for (int i = 0; i < n_angles; i++) invert_affine(M_host + 6 * (size_t)i, inv.data() + 6 * (size_t)i);
    const int wpad = (w + 31) & ~31, hpad = cdiv(h, SWW_TH) * SWW_TH;
    const int wp = cdiv(w, 32), wj = wp;
    const size_t tab_ints = (size_t)n_angles * (2 * (size_t)wpad + 2 * (size_t)hpad);
    SweepCache &cache = ctx->sweep_cache;
    const char *force = getenv("OICS_SWEEP_KERNEL");  // "general": only the general kernel; "word": no small-card kernel (A/B tests)
    const bool small = sweeps_eligible(w, h) && !(force && (force[0] == 'g' || force[0] == 'w'));
    const bool same_geometry = cache.w == w && cache.h == h && cache.small == small && cache.inv == inv && ctx->tabs.p &&
                               ctx->tabs.cap >= tab_ints * sizeof(int) && ctx->wlists.p;

    // ---- chunking: counts buffer (256 MB) and word-kernel planes (3 GB) bound a chunk
    const size_t per_item = (size_t)(h + w) * sizeof(int);
    long long chunk_items = (long long)((size_t)256 << 20) / (long long)per_item;
    if (chunk_items < 1) chunk_items = 1;
    const size_t plan_per_angle = (size_t)wj * h * sizeof(uint2);
    const long long plan_budget_angles = std::max<long long>(1, (long long)(((size_t)3 << 30) / plan_per_angle));
    int cards_per_chunk, angles_per_chunk;
    if (chunk_items >= n_angles && n_angles <= plan_budget_angles && n_angles <= 65535) {
        angles_per_chunk = n_angles;
        cards_per_chunk = (int)std::min<long long>(n_cards, chunk_items / n_angles);
        cards_per_chunk = std::max(1, std::min(cards_per_chunk, 65535 / n_angles));
    } else {
        cards_per_chunk = 1;
        angles_per_chunk = (int)std::max<long long>(1, std::min<long long>(std::min<long long>(n_angles, chunk_items),
                                                                       std::min<long long>(plan_budget_angles, 65535)));
    }

    if (!same_geometry) {
        cache.w = cache.h = 0;
        cache.plan_a0 = cache.plan_a1 = -1;
        OICS_TRY(ctx->mats.ensure(inv.size() * sizeof(double)));
        OICS_TRY(ctx->pin_small.ensure(inv.size() * sizeof(double)));
        // pin_small may still be in flight from a previous call on this stream
        OICS_CUDA_TRY(cudaStreamSynchronize(st));
        memcpy(ctx->pin_small.p, inv.data(), inv.size() * sizeof(double));
        OICS_CUDA_TRY(cudaMemcpyAsync(ctx->mats.p, ctx->pin_small.p, inv.size() * sizeof(double), cudaMemcpyHostToDevice, st));
        OICS_TRY(ctx->tabs.ensure(tab_ints * sizeof(int)));
    }
    int *ad = ctx->tabs.as<int>();
    int *bd = ad + (size_t)n_angles * wpad;
    int2 *xy = reinterpret_cast<int2 *>(bd + (size_t)n_angles * wpad);
    SweepTables T;
    T.ad = ad; T.bd = bd; T.xy = xy; T.wpad = wpad; T.hpad = hpad;
    if (!same_geometry) {
        const unsigned grid = (unsigned)cdiv(max(wpad, hpad), 256) * (unsigned)n_angles;
        sweep_tables_kernel<<<grid, 256, 0, st>>>(ctx->mats.as<double>(), w, h, wpad, hpad, ad, bd, xy);
        OICS_CUDA_TRY(cudaGetLastError());
        ctx->launches++;
        cache.small = small;
        cache.boxes_ready = false;
    }
    if (small) {
        // ---- small-card path: the whole card in one CTA, one warp per angle, std-devs fused (sweeps.cu).
        // wlists holds the per-angle sample boxes; no classification, no count buffer, no finalize pass.
        OICS_TRY(ctx->wlists.ensure((size_t)n_angles * sizeof(int4)));
        const size_t card_words = (size_t)h * wp;
        // bit planes for as many cards as 256 MB hold (a 248 x 230 card is 7 KB)
        const int cards_per_pack = (int)std::max<size_t>(1, std::min<size_t>((size_t)n_cards, ((size_t)256 << 20) / (card_words * 4)));
        OICS_TRY(ctx->bits.ensure((size_t)cards_per_pack * card_words * sizeof(uint32_t)));
        ctx->last_sweep_launches = 0;
        cudaEvent_t e0, e1;
        OICS_CUDA_TRY(cudaEventCreate(&e0));
        OICS_CUDA_TRY(cudaEventCreate(&e1));
        OICS_CUDA_TRY(cudaEventRecord(e0, st));
        for (int c0 = 0; c0 < n_cards; c0 += cards_per_pack) {
            const int nc = std::min(cards_per_pack, n_cards - c0);
            for (int z0 = 0; z0 < nc; z0 += 65535) {
                const int nz = std::min(65535, nc - z0);
                dim3 g(cdiv(wp, 128), h, nz);
                ProfScope ps(ctx, "pack_bits_kernel", (double)nz * ((double)w * h + (double)h * wp * 4));
                pack_bits_kernel<<<g, 128, 0, st>>>(dbins + (size_t)(c0 + z0) * card_stride, card_stride, stride, w, h, wp,
                                                    ctx->bits.as<uint32_t>() + (size_t)z0 * card_words, nz);
                OICS_CUDA_TRY(cudaGetLastError());
                ctx->launches++;
            }
            OICS_TRY(sweeps_launch(ctx, ctx->bits.as<uint32_t>(), card_words, w, h, T, inv.data(), n_angles, ctx->wlists.as<int4>(),
                                   nc, d_vsd, d_hsd, reinterpret_cast<long long *>(d_sums), c0 == 0 ? d_counts_out : nullptr,
                                   (size_t)c0 * n_angles, cache.boxes_ready));
            cache.boxes_ready = true;
        }
        OICS_CUDA_TRY(cudaEventRecord(e1, st));
        OICS_CUDA_TRY(cudaStreamSynchronize(st));
        float t = 0.f;
        cudaEventElapsedTime(&t, e0, e1);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        ctx->last_sweep_ms = t;
        cache.w = w;
        cache.h = h;
        cache.inv.swap(inv);
        return OICS_OK;
    }
    if (!same_geometry) {

        // ---- classify the angles on the device: word kernel (by source rows per word) or general
        const size_t meta_bytes = (size_t)n_angles * wj * SWW_META * sizeof(uint32_t);
        const size_t cls_off = (meta_bytes + 15) & ~(size_t)15;
        OICS_TRY(ctx->wmeta.ensure(cls_off + (size_t)n_angles * sizeof(SweepWClass)));
        SweepWClass *d_cls = reinterpret_cast<SweepWClass *>(ctx->wmeta.as<uint8_t>() + cls_off);
        OICS_TRY(sweepw_classify(ctx, T, n_angles, w, h, ctx->wmeta.as<uint32_t>(), d_cls));
        std::vector<SweepWClass> cls((size_t)n_angles);
        OICS_CUDA_TRY(cudaMemcpyAsync(cls.data(), d_cls, cls.size() * sizeof(SweepWClass), cudaMemcpyDeviceToHost, st));
        OICS_CUDA_TRY(cudaStreamSynchronize(st));
        const bool allow_word = !(force && force[0] == 'g');
        const char *genv = getenv("OICS_SWEEP_GROUP");    // max angles per CTA (tuning)
        const int gmax = genv ? std::max(1, std::min(SWW_GMAX, atoi(genv))) : SWW_GMAX;
        cache.general.clear();
        cache.fast.clear();
        cache.groups.clear();
        std::vector<int> kr_of;  // class of fast[k]
        for (int i = 0; i < n_angles; i++) {
            int kr = 0;
            // class = source rows per word, + 256 when the angle has two-column steps (its own kernel body)
            if (allow_word && !(cls[i].bad & 1) && cls[i].rows >= 2 && cls[i].rows <= SWW_KRMAX &&
                sweepw_group_fits(inv.data(), i, 1, cls[i].rows, w, h))
                kr = cls[i].rows | ((cls[i].bad & 2) ? 256 : 0);
            if (!kr) { cache.general.push_back(i); continue; }
            // extend the last group: consecutive angle, same class, same chunk, footprints share a tile
            bool extend = false;
            if (!cache.groups.empty()) {
                const SweepCache::Group &g = cache.groups.back();
                extend = g.count < gmax && g.kr == kr && g.first + g.count == i &&
                         g.first / angles_per_chunk == i / angles_per_chunk &&
                         sweepw_group_fits(inv.data(), g.first, g.count + 1, kr & 255, w, h);
            }
            if (extend) cache.groups.back().count++;
            else cache.groups.push_back({(int)cache.fast.size(), 1, kr, i});
            cache.fast.push_back(i);
        }
        // heaviest classes first inside a chunk: the tail of the launch is filled by the light ones
        std::stable_sort(cache.groups.begin(), cache.groups.end(), [&](const SweepCache::Group &a, const SweepCache::Group &b) {
            const int ca = a.first / angles_per_chunk, cb = b.first / angles_per_chunk;
            if (ca != cb) return ca < cb;
            return (a.kr & 255) > (b.kr & 255);
        });
        const int n_chunks = cdiv(n_angles, angles_per_chunk);
        cache.chunk_group_begin.assign((size_t)n_chunks + 1, (int)cache.groups.size());
        for (int g = (int)cache.groups.size() - 1; g >= 0; g--) cache.chunk_group_begin[cache.groups[g].first / angles_per_chunk] = g;
        for (int c = n_chunks - 1; c >= 0; c--)
            cache.chunk_group_begin[c] = std::min(cache.chunk_group_begin[c], cache.chunk_group_begin[c + 1]);
        // device copy: [err 4][kf n][fast][general][groups as int4]
        const int n_fast0 = (int)cache.fast.size();
        size_t off = 4;
        cache.kf_off = off; off += (size_t)n_angles;
        cache.fast_off = off; off += (size_t)n_fast0;
        cache.general_off = off; off += cache.general.size();
        off = (off + 3) & ~(size_t)3;  // int4 alignment
        cache.groups_off = off; off += 4 * cache.groups.size();
        std::vector<int> hostbuf(off, 0);
        for (int i = 0; i < n_angles; i++) hostbuf[cache.kf_off + i] = -1;
        for (int k = 0; k < n_fast0; k++) {
            hostbuf[cache.kf_off + cache.fast[k]] = k;
            hostbuf[cache.fast_off + k] = cache.fast[k];
        }
        std::copy(cache.general.begin(), cache.general.end(), hostbuf.begin() + cache.general_off);
        for (size_t g = 0; g < cache.groups.size(); g++) {
            hostbuf[cache.groups_off + 4 * g + 0] = cache.groups[g].start;
            hostbuf[cache.groups_off + 4 * g + 1] = cache.groups[g].count;
            hostbuf[cache.groups_off + 4 * g + 2] = cache.groups[g].kr;
            hostbuf[cache.groups_off + 4 * g + 3] = cache.groups[g].first;
        }
        OICS_TRY(ctx->wlists.ensure(off * sizeof(int)));
        OICS_CUDA_TRY(cudaMemcpyAsync(ctx->wlists.p, hostbuf.data(), off * sizeof(int), cudaMemcpyHostToDevice, st));
        OICS_CUDA_TRY(cudaStreamSynchronize(st));  // hostbuf is a temporary
        cache.angles_per_chunk = angles_per_chunk;
    }
    const std::vector<int> &fast = cache.fast, &general = cache.general;
    const int n_fast = (int)fast.size();
    int *d_err = ctx->wlists.as<int>();
    const int *d_kf = d_err + cache.kf_off, *d_fast = d_err + cache.fast_off;

    // ---- general kernel: tile shape and shared-memory budget over ITS angles
    const int CPL = pick_cpl(w);
    const int TW = 32 * CPL;
    int tile_h = SW_WARPS * SW_GROUP * 2;  // 512 rows
    if (h <= SW_WARPS * SW_GROUP) tile_h = SW_WARPS * SW_GROUP;
    size_t need_words = 0;
    for (int i : general) {
        const double *m = inv.data() + 6 * (size_t)i;
        double fw = fabs(m[0]) * TW + fabs(m[1]) * tile_h + 4.0;
        double fh = fabs(m[3]) * TW + fabs(m[4]) * tile_h + 4.0;
        if (fw > w + 64) fw = w + 64;
        if (fh > h + 2) fh = h + 2;
        size_t P = ((size_t)(fw / 32.0) + 2) | 1;
        size_t words = P * (size_t)(fh + 1);
        if (words > need_words) need_words = words;
    }
    const size_t fixed_words = 2 * (size_t)tile_h + TW;
    size_t cap_words = need_words;
    const size_t max_words = (size_t)(96 * 1024) / 4 - fixed_words;  // keep >= 2 CTAs per SM
    if (cap_words > max_words) cap_words = max_words;
    const size_t smem = (fixed_words + cap_words) * 4;
    const int tiles_x = cdiv(w, TW), tiles_y = cdiv(h, tile_h);

    {
        const void *before = ctx->counts.p;
        OICS_TRY(ctx->counts.ensure((size_t)cards_per_chunk * angles_per_chunk * per_item));
        if (ctx->counts.p != before) ctx->counts_clean = 0;
    }
    const int cards_per_pack = std::min(n_cards, std::max(cards_per_chunk, 8));
    OICS_TRY(ctx->bits.ensure((size_t)cards_per_pack * h * wp * sizeof(uint32_t)));
    if (n_fast) {
        const int max_plan = std::min<long long>(n_fast, angles_per_chunk);
        OICS_TRY(ctx->wplanes.ensure((size_t)max_plan * plan_per_angle + SWW_PLANES_PAD));
    }

    SweepArgs A;
    A.bits_card_words = (size_t)h * wp;
    A.wp = wp; A.w = w; A.h = h;
    A.ad = ad; A.bd = bd; A.xy = xy; A.wpad = wpad; A.hpad = hpad;
    A.counts = ctx->counts.as<int>();
    A.tiles_x = tiles_x;
    A.tile_h = tile_h;
    A.tile_cap = (int)cap_words;

    std::vector<cudaEvent_t> evs;
    ctx->last_sweep_launches = 0;
    int packed_first = -1, packed_n = 0;  // cards currently in ctx->bits
    for (int c0 = 0; c0 < n_cards; c0 += cards_per_chunk) {
        const int nc = std::min(cards_per_chunk, n_cards - c0);
        if (packed_first < 0 || c0 < packed_first || c0 + nc > packed_first + packed_n) {
            packed_first = c0;
            packed_n = std::min(cards_per_pack, n_cards - c0);
            dim3 g(cdiv(wp, 128), h, packed_n);
            ProfScope ps(ctx, "pack_bits_kernel", (double)packed_n * ((double)w * h + (double)h * wp * 4));
            pack_bits_kernel<<<g, 128, 0, st>>>(dbins + (size_t)packed_first * card_stride, card_stride, stride, w, h, wp,
                                                ctx->bits.as<uint32_t>(), packed_n);
            OICS_CUDA_TRY(cudaGetLastError());
            ctx->launches++;
        }
        // absolute card index addressing relative to the packed window
        const uint32_t *bits_base = ctx->bits.as<uint32_t>() - (size_t)packed_first * A.bits_card_words;
        for (int a0 = 0; a0 < n_angles; a0 += angles_per_chunk) {
            const int a1 = std::min(n_angles, a0 + angles_per_chunk), na = a1 - a0;
            // (measured and rejected: letting the finalize pass clear what it has read instead of this memset --
            // the latency-bound 64-thread CTAs take 2x as long with the extra stores, r02h; and running the
            // finalize pass of chunk k on a second high-priority stream under the sweep of chunk k + 1 -- its
            // 64-thread CTAs fill every thread slot of the idle SMs first, so nothing overlaps, r02p; nor on a LOWEST-priority
            // stream with its own half of the counts buffer: the sweep's last wave leaves no tail worth filling, 3,844
            // against 3,836 cards/s)
            OICS_CUDA_TRY(cudaMemsetAsync(ctx->counts.p, 0, (size_t)nc * na * per_item, st));
            const bool keep_counts = true;
            // word-kernel planes for the fast angles of this range
            const int f_lo = (int)(std::lower_bound(fast.begin(), fast.end(), a0) - fast.begin());
            const int f_hi = (int)(std::lower_bound(fast.begin(), fast.end(), a1) - fast.begin());
            if (f_hi > f_lo && !(cache.plan_a0 == a0 && cache.plan_a1 == a1)) {
                cache.plan_a0 = cache.plan_a1 = -1;
                OICS_TRY(sweepw_build_planes(ctx, T, ctx->wmeta.as<uint32_t>(), d_fast + f_lo, f_hi - f_lo, w, h,
                                             ctx->wplanes.as<uint2>()));
                cache.plan_a0 = a0;
                cache.plan_a1 = a1;
            }
            cudaEvent_t e0, e1;
            OICS_CUDA_TRY(cudaEventCreate(&e0));
            OICS_CUDA_TRY(cudaEventCreate(&e1));
            evs.push_back(e0);
            evs.push_back(e1);
            OICS_CUDA_TRY(cudaEventRecord(e0, st));
            {
                const int l_lo = (int)(std::lower_bound(general.begin(), general.end(), a0) - general.begin());
                const int l_hi = (int)(std::lower_bound(general.begin(), general.end(), a1) - general.begin());
                if (l_hi > l_lo) {
                    SweepItems I;
                    I.alist = d_err + cache.general_off + l_lo;
                    I.n_list = l_hi - l_lo;
                    I.a0 = a0;
                    I.na_chunk = na;
                    I.card0 = c0;
                    A.bits = bits_base;
                    A.I = I;
                    dim3 grid(tiles_x * tiles_y, (unsigned)(nc * I.n_list));
                    ProfScope ps(ctx, "sweep_kernel<general>", (double)nc * I.n_list * (double)w * h);
                    cudaError_t le = CPL == 2 ? launch_sweep<2>(ctx, A, grid, smem, st)
                                     : CPL == 3 ? launch_sweep<3>(ctx, A, grid, smem, st)
                                                : launch_sweep<4>(ctx, A, grid, smem, st);
                    OICS_CUDA_TRY(le);
                    ctx->launches++;
                    ctx->last_sweep_launches++;
                }
            }
            {
                const int ci = a0 / angles_per_chunk;
                const int g_lo = cache.chunk_group_begin[ci], g_hi = cache.chunk_group_begin[ci + 1];
                if (g_hi > g_lo) {
                    SweepWPlan P;
                    P.meta = ctx->wmeta.as<uint32_t>();
                    P.planes = ctx->wplanes.as<uint2>();
                    P.kf = d_kf;
                    P.kf0 = f_lo;
                    P.err = d_err;
                    SweepWGroups G;
                    G.alist = d_fast;
                    G.groups = reinterpret_cast<const int4 *>(d_err + cache.groups_off) + g_lo;
                    G.n_groups = g_hi - g_lo;
                    G.n_angles = f_hi - f_lo;
                    G.a0 = a0;
                    G.na_chunk = na;
                    G.card0 = c0;
                    OICS_TRY(sweepw_launch(ctx, bits_base, A.bits_card_words, wp, w, h, T, P, G, nc, ctx->counts.as<int>()));
                    ctx->last_sweep_launches++;
                }
            }
            OICS_CUDA_TRY(cudaEventRecord(e1, st));
            const size_t item0 = (size_t)c0 * n_angles + a0;
            {
                ProfScope ps(ctx, "sweep_finalize_kernel", (double)nc * na * per_item);
                sweep_finalize_kernel<<<(unsigned)(nc * na), 64, 0, st>>>(ctx->counts.as<int>(), w, h, (long long)item0, d_vsd,
                                                                         d_hsd, reinterpret_cast<long long *>(d_sums), keep_counts ? 0 : 1);
            }
            OICS_CUDA_TRY(cudaGetLastError());
            ctx->launches++;
            if (d_counts_out && c0 == 0)
                OICS_CUDA_TRY(cudaMemcpyAsync(d_counts_out + (size_t)a0 * (h + w), ctx->counts.p, (size_t)na * per_item,
                                              cudaMemcpyDeviceToDevice, st));
        }
    }
    int err_flag = 0;
    OICS_CUDA_TRY(cudaMemcpyAsync(&err_flag, d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
    OICS_CUDA_TRY(cudaStreamSynchronize(st));
    double ms = 0.0;
    for (size_t i = 0; i < evs.size(); i += 2) {
        float t = 0.f;
        cudaEventElapsedTime(&t, evs[i], evs[i + 1]);
        ms += t;
        cudaEventDestroy(evs[i]);
        cudaEventDestroy(evs[i + 1]);
    }
    ctx->last_sweep_ms = ms;
    if (err_flag) {
        cache.w = cache.h = 0;
        set_error("internal: the word sweep kernel rejected a tile its planner accepted (flag %d)", err_flag);
        return OICS_ERR_INVALID_ARG;
    }
    cache.w = w;
    cache.h = h;
    cache.inv.swap(inv);
    return OICS_OK;
}

int32_t plain_std_devs_device(oics_ctx *ctx, const uint8_t *dbin, int w, int h, size_t stride, double *d_vsd,
                              double *d_hsd)
{
    cudaStream_t st = ctx->stream;
    const int wp = cdiv(w, 32);
    OICS_TRY(ctx->bits.ensure((size_t)h * wp * sizeof(uint32_t)));
    OICS_TRY(ctx->counts.ensure((size_t)(h + w) * sizeof(int)));
    ctx->counts_clean = 0;
    dim3 g(cdiv(wp, 128), h, 1);
    pack_bits_kernel<<<g, 128, 0, st>>>(dbin, 0, stride, w, h, wp, ctx->bits.as<uint32_t>(), 1);
    OICS_CUDA_TRY(cudaGetLastError());
    OICS_CUDA_TRY(cudaMemsetAsync(ctx->counts.p, 0, (size_t)(h + w) * sizeof(int), st));
    plain_counts_kernel<<<cdiv(h, 8), 256, 0, st>>>(ctx->bits.as<uint32_t>(), wp, w, h, ctx->counts.as<int>());
    OICS_CUDA_TRY(cudaGetLastError());
    sweep_finalize_kernel<<<1, 64, 0, st>>>(ctx->counts.as<int>(), w, h, 0, d_vsd, d_hsd, nullptr, 0);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 3;
    return OICS_OK;
}

}  // namespace oics
