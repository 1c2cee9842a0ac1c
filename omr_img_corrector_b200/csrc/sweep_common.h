// Shared between the two sweep kernels (sweep.cu: one sample per instruction group,
// any affine map; sweepw.cu: one 32-pixel word per lane, near-unit-scale rotations).
#pragma once
#include <vector>

#include "common.h"

namespace oics {

// Per-angle OpenCV warpAffine tables (AB_BITS = 10), see sweep_tables_kernel.
struct SweepTables {
    const int *ad, *bd;   // [angle][wpad]  (filled for every x < wpad, not only x < w)
    const int2 *xy;       // [angle][hpad]  {X0, Y0} incl. the +512 rounding term
    int wpad, hpad;
};

// Which (card, angle) items one launch of the general kernel covers and where their counts go.
//   blockIdx.y -> card_local = by / n_list, angle = alist[by % n_list]
//   counts slot = card_local * na_chunk + (angle - a0); card = card0 + card_local
struct SweepItems {
    const int *alist;
    int n_list;
    int a0, na_chunk;
    int card0;
};

// 32x32 bit-matrix transpose across a warp: out[lane L] bit c = in[lane c] bit L.
__device__ __forceinline__ uint32_t warp_bit_transpose(uint32_t v, int lane)
{
#pragma unroll
    for (int j = 16; j >= 1; j >>= 1) {
        const uint32_t lowmask = j == 16 ? 0x0000FFFFu : j == 8 ? 0x00FF00FFu : j == 4 ? 0x0F0F0F0Fu
                                 : j == 2 ? 0x33333333u : 0x55555555u;
        // one rotate + one select per stage: the bits a plain shift would drop are masked off anyway
        const uint32_t other = __shfl_xor_sync(0xffffffffu, v, j);
        const bool up = (lane & j) != 0;
        const uint32_t keep = up ? ~lowmask : lowmask;
        v = (v & keep) | (__funnelshift_l(other, other, up ? 32 - j : j) & ~keep);
    }
    return v;
}

// ---- word kernel (sweepw.cu) ----
static constexpr int SWW_KRMAX = 11;   // source rows one output word may draw from
static constexpr int SWW_TH = 512;     // output rows per CTA
static constexpr int SWW_TJ = 8;       // output word columns per CTA (one per warp)
static constexpr int SWW_P = 17;       // staged-tile pitch in words (odd: rows spread over banks)
static constexpr int SWW_TROWS = 596;  // staged-tile rows (4 CTAs of 55.6 KB per SM)
static constexpr int SWW_META = 16;    // uint32 per (angle, word column): Q[10], -, S1, -, Aref, Bref, -
static constexpr int SWW_GMAX = 8;     // candidate angles one CTA evaluates on one staged tile
// the word kernel prefetches the planes one round (128 rows) past the last row of a column
static constexpr size_t SWW_PLANES_PAD = 256 * sizeof(uint2);

// Per-angle verdict of the meta kernel: can the word kernel take this angle, and with how
// many source rows per output word.
struct SweepWClass {
    int rows;   // max over word columns of (source rows one word draws from); 0 = not computed
    int bad;    // bit 0: a column map outside the supported range (scale, shear, |angle| too large);
                // bit 1: some word has a two-column step (the kernel then needs the third shifted copy of the window)
};

struct SweepWPlan {
    const uint32_t *meta;   // [angle][wj][SWW_META]
    const uint2 *planes;    // [slot][wj][h]  {row-carry mask C, column-step mask D0}
    const int *kf;          // [n_angles] fast index of an angle (or -1); slot = kf[angle] - kf0
    int kf0;
    int *err;               // sticky device flag
};

// Runs of consecutive candidate angles of one class, each evaluated by one CTA per tile.
struct SweepWGroups {
    const int *alist;       // fast angle list (device)
    const int4 *groups;     // {first index into alist, count, class (source rows per word), -} (device)
    int n_groups;
    int n_angles;           // candidate angles these groups hold (profiling: algorithmic bytes)
    int a0, na_chunk, card0;
};

// host side (sweepw.cu)
int32_t sweepw_classify(oics_ctx *ctx, const SweepTables &T, int n_angles, int w, int h, uint32_t *d_meta,
                        SweepWClass *d_cls);
// Can `cnt` consecutive angles starting at inv[6*first] share one staged tile for every tile of the image?
bool sweepw_group_fits(const double *inv, int first, int cnt, int kr, int w, int h);
int32_t sweepw_build_planes(oics_ctx *ctx, const SweepTables &T, const uint32_t *d_meta, const int *d_fast_angles,
                            int n_fast, int w, int h, uint2 *d_planes);
// small-card kernel (sweeps.cu): the whole card in one CTA, one warp per candidate angle, std-devs fused
bool sweeps_eligible(int w, int h);
int32_t sweeps_launch(oics_ctx *ctx, const uint32_t *bits, size_t bits_card_words, int w, int h, const SweepTables &T,
                      const double *inv_host, int n_angles, int4 *d_boxes, int n_cards, double *d_vsd, double *d_hsd,
                      long long *d_sums, int *d_counts_out, size_t item0, bool boxes_ready);
int32_t sweepw_launch(oics_ctx *ctx, const uint32_t *bits, size_t bits_card_words, int wp, int w, int h,
                      const SweepTables &T, const SweepWPlan &P, const SweepWGroups &G, int n_cards, int *counts);

}  // namespace oics
