// Shared between the two sweep kernels (sweep.cu: one sample per instruction group,
// any affine map; sweepw.cu: one 32-pixel word per lane, near-unit-scale rotations).
#pragma once
#include "common.h"

namespace oics {

// Per-angle OpenCV warpAffine tables (AB_BITS = 10), see sweep_tables_kernel.
struct SweepTables {
    const int *ad, *bd;   // [angle][wpad]
    const int2 *xy;       // [angle][hpad]  {X0, Y0} incl. the +512 rounding term
    int wpad, hpad;
};

// Which (card, angle) items one launch covers and where their counts go.
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
        uint32_t other = __shfl_xor_sync(0xffffffffu, v, j);
        if (lane & j) v = (v & ~lowmask) | ((other >> j) & lowmask);
        else v = (v & lowmask) | ((other << j) & ~lowmask);
    }
    return v;
}

// ---- word kernel (sweepw.cu) ----
static constexpr int SWW_KRMAX = 11;   // source rows one output word may draw from
static constexpr int SWW_TH = 512;     // output rows per CTA
static constexpr int SWW_TJ = 8;       // output word columns per CTA (one per warp)
static constexpr int SWW_P = 17;       // staged-tile pitch in words (odd: rows spread over banks)
static constexpr int SWW_TROWS = 608;  // staged-tile rows
static constexpr int SWW_META = 16;    // uint32 per (angle, word column): Q[11], S1, S2, bj, cj, pad

struct SweepWPlan {
    // device buffers for the fast angles of the current chunk, indexed by slot = kf[angle] - kf0
    const uint32_t *meta;   // [slot][wj][SWW_META]
    const uint32_t *pc;     // [slot][wj][h]  row-carry masks
    const uint32_t *pca;    // [slot][wj][h]  column-carry masks
    const int *kf;          // [n_angles] fast index of an angle (or -1)
    int kf0;
    int *err;               // sticky device flag
};

// host side (sweepw.cu)
// Smallest supported rho-loop length for the inverse matrix m, or 0 when the word kernel
// cannot take this angle (scale or shear too far from a small rotation).
int sweepw_class(const double m[6]);
int32_t sweepw_build_plan(oics_ctx *ctx, const SweepTables &T, const int *d_fast_angles, int n_fast, int w, int h,
                          uint32_t *d_meta, uint32_t *d_pc, uint32_t *d_pca, int *d_err);
int32_t sweepw_launch(oics_ctx *ctx, int kr, const uint32_t *bits, size_t bits_card_words, int wp, int w, int h,
                      const SweepTables &T, const SweepWPlan &P, const SweepItems &I, int n_cards, int *counts);

}  // namespace oics
