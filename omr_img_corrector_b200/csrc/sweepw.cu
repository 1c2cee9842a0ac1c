// Word-parallel projection sweep for near-unit-scale
// rotations -- the BASELINE configuration (+-15 degrees, scale 1).  Same contract as
// sweep.cu (OpenCV-exact nearest-neighbour sampling of omr.rs:165-173, zero counts of
// omr.rs:8-39), but one lane produces 32 output pixels at a time.
//
// Exact decomposition.  OpenCV samples output pixel (x, y) at
//     sx = (X0[y] + ad[x]) >> 10,   sy = (Y0[y] + bd[x]) >> 10        (AB_BITS = 10 tables).
// Take the 32 pixels x = 32 j + i of one output word and re-reference both sums to one pixel
// of the word:
//   columns, reference i = 31:   X' = X0[y] + ad[32j+31] - 31*1024
//       sx_i = (X' >> 10) + i + d_i,   d_i = (fx' + G_i) >> 10,   G_i = ad[32j+i] - ad[32j+31] + (31-i)*1024 >= 0
//       d_i is non-increasing in i and d_31 = 0, so the word reads exactly the 32 source
//       bits V = [X'>>10, X'>>10 + 31] and pixel i takes bit i + d_i of V, d_i in {0, 1, 2}.
//   rows, reference r = the end of the word where bd is smallest:   Y' = Y0[y] + bd[ref]
//       sy_i = (Y' >> 10) + rho_i,   rho_i = (fy' + H_i) >> 10,   H_i = bd[32j+i] - bd[ref] >= 0
//       rho_i in [0, KR-1] with KR = floor(31 |sin|) + 2 source rows per word.
// With G_i = 1024 g_i + fg_i and H_i = 1024 h_i + fh_i the per-(angle, j) constants are the
// masks Q_rho = {i : h_i = rho}, S1 = {i : g_i = 1} ("meta") and the carries
//     C(y)  = {i : fh_i + fy' >= 1024},   D0(y) = S1 xor {i : fg_i + fx' >= 1024}
// depend on (angle, j, y) only ("planes"); both are independent of the card, are built once
// per geometry and shared by every card of a batch.  Per output word the card-dependent
// work is then KR x (2 LDS, 1 funnel shift, 1 shift, 3 LOP3):
//     out = OR_rho  R_rho & sel(V_rho),   R_rho = (Q_rho & ~C) | (Q_(rho-1) & C),
//     sel(V) = D1 ? V << 2 : D0 ? V << 1 : V          (D1 = S1 & ~D0, only for |angle| > 14.6 deg)
//
// Layout: lane = output row, warp = one word column, CTA = 8 word columns x 512 rows for up
// to 8 CONSECUTIVE candidate angles whose source footprints (parallelograms of the 1-bit
// plane, 0.1 degree apart) share one staged shared-memory tile (odd pitch, zero outside the
// image = white border, so the hot loop has no bounds tests).  Row counts leave the lane as
// popc; column counts are accumulated as bit-sliced vertical counters (carry-save adders)
// and unpacked once per (angle, column) with warp bit-transposes.
#include <algorithm>
#include <cmath>

#include "sweep_common.h"

namespace oics {

// ---------------------------------------------------------------------------
// meta + classification: one warp per (angle, word column)
__global__ void __launch_bounds__(32) sweepw_meta_kernel(SweepTables T, int w, int wj, uint32_t *__restrict__ meta,
                                                         SweepWClass *__restrict__ cls)
{
    // 1-D grid, angle-major (the angle count may exceed the 65535 limit of grid.y)
    const int ang = blockIdx.x / wj, j = blockIdx.x - ang * wj, lane = threadIdx.x;
    const int x = 32 * j + lane;  // < wpad: the tables are filled beyond w
    const bool valid = x < w;
    const int a = T.ad[(size_t)ang * T.wpad + x], b = T.bd[(size_t)ang * T.wpad + x];
    const int a31 = __shfl_sync(0xffffffffu, a, 31);
    const int b0 = __shfl_sync(0xffffffffu, b, 0), b31 = __shfl_sync(0xffffffffu, b, 31);
    const int bref = min(b0, b31);
    const int G = a - a31 + (31 - lane) * 1024;
    const int H = b - bref;
    // a <= a31 keeps i + d_i <= 31 (the word never reads past its 32-bit window)
    const bool ok = G >= 0 && G < 2048 && a <= a31 && H >= 0 && (H >> 10) <= SWW_KRMAX - 2;
    int hmax = ok ? (H >> 10) : 0;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) hmax = max(hmax, __shfl_xor_sync(0xffffffffu, hmax, o));
    const bool all_ok = __all_sync(0xffffffffu, ok);
    uint32_t *m = meta + ((size_t)ang * wj + j) * SWW_META;
    // pixel i of the word lives at bit 31 - i
    for (int rho = 0; rho < SWW_KRMAX - 1; rho++) {
        const uint32_t q = __brev(__ballot_sync(0xffffffffu, valid && ok && (H >> 10) == rho));
        if (lane == 0) m[rho] = q;
    }
    const uint32_t s1 = __brev(__ballot_sync(0xffffffffu, ok && G >= 1024));
    if (lane == 0) {
        m[10] = 0;
        m[11] = s1;
        m[12] = 0;
        m[13] = (uint32_t)(a31 - 31 * 1024);
        m[14] = (uint32_t)bref;
        m[15] = 0;
        atomicMax(&cls[ang].rows, hmax + 2);
        if (!all_ok) atomicOr(&cls[ang].bad, 1);
        if (s1) atomicOr(&cls[ang].bad, 2);   // bit 1: some word of this angle has a two-column step (|angle| > 14.6 degrees)
    }
}

// planes: one CTA per (fast angle, word column).  Phase 1 tabulates the two masks for every
// possible fraction (1024 entries), phase 2 looks the rows up.
__global__ void __launch_bounds__(1024) sweepw_planes_kernel(SweepTables T, const uint32_t *__restrict__ meta,
                                                             const int *__restrict__ fast_angles, int h, int wj,
                                                             uint2 *__restrict__ planes)
{
    __shared__ int sfg[32], sfh[32];
    __shared__ uint32_t tabC[1024], tabD[1024];
    const int k = blockIdx.x / wj, j = blockIdx.x - k * wj, tid = threadIdx.x;
    const int ang = fast_angles[k];
    const uint32_t *m = meta + ((size_t)ang * wj + j) * SWW_META;
    const uint32_t s1 = m[11];
    const int aref = (int)m[13], bref = (int)m[14];
    if (tid < 32) {
        const int x = 32 * j + tid;
        const int a = T.ad[(size_t)ang * T.wpad + x], b = T.bd[(size_t)ang * T.wpad + x];
        sfg[tid] = (a - aref - tid * 1024) & 1023;   // G_i = a - a31 + (31 - i) * 1024
        sfh[tid] = (b - bref) & 1023;
    }
    __syncthreads();
    {
        const int f = tid;
        uint32_t c = 0, ca = 0;
#pragma unroll
        for (int i = 0; i < 32; i++) {
            c |= (uint32_t)(sfh[i] + f >= 1024) << (31 - i);
            ca |= (uint32_t)(sfg[i] + f >= 1024) << (31 - i);
        }
        tabC[f] = c;
        tabD[f] = ca ^ s1;
    }
    __syncthreads();
    const int2 *xy = T.xy + (size_t)ang * T.hpad;
    uint2 *o = planes + ((size_t)k * wj + j) * h;
    for (int y = tid; y < h; y += 1024) {
        const int2 p = xy[y];
        o[y] = make_uint2(tabC[(p.y + bref) & 1023], tabD[(p.x + aref) & 1023]);
    }
}

int32_t sweepw_classify(oics_ctx *ctx, const SweepTables &T, int n_angles, int w, int h, uint32_t *d_meta,
                        SweepWClass *d_cls)
{
    const int wj = cdiv(w, 32);
    OICS_CUDA_TRY(cudaMemsetAsync(d_cls, 0, (size_t)n_angles * sizeof(SweepWClass), ctx->stream));
    sweepw_meta_kernel<<<(unsigned)wj * (unsigned)n_angles, 32, 0, ctx->stream>>>(T, w, wj, d_meta, d_cls);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    return OICS_OK;
}

int32_t sweepw_build_planes(oics_ctx *ctx, const SweepTables &T, const uint32_t *d_meta, const int *d_fast_angles,
                            int n_fast, int w, int h, uint2 *d_planes)
{
    if (n_fast <= 0) return OICS_OK;
    const int wj = cdiv(w, 32);
    ProfScope ps(ctx, "sweepw_planes_kernel", (double)n_fast * wj * h * sizeof(uint2));
    sweepw_planes_kernel<<<(unsigned)wj * (unsigned)n_fast, 1024, 0, ctx->stream>>>(T, d_meta, d_fast_angles, h, wj, d_planes);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    return OICS_OK;
}

// Host-side footprint test (real-valued, with a margin over the 1/1024-pixel table rounding;
// the kernel re-checks with the exact integers and raises the sticky flag if this was wrong).
bool sweepw_group_fits(const double *inv, int first, int cnt, int kr, int w, int h)
{
    const int wj = cdiv(w, 32), tiles_x = cdiv(wj, SWW_TJ), tiles_y = cdiv(h, SWW_TH);
    for (int ty = 0; ty < tiles_y; ty++) {
        const double ya = ty * SWW_TH, yb = std::min(h, (ty + 1) * SWW_TH) - 1;
        for (int tx = 0; tx < tiles_x; tx++) {
            const int nj = std::min(SWW_TJ, wj - tx * SWW_TJ);
            const double xa = tx * SWW_TJ * 32, xb = xa + 32 * nj - 1;
            double cmin = 1e300, cmax = -1e300, rmin = 1e300, rmax = -1e300;
            for (int k = first; k < first + cnt; k++) {
                const double *m = inv + 6 * (size_t)k;
                for (int c = 0; c < 2; c++) {
                    const double y = c ? yb : ya;
                    cmin = std::min(cmin, m[0] * (xa + 31) + m[1] * y + m[2] - 31.0);
                    cmax = std::max(cmax, m[0] * xb + m[1] * y + m[2]);
                    for (int d = 0; d < 2; d++) {
                        const double sy = m[3] * (d ? xb : xa) + m[4] * y + m[5];
                        rmin = std::min(rmin, sy);
                        rmax = std::max(rmax, sy);
                    }
                }
            }
            const long long c0 = (long long)std::floor(cmin - 1.0), c1 = (long long)std::floor(cmax + 2.0);
            const long long r0 = (long long)std::floor(rmin - 1.0), r1 = (long long)std::floor(rmax + 2.0);
            const long long nW = (c1 >> 5) - (c0 >> 5) + 1, nR = r1 - r0 + kr;
            if (nW > SWW_P || nR > SWW_TROWS) return false;
        }
    }
    return true;
}

// ---------------------------------------------------------------------------
struct SweepWArgs {
    const uint32_t *bits;       // [card][h][wp], addressed with the absolute card index
    size_t bits_card_words;
    int wp, w, h, wj;
    SweepTables T;
    SweepWPlan P;
    SweepWGroups G;
    int *counts;                // [slot][h + w]
    int tiles_x, n_cards;
};

__device__ __forceinline__ uint32_t bsel(uint32_t m, uint32_t a, uint32_t b) { return (m & a) | (~m & b); }
__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) { return (a & b) | (a & c) | (b & c); }

// One output word.  NT2: the word may contain a two-column step (|angle| > 14.6 degrees).
template <int KR, bool NT2>
__device__ __forceinline__ uint32_t assemble_word(const uint32_t *__restrict__ src, int sh, uint32_t C, uint32_t D0,
                                                  uint32_t D1, const uint32_t (&Q)[SWW_KRMAX - 1])
{
    uint32_t out = 0;
#pragma unroll
    for (int rho = 0; rho < KR; rho++) {
        const uint32_t w0 = src[rho * SWW_P], w1 = src[rho * SWW_P + 1];
        const uint32_t V = __funnelshift_l(w1, w0, sh);
        uint32_t s = bsel(D0, V << 1, V);
        if (NT2) s = bsel(D1, V << 2, s);
        uint32_t R;
        if (rho == 0) R = Q[0] & ~C;
        else if (rho == KR - 1) R = Q[KR - 2] & C;
        else R = (Q[rho] & ~C) | (Q[rho - 1] & C);
        out |= s & R;
    }
    return out;
}

// 4 KB of {X0, Y0} rows -> shared memory, 16 bytes per thread, asynchronously
__device__ __forceinline__ void copy_rows_async(int2 *dst, const int2 *src, int tid)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst + 2 * tid);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(src + 2 * tid) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// 5 bit-sliced planes -> per-column counts.  The five 32x32 bit transposes run in lock step
// so the lane-dependent shift and mask of a butterfly stage are computed once.
__device__ __forceinline__ int unpack_columns(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t c4, int np,
                                              int lane)
{
    uint32_t c[5] = {c0, c1, c2, c3, c4};
#pragma unroll
    for (int j = 16; j >= 1; j >>= 1) {
        const uint32_t low = j == 16 ? 0x0000FFFFu : j == 8 ? 0x00FF00FFu : j == 4 ? 0x0F0F0F0Fu
                             : j == 2 ? 0x33333333u : 0x55555555u;
        const bool up = (lane & j) != 0;
        const uint32_t keep = up ? ~low : low;   // bits of v that stay
        const int rot = up ? 32 - j : j;         // partner bits move down (up) by j
#pragma unroll
        for (int k = 0; k < 5; k++) {
            if (k < np) {
                const uint32_t o = __shfl_xor_sync(0xffffffffu, c[k], j);
                c[k] = (c[k] & keep) | (__funnelshift_l(o, o, rot) & ~keep);
            }
        }
    }
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < 5; k++)
        if (k < np) cnt += __popc(c[k]) << k;
    return cnt;
}

// Four 32-row groups of one word column: lane = row.  `sxy`, `rowp`, `plp` already point at
// this lane's first row; cd[] holds the planes of these four words (prefetched one round
// ahead) and is refilled for the next round.
template <int KR, bool NT2, bool CHECK>
__device__ __forceinline__ void four_words(const uint32_t *__restrict__ tile, const int2 *__restrict__ sxy,
                                           uint8_t *__restrict__ rowp, const uint2 *__restrict__ plp, uint2 (&cd)[4],
                                           int arefp, int brefp, uint32_t S1, const uint32_t (&Q)[SWW_KRMAX - 1],
                                           int rows_total, int lane, uint32_t &c0, uint32_t &c1, uint32_t &c2, uint32_t &c3,
                                           uint32_t &c4)
{
    uint32_t o[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
        if (CHECK && 32 * u >= rows_total) {   // whole group past the tile (warp-uniform)
            o[u] = 0;
            continue;
        }
        const int2 p = sxy[32 * u];
        const uint2 m = cd[u];
        cd[u] = __ldg(plp + 128 + 32 * u);   // next round (the planes buffer is padded)
        const int xq = p.x + arefp, yq = p.y + brefp;
        const int cb = xq >> 10;
        const uint32_t *src = tile + (yq >> 10) * SWW_P + (cb >> 5);
        const int sh = cb;   // the funnel shift wraps its amount modulo 32: no mask needed
        // NT2 is a property of the whole angle (class bit): no per-word test of S1
        const uint32_t v0 = NT2 ? assemble_word<KR, true>(src, sh, m.x, m.y, S1 & ~m.y, Q)
                                : assemble_word<KR, false>(src, sh, m.x, m.y, 0u, Q);
        uint32_t v = v0;
        if (CHECK) {
            const bool valid = 32 * u + lane < rows_total;
            v = valid ? v : 0u;
            if (valid) rowp[32 * u] = (uint8_t)__popc(v);
        } else {
            rowp[32 * u] = (uint8_t)__popc(v);
        }
        o[u] = v;
    }
    // carry-save accumulation of four words into the vertical counters
    const uint32_t k0 = maj3(o[0], o[1], c0);
    c0 ^= o[0] ^ o[1];
    const uint32_t k1 = maj3(o[2], o[3], c0);
    c0 ^= o[2] ^ o[3];
    const uint32_t k2 = maj3(k0, k1, c1);
    c1 ^= k0 ^ k1;
    const uint32_t t = c2 & k2;
    c2 ^= k2;
    const uint32_t t2 = c3 & t;
    c3 ^= t;
    c4 ^= t2;
}

template <int KR, bool NT2>
__device__ __forceinline__ void sweepw_body(const SweepWArgs &A, const int4 gd, const int card_local)
{
    extern __shared__ uint32_t smem[];
    uint32_t *tile = smem;                                                           // [SWW_TROWS][SWW_P] + pad
    int2 *sXY = reinterpret_cast<int2 *>(tile + SWW_TROWS * SWW_P + 4);              // [2][SWW_TH]
    uint8_t *sRow = reinterpret_cast<uint8_t *>(sXY + 2 * SWW_TH);                   // [2][SWW_TJ][SWW_TH]
    int *sBox = reinterpret_cast<int *>(tile + SWW_TROWS * SWW_P);                   // the pad words

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_ang = gd.y;
    const int card = A.G.card0 + card_local;
    const int tx = blockIdx.x % A.tiles_x, ty = blockIdx.x / A.tiles_x;
    const int j0 = tx * SWW_TJ, x0 = j0 * 32, y0 = ty * SWW_TH;
    const int nj = min(SWW_TJ, A.wj - j0), th = min(SWW_TH, A.h - y0);
    const uint32_t *bits = A.bits + (size_t)card * A.bits_card_words;

    if (warp == 0) {
        // Source bounding box of the tile over the angles of the group: the fixed-point
        // coordinates are monotone in x and in y, so the extremes sit at the tile corners.
        // lane -> (angle = lane >> 2, corner = lane & 3)
        int cmin = 0x7fffffff, cmax = -0x7fffffff, rmin = 0x7fffffff, rmax = -0x7fffffff;
        if ((lane >> 2) < n_ang) {
            const int ang = gd.w + (lane >> 2);
            const int *ad_g = A.T.ad + (size_t)ang * A.T.wpad, *bd_g = A.T.bd + (size_t)ang * A.T.wpad;
            const int2 p = A.T.xy[(size_t)ang * A.T.hpad + ((lane & 1) ? y0 + th - 1 : y0)];
            const int xb = x0 + 32 * nj - 1;
            if (lane & 2) {
                cmax = (p.x + ad_g[xb]) >> 10;
                rmin = rmax = (p.y + bd_g[xb]) >> 10;
            } else {
                cmin = ((p.x + ad_g[x0 + 31]) >> 10) - 31;
                rmin = rmax = (p.y + bd_g[x0]) >> 10;
            }
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            cmin = min(cmin, __shfl_xor_sync(0xffffffffu, cmin, o));
            cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, o));
            rmin = min(rmin, __shfl_xor_sync(0xffffffffu, rmin, o));
            rmax = max(rmax, __shfl_xor_sync(0xffffffffu, rmax, o));
        }
        if (lane == 0) {
            const int colw0 = cmin >> 5;
            const int nW = (cmax >> 5) - colw0 + 1;
            const int nR = rmax - rmin + KR;
            sBox[0] = colw0;
            sBox[1] = rmin;
            sBox[2] = nW;
            sBox[3] = (nW > SWW_P || nR > SWW_TROWS) ? -1 : nR;
        }
    }
    __syncthreads();
    const int colw0 = sBox[0], row0 = sBox[1], nW = sBox[2], nR = sBox[3];
    if (nR < 0) {  // cannot happen for groups accepted by sweepw_group_fits; never sample out of bounds
        if (tid == 0) atomicExch(A.P.err, 2);
        return;
    }
    // stage the footprint, zero outside the image (white border = no black pixel):
    // half a warp per source row, lane = word column
    {
        const int c = lane & 15, gc = colw0 + c;
        const bool cin = c < nW && gc >= 0 && gc < A.wp;
        if (nW <= 16) {
            // asynchronous copies (LDGSTS), zero-filled outside the image: no register round trip, every row
            // of the footprint in flight at once; 32-bit offsets, one add per row
            int r = warp * 2 + (lane >> 4);
            int goff = (row0 + r) * A.wp + gc;                                    // word offset into this card's plane
            unsigned sdst = (unsigned)__cvta_generic_to_shared(tile + r * SWW_P + c);
            const int gstep = 16 * A.wp;
            if (c < nW) {
#pragma unroll 4
                for (; r < nR; r += 16, goff += gstep, sdst += 16 * SWW_P * 4) {
                    const bool ok = cin && (unsigned)(row0 + r) < (unsigned)A.h;
                    const uint32_t *src = bits + (ok ? goff : 0);   // one 32-bit select, one wide add
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(sdst), "l"(src), "r"(ok ? 4 : 0) : "memory");
                }
            }
        } else {
            for (int i = tid; i < nR * SWW_P; i += 256) {
                const int r = i / SWW_P, cc = i - r * SWW_P;
                const int gr = row0 + r, g2 = colw0 + cc;
                uint32_t v = 0;
                if (gr >= 0 && gr < A.h && g2 >= 0 && g2 < A.wp) v = __ldg(bits + (size_t)gr * A.wp + g2);
                tile[i] = v;
            }
        }
    }
    // per-row {X0, Y0} of the first angle: a 4 KB copy (the tables are padded to whole tiles,
    // rows past the image repeat its last row, so every lane computes in-bounds addresses)
    const int xoff = (colw0 * 32) << 10, yoff = row0 << 10;
    copy_rows_async(sXY, A.T.xy + (size_t)gd.w * A.T.hpad + y0, tid);
    cp_async_wait_all();
    __syncthreads();

    const int j = j0 + warp;
    const bool active = warp < nj;
    for (int ai = 0; ai < n_ang; ai++) {
        const int ang = gd.w + ai;   // the angles of a group are consecutive (and so are their places in the fast list)
        const int slot = card_local * A.G.na_chunk + (ang - A.G.a0);
        int *rows_out = A.counts + (size_t)slot * (A.h + A.w);
        int *cols_out = rows_out + A.h;
        uint8_t *rowbuf = sRow + (ai & 1) * (SWW_TJ * SWW_TH);
        if (ai + 1 < n_ang)  // next angle's rows, in flight during this angle's loop
            copy_rows_async(sXY + ((ai + 1) & 1) * SWW_TH, A.T.xy + (size_t)(ang + 1) * A.T.hpad + y0, tid);
        if (active) {
            const uint4 *m4 = reinterpret_cast<const uint4 *>(A.P.meta + ((size_t)ang * A.wj + j) * SWW_META);
            const uint4 ma = __ldg(m4), mb = KR > 5 ? __ldg(m4 + 1) : make_uint4(0, 0, 0, 0);
            const uint4 mc = __ldg(m4 + 2), md = __ldg(m4 + 3);
            const uint32_t Qall[SWW_KRMAX - 1] = {ma.x, ma.y, ma.z, ma.w, mb.x, mb.y, mb.z, mb.w, mc.x, mc.y};
            uint32_t Q[SWW_KRMAX - 1];
#pragma unroll
            for (int r = 0; r < SWW_KRMAX - 1; r++) Q[r] = r < KR - 1 ? Qall[r] : 0u;
            const uint32_t S1 = mc.w;
            const int arefp = (int)md.y - xoff, brefp = (int)md.z - yoff;
            const uint32_t kslot = (uint32_t)(gd.x + ai - A.P.kf0);
            // 32-bit element index: a chunk's planes stay below 2^32 elements (3 GB budget / 8 bytes)
            const uint2 *plp = A.P.planes + ((kslot * (uint32_t)A.wj + (uint32_t)j) * (uint32_t)A.h + (uint32_t)(y0 + lane));
            const int2 *sxy = sXY + (ai & 1) * SWW_TH + lane;
            uint8_t *rowp = rowbuf + warp * SWW_TH + lane;
            uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0;  // bit-sliced column counters (<= 16 rows per lane)
            uint2 cd[4];
#pragma unroll
            for (int u = 0; u < 4; u++) cd[u] = __ldg(plp + 32 * u);
            int done = 0;
            for (; done + 256 <= th; done += 256) {  // two rounds of four full 32-row groups per trip
                four_words<KR, NT2, false>(tile, sxy, rowp, plp, cd, arefp, brefp, S1, Q, 128, lane, c0, c1, c2, c3, c4);
                four_words<KR, NT2, false>(tile, sxy + 128, rowp + 128, plp + 128, cd, arefp, brefp, S1, Q, 128, lane, c0, c1, c2,
                                           c3, c4);
                sxy += 256; rowp += 256; plp += 256;
            }
            if (done + 128 <= th) {
                four_words<KR, NT2, false>(tile, sxy, rowp, plp, cd, arefp, brefp, S1, Q, 128, lane, c0, c1, c2, c3, c4);
                sxy += 128; rowp += 128; plp += 128;
                done += 128;
            }
            if (done < th)   // the tile's last, partial round: row 32 u + lane is valid iff < th - done
                four_words<KR, NT2, true>(tile, sxy, rowp, plp, cd, arefp, brefp, S1, Q, th - done, lane, c0, c1, c2, c3, c4);
            // unpack: lane L receives bit L of every lane's plane = pixel 31 - L of the word
            const uint32_t hi = c2 | c3 | c4;
            if (__any_sync(0xffffffffu, (c0 | c1 | hi) != 0)) {
                int cnt;
                if (!__any_sync(0xffffffffu, hi != 0)) cnt = unpack_columns(c0, c1, 0, 0, 0, 2, lane);
                else if (!__any_sync(0xffffffffu, (c3 | c4) != 0)) cnt = unpack_columns(c0, c1, c2, 0, 0, 3, lane);
                else cnt = unpack_columns(c0, c1, c2, c3, c4, 5, lane);
                const int x = j * 32 + 31 - lane;
                if (cnt && x < A.w) atomicAdd(cols_out + x, cnt);
            }
        }
        cp_async_wait_all();
        __syncthreads();
        // rows: thread t sums the 8 per-column bytes of rows 4t..4t+3 as packed 16-bit lanes
        if (tid < SWW_TH / 4 && 4 * tid < th) {
            const uint32_t *rb = reinterpret_cast<const uint32_t *>(rowbuf) + tid;
            // a byte holds at most 32, so four columns add up inside their bytes (<= 128): two packed sums,
            // widened to 16-bit lanes once
            static_assert(SWW_TJ == 8, "the row sums add the columns in two packed groups of four");
            uint32_t a = 0, b = 0;
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (k < nj) a += rb[k * (SWW_TH / 4)];
#pragma unroll
            for (int k = 4; k < 8; k++)
                if (k < nj) b += rb[k * (SWW_TH / 4)];
            const uint32_t lo = (a & 0x00FF00FFu) + (b & 0x00FF00FFu);
            const uint32_t hi = ((a >> 8) & 0x00FF00FFu) + ((b >> 8) & 0x00FF00FFu);
            int *ro = rows_out + y0 + 4 * tid;
            const int n = th - 4 * tid;   // rows of this quad inside the tile (bytes past it are stale)
            if (lo & 0xFFFFu) atomicAdd(ro, (int)(lo & 0xFFFFu));
            if (n > 1 && (hi & 0xFFFFu)) atomicAdd(ro + 1, (int)(hi & 0xFFFFu));
            if (n > 2 && (lo >> 16)) atomicAdd(ro + 2, (int)(lo >> 16));
            if (n > 3 && (hi >> 16)) atomicAdd(ro + 3, (int)(hi >> 16));
        }
    }
}

static constexpr size_t SWW_SMEM = ((size_t)SWW_TROWS * SWW_P + 4) * 4 + 2 * (size_t)SWW_TH * 8 + 2 * (size_t)SWW_TJ * SWW_TH;

// One launch covers every class: blockIdx.y -> (group, card), card fastest so that the CTAs
// in flight share the group's planes in L2; the group descriptor carries its class.
__global__ void __launch_bounds__(256, 4) sweepw_kernel(const SweepWArgs A)
{
    const int grp = blockIdx.y / A.n_cards, card_local = blockIdx.y - grp * A.n_cards;
    const int4 gd = A.G.groups[grp];
    // gd.z = source rows per word (+ 256: the angles of the group have two-column steps)
    switch (gd.z) {
    case 2: sweepw_body<2, false>(A, gd, card_local); break;
    case 3: sweepw_body<3, false>(A, gd, card_local); break;
    case 4: sweepw_body<4, false>(A, gd, card_local); break;
    case 5: sweepw_body<5, false>(A, gd, card_local); break;
    case 6: sweepw_body<6, false>(A, gd, card_local); break;
    case 7: sweepw_body<7, false>(A, gd, card_local); break;
    case 8: sweepw_body<8, false>(A, gd, card_local); break;
    case 9: sweepw_body<9, false>(A, gd, card_local); break;
    case 10: sweepw_body<10, false>(A, gd, card_local); break;
    case 11: sweepw_body<11, false>(A, gd, card_local); break;
    case 256 + 8: sweepw_body<8, true>(A, gd, card_local); break;
    case 256 + 9: sweepw_body<9, true>(A, gd, card_local); break;
    case 256 + 10: sweepw_body<10, true>(A, gd, card_local); break;
    case 256 + 11: sweepw_body<11, true>(A, gd, card_local); break;
    default:   // a two-column step with few source rows (not a plain rotation): the general body of its row class
        switch (gd.z & 255) {
        case 2: sweepw_body<2, true>(A, gd, card_local); break;
        case 3: sweepw_body<3, true>(A, gd, card_local); break;
        case 4: sweepw_body<4, true>(A, gd, card_local); break;
        case 5: sweepw_body<5, true>(A, gd, card_local); break;
        case 6: sweepw_body<6, true>(A, gd, card_local); break;
        default: sweepw_body<7, true>(A, gd, card_local); break;
        }
        break;
    }
}

int32_t sweepw_launch(oics_ctx *ctx, const uint32_t *bits, size_t bits_card_words, int wp, int w, int h,
                      const SweepTables &T, const SweepWPlan &P, const SweepWGroups &G, int n_cards, int *counts)
{
    if (G.n_groups <= 0 || n_cards <= 0) return OICS_OK;
    if (!ctx->attr_sweepw) {   // per context = per device (a process-wide flag would skip the opt-in on a second GPU)
        OICS_CUDA_TRY(cudaFuncSetAttribute(sweepw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SWW_SMEM));
        ctx->attr_sweepw = true;
    }
    SweepWArgs A;
    A.bits = bits;
    A.bits_card_words = bits_card_words;
    A.wp = wp; A.w = w; A.h = h; A.wj = cdiv(w, 32);
    A.T = T; A.P = P; A.G = G;
    A.counts = counts;
    A.tiles_x = cdiv(A.wj, SWW_TJ);
    A.n_cards = n_cards;
    dim3 grid(A.tiles_x * cdiv(h, SWW_TH), (unsigned)(n_cards * G.n_groups));
    ProfScope ps(ctx, "sweepw_kernel", (double)n_cards * G.n_angles * (double)w * h);   // 1 B per (angle, pixel) sample
    sweepw_kernel<<<grid, 256, SWW_SMEM, ctx->stream>>>(A);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    return OICS_OK;
}

}  // namespace oics
