// Word-parallel projection sweep for near-unit-scale
// rotations -- the BASELINE configuration (+-15 degrees, scale 1).  Same contract as
// sweep.cu (OpenCV-exact nearest-neighbour sampling of omr.rs:165-173, zero counts of
// omr.rs:8-39), but one lane produces 32 output pixels at a time.
//
// Exact decomposition.  With the AB_BITS = 10 tables, for output pixel x = 32 j + i of row y
//     sx = (X0[y] + ad[x]) >> 10 = qx + qa_i + [fa_i + fx >= 1024]
//     sy = (Y0[y] + bd[x]) >> 10 = qy + qb_i + [fb_i + fy >= 1024]
// (q = value >> 10, f = value & 1023).  Inside one word the integer parts move little:
//     qb_i = bmin_j + r_i,  r_i in [0, KR-2]         (a word spans <= 32 |sin| + 1 source rows)
//     qa_i = tmin_j + i + t_i,  t_i in {0, 1, 2}     (<= 32 (1 - cos) + 1 duplicated columns)
// so the word is assembled from KR source-row windows
//     out = OR_rho  R_rho & pick(window(row qy + bmin_j + rho, column qx + tmin_j), t + CA)
//     R_rho = (Q_rho & ~C) | (Q_(rho-1) & C),   Q_rho = {i : r_i = rho}
// where Q, S1 = {t >= 1}, S2 = {t >= 2}, bmin, tmin depend on (angle, j) only ("meta") and the
// carry masks C(y, j), CA(y, j) depend on (angle, j, y) only ("planes") -- both are
// independent of the card, are built once per geometry and are shared by every card of a
// batch.  The per-card work per word is then KR x (2-3 LDS, 2-5 funnel shifts, 3-5 LOP3).
//
// Layout: lane = output row, warp = one word column, CTA = 8 word columns x 512 rows whose
// source footprint (a parallelogram of the 1-bit plane) is staged once in shared memory with
// an odd pitch, zero outside the image (= white border), so the hot loop has no bounds tests.
// Row counts leave the lane as popc; column counts are accumulated as bit-sliced vertical
// counters (carry-save adders) and unpacked once per column with 5 warp bit-transposes.
#include <algorithm>
#include <cmath>

#include "sweep_common.h"

namespace oics {

int sweepw_class(const double m[6])
{
    // t range <= 2 and r range <= KR-2 over any 32 consecutive x (tables are rint()s of a
    // linear function: each difference is within 1 of the real one)
    const double tr = 31.0 * std::fabs(1.0 - m[0]) + 0.01;
    if (!(m[0] > 0) || std::floor(tr) + 1.0 > 2.0) return 0;
    const int rmax = (int)std::floor(31.0 * std::fabs(m[3]) + 0.01) + 1;
    const int need = rmax + 2;
    if (need > SWW_KRMAX) return 0;
    // staged footprint of a 256 x 512 tile must fit SWW_P words x SWW_TROWS rows
    const double fw = 255.0 * std::fabs(m[0]) + 511.0 * std::fabs(m[1]);
    const double fh = 511.0 * std::fabs(m[4]) + 255.0 * std::fabs(m[3]);
    if ((fw + 5.0) / 32.0 + 4.0 > (double)SWW_P) return 0;
    if (fh + 4.0 + SWW_KRMAX > (double)SWW_TROWS) return 0;
    return need <= 3 ? 3 : need <= 5 ? 5 : need <= 8 ? 8 : 11;
}

// ---------------------------------------------------------------------------
// meta: one warp per (fast angle, word column)
__global__ void __launch_bounds__(32) sweepw_meta_kernel(SweepTables T, const int *__restrict__ fast_angles, int w,
                                                         int wj, uint32_t *__restrict__ meta, int *err)
{
    const int j = blockIdx.x, k = blockIdx.y, lane = threadIdx.x;
    const int ang = fast_angles[k];
    const int x = 32 * j + lane;
    const bool valid = x < w;
    const int xc = valid ? x : w - 1;
    const int a = T.ad[(size_t)ang * T.wpad + xc], b = T.bd[(size_t)ang * T.wpad + xc];
    int qa = (a >> 10) - lane, qb = b >> 10;
    int tmin = valid ? qa : 0x7fffffff, bmin = valid ? qb : 0x7fffffff;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        tmin = min(tmin, __shfl_xor_sync(0xffffffffu, tmin, o));
        bmin = min(bmin, __shfl_xor_sync(0xffffffffu, bmin, o));
    }
    const int t = qa - tmin, r = qb - bmin;
    if (valid && (t > 2 || r > SWW_KRMAX - 2)) atomicExch(err, 1);
    uint32_t *m = meta + ((size_t)k * wj + j) * SWW_META;
    // pixel i of the word lives at bit 31 - i
    for (int rho = 0; rho < SWW_KRMAX; rho++) {
        const uint32_t q = __brev(__ballot_sync(0xffffffffu, valid && r == rho));
        if (lane == 0) m[rho] = q;
    }
    const uint32_t s1 = __brev(__ballot_sync(0xffffffffu, valid && t >= 1));
    const uint32_t s2 = __brev(__ballot_sync(0xffffffffu, valid && t >= 2));
    if (lane == 0) {
        m[11] = s1;
        m[12] = s2;
        m[13] = (uint32_t)bmin;
        m[14] = (uint32_t)tmin;
        m[15] = 0;
    }
}

// planes: one CTA per (fast angle, word column).  Phase 1 tabulates the carry mask for every
// possible fraction (1024 entries), phase 2 looks the rows up.
__global__ void __launch_bounds__(1024) sweepw_planes_kernel(SweepTables T, const int *__restrict__ fast_angles, int w,
                                                             int h, int wj, uint32_t *__restrict__ pc,
                                                             uint32_t *__restrict__ pca)
{
    __shared__ int sfa[32], sfb[32];
    __shared__ uint32_t tabC[1024], tabA[1024];
    const int j = blockIdx.x, k = blockIdx.y, tid = threadIdx.x;
    const int ang = fast_angles[k];
    if (tid < 32) {
        const int x = 32 * j + tid;
        // beyond the image: fraction 0 never carries (and those bits are masked by Q = 0 anyway)
        sfa[tid] = x < w ? (T.ad[(size_t)ang * T.wpad + x] & 1023) : 0;
        sfb[tid] = x < w ? (T.bd[(size_t)ang * T.wpad + x] & 1023) : 0;
    }
    __syncthreads();
    {
        const int f = tid;
        uint32_t c = 0, ca = 0;
#pragma unroll
        for (int i = 0; i < 32; i++) {
            c |= (uint32_t)(sfb[i] + f >= 1024) << (31 - i);
            ca |= (uint32_t)(sfa[i] + f >= 1024) << (31 - i);
        }
        tabC[f] = c;
        tabA[f] = ca;
    }
    __syncthreads();
    const int2 *xy = T.xy + (size_t)ang * T.hpad;
    uint32_t *oc = pc + ((size_t)k * wj + j) * h, *oa = pca + ((size_t)k * wj + j) * h;
    for (int y = tid; y < h; y += 1024) {
        const int2 p = xy[y];
        oc[y] = tabC[p.y & 1023];
        oa[y] = tabA[p.x & 1023];
    }
}

int32_t sweepw_build_plan(oics_ctx *ctx, const SweepTables &T, const int *d_fast_angles, int n_fast, int w, int h,
                          uint32_t *d_meta, uint32_t *d_pc, uint32_t *d_pca, int *d_err)
{
    if (n_fast <= 0) return OICS_OK;
    const int wj = cdiv(w, 32);
    dim3 grid(wj, n_fast);
    sweepw_meta_kernel<<<grid, 32, 0, ctx->stream>>>(T, d_fast_angles, w, wj, d_meta, d_err);
    OICS_CUDA_TRY(cudaGetLastError());
    sweepw_planes_kernel<<<grid, 1024, 0, ctx->stream>>>(T, d_fast_angles, w, h, wj, d_pc, d_pca);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 2;
    return OICS_OK;
}


// ---------------------------------------------------------------------------
struct SweepWArgs {
    const uint32_t *bits;       // [card][h][wp], addressed with the absolute card index
    size_t bits_card_words;
    int wp, w, h, wj;
    SweepTables T;
    SweepWPlan P;
    SweepItems I;
    int *counts;                // [slot][h + w]
    int tiles_x;
};

__device__ __forceinline__ uint32_t bsel(uint32_t m, uint32_t a, uint32_t b) { return (m & a) | (~m & b); }
__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) { return (a & b) | (a & c) | (b & c); }

// One output word.  NT = number of duplicated-column steps the word can contain (t range).
template <int KR, int NT>
__device__ __forceinline__ uint32_t assemble_word(const uint32_t *__restrict__ src, int sh, uint32_t C, uint32_t CA,
                                                  const uint32_t (&Q)[SWW_KRMAX], uint32_t S1, uint32_t S2)
{
    uint32_t out = 0;
    uint32_t d0, d1;  // delta = t + CA as two bit planes
    if (NT == 0) { d0 = CA; d1 = 0; }
    else if (NT == 1) { d0 = S1 ^ CA; d1 = S1 & CA; }
    else { d0 = S1 ^ S2 ^ CA; d1 = maj3(S1, S2, CA); }
#pragma unroll
    for (int rho = 0; rho < KR; rho++) {
        const uint32_t qa = rho < KR - 1 ? Q[rho] : 0u, qb = rho > 0 ? Q[rho - 1] : 0u;
        const uint32_t R = (qa & ~C) | (qb & C);
        const uint32_t *p = src + rho * SWW_P;
        const uint32_t w0 = p[0], w1 = p[1];
        uint32_t s;
        if (NT == 0) {
            const uint32_t f0 = __funnelshift_l(w1, w0, sh);
            const uint32_t f1 = __funnelshift_lc(w1, w0, sh + 1);
            s = bsel(d0, f1, f0);
        } else {
            const uint32_t w2 = p[2];
            const uint32_t vh = __funnelshift_l(w1, w0, sh), vl = __funnelshift_l(w2, w1, sh);
            const uint32_t f1 = __funnelshift_l(vl, vh, 1), f2 = __funnelshift_l(vl, vh, 2);
            const uint32_t lo = bsel(d0, f1, vh);
            if (NT == 1) {
                s = bsel(d1, f2, lo);
            } else {
                const uint32_t f3 = __funnelshift_l(vl, vh, 3);
                s = bsel(d1, bsel(d0, f3, f2), lo);
            }
        }
        out |= s & R;
    }
    return out;
}

template <int KR>
__global__ void __launch_bounds__(256) sweepw_kernel(const SweepWArgs A)
{
    extern __shared__ uint32_t smem[];
    uint32_t *tile = smem;                                           // [SWW_TROWS][SWW_P]
    int2 *sXY = reinterpret_cast<int2 *>(tile + SWW_TROWS * SWW_P);  // [SWW_TH]
    uint8_t *sRow = reinterpret_cast<uint8_t *>(sXY + SWW_TH);       // [SWW_TJ][SWW_TH]
    __shared__ int sBox[4];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int card_local = blockIdx.y / A.I.n_list;
    const int ang = A.I.alist[blockIdx.y - card_local * A.I.n_list];
    const int slot = card_local * A.I.na_chunk + (ang - A.I.a0);
    const int card = A.I.card0 + card_local;
    const int tx = blockIdx.x % A.tiles_x, ty = blockIdx.x / A.tiles_x;
    const int j0 = tx * SWW_TJ, x0 = j0 * 32, y0 = ty * SWW_TH;
    const int tw = min(SWW_TJ * 32, A.w - x0), th = min(SWW_TH, A.h - y0);

    const int *ad_g = A.T.ad + (size_t)ang * A.T.wpad;
    const int *bd_g = A.T.bd + (size_t)ang * A.T.wpad;
    const int2 *xy_g = A.T.xy + (size_t)ang * A.T.hpad;
    const uint32_t *bits = A.bits + (size_t)card * A.bits_card_words;

    if (tid == 0) {
        // source bounding box of the tile: the fixed-point coordinates are monotone in x and y
        const int ax0 = ad_g[x0], ax1 = ad_g[x0 + tw - 1], bx0 = bd_g[x0], bx1 = bd_g[x0 + tw - 1];
        const int2 p0 = xy_g[y0], p1 = xy_g[y0 + th - 1];
        const int sxa = (p0.x + ax0) >> 10, sxb = (p0.x + ax1) >> 10, sxc = (p1.x + ax0) >> 10, sxd = (p1.x + ax1) >> 10;
        const int sya = (p0.y + bx0) >> 10, syb = (p0.y + bx1) >> 10, syc = (p1.y + bx0) >> 10, syd = (p1.y + bx1) >> 10;
        const int sxmin = min(min(sxa, sxb), min(sxc, sxd)), sxmax = max(max(sxa, sxb), max(sxc, sxd));
        const int symin = min(min(sya, syb), min(syc, syd)), symax = max(max(sya, syb), max(syc, syd));
        const int colw0 = (sxmin - 3) >> 5;                       // first staged word column
        const int nW = (sxmax >> 5) + 2 - colw0 + 1;              // windows read 3 words
        const int row0 = symin - 1;
        const int nR = symax + SWW_KRMAX - row0 + 1;
        sBox[0] = colw0;
        sBox[1] = row0;
        sBox[2] = nR;
        sBox[3] = (nW > SWW_P || nR > SWW_TROWS) ? 1 : 0;
    }
    for (int i = tid; i < SWW_TJ * SWW_TH; i += 256) sRow[i] = 0;
    __syncthreads();
    const int colw0 = sBox[0], row0 = sBox[1], nR = sBox[2];
    if (sBox[3]) {  // cannot happen for angles accepted by sweepw_class; never sample out of bounds
        if (tid == 0) atomicExch(A.P.err, 2);
        return;
    }
    // stage the footprint, zero outside the image (white border = no black pixel)
    for (int i = tid; i < nR * SWW_P; i += 256) {
        const int r = i / SWW_P, c = i - r * SWW_P;
        const int gr = row0 + r, gc = colw0 + c;
        uint32_t v = 0;
        if (gr >= 0 && gr < A.h && gc >= 0 && gc < A.wp) v = __ldg(bits + (size_t)gr * A.wp + gc);
        tile[i] = v;
    }
    for (int i = tid; i < th; i += 256) {
        const int2 p = xy_g[y0 + i];
        sXY[i] = make_int2((p.x >> 10) - colw0 * 32, (p.y >> 10) - row0);
    }
    __syncthreads();

    const int j = j0 + warp;
    int *rows_out = A.counts + (size_t)slot * (A.h + A.w);
    int *cols_out = rows_out + A.h;
    if (j < A.wj) {
        const int kslot = A.P.kf[ang] - A.P.kf0;
        const uint32_t *m = A.P.meta + ((size_t)kslot * A.wj + j) * SWW_META;
        uint32_t Q[SWW_KRMAX];
#pragma unroll
        for (int r = 0; r < SWW_KRMAX; r++) Q[r] = r < KR - 1 ? __ldg(m + r) : 0u;
        const uint32_t S1 = __ldg(m + 11), S2 = __ldg(m + 12);
        const int bj = (int)__ldg(m + 13), cj = (int)__ldg(m + 14);
        const uint32_t *pc = A.P.pc + ((size_t)kslot * A.wj + j) * A.h + y0;
        const uint32_t *pca = A.P.pca + ((size_t)kslot * A.wj + j) * A.h + y0;
        uint8_t *myRow = sRow + warp * SWW_TH;
        uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0;  // bit-sliced column counters (<= 16 rows per lane)
        const int ngroups = (th + 31) >> 5;
        for (int g4 = 0; g4 < ngroups; g4 += 4) {
            uint32_t o[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int r = (g4 + u) * 32 + lane;
                const bool valid = r < th;
                const int rr = valid ? r : th - 1;
                const int2 p = sXY[rr];
                const uint32_t C = __ldg(pc + rr), CA = __ldg(pca + rr);
                const int sc = p.x + cj;
                const uint32_t *src = tile + (p.y + bj) * SWW_P + (sc >> 5);
                const int sh = sc & 31;
                uint32_t v;
                if (S1 == 0) v = assemble_word<KR, 0>(src, sh, C, CA, Q, S1, S2);
                else if (S2 == 0) v = assemble_word<KR, 1>(src, sh, C, CA, Q, S1, S2);
                else v = assemble_word<KR, 2>(src, sh, C, CA, Q, S1, S2);
                v = valid ? v : 0u;
                if (valid) myRow[r] = (uint8_t)__popc(v);
                o[u] = v;
            }
            // carry-save accumulation of four words into the vertical counters
            const uint32_t k0 = maj3(o[0], o[1], c0);
            c0 ^= o[0] ^ o[1];
            const uint32_t k1 = maj3(o[2], o[3], c0);
            c0 ^= o[2] ^ o[3];
            const uint32_t k2 = maj3(k0, k1, c1);
            c1 ^= k0 ^ k1;
            uint32_t t = c2 & k2;
            c2 ^= k2;
            uint32_t t2 = c3 & t;
            c3 ^= t;
            c4 ^= t2;
        }
        // unpack: lane L receives bit L of every lane's plane = pixel 31 - L of the word
        int cnt = __popc(warp_bit_transpose(c0, lane)) + 2 * __popc(warp_bit_transpose(c1, lane)) +
                  4 * __popc(warp_bit_transpose(c2, lane)) + 8 * __popc(warp_bit_transpose(c3, lane)) +
                  16 * __popc(warp_bit_transpose(c4, lane));
        const int x = j * 32 + 31 - lane;
        if (cnt && x < A.w) atomicAdd(cols_out + x, cnt);
    }
    __syncthreads();
    for (int i = tid; i < th; i += 256) {
        int s = 0;
#pragma unroll
        for (int k = 0; k < SWW_TJ; k++) s += sRow[k * SWW_TH + i];
        if (s) atomicAdd(rows_out + y0 + i, s);
    }
}

static constexpr size_t SWW_SMEM = (size_t)SWW_TROWS * SWW_P * 4 + (size_t)SWW_TH * 8 + (size_t)SWW_TJ * SWW_TH;

template <int KR>
static cudaError_t launch_w(const SweepWArgs &A, dim3 grid, cudaStream_t st)
{
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(sweepw_kernel<KR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SWW_SMEM);
        if (e != cudaSuccess) return e;
        attr = true;
    }
    sweepw_kernel<KR><<<grid, 256, SWW_SMEM, st>>>(A);
    return cudaGetLastError();
}

int32_t sweepw_launch(oics_ctx *ctx, int kr, const uint32_t *bits, size_t bits_card_words, int wp, int w, int h,
                      const SweepTables &T, const SweepWPlan &P, const SweepItems &I, int n_cards, int *counts)
{
    if (I.n_list <= 0 || n_cards <= 0) return OICS_OK;
    SweepWArgs A;
    A.bits = bits;
    A.bits_card_words = bits_card_words;
    A.wp = wp; A.w = w; A.h = h; A.wj = cdiv(w, 32);
    A.T = T; A.P = P; A.I = I;
    A.counts = counts;
    A.tiles_x = cdiv(A.wj, SWW_TJ);
    dim3 grid(A.tiles_x * cdiv(h, SWW_TH), (unsigned)(n_cards * I.n_list));
    cudaError_t e = kr == 3 ? launch_w<3>(A, grid, ctx->stream) : kr == 5 ? launch_w<5>(A, grid, ctx->stream)
                    : kr == 8 ? launch_w<8>(A, grid, ctx->stream) : launch_w<11>(A, grid, ctx->stream);
    OICS_CUDA_TRY(e);
    ctx->launches++;
    return OICS_OK;
}

}  // namespace oics
