// H2: the Hough stage.
//
//  hough_accumulate_device   the standard (rho, theta) accumulator over ALL edge pixels
//      with HoughLinesP's vote formula r = rint_f32(x*cos_f32 + y*sin_f32) + (numrho-1)/2.
//      One CTA per (theta, slice of the edge list): the numrho-bin histogram of that theta
//      lives in shared memory (privatised atomics), then is flushed once to HBM.
//  hough_lines_p_device      cv::HoughLinesP = the progressive probabilistic transform the
//      reference calls (hough.rs:31-42, omr.rs:245-253, fft.rs:183-191).  The algorithm is
//      sequential by definition (RNG-ordered, each accepted segment erases pixels and votes
//      that later points depend on), so one CTA walks one image:
//        * the visiting order depends only on the NUMBER of edge pixels (cv::RNG seeded
//          with (uint64)-1, swap-remove), not on the image: the host expands it from N and
//          a gather kernel lays the points out in visiting order, so the CTA streams them
//          and tests 256 upcoming points per L2 round trip;
//        * the next four live points are processed speculatively at once: all vote in one L2
//          round trip (one thread per theta, atomics issued in visiting order, so each point
//          sees exactly the sequential counts), one warp walks each point read-only, and the
//          results commit in visiting order; a point that an earlier commit could have
//          influenced is rolled back (its votes subtracted) and retried;
//        * the arg-max keeps the first maximal theta, like the sequential `max_val < val`;
//        * the two line walks test 32 steps each per round trip; gap logic runs on ballots;
//        * a commit clears the step range and, for accepted segments, un-votes its live pixels.
//      Latency-bound per image (like the CPU), throughput comes from images in flight.
#include <algorithm>
#include <cmath>

#include "common.h"

namespace oics {

// ---------------------------------------------------------------------------
// ordered compaction of the edge pixels (row-major, as cv::HoughLinesP collects them)
__global__ void edge_row_count_kernel(const uint8_t *__restrict__ edges, int w, int h, size_t stride,
                                      int *__restrict__ rowcnt)
{
    const int lane = threadIdx.x & 31;
    const int y = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (y >= h) return;
    const uint8_t *row = edges + (size_t)y * stride;
    int c = 0;
    for (int x = lane; x < w; x += 32) c += row[x] != 0;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) rowcnt[y] = c;
}

// exclusive scan of rowcnt[h] in place -> row offsets; total to *total.  One CTA.
__global__ void __launch_bounds__(1024) edge_row_scan_kernel(int *__restrict__ rowcnt, int h, int *__restrict__ total)
{
    __shared__ int warp_sums[32];
    __shared__ int carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < h; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < h ? rowcnt[i] : 0;
        int s = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += t;
        }
        if (lane == 31) warp_sums[warp] = s;
        __syncthreads();
        if (warp == 0) {
            int ws = warp_sums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, ws, o);
                if (lane >= o) ws += t;
            }
            warp_sums[lane] = ws;
        }
        __syncthreads();
        const int before = carry + (warp ? warp_sums[warp - 1] : 0) + s - v;
        if (i < h) rowcnt[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}

// points packed x | y << 16; mask (nullable) = edges != 0, tightly packed
__global__ void edge_row_write_kernel(const uint8_t *__restrict__ edges, int w, int h, size_t stride,
                                      const int *__restrict__ rowoff, uint32_t *__restrict__ pts,
                                      uint8_t *__restrict__ mask)
{
    const int lane = threadIdx.x & 31;
    const int y = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (y >= h) return;
    const uint8_t *row = edges + (size_t)y * stride;
    int off = rowoff[y];
    for (int x0 = 0; x0 < w; x0 += 32) {
        const int x = x0 + lane;
        const bool e = x < w && row[x] != 0;
        if (mask && x < w) mask[(size_t)y * w + x] = e ? 1 : 0;
        const uint32_t b = __ballot_sync(0xffffffffu, e);
        if (e) pts[off + __popc(b & ((1u << lane) - 1u))] = (uint32_t)x | ((uint32_t)y << 16);
        off += __popc(b);
    }
}

__device__ __forceinline__ int vote_r(float x, float y, float c, float s, int half)
{
    return __float2int_rn(__fadd_rn(__fmul_rn(x, c), __fmul_rn(y, s))) + half;
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) hough_accumulate_kernel(const uint32_t *__restrict__ pts,
                                                               const int *__restrict__ npts_ptr,
                                                               const float *__restrict__ ttab, int numrho,
                                                               int slices, int use_smem, int32_t *__restrict__ accum)
{
    extern __shared__ int hist[];
    const int n = blockIdx.x, slice = blockIdx.y;
    const int npts = *npts_ptr;
    const float c = ttab[2 * n], s = ttab[2 * n + 1];
    const int half = (numrho - 1) / 2;
    int32_t *out = accum + (size_t)n * numrho;
    const int per = (npts + slices - 1) / slices;
    const int i0 = slice * per, i1 = min(npts, i0 + per);
    if (use_smem) {
        for (int i = threadIdx.x; i < numrho; i += blockDim.x) hist[i] = 0;
        __syncthreads();
        for (int i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
            const uint32_t p = pts[i];
            atomicAdd(&hist[vote_r((float)(p & 0xffffu), (float)(p >> 16), c, s, half)], 1);
        }
        __syncthreads();
        for (int i = threadIdx.x; i < numrho; i += blockDim.x) {
            const int v = hist[i];
            if (v) {
                if (slices == 1) out[i] = v;
                else atomicAdd(out + i, v);
            }
        }
    } else {
        for (int i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
            const uint32_t p = pts[i];
            atomicAdd(out + vote_r((float)(p & 0xffffu), (float)(p >> 16), c, s, half), 1);
        }
    }
}

static int32_t compact_edges(oics_ctx *ctx, const uint8_t *dedges, int w, int h, size_t stride, bool want_mask,
                             int **d_total, uint32_t **d_pts, uint8_t **d_mask)
{
    cudaStream_t st = ctx->stream;
    if (w > 32767 || h > 32767) {
        set_error("hough: image larger than 32767 px per side is not supported");
        return OICS_ERR_UNSUPPORTED;
    }
    const size_t n = (size_t)w * h;
    // lines: [total (16 B)][rowoff h ints] ; labels: points ; mag: mask
    OICS_TRY(ctx->lines.ensure(16 + sizeof(int) * (size_t)h));
    OICS_TRY(ctx->labels.ensure(sizeof(uint32_t) * n));
    if (want_mask) OICS_TRY(ctx->mag.ensure(n));
    int *total = ctx->lines.as<int>();
    int *rowoff = total + 4;
    uint32_t *pts = ctx->labels.as<uint32_t>();
    uint8_t *mask = want_mask ? ctx->mag.as<uint8_t>() : nullptr;
    edge_row_count_kernel<<<cdiv(h, 8), 256, 0, st>>>(dedges, w, h, stride, rowoff);
    edge_row_scan_kernel<<<1, 1024, 0, st>>>(rowoff, h, total);
    edge_row_write_kernel<<<cdiv(h, 8), 256, 0, st>>>(dedges, w, h, stride, rowoff, pts, mask);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 3;
    *d_total = total;
    *d_pts = pts;
    if (d_mask) *d_mask = mask;
    return OICS_OK;
}

int32_t hough_accumulate_device(oics_ctx *ctx, const uint8_t *dedges, int w, int h, size_t stride, double rho,
                                double theta, int32_t *daccum)
{
    cudaStream_t st = ctx->stream;
    int numangle, numrho;
    std::vector<float> ttab;
    hough_geometry(w, h, rho, theta, &numangle, &numrho, &ttab);
    int *d_total;
    uint32_t *d_pts;
    OICS_TRY(compact_edges(ctx, dedges, w, h, stride, false, &d_total, &d_pts, nullptr));
    OICS_TRY(ctx->twid.ensure(ttab.size() * sizeof(float)));
    OICS_CUDA_TRY(cudaMemcpyAsync(ctx->twid.p, ttab.data(), ttab.size() * sizeof(float), cudaMemcpyHostToDevice, st));
    const size_t acc_bytes = (size_t)numangle * numrho * sizeof(int32_t);
    OICS_CUDA_TRY(cudaMemsetAsync(daccum, 0, acc_bytes, st));
    const size_t smem = (size_t)numrho * sizeof(int);
    const int use_smem = smem <= (size_t)ctx->smem_optin - 1024;
    if (use_smem && smem > 48 * 1024)
        OICS_CUDA_TRY(cudaFuncSetAttribute(hough_accumulate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // enough CTAs to fill the machine: numangle x slices >= ~4 waves
    int slices = std::max(1, (4 * ctx->sm_count + numangle - 1) / numangle);
    dim3 grid(numangle, slices);
    hough_accumulate_kernel<<<grid, 256, use_smem ? smem : 0, st>>>(d_pts, d_total, ctx->twid.as<float>(), numrho, slices,
                                                                    use_smem, daccum);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    // ttab is a host temporary: the async copy above must have consumed it
    OICS_CUDA_TRY(cudaStreamSynchronize(st));
    return OICS_OK;
}

// ---------------------------------------------------------------------------
// PPHT
__global__ void gather_points_kernel(const uint32_t *__restrict__ pts, const int *__restrict__ order, int n,
                                     uint32_t *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = pts[order[i]];
}

struct PphtArgs {
    const uint32_t *seq;  // points in visiting order
    int npts;
    uint8_t *mask;        // w*h, 1 = live edge pixel
    int w, h;
    int32_t *accum;       // numangle * numrho, zeroed
    const float *ttab;    // cos, sin per theta (already divided by rho)
    const int *dirs;      // per theta: xflag, dx0, dy0, pad
    int numangle, numrho;
    int threshold, line_length, line_gap;
    int32_t *lines;       // cap * 4
    int cap;
    int *n_lines;
    uint32_t *list;       // accepted pixels of the current segment: 2 * max(w, h) + 2 entries
};

// step t of the walk from (x0, y0) with (dx, dy): pixel (j1, i1)
__device__ __forceinline__ void walk_pixel(int xflag, int x0, int y0, int dx, int dy, int t, int *j1, int *i1)
{
    const int x = x0 + t * dx, y = y0 + t * dy;
    if (xflag) { *j1 = x; *i1 = y >> 16; }
    else { *j1 = x >> 16; *i1 = y; }
}

// cv::HoughLinesProbabilistic, one CTA of PPHT_K warps per image, PPHT_K points in flight.
//
// The algorithm is sequential by definition (RNG visiting order; every accepted segment erases
// pixels and votes that later points see), so one image is latency-bound: a live point costs one
// L2 round trip for its 180 vote atomics plus one per 16 walk steps.  The kernel keeps the exact
// sequential semantics and hides that latency by SPECULATING on the next PPHT_K live points:
//   vote    all points of the group vote in one round trip.  Theta n is always voted (and later
//           un-voted) by the same thread, and a thread issues its atomics in visiting order, so
//           same-address ordering gives every point exactly the counts the sequential loop
//           would have seen (the earlier points of the group included, the later ones not);
//   walk    warp q walks point q in both directions, READ-ONLY (no pixel is cleared while any
//           walk of the group is in flight);
//   commit  in visiting order, uniformly on every warp: a point commits if no earlier segment
//           of the group erased it and its scanned path cannot meet an earlier erased range
//           (bounding boxes); committing clears the whole step range [-e1, e0] (cv clears the
//           live pixels of it, the others are zero already) and kills the pending candidates
//           that sat on it.  The first point that fails, and everything after it, is rolled
//           back: speculative state is votes only, they are subtracted by the threads that added
//           them, and the points return to the pending set.  An accepted ("good") segment
//           un-votes its pixels, so it also ends the group.  The first point of a group always
//           commits, so every group makes progress.
constexpr int PPHT_K = 4;
constexpr int PPHT_THREADS = 32 * PPHT_K;
constexpr int PPHT_NT = 2;    // thetas per thread held in registers (numangle <= 256)

struct PphtSeg { int xflag, x0, y0, dx, dy, e0, e1; };   // (dx, dy) = forward direction

__device__ __forceinline__ void seg_box(const PphtSeg &s, int a, int b, int *bx0, int *by0, int *bx1, int *by1)
{
    int ax, ay, cx, cy;
    walk_pixel(s.xflag, s.x0, s.y0, s.dx, s.dy, a, &ax, &ay);
    walk_pixel(s.xflag, s.x0, s.y0, s.dx, s.dy, -b, &cx, &cy);
    *bx0 = min(ax, cx) - 1; *bx1 = max(ax, cx) + 1;
    *by0 = min(ay, cy) - 1; *by1 = max(ay, cy) + 1;
}

__global__ void __launch_bounds__(PPHT_THREADS) ppht_kernel(const PphtArgs A)
{
    __shared__ unsigned long long sKey[PPHT_K][PPHT_K];   // [warp][point] partial arg-max keys
    __shared__ int sRes[PPHT_K][2];                       // [point] last accepted step forwards / backwards
    __shared__ uint32_t sList[2 * PPHT_THREADS];          // accepted-segment pixels, one chunk

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int half = (A.numrho - 1) / 2;
    const int d = lane >> 4, s16 = lane & 15;            // walk direction and step slot of this lane
    const int nth = (A.numangle + PPHT_THREADS - 1) / PPHT_THREADS;   // thetas per thread: n = tid + 128 k
    float tc[PPHT_NT], ts[PPHT_NT];
#pragma unroll
    for (int k = 0; k < PPHT_NT; k++) {
        const int n = tid + PPHT_THREADS * k;
        tc[k] = n < A.numangle ? A.ttab[2 * n] : 0.f;
        ts[k] = n < A.numangle ? A.ttab[2 * n + 1] : 0.f;
    }
    int nlines = 0;
    uint32_t pnext = lane < A.npts ? A.seq[lane] : 0u;
    for (int t0 = 0; t0 < A.npts; t0 += 32) {
        // every warp holds the same 32-point window (lane = point); the next one is prefetched
        const int t = t0 + lane;
        const uint32_t p = pnext;
        pnext = t + 32 < A.npts ? A.seq[t + 32] : 0u;
        const int px = (int)(p & 0xffffu), py = (int)(p >> 16);
        const bool live0 = t < A.npts && __ldcg(A.mask + (size_t)py * A.w + px) != 0;
        uint32_t pending = __ballot_sync(0xffffffffu, live0);
        while (pending) {
            // ---- the group: the next PPHT_K pending points in visiting order
            int gi[PPHT_K], gx[PPHT_K], gy[PPHT_K], gn = 0;
            {
                uint32_t rest = pending;
#pragma unroll
                for (int q = 0; q < PPHT_K; q++) {
                    gi[q] = rest ? __ffs(rest) - 1 : 0;
                    if (rest) gn = q + 1;
                    rest &= rest - 1;
                    gx[q] = __shfl_sync(0xffffffffu, px, gi[q]);
                    gy[q] = __shfl_sync(0xffffffffu, py, gi[q]);
                }
            }
            // ---- vote: first maximal theta wins
            unsigned long long key[PPHT_K];
            if (nth <= PPHT_NT) {
                int val[PPHT_K][PPHT_NT];
#pragma unroll
                for (int q = 0; q < PPHT_K; q++) {
#pragma unroll
                    for (int k = 0; k < PPHT_NT; k++) {
                        const int n = tid + PPHT_THREADS * k;
                        val[q][k] = -0x40000000;
                        if (q < gn && n < A.numangle) {
                            int32_t *bin = A.accum + (size_t)n * A.numrho + vote_r((float)gx[q], (float)gy[q], tc[k], ts[k], half);
                            asm volatile("atom.global.add.s32 %0, [%1], 1;" : "=r"(val[q][k]) : "l"(bin) : "memory");
                        }
                    }
                }
#pragma unroll
                for (int q = 0; q < PPHT_K; q++) {
                    int bv = A.threshold - 1, bn = 0;
#pragma unroll
                    for (int k = 0; k < PPHT_NT; k++) {
                        const int n = tid + PPHT_THREADS * k;
                        if (n < A.numangle && bv < val[q][k] + 1) { bv = val[q][k] + 1; bn = n; }
                    }
                    key[q] = ((unsigned long long)(uint32_t)(bv + 0x40000000) << 32) | (uint32_t)(0x7fffffff - bn);
                }
            } else {
#pragma unroll
                for (int q = 0; q < PPHT_K; q++) {
                    int bv = A.threshold - 1, bn = 0;
                    if (q < gn)
                        for (int n = tid; n < A.numangle; n += PPHT_THREADS) {
                            const int v = atomicAdd(A.accum + (size_t)n * A.numrho +
                                                    vote_r((float)gx[q], (float)gy[q], A.ttab[2 * n], A.ttab[2 * n + 1], half), 1) + 1;
                            if (bv < v) { bv = v; bn = n; }
                        }
                    key[q] = ((unsigned long long)(uint32_t)(bv + 0x40000000) << 32) | (uint32_t)(0x7fffffff - bn);
                }
            }
            // key: larger value first, then smaller theta index
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
#pragma unroll
                for (int q = 0; q < PPHT_K; q++) {
                    const unsigned long long other = __shfl_xor_sync(0xffffffffu, key[q], o);
                    if (other > key[q]) key[q] = other;
                }
            }
            if (lane == 0) {
#pragma unroll
                for (int q = 0; q < PPHT_K; q++) sKey[warp][q] = key[q];
            }
            __syncthreads();
            int mval[PPHT_K], mn[PPHT_K];
#pragma unroll
            for (int q = 0; q < PPHT_K; q++) {
                unsigned long long kq = sKey[0][q];
#pragma unroll
                for (int w2 = 1; w2 < PPHT_K; w2++) kq = sKey[w2][q] > kq ? sKey[w2][q] : kq;
                mval[q] = (int)(uint32_t)(kq >> 32) - 0x40000000;
                mn[q] = 0x7fffffff - (int)(uint32_t)(kq & 0xffffffffu);
            }
            // ---- walk: warp q walks point q, both directions at once (lanes 0-15 forwards,
            // 16-31 backwards, 32 steps a round), reading only
            PphtSeg seg[PPHT_K];
#pragma unroll
            for (int q = 0; q < PPHT_K; q++) {
                const int4 dr = *reinterpret_cast<const int4 *>(A.dirs + 4 * mn[q]);
                seg[q].xflag = dr.x; seg[q].dx = dr.y; seg[q].dy = dr.z;
                seg[q].x0 = dr.x ? gx[q] : (gx[q] << 16) + 32768;
                seg[q].y0 = dr.x ? (gy[q] << 16) + 32768 : gy[q];
                seg[q].e0 = seg[q].e1 = 0;
            }
#pragma unroll
            for (int q = 0; q < PPHT_K; q++) {
                if (q == warp && q < gn && mval[q] >= A.threshold) {
                    const int dx = d ? -seg[q].dx : seg[q].dx, dy = d ? -seg[q].dy : seg[q].dy;
                    int last = 0;          // last accepted step of this lane's direction (step 0 = the point itself)
                    bool done = false;     // uniform inside each half-warp
                    for (int base = 1; __any_sync(0xffffffffu, !done); base += 32) {   // a round = steps base .. base + 31
                        // two loads per lane in flight: steps base + s16 and base + 16 + s16
                        int ja, ia, jb, ib;
                        walk_pixel(seg[q].xflag, seg[q].x0, seg[q].y0, dx, dy, base + s16, &ja, &ia);
                        walk_pixel(seg[q].xflag, seg[q].x0, seg[q].y0, dx, dy, base + 16 + s16, &jb, &ib);
                        const bool inba = ja >= 0 && ja < A.w && ia >= 0 && ia < A.h;
                        const bool inbb = jb >= 0 && jb < A.w && ib >= 0 && ib < A.h;
                        uint8_t va = 0, vb = 0;
                        if (!done && inba) va = __ldcg(A.mask + (size_t)ia * A.w + ja);
                        if (!done && inbb) vb = __ldcg(A.mask + (size_t)ib * A.w + jb);
                        const uint32_t bma = __ballot_sync(0xffffffffu, va != 0), bmb = __ballot_sync(0xffffffffu, vb != 0);
                        const uint32_t boa = __ballot_sync(0xffffffffu, !inba), bob = __ballot_sync(0xffffffffu, !inbb);
                        if (!done) {
                            uint32_t bits = ((bma >> (16 * d)) & 0xffffu) | (((bmb >> (16 * d)) & 0xffffu) << 16);
                            const uint32_t ob = ((boa >> (16 * d)) & 0xffffu) | (((bob >> (16 * d)) & 0xffffu) << 16);
                            const int first_oob = ob ? __ffs(ob) - 1 : 32;
                            if (first_oob < 32) bits &= (1u << first_oob) - 1u;
                            while (bits) {
                                const int sb = base + __ffs(bits) - 1;
                                bits &= bits - 1;
                                if (sb - last > A.line_gap + 1) { done = true; break; }
                                last = sb;
                            }
                            if (!done && first_oob < 32) done = true;      // left the image: line_end = last live step
                            if (!done && (base + 31) - last > A.line_gap) done = true;
                        }
                    }
                    if (s16 == 0) sRes[q][d] = last;
                }
            }
            __syncthreads();
            // ---- commit in visiting order (every warp runs this identically)
            int cbox[PPHT_K][4], ncommit = 0;    // erased ranges of the segments committed in this group
            int undo_from = gn;
#pragma unroll
            for (int q = 0; q < PPHT_K; q++) {
                if (q < gn && q < undo_from) {
                    const uint32_t bit = 1u << gi[q];
                    const bool walked = mval[q] >= A.threshold;
                    bool ok = (pending & bit) != 0;          // else an earlier segment of the group erased it
                    if (ok && walked) {
                        seg[q].e0 = sRes[q][0];
                        seg[q].e1 = sRes[q][1];
                        // every pixel the walk looked at lies within line_gap + 1 steps of its ends
                        int bx0, by0, bx1, by1;
                        seg_box(seg[q], seg[q].e0 + A.line_gap + 1, seg[q].e1 + A.line_gap + 1, &bx0, &by0, &bx1, &by1);
#pragma unroll
                        for (int c = 0; c < PPHT_K; c++)
                            if (c < ncommit && bx0 <= cbox[c][2] && cbox[c][0] <= bx1 && by0 <= cbox[c][3] && cbox[c][1] <= by1)
                                ok = false;                  // may have read a pixel erased since: walk again
                    }
                    if (!ok) {
                        undo_from = q;
                    } else {
                        pending &= ~bit;
                        if (walked) {
                            const PphtSeg s = seg[q];
                            int ax, ay, bx, by;
                            walk_pixel(s.xflag, s.x0, s.y0, s.dx, s.dy, s.e0, &ax, &ay);
                            walk_pixel(s.xflag, s.x0, s.y0, s.dx, s.dy, -s.e1, &bx, &by);
                            const bool good = abs(bx - ax) >= A.line_length || abs(by - ay) >= A.line_length;
                            const int len = s.e0 + s.e1 + 1;
                            if (good) {
                                // un-vote the live pixels of [-e1, e0]; theta n stays with its thread
                                for (int c0 = 0; c0 < len; c0 += 2 * PPHT_THREADS) {
#pragma unroll
                                    for (int u = 0; u < 2; u++) {
                                        const int idx = c0 + tid + PPHT_THREADS * u;
                                        uint32_t e = 0xffffffffu;
                                        if (idx < len) {
                                            int j1, i1;
                                            walk_pixel(s.xflag, s.x0, s.y0, s.dx, s.dy, idx - s.e1, &j1, &i1);
                                            if (__ldcg(A.mask + (size_t)i1 * A.w + j1) != 0) e = (uint32_t)j1 | ((uint32_t)i1 << 16);
                                        }
                                        sList[tid + PPHT_THREADS * u] = e;
                                    }
                                    __syncthreads();
                                    const int cn = min(2 * PPHT_THREADS, len - c0);
                                    for (int i = 0; i < cn; i++) {
                                        const uint32_t e = sList[i];
                                        if (e == 0xffffffffu) continue;
                                        const float fx = (float)(e & 0xffffu), fy = (float)(e >> 16);
                                        if (nth <= PPHT_NT) {
#pragma unroll
                                            for (int k = 0; k < PPHT_NT; k++) {
                                                const int n = tid + PPHT_THREADS * k;
                                                if (n < A.numangle)
                                                    atomicSub(A.accum + (size_t)n * A.numrho + vote_r(fx, fy, tc[k], ts[k], half), 1);
                                            }
                                        } else {
                                            for (int n = tid; n < A.numangle; n += PPHT_THREADS)
                                                atomicSub(A.accum + (size_t)n * A.numrho + vote_r(fx, fy, A.ttab[2 * n], A.ttab[2 * n + 1], half), 1);
                                        }
                                    }
                                    __syncthreads();
                                }
                                if (tid == 0 && nlines < A.cap) {
                                    A.lines[4 * nlines + 0] = ax;
                                    A.lines[4 * nlines + 1] = ay;
                                    A.lines[4 * nlines + 2] = bx;
                                    A.lines[4 * nlines + 3] = by;
                                }
                                nlines++;
                                undo_from = q + 1;           // the counts later points saw are stale now
                            }
                            // erase the range (cv clears its live pixels; the others are zero already)
                            for (int idx = tid; idx < len; idx += PPHT_THREADS) {
                                int j1, i1;
                                walk_pixel(s.xflag, s.x0, s.y0, s.dx, s.dy, idx - s.e1, &j1, &i1);
                                __stcg(A.mask + (size_t)i1 * A.w + j1, (uint8_t)0);
                            }
                            // pending candidates of this window that sat on the range are dead now
                            {
                                bool dead = false;
                                if ((pending >> lane) & 1u) {
                                    const int tt = s.xflag ? (px - gx[q]) * s.dx : (py - gy[q]) * s.dy;   // +-1 on the major axis
                                    int cj, ci;
                                    walk_pixel(s.xflag, s.x0, s.y0, s.dx, s.dy, tt, &cj, &ci);
                                    dead = cj == px && ci == py && tt <= s.e0 && -tt <= s.e1;
                                }
                                pending &= ~__ballot_sync(0xffffffffu, dead);
                            }
                            seg_box(s, s.e0, s.e1, &cbox[ncommit][0], &cbox[ncommit][1], &cbox[ncommit][2], &cbox[ncommit][3]);
                            ncommit++;
                        }
                    }
                }
            }
            // ---- roll back the votes of the points that did not commit (same thread, same bins)
#pragma unroll
            for (int q = 0; q < PPHT_K; q++) {
                if (q >= undo_from && q < gn) {
                    if (nth <= PPHT_NT) {
#pragma unroll
                        for (int k = 0; k < PPHT_NT; k++) {
                            const int n = tid + PPHT_THREADS * k;
                            if (n < A.numangle)
                                atomicSub(A.accum + (size_t)n * A.numrho + vote_r((float)gx[q], (float)gy[q], tc[k], ts[k], half), 1);
                        }
                    } else {
                        for (int n = tid; n < A.numangle; n += PPHT_THREADS)
                            atomicSub(A.accum + (size_t)n * A.numrho +
                                      vote_r((float)gx[q], (float)gy[q], A.ttab[2 * n], A.ttab[2 * n + 1], half), 1);
                    }
                }
            }
        }
        // bar.sync orders this window's erasing stores (st.cg) before the next window's liveness loads (ld.cg)
        __syncthreads();
    }
    if (tid == 0) *A.n_lines = nlines;
}

// The visiting order of cv::HoughLinesProbabilistic for n points: count = n..1,
// idx = rng % count, visit arr[idx], arr[idx] = arr[count-1]; cv::RNG(-1) is the
// multiply-with-carry generator state = lo32 * 4164903690 + hi32.
static void ppht_order(int n, std::vector<int> &order)
{
    std::vector<int> arr((size_t)n);
    for (int i = 0; i < n; i++) arr[i] = i;
    order.resize((size_t)n);
    uint64_t state = (uint64_t)-1;
    int k = 0;
    for (int count = n; count > 0; count--) {
        state = (uint64_t)(uint32_t)state * 4164903690u + (uint32_t)(state >> 32);
        const int idx = (int)((uint32_t)state % (uint32_t)count);
        order[k++] = arr[idx];
        arr[idx] = arr[count - 1];
    }
}

int32_t hough_lines_p_device(oics_ctx *ctx, const uint8_t *dedges, int w, int h, size_t stride, double rho,
                             double theta, int threshold, int line_length, int line_gap, int32_t *h_lines,
                             int cap, int *n_lines)
{
    cudaStream_t st = ctx->stream;
    int numangle, numrho;
    std::vector<float> ttab;
    hough_geometry(w, h, rho, theta, &numangle, &numrho, &ttab);
    // per-theta walk direction (cv::HoughLinesProbabilistic): a = -sin, b = cos
    std::vector<int> dirs(4 * (size_t)numangle);
    for (int n = 0; n < numangle; n++) {
        const float a = -ttab[2 * n + 1], b = ttab[2 * n];
        int xflag, dx0, dy0;
        if (std::fabs(a) > std::fabs(b)) {
            xflag = 1;
            dx0 = a > 0 ? 1 : -1;
            dy0 = (int)std::lrint(b * (1 << 16) / std::fabs((double)a));
        } else {
            xflag = 0;
            dy0 = b > 0 ? 1 : -1;
            dx0 = (int)std::lrint(a * (1 << 16) / std::fabs((double)b));
        }
        dirs[4 * n] = xflag; dirs[4 * n + 1] = dx0; dirs[4 * n + 2] = dy0; dirs[4 * n + 3] = 0;
    }
    int *d_total;
    uint32_t *d_pts;
    uint8_t *d_mask;
    OICS_TRY(compact_edges(ctx, dedges, w, h, stride, true, &d_total, &d_pts, &d_mask));
    int npts = 0;
    OICS_CUDA_TRY(cudaMemcpyAsync(&npts, d_total, sizeof(int), cudaMemcpyDeviceToHost, st));
    OICS_CUDA_TRY(cudaStreamSynchronize(st));
    *n_lines = 0;
    if (npts == 0) return OICS_OK;

    std::vector<int> order;
    ppht_order(npts, order);
    const size_t acc_bytes = (size_t)numangle * numrho * sizeof(int32_t);
    OICS_TRY(ctx->accum.ensure(acc_bytes));
    // misc: [order npts][seq npts][ttab][dirs][n_lines][lines cap*4]
    const size_t off_seq = sizeof(int) * (size_t)npts;
    const size_t off_ttab = (off_seq + sizeof(uint32_t) * (size_t)npts + 15) & ~(size_t)15;
    const size_t off_dirs = (off_ttab + sizeof(float) * ttab.size() + 15) & ~(size_t)15;   // int4 loads
    const size_t off_nl = off_dirs + sizeof(int) * dirs.size();
    const size_t off_lines = off_nl + 16;
    const int dcap = std::max(cap, 1);
    const size_t off_list = off_lines + sizeof(int32_t) * 4 * (size_t)dcap;
    OICS_TRY(ctx->misc.ensure(off_list + sizeof(uint32_t) * (2 * (size_t)std::max(w, h) + 2)));
    uint8_t *base = ctx->misc.as<uint8_t>();
    OICS_CUDA_TRY(cudaMemcpyAsync(base, order.data(), sizeof(int) * (size_t)npts, cudaMemcpyHostToDevice, st));
    OICS_CUDA_TRY(cudaMemcpyAsync(base + off_ttab, ttab.data(), sizeof(float) * ttab.size(), cudaMemcpyHostToDevice, st));
    OICS_CUDA_TRY(cudaMemcpyAsync(base + off_dirs, dirs.data(), sizeof(int) * dirs.size(), cudaMemcpyHostToDevice, st));
    OICS_CUDA_TRY(cudaMemsetAsync(ctx->accum.p, 0, acc_bytes, st));
    gather_points_kernel<<<cdiv(npts, 256), 256, 0, st>>>(d_pts, reinterpret_cast<const int *>(base), npts,
                                                         reinterpret_cast<uint32_t *>(base + off_seq));
    PphtArgs A;
    A.seq = reinterpret_cast<const uint32_t *>(base + off_seq);
    A.npts = npts;
    A.mask = d_mask;
    A.w = w; A.h = h;
    A.accum = ctx->accum.as<int32_t>();
    A.ttab = reinterpret_cast<const float *>(base + off_ttab);
    A.dirs = reinterpret_cast<const int *>(base + off_dirs);
    A.numangle = numangle; A.numrho = numrho;
    A.threshold = threshold; A.line_length = line_length; A.line_gap = line_gap;
    A.lines = reinterpret_cast<int32_t *>(base + off_lines);
    A.cap = cap;
    A.n_lines = reinterpret_cast<int *>(base + off_nl);
    A.list = reinterpret_cast<uint32_t *>(base + off_list);
    ppht_kernel<<<1, PPHT_THREADS, 0, st>>>(A);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 2;
    int n = 0;
    OICS_CUDA_TRY(cudaMemcpyAsync(&n, A.n_lines, sizeof(int), cudaMemcpyDeviceToHost, st));
    OICS_CUDA_TRY(cudaStreamSynchronize(st));
    *n_lines = n;
    const int ncopy = std::min(n, cap);
    if (ncopy > 0) {
        if (is_device_ptr(h_lines))
            OICS_CUDA_TRY(cudaMemcpyAsync(h_lines, A.lines, sizeof(int32_t) * 4 * (size_t)ncopy, cudaMemcpyDeviceToDevice, st));
        else
            OICS_CUDA_TRY(cudaMemcpyAsync(h_lines, A.lines, sizeof(int32_t) * 4 * (size_t)ncopy, cudaMemcpyDeviceToHost, st));
        OICS_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return OICS_OK;
}

}  // namespace oics
