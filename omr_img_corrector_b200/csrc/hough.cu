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
//        * the next sixteen live points are processed speculatively at once on a cluster of
//          four CTAs: all vote in one L2 round trip (one thread per theta, thetas spread over
//          the four SMs, atomics issued in visiting order, so each point sees exactly the
//          sequential counts), one warp walks each point read-only, and the results commit in
//          visiting order; a point that an earlier commit could have influenced is rolled
//          back (its votes subtracted) and retried;
//        * the arg-max keeps the first maximal theta, like the sequential `max_val < val`;
//        * the two line walks test 32 steps each per round trip; gap logic runs on ballots;
//        * a commit clears the step range and, for accepted segments, un-votes its live pixels.
//      Latency-bound per image (like the CPU), throughput comes from images in flight.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <thread>

#include <cooperative_groups.h>

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
        // eight independent point loads in flight per thread before their votes (the loop is
        // latency-bound otherwise: one L2 round trip per vote)
        for (int base = i0 + threadIdx.x; base < i1; base += 8 * blockDim.x) {
            uint32_t p[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int i = base + u * blockDim.x;
                p[u] = i < i1 ? __ldg(pts + i) : 0xffffffffu;
            }
#pragma unroll
            for (int u = 0; u < 8; u++)
                if (p[u] != 0xffffffffu) atomicAdd(&hist[vote_r((float)(p[u] & 0xffffu), (float)(p[u] >> 16), c, s, half)], 1);
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
    // one full wave, no tail: 4 CTAs fit an SM (48 KB histogram each), numangle x slices <= 4 x SMs
    int slices = std::max(1, (4 * ctx->sm_count) / numangle);
    dim3 grid(numangle, slices);
    {
        ProfScope ps(ctx, "hough_accumulate_kernel", (double)w * h + (double)acc_bytes);
        hough_accumulate_kernel<<<grid, 256, use_smem ? smem : 0, st>>>(d_pts, d_total, ctx->twid.as<float>(), numrho, slices,
                                                                        use_smem, daccum);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    // ttab is a host temporary: the async copy above must have consumed it
    OICS_CUDA_TRY(cudaStreamSynchronize(st));
    return OICS_OK;
}

// ---------------------------------------------------------------------------
// Peak extraction over the standard accumulator: cv::HoughLines' findLocalMaximums rule
// (count > threshold, > left and upper neighbour, >= right and lower neighbour; outside = 0) followed
// by its ordering (count descending, then cell index ascending).  One thread per cell, loads coalesced
// along rho; peaks leave as 64-bit keys (count << 32 | ~cell) through a warp-aggregated append, so the
// order of the appends does not matter: the host sorts the keys, which are unique.
__global__ void __launch_bounds__(256) hough_peaks_kernel(const int32_t *__restrict__ accum, int numangle, int numrho,
                                                          int threshold, unsigned long long *__restrict__ keys,
                                                          unsigned int cap, unsigned int *__restrict__ count)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x, n = blockIdx.y;
    bool peak = false;
    int v = 0;
    if (r < numrho) {
        const int32_t *row = accum + (size_t)n * numrho;
        v = row[r];
        if (v > threshold) {
            const int left = r > 0 ? row[r - 1] : 0, right = r + 1 < numrho ? row[r + 1] : 0;
            const int up = n > 0 ? row[r - numrho] : 0, down = n + 1 < numangle ? row[r + numrho] : 0;
            peak = v > left && v >= right && v > up && v >= down;
        }
    }
    const uint32_t m = __ballot_sync(0xffffffffu, peak);
    if (!m) return;
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    unsigned int base = 0;
    if (lane == leader) base = atomicAdd(count, (unsigned int)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (peak) {
        const unsigned int slot = base + __popc(m & ((1u << lane) - 1u));
        const uint32_t cell = (uint32_t)n * (uint32_t)numrho + (uint32_t)r;
        if (slot < cap) keys[slot] = ((unsigned long long)(uint32_t)v << 32) | (uint32_t)~cell;
    }
}

int32_t hough_peaks_device(oics_ctx *ctx, const int32_t *daccum, int numangle, int numrho, int threshold,
                           std::vector<int32_t> *peaks)
{
    cudaStream_t st = ctx->stream;
    const size_t cells = (size_t)numangle * numrho;
    if (cells == 0 || cells > 0x7fffffffull) {
        set_error("hough peaks: accumulator of %d x %d cells is not supported", numangle, numrho);
        return OICS_ERR_INVALID_ARG;
    }
    // a peak is strictly greater than its left neighbour: at most every second cell of a row
    const size_t cap = (size_t)numangle * ((size_t)numrho / 2 + 1);
    OICS_TRY(ctx->lines.ensure(16 + cap * sizeof(unsigned long long)));
    unsigned int *d_count = ctx->lines.as<unsigned int>();
    unsigned long long *d_keys = reinterpret_cast<unsigned long long *>(ctx->lines.as<uint8_t>() + 16);
    OICS_CUDA_TRY(cudaMemsetAsync(d_count, 0, 16, st));
    dim3 grid(cdiv(numrho, 256), numangle);
    {
        ProfScope ps(ctx, "hough_peaks_kernel", (double)cells * sizeof(int32_t));
        hough_peaks_kernel<<<grid, 256, 0, st>>>(daccum, numangle, numrho, threshold, d_keys, (unsigned int)cap, d_count);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    unsigned int n = 0;
    OICS_CUDA_TRY(cudaMemcpyAsync(&n, d_count, sizeof(n), cudaMemcpyDeviceToHost, st));
    OICS_CUDA_TRY(cudaStreamSynchronize(st));
    std::vector<unsigned long long> keys(n);
    if (n) {
        OICS_CUDA_TRY(cudaMemcpyAsync(keys.data(), d_keys, (size_t)n * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        OICS_CUDA_TRY(cudaStreamSynchronize(st));
    }
    std::sort(keys.begin(), keys.end(), [](unsigned long long a, unsigned long long b) { return a > b; });
    peaks->resize((size_t)n * 3);
    for (unsigned int i = 0; i < n; i++) {
        const uint32_t cell = ~(uint32_t)(keys[i] & 0xffffffffull);
        (*peaks)[3 * (size_t)i + 0] = (int32_t)(cell / (uint32_t)numrho);
        (*peaks)[3 * (size_t)i + 1] = (int32_t)(cell % (uint32_t)numrho);
        (*peaks)[3 * (size_t)i + 2] = (int32_t)(keys[i] >> 32);
    }
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

// cv::HoughLinesProbabilistic, one thread-block CLUSTER per image, PPHT_K points in flight.
//
// The algorithm is sequential by definition (RNG visiting order; every accepted segment erases
// pixels and votes that later points see), so one image is latency-bound: a live point costs one
// L2 round trip for its 180 vote atomics plus one per 32 walk steps, and ONE SM can only issue a
// scattered returning atomic every ~4 cycles.  The kernel keeps the exact sequential semantics and
// hides both by SPECULATING on the next PPHT_K live points with the work spread over the
// PPHT_C SMs of a cluster:
//   vote    theta n belongs to ONE thread of ONE CTA (n mod PPHT_C picks the CTA) for every vote,
//           un-vote and roll-back, and that thread issues its atomics in visiting order, so
//           same-address program order gives every point exactly the counts the sequential loop
//           would have seen (the earlier points of the group included, the later ones not).  Each
//           CTA reduces its thetas' arg-max per point (REDUX on a packed 32-bit key) and writes it
//           into every CTA's shared memory (DSMEM); barrier.cluster #1;
//   walk    point q is walked by warp q / PPHT_C of CTA q mod PPHT_C, both directions, 64 steps
//           each per round trip, READ-ONLY (no pixel is cleared while any walk of the group is in
//           flight); the two end steps go to every CTA's shared memory; barrier.cluster #2;
//   commit  in visiting order, identically on every warp of every CTA, resolved with ballots (lane
//           q prepares point q, every lane tests its points against all segments of the group): a
//           point commits if no earlier segment of the group erased it and its scanned path cannot
//           meet an earlier erased range (bounding boxes); committing clears the whole step range [-e1, e0] (cv
//           clears the live pixels of it, the others are zero already; the stores are shared out
//           over the cluster) and kills the pending candidates that sat on it.  The first point
//           that fails, and everything after it, is rolled back: speculative state is votes only,
//           they are subtracted by the threads that added them, and the points return to the
//           pending set.  An accepted ("good") segment un-votes its pixels, so it also ends the
//           group.  The first point of a group always commits, so every group makes progress.
// barrier.cluster (release/acquire at cluster scope) orders a group's erasing stores before the
// next group's walk loads; the accumulator needs no fence at all (one thread per theta).
constexpr int PPHT_C = 4;                     // CTAs (SMs) per image
constexpr int PPHT_W = 4;                     // warps per CTA
constexpr int PPHT_T = 32 * PPHT_W;           // threads per CTA
constexpr int PPHT_K = PPHT_C * PPHT_W;       // points in flight

struct PphtSeg { int xflag, x0, y0, dx, dy, e0, e1; };   // (dx, dy) = forward direction

__device__ __forceinline__ void seg_box(const PphtSeg &s, int a, int b, int *bx0, int *by0, int *bx1, int *by1)
{
    int ax, ay, cx, cy;
    walk_pixel(s.xflag, s.x0, s.y0, s.dx, s.dy, a, &ax, &ay);
    walk_pixel(s.xflag, s.x0, s.y0, s.dx, s.dy, -b, &cx, &cy);
    *bx0 = min(ax, cx) - 1; *bx1 = max(ax, cx) + 1;
    *by0 = min(ay, cy) - 1; *by1 = max(ay, cy) + 1;
}

__device__ __forceinline__ unsigned long long ppht_key(int bv, int bn)   // larger value first, then smaller theta
{
    return ((unsigned long long)(uint32_t)(bv + 0x40000000) << 32) | (uint32_t)(0x7fffffff - bn);
}

__global__ void __cluster_dims__(PPHT_C, 1, 1) __launch_bounds__(PPHT_T) ppht_kernel(const PphtArgs *__restrict__ images)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    // one cluster per image of the batch (images in flight = the reference's one thread per card,
    // thread_pool.rs:41-73); the image's arguments are staged in shared memory once
    __shared__ PphtArgs sArgs;
    if (threadIdx.x == 0) sArgs = images[blockIdx.x / PPHT_C];
    __syncthreads();
    const PphtArgs &A = sArgs;
    __shared__ unsigned long long sKey[2][PPHT_C][PPHT_K];   // [group parity][source CTA][point] arg-max keys
    __shared__ unsigned long long sPart[PPHT_W][PPHT_K];     // [warp][point] partial keys of this CTA
    __shared__ int sRes[PPHT_K][2];                          // [point] last accepted step forwards / backwards
    __shared__ uint32_t sList[2 * PPHT_T];                   // accepted-segment pixels, one chunk
    __shared__ int4 sDirs[PPHT_C * PPHT_T];                  // per-theta walk direction (fast path: numangle <= 512)
    __shared__ int4 sSeg[PPHT_W][PPHT_K][3];                 // [warp][point] walk geometry, ends, erased box (per-warp copy)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cta = (int)cluster.block_rank();
    const int half = (A.numrho - 1) / 2;
    const int d = lane >> 4, s16 = lane & 15;                // walk direction and step slot of this lane
    // thetas of this thread: n = cta + PPHT_C * (tid + PPHT_T * k)
    const int n0 = cta + PPHT_C * tid;
    const bool fast = A.numangle <= PPHT_C * PPHT_T && A.numangle <= 1024 && A.threshold >= 0;
    const bool voter = n0 < A.numangle;
    const bool warp_votes = cta + PPHT_C * (warp * 32) < A.numangle;   // any voter in this warp
    const float tc0 = voter ? A.ttab[2 * n0] : 0.f, ts0 = voter ? A.ttab[2 * n0 + 1] : 0.f;
    if (fast) {
        for (int n = tid; n < A.numangle; n += PPHT_T) sDirs[n] = *reinterpret_cast<const int4 *>(A.dirs + 4 * n);
        __syncthreads();
    }
    int nlines = 0, parity = 0;
    // every warp of every CTA holds the same 64-point window: lane l keeps points l (half A) and
    // 32 + l (half B) of it; the next window is prefetched
    uint32_t pnextA = lane < A.npts ? A.seq[lane] : 0u, pnextB = lane + 32 < A.npts ? A.seq[lane + 32] : 0u;
    for (int t0 = 0; t0 < A.npts; t0 += 64) {
        const int t = t0 + lane;
        const uint32_t pA = pnextA, pB = pnextB;
        pnextA = t + 64 < A.npts ? A.seq[t + 64] : 0u;
        pnextB = t + 96 < A.npts ? A.seq[t + 96] : 0u;
        const int pxA = (int)(pA & 0xffffu), pyA = (int)(pA >> 16), pxB = (int)(pB & 0xffffu), pyB = (int)(pB >> 16);
        const bool liveA = t < A.npts && __ldcg(A.mask + (size_t)pyA * A.w + pxA) != 0;
        const bool liveB = t + 32 < A.npts && __ldcg(A.mask + (size_t)pyB * A.w + pxB) != 0;
        uint32_t pendA = __ballot_sync(0xffffffffu, liveA), pendB = __ballot_sync(0xffffffffu, liveB);
        while (pendA | pendB) {
            // ---- the group: the next PPHT_K pending points in visiting order (half A before half B);
            // lane q keeps point q: my_i in [0, 64) is its place in the window
            const int nA = __popc(pendA);
            const int gn = min(PPHT_K, nA + __popc(pendB));
            int my_i = 0;
            if (lane < gn) my_i = lane < nA ? (int)__fns(pendA, 0, lane + 1) : 32 + (int)__fns(pendB, 0, lane - nA + 1);
            const int xa = __shfl_sync(0xffffffffu, pxA, my_i & 31), ya = __shfl_sync(0xffffffffu, pyA, my_i & 31);
            const int xb = __shfl_sync(0xffffffffu, pxB, my_i & 31), yb = __shfl_sync(0xffffffffu, pyB, my_i & 31);
            const int my_x = my_i < 32 ? xa : xb, my_y = my_i < 32 ? ya : yb;
            // ---- vote: first maximal theta wins
            if (fast) {
                int val[PPHT_K];
#pragma unroll
                for (int q = 0; q < PPHT_K; q++) {
                    const int x = __shfl_sync(0xffffffffu, my_x, q), y = __shfl_sync(0xffffffffu, my_y, q);
                    val[q] = -1;
                    if (q < gn && voter) {
                        int32_t *bin = A.accum + (size_t)n0 * A.numrho + vote_r((float)x, (float)y, tc0, ts0, half);
                        asm volatile("atom.global.add.s32 %0, [%1], 1;" : "=r"(val[q]) : "l"(bin) : "memory");
                    }
                }
                if (warp_votes) {
#pragma unroll
                    for (int q = 0; q < PPHT_K; q++) {
                        // packed key: count << 10 | (1023 - theta); "no vote reaches the threshold" packs below every real one
                        uint32_t k32 = ((uint32_t)A.threshold << 10) | 1023u;    // (threshold - 1) + 1, theta 0
                        if (voter && q < gn && val[q] + 1 >= A.threshold) k32 = ((uint32_t)(val[q] + 2) << 10) | (uint32_t)(1023 - n0);
                        k32 = __reduce_max_sync(0xffffffffu, k32);
                        if (lane == q) sPart[warp][q] = ppht_key((int)(k32 >> 10) - 1, 1023 - (int)(k32 & 1023u));
                    }
                } else if (lane < PPHT_K) {
                    sPart[warp][lane] = ppht_key(A.threshold - 1, 0);
                }
            } else {
                for (int q = 0; q < PPHT_K; q++) {
                    const int x = __shfl_sync(0xffffffffu, my_x, q), y = __shfl_sync(0xffffffffu, my_y, q);
                    int bv = A.threshold - 1, bn = 0;
                    if (q < gn)
                        for (int n = n0; n < A.numangle; n += PPHT_C * PPHT_T) {
                            const int v = atomicAdd(A.accum + (size_t)n * A.numrho +
                                                    vote_r((float)x, (float)y, A.ttab[2 * n], A.ttab[2 * n + 1], half), 1) + 1;
                            if (bv < v) { bv = v; bn = n; }
                        }
                    unsigned long long key = ppht_key(bv, bn);
#pragma unroll
                    for (int o = 16; o >= 1; o >>= 1) {
                        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
                        if (other > key) key = other;
                    }
                    if (lane == q) sPart[warp][q] = key;
                }
            }
            __syncthreads();
            if (tid < PPHT_K) {   // this CTA's arg-max per point -> every CTA of the cluster
                unsigned long long kq = sPart[0][tid];
#pragma unroll
                for (int w2 = 1; w2 < PPHT_W; w2++) kq = sPart[w2][tid] > kq ? sPart[w2][tid] : kq;
#pragma unroll
                for (int c = 0; c < PPHT_C; c++) *cluster.map_shared_rank(&sKey[parity][cta][tid], c) = kq;
            }
            cluster.sync();
            // lane q: arg-max of point q over the cluster, and its walk geometry
            int my_val = A.threshold - 1, my_n = 0;
            PphtSeg ms = {0, 0, 0, 0, 0, 0, 0};
            if (lane < gn) {
                unsigned long long kq = sKey[parity][0][lane];
#pragma unroll
                for (int c = 1; c < PPHT_C; c++) kq = sKey[parity][c][lane] > kq ? sKey[parity][c][lane] : kq;
                my_val = (int)(uint32_t)(kq >> 32) - 0x40000000;
                my_n = 0x7fffffff - (int)(uint32_t)(kq & 0xffffffffu);
                const int4 dr = fast ? sDirs[my_n] : *reinterpret_cast<const int4 *>(A.dirs + 4 * my_n);
                ms.xflag = dr.x; ms.dx = dr.y; ms.dy = dr.z;
                ms.x0 = dr.x ? my_x : (my_x << 16) + 32768;
                ms.y0 = dr.x ? (my_y << 16) + 32768 : my_y;
            }
            const uint32_t walked_mask = __ballot_sync(0xffffffffu, lane < gn && my_val >= A.threshold);
            // ---- walk: point q on warp q / PPHT_C of CTA q % PPHT_C, both directions at once (lanes 0-15
            // forwards, 16-31 backwards, 64 steps a round), reading only
            {
                const int q = warp * PPHT_C + cta;
                if (q < gn && ((walked_mask >> q) & 1u)) {
                    const int xflag = __shfl_sync(0xffffffffu, ms.xflag, q), x0 = __shfl_sync(0xffffffffu, ms.x0, q);
                    const int y0 = __shfl_sync(0xffffffffu, ms.y0, q);
                    const int dxf = __shfl_sync(0xffffffffu, ms.dx, q), dyf = __shfl_sync(0xffffffffu, ms.dy, q);
                    const int dx = d ? -dxf : dxf, dy = d ? -dyf : dyf;
                    int last = 0;          // last accepted step of this lane's direction (step 0 = the point itself)
                    bool done = false;     // uniform inside each half-warp
                    for (int base = 1; __any_sync(0xffffffffu, !done); base += 64) {   // a round = steps base .. base + 63
                        // four loads per lane in flight: steps base + 16 u + s16
                        unsigned long long bits = 0, ob = 0;
                        uint8_t v[4];
                        bool inb[4];
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            int j1, i1;
                            walk_pixel(xflag, x0, y0, dx, dy, base + 16 * u + s16, &j1, &i1);
                            inb[u] = j1 >= 0 && j1 < A.w && i1 >= 0 && i1 < A.h;
                            v[u] = 0;
                            if (!done && inb[u]) v[u] = __ldcg(A.mask + (size_t)i1 * A.w + j1);
                        }
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            const uint32_t bm = __ballot_sync(0xffffffffu, v[u] != 0), bo = __ballot_sync(0xffffffffu, !inb[u]);
                            bits |= (unsigned long long)((bm >> (16 * d)) & 0xffffu) << (16 * u);
                            ob |= (unsigned long long)((bo >> (16 * d)) & 0xffffu) << (16 * u);
                        }
                        if (!done) {
                            const int first_oob = ob ? __ffsll((long long)ob) - 1 : 64;
                            if (first_oob < 64) bits &= (1ull << first_oob) - 1ull;
                            while (bits) {
                                const int sb = base + __ffsll((long long)bits) - 1;
                                bits &= bits - 1;
                                if (sb - last > A.line_gap + 1) { done = true; break; }
                                last = sb;
                            }
                            if (!done && first_oob < 64) done = true;      // left the image: line_end = last live step
                            if (!done && (base + 63) - last > A.line_gap) done = true;
                        }
                    }
                    if (s16 == 0) {
#pragma unroll
                        for (int c = 0; c < PPHT_C; c++) *cluster.map_shared_rank(&sRes[q][d], c) = last;
                    }
                }
            }
            if (walked_mask) cluster.sync();     // uniform over the cluster
            // ---- commit in visiting order, resolved in parallel (every warp of every CTA runs this
            // identically).  Lane q prepares point q; then every lane tests its points against all
            // walked segments of the group; the first point that fails is found with ballots.
            const bool walked_me = (walked_mask >> lane) & 1u;
            int ax = 0, ay = 0, bx = 0, by = 0, len = 0;
            bool good_me = false;
            int sb0 = 0, sb1 = 0, sb2 = 0, sb3 = 0;        // what this point's walk looked at
            if (walked_me) {
                ms.e0 = sRes[lane][0];
                ms.e1 = sRes[lane][1];
                walk_pixel(ms.xflag, ms.x0, ms.y0, ms.dx, ms.dy, ms.e0, &ax, &ay);
                walk_pixel(ms.xflag, ms.x0, ms.y0, ms.dx, ms.dy, -ms.e1, &bx, &by);
                good_me = abs(bx - ax) >= A.line_length || abs(by - ay) >= A.line_length;
                len = ms.e0 + ms.e1 + 1;
                // every pixel the walk looked at lies within line_gap + 1 steps of its ends
                seg_box(ms, ms.e0 + A.line_gap + 1, ms.e1 + A.line_gap + 1, &sb0, &sb1, &sb2, &sb3);
                int4 *o = sSeg[warp][lane];
                o[0] = make_int4(ms.xflag, ms.x0, ms.y0, ms.dx);
                o[1] = make_int4(ms.dy, ms.e0, ms.e1, len);
                o[2] = make_int4(min(ax, bx) - 1, min(ay, by) - 1, max(ax, bx) + 1, max(ay, by) + 1);   // erased range
            }
            __syncwarp();
            uint32_t deadA = 0, deadB = 0, failG = 0;     // bit c: segment c erases my window points / defeats my group point
            const bool pendWA = (pendA >> lane) & 1u, pendWB = (pendB >> lane) & 1u;
            for (uint32_t rest = walked_mask; rest; rest &= rest - 1) {
                const int c = __ffs(rest) - 1;
                const int4 *sp = sSeg[warp][c];
                const int4 s2 = sp[2];                           // erased range of segment c, as a box
                const bool member = lane < gn && c < lane;
                if (member && walked_me && sb0 <= s2.z && s2.x <= sb2 && sb1 <= s2.w && s2.y <= sb3)
                    failG |= 1u << c;                            // may have read an erased pixel: walk again
                const bool nearA = pendWA && pxA >= s2.x && pxA <= s2.z && pyA >= s2.y && pyA <= s2.w;
                const bool nearB = pendWB && pxB >= s2.x && pxB <= s2.z && pyB >= s2.y && pyB <= s2.w;
                const bool nearG = member && my_x >= s2.x && my_x <= s2.z && my_y >= s2.y && my_y <= s2.w;
                if (!__any_sync(0xffffffffu, nearA || nearB || nearG)) continue;      // nobody near it: the usual case
                const int4 s0 = sp[0], s1 = sp[1];
                const int xflag = s0.x, x0 = s0.y, y0 = s0.z, dx = s0.w, dy = s1.x, e0 = s1.y, e1 = s1.z;
                const int gxc = xflag ? x0 : x0 >> 16, gyc = xflag ? y0 >> 16 : y0;
                int cj, ci;
                if (nearA) {
                    const int tt = xflag ? (pxA - gxc) * dx : (pyA - gyc) * dy;      // +-1 on the major axis
                    walk_pixel(xflag, x0, y0, dx, dy, tt, &cj, &ci);
                    if (cj == pxA && ci == pyA && tt <= e0 && -tt <= e1) deadA |= 1u << c;
                }
                if (nearB) {
                    const int tt = xflag ? (pxB - gxc) * dx : (pyB - gyc) * dy;
                    walk_pixel(xflag, x0, y0, dx, dy, tt, &cj, &ci);
                    if (cj == pxB && ci == pyB && tt <= e0 && -tt <= e1) deadB |= 1u << c;
                }
                if (nearG) {
                    const int tt = xflag ? (my_x - gxc) * dx : (my_y - gyc) * dy;
                    walk_pixel(xflag, x0, y0, dx, dy, tt, &cj, &ci);
                    if (cj == my_x && ci == my_y && tt <= e0 && -tt <= e1) failG |= 1u << c;      // erased before its turn
                }
            }
            const uint32_t failmask = __ballot_sync(0xffffffffu, failG != 0);
            const uint32_t goodmask = __ballot_sync(0xffffffffu, walked_me && good_me);
            int F = failmask ? __ffs(failmask) - 1 : gn;          // first point that does not commit
            const int G = goodmask ? __ffs(goodmask) - 1 : 32;    // first accepted segment
            const bool has_good = G < F;
            if (has_good) F = G + 1;                              // the counts later points saw are stale
            const uint32_t committed = F >= 32 ? 0xffffffffu : (1u << F) - 1u;
            const uint32_t commit_walked = walked_mask & committed;
            pendA &= ~__reduce_or_sync(0xffffffffu, lane < F && my_i < 32 ? 1u << my_i : 0u);
            pendB &= ~__reduce_or_sync(0xffffffffu, lane < F && my_i >= 32 ? 1u << (my_i - 32) : 0u);
            pendA &= ~__ballot_sync(0xffffffffu, (deadA & commit_walked) != 0);
            pendB &= ~__ballot_sync(0xffffffffu, (deadB & commit_walked) != 0);
            if (has_good) {
                // un-vote the live pixels of [-e1, e0] of segment G; theta n stays with its thread
                const int4 *sp = sSeg[warp][G];
                const int4 s0 = sp[0], s1 = sp[1];
                const int glen = s1.w;
                for (int c0 = 0; c0 < glen; c0 += 2 * PPHT_T) {
#pragma unroll
                    for (int u = 0; u < 2; u++) {
                        const int idx = c0 + tid + PPHT_T * u;
                        uint32_t e = 0xffffffffu;
                        if (idx < glen) {
                            int j1, i1;
                            walk_pixel(s0.x, s0.y, s0.z, s0.w, s1.x, idx - s1.z, &j1, &i1);
                            if (__ldcg(A.mask + (size_t)i1 * A.w + j1) != 0) e = (uint32_t)j1 | ((uint32_t)i1 << 16);
                        }
                        sList[tid + PPHT_T * u] = e;
                    }
                    __syncthreads();
                    const int cn = min(2 * PPHT_T, glen - c0);
                    for (int i = 0; i < cn; i++) {
                        const uint32_t e = sList[i];
                        if (e == 0xffffffffu) continue;
                        const float fx = (float)(e & 0xffffu), fy = (float)(e >> 16);
                        if (fast) {
                            if (voter) atomicSub(A.accum + (size_t)n0 * A.numrho + vote_r(fx, fy, tc0, ts0, half), 1);
                        } else {
                            for (int n = n0; n < A.numangle; n += PPHT_C * PPHT_T)
                                atomicSub(A.accum + (size_t)n * A.numrho + vote_r(fx, fy, A.ttab[2 * n], A.ttab[2 * n + 1], half), 1);
                        }
                    }
                    __syncthreads();
                }
                cluster.sync();     // no CTA may erase while another still reads the segment's pixels
                const int lx0 = __shfl_sync(0xffffffffu, ax, G), ly0 = __shfl_sync(0xffffffffu, ay, G);
                const int lx1 = __shfl_sync(0xffffffffu, bx, G), ly1 = __shfl_sync(0xffffffffu, by, G);
                if (cta == 0 && tid == 0 && nlines < A.cap) {
                    A.lines[4 * nlines + 0] = lx0;
                    A.lines[4 * nlines + 1] = ly0;
                    A.lines[4 * nlines + 2] = lx1;
                    A.lines[4 * nlines + 3] = ly1;
                }
                nlines++;
            }
            // erase the committed ranges (cv clears their live pixels; the others are zero already).
            // Every CTA stores the same zeros, so each one's later loads are ordered behind its own stores.
            for (uint32_t rest = commit_walked; rest; rest &= rest - 1) {
                const int c = __ffs(rest) - 1;
                const int4 *sp = sSeg[warp][c];
                const int4 s1 = sp[1];
                if (s1.w <= 32 && (c % PPHT_W) != warp) continue;    // short ranges: one warp each
                const int4 s0 = sp[0];
                const int first = s1.w <= 32 ? lane : tid, step = s1.w <= 32 ? 32 : PPHT_T;
                for (int idx = first; idx < s1.w; idx += step) {
                    int j1, i1;
                    walk_pixel(s0.x, s0.y, s0.z, s0.w, s1.x, idx - s1.z, &j1, &i1);
                    __stcg(A.mask + (size_t)i1 * A.w + j1, (uint8_t)0);
                }
            }
            // ---- roll back the votes of the points that did not commit (same thread, same bins)
            for (int q = F; q < gn; q++) {
                const int x = __shfl_sync(0xffffffffu, my_x, q), y = __shfl_sync(0xffffffffu, my_y, q);
                if (fast) {
                    if (voter) atomicSub(A.accum + (size_t)n0 * A.numrho + vote_r((float)x, (float)y, tc0, ts0, half), 1);
                } else {
                    for (int n = n0; n < A.numangle; n += PPHT_C * PPHT_T)
                        atomicSub(A.accum + (size_t)n * A.numrho + vote_r((float)x, (float)y, A.ttab[2 * n], A.ttab[2 * n + 1], half), 1);
                }
            }
            parity ^= 1;
        }
        // bar.sync orders this window's erasing stores (st.cg) before the next window's liveness loads (ld.cg)
        __syncthreads();
    }
    if (cta == 0 && tid == 0) *A.n_lines = nlines;
}

// The visiting order of cv::HoughLinesProbabilistic for n points: count = n..1,
// idx = rng % count, visit arr[idx], arr[idx] = arr[count-1]; cv::RNG(-1) is the
// multiply-with-carry generator state = lo32 * 4164903690 + hi32.
static void ppht_order(int n, int *order, int *arr)
{
    for (int i = 0; i < n; i++) arr[i] = i;
    uint64_t state = (uint64_t)-1;
    int k = 0;
    for (int count = n; count > 0; count--) {
        state = (uint64_t)(uint32_t)state * 4164903690u + (uint32_t)(state >> 32);
        const int idx = (int)((uint32_t)state % (uint32_t)count);
        order[k++] = arr[idx];
        arr[idx] = arr[count - 1];
    }
}

// cv::HoughLinesP for a BATCH of equally sized edge maps, all images in flight at once: per image the ordered
// edge compaction, then ONE host round trip for the point counts (the visiting orders are expanded on host
// threads meanwhile -- they depend on the counts only), one gather per image and ONE launch of ppht_kernel
// with a cluster per image.  lines_out[i] receives image i's segments (x0, y0, x1, y1 each).
static int32_t hough_lines_p_batch_once(oics_ctx *ctx, const uint8_t *const *dedges, int n_img, int w, int h, size_t stride,
                                        double rho, double theta, int threshold, int line_length, int line_gap, int cap,
                                        std::vector<std::vector<int32_t>> *lines_out, int *max_found)
{
    cudaStream_t st = ctx->stream;
    // OICS_HOUGH_TRACE=1: host-side wall times of the phases, on stderr (diagnostics)
    static const bool trace = getenv("OICS_HOUGH_TRACE") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = trace ? now() : 0.0;
    double t1 = 0.0, t2 = 0.0;
    *max_found = 0;
    lines_out->assign((size_t)n_img, std::vector<int32_t>());
    if (n_img <= 0) return OICS_OK;
    if (w > 32767 || h > 32767) {
        set_error("hough: image larger than 32767 px per side is not supported");
        return OICS_ERR_UNSUPPORTED;
    }
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
    // ---- per-image scratch (hough_img): [total 16 B][rowoff h ints][pts w*h u32][mask w*h]; accum separately
    const size_t npx = (size_t)w * h;
    const size_t off_pts = (16 + sizeof(int) * (size_t)h + 15) & ~(size_t)15;
    const size_t off_mask = off_pts + sizeof(uint32_t) * npx;
    const size_t img_bytes = (off_mask + npx + 255) & ~(size_t)255;
    const size_t acc_bytes = ((size_t)numangle * numrho * sizeof(int32_t) + 255) & ~(size_t)255;
    OICS_TRY(ctx->labels.ensure((size_t)n_img * img_bytes));
    OICS_TRY(ctx->accum.ensure((size_t)n_img * acc_bytes));
    uint8_t *ibase = ctx->labels.as<uint8_t>();
    for (int i = 0; i < n_img; i++) {
        int *total = reinterpret_cast<int *>(ibase + (size_t)i * img_bytes);
        int *rowoff = total + 4;
        uint32_t *pts = reinterpret_cast<uint32_t *>(ibase + (size_t)i * img_bytes + off_pts);
        uint8_t *mask = ibase + (size_t)i * img_bytes + off_mask;
        edge_row_count_kernel<<<cdiv(h, 8), 256, 0, st>>>(dedges[i], w, h, stride, rowoff);
        edge_row_scan_kernel<<<1, 1024, 0, st>>>(rowoff, h, total);
        edge_row_write_kernel<<<cdiv(h, 8), 256, 0, st>>>(dedges[i], w, h, stride, rowoff, pts, mask);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 3 * n_img;
    // the point counts: one read-back (strided: the first int of every image block)
    OICS_TRY(ctx->pin_small.ensure(sizeof(int) * (size_t)n_img));
    OICS_CUDA_TRY(cudaMemcpy2DAsync(ctx->pin_small.p, sizeof(int), ibase, img_bytes, sizeof(int), (size_t)n_img, cudaMemcpyDeviceToHost, st));
    OICS_CUDA_TRY(cudaStreamSynchronize(st));
    std::vector<int> npts((size_t)n_img);
    std::vector<size_t> off((size_t)n_img + 1, 0);
    for (int i = 0; i < n_img; i++) {
        npts[(size_t)i] = ctx->pin_small.as<int>()[i];
        off[(size_t)i + 1] = off[(size_t)i] + (size_t)npts[(size_t)i];
    }
    const size_t total_pts = off[(size_t)n_img];
    if (total_pts == 0) return OICS_OK;
    if (trace) t1 = now();

    // ---- visiting orders, expanded in pinned memory on host threads (data-independent, like the trig table)
    OICS_TRY(ctx->pin_in.ensure(2 * sizeof(int) * total_pts));
    int *order = ctx->pin_in.as<int>();
    {
        const int nthr = std::max(1, std::min<int>(n_img, (int)std::thread::hardware_concurrency()));
        std::atomic<int> next(0);
        auto work = [&]() {
            for (int i = next.fetch_add(1); i < n_img; i = next.fetch_add(1))
                if (npts[(size_t)i] > 0) ppht_order(npts[(size_t)i], order + off[(size_t)i], order + total_pts + off[(size_t)i]);
        };
        if (nthr == 1) {
            work();
        } else {
            std::vector<std::thread> pool;
            for (int t = 0; t < nthr; t++) pool.emplace_back(work);
            for (std::thread &t : pool) t.join();
        }
    }
    if (trace) t2 = now();
    // misc: [order total_pts][seq total_pts][ttab][dirs][PphtArgs n][n_lines n x 16 B][lines n x cap x 4]
    const size_t off_seq = sizeof(int) * total_pts;
    const size_t off_ttab = (off_seq + sizeof(uint32_t) * total_pts + 15) & ~(size_t)15;
    const size_t off_dirs = (off_ttab + sizeof(float) * ttab.size() + 15) & ~(size_t)15;   // int4 loads
    const size_t off_args = (off_dirs + sizeof(int) * dirs.size() + 15) & ~(size_t)15;
    const size_t off_nl = (off_args + sizeof(PphtArgs) * (size_t)n_img + 15) & ~(size_t)15;
    const size_t off_lines = off_nl + 16 * (size_t)n_img;
    const size_t misc_bytes = off_lines + sizeof(int32_t) * 4 * (size_t)cap * n_img;
    OICS_TRY(ctx->misc.ensure(misc_bytes));
    uint8_t *base = ctx->misc.as<uint8_t>();
    OICS_CUDA_TRY(cudaMemcpyAsync(base, order, sizeof(int) * total_pts, cudaMemcpyHostToDevice, st));
    OICS_CUDA_TRY(cudaMemcpyAsync(base + off_ttab, ttab.data(), sizeof(float) * ttab.size(), cudaMemcpyHostToDevice, st));
    OICS_CUDA_TRY(cudaMemcpyAsync(base + off_dirs, dirs.data(), sizeof(int) * dirs.size(), cudaMemcpyHostToDevice, st));
    OICS_CUDA_TRY(cudaMemsetAsync(ctx->accum.p, 0, (size_t)n_img * acc_bytes, st));
    OICS_CUDA_TRY(cudaMemsetAsync(base + off_nl, 0, 16 * (size_t)n_img, st));
    std::vector<PphtArgs> args((size_t)n_img);
    for (int i = 0; i < n_img; i++) {
        PphtArgs &A = args[(size_t)i];
        const uint32_t *pts = reinterpret_cast<const uint32_t *>(ibase + (size_t)i * img_bytes + off_pts);
        uint32_t *seq = reinterpret_cast<uint32_t *>(base + off_seq) + off[(size_t)i];
        if (npts[(size_t)i] > 0)
            gather_points_kernel<<<cdiv(npts[(size_t)i], 256), 256, 0, st>>>(pts, reinterpret_cast<const int *>(base) + off[(size_t)i],
                                                                            npts[(size_t)i], seq);
        A.seq = seq;
        A.npts = npts[(size_t)i];
        A.mask = ibase + (size_t)i * img_bytes + off_mask;
        A.w = w; A.h = h;
        A.accum = reinterpret_cast<int32_t *>(ctx->accum.as<uint8_t>() + (size_t)i * acc_bytes);
        A.ttab = reinterpret_cast<const float *>(base + off_ttab);
        A.dirs = reinterpret_cast<const int *>(base + off_dirs);
        A.numangle = numangle; A.numrho = numrho;
        A.threshold = threshold; A.line_length = line_length; A.line_gap = line_gap;
        A.lines = reinterpret_cast<int32_t *>(base + off_lines) + 4 * (size_t)cap * i;
        A.cap = cap;
        A.n_lines = reinterpret_cast<int *>(base + off_nl + 16 * (size_t)i);
        A.list = nullptr;
    }
    OICS_CUDA_TRY(cudaGetLastError());
    OICS_CUDA_TRY(cudaMemcpyAsync(base + off_args, args.data(), sizeof(PphtArgs) * (size_t)n_img, cudaMemcpyHostToDevice, st));
    {
        ProfScope ps(ctx, "ppht_kernel", 0.0);
        ppht_kernel<<<PPHT_C * n_img, PPHT_T, 0, st>>>(reinterpret_cast<const PphtArgs *>(base + off_args));   // a cluster per image
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 1 + n_img;
    // results: counts and segment lists in one read-back each
    std::vector<int> nl(4 * (size_t)n_img);
    std::vector<int32_t> all(4 * (size_t)cap * n_img);
    OICS_CUDA_TRY(cudaMemcpyAsync(nl.data(), base + off_nl, 16 * (size_t)n_img, cudaMemcpyDeviceToHost, st));
    OICS_CUDA_TRY(cudaMemcpyAsync(all.data(), base + off_lines, sizeof(int32_t) * all.size(), cudaMemcpyDeviceToHost, st));
    OICS_CUDA_TRY(cudaStreamSynchronize(st));   // also: args / ttab / dirs (host temporaries) have been consumed
    if (trace)
        fprintf(stderr, "[oics hough] %d images, %zu edge points: edge lists (waits for Canny) %.2f ms | visiting orders (host) %.2f | "
                        "gather + PPHT + read-back %.2f\n", n_img, total_pts, t1 - t0, t2 - t1, now() - t2);
    for (int i = 0; i < n_img; i++) {
        const int n = nl[4 * (size_t)i];
        *max_found = std::max(*max_found, n);
        const int keep = std::min(n, cap);
        (*lines_out)[(size_t)i].assign(all.begin() + 4 * (size_t)cap * i, all.begin() + 4 * (size_t)cap * i + 4 * (size_t)keep);
    }
    return OICS_OK;
}

int32_t hough_lines_p_batch_device(oics_ctx *ctx, const uint8_t *const *dedges, int n_img, int w, int h, size_t stride,
                                   double rho, double theta, int threshold, int line_length, int line_gap,
                                   std::vector<std::vector<int32_t>> *lines_out)
{
    // 4096 segments per image fit the device list; the count the kernel reports is exact, so a longer list
    // (threshold 0 on noise) re-runs the batch once with room for it -- the edge maps are untouched by a run
    int cap = 4096, found = 0;
    OICS_TRY(hough_lines_p_batch_once(ctx, dedges, n_img, w, h, stride, rho, theta, threshold, line_length, line_gap, cap,
                                      lines_out, &found));
    if (found > cap)
        OICS_TRY(hough_lines_p_batch_once(ctx, dedges, n_img, w, h, stride, rho, theta, threshold, line_length, line_gap, found,
                                          lines_out, &found));
    return OICS_OK;
}

int32_t hough_lines_p_device(oics_ctx *ctx, const uint8_t *dedges, int w, int h, size_t stride, double rho,
                             double theta, int threshold, int line_length, int line_gap, int32_t *h_lines,
                             int cap, int *n_lines)
{
    std::vector<std::vector<int32_t>> out;
    OICS_TRY(hough_lines_p_batch_device(ctx, &dedges, 1, w, h, stride, rho, theta, threshold, line_length, line_gap, &out));
    const int n = (int)(out[0].size() / 4);
    *n_lines = n;
    const int ncopy = std::min(n, cap);
    if (ncopy > 0) {
        if (is_device_ptr(h_lines))
            OICS_CUDA_TRY(cudaMemcpy(h_lines, out[0].data(), sizeof(int32_t) * 4 * (size_t)ncopy, cudaMemcpyHostToDevice));
        else
            memcpy(h_lines, out[0].data(), sizeof(int32_t) * 4 * (size_t)ncopy);
    }
    return OICS_OK;
}

}  // namespace oics
