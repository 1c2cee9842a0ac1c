// H1: cv::Canny(src, lo, hi, apertureSize = 3, L2gradient = false) as the reference
// calls it at hough.rs:27 (gray), omr.rs:239 (3-channel BGR) and fft.rs:164-171.
// Integer-exact restatement:
//   Sobel 3x3, BORDER_REPLICATE, int16; L1 magnitude; for 3 channels the FIRST channel
//   with the maximal magnitude per pixel; magnitude image zero-padded by one pixel;
//   TG22 = 13573 fixed-point sector test; candidate iff m > low and the directional
//   maximum test passes; strong iff candidate and m > high; output 255 on every
//   candidate 8-connected (through candidates) to a strong pixel.
//
// Two kernels:
//   canny_nms_kernel   one pass over the source: the tile and a 2-pixel halo are staged
//                      in shared memory, Sobel + magnitude are evaluated once per pixel
//                      of tile + 1 halo, the non-maximum test runs on the shared
//                      magnitudes.  Writes the state map (0 weak candidate, 1 no edge,
//                      2 edge) and appends strong pixels to the frontier queue.
//   canny_hysteresis_kernel  cooperative, persistent (one wave of CTAs): breadth-first
//                      propagation over the frontier queue with a grid barrier per
//                      level; each weak pixel is claimed with one atomicOr, so the
//                      result is order-independent (= the CPU stack walk) and every
//                      pixel enters the queue at most once.  Ends by writing the 0/255
//                      edge map.
#include <cooperative_groups.h>

#include "common.h"

namespace cg = cooperative_groups;

namespace oics {

static constexpr int CN_TX = 64, CN_TY = 16;  // output tile
static constexpr int CN_SW = CN_TX + 4, CN_SH = CN_TY + 4;  // source tile (halo 2)
static constexpr int CN_MW = CN_TX + 2, CN_MH = CN_TY + 2;  // magnitude tile (halo 1)

template <int CH>
__global__ void __launch_bounds__(256) canny_nms_kernel(const uint8_t *__restrict__ src, int w, int h, size_t stride,
                                                        int low, int high, uint8_t *__restrict__ map,
                                                        int *__restrict__ queue, int *__restrict__ qtail)
{
    __shared__ uint8_t s[CN_SH][CN_SW * CH + 4];      // interleaved channels, as in the source
    __shared__ uint16_t smag[CN_MH][CN_MW];
    __shared__ short2 sgrad[CN_MH][CN_MW];
    __shared__ int s_cnt, s_base;
    __shared__ int s_strong[CN_TX * CN_TY];

    const int bx = blockIdx.x * CN_TX, by = blockIdx.y * CN_TY;
    if (threadIdx.x == 0) s_cnt = 0;
    for (int i = threadIdx.x; i < CN_SW * CN_SH; i += blockDim.x) {
        const int ly = i / CN_SW, lx = i - ly * CN_SW;
        const int gx = min(max(bx + lx - 2, 0), w - 1), gy = min(max(by + ly - 2, 0), h - 1);  // BORDER_REPLICATE
        const uint8_t *p = src + (size_t)gy * stride + (size_t)gx * CH;
#pragma unroll
        for (int c = 0; c < CH; c++) s[ly][lx * CH + c] = p[c];
    }
    __syncthreads();
    // Sobel + L1 magnitude, the channel with the largest magnitude wins (first on ties).  A thread
    // walks CN_SEG consecutive rows of one column with a three-row window of the horizontal
    // difference d = right - left and the horizontal smooth sm = left + 2 mid + right, so a
    // magnitude costs (CN_SEG + 2) / CN_SEG row reads of three bytes per channel instead of eight taps.
    constexpr int CN_SEG = CN_MH / 3;                 // 18 rows = 3 segments of 6
    static_assert(CN_SEG * 3 == CN_MH && CN_MW * 3 <= 256, "canny tile / thread mapping");
    if (threadIdx.x < CN_MW * 3) {
        const int seg = threadIdx.x / CN_MW, lx = threadIdx.x - seg * CN_MW, ly0 = seg * CN_SEG;
        const int gx = bx + lx - 1;
        int bm[CN_SEG], bdx[CN_SEG], bdy[CN_SEG];
#pragma unroll
        for (int r = 0; r < CN_SEG; r++) { bm[r] = -1; bdx[r] = 0; bdy[r] = 0; }
#pragma unroll
        for (int c = 0; c < CH; c++) {
            int d[CN_SEG + 2], sm[CN_SEG + 2];
#pragma unroll
            for (int r = 0; r < CN_SEG + 2; r++) {
                // source tile coordinates of magnitude (lx, ly) are (lx + 1, ly + 1): rows ly .. ly + 2
                const uint8_t *row = &s[ly0 + r][lx * CH + c];
                const int l = row[0], m = row[CH], rr = row[2 * CH];
                d[r] = rr - l;
                sm[r] = l + 2 * m + rr;
            }
#pragma unroll
            for (int r = 0; r < CN_SEG; r++) {
                const int gxv = d[r] + 2 * d[r + 1] + d[r + 2];
                const int gyv = sm[r + 2] - sm[r];
                const int m = abs(gxv) + abs(gyv);
                if (m > bm[r]) { bm[r] = m; bdx[r] = gxv; bdy[r] = gyv; }
            }
        }
#pragma unroll
        for (int r = 0; r < CN_SEG; r++) {
            const int gy = by + ly0 + r - 1;
            const bool in = gx >= 0 && gx < w && gy >= 0 && gy < h;
            // outside the image the magnitude is the zero padding
            smag[ly0 + r][lx] = (uint16_t)(in ? bm[r] : 0);
            sgrad[ly0 + r][lx] = in ? make_short2((short)bdx[r], (short)bdy[r]) : make_short2(0, 0);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < CN_TX * CN_TY; i += blockDim.x) {
        const int ly = i / CN_TX, lx = i - ly * CN_TX;
        const int gx = bx + lx, gy = by + ly;
        if (gx >= w || gy >= h) continue;
        const int my = ly + 1, mx = lx + 1;
        const int m = smag[my][mx];
        bool keep = false;
        if (m > low) {
            const short2 g = sgrad[my][mx];
            const int xs = g.x, ys = g.y;
            const int ax = abs(xs), ay = abs(ys) << 15;
            const int tg22x = ax * 13573;
            if (ay < tg22x) {
                keep = m > smag[my][mx - 1] && m >= smag[my][mx + 1];
            } else {
                const int tg67x = tg22x + (ax << 16);
                if (ay > tg67x) {
                    keep = m > smag[my - 1][mx] && m >= smag[my + 1][mx];
                } else {
                    const int sgn = (xs ^ ys) < 0 ? -1 : 1;
                    keep = m > smag[my - 1][mx - sgn] && m > smag[my + 1][mx + sgn];
                }
            }
        }
        const size_t gi = (size_t)gy * w + gx;
        uint8_t st = 1;
        if (keep) {
            if (m > high) {
                st = 2;
                s_strong[atomicAdd(&s_cnt, 1)] = (int)gi;
            } else
                st = 0;
        }
        map[gi] = st;
    }
    __syncthreads();
    const int n = s_cnt;
    if (n == 0) return;
    if (threadIdx.x == 0) s_base = atomicAdd(qtail, n);
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) queue[s_base + i] = s_strong[i];
}

__global__ void __launch_bounds__(256) canny_hysteresis_kernel(uint8_t *__restrict__ map, int w, int h,
                                                               int *__restrict__ queue, int *qtail,
                                                               uint8_t *__restrict__ edges)
{
    cg::grid_group grid = cg::this_grid();
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
    uint32_t *map32 = reinterpret_cast<uint32_t *>(map);
    int start = 0;
    int end = *reinterpret_cast<volatile int *>(qtail);
    grid.sync();  // everyone has read the level-0 end before anyone pushes
    while (end > start) {
        for (int i = start + tid; i < end; i += nthreads) {
            const int p = queue[i];
            const int y = p / w, x = p - y * w;
#pragma unroll
            for (int dy = -1; dy <= 1; dy++) {
                const int yy = y + dy;
                if (yy < 0 || yy >= h) continue;
#pragma unroll
                for (int dx = -1; dx <= 1; dx++) {
                    const int xx = x + dx;
                    if ((dx | dy) == 0 || xx < 0 || xx >= w) continue;
                    const int j = yy * w + xx;
                    if (map[j] != 0) continue;
                    const int sh = (j & 3) * 8;
                    const uint32_t old = atomicOr(map32 + (j >> 2), 2u << sh);
                    if (((old >> sh) & 0xffu) == 0) queue[atomicAdd(qtail, 1)] = j;
                }
            }
        }
        __threadfence();
        grid.sync();
        start = end;
        end = *reinterpret_cast<volatile int *>(qtail);
        grid.sync();
    }
    const int n = w * h;
    for (int i = tid; i < n; i += nthreads) edges[i] = map[i] == 2 ? 255 : 0;
}

int32_t canny_device(oics_ctx *ctx, const uint8_t *dsrc, int ch, int w, int h, size_t stride, int low, int high,
                     uint8_t *dedges)
{
    cudaStream_t st = ctx->stream;
    if (ch != 1 && ch != 3) {
        set_error("canny: %d channels not supported (1 or 3)", ch);
        return OICS_ERR_UNSUPPORTED;
    }
    if ((long long)w * h > 0x7fffffffLL) {
        set_error("canny: image too large");
        return OICS_ERR_UNSUPPORTED;
    }
    const size_t n = (size_t)w * h;
    // labels: [qtail + pad (16 B)][queue n ints]; mag: state map, 4-byte aligned and padded
    OICS_TRY(ctx->labels.ensure(16 + n * sizeof(int)));
    OICS_TRY(ctx->mag.ensure(n + 16));
    int *qtail = ctx->labels.as<int>();
    int *queue = qtail + 4;
    uint8_t *map = ctx->mag.as<uint8_t>();
    OICS_CUDA_TRY(cudaMemsetAsync(qtail, 0, 16, st));
    dim3 grid(cdiv(w, CN_TX), cdiv(h, CN_TY));
    {
        ProfScope ps(ctx, ch == 1 ? "canny_nms_kernel<1>" : "canny_nms_kernel<3>", (double)n * (ch + 1));
        if (ch == 1)
            canny_nms_kernel<1><<<grid, 256, 0, st>>>(dsrc, w, h, stride, low, high, map, queue, qtail);
        else
            canny_nms_kernel<3><<<grid, 256, 0, st>>>(dsrc, w, h, stride, low, high, map, queue, qtail);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    int &blocks_per_sm = ctx->canny_blocks_per_sm;   // per context = per device, under ctx->mu
    if (!blocks_per_sm) {
        OICS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, canny_hysteresis_kernel, 256, 0));
        if (blocks_per_sm < 1) {
            set_error("canny: hysteresis kernel cannot be made resident");
            return OICS_ERR_CUDA;
        }
        if (blocks_per_sm > 4) blocks_per_sm = 4;
    }
    int nblocks = ctx->sm_count * blocks_per_sm;
    void *args[] = {(void *)&map, (void *)&w, (void *)&h, (void *)&queue, (void *)&qtail, (void *)&dedges};
    {
        ProfScope ps(ctx, "canny_hysteresis_kernel", (double)n * 2);
        OICS_CUDA_TRY(cudaLaunchCooperativeKernel((const void *)canny_hysteresis_kernel, dim3(nblocks), dim3(256), args, 0, st));
    }
    ctx->launches += 2;
    return OICS_OK;
}

}  // namespace oics
