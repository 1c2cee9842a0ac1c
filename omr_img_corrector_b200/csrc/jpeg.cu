// N2 (output side): TransformableMatrix::im_write(path, JPEG, 100) (transfer.rs:162-179,
// omr.rs:447-448) = cv::imwrite = libjpeg's baseline encoder with OpenCV's defaults (JFIF, 4:2:0,
// standard Huffman tables, one scan), byte for byte, on the GPU -- so a deskewed card leaves the
// device as the few-MB file the reference writes instead of 30 MB of pixels.
//
//   jpeg_dct_kernel    8 threads per 8x8 block (a thread = a row, then a column): BGR -> Y / Cb / Cr
//                      (jccolor.c fixed point), h2v2 chroma box average with the alternating bias
//                      (jcsample.c), edge replication (jcprepct.c), jfdctint.c forward DCT and the
//                      jcdctmgr.c quantiser; coefficients are stored zigzag-major [k][block] so the
//                      entropy kernels read them coalesced (lane = block).
//   jpeg_bits_kernel   one thread per block: resolves the dummy blocks of the right / bottom edge
//                      (jccoefct.c: AC zero, DC of the previous block), the DC prediction of its
//                      component, and counts the bits of jchuff.c's encode_one_block.
//   scan               exclusive prefix sum of the bit counts -> bit offset of every block.
//   jpeg_emit_kernel   one thread per block: the same walk again, now OR-ing the codes into the
//                      (zeroed) unstuffed bit stream at the block's offset.
//   jpeg_stuff_*       0xFF -> 0xFF 0x00: count per 32-byte chunk, scan, scatter behind the header.
// The header (markers, quantisation and Huffman tables) is built on the host.
#include <algorithm>
#include <mutex>

#include "common.h"

namespace oics {

__constant__ uint8_t c_izz[64];        // natural index -> zigzag position
__constant__ uint16_t c_hcode[4][256]; // DC luma, AC luma, DC chroma, AC chroma
__constant__ uint8_t c_hlen[4][256];

static const uint8_t h_zz[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                                 41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
static const uint8_t h_ql[64] = {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57,
                                 69, 56, 14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64,
                                 81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
static const uint8_t h_qc[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99,
                                 99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                                 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
static const uint8_t h_dc_l_bits[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
static const uint8_t h_dc_c_bits[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
static const uint8_t h_dc_vals[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
static const uint8_t h_ac_l_bits[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
static const uint8_t h_ac_l_vals[162] = {
    0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71, 0x14, 0x32,
    0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0a, 0x16,
    0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45,
    0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69,
    0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94,
    0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6,
    0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8,
    0xd9, 0xda, 0xe1, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
    0xf9, 0xfa};
static const uint8_t h_ac_c_bits[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
static const uint8_t h_ac_c_vals[162] = {
    0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22, 0x32, 0x81,
    0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16, 0x24, 0x34,
    0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44,
    0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68,
    0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92,
    0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4,
    0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6,
    0xd7, 0xd8, 0xd9, 0xda, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
    0xf9, 0xfa};

struct JpegGeom {
    int w, h, mcux, mcuy, wibY, hibY, nblocks;   // nblocks = 6 * mcux * mcuy, MCU order: Y00 Y01 Y10 Y11 Cb Cr
    size_t cstride;                                // blocks per coefficient plane (padded)
};

#define JFIX13(x) ((int)((x) * 8192.0 + 0.5))
#define JFIX16(x) ((int)((x) * 65536.0 + 0.5))

// jfdctint.c, one row (first pass) or column (second pass)
__device__ __forceinline__ void fdct_1d(const int (&x)[8], int (&o)[8], bool first)
{
    const int t0 = x[0] + x[7], t7 = x[0] - x[7], t1 = x[1] + x[6], t6 = x[1] - x[6];
    const int t2 = x[2] + x[5], t5 = x[2] - x[5], t3 = x[3] + x[4], t4 = x[3] - x[4];
    const int t10 = t0 + t3, t13 = t0 - t3, t11 = t1 + t2, t12 = t1 - t2;
    const int sh = first ? 13 - 2 : 13 + 2;
    const int rnd = 1 << (sh - 1);
    if (first) {
        o[0] = (t10 + t11) * 4;
        o[4] = (t10 - t11) * 4;
    } else {
        o[0] = (t10 + t11 + 2) >> 2;
        o[4] = (t10 - t11 + 2) >> 2;
    }
    int z1 = (t12 + t13) * JFIX13(0.541196100);
    o[2] = (z1 + t13 * JFIX13(0.765366865) + rnd) >> sh;
    o[6] = (z1 + t12 * (-JFIX13(1.847759065)) + rnd) >> sh;
    z1 = t4 + t7;
    int z2 = t5 + t6, z3 = t4 + t6, z4 = t5 + t7;
    const int z5 = (z3 + z4) * JFIX13(1.175875602);
    const int u4 = t4 * JFIX13(0.298631336), u5 = t5 * JFIX13(2.053119869);
    const int u6 = t6 * JFIX13(3.072711026), u7 = t7 * JFIX13(1.501321110);
    z1 *= -JFIX13(0.899976223);
    z2 *= -JFIX13(2.562915447);
    z3 = z3 * (-JFIX13(1.961570560)) + z5;
    z4 = z4 * (-JFIX13(0.390180644)) + z5;
    o[7] = (u4 + z1 + z3 + rnd) >> sh;
    o[5] = (u5 + z2 + z4 + rnd) >> sh;
    o[3] = (u6 + z2 + z3 + rnd) >> sh;
    o[1] = (u7 + z1 + z4 + rnd) >> sh;
}

// One CTA = JD_MCU consecutive MCUs of one MCU row (16 rows x 128 pixels), 8 threads per 8x8 block.
//   1. the 16 source rows are staged with aligned 32-bit loads (every pixel is read from HBM once);
//   2. one thread per 2x2 pixel quad: four Y samples and the box-averaged Cb / Cr sample -> int8 planes;
//   3. thread = a row of its block, then a column (jfdctint.c), quantiser, zigzag;
//   4. the CTA's coefficients leave as 32-bit stores, 96 contiguous bytes per zigzag plane.
#define JD_MCU 8
#define JD_BLOCKS (JD_MCU * 6)
#define JD_T (JD_BLOCKS * 8)
#define JD_ROWW (JD_MCU * 12 + 1)   // words per staged row: 384 bytes + up to 3 bytes of misalignment

struct JpegQuant {
    uint8_t q[2][64];     // natural order: luma, chroma
    float rcp[2][64];     // 1 / (8 q)
};

__global__ void __launch_bounds__(JD_T) jpeg_dct_kernel(const uint8_t *__restrict__ bgr, size_t stride, JpegGeom G,
                                                        JpegQuant Q, int strips, int16_t *__restrict__ coef)
{
    __shared__ uint32_t sraw[16][JD_ROWW];
    __shared__ int smis[16];
    // row pitches chosen so that the 8 rows a block's threads read fall into different banks
    __shared__ __align__(8) int8_t sY[16][JD_MCU * 16 + 16];
    __shared__ __align__(8) int8_t sC[2][8][JD_MCU * 8 + 8];
    __shared__ int sT[JD_BLOCKS][8][9];
    __shared__ __align__(4) int16_t sO[64][JD_BLOCKS + 2];
    __shared__ float sRcp[2][64];
    __shared__ uint16_t sD[2][64];
    __shared__ uint8_t sIzz[64];
    const int tid = threadIdx.x;
    if (tid < 128) {
        sRcp[tid >> 6][tid & 63] = Q.rcp[tid >> 6][tid & 63];
        sD[tid >> 6][tid & 63] = (uint16_t)((int)Q.q[tid >> 6][tid & 63] << 3);
        if (tid < 64) sIzz[tid] = c_izz[tid];
    }
    const int my = blockIdx.x / strips, mx0 = (blockIdx.x - my * strips) * JD_MCU;
    const int nm = min(JD_MCU, G.mcux - mx0);
    // 1. stage
    {
        const uintptr_t last_word = ((uintptr_t)bgr + (size_t)(G.h - 1) * stride + 3 * (size_t)G.w - 1) & ~(uintptr_t)3;
#pragma unroll
        for (int it = 0; it < (16 * JD_ROWW + JD_T - 1) / JD_T; it++) {
            const int t = tid + it * JD_T;
            if (t >= 16 * JD_ROWW) break;
            const int rr = t / JD_ROWW, j = t - rr * JD_ROWW;
            const uintptr_t a = (uintptr_t)bgr + (size_t)min(my * 16 + rr, G.h - 1) * stride + (size_t)mx0 * 48;
            if (j == 0) smis[rr] = (int)(a & 3);
            sraw[rr][j] = __ldg((const uint32_t *)min((a & ~(uintptr_t)3) + 4 * (uintptr_t)j, last_word));
        }
    }
    __syncthreads();
    // 2. colour conversion (jccolor.c) and h2v2 downsampling (jcsample.c); edges by index clamping
    for (int qd = tid; qd < JD_MCU * 64; qd += JD_T) {
        const int qy = qd / (JD_MCU * 8), qx = qd - qy * (JD_MCU * 8);
        const int lc0 = min(mx0 * 16 + 2 * qx, G.w - 1) - mx0 * 16, lc1 = min(mx0 * 16 + 2 * qx + 1, G.w - 1) - mx0 * 16;
        const int cyy = min(my * 8 + qy, (G.h + 1) / 2 - 1);
        const int ly0 = min(2 * cyy, G.h - 1) - my * 16, ly1 = min(2 * cyy + 1, G.h - 1) - my * 16;
        int cb = 0, cr = 0;
        auto fetch = [&](int ly, int lc, int &R, int &Gc, int &B) {
            const uint8_t *p = (const uint8_t *)sraw[ly] + smis[ly] + 3 * lc;
            B = p[0]; Gc = p[1]; R = p[2];
        };
        auto luma = [](int R, int Gc, int B) {
            return ((JFIX16(0.29900) * R + JFIX16(0.58700) * Gc + JFIX16(0.11400) * B + 32768) >> 16) - 128;
        };
        auto chroma = [&](int R, int Gc, int B) {
            cb += (-JFIX16(0.16874) * R - JFIX16(0.33126) * Gc + JFIX16(0.50000) * B + (128 << 16) + 32767) >> 16;
            cr += (JFIX16(0.50000) * R - JFIX16(0.41869) * Gc - JFIX16(0.08131) * B + (128 << 16) + 32767) >> 16;
        };
        const bool same = ly0 == 2 * qy && ly1 == 2 * qy + 1;   // false only in MCU rows padded past the image
#pragma unroll
        for (int k = 0; k < 4; k++) {
            int R, Gc, B;
            fetch(2 * qy + (k >> 1), (k & 1) ? lc1 : lc0, R, Gc, B);
            sY[2 * qy + (k >> 1)][2 * qx + (k & 1)] = (int8_t)luma(R, Gc, B);
            if (same) chroma(R, Gc, B);
        }
        if (!same) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                int R, Gc, B;
                fetch((k >> 1) ? ly1 : ly0, (k & 1) ? lc1 : lc0, R, Gc, B);
                chroma(R, Gc, B);
            }
        }
        const int bias = (qx & 1) ? 2 : 1;                      // mx0 * 8 is even: the parity of the global column
        sC[0][qy][qx] = (int8_t)(((cb + bias) >> 2) - 128);
        sC[1][qy][qx] = (int8_t)(((cr + bias) >> 2) - 128);
    }
    __syncthreads();
    // 3. forward DCT of block lb = (MCU m, component block b)
    const int lb = tid >> 3, r = tid & 7;
    const int m = lb / 6, b = lb - m * 6;
    int x[8], o[8];
    bool real = m < nm;
    if (b < 4) {
        real = real && my * 2 + (b >> 1) < G.hibY && (mx0 + m) * 2 + (b & 1) < G.wibY;   // else a dummy block
        const int2 v = *(const int2 *)&sY[(b >> 1) * 8 + r][m * 16 + (b & 1) * 8];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            x[i] = (int)(int8_t)(v.x >> (8 * i));
            x[4 + i] = (int)(int8_t)(v.y >> (8 * i));
        }
    } else {
        const int2 v = *(const int2 *)&sC[b - 4][r][m * 8];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            x[i] = (int)(int8_t)(v.x >> (8 * i));
            x[4 + i] = (int)(int8_t)(v.y >> (8 * i));
        }
    }
    fdct_1d(x, o, true);
#pragma unroll
    for (int i = 0; i < 8; i++) sT[lb][r][i] = o[i];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = sT[lb][i][r];      // column r
    fdct_1d(x, o, false);
    const int qt = b < 4 ? 0 : 1;
#pragma unroll
    for (int i = 0; i < 8; i++) {                          // coefficient (row i, column r)
        const int n = 8 * i + r, d = sD[qt][n], v = o[i];
        // (|v| + d / 2) / d by reciprocal, corrected to the exact quotient
        const unsigned a = (unsigned)abs(v) + (unsigned)(d >> 1);
        unsigned qv = (unsigned)((float)a * sRcp[qt][n]);
        int rem = (int)a - (int)qv * d;
        if (rem < 0) { qv--; rem += d; }
        if (rem >= d) qv++;
        sO[sIzz[n]][lb] = real ? (int16_t)(v < 0 ? -(int)qv : (int)qv) : (int16_t)0;
    }
    __syncthreads();
    // 4. zigzag-major planes [k][block]
    const size_t g0 = ((size_t)my * G.mcux + mx0) * 6;
    for (int t = tid; t < 64 * (JD_BLOCKS / 2); t += JD_T) {
        const int k = t / (JD_BLOCKS / 2), j = t - k * (JD_BLOCKS / 2);
        if (2 * j < nm * 6) *(uint32_t *)(coef + (size_t)k * G.cstride + g0 + 2 * j) = ((const uint32_t *)sO[k])[j];
    }
}

// The resolved DC of block (mcu, b): jccoefct.c fills whole blocks past the right / bottom edge of the
// luma plane with zero AC and the DC of the previous block of the MCU buffer.
__device__ __forceinline__ int resolved_dc(const int16_t *__restrict__ coef, const JpegGeom &G, int mcu, int b)
{
    if (b < 4) {
        const int my = mcu / G.mcux, mx = mcu - my * G.mcux;
        const bool row1_dummy = my * 2 + 1 >= G.hibY, col1_dummy = mx * 2 + 1 >= G.wibY;
        if ((b >> 1) && row1_dummy) b = 1;                 // dummy row: the last block of the row above
        if ((b & 1) && col1_dummy) b -= 1;                 // dummy at the right edge: its left neighbour
    }
    return coef[(size_t)mcu * 6 + b];                      // zigzag position 0 plane
}

__device__ __forceinline__ int bit_length(int v) { return 32 - __clz(v); }

// jchuff.c encode_one_block of block g; SINK(code, size) receives every Huffman code and value field
template <typename SINK>
__device__ __forceinline__ void walk_block(const int16_t *__restrict__ coef, const JpegGeom &G, int g, SINK &sink)
{
    const int mcu = g / 6, b = g - mcu * 6;
    bool dummy = false;
    if (b < 4) {
        const int my = mcu / G.mcux, mx = mcu - my * G.mcux;
        dummy = my * 2 + (b >> 1) >= G.hibY || mx * 2 + (b & 1) >= G.wibY;
    }
    // DC prediction: the previous block of the same component in scan order
    int pred = 0;
    if (b < 4) {
        if (b > 0) pred = resolved_dc(coef, G, mcu, b - 1);
        else if (mcu > 0) pred = resolved_dc(coef, G, mcu - 1, 3);
    } else if (mcu > 0) {
        pred = resolved_dc(coef, G, mcu - 1, b);
    }
    const int tdc = b < 4 ? 0 : 2, tac = tdc + 1;
    int t = resolved_dc(coef, G, mcu, b) - pred, t2 = t;
    if (t < 0) { t = -t; t2--; }
    int nb = bit_length(t);
    // a Huffman code (<= 16 bits) and the value bits behind it (<= 11) leave as ONE field
    sink(((unsigned)c_hcode[tdc][nb] << nb) | ((unsigned)t2 & ((1u << nb) - 1u)), c_hlen[tdc][nb] + nb);
    if (dummy) {                                            // all AC zero
        sink(c_hcode[tac][0], c_hlen[tac][0]);
        return;
    }
    int r = 0;
    for (int k0 = 0; k0 < 64; k0 += 8) {
        int vv[8];
#pragma unroll
        for (int j = 0; j < 8; j++) vv[j] = coef[(size_t)(k0 + j) * G.cstride + g];   // eight loads in flight
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (k0 + j == 0) continue;
            int v = vv[j];
            if (v == 0) { r++; continue; }
            while (r > 15) { sink(c_hcode[tac][0xF0], c_hlen[tac][0xF0]); r -= 16; }
            t2 = v;
            if (v < 0) { v = -v; t2--; }
            nb = bit_length(v);
            sink(((unsigned)c_hcode[tac][(r << 4) + nb] << nb) | ((unsigned)t2 & ((1u << nb) - 1u)), c_hlen[tac][(r << 4) + nb] + nb);
            r = 0;
        }
    }
    if (r > 0) sink(c_hcode[tac][0], c_hlen[tac][0]);
}

struct CountSink {
    unsigned n = 0;
    __device__ __forceinline__ void operator()(unsigned, int size) { n += size; }
};

__global__ void __launch_bounds__(256) jpeg_bits_kernel(const int16_t *__restrict__ coef, JpegGeom G,
                                                        unsigned long long *__restrict__ bits)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G.nblocks) return;
    CountSink s;
    walk_block(coef, G, g, s);
    bits[g] = s.n;
}

// MSB-first bit stream in 32-bit words (word i holds stream bits 32 i .. 32 i + 31, bit 31 first)
struct EmitSink {
    uint32_t *words;
    unsigned long long pos;     // next stream bit
    unsigned long long acc = 0; // pending bits, right-aligned
    int n = 0;
    __device__ __forceinline__ void flush_word()
    {
        // the top (32 - pos % 32) pending bits complete word pos / 32
        const int room = 32 - (int)(pos & 31);
        const int take = min(room, n);
        const uint32_t chunk = (uint32_t)((acc >> (n - take)) & ((1ull << take) - 1ull));
        atomicOr(words + (pos >> 5), chunk << (room - take));
        pos += take;
        n -= take;
    }
    __device__ __forceinline__ void operator()(unsigned code, int size)
    {
        if (!size) return;
        acc = (acc << size) | (code & ((1u << size) - 1u));
        n += size;
        while (n >= 32) flush_word();
    }
    __device__ __forceinline__ void finish()
    {
        while (n > 0) flush_word();
    }
};

__global__ void __launch_bounds__(256) jpeg_emit_kernel(const int16_t *__restrict__ coef, JpegGeom G,
                                                        const unsigned long long *__restrict__ offs,
                                                        const unsigned long long *__restrict__ offs_base,
                                                        uint32_t *__restrict__ words)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G.nblocks) return;
    EmitSink s;
    s.words = words;
    s.pos = offs[g] + offs_base[g >> 10];
    walk_block(coef, G, g, s);
    s.finish();
}

// ---------------------------------------------------------------------------
// exclusive prefix sum of n 64-bit values (in place), total to *total: 1024 values per CTA.  With
// `tb` the count is not known on the host: n = the 32-byte chunks of a stream of *tb bits (the grid
// covers the host's bound, CTAs past the real count leave at once).
__device__ __forceinline__ size_t stream_chunks(const unsigned long long *tb) { return (size_t)((((*tb + 7) >> 3) + 31) >> 5); }

__global__ void __launch_bounds__(1024) scan_block_kernel(unsigned long long *__restrict__ v, size_t n,
                                                          unsigned long long *__restrict__ sums,
                                                          const unsigned long long *__restrict__ tb)
{
    __shared__ unsigned long long ws[32];
    if (tb) n = stream_chunks(tb);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // a CTA takes tiles blockIdx.x, + gridDim.x, ...: with `tb` the grid is sized for the host's bound, the real count
    // is a few per cent of it
    for (size_t tile = blockIdx.x; tile * 1024 < n; tile += gridDim.x) {
    const size_t i = tile * 1024 + threadIdx.x;
    const unsigned long long x = i < n ? v[i] : 0ull;
    unsigned long long s = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, s, o);
        if (lane >= o) s += t;
    }
    if (lane == 31) ws[warp] = s;
    __syncthreads();
    if (warp == 0) {
        unsigned long long w = ws[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        ws[lane] = w;
    }
    __syncthreads();
    const unsigned long long before = (warp ? ws[warp - 1] : 0ull) + s - x;
    if (i < n) v[i] = before;
    if (threadIdx.x == 1023) sums[tile] = before + x;
    __syncthreads();                                 // ws is reused by the next tile
    }
}

// one CTA: exclusive scan of the per-CTA sums (any count), total out
__global__ void __launch_bounds__(1024) scan_sums_kernel(unsigned long long *__restrict__ sums, int m,
                                                         unsigned long long *__restrict__ total,
                                                         const unsigned long long *__restrict__ tb)
{
    __shared__ unsigned long long ws[32];
    __shared__ unsigned long long carry;
    if (tb) m = (int)((stream_chunks(tb) + 1023) / 1024);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < m; base += 1024) {
        const int i = base + threadIdx.x;
        const unsigned long long x = i < m ? sums[i] : 0ull;
        unsigned long long s = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += t;
        }
        if (lane == 31) ws[warp] = s;
        __syncthreads();
        if (warp == 0) {
            unsigned long long w = ws[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            ws[lane] = w;
        }
        __syncthreads();
        const unsigned long long before = carry + (warp ? ws[warp - 1] : 0ull) + s - x;
        if (i < m) sums[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}

// Two launches: afterwards the exclusive prefix of element i is  v[i] + sums[i >> 10]  (the consumers add
// the per-CTA base themselves; a third pass over the data would only do that addition).
int32_t exclusive_scan(oics_ctx *ctx, cudaStream_t st, unsigned long long *v, size_t n, unsigned long long *sums,
                       unsigned long long *total, const unsigned long long *tb)
{
    const int m = (int)((n + 1023) / 1024);
    scan_block_kernel<<<tb ? std::min(m, 296) : m, 1024, 0, st>>>(v, n, sums, tb);
    scan_sums_kernel<<<1, 1024, 0, st>>>(sums, m, total, tb);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 2;
    return OICS_OK;
}

// ---------------------------------------------------------------------------
// byte stuffing: stream byte j of the MSB-first words; the last byte is padded with ones
__device__ __forceinline__ uint32_t stream_byte(const uint32_t *__restrict__ words, unsigned long long j,
                                                unsigned long long nbytes, int pad_bits)
{
    uint32_t b = (words[j >> 2] >> (24 - 8 * (int)(j & 3))) & 0xFFu;
    if (j == nbytes - 1) b |= (1u << pad_bits) - 1u;
    return b;
}

// The grids of the kernels below cover the host's bound on the stream (208 bytes per block); the real
// size is read from *total_bits on the device, so that no host round trip sits between the kernels.
__global__ void __launch_bounds__(256) jpeg_zero_kernel(uint4 *__restrict__ words, const unsigned long long *__restrict__ total_bits)
{
    const size_t n = ((*total_bits + 7) >> 3) / 16 + 2;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        words[i] = make_uint4(0u, 0u, 0u, 0u);
}

// One LANE per 32-bit word of the stream, eight lanes per 32-byte chunk (the unit of the offsets' scan).
__device__ __forceinline__ uint32_t stream_word(const uint32_t *__restrict__ words, size_t wi, unsigned long long nbytes, int pad_bits)
{
    // bytes at and behind nbytes are zero in `words` (jpeg_zero_kernel); the last real byte is padded with ones
    uint32_t v = 4 * wi < nbytes ? words[wi] : 0u;
    if ((nbytes - 1) >> 2 == wi) v |= ((1u << pad_bits) - 1u) << (24 - 8 * (int)((nbytes - 1) & 3));
    return v;
}
__device__ __forceinline__ int ff_bytes(uint32_t v, int nvalid)      // 0xFF bytes among the first nvalid (MSB first)
{
    uint32_t m = __vcmpeq4(v, 0xFFFFFFFFu) & 0x01010101u;
    if (nvalid < 4) m &= nvalid <= 0 ? 0u : ~(0xFFFFFFFFu >> (8 * nvalid));
    return __popc(m);
}

__global__ void __launch_bounds__(256) jpeg_stuff_count_kernel(const uint32_t *__restrict__ words,
                                                               const unsigned long long *__restrict__ total_bits,
                                                               unsigned long long *__restrict__ cnt)
{
    const size_t nch = stream_chunks(total_bits);
    const unsigned long long T = *total_bits, nbytes = (T + 7) >> 3;
    const int pad = (int)(8 * nbytes - T);
    const size_t nw = nch * 8;                       // whole groups of eight lanes: the shuffles below are full-warp
    // (the trip count is the same for the 32 lanes of a warp: the loop runs on the warp's first word)
    for (size_t w0 = (size_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31u); w0 < nw; w0 += (size_t)gridDim.x * blockDim.x) {
        const size_t wi = w0 + (threadIdx.x & 31);
        const long long left = (long long)nbytes - 4 * (long long)wi;
        int c = wi < nw ? ff_bytes(stream_word(words, wi, nbytes, pad), (int)min(4ll, max(0ll, left))) : 0;
        c += __shfl_xor_sync(0xffffffffu, c, 1);
        c += __shfl_xor_sync(0xffffffffu, c, 2);
        c += __shfl_xor_sync(0xffffffffu, c, 4);
        if (wi < nw && (wi & 7) == 0) cnt[wi >> 3] = (unsigned long long)c;
    }
}

// header (cached on the device, SOF0 height / width patched in) + stuffed stream + EOI
__global__ void __launch_bounds__(256) jpeg_stuff_write_kernel(const uint32_t *__restrict__ words,
                                                               const unsigned long long *__restrict__ total_bits,
                                                               const unsigned long long *__restrict__ off,
                                                               const unsigned long long *__restrict__ off_base,
                                                               const uint8_t *__restrict__ header, int hdr, int sof, int w,
                                                               int h, uint8_t *__restrict__ file, size_t file_cap)
{
    if (blockIdx.x == 0) {
        for (int t = threadIdx.x; t < hdr; t += blockDim.x) {
            uint8_t v = header[t];
            if (t == sof) v = (uint8_t)(h >> 8);
            if (t == sof + 1) v = (uint8_t)(h & 255);
            if (t == sof + 2) v = (uint8_t)(w >> 8);
            if (t == sof + 3) v = (uint8_t)(w & 255);
            if ((size_t)t < file_cap) file[t] = v;
        }
    }
    const size_t nch = stream_chunks(total_bits);
    uint8_t *out = file + hdr;
    const size_t cap = file_cap > (size_t)hdr ? file_cap - hdr : 0;
    const unsigned long long T = *total_bits, nbytes = (T + 7) >> 3;
    const int pad = (int)(8 * nbytes - T);
    const size_t nw = nch * 8;
    const int lane = threadIdx.x & 31;
    for (size_t w0 = (size_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31u); w0 < nw; w0 += (size_t)gridDim.x * blockDim.x) {
        const size_t wi = w0 + lane;
        const long long left = (long long)nbytes - 4 * (long long)wi;
        const int nvalid = wi < nw ? (int)min(4ll, max(0ll, left)) : 0;
        const uint32_t v = wi < nw ? stream_word(words, wi, nbytes, pad) : 0u;
        const int c = ff_bytes(v, nvalid);
        // 0xFF bytes in the words of this chunk in front of this one: inclusive scan over the group of eight lanes
        int inc = c;
#pragma unroll
        for (int d = 1; d < 8; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, d);
            if ((lane & 7) >= d) inc += t;
        }
        if (wi >= nw) continue;                      // (after the shuffles)
        const size_t ch = wi >> 3;
        size_t o = 4 * wi + (size_t)(off[ch] + off_base[ch >> 10]) + (size_t)(inc - c);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k < nvalid) {
                const uint32_t byte = (v >> (24 - 8 * k)) & 0xFFu;
                if (o < cap) out[o] = (uint8_t)byte;
                o++;
                if (byte == 0xFFu) {
                    if (o < cap) out[o] = 0;
                    o++;
                }
            }
        }
        if (nvalid > 0 && left <= 4) {   // EOI behind the last data byte
            if (o < cap) out[o] = 0xFF;
            if (o + 1 < cap) out[o + 1] = 0xD9;
        }
    }
}

// ---------------------------------------------------------------------------
static void host_quant(const uint8_t *base, int quality, uint8_t *out)   // jcparam.c, force_baseline
{
    quality = std::min(100, std::max(1, quality));
    const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
    for (int i = 0; i < 64; i++) out[i] = (uint8_t)std::min(255L, std::max(1L, ((long)base[i] * scale + 50L) / 100L));
}

static void host_huff(const uint8_t *bits, const uint8_t *vals, uint16_t *code, uint8_t *len)
{
    memset(code, 0, 256 * sizeof(uint16_t));
    memset(len, 0, 256);
    unsigned c = 0;
    int k = 0;
    for (int l = 1; l <= 16; l++) {
        for (int i = 0; i < bits[l - 1]; i++, k++) {
            code[vals[k]] = (uint16_t)c++;
            len[vals[k]] = (uint8_t)l;
        }
        c <<= 1;
    }
}

static size_t host_header(const uint8_t *ql, const uint8_t *qc, uint8_t *o, int *sof)
{
    size_t n = 0;
    auto put = [&](std::initializer_list<int> b) { for (int v : b) o[n++] = (uint8_t)v; };
    put({0xFF, 0xD8, 0xFF, 0xE0, 0, 16, 'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0});
    for (int t = 0; t < 2; t++) {
        put({0xFF, 0xDB, 0, 67, t});
        for (int k = 0; k < 64; k++) o[n++] = (t ? qc : ql)[h_zz[k]];
    }
    *sof = (int)n + 5;   // height (2 bytes), width (2 bytes): written by the kernel
    put({0xFF, 0xC0, 0, 17, 8, 0, 0, 0, 0, 3, 1, 0x22, 0, 2, 0x11, 1, 3, 0x11, 1});
    const struct { int id; const uint8_t *bits, *vals; int nv; } dht[4] = {
        {0x00, h_dc_l_bits, h_dc_vals, 12}, {0x10, h_ac_l_bits, h_ac_l_vals, 162},
        {0x01, h_dc_c_bits, h_dc_vals, 12}, {0x11, h_ac_c_bits, h_ac_c_vals, 162}};
    for (const auto &d : dht) {
        const int L = 2 + 1 + 16 + d.nv;
        put({0xFF, 0xC4, L >> 8, L & 255, d.id});
        for (int i = 0; i < 16; i++) o[n++] = d.bits[i];
        for (int i = 0; i < d.nv; i++) o[n++] = d.vals[i];
    }
    put({0xFF, 0xDA, 0, 12, 3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 0x3F, 0});
    return n;
}

// The tables that do not depend on the quality: once per device.
static int32_t fixed_tables(int device)
{
    static std::mutex mu;
    static bool done[64] = {false};
    std::lock_guard<std::mutex> lk(mu);
    if (device >= 0 && device < 64 && done[device]) return OICS_OK;
    uint8_t izz[64], hlen[4][256];
    uint16_t hcode[4][256];
    for (int k = 0; k < 64; k++) izz[h_zz[k]] = (uint8_t)k;
    host_huff(h_dc_l_bits, h_dc_vals, hcode[0], hlen[0]);
    host_huff(h_ac_l_bits, h_ac_l_vals, hcode[1], hlen[1]);
    host_huff(h_dc_c_bits, h_dc_vals, hcode[2], hlen[2]);
    host_huff(h_ac_c_bits, h_ac_c_vals, hcode[3], hlen[3]);
    OICS_CUDA_TRY(cudaMemcpyToSymbol(c_izz, izz, sizeof(izz)));
    OICS_CUDA_TRY(cudaMemcpyToSymbol(c_hcode, hcode, sizeof(hcode)));
    OICS_CUDA_TRY(cudaMemcpyToSymbol(c_hlen, hlen, sizeof(hlen)));
    if (device >= 0 && device < 64) done[device] = true;
    return OICS_OK;
}

// Enqueue the encoder on ctx->stream without a host round trip: the file lands in d_out (writes past
// d_cap are dropped) and d_totals[0..1] (device) receive the entropy-coded bits and the stuffed 0xFF
// count; the file size is jpeg_file_size(*hdr_len, totals) once the caller has read them back.
int32_t jpeg_lane_get(oics_ctx *ctx, int i, JpegLane **lane)
{
    JpegLane &L = ctx->jpeg_lanes[i];
    if (!L.st) {
        int lo = 0, hi = 0;   // beside the context's own stream, above the decode streams
        OICS_CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        static const int prio_env = [] { const char *e = getenv("OICS_LANE_PRIORITY"); return e ? atoi(e) : 1000; }();
        OICS_CUDA_TRY(cudaStreamCreateWithPriority(&L.st, cudaStreamNonBlocking, prio_env == 1000 ? hi : std::max(hi, std::min(lo, prio_env))));
        OICS_CUDA_TRY(cudaEventCreateWithFlags(&L.done, cudaEventDisableTiming));
    }
    *lane = &L;
    return OICS_OK;
}

void jpeg_lanes_release(oics_ctx *ctx)
{
    for (JpegLane &L : ctx->jpeg_lanes) {
        if (L.st) { cudaStreamSynchronize(L.st); cudaStreamDestroy(L.st); L.st = nullptr; }
        if (L.done) { cudaEventDestroy(L.done); L.done = nullptr; }
        for (DevBuf *b : {&L.tabs, &L.raw, &L.coef, &L.bits, &L.words, &L.cnt}) b->release();
    }
}

// `lane`: encode on that lane's stream with its scratch (the caller orders the lane against ctx->stream);
// null = ctx->stream and the context's scratch.
int32_t jpeg_encode_enqueue(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, int quality, uint8_t *d_out,
                            size_t d_cap, unsigned long long *d_totals, size_t *hdr_len, JpegLane *lane)
{
    cudaStream_t st = lane ? lane->st : ctx->stream;
    if (w <= 0 || h <= 0 || w > 65535 || h > 65535) {
        set_error("jpeg: image size %dx%d is outside 1..65535", w, h);
        return OICS_ERR_INVALID_ARG;
    }
    JpegGeom G;
    G.w = w; G.h = h;
    G.mcux = cdiv(w, 16); G.mcuy = cdiv(h, 16);
    G.wibY = cdiv(w, 8); G.hibY = cdiv(h, 8);
    G.nblocks = 6 * G.mcux * G.mcuy;
    G.cstride = ((size_t)G.nblocks + 63) & ~(size_t)63;
    OICS_TRY(fixed_tables(ctx->device));
    quality = std::min(100, std::max(1, quality));
    JpegQuant Q;
    host_quant(h_ql, quality, Q.q[0]);
    host_quant(h_qc, quality, Q.q[1]);
    for (int t = 0; t < 2; t++)
        for (int k = 0; k < 64; k++) Q.rcp[t][k] = 1.0f / (float)((int)Q.q[t][k] << 3);
    if (ctx->jpeg_hdr_quality != quality) {                 // the header of this quality, cached on the device
        ctx->jpeg_hdr_len = host_header(Q.q[0], Q.q[1], ctx->jpeg_hdr_host, &ctx->jpeg_hdr_sof);
        OICS_TRY(ctx->jpeg_hdr.ensure(1024));
        OICS_CUDA_TRY(cudaDeviceSynchronize());             // lanes may still read the header of the last quality
        OICS_CUDA_TRY(cudaMemcpyAsync(ctx->jpeg_hdr.p, ctx->jpeg_hdr_host, ctx->jpeg_hdr_len, cudaMemcpyHostToDevice, st));
        OICS_CUDA_TRY(cudaStreamSynchronize(st));
        ctx->jpeg_hdr_quality = quality;
    }
    const size_t hdr = ctx->jpeg_hdr_len;
    *hdr_len = hdr;
    // scratch: cplx = coefficients, labels = bit counts / offsets + scan sums, cplx2 = bit stream,
    // mag = 0xFF counts per 32-byte chunk + scan sums
    const size_t nb = (size_t)G.nblocks;
    const size_t bound = nb * 208 + 64;                      // bytes: 20 DC + 63 * 26 AC bits per block at most
    const size_t nchunks = (bound + 31) / 32;
    const size_t nsums = (nb + 1023) / 1024 + 8, nsums2 = (nchunks + 1023) / 1024 + 8;
    DevBuf &b_coef = lane ? lane->coef : ctx->cplx, &b_bits = lane ? lane->bits : ctx->labels;
    DevBuf &b_words = lane ? lane->words : ctx->cplx2, &b_cnt = lane ? lane->cnt : ctx->mag;
    OICS_TRY(b_coef.ensure(64 * G.cstride * sizeof(int16_t)));
    OICS_TRY(b_bits.ensure((nb + nsums) * sizeof(unsigned long long)));
    OICS_TRY(b_words.ensure(bound + 64));
    OICS_TRY(b_cnt.ensure((nchunks + nsums2) * sizeof(unsigned long long)));
    int16_t *coef = b_coef.as<int16_t>();
    unsigned long long *bits = b_bits.as<unsigned long long>(), *sums = bits + nb;
    unsigned long long *cnt = b_cnt.as<unsigned long long>(), *sums2 = cnt + nchunks;
    uint32_t *words = b_words.as<uint32_t>();
    const int strips = cdiv(G.mcux, JD_MCU);
    {   // algorithmic bytes: the BGR image in, 16-bit coefficients of the 4:2:0 planes out (1.5 x 2 B per pixel)
        ProfScope ps(ctx, "jpeg_dct_kernel", 3.0 * w * h + 2.0 * 64 * nb, st);
        jpeg_dct_kernel<<<G.mcuy * strips, JD_T, 0, st>>>(dbgr, stride, G, Q, strips, coef);
    }
    {
        ProfScope ps(ctx, "jpeg_bits_kernel", 2.0 * 64 * nb + 8.0 * nb, st);
        jpeg_bits_kernel<<<cdiv(G.nblocks, 256), 256, 0, st>>>(coef, G, bits);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 2;
    OICS_TRY(exclusive_scan(ctx, st, bits, nb, sums, d_totals, nullptr));
    // the grids below cover the real stream with a stride loop: the host only knows the bound (208 bytes per block,
    // ~25 x the stream of a quality-100 card), and CTAs that find nothing to do are not free
    const unsigned cap_grid = (unsigned)(ctx->sm_count * 8);
    jpeg_zero_kernel<<<std::min<unsigned>(cap_grid, (unsigned)((bound / 16 + 2 + 255) / 256)), 256, 0, st>>>((uint4 *)words, d_totals);
    {
        ProfScope ps(ctx, "jpeg_emit_kernel", 2.0 * 64 * nb + 8.0 * nb, st);
        jpeg_emit_kernel<<<cdiv(G.nblocks, 256), 256, 0, st>>>(coef, G, bits, sums, words);
    }
    jpeg_stuff_count_kernel<<<std::min<unsigned>(cap_grid, (unsigned)((nchunks + 255) / 256)), 256, 0, st>>>(words, d_totals, cnt);
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches += 3;
    OICS_TRY(exclusive_scan(ctx, st, cnt, nchunks, sums2, d_totals + 1, d_totals));
    {
        ProfScope ps(ctx, "jpeg_stuff_write_kernel", 0.0, st);
        jpeg_stuff_write_kernel<<<std::min<unsigned>(cap_grid, (unsigned)((nchunks + 255) / 256)), 256, 0, st>>>(
            words, d_totals, cnt, sums2, ctx->jpeg_hdr.as<uint8_t>(), (int)hdr, ctx->jpeg_hdr_sof, w, h, d_out, d_cap);
    }
    OICS_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    return OICS_OK;
}

size_t jpeg_file_size(size_t hdr_len, const unsigned long long *totals)
{
    return hdr_len + (size_t)((totals[0] + 7) >> 3) + (size_t)totals[1] + 2;
}

// BGR image on the device -> JPEG file bytes on the device (d_out, capacity d_cap); *size = file size
// (OICS_ERR_CAPACITY with the needed size if it does not fit).
int32_t jpeg_encode_device(oics_ctx *ctx, const uint8_t *dbgr, int w, int h, size_t stride, int quality, uint8_t *d_out,
                           size_t d_cap, size_t *size)
{
    OICS_TRY(ctx->jpeg_totals.ensure(2 * sizeof(unsigned long long)));
    size_t hdr = 0;
    OICS_TRY(jpeg_encode_enqueue(ctx, dbgr, w, h, stride, quality, d_out, d_cap, ctx->jpeg_totals.as<unsigned long long>(), &hdr));
    unsigned long long totals[2] = {0, 0};
    OICS_CUDA_TRY(cudaMemcpyAsync(totals, ctx->jpeg_totals.p, sizeof(totals), cudaMemcpyDeviceToHost, ctx->stream));
    OICS_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    *size = jpeg_file_size(hdr, totals);
    if (*size > d_cap) {
        set_error("jpeg: the file needs %zu bytes, capacity %zu", *size, d_cap);
        return OICS_ERR_CAPACITY;
    }
    return OICS_OK;
}

}  // namespace oics
