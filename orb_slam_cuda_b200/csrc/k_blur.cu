// csrc/k_blur.cu — cv::GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) on 8-bit planes
// (reference call site src/ORBextractor.cc:1746-1747).  OpenCV's CV_8U path is a bit-exact 8.8 fixed-point
// separable filter: kernel [18,34,48,56,48,34,18]/256, rows then columns, result = (sum + 32768) >> 16.
//
// One CTA produces a 128 x 32 output tile.  The 134 x 38 input window (3-px halo) is staged with one TMA box load of
// 160 x 38 bytes from the padded plane (TMA needs the box origin on a 16-byte boundary of the row: measured on B200,
// tools/tma_probe.cu; the window starts at byte 13 of the box) — the plane's 19-px border IS the reflect-101
// extension of the level, so no border logic is needed here.  The horizontal pass writes 16-bit partial sums
// (<= 255*256) to shared memory, the vertical pass emits 4 bytes per thread per row with 32-bit stores.
//
// Roofline: HBM-bound (reads each level once, writes it once).
#include "orbx_internal.h"
#include "tma.cuh"

namespace orbx {

constexpr int BLUR_TW = 128, BLUR_TH = 32, BLUR_BW = 160, BLUR_BH = 38, BLUR_THREADS = 256;

__global__ void __launch_bounds__(BLUR_THREADS) k_blur7(const __grid_constant__ FrameGeom g,
                                                        const __grid_constant__ TmaMaps maps) {
    __shared__ __align__(128) uint8_t tile[BLUR_BH * BLUR_BW];
    __shared__ __align__(16) uint16_t hsum[BLUR_BH * BLUR_TW];
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, f = blockIdx.y;
    int t = blockIdx.x, level = 0;
    while (level + 1 < g.nlevels && t >= g.lv[level + 1].blurTileBase) ++level;
    const LevelGeom& L = g.lv[level];
    t -= L.blurTileBase;
    const int ty = t / L.blurTilesX, tx = t - ty * L.blurTilesX;
    const int x0 = tx * BLUR_TW, y0 = ty * BLUR_TH;
    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(&bar, BLUR_BH * BLUR_BW);
        tma_load_3d(tile, &maps.m[level], XOFF + x0 - 16, EDGE + y0 - 3, f, &bar);  // window = box columns 13..146
    }
    mbar_wait(&bar, 0);

    // horizontal pass: item = (row, group of 4 output columns)
    for (int it = tid; it < BLUR_BH * (BLUR_TW / 4); it += BLUR_THREADS) {
        const int r = it / (BLUR_TW / 4), q = it - r * (BLUR_TW / 4);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(tile + r * BLUR_BW + 12 + 4 * q);
        const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];  // window bytes 4q..4q+9 are bytes 1..10 of these words
        int p[10];
#pragma unroll
        for (int k = 0; k < 3; ++k) p[k] = (w0 >> (8 * (k + 1))) & 0xFF;
#pragma unroll
        for (int k = 0; k < 4; ++k) p[3 + k] = (w1 >> (8 * k)) & 0xFF;
#pragma unroll
        for (int k = 0; k < 3; ++k) p[7 + k] = (w2 >> (8 * k)) & 0xFF;
        uint32_t s[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
            s[k] = 18 * (p[k] + p[k + 6]) + 34 * (p[k + 1] + p[k + 5]) + 48 * (p[k + 2] + p[k + 4]) + 56 * p[k + 3];
        uint2 o;
        o.x = s[0] | (s[1] << 16);
        o.y = s[2] | (s[3] << 16);
        *reinterpret_cast<uint2*>(hsum + r * BLUR_TW + 4 * q) = o;
    }
    __syncthreads();
    // vertical pass: item = (group of 4 rows, group of 4 columns)
    uint8_t* out = L.blur + (long long)f * L.blurStride;
    for (int it = tid; it < (BLUR_TH / 4) * (BLUR_TW / 4); it += BLUR_THREADS) {
        const int rg = it / (BLUR_TW / 4), q = it - rg * (BLUR_TW / 4);
        uint32_t c[10][4];
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            const uint2 v = *reinterpret_cast<const uint2*>(hsum + (4 * rg + r) * BLUR_TW + 4 * q);
            c[r][0] = v.x & 0xFFFF; c[r][1] = v.x >> 16; c[r][2] = v.y & 0xFFFF; c[r][3] = v.y >> 16;
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int y = y0 + 4 * rg + r, x = x0 + 4 * q;
            uint32_t word = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t acc = 18u * (c[r][k] + c[r + 6][k]) + 34u * (c[r + 1][k] + c[r + 5][k]) +
                                     48u * (c[r + 2][k] + c[r + 4][k]) + 56u * c[r + 3][k];
                word |= ((acc + 32768u) >> 16) << (8 * k);
            }
            if (y < L.h && x < L.bpitch) *reinterpret_cast<uint32_t*>(out + (long long)y * L.bpitch + x) = word;
        }
    }
}

int launch_blur(const FrameGeom& g, const TmaMaps& maps, int batch, cudaStream_t s) {
    dim3 grid(g.blurTilesTotal, batch);
    k_blur7<<<grid, BLUR_THREADS, 0, s>>>(g, maps);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
