// csrc/k_blur.cu — cv::GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) on 8-bit planes
// (reference call site src/ORBextractor.cc:1746-1747).  OpenCV's CV_8U path is a bit-exact 8.8 fixed-point
// separable filter: kernel [18,34,48,56,48,34,18]/256, rows then columns, result = (sum + 32768) >> 16.
//
// One CTA produces a 128 x 32 output tile.  The 134 x 38 input window (3-px halo) is staged with one TMA box load of
// 160 x 38 bytes from the padded plane (TMA needs the box origin on a 16-byte boundary of the row: measured on B200,
// tools/tma_probe.cu; the window starts at byte 13 of the box) — the plane's 19-px border IS the reflect-101
// extension of the level, so no border logic is needed here.  Both passes run in registers (see the kernel comment);
// a thread emits 4 bytes per row with 32-bit stores.
//
// Roofline: HBM-bound (reads each level once, writes it once).
#include "orbx_internal.h"
#include "tma.cuh"

namespace orbx {

#ifndef ORBX_BLUR_BAND
#define ORBX_BLUR_BAND 32  // = the tile height: ONE warp per CTA (measured on B200: 0.41 ms against 0.54 ms with two warps of 16 rows; profiles/r02_variants.txt)
#endif
constexpr int BLUR_TW = 128, BLUR_TH = BLUR_TILE_H, BLUR_BW = 160, BLUR_BH = BLUR_TH + 6;
constexpr int BLUR_BAND = ORBX_BLUR_BAND;            // output rows per thread (a thread walks BLUR_BAND + 6 input rows)
constexpr int BLUR_THREADS = 32 * (BLUR_TH / BLUR_BAND);

// One CTA = one 128 x 32 output tile; one thread = 4 output columns x 16 output rows.  Walking down its band, a thread
// forms the horizontal 7-tap sums of the arriving input row with dp4a (pixels realigned by funnel shifts; coefficient
// bytes (18,34,48,56) and (48,34,18,0)), keeps the last rows as vertically packed 16-bit pairs in registers and forms the
// vertical sum with three dp2a + one multiply-add.  The horizontal sums (<= 255 * 256) fit 16 bits, the vertical sum
// + 32768 fits 32 bits, and the result byte is bits 16..23 of the accumulator (picked with PRMT, no shift).
template <int BAND>  // output rows per warp: BLUR_BAND (= the tile height: one warp per CTA) in batches, 8 (four warps per CTA) for single frames
__global__ void __launch_bounds__(32 * (BLUR_TH / BAND)) k_blur7(const __grid_constant__ FrameGeom g,
                                                        const __grid_constant__ TmaMaps maps) {
    __shared__ __align__(128) uint8_t tile[BLUR_BH * BLUR_BW];
    __shared__ __align__(8) uint64_t bar;
    if (g.pdl) griddep_launch_dependents();
    const int tid = threadIdx.x, f = blockIdx.y + g.f0;  // scratch frame
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
    if (g.pdl) griddep_wait();  // the pyramid comes from kernels further up the chain: complete once the predecessor is
    if (tid == 0) {
        mbar_expect_tx(&bar, BLUR_BH * BLUR_BW);
        tma_load_3d(tile, &maps.m[level], XOFF + x0 - 16, EDGE + y0 - 3, f, &bar);  // window = box columns 13..146
    }
    const int cg = tid & 31, band = tid >> 5;
    const int xo = x0 + 4 * cg, yb = y0 + band * BAND;
    uint8_t* out = L.blur + (long long)f * L.blurStride + xo;
    const bool colOk = xo < L.bpitch;
    const uint32_t K0 = 18u | (34u << 8) | (48u << 16) | (56u << 24), K1 = 48u | (34u << 8) | (18u << 16);
    const uint32_t V0 = 18u | (34u << 8), V1 = 48u | (56u << 8), V2 = 48u | (34u << 8);
    mbar_wait(&bar, 0);

    const uint32_t* rowp = reinterpret_cast<const uint32_t*>(tile + band * BAND * BLUR_BW + 12) + cg;
    uint32_t pr[8][4];  // pr[i & 7][c] = (h(i), h(i + 1)) of column c as two 16-bit halves
    uint32_t hp[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < BAND + 6; ++i) {
        // window bytes 4cg .. 4cg + 9 of input row i are bytes 1 .. 10 of these three words
        const uint32_t w0 = rowp[i * (BLUR_BW / 4)], w1 = rowp[i * (BLUR_BW / 4) + 1], w2 = rowp[i * (BLUR_BW / 4) + 2];
        uint32_t h[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t a = k == 3 ? w1 : __funnelshift_r(w0, w1, 8 * (k + 1));
            const uint32_t b = k == 3 ? w2 : __funnelshift_r(w1, w2, 8 * (k + 1));
            h[k] = __dp4a(a, K0, __dp4a(b, K1, 0u));
        }
        if (i >= 1) {
#pragma unroll
            for (int k = 0; k < 4; ++k) pr[(i - 1) & 7][k] = hp[k] | (h[k] << 16);
        }
        if (i >= 6) {
            const int o = i - 6;  // output row of the band: inputs o .. o + 6
            uint32_t acc[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                uint32_t v = __dp2a_lo(pr[o & 7][k], V0, 32768u);
                v = __dp2a_lo(pr[(o + 2) & 7][k], V1, v);
                v = __dp2a_lo(pr[(o + 4) & 7][k], V2, v);
                acc[k] = v + 18u * h[k];
            }
            const uint32_t word = __byte_perm(__byte_perm(acc[0], acc[1], 0x0062), __byte_perm(acc[2], acc[3], 0x0062), 0x5410);
            if (colOk && yb + o < L.h) *reinterpret_cast<uint32_t*>(out + (long long)(yb + o) * L.bpitch) = word;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) hp[k] = h[k];
    }
}

int launch_blur(const FrameGeom& g, const TmaMaps& maps, int batch, cudaStream_t s) {
    dim3 grid(g.blurTilesTotal, batch);
    if (g.pdl)  // single frames: four warps share a tile's box (the row walk of a warp is the critical path on an empty GPU)
        ORBX_CUDA(launch_kernel(k_blur7<8>, grid, dim3(32 * (BLUR_TH / 8)), 0, s, !g.plainLaunch, g, maps));
    else
        ORBX_CUDA(launch_kernel(k_blur7<BLUR_BAND>, grid, dim3(BLUR_THREADS), 0, s, false, g, maps));
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
