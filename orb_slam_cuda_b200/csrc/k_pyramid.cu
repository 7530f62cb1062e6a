// csrc/k_pyramid.cu — ORBextractor::ComputePyramid (reference src/ORBextractor.cc:1837-1862, CPU branch).
//
// Level 0 is the input frame with a 19-px BORDER_REFLECT_101 frame around it (copyMakeBorder, :1859-1860); level l
// is cv::resize(level l-1, INTER_LINEAR) followed by the same border (:1852-1855).  cv::resize on CV_8U is an
// 11-bit fixed-point separable bilinear filter: the per-column / per-row source indices and coefficients are
// computed on the host exactly as OpenCV does (double scale, float fx, cvRound(f*2048)) and handed in as tables;
// the kernel does the integer blend ((b0*(r0>>4))>>16) + ((b1*(r1>>4))>>16) + 2) >> 2.  Border bytes are produced
// by evaluating the interior formula at the reflected coordinate, so the plane is written exactly once and no second
// border pass is needed.
//
// Roofline: HBM-bound streaming (reads level l-1 once through L1/L2, writes level l once).
#include <algorithm>

#include "orbx_internal.h"
#include "tma.cuh"

namespace orbx {

__device__ __forceinline__ int reflect101(int p, int n) {
    // cv::borderInterpolate(p, n, BORDER_REFLECT_101); |p| excursions here are <= 19 but n may be tiny
    if ((unsigned)p < (unsigned)n) return p;
    if (n == 1) return 0;
    do {
        if (p < 0) p = -p;
        else p = 2 * n - 2 - p;
    } while ((unsigned)p >= (unsigned)n);
    return p;
}

// Level 0, interior: a CTA of 8 warps writes a 512-byte x rowsPerCta tile of the padded plane; a thread owns 16 consecutive
// bytes (one 128-bit store per row).  A frame row starts at any byte alignment (w is odd for KITTI), so a lane loads the
// ALIGNED 16-byte chunk that holds its first byte (a warp reads 512 consecutive bytes), takes the following chunk from its
// neighbour lane by shuffle (the misalignment is the same for every lane of a row), and realigns the 32 bytes with four
// funnel shifts.  Only groups that lie fully inside the interior are written here; the two ends of a row (reflected border,
// the last partial interior group, row padding) are k_pyramid_base_edges' job, so no warp of this kernel diverges.
// grid: (ceil(pitch / 512), ceil(rows / rowsPerCta), batch); block 256
constexpr int RZ_ROWS = 64;  // minimum rows per CTA; launchers raise it (rowsPerCta) while the grid still fills the GPU
__global__ void __launch_bounds__(256) k_pyramid_base(uint8_t* __restrict__ plane, int w, int h, int pitch, int rows,
                                                      long long planeStride, const uint8_t* __restrict__ frames,
                                                      long long stride, long long frameStride, int rowsPerCta, int pdl) {
    if (pdl) griddep_launch_dependents();  // the edge kernel reads the same frames and writes other bytes: it runs beside this one
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bx = blockIdx.x * 512 + lane * 16;
    const int f = blockIdx.z;
    const int px0 = bx - XOFF;                                   // interior x of byte 0 of the group (a multiple of 16)
    const bool run = bx < pitch && px0 >= 0 && px0 + 16 <= w;    // 16 consecutive interior source bytes
    const bool nextRun = lane < 31 && px0 + 16 >= 0 && px0 + 32 <= w;  // lane + 1 loads the chunk after mine
    const uint8_t* __restrict__ fbase = frames + (long long)f * frameStride + px0;
    uint8_t* __restrict__ dbase = plane + (long long)f * planeStride + bx;
    const int rowEnd = min(blockIdx.y * rowsPerCta + rowsPerCta, rows);
    for (int py = blockIdx.y * rowsPerCta + warp; py < rowEnd; py += 8) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(fbase + (long long)reflect101(py - EDGE, h) * stride);
        const uint4* c0 = reinterpret_cast<const uint4*>(a & ~(uintptr_t)15);
        const uint32_t mis = (uint32_t)(a & 15);  // the same in every lane
        uint4 A = make_uint4(0u, 0u, 0u, 0u);
        if (run) A = __ldg(c0);
        uint4 N;
        N.x = __shfl_down_sync(0xffffffffu, A.x, 1);
        N.y = __shfl_down_sync(0xffffffffu, A.y, 1);
        N.z = __shfl_down_sync(0xffffffffu, A.z, 1);
        N.w = __shfl_down_sync(0xffffffffu, A.w, 1);
        if (!run) continue;
        // the following chunk is read by this lane only when no neighbour has it, and only when it holds a needed byte
        // (mis != 0), so no load passes the 16-byte chunk of the frame's last byte
        if (!nextRun) N = mis ? __ldg(c0 + 1) : make_uint4(0u, 0u, 0u, 0u);
        const uint32_t sh = (mis & 3u) * 8u;
        uint4 v;
        switch (mis >> 2) {
            case 0: v = make_uint4(__funnelshift_r(A.x, A.y, sh), __funnelshift_r(A.y, A.z, sh), __funnelshift_r(A.z, A.w, sh), __funnelshift_r(A.w, N.x, sh)); break;
            case 1: v = make_uint4(__funnelshift_r(A.y, A.z, sh), __funnelshift_r(A.z, A.w, sh), __funnelshift_r(A.w, N.x, sh), __funnelshift_r(N.x, N.y, sh)); break;
            case 2: v = make_uint4(__funnelshift_r(A.z, A.w, sh), __funnelshift_r(A.w, N.x, sh), __funnelshift_r(N.x, N.y, sh), __funnelshift_r(N.y, N.z, sh)); break;
            default: v = make_uint4(__funnelshift_r(A.w, N.x, sh), __funnelshift_r(N.x, N.y, sh), __funnelshift_r(N.y, N.z, sh), __funnelshift_r(N.z, N.w, sh)); break;
        }
        *reinterpret_cast<uint4*>(dbase + (long long)py * pitch) = v;
    }
}

// Level 0, row ends: one thread per (row, 16-byte group outside the fully interior range) gathers its bytes one by one at
// the reflected coordinate (BORDER_REFLECT_101) and writes zeros into the row padding.  nEdge = groups per row it owns:
// groups 0, 1 (left border) and the groups from the first not fully interior one to the end of the row.
__device__ __forceinline__ void pyramid_base_edges_body(uint8_t* __restrict__ plane, int w, int h, int pitch, int rows,
                                                            long long planeStride, const uint8_t* __restrict__ frames,
                                                            long long stride, long long frameStride, int firstRight, int nEdge) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= rows * nEdge) return;
    const int py = i / nEdge, e = i - py * nEdge;
    const int bx = (e < 2 ? e : firstRight + (e - 2)) * 16;
    const int f = blockIdx.y;
    const uint8_t* __restrict__ src = frames + (long long)f * frameStride + (long long)reflect101(py - EDGE, h) * stride;
    uint32_t q[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const int px = bx + j - PADL;
        if (px >= 0 && px < w + 2 * EDGE) q[j >> 2] |= (uint32_t)__ldg(src + reflect101(px - EDGE, w)) << (8 * (j & 3));
    }
    *reinterpret_cast<uint4*>(plane + (long long)f * planeStride + (long long)py * pitch + bx) = make_uint4(q[0], q[1], q[2], q[3]);
}

__global__ void __launch_bounds__(256) k_pyramid_base_edges(uint8_t* __restrict__ plane, int w, int h, int pitch, int rows,
                                                            long long planeStride, const uint8_t* __restrict__ frames,
                                                            long long stride, long long frameStride, int firstRight, int nEdge, int pdl) {
    // pdl: launched beside k_pyramid_base (same input, disjoint output).  Its successor waits for THIS grid only, so this grid
    // does not end before k_pyramid_base has: every thread passes griddep_wait() before it exits.
    if (pdl) griddep_launch_dependents();
    pyramid_base_edges_body(plane, w, h, pitch, rows, planeStride, frames, stride, frameStride, firstRight, nEdge);
    if (pdl) griddep_wait();
}

// Resize: a CTA of 8 warps produces a 128-px x 64-row tile of the padded destination plane.  A thread owns 4
// consecutive destination pixels (one 32-bit store per row) and keeps their four x-table entries in registers for
// the 8 rows its warp walks, so the tables cost nothing per pixel; a warp reads ~154 consecutive source bytes per
// source row through L1 (read-only path).  grid: (ceil(dpitch / 128), ceil(drows / 64), batch); block 256
__global__ void __launch_bounds__(256) k_pyramid_resize(uint8_t* __restrict__ dplane, int dw, int dh, int dpitch, int drows,
                                                        long long dStride, const uint8_t* __restrict__ splane, int spitch,
                                                        long long sStride, const ResizeTab* __restrict__ xtab,
                                                        const ResizeTab* __restrict__ ytab, int areaFast, int rowsPerCta) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bx = blockIdx.x * 128 + lane * 4;  // byte column of this thread's 4 pixels
    if (bx >= dpitch) return;
    const int f = blockIdx.z;
    int s0[4], s1[4], c0[4], c1[4];
    bool valid[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int px = bx + j - PADL;
        valid[j] = px >= 0 && px < dw + 2 * EDGE;
        const ResizeTab xt = xtab[valid[j] ? reflect101(px - EDGE, dw) : 0];
        s0[j] = xt.s0; s1[j] = xt.s1; c0[j] = xt.c0; c1[j] = xt.c1;
    }
    const uint8_t* __restrict__ sInt = splane + (long long)f * sStride + (long long)EDGE * spitch + XOFF;
    uint8_t* __restrict__ dbase = dplane + (long long)f * dStride + bx;
    const int rowEnd = min(blockIdx.y * rowsPerCta + rowsPerCta, drows);
    for (int py = blockIdx.y * rowsPerCta + warp; py < rowEnd; py += 8) {
        const ResizeTab yt = ytab[reflect101(py - EDGE, dh)];
        const uint8_t* __restrict__ r0p = sInt + (long long)yt.s0 * spitch;
        const uint8_t* __restrict__ r1p = sInt + (long long)yt.s1 * spitch;
        const int b0 = yt.c0, b1 = yt.c1;
        uint32_t v = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int p00 = __ldg(r0p + s0[j]), p01 = __ldg(r0p + s1[j]);
            const int p10 = __ldg(r1p + s0[j]), p11 = __ldg(r1p + s1[j]);
            int val;
            if (areaFast) {
                val = (p00 + p01 + p10 + p11 + 2) >> 2;
            } else {
                const int h0 = p00 * c0[j] + p01 * c1[j];
                const int h1 = p10 * c0[j] + p11 * c1[j];
                val = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
                val = min(max(val, 0), 255);
            }
            v |= (valid[j] ? (uint32_t)val : 0u) << (8 * j);
        }
        *reinterpret_cast<uint32_t*>(dbase + (long long)py * dpitch) = v;
    }
}

// Word-based resize (the fast path; used when the 4 destination pixels of a thread read at most 8 consecutive source
// bytes per row, i.e. for scale factors up to 2).  Per source row a thread loads three aligned words, realigns them to
// an 8-byte window with two funnel shifts, picks each pixel's (left, right) byte pair with one PRMT whose selector is
// fixed for the thread, and forms the horizontal blend with one dp2a against the packed 16-bit coefficient pair.
// The vertical blend keeps OpenCV's truncation order: ((b * (h >> 4)) >> 16) is one multiply-high by (b << 16).
// Exact 2x decimation (cv::resize's INTER_AREA shortcut) is the same formula with all four coefficients = 1024.
__global__ void __launch_bounds__(256) k_pyramid_resize4(uint8_t* __restrict__ dplane, int dw, int dh, int dpitch, int drows,
                                                         long long dStride, const uint8_t* __restrict__ splane, int spitch,
                                                         long long sStride, const ResizeTab* __restrict__ xtab,
                                                         const ResizeTab* __restrict__ ytab, int rowsPerCta) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bx = blockIdx.x * 128 + lane * 4;  // byte column of this thread's 4 pixels
    const bool inRow = bx < dpitch;  // (every thread helps to fill the row table below)
    const int f = blockIdx.z;
    int s0[4], s1[4];
    uint32_t coef[4], vmask = 0;
    int smin = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int px = bx + j - PADL;
        const bool valid = px >= 0 && px < dw + 2 * EDGE;
        const ResizeTab xt = xtab[valid ? reflect101(px - EDGE, dw) : 0];
        s0[j] = xt.s0; s1[j] = xt.s1;
        coef[j] = (uint32_t)(uint16_t)xt.c0 | ((uint32_t)(uint16_t)xt.c1 << 16);
        if (valid) {
            vmask |= 0xFFu << (8 * j);
            smin = min(smin, (int)xt.s0);
        }
    }
    if (vmask == 0) smin = 0;
    uint32_t sel[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const bool valid = (vmask >> (8 * j)) & 1;
        sel[j] = valid ? (uint32_t)(s0[j] - smin) | ((uint32_t)(s1[j] - smin) << 4) : 0u;
    }
    const int q0 = smin >> 2;
    const uint32_t sh = (uint32_t)(smin & 3) * 8;
    // per-row constants of the CTA's rows, computed once: word offsets of the two source rows and the vertical coefficients
    // (<< 16 for the multiply-high) — the row loop then costs one 128-bit shared load instead of a reflect, a table load and
    // two 64-bit address computations per row and thread
    extern __shared__ uint4 rowTab[];
    const int row0 = blockIdx.y * rowsPerCta, rowEnd = min(row0 + rowsPerCta, drows);
    const int spq = spitch >> 2;
    for (int r = row0 + threadIdx.x; r < rowEnd; r += blockDim.x) {
        const ResizeTab yt = ytab[reflect101(r - EDGE, dh)];
        rowTab[r - row0] = make_uint4((uint32_t)((int)yt.s0 * spq), (uint32_t)((int)yt.s1 * spq), (uint32_t)(uint16_t)yt.c0 << 16,
                                      (uint32_t)(uint16_t)yt.c1 << 16);
    }
    __syncthreads();
    if (!inRow) return;
    const uint32_t* __restrict__ sInt =
        reinterpret_cast<const uint32_t*>(splane + (long long)f * sStride + (long long)EDGE * spitch + XOFF) + q0;
    uint8_t* __restrict__ dbase = dplane + (long long)f * dStride + bx;
    // software pipeline: the six source words of the NEXT row are in flight while the current row is blended and stored
    int py = row0 + warp;
    if (py >= rowEnd) return;
    uint4 rt = rowTab[py - row0];
    uint32_t a0 = __ldg(sInt + rt.x), a1 = __ldg(sInt + rt.x + 1), a2 = __ldg(sInt + rt.x + 2);
    uint32_t b0w = __ldg(sInt + rt.y), b1w = __ldg(sInt + rt.y + 1), b2w = __ldg(sInt + rt.y + 2);
    for (; py < rowEnd; py += 8) {
        const int pn = py + 8 < rowEnd ? py + 8 : py;  // (the last row loads itself again: harmless)
        const uint4 rn = rowTab[pn - row0];
        const uint32_t na0 = __ldg(sInt + rn.x), na1 = __ldg(sInt + rn.x + 1), na2 = __ldg(sInt + rn.x + 2);
        const uint32_t nb0 = __ldg(sInt + rn.y), nb1 = __ldg(sInt + rn.y + 1), nb2 = __ldg(sInt + rn.y + 2);
        const uint32_t A0 = __funnelshift_r(a0, a1, sh), A1 = __funnelshift_r(a1, a2, sh);
        const uint32_t B0 = __funnelshift_r(b0w, b1w, sh), B1 = __funnelshift_r(b1w, b2w, sh);
        const uint32_t m0 = rt.z, m1 = rt.w;
        uint32_t v = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t h0 = __dp2a_lo(coef[j], __byte_perm(A0, A1, sel[j]), 0u);
            const uint32_t h1 = __dp2a_lo(coef[j], __byte_perm(B0, B1, sel[j]), 0u);
            const uint32_t val = (__umulhi(m0, h0 >> 4) + __umulhi(m1, h1 >> 4) + 2u) >> 2;  // <= 255
            v |= val << (8 * j);
        }
        *reinterpret_cast<uint32_t*>(dbase + (uint32_t)py * (uint32_t)dpitch) = v & vmask;
        rt = rn;
        a0 = na0; a1 = na1; a2 = na2;
        b0w = nb0; b1w = nb1; b2w = nb2;
    }
}


// TMA-tiled resize (the fast path for scale factors whose tile windows fit the box; 1.2 does): ONE warp per CTA produces a 128-byte
// x RZT_TH-row tile of the padded destination plane.  The source window of the tile (about 1.2 x the tile in both directions; its
// origin per tile column / tile row comes from host tables, 16-byte aligned in x as TMA wants) is staged in shared memory by one
// TMA box load, so no thread waits on a global load, and many one-warp CTAs per SM keep many boxes in flight (the layout that
// made the blur kernel 24 % faster).  A thread owns 4 destination columns and walks the tile's rows top to bottom: per SOURCE row
// it reads three words, realigns them and forms the four horizontal blends (PRMT + dp2a, as k_pyramid_resize4); the lower source
// row of one destination row is the upper one of the next four times out of five at 1.2, so its blends are kept and reused —
// a destination row costs about 1.2 source rows instead of 2.  The vertical blend keeps OpenCV's truncation order.
__global__ void __launch_bounds__(32) k_pyramid_resize_tma(const __grid_constant__ TmaMaps maps, int srcLevel, uint8_t* __restrict__ dplane,
                                                           int dw, int dh, int dpitch, int drows, long long dStride,
                                                           const ResizeTab* __restrict__ xtab, const ResizeTab* __restrict__ ytab,
                                                           const int* __restrict__ tileX0, const int* __restrict__ tileY0, int f0, int pdl,
                                                           int split) {
    // pdl (single-frame latency path): this grid was launched while the level above was still being written, and lets the next
    // level start in turn; everything up to the TMA load (tables, constants) overlaps the predecessor, the load waits for it
    if (pdl) griddep_launch_dependents();
    __shared__ __align__(128) uint8_t tile[RZT_BH * RZT_BW];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint4 rowTab[RZT_TH];
    const int lane = threadIdx.x;
    const int bx = blockIdx.x * 128 + lane * 4;  // byte column of this thread's 4 pixels
    const int f = blockIdx.z;
    // split > 1 (single frames: the GPU is empty and a tile's row loop is the critical path): `split` CTAs share one tile, each
    // loads the tile's box and produces RZT_TH / split of its rows
    const int tileY = blockIdx.y / split, part = blockIdx.y - tileY * split;
    const int row0 = tileY * RZT_TH + part * (RZT_TH / split), rowEnd = min(row0 + RZT_TH / split, min(tileY * RZT_TH + RZT_TH, drows));
    if (row0 >= rowEnd) return;
    const int bX0 = tileX0[blockIdx.x], bY0 = tileY0[tileY];  // box origin in the source plane (byte column, row)
    if (lane == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
        mbar_expect_tx(&bar, RZT_BH * RZT_BW);
        if (!pdl) tma_load_3d(tile, &maps.m[srcLevel], bX0, bY0, f + f0, &bar);
    }
    if (row0 + lane < rowEnd) {  // per-row constants: tile rows of the two source rows, vertical coefficients << 16
        const ResizeTab yt = ytab[reflect101(row0 + lane - EDGE, dh)];
        rowTab[lane] = make_uint4((uint32_t)(EDGE + (int)yt.s0 - bY0), (uint32_t)(EDGE + (int)yt.s1 - bY0), (uint32_t)(uint16_t)yt.c0 << 16,
                                  (uint32_t)(uint16_t)yt.c1 << 16);
    }
    int s0[4], s1[4];
    uint32_t coef[4], vmask = 0;
    int smin = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int px = bx + j - PADL;
        const bool valid = bx < dpitch && px >= 0 && px < dw + 2 * EDGE;
        const ResizeTab xt = xtab[valid ? reflect101(px - EDGE, dw) : 0];
        s0[j] = xt.s0; s1[j] = xt.s1;
        coef[j] = (uint32_t)(uint16_t)xt.c0 | ((uint32_t)(uint16_t)xt.c1 << 16);
        if (valid) {
            vmask |= 0xFFu << (8 * j);
            smin = min(smin, (int)xt.s0);
        }
    }
    if (vmask == 0) smin = max(bX0 - XOFF, 0);
    uint32_t sel[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const bool valid = (vmask >> (8 * j)) & 1;
        sel[j] = valid ? (uint32_t)(s0[j] - smin) | ((uint32_t)(s1[j] - smin) << 4) : 0u;
    }
    const int colByte = XOFF + smin - bX0;  // byte column of the thread's window inside the box
    const uint32_t sh = (uint32_t)(colByte & 3) * 8;
    const uint32_t* __restrict__ tq = reinterpret_cast<const uint32_t*>(tile) + (colByte >> 2);
    __syncwarp();
    if (pdl) {
        griddep_wait();
        if (lane == 0) tma_load_3d(tile, &maps.m[srcLevel], bX0, bY0, f + f0, &bar);
    }
    mbar_wait(&bar, 0);
    if (bx >= dpitch) return;
    uint8_t* __restrict__ dbase = dplane + (long long)(f + f0) * dStride + bx;
    uint32_t hKeep[4] = {0u, 0u, 0u, 0u};
    uint32_t keepRow = 0xFFFFFFFFu;
    for (int py = row0; py < rowEnd; ++py) {
        const uint4 rt = rowTab[py - row0];
        uint32_t h0[4], h1[4];
        if (rt.x == keepRow) {
#pragma unroll
            for (int j = 0; j < 4; ++j) h0[j] = hKeep[j];
        } else {
            const uint32_t* w = tq + rt.x * (RZT_BW / 4);
            const uint32_t a0 = w[0], a1 = w[1], a2 = w[2];
            const uint32_t A0 = __funnelshift_r(a0, a1, sh), A1 = __funnelshift_r(a1, a2, sh);
#pragma unroll
            for (int j = 0; j < 4; ++j) h0[j] = __dp2a_lo(coef[j], __byte_perm(A0, A1, sel[j]), 0u);
        }
        if (rt.y == rt.x) {  // clamped rows at the top / bottom of the level: both source rows are the same
#pragma unroll
            for (int j = 0; j < 4; ++j) h1[j] = h0[j];
        } else {
            const uint32_t* w = tq + rt.y * (RZT_BW / 4);
            const uint32_t b0 = w[0], b1 = w[1], b2 = w[2];
            const uint32_t B0 = __funnelshift_r(b0, b1, sh), B1 = __funnelshift_r(b1, b2, sh);
#pragma unroll
            for (int j = 0; j < 4; ++j) h1[j] = __dp2a_lo(coef[j], __byte_perm(B0, B1, sel[j]), 0u);
        }
        uint32_t v = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t val = (__umulhi(rt.z, h0[j] >> 4) + __umulhi(rt.w, h1[j] >> 4) + 2u) >> 2;  // <= 255
            v |= val << (8 * j);
            hKeep[j] = h1[j];
        }
        keepRow = rt.y;
        *reinterpret_cast<uint32_t*>(dbase + (uint32_t)py * (uint32_t)dpitch) = v & vmask;
    }
}

// Rows one CTA walks: the per-thread column set-up (x-table entries, PRMT selectors) is paid once per CTA, so a CTA
// takes as many rows as possible while the grid keeps >= 2 full waves of 148 SMs x 8 CTAs.
static int rows_per_cta(int rows, int gx, int batch, int minRows = RZ_ROWS) {
    int rpc = (rows + RZ_ROWS - 1) / RZ_ROWS * RZ_ROWS;
    while (rpc > minRows && (long long)gx * ((rows + rpc - 1) / rpc) * batch < 2 * 148 * 8) rpc = (rpc / 2 + 7) / 8 * 8;
    return rpc < minRows ? minRows : rpc;
}

int launch_pyramid_base(const FrameGeom& g, const uint8_t* dFrames, int batch, size_t stride, size_t frameStride,
                        cudaStream_t s) {
    const LevelGeom& L = g.lv[0];
    // single frames (pdl): one row per warp, so that no warp walks a chain of dependent row loads
    const int gx = (L.pitch + 511) / 512, rpc = rows_per_cta(L.rows, gx, batch, g.pdl ? 8 : RZ_ROWS);
    dim3 grid(gx, (L.rows + rpc - 1) / rpc, batch);
    uint8_t* plane = L.plane + (long long)g.f0 * L.planeStride;
    k_pyramid_base<<<grid, 256, 0, s>>>(plane, L.w, L.h, L.pitch, L.rows, L.planeStride, dFrames, (long long)stride, (long long)frameStride, rpc,
                                        g.pdl);
    count_launch();
    // row ends: groups 0, 1 and everything from the first group that is not fully interior (interior x0 + 16 > w)
    const int firstRight = std::max(2, XOFF / 16 + L.w / 16), nEdge = 2 + (L.pitch / 16 - firstRight);
    dim3 egrid((L.rows * nEdge + 255) / 256, batch);
    ORBX_CUDA(launch_kernel(k_pyramid_base_edges, egrid, dim3(256), 0, s, g.pdl != 0, plane, L.w, L.h, L.pitch, L.rows, L.planeStride, dFrames,
                            (long long)stride, (long long)frameStride, firstRight, nEdge, g.pdl));
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

int launch_pyramid_level(const FrameGeom& g, int level, const ResizeTab* xtab, const ResizeTab* ytab, int areaFast,
                         int batch, cudaStream_t s) {
    const LevelGeom& D = g.lv[level];
    const LevelGeom& S = g.lv[level - 1];
    uint8_t* dPlane = D.plane + (long long)g.f0 * D.planeStride;
    const uint8_t* sPlane = S.plane + (long long)g.f0 * S.planeStride;
    const int gx = (D.pitch + 127) / 128, rpc = rows_per_cta(D.rows, gx, batch);
    dim3 grid(gx, (D.rows + rpc - 1) / rpc, batch);
    if (areaFast & 2)  // bit 1: every thread's source window fits 8 bytes -> word-based kernel
        k_pyramid_resize4<<<grid, 256, (size_t)rpc * sizeof(uint4), s>>>(dPlane, D.w, D.h, D.pitch, D.rows, D.planeStride, sPlane, S.pitch,
                                                 S.planeStride, xtab, ytab, rpc);
    else
        k_pyramid_resize<<<grid, 256, 0, s>>>(dPlane, D.w, D.h, D.pitch, D.rows, D.planeStride, sPlane, S.pitch,
                                                S.planeStride, xtab, ytab, areaFast & 1, rpc);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

// the TMA-tiled kernel (areaFast bit 2: the tile windows of this level fit the box); `maps` are the resize maps of the SOURCE levels
int launch_pyramid_level_tma(const FrameGeom& g, const TmaMaps& maps, int level, const ResizeTab* xtab, const ResizeTab* ytab,
                             const int* tileX0, const int* tileY0, int batch, cudaStream_t s, bool pdl) {
    const LevelGeom& D = g.lv[level];
    dim3 grid((D.pitch + 127) / 128, (D.rows + RZT_TH - 1) / RZT_TH, batch);
    static const int splitPref = []() { const char* e = getenv("ORBX_RZ_SPLIT"); return e ? atoi(e) : 6; }();
    const int split = pdl && splitPref >= 1 && RZT_TH % splitPref == 0 ? splitPref : 1;
    grid.y *= split;
    if (pdl) {  // programmatic dependent launch: may start while the previous kernel of the stream still runs (the kernel waits itself)
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = grid;
        cfg.blockDim = dim3(32);
        cfg.stream = s;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        ORBX_CUDA(cudaLaunchKernelEx(&cfg, k_pyramid_resize_tma, maps, level - 1, D.plane, D.w, D.h, D.pitch, D.rows, D.planeStride, xtab, ytab,
                                     tileX0, tileY0, g.f0, 1, split));
    } else {
        k_pyramid_resize_tma<<<grid, 32, 0, s>>>(maps, level - 1, D.plane, D.w, D.h, D.pitch, D.rows, D.planeStride, xtab, ytab, tileX0, tileY0,
                                                  g.f0, 0, 1);
    }
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
