// csrc/k_pyramid.cu — ORBextractor::ComputePyramid (reference src/ORBextractor.cc:1837-1862, CPU branch).
//
// Level 0 is the input frame with a 19-px BORDER_REFLECT_101 frame around it (copyMakeBorder, :1859-1860); level l
// is cv::resize(level l-1, INTER_LINEAR) followed by the same border (:1852-1855).  cv::resize on CV_8U is an
// 11-bit fixed-point separable bilinear filter: the per-column / per-row source indices and coefficients are
// computed on the host exactly as OpenCV does (double scale, float fx, cvRound(f*2048)) and handed in as tables;
// the kernel does the integer blend ((b0*(r0>>4))>>16) + ((b1*(r1>>4))>>16) + 2) >> 2.  Every thread produces 16
// bytes of one padded row (one 16-byte store); border bytes are produced by evaluating the interior formula at
// the reflected coordinate, so the plane is written exactly once and no second border pass is needed.
//
// Roofline: HBM-bound streaming (reads level l-1 once through L1/L2, writes level l once).
#include "orbx_internal.h"

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

// grid: (ceil(pitch/16 / 32), ceil(rows / 8), batch); block (32, 8)
__global__ void __launch_bounds__(256) k_pyramid_base(uint8_t* __restrict__ plane, int w, int h, int pitch, int rows,
                                                      long long planeStride, const uint8_t* __restrict__ frames,
                                                      long long stride, long long frameStride) {
    const int cg = blockIdx.x * 32 + threadIdx.x;
    const int py = blockIdx.y * 8 + threadIdx.y;
    if (cg * 16 >= pitch || py >= rows) return;
    const int f = blockIdx.z;
    const int iy = reflect101(py - EDGE, h);
    const uint8_t* __restrict__ src = frames + (long long)f * frameStride + (long long)iy * stride;
    const int px0 = cg * 16 - PADL;  // padded x of byte 0 of this group
    uint32_t out[4];
    if (px0 - EDGE >= 0 && px0 - EDGE + 15 < w) {  // fully interior: 16 consecutive source bytes
        const uint8_t* p = src + (px0 - EDGE);
        if ((((uintptr_t)p) & 3) == 0) {
            const uint32_t* p4 = reinterpret_cast<const uint32_t*>(p);
#pragma unroll
            for (int k = 0; k < 4; ++k) out[k] = __ldg(p4 + k);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                out[k] = (uint32_t)__ldg(p + 4 * k) | ((uint32_t)__ldg(p + 4 * k + 1) << 8) |
                         ((uint32_t)__ldg(p + 4 * k + 2) << 16) | ((uint32_t)__ldg(p + 4 * k + 3) << 24);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint32_t v = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int px = px0 + 4 * k + j;
                uint32_t b = 0;
                if (px >= 0 && px < w + 2 * EDGE) b = __ldg(src + reflect101(px - EDGE, w));
                v |= b << (8 * j);
            }
            out[k] = v;
        }
    }
    uint4* dst = reinterpret_cast<uint4*>(plane + (long long)f * planeStride + (long long)py * pitch + cg * 16);
    *dst = make_uint4(out[0], out[1], out[2], out[3]);
}

// grid: (ceil(dpitch/16 / 32), ceil(drows / 8), batch); block (32, 8)
__global__ void __launch_bounds__(256) k_pyramid_resize(uint8_t* __restrict__ dplane, int dw, int dh, int dpitch, int drows,
                                                        long long dStride, const uint8_t* __restrict__ splane, int spitch,
                                                        long long sStride, const ResizeTab* __restrict__ xtab,
                                                        const ResizeTab* __restrict__ ytab, int areaFast) {
    const int cg = blockIdx.x * 32 + threadIdx.x;
    const int py = blockIdx.y * 8 + threadIdx.y;
    if (cg * 16 >= dpitch || py >= drows) return;
    const int f = blockIdx.z;
    const int iy = reflect101(py - EDGE, dh);
    const ResizeTab yt = ytab[iy];
    const uint8_t* __restrict__ sInt = splane + (long long)f * sStride + (long long)EDGE * spitch + XOFF;
    const uint8_t* __restrict__ r0p = sInt + (long long)yt.s0 * spitch;
    const uint8_t* __restrict__ r1p = sInt + (long long)yt.s1 * spitch;
    const int b0 = yt.c0, b1 = yt.c1;
    const int px0 = cg * 16 - PADL;
    uint32_t out[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        uint32_t v = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int px = px0 + 4 * k + j;
            uint32_t b = 0;
            if (px >= 0 && px < dw + 2 * EDGE) {
                const int ix = reflect101(px - EDGE, dw);
                const ResizeTab xt = xtab[ix];
                const int p00 = __ldg(r0p + xt.s0), p01 = __ldg(r0p + xt.s1);
                const int p10 = __ldg(r1p + xt.s0), p11 = __ldg(r1p + xt.s1);
                int val;
                if (areaFast) {
                    val = (p00 + p01 + p10 + p11 + 2) >> 2;
                } else {
                    const int h0 = p00 * xt.c0 + p01 * xt.c1;
                    const int h1 = p10 * xt.c0 + p11 * xt.c1;
                    val = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
                    val = min(max(val, 0), 255);
                }
                b = (uint32_t)val;
            }
            v |= b << (8 * j);
        }
        out[k] = v;
    }
    uint4* dst = reinterpret_cast<uint4*>(dplane + (long long)f * dStride + (long long)py * dpitch + cg * 16);
    *dst = make_uint4(out[0], out[1], out[2], out[3]);
}

int launch_pyramid_base(const FrameGeom& g, const uint8_t* dFrames, int batch, size_t stride, size_t frameStride,
                        cudaStream_t s) {
    const LevelGeom& L = g.lv[0];
    dim3 block(32, 8), grid((L.pitch / 16 + 31) / 32, (L.rows + 7) / 8, batch);
    k_pyramid_base<<<grid, block, 0, s>>>(L.plane, L.w, L.h, L.pitch, L.rows, L.planeStride, dFrames, (long long)stride,
                                          (long long)frameStride);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

int launch_pyramid_level(const FrameGeom& g, int level, const ResizeTab* xtab, const ResizeTab* ytab, int areaFast,
                         int batch, cudaStream_t s) {
    const LevelGeom& D = g.lv[level];
    const LevelGeom& S = g.lv[level - 1];
    dim3 block(32, 8), grid((D.pitch / 16 + 31) / 32, (D.rows + 7) / 8, batch);
    k_pyramid_resize<<<grid, block, 0, s>>>(D.plane, D.w, D.h, D.pitch, D.rows, D.planeStride, S.plane, S.pitch,
                                            S.planeStride, xtab, ytab, areaFast);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
