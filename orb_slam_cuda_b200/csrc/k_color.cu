// csrc/k_color.cu — cvtColor(..., CV_RGB2GRAY / CV_BGR2GRAY / CV_RGBA2GRAY / CV_BGRA2GRAY) as Tracking::GrabImage* applies it before
// the extractor (reference src/Tracking.cc:221-246, :261-274, :294-307).  SURVEY.md §8f rank 4.
//
// 8-bit fixed point, as OpenCV computes it: gray = (R*cr + G*cg + B*cb + (1 << (shift-1))) >> shift with
//   shift 15: cr, cg, cb = 9798, 19235, 3735   (OpenCV 4.x, what cv2 4.13 in this image does — the oracle is pinned to it)
//   shift 14: cr, cg, cb = 4899,  9617, 1868   (OpenCV 2.4 / 3.x, the versions the reference's CMakeLists asks for)
// The two differ by one grey level on ~0.3 % of the pixels; the caller picks.  HBM bound: 3-4 bytes in, 1 byte out per pixel; one
// thread converts 4 pixels (12 or 16 contiguous source bytes, one 32-bit store).
#include "orbx_internal.h"

namespace orbx {

__global__ void __launch_bounds__(256) k_gray(const uint8_t* __restrict__ src, long long srcStride, int w, int h, int channels, int bgr,
                                               int shift, uint8_t* __restrict__ dst, long long dstStride) {
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x4 >= w || y >= h) return;
    const int cr = shift == 15 ? 9798 : 4899, cg = shift == 15 ? 19235 : 9617, cb = shift == 15 ? 3735 : 1868;
    const int c0 = bgr ? cb : cr, c2 = bgr ? cr : cb, half = 1 << (shift - 1);
    const uint8_t* s = src + (size_t)y * srcStride + (size_t)x4 * channels;
    uint8_t* d = dst + (size_t)y * dstStride + x4;
    uint32_t out = 0;
    const int n = min(4, w - x4);
    for (int k = 0; k < n; ++k) {
        const uint8_t* p = s + k * channels;
        out |= (uint32_t)((p[0] * c0 + p[1] * cg + p[2] * c2 + half) >> shift) << (8 * k);
    }
    if (n == 4 && ((reinterpret_cast<uintptr_t>(d) & 3) == 0)) {
        *reinterpret_cast<uint32_t*>(d) = out;
    } else {
        for (int k = 0; k < n; ++k) d[k] = (uint8_t)(out >> (8 * k));
    }
}

int launch_gray(const uint8_t* src, long long srcStride, int w, int h, int channels, int bgr, int shift, uint8_t* dst, long long dstStride,
                cudaStream_t s) {
    if (w <= 0 || h <= 0) return ORBX_OK;
    const int threads = 256, perRow = (w + 3) / 4;
    k_gray<<<dim3((perRow + threads - 1) / threads, h), threads, 0, s>>>(src, srcStride, w, h, channels, bgr, shift, dst, dstStride);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
