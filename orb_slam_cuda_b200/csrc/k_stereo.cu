// csrc/k_stereo.cu — Frame::ComputeStereoMatches (reference src/Frame.cc:465-639).
//
// k_stereo_match (one warp per left keypoint):
//   1. candidates = right keypoints whose row band [floor(y - r), ceil(y + r)], r = 2 * mvScaleFactors[octave]
//      (:481-491) holds row (int)vL, with octave within +-1 of the left octave and uR in [uL - maxD, uL] (:515-527);
//      the vRowIndices table of the reference lists them in ascending iR, and `dist < bestDist` is strict, so the
//      winner is the smallest (distance, iR) pair below TH_HIGH — found here by lanes striding over ALL right
//      keypoints (a band test is a handful of compares; no table is built) and a warp min-reduction.
//   2. if bestDist < thOrbDist = 75: 11 x 11 centre-subtracted L1 SAD of the left patch against the right level at the
//      11 shifts -5..+5 (:551-591).  cv::norm(NORM_L1) of integer-valued float patches is an exact integer sum
//      (<= 121 * 510), evaluated in int32; first minimal shift wins (strict '<' against the int-truncated best).
//   3. parabola sub-pixel fit, disparity and depth in float32 with individually rounded operations (:593-621).
// k_stereo_median (one CTA): thDist = 1.5f * 1.4f * median of the accepted SADs, the median being element size/2 of
// the (SAD, iL)-sorted list (:625-638); matches with SAD >= thDist are reset to -1.
#include "orbx_internal.h"

namespace orbx {

__device__ __forceinline__ int hamming32B_s(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

constexpr int ST_WARPS = 8;

__global__ void __launch_bounds__(ST_WARPS * 32) k_stereo_match(const __grid_constant__ StereoLevels lv,
                                                                 const orbx_keypoint* __restrict__ kL, int nL,
                                                                 const uint4* __restrict__ dL,
                                                                 const orbx_keypoint* __restrict__ kR, int nR,
                                                                 const uint4* __restrict__ dR, float mbf, float maxD,
                                                                 float* __restrict__ uRight, float* __restrict__ depth,
                                                                 int* __restrict__ sad) {
    const int lane = threadIdx.x & 31;
    const int iL = blockIdx.x * ST_WARPS + (threadIdx.x >> 5);
    if (iL >= nL) return;
    float outU = -1.0f, outZ = -1.0f;
    int outSad = -1;
    const orbx_keypoint kp = kL[iL];
    const int levelL = kp.octave;
    const float vL = kp.y, uL = kp.x;
    const float minU = __fsub_rn(uL, maxD), maxU = uL;  // minD = 0
    const int row = (int)vL;
    uint32_t best = (100u << 16) | 0xFFFFu;  // TH_HIGH, no index
    if (!(maxU < 0.f) && levelL >= 0 && levelL < lv.nlevels) {
        const uint4 a0 = dL[2 * iL], a1 = dL[2 * iL + 1];
        for (int iR = lane; iR < nR; iR += 32) {
            const orbx_keypoint kr = kR[iR];
            if (kr.octave < levelL - 1 || kr.octave > levelL + 1 || kr.octave < 0 || kr.octave >= lv.nlevels) continue;
            const float r = __fmul_rn(2.0f, lv.scale[kr.octave]);
            const int maxr = (int)ceilf(__fadd_rn(kr.y, r)), minr = (int)floorf(__fsub_rn(kr.y, r));
            if (row < minr || row > maxr) continue;
            if (!(kr.x >= minU && kr.x <= maxU)) continue;
            const uint32_t dist = (uint32_t)hamming32B_s(a0, a1, dR[2 * iR], dR[2 * iR + 1]);
            best = min(best, (dist << 16) | (uint32_t)iR);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    const int bestDist = (int)(best >> 16);
    if (bestDist < 75) {  // thOrbDist = (TH_HIGH + TH_LOW) / 2
        const int bestIdxR = (int)(best & 0xFFFFu);
        const float uR0 = kR[bestIdxR].x;
        const float sf = lv.invScale[levelL];
        const float scaleduL = roundf(__fmul_rn(kp.x, sf)), scaledvL = roundf(__fmul_rn(kp.y, sf));
        const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
        const int w = 5, L = 5;
        const float iniu = __fsub_rn(__fadd_rn(scaleduR0, (float)L), (float)w);
        const float endu = __fadd_rn(__fadd_rn(__fadd_rn(scaleduR0, (float)L), (float)w), 1.0f);
        if (!(iniu < 0.f || endu >= (float)lv.widthR[levelL])) {
            const uint8_t* PL = lv.left[levelL];
            const uint8_t* PR = lv.right[levelL];
            const int sL = lv.pitchL[levelL], sR = lv.pitchR[levelL];
            const int y0 = (int)__fsub_rn(scaledvL, (float)w), xl0 = (int)__fsub_rn(scaleduL, (float)w);
            const int cL = PL[(long long)(y0 + w) * sL + xl0 + w];
            int pl[4], py[4], px[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int i = lane + 32 * k;
                py[k] = i / 11;
                px[k] = i - py[k] * 11;
                pl[k] = i < 121 ? (int)PL[(long long)(y0 + py[k]) * sL + xl0 + px[k]] - cL : 0;
            }
            int bestSad = 0x7fffffff, bestInc = 0, d[11];
#pragma unroll
            for (int inc = -L; inc <= L; ++inc) {
                const int xr0 = (int)__fsub_rn(__fadd_rn(scaleduR0, (float)inc), (float)w);
                const int cR = PR[(long long)(y0 + w) * sR + xr0 + w];
                int s = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (lane + 32 * k < 121) s += abs(pl[k] - ((int)PR[(long long)(y0 + py[k]) * sR + xr0 + px[k]] - cR));
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                d[inc + L] = s;
                if (s < bestSad) { bestSad = s; bestInc = inc; }
            }
            if (bestInc != -L && bestInc != L) {
                float dist1 = 0.f, dist2 = 0.f, dist3 = 0.f;
#pragma unroll
                for (int k = 1; k < 10; ++k)
                    if (k == bestInc + L) { dist1 = (float)d[k - 1]; dist2 = (float)d[k]; dist3 = (float)d[k + 1]; }
                const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2)));
                const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3), den);
                if (!(deltaR < -1.f || deltaR > 1.f)) {
                    float bestuR = __fmul_rn(lv.scale[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestInc), deltaR));
                    float disparity = __fsub_rn(uL, bestuR);
                    if (disparity >= 0.f && disparity < maxD) {
                        if (disparity <= 0.f) {
                            disparity = 0.01f;                                   // disparity = 0.01 (double -> float)
                            bestuR = (float)__dsub_rn((double)uL, 0.01);         // bestuR = uL - 0.01 (in double)
                        }
                        outZ = __fdiv_rn(mbf, disparity);
                        outU = bestuR;
                        outSad = bestSad;
                    }
                }
            }
        }
    }
    if (lane == 0) {
        uRight[iL] = outU;
        depth[iL] = outZ;
        sad[iL] = outSad;
    }
}

// One CTA.  sad[i] >= 0 marks an accepted match.
__global__ void __launch_bounds__(1024) k_stereo_median(int n, const int* __restrict__ sad, float* __restrict__ uRight,
                                                         float* __restrict__ depth) {
    __shared__ int m, med;
    if (threadIdx.x == 0) { m = 0; med = -1; }
    __syncthreads();
    int mine = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) mine += sad[i] >= 0 ? 1 : 0;
    if (mine) atomicAdd(&m, mine);
    __syncthreads();
    const int cnt = m;
    if (cnt == 0) return;
    const int target = cnt / 2;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int s = sad[i];
        if (s < 0) continue;
        int rank = 0;
        for (int j = 0; j < n; ++j) {
            const int t = sad[j];
            rank += (t >= 0 && (t < s || (t == s && j < i))) ? 1 : 0;
        }
        if (rank == target) med = s;
    }
    __syncthreads();
    const float thDist = __fmul_rn(1.5f * 1.4f, (float)med);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int s = sad[i];
        if (s >= 0 && !((float)s < thDist)) {
            uRight[i] = -1.f;
            depth[i] = -1.f;
        }
    }
}

int launch_stereo(const StereoLevels& lv, const orbx_keypoint* kL, int nL, const uint8_t* dL, const orbx_keypoint* kR, int nR,
                  const uint8_t* dR, float mbf, float maxD, float* uRight, float* depth, int* sad, cudaStream_t s) {
    if (nL <= 0) return ORBX_OK;
    k_stereo_match<<<(nL + ST_WARPS - 1) / ST_WARPS, ST_WARPS * 32, 0, s>>>(lv, kL, nL, reinterpret_cast<const uint4*>(dL), kR, nR,
                                                                             reinterpret_cast<const uint4*>(dR), mbf, maxD, uRight,
                                                                             depth, sad);
    count_launch();
    k_stereo_median<<<1, 1024, 0, s>>>(nL, sad, uRight, depth);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
