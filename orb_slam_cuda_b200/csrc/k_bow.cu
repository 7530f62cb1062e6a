// csrc/k_bow.cu — ORBmatcher::SearchByBoW, both overloads (reference src/ORBmatcher.cc:159-288 KeyFrame vs Frame, :522-655
// KeyFrame vs KeyFrame).  The DBoW2 FeatureVector merge stays on the host (it is a walk over two std::maps); what arrives
// here is the list of common vocabulary nodes, node k holding feature indices idx1[start1[k] .. start1[k+1]) of the first
// and idx2[start2[k] ..) of the second argument.
//
// Same decomposition as k_search.cu: the Hamming distances between the features of a node do not depend on the loop
// order, the assignment does (a second-argument feature that received a match is skipped from then on, :205-206 /
// :579), so
//   k_bow_candidates (warp per first-argument feature, in the reference's iteration order): distances to the valid
//                    second-argument features of its node, stored in visiting order as (dist << 16 | i2)
//   k_bow_resolve    (1 warp): replays the loop — two smallest free candidates, threshold, ratio test against 256 when
//                    there is no second (:200-202), assignment, rotation histogram, ComputeThreeMaxima, removals; 32 features
//                    are scanned in parallel, then decided in order (resolve_common.cuh).
#include <algorithm>

#include "orbx_internal.h"
#include "resolve_common.cuh"

namespace orbx {

__device__ __forceinline__ int hamming32(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// item = position in idx1 (node-major).  candOff[item] = first slot of the item's candidate list (prefix of the node sizes).
__global__ void __launch_bounds__(256) k_bow_candidates(int nItems, const int* __restrict__ nodeOf, const uint32_t* __restrict__ idx1,
                                                         const int* __restrict__ start2, const uint32_t* __restrict__ idx2,
                                                         const uint8_t* __restrict__ valid1, const uint4* __restrict__ desc1,
                                                         const uint8_t* __restrict__ valid2, const uint4* __restrict__ desc2,
                                                         const int* __restrict__ candOff, uint32_t* __restrict__ cand,
                                                         int* __restrict__ candCnt) {
    const int lane = threadIdx.x & 31;
    const int item = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (item >= nItems) return;
    const int i1 = (int)idx1[item];
    int count = 0;
    TopK mine;
    mine.clear();
    if (valid1[i1]) {
        const int node = nodeOf[item];
        const uint4 a0 = desc1[2 * i1], a1 = desc1[2 * i1 + 1];
        uint32_t* out = cand + candOff[item];
        const int b = start2[node], e = start2[node + 1];
        for (int base = b; base < e; base += 32) {
            const int p = base + lane;
            bool ok = false;
            int i2 = 0, dist = 0;
            if (p < e) {
                i2 = (int)idx2[p];
                ok = valid2[i2] != 0;
                if (ok) dist = hamming32(a0, a1, desc2[2 * i2], desc2[2 * i2 + 1]);
            }
            const uint32_t m = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                const int o = count + __popc(m & ((1u << lane) - 1));
                out[o] = ((uint32_t)dist << 16) | (uint32_t)i2;
                mine.push(((uint32_t)dist << 16) | (uint32_t)o, (uint32_t)i2, true);
            }
            count += __popc(m);
        }
    }
    if (lane == 0) candCnt[item] = count;
    warp_store_top(mine, lane, candCnt, nItems, item);
}

// One CTA: warp 0 replays the loop on shared memory only (taken[n2] uint8, evI2[nItems] uint16 = the keypoint an item
// accepted); all warps clear the outputs before and, after it, write match12 / match21, bin the rotations and apply
// ComputeThreeMaxima.  Dynamic shared memory: hist[32] int, claim[n2] int, evI2[nItems] uint16, taken[n2] uint8, evBin[nItems] int8.
constexpr int BOW_RESOLVE_THREADS = 256;
__global__ void __launch_bounds__(BOW_RESOLVE_THREADS) k_bow_resolve(int nItems, const uint32_t* __restrict__ idx1,
                                                                     const float* __restrict__ angle1, const float* __restrict__ angle2,
                                                                     int n1, int n2, const int* __restrict__ candOff,
                                                                     const uint32_t* __restrict__ cand, const int* __restrict__ candCnt,
                                                                     int thDist, float nnratio, int checkOri, int* __restrict__ match12,
                                                                     int* __restrict__ match21, int* __restrict__ nmatchesOut) {
    extern __shared__ __align__(16) uint8_t sm[];
    int* hist = reinterpret_cast<int*>(sm);
    int* claim = hist + 32;
    uint16_t* evI2 = reinterpret_cast<uint16_t*>(claim + n2);
    uint8_t* taken = reinterpret_cast<uint8_t*>(evI2 + nItems + (nItems & 1));
    signed char* evBin = reinterpret_cast<signed char*>(taken + n2);
    const int tid = threadIdx.x, lane = tid & 31;
    for (int i = tid; i < n2; i += BOW_RESOLVE_THREADS) { taken[i] = 0; claim[i] = RS_CLAIM_FREE; match21[i] = -1; }
    for (int i = tid; i < n1; i += BOW_RESOLVE_THREADS) match12[i] = -1;
    for (int i = tid; i < nItems; i += BOW_RESOLVE_THREADS) { evI2[i] = 0xFFFFu; evBin[i] = -1; }
    if (tid < 32) hist[tid] = 0;
    __syncthreads();
    if (tid < 32) {
        auto isFree = [&](uint32_t i2, uint32_t) { return taken[i2] == 0; };
        int cntNext;
        TopK topNext;
        fetch_feature(candCnt, nItems, lane, cntNext, topNext);
        for (int base = 0; base < nItems; base += 32) {
            // lane = feature base + lane; its best candidates come from k_bow_candidates (resolve_common.cuh)
            const int item = base + lane;
            const int cnt = cntNext;
            const TopK top = topNext;
            fetch_feature(candCnt, nItems, item + 32, cntNext, topNext);
            const uint32_t* c = cand;  // only the rescan needs the list
            // decide + commit rounds, shared memory only
            resolve_rounds(
                __ballot_sync(0xffffffffu, cnt > 0), lane, claim,
                [&](uint32_t& b1, uint32_t& i2a, uint32_t& b2, uint32_t& i2b) {
                    if (!top.pick(isFree, 2, b1, i2a, b2, i2b)) rescan_list(c + candOff[item], cnt, isFree, b1, i2a, b2, i2b);
                    const int best = (int)(b1 >> 16);
                    const int second = b2 == RS_NONE ? 256 : (int)(b2 >> 16);
                    const bool near = b1 != RS_NONE && best <= thDist;
                    const bool acc = near && (float)best < __fmul_rn(nnratio, (float)second);
                    // what an earlier lane of the round can still change: nothing once the best is too far (it only gets
                    // farther); not the second of an accepted match (the second only grows, the ratio test keeps passing)
                    if (!near) b1 = RS_NONE;
                    if (!near || acc) b2 = RS_NONE;
                    return acc;
                },
                [&](uint32_t, uint32_t i2a, uint32_t, uint32_t) {
                    taken[i2a] = 1;
                    evI2[item] = (uint16_t)i2a;
                });
        }
    }
    __syncthreads();
    // every keypoint is accepted at most once (taken), so the writes below do not collide
    const float factor = 1.0f / 30.0f;
    int total = 0;
    for (int item = tid; item < nItems; item += BOW_RESOLVE_THREADS) {
        const int i2 = evI2[item];
        if (i2 == 0xFFFF) continue;
        const int i1 = (int)idx1[item];
        match12[i1] = i2;
        match21[i2] = i1;
        ++total;
        if (checkOri) {
            float rot = __fsub_rn(angle1[i1], angle2[i2]);
            if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
            int bin = (int)roundf(__fmul_rn(rot, factor));
            if (bin == 30) bin = 0;
            bin = min(max(bin, 0), 29);
            evBin[item] = (signed char)bin;
            atomicAdd(&hist[bin], 1);
        }
    }
    __syncthreads();
    if (checkOri) {
        int ind1, ind2, ind3;
        three_maxima(hist, ind1, ind2, ind3);
        for (int item = tid; item < nItems; item += BOW_RESOLVE_THREADS) {
            const int b = evBin[item];
            if (b >= 0 && b != ind1 && b != ind2 && b != ind3) {
                match12[(int)idx1[item]] = -2;
                match21[evI2[item]] = -2;
                --total;
            }
        }
    }
    total = __reduce_add_sync(0xffffffffu, total);
    __syncthreads();  // hist is reused for the count
    if (tid == 0) hist[0] = 0;
    __syncthreads();
    if (lane == 0 && total) atomicAdd(&hist[0], total);
    __syncthreads();
    if (tid == 0) *nmatchesOut = hist[0];
}

// ---- SearchForTriangulation (src/ORBmatcher.cc:657-823; LocalMapping::CreateNewMapPoints, src/LocalMapping.cc:248ff) ------
// vbMatched2 is declared (:683) but never set, so the iterations are independent: feature idx1 takes, among the features of its
// vocabulary node in the other keyframe that are free (no map point, stereo filter), within TH_LOW, far enough from the epipole
// (:739-745) and close enough to the epipolar line (CheckDistEpipolarLine, :140-157), the one at the smallest distance — the LAST
// such one in visiting order, because `dist > bestDist` lets an equal distance replace the best (:734, :751-755).
struct SftParams {
    float f[9];  // F12, row major
    float ex, ey;
    int checkOri;
};

__global__ void __launch_bounds__(256) k_sft_match(int nItems, const int* __restrict__ nodeOf, const uint32_t* __restrict__ idx1,
                                                    const int* __restrict__ start2, const uint32_t* __restrict__ idx2,
                                                    const orbx_keypoint* __restrict__ k1, const uint4* __restrict__ desc1,
                                                    const uint8_t* __restrict__ valid1, const uint8_t* __restrict__ stereo1,
                                                    const orbx_keypoint* __restrict__ k2, const uint4* __restrict__ desc2,
                                                    const uint8_t* __restrict__ valid2, const uint8_t* __restrict__ stereo2,
                                                    const float* __restrict__ scaleFactors2, const float* __restrict__ levelSigma2,
                                                    const SftParams P, int* __restrict__ match12, signed char* __restrict__ evBin,
                                                    int* __restrict__ hist) {
    const int lane = threadIdx.x & 31;
    const int item = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (item >= nItems) return;
    const int i1 = (int)idx1[item];
    uint32_t best = RS_NONE;  // dist << 16 | (0xFFFF - position): smallest distance, last position
    uint32_t bestI2 = 0;
    if (valid1[i1]) {
        const orbx_keypoint kp1 = k1[i1];
        const bool st1 = stereo1[i1] != 0;
        // epipolar line in the second image l = x1' F12 = [a b c] (:143-145)
        const float a = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, P.f[0]), __fmul_rn(kp1.y, P.f[3])), P.f[6]);
        const float b = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, P.f[1]), __fmul_rn(kp1.y, P.f[4])), P.f[7]);
        const float c = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, P.f[2]), __fmul_rn(kp1.y, P.f[5])), P.f[8]);
        const float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
        const uint4 a0 = desc1[2 * i1], a1 = desc1[2 * i1 + 1];
        const int node = nodeOf[item];
        const int s = start2[node], e = start2[node + 1];
        for (int p = s + lane; p < e; p += 32) {
            const int i2 = (int)idx2[p];
            if (!valid2[i2]) continue;
            const int dist = hamming32(a0, a1, desc2[2 * i2], desc2[2 * i2 + 1]);
            if (dist > 50) continue;  // TH_LOW
            const orbx_keypoint kp2 = k2[i2];
            if (!st1 && !stereo2[i2]) {
                const float dx = __fsub_rn(P.ex, kp2.x), dy = __fsub_rn(P.ey, kp2.y);
                if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.0f, scaleFactors2[kp2.octave])) continue;
            }
            if (den == 0.0f) continue;
            const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, kp2.x), __fmul_rn(b, kp2.y)), c);
            const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
            if (!((double)dsqr < __dmul_rn(3.84, (double)levelSigma2[kp2.octave]))) continue;
            const uint32_t key = ((uint32_t)dist << 16) | (uint32_t)(0xFFFF - (p - s));
            if (key < best) { best = key; bestI2 = (uint32_t)i2; }
        }
    }
    const uint32_t m = __reduce_min_sync(0xffffffffu, best);
    const unsigned who = __ballot_sync(0xffffffffu, best == m && m != RS_NONE);
    if (!who) return;
    const int i2 = (int)__shfl_sync(0xffffffffu, bestI2, __ffs(who) - 1);
    if (lane == 0) {
        match12[i1] = i2;
        if (P.checkOri) {
            const int bin = rotation_bin(k1[i1].angle, k2[i2].angle);
            evBin[item] = (signed char)bin;
            atomicAdd(&hist[bin], 1);
        }
    }
}

// ComputeThreeMaxima + removals (:789-806) and the count; one CTA.
__global__ void __launch_bounds__(256) k_sft_finish(int nItems, const uint32_t* __restrict__ idx1, const signed char* __restrict__ evBin,
                                                     const int* __restrict__ hist, int checkOri, int n1, int* __restrict__ match12,
                                                     int* __restrict__ nmatchesOut) {
    __shared__ int total;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    if (checkOri) {
        int ind1, ind2, ind3;
        three_maxima(hist, ind1, ind2, ind3);
        for (int item = threadIdx.x; item < nItems; item += blockDim.x) {
            const int b = evBin[item];
            if (b >= 0 && b != ind1 && b != ind2 && b != ind3) match12[(int)idx1[item]] = -1;
        }
    }
    __syncthreads();
    int mine = 0;
    for (int i = threadIdx.x; i < n1; i += blockDim.x) mine += match12[i] >= 0 ? 1 : 0;
    mine = __reduce_add_sync(0xffffffffu, mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&total, mine);
    __syncthreads();
    if (threadIdx.x == 0) *nmatchesOut = total;
}

int launch_search_for_triangulation(int nItems, const int* nodeOf, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                                    const orbx_keypoint* k1, const uint8_t* desc1, const uint8_t* valid1, const uint8_t* stereo1, int n2,
                                    const orbx_keypoint* k2, const uint8_t* desc2, const uint8_t* valid2, const uint8_t* stereo2,
                                    const float* scaleFactors2, const float* levelSigma2, const float F12[9], float ex, float ey,
                                    int checkOri, int* match12, signed char* evBin, int* hist, int* nmatches, cudaStream_t s) {
    (void)n2;
    SftParams P;
    for (int i = 0; i < 9; ++i) P.f[i] = F12[i];
    P.ex = ex; P.ey = ey; P.checkOri = checkOri;
    ORBX_CUDA(cudaMemsetAsync(match12, 0xFF, (size_t)n1 * 4, s));
    ORBX_CUDA(cudaMemsetAsync(evBin, 0xFF, (size_t)std::max(nItems, 1), s));
    ORBX_CUDA(cudaMemsetAsync(hist, 0, 32 * 4, s));
    if (nItems > 0) {
        k_sft_match<<<(nItems + 7) / 8, 256, 0, s>>>(nItems, nodeOf, idx1, start2, idx2, k1, reinterpret_cast<const uint4*>(desc1), valid1,
                                                       stereo1, k2, reinterpret_cast<const uint4*>(desc2), valid2, stereo2, scaleFactors2,
                                                       levelSigma2, P, match12, evBin, hist);
        count_launch();
    }
    k_sft_finish<<<1, 256, 0, s>>>(nItems, idx1, evBin, hist, checkOri, n1, match12, nmatches);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

int launch_search_by_bow(int nItems, const int* nodeOf, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                         const uint8_t* valid1, const uint8_t* desc1, const float* angle1, int n2, const uint8_t* valid2,
                         const uint8_t* desc2, const float* angle2, const int* candOff, uint32_t* cand, int* candCnt, int thDist,
                         float nnratio, int checkOri, int* match12, int* match21, int* nmatches, cudaStream_t s) {
    const size_t smem = 128 + (size_t)n2 * 4 + (size_t)2 * (nItems + 1) + (size_t)n2 + (size_t)nItems + 16;
    if (smem > 200 * 1024) {
        set_error("SearchByBoW: %d + %d features do not fit the resolve kernel's shared memory", nItems, n2);
        return ORBX_E_INVALID;
    }
    static size_t configured[ORBX_MAX_DEVICES] = {};
    if (int rcAttr = ensure_dynamic_smem((const void*)k_bow_resolve, smem, configured)) return rcAttr;
    if (nItems > 0) {
        k_bow_candidates<<<(nItems + 7) / 8, 256, 0, s>>>(nItems, nodeOf, idx1, start2, idx2, valid1, reinterpret_cast<const uint4*>(desc1),
                                                            valid2, reinterpret_cast<const uint4*>(desc2), candOff, cand, candCnt);
        count_launch();
    }
    k_bow_resolve<<<1, BOW_RESOLVE_THREADS, smem, s>>>(nItems, idx1, angle1, angle2, n1, n2, candOff, cand, candCnt, thDist, nnratio, checkOri, match12,
                                        match21, nmatches);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
