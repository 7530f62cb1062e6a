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
//                    there is no second (:200-202), assignment, rotation histogram, ComputeThreeMaxima, removals.
#include "orbx_internal.h"

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
            if (ok) out[count + __popc(m & ((1u << lane) - 1))] = ((uint32_t)dist << 16) | (uint32_t)i2;
            count += __popc(m);
        }
    }
    if (lane == 0) candCnt[item] = count;
}

// Dynamic shared memory: taken[n2] (uint8), evBin[nItems] (int8).  match12 / match21 live in global memory.
__global__ void __launch_bounds__(32) k_bow_resolve(int nItems, const uint32_t* __restrict__ idx1, const float* __restrict__ angle1,
                                                     const float* __restrict__ angle2, int n1, int n2,
                                                     const int* __restrict__ candOff, const uint32_t* __restrict__ cand,
                                                     const int* __restrict__ candCnt, int thDist, float nnratio, int checkOri,
                                                     int* __restrict__ match12, int* __restrict__ match21,
                                                     int* __restrict__ nmatchesOut) {
    extern __shared__ __align__(16) uint8_t sm[];
    uint8_t* taken = sm;
    signed char* evBin = reinterpret_cast<signed char*>(sm + n2);
    const int lane = threadIdx.x;
    for (int i = lane; i < n2; i += 32) { taken[i] = 0; match21[i] = -1; }
    for (int i = lane; i < n1; i += 32) match12[i] = -1;
    for (int i = lane; i < nItems; i += 32) evBin[i] = -1;
    __syncwarp();
    const float factor = 1.0f / 30.0f;
    int total = 0;
    for (int item = 0; item < nItems; ++item) {
        const int cnt = candCnt[item];
        if (cnt == 0) continue;
        const uint32_t* c = cand + candOff[item];
        uint32_t b1 = 0xFFFFFFFFu, b2 = 0xFFFFFFFFu;
        for (int p = lane; p < cnt; p += 32) {
            const uint32_t e = c[p];
            if (taken[e & 0xFFFFu]) continue;
            const uint32_t key = ((e >> 16) << 16) | (uint32_t)p;
            if (key < b1) { b2 = b1; b1 = key; }
            else if (key < b2) b2 = key;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const uint32_t o1 = __shfl_xor_sync(0xffffffffu, b1, o), o2 = __shfl_xor_sync(0xffffffffu, b2, o);
            const uint32_t lo = min(b1, o1), hi = max(b1, o1);
            b2 = min(hi, min(b2, o2));
            b1 = lo;
        }
        if (b1 == 0xFFFFFFFFu) continue;
        const int best = (int)(b1 >> 16);
        if (best > thDist) continue;
        const int second = b2 == 0xFFFFFFFFu ? 256 : (int)(b2 >> 16);
        if (!((float)best < __fmul_rn(nnratio, (float)second))) continue;
        if (lane == 0) {
            const int i1 = (int)idx1[item], i2 = (int)(c[b1 & 0xFFFFu] & 0xFFFFu);
            match12[i1] = i2;
            match21[i2] = i1;
            taken[i2] = 1;
            if (checkOri) {
                float rot = __fsub_rn(angle1[i1], angle2[i2]);
                if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
                int bin = (int)roundf(__fmul_rn(rot, factor));
                if (bin == 30) bin = 0;
                evBin[item] = (signed char)min(max(bin, 0), 29);
            } else {
                evBin[item] = 0;
            }
        }
        ++total;
        __syncwarp();
    }
    __syncwarp();
    if (checkOri) {
        int h = 0;
        for (int i = 0; i < nItems; ++i) h += (evBin[i] == lane) ? 1 : 0;
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int b = 0; b < 30; ++b) {  // ComputeThreeMaxima (:1601-1642)
            const int s = __shfl_sync(0xffffffffu, h, b);
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = b; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = b; }
            else if (s > max3) { max3 = s; ind3 = b; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
        int removed = 0;
        for (int i = lane; i < nItems; i += 32) {
            const int b = evBin[i];
            if (b >= 0 && b != ind1 && b != ind2 && b != ind3) {
                const int i1 = (int)idx1[i];
                match21[match12[i1]] = -2;
                match12[i1] = -2;
                ++removed;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) removed += __shfl_xor_sync(0xffffffffu, removed, o);
        total -= removed;
    }
    if (lane == 0) *nmatchesOut = total;
}

int launch_search_by_bow(int nItems, const int* nodeOf, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                         const uint8_t* valid1, const uint8_t* desc1, const float* angle1, int n2, const uint8_t* valid2,
                         const uint8_t* desc2, const float* angle2, const int* candOff, uint32_t* cand, int* candCnt, int thDist,
                         float nnratio, int checkOri, int* match12, int* match21, int* nmatches, cudaStream_t s) {
    const size_t smem = (size_t)n2 + (size_t)nItems + 16;
    if (smem > 200 * 1024) {
        set_error("SearchByBoW: %d + %d features do not fit the resolve kernel's shared memory", nItems, n2);
        return ORBX_E_INVALID;
    }
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        ORBX_CUDA(cudaFuncSetAttribute(k_bow_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    if (nItems > 0) {
        k_bow_candidates<<<(nItems + 7) / 8, 256, 0, s>>>(nItems, nodeOf, idx1, start2, idx2, valid1, reinterpret_cast<const uint4*>(desc1),
                                                            valid2, reinterpret_cast<const uint4*>(desc2), candOff, cand, candCnt);
        count_launch();
    }
    k_bow_resolve<<<1, 32, smem, s>>>(nItems, idx1, angle1, angle2, n1, n2, candOff, cand, candCnt, thDist, nnratio, checkOri, match12,
                                        match21, nmatches);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
