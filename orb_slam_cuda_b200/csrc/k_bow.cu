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
    Top4 mine;
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

int launch_search_by_bow(int nItems, const int* nodeOf, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                         const uint8_t* valid1, const uint8_t* desc1, const float* angle1, int n2, const uint8_t* valid2,
                         const uint8_t* desc2, const float* angle2, const int* candOff, uint32_t* cand, int* candCnt, int thDist,
                         float nnratio, int checkOri, int* match12, int* match21, int* nmatches, cudaStream_t s) {
    const size_t smem = 128 + (size_t)n2 * 4 + (size_t)2 * (nItems + 1) + (size_t)n2 + (size_t)nItems + 16;
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
    k_bow_resolve<<<1, BOW_RESOLVE_THREADS, smem, s>>>(nItems, idx1, angle1, angle2, n1, n2, candOff, cand, candCnt, thDist, nnratio, checkOri, match12,
                                        match21, nmatches);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
