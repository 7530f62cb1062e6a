// csrc/k_search.cu — ORBmatcher::SearchForInitialization (reference src/ORBmatcher.cc:405-520) with
// Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea (src/Frame.cc:229-244, :381-391, :326-379) and
// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:1601-1642).
//
// The reference loop is sequential and stateful (vMatchedDistance[i2] <= dist skips a candidate, :444; an accepted
// match steals i2 from its previous owner, :463-467), so the result depends on the i1 order.  What does NOT depend on
// the order is the expensive part — which frame-2 keypoints fall into the window of i1, in which order the grid walk
// visits them, and their Hamming distances.  Hence three kernels:
//   k_sfi_grid       (1 CTA)   frame-2 keypoints binned into the 64 x 48 grid (cell = ix*48 + iy, keypoints of a
//                              cell in index order = push_back order), as a counting sort
//   k_sfi_candidates (warp/i1) for every octave-0 keypoint of frame 1: the window's cells are contiguous runs of the
//                              sorted array per grid column, lanes test box + octave and compute DescriptorDistance,
//                              survivors are appended in the reference's visiting order as (dist << 16 | i2)
//   k_sfi_resolve    (1 warp)  replays the i1 loop over the precomputed candidate lists with the mutable state
//                              (vMatchedDistance, vnMatches21) in shared memory, then the rotation histogram, the
//                              three maxima, the removal of the other bins and the vbPrevMatched update.
// best = first candidate (visiting order) of minimal distance, second = second smallest distance of the multiset:
// both are what the strict-'<' updates of :446-456 produce, and both are order independent given the skip set.
#include <cstring>

#include "orbx_internal.h"
#include "resolve_common.cuh"

namespace orbx {

constexpr int GRID_COLS = 64, GRID_ROWS = 48, GRID_CELLS = GRID_COLS * GRID_ROWS;  // include/Frame.h:37-38

__device__ __forceinline__ int hamming32B(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// bounds = {mnMinX, mnMaxX, mnMinY, mnMaxY}; sorted[] receives the keypoint indices ordered by (cell, index),
// cellStart[GRID_CELLS + 1] the run boundaries.  cellOf is scratch (n ints).  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024) k_sfi_grid(const orbx_keypoint* __restrict__ k2, int n, float minX, float minY,
                                                    float invW, float invH, int* __restrict__ cellOf,
                                                    int* __restrict__ sorted, int* __restrict__ cellStart) {
    __shared__ int cnt[GRID_CELLS + 1];
    __shared__ int fill[GRID_CELLS];
    __shared__ int warpSums[32];
    const int tid = threadIdx.x;
    for (int c = tid; c <= GRID_CELLS; c += blockDim.x) cnt[c] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) {
        // PosInGrid (src/Frame.cc:381-391): round((kp.pt.x - mnMinX) * mfGridElementWidthInv)
        const int px = (int)roundf(__fmul_rn(__fsub_rn(k2[i].x, minX), invW));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(k2[i].y, minY), invH));
        const bool in = px >= 0 && px < GRID_COLS && py >= 0 && py < GRID_ROWS;
        const int c = in ? px * GRID_ROWS + py : -1;
        cellOf[i] = c;
        if (in) atomicAdd(&cnt[c], 1);
    }
    __syncthreads();
    // exclusive scan of cnt[0..GRID_CELLS] (3073 entries, 1024 threads -> chunks of 4)
    {
        const int lane = tid & 31, wid = tid >> 5;
        const int chunk = (GRID_CELLS + 1 + blockDim.x - 1) / blockDim.x;
        const int b = min(tid * chunk, GRID_CELLS + 1), e = min(b + chunk, GRID_CELLS + 1);
        int local = 0;
        for (int i = b; i < e; ++i) local += cnt[i];
        int incl = local;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) warpSums[wid] = incl;
        __syncthreads();
        int off = 0;
        for (int w = 0; w < wid; ++w) off += warpSums[w];
        int run = off + incl - local;
        for (int i = b; i < e; ++i) {
            const int v = cnt[i];
            cnt[i] = run;
            run += v;
        }
        __syncthreads();
    }
    for (int c = tid; c <= GRID_CELLS; c += blockDim.x) cellStart[c] = cnt[c];
    // placement: any order inside the cell first (one atomic per keypoint), then every cell run is sorted by index, which is
    // the order Frame::AssignFeaturesToGrid pushes (src/Frame.cc:229-244) — runs hold a handful of keypoints
    for (int c = tid; c < GRID_CELLS; c += blockDim.x) fill[c] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) {
        const int c = cellOf[i];
        if (c >= 0) sorted[cnt[c] + atomicAdd(&fill[c], 1)] = i;
    }
    __syncthreads();
    for (int c = tid; c < GRID_CELLS; c += blockDim.x) {
        const int b = cnt[c], e = cnt[c + 1];
        for (int i = b + 1; i < e; ++i) {
            const int v = sorted[i];
            int j = i - 1;
            while (j >= b && sorted[j] > v) { sorted[j + 1] = sorted[j]; --j; }
            sorted[j + 1] = v;
        }
    }
}

// One warp per frame-1 keypoint.  cand[i1 * candCap + k] = (dist << 16) | i2 in visiting order, candCnt[i1] = count
// (0 for keypoints of octave > 0, which the reference skips at :421-423).
__global__ void __launch_bounds__(256) k_sfi_candidates(const orbx_keypoint* __restrict__ k1, int n1,
                                                         const uint4* __restrict__ desc1,
                                                         const orbx_keypoint* __restrict__ k2,
                                                         const uint4* __restrict__ desc2, const float* __restrict__ prev,
                                                         float minX, float minY, float invW, float invH, float r,
                                                         const int* __restrict__ sorted, const int* __restrict__ cellStart,
                                                         uint32_t* __restrict__ cand, int candCap, int* __restrict__ candCnt,
                                                         int* __restrict__ dStatus) {
    const int lane = threadIdx.x & 31;
    const int i1 = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i1 >= n1) return;
    int count = 0;
    TopK mine;
    mine.clear();
    if (k1[i1].octave <= 0) {
        const float x = prev[2 * i1], y = prev[2 * i1 + 1];
        // GetFeaturesInArea(x, y, r, 0, 0) (src/Frame.cc:326-379)
        const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, minX), r), invW)));
        const int cx1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, minX), r), invW)));
        const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, minY), r), invH)));
        const int cy1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, minY), r), invH)));
        if (cx0 < GRID_COLS && cx1 >= 0 && cy0 < GRID_ROWS && cy1 >= 0) {
            const uint4 a0 = desc1[2 * i1], a1 = desc1[2 * i1 + 1];
            uint32_t* out = cand + (size_t)i1 * candCap;
            for (int ix = cx0; ix <= cx1; ++ix) {
                const int b = cellStart[ix * GRID_ROWS + cy0], e = cellStart[ix * GRID_ROWS + cy1 + 1];
                for (int base = b; base < e; base += 32) {
                    const int p = base + lane;
                    bool ok = false;
                    int i2 = 0, dist = 0;
                    if (p < e) {
                        i2 = sorted[p];
                        const orbx_keypoint kp = k2[i2];
                        // bCheckLevels is true for (0, 0): octave must be 0; then |dx| < r && |dy| < r
                        ok = kp.octave == 0 && fabsf(__fsub_rn(kp.x, x)) < r && fabsf(__fsub_rn(kp.y, y)) < r;
                        if (ok) dist = hamming32B(a0, a1, desc2[2 * i2], desc2[2 * i2 + 1]);
                    }
                    const uint32_t m = __ballot_sync(0xffffffffu, ok);
                    if (ok) {
                        const int o = count + __popc(m & ((1u << lane) - 1));
                        if (o < candCap) out[o] = ((uint32_t)dist << 16) | (uint32_t)i2;
                        mine.push(((uint32_t)dist << 16) | (uint32_t)o, (uint32_t)i2, o < candCap);
                    }
                    count += __popc(m);
                }
            }
        }
    }
    warp_store_top(mine, lane, candCnt, n1, i1);
    if (lane == 0) {
        if (count > candCap) {
            atomicExch(dStatus, ORBX_E_OVERFLOW);
            count = candCap;
        }
        candCnt[i1] = count;
    }
}

// Warp 0 replays the sequential loop on shared memory only; all warps initialise, bin the rotations and write the result.
// Dynamic shared memory: hist[32] (int32), claim[n2] (int32), matches21[n2] (int32), matches12[n1] (int32), matchedDist[n2] (uint16, 0xFFFF =
// INT_MAX), ev2[n1] (uint16: the keypoint accepted by iteration i1, 0xFFFF = none), bin[n1] (int8).
__global__ void __launch_bounds__(RESOLVE_THREADS) k_sfi_resolve(const orbx_keypoint* __restrict__ k1, int n1,
                                                                 const orbx_keypoint* __restrict__ k2, int n2,
                                                                 const uint32_t* __restrict__ cand, int candCap,
                                                                 const int* __restrict__ candCnt, float nnratio, int checkOri,
                                                                 float* __restrict__ prev, int* __restrict__ matches12Out,
                                                                 int* __restrict__ nmatchesOut) {
    extern __shared__ __align__(16) uint8_t sm[];
    int* hist = reinterpret_cast<int*>(sm);
    int* claim = hist + 32;
    int* matches21 = claim + n2;
    int* matches12 = matches21 + n2;
    uint16_t* matchedDist = reinterpret_cast<uint16_t*>(matches12 + n1);
    uint16_t* ev2 = matchedDist + n2;
    signed char* binOf = reinterpret_cast<signed char*>(ev2 + n1);
    const int tid = threadIdx.x, lane = tid & 31;
    for (int i = tid; i < n2; i += RESOLVE_THREADS) { matches21[i] = -1; matchedDist[i] = 0xFFFFu; claim[i] = RS_CLAIM_FREE; }
    for (int i = tid; i < n1; i += RESOLVE_THREADS) { matches12[i] = -1; ev2[i] = 0xFFFFu; binOf[i] = -1; }
    if (tid < 32) hist[tid] = 0;
    __syncthreads();
    if (tid < 32) {
        // a candidate is skipped when the keypoint already holds a match at least as close (:444; 0xFFFF stands for INT_MAX,
        // dist <= 256); matchedDist only decreases, so the skipped set only grows (resolve_common.cuh)
        auto isFree = [&](uint32_t i2, uint32_t dist) { return (uint32_t)matchedDist[i2] > dist; };
        int cntNext;
        TopK topNext;
        fetch_feature(candCnt, n1, lane, cntNext, topNext);
        for (int base = 0; base < n1; base += 32) {
            const int i1 = base + lane;
            const int cnt = cntNext;
            const TopK top = topNext;  // key = dist << 16 | position: minimal key = minimal distance, first in visiting order
            fetch_feature(candCnt, n1, i1 + 32, cntNext, topNext);
            const uint32_t* c = cand + (size_t)min(i1, n1 - 1) * candCap;
            resolve_rounds(
                __ballot_sync(0xffffffffu, cnt > 0), lane, claim,
                [&](uint32_t& b1, uint32_t& i2a, uint32_t& b2, uint32_t& i2b) {
                    if (!top.pick(isFree, 2, b1, i2a, b2, i2b)) rescan_list(c, cnt, isFree, b1, i2a, b2, i2b);
                    const int bestDist = (int)(b1 >> 16);  // every candidate skipped: bestDist stays INT_MAX
                    const float second = b2 == RS_NONE ? 2147483648.0f /* (float)INT_MAX */ : (float)(int)(b2 >> 16);
                    const bool near = b1 != RS_NONE && bestDist <= 50;  // TH_LOW (:459)
                    const bool acc = near && (float)bestDist < __fmul_rn(second, nnratio);
                    // what an earlier lane of the round can still change: nothing once the best is too far (it only gets
                    // farther); not the second of an accepted match (the second only grows, the ratio test keeps passing)
                    if (!near) b1 = RS_NONE;
                    if (!near || acc) b2 = RS_NONE;
                    return acc;
                },
                [&](uint32_t b1, uint32_t i2a, uint32_t, uint32_t) {
                    const int owner = matches21[i2a];
                    if (owner >= 0) matches12[owner] = -1;
                    matches12[i1] = (int)i2a;
                    matches21[i2a] = i1;
                    matchedDist[i2a] = (uint16_t)(b1 >> 16);
                    ev2[i1] = (uint16_t)i2a;
                });
        }
    }
    __syncthreads();
    if (checkOri) {
        // histogram sizes count acceptance events, including matches stolen later (rotHist[bin].push_back(i1), :478)
        for (int i = tid; i < n1; i += RESOLVE_THREADS) {
            const int e = ev2[i];
            if (e == 0xFFFF) continue;
            const int bin = rotation_bin(k1[i].angle, k2[e].angle);
            binOf[i] = (signed char)bin;
            atomicAdd(&hist[bin], 1);
        }
        __syncthreads();
        int ind1, ind2, ind3;
        three_maxima(hist, ind1, ind2, ind3);
        for (int i = tid; i < n1; i += RESOLVE_THREADS) {
            const int b = binOf[i];
            if (b >= 0 && b != ind1 && b != ind2 && b != ind3) matches12[i] = -1;
        }
        __syncthreads();
    }
    int total = 0;
    for (int i = tid; i < n1; i += RESOLVE_THREADS) {
        const int m = matches12[i];
        matches12Out[i] = m;
        if (m >= 0) {
            ++total;
            prev[2 * i] = k2[m].x;  // vbPrevMatched update (:514-517)
            prev[2 * i + 1] = k2[m].y;
        }
    }
    total = __reduce_add_sync(0xffffffffu, total);
    __syncthreads();
    if (tid == 0) hist[0] = 0;
    __syncthreads();
    if (lane == 0 && total) atomicAdd(&hist[0], total);
    __syncthreads();
    // the reference's running nmatches (++ on accept, -- on steal / histogram removal) = surviving matches
    if (tid == 0) *nmatchesOut = hist[0];
}

size_t sfi_resolve_smem(int n1, int n2) { return 128 + (size_t)n2 * 8 + (size_t)n1 * 4 + (size_t)n2 * 2 + (size_t)n1 * 2 + (size_t)n1 + 16; }

int launch_search_for_initialization(const orbx_keypoint* k1, int n1, const uint8_t* desc1, const orbx_keypoint* k2, int n2,
                                     const uint8_t* desc2, const float bounds[4], float* prev, int* matches12, int window,
                                     float nnratio, int checkOri, int* cellOf, int* sorted, int* cellStart, uint32_t* cand,
                                     int candCap, int* candCnt, int* nmatches, int* dStatus, cudaStream_t s) {
    const float minX = bounds[0], maxX = bounds[1], minY = bounds[2], maxY = bounds[3];
    // mfGridElementWidthInv / HeightInv (src/Frame.cc:104-105)
    const float invW = (float)GRID_COLS / (maxX - minX), invH = (float)GRID_ROWS / (maxY - minY);
    const size_t smem = sfi_resolve_smem(n1, n2);
    if (smem > 200 * 1024) {
        set_error("SearchForInitialization: %d + %d keypoints do not fit the resolve kernel's shared memory", n1, n2);
        return ORBX_E_INVALID;
    }
    static size_t configured[ORBX_MAX_DEVICES] = {};
    if (int rcAttr = ensure_dynamic_smem((const void*)k_sfi_resolve, smem, configured)) return rcAttr;
    k_sfi_grid<<<1, 1024, 0, s>>>(k2, n2, minX, minY, invW, invH, cellOf, sorted, cellStart);
    count_launch();
    k_sfi_candidates<<<(n1 + 7) / 8, 256, 0, s>>>(k1, n1, reinterpret_cast<const uint4*>(desc1), k2,
                                                    reinterpret_cast<const uint4*>(desc2), prev, minX, minY, invW, invH,
                                                    (float)window, sorted, cellStart, cand, candCap, candCnt, dStatus);
    count_launch();
    k_sfi_resolve<<<1, RESOLVE_THREADS, smem, s>>>(k1, n1, k2, n2, cand, candCap, candCnt, nnratio, checkOri, prev, matches12, nmatches);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

// ---- ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono)  (reference src/ORBmatcher.cc:1328-1470) ----
// Same decomposition as above: the window walk and the Hamming distances do not depend on the loop order, the
// assignment does (a current keypoint that received a map point with observations is skipped by later iterations,
// :1404-1406), so a warp replays the loop over precomputed candidate lists.
struct SbpParams {
    float fx, fy, cx, cy, mbf, th;
    float minX, maxX, minY, maxY, invW, invH;
    int direction;  // 0: octave-1 .. octave+1, 1: bForward (>= octave), 2: bBackward (0 .. octave)   (:1381-1386)
    int hasRight;
    int noDepthTest;  // relocalisation variant (:1472-1599): no `invzc < 0` test
    int thDist;       // TH_HIGH (:1427) or the caller's ORBdist (:1552)
    int mapMode;    // 1: SearchByProjection(F, vpMapPoints, th) (:45-129): x3Dc holds (mTrackProjX, mTrackProjY, mTrackProjXR),
                    //    radius = RadiusByViewingCos * th * scale, levels level-1 .. level, no image-bounds test
};

// One warp per last-frame keypoint i.  cand[i * candCap + k] = (dist << 16) | i2 in GetFeaturesInArea's visiting order.
__global__ void __launch_bounds__(256) k_sbp_candidates(int n1, const float* __restrict__ x3Dc, const uint8_t* __restrict__ valid1,
                                                         const uint4* __restrict__ mpDesc1, const int* __restrict__ octave1,
                                                         const orbx_keypoint* __restrict__ k2, const uint4* __restrict__ desc2,
                                                         const float* __restrict__ uRight2, const uint8_t* __restrict__ occupied2,
                                                         const float* __restrict__ scaleFactors, const float* __restrict__ viewCos,
                                                         const SbpParams P,
                                                         const int* __restrict__ sorted, const int* __restrict__ cellStart,
                                                         uint32_t* __restrict__ cand, int candCap, int* __restrict__ candCnt,
                                                         int* __restrict__ dStatus) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= n1) return;
    int count = 0;
    TopK mine;
    mine.clear();
    if (valid1[i]) {
        float u, v, ur, r = 0.f;
        bool inImage = true;
        int minLevel, maxLevel;
        const int oct = octave1[i];
        if (P.mapMode) {
            u = x3Dc[3 * i];
            v = x3Dc[3 * i + 1];
            ur = x3Dc[3 * i + 2];
            if (P.mapMode == 2) {  // Sim3 variant (:290-403): the third value is the radius th*mvScaleFactors[level] itself
                r = ur;
            } else {
                r = (double)viewCos[i] > 0.998 ? 2.5f : 4.0f;              // RadiusByViewingCos (:131-137)
                if ((double)P.th != 1.0) r = __fmul_rn(r, P.th);           // bFactor (:49, :65-66)
                r = __fmul_rn(r, scaleFactors[oct]);
            }
            minLevel = oct - 1;
            maxLevel = oct;
        } else {
            const float xc = x3Dc[3 * i], yc = x3Dc[3 * i + 1];
            const float invzc = __double2float_rn(__ddiv_rn(1.0, (double)x3Dc[3 * i + 2]));  // const float invzc = 1.0/z  (:1367)
            u = __fadd_rn(__fmul_rn(__fmul_rn(P.fx, xc), invzc), P.cx);                       // fx*xc*invzc+cx        (:1372)
            v = __fadd_rn(__fmul_rn(__fmul_rn(P.fy, yc), invzc), P.cy);
            inImage = (P.noDepthTest || !(invzc < 0.f)) && !(u < P.minX || u > P.maxX) && !(v < P.minY || v > P.maxY);
            ur = __fsub_rn(u, __fmul_rn(P.mbf, invzc));
            r = __fmul_rn(P.th, scaleFactors[oct]);
            minLevel = P.direction == 1 ? oct : (P.direction == 2 ? 0 : oct - 1);
            maxLevel = P.direction == 1 ? -1 : (P.direction == 2 ? oct : oct + 1);
        }
        if (inImage) {
            const bool checkLevels = minLevel > 0 || maxLevel >= 0;
            // GetFeaturesInArea(u, v, r, minLevel, maxLevel)  (src/Frame.cc:326-379)
            const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(u, P.minX), r), P.invW)));
            const int cx1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(u, P.minX), r), P.invW)));
            const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(v, P.minY), r), P.invH)));
            const int cy1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(v, P.minY), r), P.invH)));
            if (cx0 < GRID_COLS && cx1 >= 0 && cy0 < GRID_ROWS && cy1 >= 0) {
                const uint4 a0 = mpDesc1[2 * i], a1 = mpDesc1[2 * i + 1];
                uint32_t* out = cand + (size_t)i * candCap;
                for (int ix = cx0; ix <= cx1; ++ix) {
                    const int b = cellStart[ix * GRID_ROWS + cy0], e = cellStart[ix * GRID_ROWS + cy1 + 1];
                    for (int base = b; base < e; base += 32) {
                        const int p = base + lane;
                        bool ok = false;
                        int i2 = 0, dist = 0;
                        if (p < e) {
                            i2 = sorted[p];
                            const orbx_keypoint kp = k2[i2];
                            ok = !(checkLevels && (kp.octave < minLevel || (maxLevel >= 0 && kp.octave > maxLevel))) &&
                                 fabsf(__fsub_rn(kp.x, u)) < r && fabsf(__fsub_rn(kp.y, v)) < r;
                            // keypoints that hold a map point with observations on entry are skipped by every iteration
                            ok = ok && !occupied2[i2];
                            if (ok && P.hasRight) {
                                const float uR = uRight2[i2];
                                if (uR > 0.f && fabsf(__fsub_rn(ur, uR)) > r) ok = false;  // :1408-1414
                            }
                            if (ok) dist = hamming32B(a0, a1, desc2[2 * i2], desc2[2 * i2 + 1]);
                        }
                        const uint32_t m = __ballot_sync(0xffffffffu, ok);
                        if (ok) {
                            const int o = count + __popc(m & ((1u << lane) - 1));
                            if (o < candCap) out[o] = ((uint32_t)dist << 16) | (uint32_t)i2;
                            mine.push(((uint32_t)dist << 16) | (uint32_t)o, (uint32_t)i2, o < candCap);
                        }
                        count += __popc(m);
                    }
                }
            }
        }
    }
    warp_store_top(mine, lane, candCnt, n1, i);
    if (lane == 0) {
        if (count > candCap) {
            atomicExch(dStatus, ORBX_E_OVERFLOW);
            count = candCap;
        }
        candCnt[i] = count;
    }
}

// One CTA: warp 0 replays the loop on shared memory and registers only; all warps initialise, bin the rotations and write the
// result.  Dynamic shared memory: hist[32] (int32), claim[n2] (int32), owner[n2] (int32), evIdx[n1] (int32: bestIdx2 of the acceptance event of
// iteration i, -1 = none), taken[n2] (uint8), evBin[n1] (int8).
__global__ void __launch_bounds__(RESOLVE_THREADS) k_sbp_resolve(int n1, const uint8_t* __restrict__ obs1, const float* __restrict__ angle1,
                                                                 const orbx_keypoint* __restrict__ k2, int n2,
                                                                 const uint32_t* __restrict__ cand, int candCap,
                                                                 const int* __restrict__ candCnt, int checkOri, int thDist,
                                                                 int* __restrict__ match21, int* __restrict__ nmatchesOut) {
    extern __shared__ __align__(16) uint8_t sm[];
    int* hist = reinterpret_cast<int*>(sm);
    int* claim = hist + 32;
    int* owner = claim + n2;
    int* evIdx = owner + n2;
    uint8_t* taken = reinterpret_cast<uint8_t*>(evIdx + n1);
    signed char* evBin = reinterpret_cast<signed char*>(taken + n2);
    const int tid = threadIdx.x, lane = tid & 31;
    for (int i = tid; i < n2; i += RESOLVE_THREADS) { owner[i] = -1; taken[i] = 0; claim[i] = RS_CLAIM_FREE; }
    for (int i = tid; i < n1; i += RESOLVE_THREADS) { evIdx[i] = -1; evBin[i] = -1; }
    if (tid < 32) hist[tid] = 0;
    __syncthreads();
    if (tid < 32) {
        auto isFree = [&](uint32_t i2, uint32_t) { return taken[i2] == 0; };
        int cntNext;
        TopK topNext;
        fetch_feature(candCnt, n1, lane, cntNext, topNext);
        uint8_t obsNext = lane < n1 ? obs1[lane] : 0;
        for (int base = 0; base < n1; base += 32) {
            // 32 points at a time, their best candidates from k_sbp_candidates (resolve_common.cuh); key = dist << 16 | position: minimal
            // distance, first in visiting order (strict '<', :1420)
            const int i = base + lane;
            const int cnt = cntNext;
            const uint8_t obs = obsNext;
            const TopK top = topNext;
            fetch_feature(candCnt, n1, i + 32, cntNext, topNext);
            obsNext = i + 32 < n1 ? obs1[i + 32] : 0;
            const uint32_t* c = cand + (size_t)min(i, n1 - 1) * candCap;
            resolve_rounds(
                __ballot_sync(0xffffffffu, cnt > 0), lane, claim,
                [&](uint32_t& b1, uint32_t& i2a, uint32_t& b2, uint32_t& i2b) {
                    if (!top.pick(isFree, 1, b1, i2a, b2, i2b)) rescan_list(c, cnt, isFree, b1, i2a, b2, i2b);
                    b2 = RS_NONE;  // only the best matters here
                    return b1 != RS_NONE && (int)(b1 >> 16) <= thDist;  // TH_HIGH (:1427) / ORBdist (:1552)
                },
                [&](uint32_t, uint32_t i2a, uint32_t, uint32_t) {
                    owner[i2a] = i;
                    taken[i2a] = obs;
                    evIdx[i] = (int)i2a;
                });
        }
    }
    __syncthreads();
    int total = 0;
    for (int i = tid; i < n1; i += RESOLVE_THREADS) {
        const int idx = evIdx[i];
        if (idx < 0) continue;
        ++total;
        if (checkOri) {
            const int bin = rotation_bin(angle1[i], k2[idx].angle);
            evBin[i] = (signed char)bin;
            atomicAdd(&hist[bin], 1);
        }
    }
    __syncthreads();
    if (checkOri) {
        int ind1, ind2, ind3;
        three_maxima(hist, ind1, ind2, ind3);
        for (int i = tid; i < n1; i += RESOLVE_THREADS) {
            const int b = evBin[i];
            if (b >= 0 && b != ind1 && b != ind2 && b != ind3) {
                owner[evIdx[i]] = -2;  // CurrentFrame.mvpMapPoints[idx] = NULL; nmatches--  (:1460-1461)
                --total;
            }
        }
    }
    total = __reduce_add_sync(0xffffffffu, total);
    __syncthreads();
    if (tid == 0) hist[0] = 0;
    __syncthreads();
    if (lane == 0 && total) atomicAdd(&hist[0], total);
    for (int i = tid; i < n2; i += RESOLVE_THREADS) match21[i] = owner[i];
    __syncthreads();
    if (tid == 0) *nmatchesOut = hist[0];
}

int launch_search_by_projection(int n1, const float* x3Dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mpDesc1,
                                const int* octave1, const float* angle1, const orbx_keypoint* k2, int n2, const uint8_t* desc2,
                                const float* uRight2, const uint8_t* occupied2, const float* scaleFactors, const float cam[4],
                                float mbf, const float bounds[4], float th, int direction, int checkOri, int thDist,
                                int noDepthTest, int* cellOf, int* sorted, int* cellStart, uint32_t* cand, int candCap,
                                int* candCnt, int* match21, int* nmatches, int* dStatus, cudaStream_t s) {
    SbpParams P;
    P.fx = cam[0]; P.fy = cam[1]; P.cx = cam[2]; P.cy = cam[3];
    P.mbf = mbf; P.th = th;
    P.minX = bounds[0]; P.maxX = bounds[1]; P.minY = bounds[2]; P.maxY = bounds[3];
    P.invW = (float)GRID_COLS / (P.maxX - P.minX);  // mfGridElementWidthInv (src/Frame.cc:104-105)
    P.invH = (float)GRID_ROWS / (P.maxY - P.minY);
    P.direction = direction;
    P.hasRight = uRight2 != nullptr;
    P.mapMode = 0;
    P.thDist = thDist;
    P.noDepthTest = noDepthTest;
    const size_t smem = 128 + (size_t)n2 * 8 + (size_t)n1 * 4 + (size_t)n2 + (size_t)n1 + 16;
    if (smem > 200 * 1024) {
        set_error("SearchByProjection: %d + %d keypoints do not fit the resolve kernel's shared memory", n1, n2);
        return ORBX_E_INVALID;
    }
    static size_t configured[ORBX_MAX_DEVICES] = {};
    if (int rcAttr = ensure_dynamic_smem((const void*)k_sbp_resolve, smem, configured)) return rcAttr;
    if (cellOf) {  // null: sorted / cellStart hold the frame's grid already (orbx_frame)
        k_sfi_grid<<<1, 1024, 0, s>>>(k2, n2, P.minX, P.minY, P.invW, P.invH, cellOf, sorted, cellStart);
        count_launch();
    }
    k_sbp_candidates<<<(n1 + 7) / 8, 256, 0, s>>>(n1, x3Dc, valid1, reinterpret_cast<const uint4*>(mpDesc1), octave1, k2,
                                                    reinterpret_cast<const uint4*>(desc2), uRight2, occupied2, scaleFactors, nullptr, P,
                                                    sorted, cellStart, cand, candCap, candCnt, dStatus);
    count_launch();
    k_sbp_resolve<<<1, RESOLVE_THREADS, smem, s>>>(n1, obs1, angle1, k2, n2, cand, candCap, candCnt, checkOri, thDist, match21, nmatches);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

// SearchByProjection(F, vpMapPoints, th) (:45-129): warp 0 replays the loop; best / second-best with their octaves, ratio test
// only when both lie on the same level (:113-117).  Dynamic shared memory: claim[n2], owner[n2] (int32), taken[n2] (uint8), oct2[n2] (uint8).
__global__ void __launch_bounds__(RESOLVE_THREADS) k_sbpm_resolve(int n1, const uint8_t* __restrict__ obs1,
                                                                  const orbx_keypoint* __restrict__ k2, int n2,
                                                                  const uint32_t* __restrict__ cand, int candCap,
                                                                  const int* __restrict__ candCnt, float nnratio, int* __restrict__ match21,
                                                                  int* __restrict__ nmatchesOut) {
    extern __shared__ __align__(16) uint8_t sm[];
    int* claim = reinterpret_cast<int*>(sm);
    int* owner = claim + n2;
    uint8_t* taken = reinterpret_cast<uint8_t*>(owner + n2);
    uint8_t* oct2 = taken + n2;
    const int tid = threadIdx.x, lane = tid & 31;
    for (int i = tid; i < n2; i += RESOLVE_THREADS) { owner[i] = -1; taken[i] = 0; claim[i] = RS_CLAIM_FREE; oct2[i] = (uint8_t)k2[i].octave; }
    __syncthreads();
    if (tid < 32) {
        int nmatches = 0;  // lane-local
        auto isFree = [&](uint32_t i2, uint32_t) { return taken[i2] == 0; };
        int cntNext;
        TopK topNext;
        fetch_feature(candCnt, n1, lane, cntNext, topNext);
        uint8_t obsNext = lane < n1 ? obs1[lane] : 0;
        for (int base = 0; base < n1; base += 32) {
            const int i = base + lane;
            const int cnt = cntNext;
            const uint8_t obs = obsNext;
            const TopK top = topNext;  // the smallest (dist << 16 | position) keys of the list
            fetch_feature(candCnt, n1, i + 32, cntNext, topNext);
            obsNext = i + 32 < n1 ? obs1[i + 32] : 0;
            const uint32_t* c = cand + (size_t)min(i, n1 - 1) * candCap;
            resolve_rounds(
                __ballot_sync(0xffffffffu, cnt > 0), lane, claim,
                [&](uint32_t& b1, uint32_t& i2a, uint32_t& b2, uint32_t& i2b) {
                    if (!top.pick(isFree, 2, b1, i2a, b2, i2b)) rescan_list(c, cnt, isFree, b1, i2a, b2, i2b);
                    const int bestDist = (int)(b1 >> 16);
                    bool accept = b1 != RS_NONE && bestDist <= 100;  // TH_HIGH (:111)
                    // without a second candidate bestLevel2 stays -1 and the ratio test is skipped
                    if (accept && b2 != RS_NONE && oct2[i2a] == oct2[i2b] && (float)bestDist > __fmul_rn(nnratio, (float)(int)(b2 >> 16)))
                        accept = false;
                    return accept;
                },
                [&](uint32_t, uint32_t i2a, uint32_t, uint32_t) {
                    owner[i2a] = i;
                    taken[i2a] = obs;
                    ++nmatches;
                });
        }
        nmatches = __reduce_add_sync(0xffffffffu, nmatches);
        if (lane == 0) *nmatchesOut = nmatches;
    }
    __syncthreads();
    for (int i = tid; i < n2; i += RESOLVE_THREADS) match21[i] = owner[i];
}

int launch_search_by_projection_map(int n1, const float* proj, const float* viewCos, const int* level1, const uint8_t* valid1,
                                    const uint8_t* obs1, const uint8_t* mpDesc1, const orbx_keypoint* k2, int n2,
                                    const uint8_t* desc2, const float* uRight2, const uint8_t* occupied2, const float* scaleFactors,
                                    const float bounds[4], float th, float nnratio, int* cellOf, int* sorted, int* cellStart,
                                    uint32_t* cand, int candCap, int* candCnt, int* match21, int* nmatches, int* dStatus,
                                    cudaStream_t s) {
    SbpParams P;
    memset(&P, 0, sizeof P);
    P.th = th;
    P.minX = bounds[0]; P.maxX = bounds[1]; P.minY = bounds[2]; P.maxY = bounds[3];
    P.invW = (float)GRID_COLS / (P.maxX - P.minX);
    P.invH = (float)GRID_ROWS / (P.maxY - P.minY);
    P.hasRight = uRight2 != nullptr;
    P.mapMode = 1;
    const size_t smem = (size_t)n2 * 8 + (size_t)n2 * 2 + 16;
    if (smem > 200 * 1024) {
        set_error("SearchByProjection: %d keypoints do not fit the resolve kernel's shared memory", n2);
        return ORBX_E_INVALID;
    }
    static size_t configured[ORBX_MAX_DEVICES] = {};
    if (int rcAttr = ensure_dynamic_smem((const void*)k_sbpm_resolve, smem, configured)) return rcAttr;
    if (cellOf) {  // null: sorted / cellStart hold the frame's grid already (orbx_frame)
        k_sfi_grid<<<1, 1024, 0, s>>>(k2, n2, P.minX, P.minY, P.invW, P.invH, cellOf, sorted, cellStart);
        count_launch();
    }
    k_sbp_candidates<<<(n1 + 7) / 8, 256, 0, s>>>(n1, proj, valid1, reinterpret_cast<const uint4*>(mpDesc1), level1, k2,
                                                    reinterpret_cast<const uint4*>(desc2), uRight2, occupied2, scaleFactors, viewCos,
                                                    P, sorted, cellStart, cand, candCap, candCnt, dStatus);
    count_launch();
    k_sbpm_resolve<<<1, RESOLVE_THREADS, smem, s>>>(n1, obs1, k2, n2, cand, candCap, candCnt, nnratio, match21, nmatches);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

int launch_search_by_projection_sim3(int n1, const float* uvr, const int* level1, const uint8_t* valid1, const uint8_t* ones1,
                                     const uint8_t* mpDesc1, const orbx_keypoint* k2, int n2, const uint8_t* desc2,
                                     const uint8_t* occupied2, const float bounds[4], int thDist, int* cellOf, int* sorted,
                                     int* cellStart, uint32_t* cand, int candCap, int* candCnt, int* match21, int* nmatches,
                                     int* dStatus, cudaStream_t s) {
    SbpParams P;
    memset(&P, 0, sizeof P);
    P.minX = bounds[0]; P.maxX = bounds[1]; P.minY = bounds[2]; P.maxY = bounds[3];
    P.invW = (float)GRID_COLS / (P.maxX - P.minX);
    P.invH = (float)GRID_ROWS / (P.maxY - P.minY);
    P.mapMode = 2;
    const size_t smem = 128 + (size_t)n2 * 8 + (size_t)n1 * 4 + (size_t)n2 + (size_t)n1 + 16;
    if (smem > 200 * 1024) {
        set_error("SearchByProjection: %d + %d keypoints do not fit the resolve kernel's shared memory", n1, n2);
        return ORBX_E_INVALID;
    }
    if (smem > 48 * 1024) ORBX_CUDA(cudaFuncSetAttribute(k_sbp_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    k_sfi_grid<<<1, 1024, 0, s>>>(k2, n2, P.minX, P.minY, P.invW, P.invH, cellOf, sorted, cellStart);
    count_launch();
    k_sbp_candidates<<<(n1 + 7) / 8, 256, 0, s>>>(n1, uvr, valid1, reinterpret_cast<const uint4*>(mpDesc1), level1, k2,
                                                    reinterpret_cast<const uint4*>(desc2), nullptr, occupied2, nullptr, nullptr, P,
                                                    sorted, cellStart, cand, candCap, candCnt, dStatus);
    count_launch();
    // best candidate only, every assigned keypoint is taken (ones1), no rotation pass
    k_sbp_resolve<<<1, RESOLVE_THREADS, smem, s>>>(n1, ones1, nullptr, k2, n2, cand, candCap, candCnt, 0, thDist, match21, nmatches);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

// ---- Fuse (both overloads): the candidate loop (src/ORBmatcher.cc:893-945 and :1055-1075) ----------------------------------
// Independent per map point: KeyFrame::GetFeaturesInArea(u, v, radius), level gate, optional chi-square gate on the reprojection
// error (stereo 7.8 / mono 5.99, float product compared in double), smallest Hamming distance, first in visiting order.
__global__ void __launch_bounds__(256) k_fuse_match(int n1, const float* __restrict__ uvr, const float* __restrict__ ur1,
                                                     const int* __restrict__ level1, const uint8_t* __restrict__ valid1,
                                                     const uint4* __restrict__ mpDesc1, const orbx_keypoint* __restrict__ k2,
                                                     const uint4* __restrict__ desc2, const float* __restrict__ uRight2,
                                                     const float* __restrict__ invLevelSigma2, float minX, float minY, float invW,
                                                     float invH, const int* __restrict__ sorted, const int* __restrict__ cellStart,
                                                     int* __restrict__ bestIdx, int* __restrict__ bestDist) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= n1) return;
    uint32_t best = RS_NONE, bestI2 = 0;
    if (valid1[i]) {
        const float u = uvr[3 * i], v = uvr[3 * i + 1], r = uvr[3 * i + 2];
        const int pred = level1[i];
        const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(u, minX), r), invW)));
        const int cx1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(u, minX), r), invW)));
        const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(v, minY), r), invH)));
        const int cy1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(v, minY), r), invH)));
        if (cx0 < GRID_COLS && cx1 >= 0 && cy0 < GRID_ROWS && cy1 >= 0) {
            const uint4 a0 = mpDesc1[2 * i], a1 = mpDesc1[2 * i + 1];
            int visited = 0;
            for (int ix = cx0; ix <= cx1; ++ix) {
                const int b = cellStart[ix * GRID_ROWS + cy0], e = cellStart[ix * GRID_ROWS + cy1 + 1];
                for (int base = b; base < e; base += 32) {
                    const int p = base + lane;
                    if (p < e) {
                        const int i2 = sorted[p];
                        const orbx_keypoint kp = k2[i2];
                        bool ok = fabsf(__fsub_rn(kp.x, u)) < r && fabsf(__fsub_rn(kp.y, v)) < r && !(kp.octave < pred - 1 || kp.octave > pred);
                        if (ok && ur1) {
                            const float ex = __fsub_rn(u, kp.x), ey = __fsub_rn(v, kp.y);
                            float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                            double lim = 5.99;
                            const float uR = uRight2 ? uRight2[i2] : -1.0f;
                            if (uR >= 0.0f) {
                                const float er = __fsub_rn(ur1[i], uR);
                                e2 = __fadd_rn(e2, __fmul_rn(er, er));
                                lim = 7.8;
                            }
                            ok = !((double)__fmul_rn(e2, invLevelSigma2[kp.octave]) > lim);
                        }
                        if (ok) {
                            const uint32_t dist = (uint32_t)hamming32B(a0, a1, desc2[2 * i2], desc2[2 * i2 + 1]);
                            const uint32_t key = (dist << 20) | (uint32_t)min(visited + (p - base), 0xFFFFF);
                            if (key < best) { best = key; bestI2 = (uint32_t)i2; }
                        }
                    }
                    visited += min(32, e - base);
                }
            }
        }
    }
    const uint32_t m = __reduce_min_sync(0xffffffffu, best);
    const unsigned who = __ballot_sync(0xffffffffu, best == m && m != RS_NONE);
    const int src = who ? __ffs(who) - 1 : 0;
    const uint32_t i2 = __shfl_sync(0xffffffffu, bestI2, src);
    if (lane == 0) {
        bestIdx[i] = who ? (int)i2 : -1;
        bestDist[i] = who ? (int)(m >> 20) : 256;
    }
}

int launch_fuse_match(int n1, const float* uvr, const float* ur1, const int* level1, const uint8_t* valid1, const uint8_t* mpDesc1,
                      const orbx_keypoint* k2, int n2, const uint8_t* desc2, const float* uRight2, const float* invLevelSigma2,
                      const float bounds[4], int* cellOf, int* sorted, int* cellStart, int* bestIdx, int* bestDist, cudaStream_t s) {
    const float invW = (float)GRID_COLS / (bounds[1] - bounds[0]), invH = (float)GRID_ROWS / (bounds[3] - bounds[2]);
    k_sfi_grid<<<1, 1024, 0, s>>>(k2, n2, bounds[0], bounds[2], invW, invH, cellOf, sorted, cellStart);
    count_launch();
    k_fuse_match<<<(n1 + 7) / 8, 256, 0, s>>>(n1, uvr, ur1, level1, valid1, reinterpret_cast<const uint4*>(mpDesc1), k2,
                                                reinterpret_cast<const uint4*>(desc2), uRight2, invLevelSigma2, bounds[0], bounds[2], invW,
                                                invH, sorted, cellStart, bestIdx, bestDist);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

// Frame::AssignFeaturesToGrid (src/Frame.cc:229-244) on its own: the 64 x 48 grid of a frame that stays on the device (orbx_frame)
int launch_frame_grid(const orbx_keypoint* k, int n, const float bounds[4], int* cellOf, int* sorted, int* cellStart, cudaStream_t s) {
    const float invW = (float)GRID_COLS / (bounds[1] - bounds[0]), invH = (float)GRID_ROWS / (bounds[3] - bounds[2]);
    k_sfi_grid<<<1, 1024, 0, s>>>(k, n, bounds[0], bounds[2], invW, invH, cellOf, sorted, cellStart);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
