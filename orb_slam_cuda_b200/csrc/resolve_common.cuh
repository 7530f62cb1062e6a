// csrc/resolve_common.cuh — shared by the "replay the sequential loop" kernels of k_search.cu and k_bow.cu.
//
// Every matcher of src/ORBmatcher.cc walks its first-argument features in order; feature i looks at a candidate list, skips
// the candidates that earlier iterations blocked (a keypoint that already holds a match), takes the best (and second best) of
// the rest and may block one more keypoint.  The blocked set only grows, so the replay works on 32 consecutive features at a
// time, one per lane:
//   scan    done ahead, in the candidate kernels (one warp per feature, all features in parallel): next to the list they store
//           its 8 smallest keys (distance << 16 | position in visiting order) — the best candidates while nothing is blocked yet;
//           the resolve lane loads those 32 bytes;
//   decide  all undecided lanes re-filter their cached 8 against the blocked set as it is by then (shared memory).  Whatever
//           became blocked since the scan can only remove entries, and every uncached candidate has a larger key than all cached
//           ones, so the first one / two free cached entries ARE the best / second best — unless too few are left while the
//           list had more than 8 candidates; then that lane rescans its list;
//   commit  a lane's decision equals the sequential one iff no EARLIER lane of the same round accepts the keypoint that is this
//           lane's best or second best (nothing else it looked at changes rank); decide() clears b1 / b2 where even that cannot
//           change the outcome.  Accepting lanes claim their keypoint with
//           atomicMin(lane); lanes before the first conflicting lane commit, the rest decide again in the next round.  The
//           lowest undecided lane never conflicts, so every round retires at least one feature; without conflicts, all 32.
// The global-memory latency is paid once per 32 features, and the decisions run 32 wide except where the order matters.
#pragma once
#include <cstdint>

namespace orbx {

constexpr uint32_t RS_NONE = 0xFFFFFFFFu;
constexpr int RESOLVE_THREADS = 256;  // one CTA per call: warp 0 decides, all warps run the parallel passes around it

constexpr int RS_K = 8;  // cached candidates per feature

struct TopK {
    uint32_t key[RS_K], idx[RS_K];
    int kept;  // candidates of the list (may exceed RS_K)
    __device__ __forceinline__ void clear() {
#pragma unroll
        for (int s = 0; s < RS_K; ++s) { key[s] = RS_NONE; idx[s] = 0; }
        kept = 0;
    }
    // Branch-free sorted insert; `take` = the candidate counts.
    __device__ __forceinline__ void push(uint32_t k, uint32_t i2, bool take) {
        kept += take ? 1 : 0;
        k = take ? k : RS_NONE;
        bool l[RS_K];  // l[0] => l[1] => ... (the keys are sorted)
#pragma unroll
        for (int s = 0; s < RS_K; ++s) l[s] = k < key[s];
#pragma unroll
        for (int s = RS_K - 1; s > 0; --s) {
            key[s] = l[s - 1] ? key[s - 1] : (l[s] ? k : key[s]);
            idx[s] = l[s - 1] ? idx[s - 1] : (l[s] ? i2 : idx[s]);
        }
        key[0] = l[0] ? k : key[0];
        idx[0] = l[0] ? i2 : idx[0];
    }
    // First two cached entries that are still free.  Returns false when the answer needs a rescan.
    template <class Free>
    __device__ __forceinline__ bool pick(Free free, int need, uint32_t& b1, uint32_t& i1, uint32_t& b2, uint32_t& i2) const {
        b1 = b2 = RS_NONE;
        i1 = i2 = 0;
        int found = 0;
#pragma unroll
        for (int s = 0; s < RS_K; ++s) {
            const bool ok = key[s] != RS_NONE && free(idx[s], key[s] >> 16);
            if (ok && found == 0) { b1 = key[s]; i1 = idx[s]; }
            if (ok && found == 1) { b2 = key[s]; i2 = idx[s]; }
            found += ok ? 1 : 0;
        }
        return found >= need || kept <= RS_K;
    }
};

// Candidate kernels: every lane pushed the candidates it wrote into its own TopK; the warp merges them into the feature's RS_K
// smallest keys (keys are unique: the position is part of the key) and lane 0 stores them.
__device__ __forceinline__ uint32_t* top_keys(int* candCnt, int n) { return reinterpret_cast<uint32_t*>(candCnt + (n + 3) / 4 * 4); }
__device__ __forceinline__ const uint32_t* top_keys(const int* candCnt, int n) {
    return reinterpret_cast<const uint32_t*>(candCnt + (n + 3) / 4 * 4);
}
// layout behind candCnt[n] (16-byte aligned): key[n][RS_K], then idx[n][RS_K]
__device__ __forceinline__ void warp_store_top(TopK& mine, int lane, int* candCnt, int n, int item) {
    uint32_t myK = RS_NONE, myV = 0;  // lane s < RS_K keeps the s-th smallest
#pragma unroll
    for (int s = 0; s < RS_K; ++s) {
        const uint32_t m = __reduce_min_sync(0xffffffffu, mine.key[0]);
        const unsigned who = __ballot_sync(0xffffffffu, mine.key[0] == m && m != RS_NONE);
        const int src = who ? __ffs(who) - 1 : 0;
        const uint32_t vi = __shfl_sync(0xffffffffu, mine.idx[0], src);
        if (lane == s) { myK = m; myV = who ? vi : 0u; }
        if (who && lane == src) {
#pragma unroll
            for (int t = 0; t < RS_K - 1; ++t) { mine.key[t] = mine.key[t + 1]; mine.idx[t] = mine.idx[t + 1]; }
            mine.key[RS_K - 1] = RS_NONE;
        }
    }
    if (lane < RS_K) {
        uint32_t* tk = top_keys(candCnt, n);
        tk[(size_t)item * RS_K + lane] = myK;
        tk[(size_t)(n + item) * RS_K + lane] = myV;
    }
}
// Resolve kernels: the lane's cache for feature i with cnt candidates.
__device__ __forceinline__ void load_top(TopK& t, const int* candCnt, int n, int i, int cnt) {
    const uint4* tk = reinterpret_cast<const uint4*>(top_keys(candCnt, n));
#pragma unroll
    for (int q = 0; q < RS_K / 4; ++q) {
        const uint4 k = tk[(size_t)i * (RS_K / 4) + q], v = tk[(size_t)(n + i) * (RS_K / 4) + q];
        t.key[4 * q] = k.x; t.key[4 * q + 1] = k.y; t.key[4 * q + 2] = k.z; t.key[4 * q + 3] = k.w;
        t.idx[4 * q] = v.x; t.idx[4 * q + 1] = v.y; t.idx[4 * q + 2] = v.z; t.idx[4 * q + 3] = v.w;
    }
    t.kept = cnt;
}
// cnt and cache of feature i (nothing for i >= n); issued one batch ahead so that the loads fly during the rounds
__device__ __forceinline__ void fetch_feature(const int* candCnt, int n, int i, int& cnt, TopK& t) {
    t.clear();
    cnt = 0;
    if (i < n) {
        cnt = candCnt[i];
        load_top(t, candCnt, n, i, 0);
        t.kept = cnt;
    }
}

// The rescan of the slow path: a lane-local walk of the candidate list (entries dist << 16 | i2, in visiting order) for the best /
// second best among the entries that are still free.
template <class Free>
__device__ __forceinline__ void rescan_list(const uint32_t* __restrict__ c, int cnt, Free free, uint32_t& b1, uint32_t& i1, uint32_t& b2,
                                            uint32_t& i2) {
    b1 = b2 = RS_NONE;
    i1 = i2 = 0;
    for (int p = 0; p < cnt; ++p) {
        const uint32_t e = __ldg(c + p);
        if (!free(e & 0xFFFFu, e >> 16)) continue;
        const uint32_t key = ((e >> 16) << 16) | (uint32_t)p;
        if (key < b1) { b2 = b1; i2 = i1; b1 = key; i1 = e & 0xFFFFu; }
        else if (key < b2) { b2 = key; i2 = e & 0xFFFFu; }
    }
}

// ComputeThreeMaxima (src/ORBmatcher.cc:1601-1642) over hist[30]; evaluated redundantly by every thread.
__device__ __forceinline__ void three_maxima(const int* hist, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int b = 0; b < 30; ++b) {
        const int s = hist[b];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = b; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = b; }
        else if (s > max3) { max3 = s; ind3 = b; }
    }
    if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
}

// The decide / commit rounds of one batch of 32 features (see the header comment).  todo = lanes holding a feature with
// candidates; claim[n2] is all RS_CLAIM_FREE on entry and on exit.  decide(b1, i2a, b2, i2b) -> accept; commit(b1, i2a, b2, i2b).
constexpr int RS_CLAIM_FREE = 0x7FFFFFFF;
template <class Decide, class Commit>
__device__ __forceinline__ void resolve_rounds(unsigned todo, int lane, int* claim, Decide decide, Commit commit) {
    while (todo) {
        const bool mine = (todo >> lane) & 1u;
        uint32_t b1 = RS_NONE, i2a = 0, b2 = RS_NONE, i2b = 0;
        bool acc = false;
        if (mine) acc = decide(b1, i2a, b2, i2b);
        if (acc) atomicMin(&claim[i2a], lane);
        __syncwarp();
        const bool conflict = mine && ((b1 != RS_NONE && claim[i2a] < lane) || (b2 != RS_NONE && claim[i2b] < lane));
        const unsigned cm = __ballot_sync(0xffffffffu, conflict);
        const int fc = cm ? __ffs(cm) - 1 : 32;
        if (acc && lane < fc) commit(b1, i2a, b2, i2b);
        if (acc) claim[i2a] = RS_CLAIM_FREE;
        __syncwarp();
        todo = fc >= 32 ? 0u : (todo & ~((1u << fc) - 1u));
    }
}

// Rotation-histogram bin of an accepted pair (src/ORBmatcher.cc:238-244 and the same lines of every matcher).
__device__ __forceinline__ int rotation_bin(float angle1, float angle2) {
    float rot = __fsub_rn(angle1, angle2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 1.0f / 30.0f));  // factor = 1.0f / HISTO_LENGTH
    if (bin == 30) bin = 0;
    return min(max(bin, 0), 29);
}

}  // namespace orbx
