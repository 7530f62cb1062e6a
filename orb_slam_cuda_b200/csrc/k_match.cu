// csrc/k_match.cu — Hamming matching kernels: ORBmatcher::DescriptorDistance (reference src/ORBmatcher.cc:1647-1663)
// as XOR (LOP3) + POPC over 8 x u32 per descriptor, and the best / second-best loops built on it
// (inner loops of SearchByBoW, :187-232; N x N distance tables of MapPoint::ComputeDistinctiveDescriptors,
// src/MapPoint.cc:280-305).
//
// Sequential semantics being reproduced:  best = smallest distance, bestIdx = FIRST index attaining it (updates use
// a strict '<'), second = second smallest value of the multiset (so second == best when the minimum occurs twice).
// Both are order independent, so partial results merge exactly:
//     best = min(b, c) on (distance, index);  second = min(max(b1, c1), min(b2, c2)).
//
// Roofline: integer pipe — 8 POPC per descriptor pair; descriptors are tiny (32 B) and stay in shared memory / L2.
#include <algorithm>

#include "orbx_internal.h"

namespace orbx {

__device__ __forceinline__ int hamming256(const uint32_t (&a)[8], const uint4 b0, const uint4 b1) {
    return __popc(a[0] ^ b0.x) + __popc(a[1] ^ b0.y) + __popc(a[2] ^ b0.z) + __popc(a[3] ^ b0.w) + __popc(a[4] ^ b1.x) +
           __popc(a[5] ^ b1.y) + __popc(a[6] ^ b1.z) + __popc(a[7] ^ b1.w);
}

// The same distance with HALF the POPC instructions, for the throughput kernels.  POPC issues at 0.5 warp instructions / clk /
// SM on B200 and LOP3 at 2 (tools/pipe_rates.cu, profiles/r02_pipe_rates.json), so 8 POPC per pair bound the plain form at
// 16 clk per warp-pair.  Four carry-save adders (sum = x ^ y ^ z: LUT 0x96, carry = majority: LUT 0xE8) fold seven of the
// eight XOR words into a ones, a twos and a fours plane: 4 POPC (8 clk on the POPC pipe) + 16 LOP3 (8 clk on the ALU pipe).
template <int LUT>
__device__ __forceinline__ uint32_t lop3m(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(LUT));
    return r;
}
__device__ __forceinline__ uint32_t hamming256_csa(const uint32_t (&a)[8], const uint4 b0, const uint4 b1) {
    const uint32_t x0 = a[0] ^ b0.x, x1 = a[1] ^ b0.y, x2 = a[2] ^ b0.z, x3 = a[3] ^ b0.w;
    const uint32_t x4 = a[4] ^ b1.x, x5 = a[5] ^ b1.y, x6 = a[6] ^ b1.z, x7 = a[7] ^ b1.w;
    const uint32_t s1 = lop3m<0x96>(x0, x1, x2), c1 = lop3m<0xE8>(x0, x1, x2);
    const uint32_t s2 = lop3m<0x96>(s1, x3, x4), c2 = lop3m<0xE8>(s1, x3, x4);
    const uint32_t s3 = lop3m<0x96>(s2, x5, x6), c3 = lop3m<0xE8>(s2, x5, x6);
    const uint32_t t = lop3m<0x96>(c1, c2, c3), f = lop3m<0xE8>(c1, c2, c3);
    return (uint32_t)(__popc(s3) + __popc(x7)) + 2u * (uint32_t)__popc(t) + 4u * (uint32_t)__popc(f);
}

__global__ void k_hamming_pairs(const uint4* __restrict__ a, const uint4* __restrict__ b, int n, int* __restrict__ dist) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 a0 = a[2 * i], a1 = a[2 * i + 1];
    const uint32_t q[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    dist[i] = hamming256(q, b[2 * i], b[2 * i + 1]);
}

constexpr int BF_THREADS = 128;  // queries per CTA
constexpr int BF_TILE = 64;      // database descriptors staged per tile (2 KB): small tiles give a 2000 x 2000 call 512 CTAs
constexpr uint32_t BF_NOIDX = 0x3FFFFFu;

// grid (ceil(nq / BF_THREADS), nsplit): every CTA matches 128 queries against one slice of the database and writes
// a partial (best = dist << 22 | idx, second) per query; the CTA that finishes a query block LAST (a counter per query
// block, left at zero for the next call) merges the partials and applies the acceptance test — one launch per call.
__global__ void __launch_bounds__(BF_THREADS) k_bruteforce(const uint4* __restrict__ q, int nq, const uint4* __restrict__ d, int nd,
                                                            int sliceLen, unsigned long long* __restrict__ partial,
                                                            unsigned int* __restrict__ counters, int th, float ratio,
                                                            int* __restrict__ idx, int* __restrict__ d1, int* __restrict__ d2,
                                                            int* __restrict__ match) {
    __shared__ uint4 tile[BF_TILE * 2];
    __shared__ bool isLast;
    const int qi = blockIdx.x * BF_THREADS + threadIdx.x;
    const int begin = blockIdx.y * sliceLen, end = min(begin + sliceLen, nd);
    uint32_t a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (qi < nq) {
        const uint4 a0 = q[2 * qi], a1 = q[2 * qi + 1];
        a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
    }
    // best and second are both kept as packed (distance << 22 | index) keys: the larger of two keys carries the larger
    // distance, so second's distance field is the second smallest value of the multiset
    uint32_t best = (256u << 22) | BF_NOIDX, second2 = (256u << 22) | BF_NOIDX;
    for (int t0 = begin; t0 < end; t0 += BF_TILE) {
        const int m = min(BF_TILE, end - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < 2 * m; i += BF_THREADS) tile[i] = d[2 * t0 + i];
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < m; ++j) {
            const uint32_t cand = (hamming256_csa(a, tile[2 * j], tile[2 * j + 1]) << 22) + (uint32_t)(t0 + j);
            second2 = min(second2, max(cand, best));
            best = min(best, cand);
        }
    }
    int second = (int)(second2 >> 22);
    const int nsplit = gridDim.y;
    if (nsplit > 1) {
        if (qi < nq) partial[(size_t)blockIdx.y * nq + qi] = ((unsigned long long)best << 32) | (uint32_t)second;
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) isLast = atomicAdd(&counters[blockIdx.x], 1u) == (unsigned)(nsplit - 1);
        __syncthreads();
        if (!isLast) return;
        __threadfence();
        if (threadIdx.x == 0) counters[blockIdx.x] = 0;
        if (qi >= nq) return;
        best = (256u << 22) | BF_NOIDX;
        second = 256;
        for (int s = 0; s < nsplit; ++s) {
            const unsigned long long p = __ldcg(&partial[(size_t)s * nq + qi]);
            const uint32_t b = (uint32_t)(p >> 32);
            const int s2 = (int)(uint32_t)p;
            second = min(max((int)(best >> 22), (int)(b >> 22)), min(second, s2));
            best = min(best, b);
        }
    }
    if (qi >= nq) return;
    const int b1 = (int)(best >> 22);
    const int bi = b1 >= 256 ? -1 : (int)(best & BF_NOIDX);  // a distance of 256 never passes the strict '<' against 256
    if (idx) idx[qi] = bi;
    if (d1) d1[qi] = b1;
    if (d2) d2[qi] = second;
    if (match) match[qi] = (b1 <= th && (float)b1 < __fmul_rn(ratio, (float)second)) ? bi : -1;
}

size_t bruteforce_scratch_bytes(int nq, int nd) {
    (void)nd;
    return (size_t)64 * (size_t)nq * sizeof(unsigned long long);
}
size_t bruteforce_counter_bytes(int nq) { return (size_t)((nq + BF_THREADS - 1) / BF_THREADS) * sizeof(unsigned int); }

int launch_hamming_pairs(const uint8_t* a, const uint8_t* b, int n, int* dist, cudaStream_t s) {
    if (n <= 0) return ORBX_OK;
    k_hamming_pairs<<<(n + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b), n, dist);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

int launch_bruteforce(const uint8_t* q, int nq, const uint8_t* d, int nd, int th, float ratio, int* idx, int* d1, int* d2,
                      int* match, unsigned long long* dPartial, unsigned int* dCounters, cudaStream_t s) {
    if (nq <= 0) return ORBX_OK;
    const int qBlocks = (nq + BF_THREADS - 1) / BF_THREADS;
    // enough CTAs to fill 148 SMs a few times over, at least one tile per slice, at most 64 slices
    int nsplit = std::max(1, std::min(64, (4 * 148 + qBlocks - 1) / qBlocks));
    nsplit = std::max(1, std::min(nsplit, (nd + BF_TILE - 1) / BF_TILE));
    const int sliceLen = nd > 0 ? ((nd + nsplit - 1) / nsplit + BF_TILE - 1) / BF_TILE * BF_TILE : BF_TILE;
    nsplit = nd > 0 ? (nd + sliceLen - 1) / sliceLen : 1;
    k_bruteforce<<<dim3(qBlocks, nsplit), BF_THREADS, 0, s>>>(reinterpret_cast<const uint4*>(q), nq, reinterpret_cast<const uint4*>(d),
                                                               nd, sliceLen, dPartial, dCounters, th, ratio, idx, d1, d2, match);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

// ---- all-vs-all keyframe matching ---------------------------------------------------------------------------------
// One CTA per (query keyframe i, database keyframe j): 256 threads x AP_Q queries in registers, the database keyframe
// streams through shared memory in tiles; every loaded database descriptor is compared with AP_Q queries, so the
// shared-memory traffic per pair is 32 B / AP_Q.  score[i][j] = number of queries whose best match passes the test.
constexpr int AP_THREADS = 256, AP_Q = 4, AP_TILE = 256;

// colMap (may be null): database keyframe j goes to column colMap[j] of the score row (row pitch ldScore); a negative entry
// skips the keyframe (padding of a gathered chunk, or a block that was scored already).
__global__ void __launch_bounds__(AP_THREADS) k_allpairs(const uint4* __restrict__ q, const uint4* __restrict__ db, int n, int ldScore,
                                                          const int* __restrict__ colMap, int th, float ratio, int* __restrict__ score) {
    __shared__ uint4 tile[AP_TILE * 2];
    __shared__ int total;
    const int j = blockIdx.x, i = blockIdx.y;
    const int col = colMap ? colMap[j] : j;
    if (col < 0) return;
    const uint4* __restrict__ qk = q + (size_t)i * n * 2;
    const uint4* __restrict__ dk = db + (size_t)j * n * 2;
    if (threadIdx.x == 0) total = 0;
    int accepted = 0;
    for (int q0 = 0; q0 < n; q0 += AP_THREADS * AP_Q) {  // query chunk of 1024
        uint32_t a[AP_Q][8];
        uint32_t best[AP_Q], second[AP_Q];  // packed (distance << 22 | index) keys, see k_bruteforce
#pragma unroll
        for (int k = 0; k < AP_Q; ++k) {
            const int qi = q0 + k * AP_THREADS + threadIdx.x;
            uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
            if (qi < n) { a0 = qk[2 * qi]; a1 = qk[2 * qi + 1]; }
            a[k][0] = a0.x; a[k][1] = a0.y; a[k][2] = a0.z; a[k][3] = a0.w;
            a[k][4] = a1.x; a[k][5] = a1.y; a[k][6] = a1.z; a[k][7] = a1.w;
            best[k] = second[k] = (256u << 22) | BF_NOIDX;
        }
        for (int t0 = 0; t0 < n; t0 += AP_TILE) {
            const int m = min(AP_TILE, n - t0);
            __syncthreads();
            for (int x = threadIdx.x; x < 2 * m; x += AP_THREADS) tile[x] = dk[2 * t0 + x];
            __syncthreads();
#pragma unroll 2
            for (int jj = 0; jj < m; ++jj) {
                const uint4 b0 = tile[2 * jj], b1 = tile[2 * jj + 1];
#pragma unroll
                for (int k = 0; k < AP_Q; ++k) {
                    const uint32_t cand = (hamming256_csa(a[k], b0, b1) << 22) + (uint32_t)(t0 + jj);
                    second[k] = min(second[k], max(cand, best[k]));
                    best[k] = min(best[k], cand);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < AP_Q; ++k) {
            const int qi = q0 + k * AP_THREADS + threadIdx.x;
            const int b1 = (int)(best[k] >> 22);
            if (qi < n && b1 <= th && (float)b1 < __fmul_rn(ratio, (float)(second[k] >> 22))) ++accepted;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) accepted += __shfl_xor_sync(0xffffffffu, accepted, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0 && accepted) atomicAdd(&total, accepted);
    __syncthreads();
    if (threadIdx.x == 0) score[(size_t)i * ldScore + col] = total;
}

// ---- MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:246-312) ---------------------------------------------------
// One warp per map point; lane i takes observation i (then i + 32, ...): its row of the N x N distance table, the row's median
// vDists[0.5 * (N - 1)] (:297-298) by bisection on the value (distances are 0..256: 9 counting passes over the N descriptors,
// which every lane reads at the same address), then the warp keeps the smallest (median, index): strict '<' = first index wins.
__global__ void __launch_bounds__(256) k_distinctive(int nPoints, const int* __restrict__ start, const uint4* __restrict__ desc,
                                                      int* __restrict__ bestIdx, int* __restrict__ bestMedian) {
    const int lane = threadIdx.x & 31;
    const int pt = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (pt >= nPoints) return;
    const int b = start[pt], n = start[pt + 1] - b;
    const int k = (n - 1) / 2;  // (size_t)(0.5 * (N - 1))
    uint32_t best = 0xFFFFFFFFu;
    for (int i = lane; i < n; i += 32) {
        const uint4 a0 = desc[2 * (b + i)], a1 = desc[2 * (b + i) + 1];
        const uint32_t a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        int lo = 0, hi = 256;  // smallest v with #(d <= v) >= k + 1
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int cnt = 0;
            for (int j = 0; j < n; ++j) cnt += hamming256(a, __ldg(&desc[2 * (b + j)]), __ldg(&desc[2 * (b + j) + 1])) <= mid ? 1 : 0;
            if (cnt >= k + 1) hi = mid;
            else lo = mid + 1;
        }
        best = min(best, ((uint32_t)lo << 16) | (uint32_t)min(i, 0xFFFF));
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if (lane == 0) {
        bestIdx[pt] = n > 0 ? (int)(best & 0xFFFFu) : -1;
        bestMedian[pt] = n > 0 ? (int)(best >> 16) : -1;
    }
}

int launch_distinctive(int nPoints, const int* start, const uint8_t* desc, int* bestIdx, int* bestMedian, cudaStream_t s) {
    if (nPoints <= 0) return ORBX_OK;
    k_distinctive<<<(nPoints + 7) / 8, 256, 0, s>>>(nPoints, start, reinterpret_cast<const uint4*>(desc), bestIdx, bestMedian);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

int launch_allpairs(const uint8_t* q, int kq, const uint8_t* db, int kd, int n, int th, float ratio, int* score, int ldScore,
                    const int* colMap, cudaStream_t s) {
    if (kq <= 0 || kd <= 0) return ORBX_OK;
    k_allpairs<<<dim3(kd, kq), AP_THREADS, 0, s>>>(reinterpret_cast<const uint4*>(q), reinterpret_cast<const uint4*>(db), n, ldScore,
                                                   colMap, th, ratio, score);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
