// csrc/k_quadtree.cu — ORBextractor::DistributeOctTree as a deterministic data-parallel quadtree
// (reference src/ORBextractor.cc:831-887 DivideNode, :889-1120 DistributeOctTree incl. the fork's early exit at
// :959-963).  One CTA per (frame, level).
//
// The reference walks a std::list serially.  The same result is obtained level-synchronously:
//   sweep:  every expandable node of the list is split independently (a warp per node does a stable 4-way
//           partition of the node's key segment into the other ping-pong buffer); a prefix sum of the size deltas in
//           list order finds the node at which `lNodes.size() >= N` would have stopped the walk (:959-963); the
//           new list is reverse(children pushed by the processed nodes) ++ survivors, which is what push_front /
//           erase produce (:974-1020).
//   final:  when size + 3*nToExpand > N (:1030) the expandable children of the last pass are ordered by
//           (key count, creation order) descending — the reference sorts pair<int, node*> ascending and walks it
//           from the back (:1041-1043); the pointer tie-break is allocator-dependent there and is defined as
//           creation order here and in the oracle — and a prefix sum finds where size >= N breaks the walk (:1087).
//   keep:   per final node the first key of maximal response (:1098-1117), emitted in list order.
// tools/quadtree_proto.py checks this formulation against the CPU oracle on random inputs.
//
// Keys are 32-bit packed (resp << 24 | y << 12 | x); a node owns a contiguous segment of one of the two key
// buffers, children own sub-segments of the other buffer, so no key is ever copied except when its node splits.
#include <algorithm>
#include <cstdlib>

#include "orbx_internal.h"
#include "tma.cuh"

namespace orbx {

struct QNode {
    uint16_t x0, y0, x1, y1;
    uint32_t begin;  // bit 31: key buffer (0/1), bits 0..30: first key
    uint32_t count;
};

struct QtShared {  // carved out of dynamic shared memory
    QNode* list[2];
    uint4* cc;      // child key counts per list position
    int* sA;        // scan scratch (delta / pushed)
    int* sB;        // scan scratch (survivors)
    int* sC;        // processing order of the final phase
    int* cellOff;   // exclusive prefix of the per-cell counts
    int* misc;      // [0..15] scalars, [16..47] warp sums, [48..48+MAX_ROOTS] root counters
    int* rootW;     // [warps][nIni] per-warp-chunk root counts, then running offsets (root partition)
    uint4* bigCnt;  // [QT_MAXBIG][warps] child counts of the warp chunks of the nodes split by the whole block
    int* bigList;   // [QT_MAXBIG] list positions of those nodes
};

constexpr int QT_THREADS_SMALL = 512;  // small batches (single-frame latency): the SMs are idle anyway, so a CTA gets 16 warps
constexpr int QT_MAXBIG = 32;       // nodes per sweep that the whole block splits together
constexpr uint32_t QT_BIG = 768;    // ... when they hold more keys than this (else: one warp per node)

__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    return v;
}

// In-place exclusive prefix sum of a[0..n) in shared memory by the whole block; returns the total.
// Every thread owns a contiguous chunk, so the result is independent of the thread count.
__device__ __forceinline__ int block_exscan(int* a, int n, int* warpSums) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = blockDim.x >> 5;
    const int chunk = (n + blockDim.x - 1) / blockDim.x;
    const int b = min(tid * chunk, n), e = min(b + chunk, n);
    int local = 0;
    for (int i = b; i < e; ++i) local += a[i];
    const int incl = warp_incl_scan(local, lane);
    if (lane == 31) warpSums[wid] = incl;
    __syncthreads();
    int off = 0, total = 0;
    for (int w = 0; w < nw; ++w) {
        const int s = warpSums[w];
        if (w < wid) off += s;
        total += s;
    }
    int run = off + incl - local;
    for (int i = b; i < e; ++i) {
        const int v = a[i];
        a[i] = run;
        run += v;
    }
    __syncthreads();
    return total;
}

__device__ __forceinline__ int child_of(uint32_t key, int mx, int my) {
    return (key_x(key) < mx ? 0 : 1) + (key_y(key) < my ? 0 : 2);
}

// warp: count the keys k[0..count) falling into each of the four children.  Four loads per lane are in flight per trip (the
// keys live in global memory and a trip is otherwise one exposed L1 / L2 latency); long ranges count per lane and reduce once.
__device__ __forceinline__ uint4 warp_count_range(const uint32_t* k, uint32_t count, int mx, int my, int lane) {
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    if (count <= 128u) {
        int c[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t i = 32u * u + lane;
            c[u] = (i < count) ? child_of(k[i], mx, my) : 4;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (32u * u >= count) break;
            c0 += __popc(__ballot_sync(0xffffffffu, c[u] == 0));
            c1 += __popc(__ballot_sync(0xffffffffu, c[u] == 1));
            c2 += __popc(__ballot_sync(0xffffffffu, c[u] == 2));
            c3 += __popc(__ballot_sync(0xffffffffu, c[u] == 3));
        }
        return make_uint4(c0, c1, c2, c3);
    }
    for (uint32_t base = 0; base < count; base += 128) {
        uint32_t key[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t i = base + 32u * u + lane;
            key[u] = (i < count) ? k[i] : 0xFFFFFFFFu;  // x = y = 4095: never a key (coordinates stay below 4096 - border)
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const bool ok = base + 32u * u + lane < count;
            const int c = child_of(key[u], mx, my);
            c0 += (ok && c == 0); c1 += (ok && c == 1); c2 += (ok && c == 2); c3 += (ok && c == 3);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        c0 += __shfl_xor_sync(0xffffffffu, c0, o);
        c1 += __shfl_xor_sync(0xffffffffu, c1, o);
        c2 += __shfl_xor_sync(0xffffffffu, c2, o);
        c3 += __shfl_xor_sync(0xffffffffu, c3, o);
    }
    return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ void node_mid(const QNode& nd, int& mx, int& my) {
    mx = nd.x0 + ((nd.x1 - nd.x0 + 1) >> 1);
    my = nd.y0 + ((nd.y1 - nd.y0 + 1) >> 1);
}
__device__ __forceinline__ const uint32_t* node_keys(const QNode& nd, const uint32_t* bufs, int candCap) {
    return bufs + ((nd.begin >> 31) ? candCap : 0) + (nd.begin & 0x7FFFFFFFu);
}

__device__ __forceinline__ uint4 warp_count_children(const QNode& nd, const uint32_t* bufs, int candCap, int lane) {
    int mx, my;
    node_mid(nd, mx, my);
    return warp_count_range(node_keys(nd, bufs, candCap), nd.count, mx, my, lane);
}

// lanes 0..3: lane c writes the record of child c of `nd` (the k-th non-empty child goes to list position firstPos - k)
__device__ __forceinline__ void write_children(const QNode& nd, uint4 cc, QNode* outList, int firstPos, int lane) {
    if (lane >= 4) return;
    int mx, my;
    node_mid(nd, mx, my);
    const uint32_t srcBuf = nd.begin >> 31, first = nd.begin & 0x7FFFFFFFu;
    const int c = lane;
    const uint32_t cnt = c == 0 ? cc.x : c == 1 ? cc.y : c == 2 ? cc.z : cc.w;
    const uint32_t off = c == 0 ? 0u : c == 1 ? cc.x : c == 2 ? cc.x + cc.y : cc.x + cc.y + cc.z;
    const int k = (c > 0 && cc.x != 0) + (c > 1 && cc.y != 0) + (c > 2 && cc.z != 0);  // non-empty children before c
    if (cnt == 0) return;
    QNode ch;
    ch.x0 = (c & 1) ? (uint16_t)mx : nd.x0;
    ch.x1 = (c & 1) ? nd.x1 : (uint16_t)mx;
    ch.y0 = (c & 2) ? (uint16_t)my : nd.y0;
    ch.y1 = (c & 2) ? nd.y1 : (uint16_t)my;
    ch.begin = ((srcBuf ^ 1u) << 31) | (first + off);
    ch.count = cnt;
    outList[firstPos - k] = ch;
}

// warp: move src[0..count) stably into the four child runs of dst that start at run0..run3 (four loads in flight per trip)
__device__ __forceinline__ void warp_split_range(const uint32_t* src, uint32_t count, uint32_t* dst, uint32_t run0, uint32_t run1, uint32_t run2,
                                 uint32_t run3, int mx, int my, int lane) {
    const uint32_t below = (1u << lane) - 1;
    for (uint32_t base = 0; base < count; base += 128) {
        uint32_t key[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t i = base + 32u * u + lane;
            key[u] = (i < count) ? src[i] : 0u;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (base + 32u * u >= count) break;
            const int c = (base + 32u * u + lane < count) ? child_of(key[u], mx, my) : 4;
            const uint32_t b0 = __ballot_sync(0xffffffffu, c == 0), b1 = __ballot_sync(0xffffffffu, c == 1);
            const uint32_t b2 = __ballot_sync(0xffffffffu, c == 2), b3 = __ballot_sync(0xffffffffu, c == 3);
            if (c == 0) dst[run0 + __popc(b0 & below)] = key[u];
            else if (c == 1) dst[run1 + __popc(b1 & below)] = key[u];
            else if (c == 2) dst[run2 + __popc(b2 & below)] = key[u];
            else if (c == 3) dst[run3 + __popc(b3 & below)] = key[u];
            run0 += __popc(b0); run1 += __popc(b1); run2 += __popc(b2); run3 += __popc(b3);
        }
    }
}

// warp: write the child node records of `nd` and move its keys, stably, into the other buffer
__device__ __forceinline__ void warp_split(const QNode& nd, uint4 cc, uint32_t* bufs, int candCap, QNode* outList, int firstPos, int lane) {
    int mx, my;
    node_mid(nd, mx, my);
    const uint32_t srcBuf = nd.begin >> 31, first = nd.begin & 0x7FFFFFFFu;
    const uint32_t* src = bufs + (srcBuf ? candCap : 0) + first;
    uint32_t* dst = bufs + (srcBuf ? 0 : candCap) + first;
    write_children(nd, cc, outList, firstPos, lane);
    warp_split_range(src, nd.count, dst, 0, cc.x, cc.x + cc.y, cc.x + cc.y + cc.z, mx, my, lane);
}

__device__ __forceinline__ int nonzero4(uint4 c) { return (c.x != 0) + (c.y != 0) + (c.z != 0) + (c.w != 0); }

// Core: keys[0..K) sit in buffer 0 of `bufs` in the reference's vToDistributeKeys order.  Returns the number of
// kept keys (written to outKeys with (addX, addY) added to the coordinates).
__device__ __forceinline__ int quadtree_core(uint32_t* bufs, int candCap, int K, int W, int H, int nIni, float hX, int N, int addX,
                             int addY, uint32_t* outKeys, int outCap, const QtShared& S) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = blockDim.x >> 5;
    int* misc = S.misc;
    int* warpSums = misc + 16;
    int* rootCnt = misc + 48;
    if (K == 0 || nIni < 1) return 0;
    if (nIni > MAX_ROOTS) nIni = MAX_ROOTS;  // rejected on the host; guard only

    // ---- roots: stable partition of the candidates by kp.pt.x / hX (:917-921) into buffer 1 ----
    // Warp w owns the contiguous chunk w of the candidates: it counts its keys per root, a prefix over (root, warp) gives
    // every (warp, root) its first slot, and the warp scatters its chunk in order — stable, and each key is read twice.
    const int chunk = ((K + nw - 1) / nw + 31) & ~31;
    const int cb = min(wid * chunk, K), ce = min(cb + chunk, K);
    int* myRoot = S.rootW + wid * nIni;
    for (int i = tid; i < nw * nIni; i += blockDim.x) S.rootW[i] = 0;
    __syncthreads();
    const uint32_t laneBelow = (1u << lane) - 1;
    for (int base = cb; base < ce; base += 128) {
        int r[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = base + 32 * u + lane;
            r[u] = (i < ce) ? min((int)__fdiv_rn((float)key_x(bufs[i]), hX), nIni - 1) : -1;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (base + 32 * u >= ce) break;
            const uint32_t act = __ballot_sync(0xffffffffu, r[u] >= 0);
            if (r[u] >= 0) {
                const uint32_t peers = __match_any_sync(act, r[u]);
                if ((peers & laneBelow) == 0) myRoot[r[u]] += __popc(peers);  // one leader per root present in this row
            }
            __syncwarp();
        }
    }
    __syncthreads();
    for (int r = tid; r < nIni; r += blockDim.x) {  // per root: exclusive prefix over the warps, total into rootCnt
        int run = 0;
        for (int w = 0; w < nw; ++w) {
            const int c = S.rootW[w * nIni + r];
            S.rootW[w * nIni + r] = run;
            run += c;
        }
        rootCnt[r] = run;
    }
    __syncthreads();
    if (tid == 0) {
        int run = 0, n = 0;
        for (int r = 0; r < nIni; ++r) {
            const int c = rootCnt[r];
            if (c > 0) {
                QNode nd;
                nd.x0 = (uint16_t)(int)__fmul_rn(hX, (float)r);
                nd.x1 = (uint16_t)(int)__fmul_rn(hX, (float)(r + 1));
                nd.y0 = 0;
                nd.y1 = (uint16_t)H;
                nd.begin = (1u << 31) | (uint32_t)run;
                nd.count = (uint32_t)c;
                S.list[0][n++] = nd;
            }
            rootCnt[r] = run;  // becomes the segment start
            run += c;
        }
        misc[0] = n;
    }
    __syncthreads();
    {
        uint32_t* dst = bufs + candCap;
        for (int base = cb; base < ce; base += 128) {
            uint32_t key[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = base + 32 * u + lane;
                key[u] = (i < ce) ? bufs[i] : 0u;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (base + 32 * u >= ce) break;
                const bool ok = base + 32 * u + lane < ce;
                const int r = ok ? min((int)__fdiv_rn((float)key_x(key[u]), hX), nIni - 1) : -1;
                const uint32_t act = __ballot_sync(0xffffffffu, ok);
                uint32_t peers = 0;
                if (ok) {
                    peers = __match_any_sync(act, r);
                    dst[rootCnt[r] + myRoot[r] + __popc(peers & laneBelow)] = key[u];
                }
                __syncwarp();
                if (ok && (peers & laneBelow) == 0) myRoot[r] += __popc(peers);
                __syncwarp();
            }
        }
    }
    __syncthreads();
    (void)W;

    int cur = 0;          // which list buffer is current
    int n = misc[0];      // list size
    int P = 0;            // the first P list positions are the children pushed by the last pass
    bool finalPhase = false;
    while (true) {
        // (constant indices + a select: indexing S.list with `cur` would put the whole struct into local memory)
        QNode* L = cur ? S.list[1] : S.list[0];
        QNode* Lnew = cur ? S.list[0] : S.list[1];
        const int prev = n;
        if (!finalPhase) {
            // ---- sweep ----
            // nodes holding more than QT_BIG keys (the first passes: a handful of nodes own thousands of keys each) are
            // counted and split by the whole block, warp w taking chunk w of the node's segment; all others by one warp each
            if (tid == 0) {
                misc[1] = n;  // cut
                misc[3] = 0;  // nodes for the whole block
            }
            __syncthreads();
            for (int i = tid; i < n; i += blockDim.x) {
                int slot = -1;
                if (L[i].count > QT_BIG) {
                    slot = atomicAdd(&misc[3], 1);
                    if (slot < QT_MAXBIG) S.bigList[slot] = i;
                }
                S.sC[i] = slot;
            }
            __syncthreads();
            const int nBig = misc[3] <= QT_MAXBIG ? misc[3] : 0;  // more than the table holds: all of them go one warp each
            for (int i = wid; i < n; i += nw) {
                const QNode nd = L[i];
                if (nBig && S.sC[i] >= 0) continue;
                uint4 cc = make_uint4(0, 0, 0, 0);
                if (nd.count > 1) cc = warp_count_children(nd, bufs, candCap, lane);
                if (lane == 0) {
                    S.cc[i] = cc;
                    S.sA[i] = nd.count > 1 ? nonzero4(cc) - 1 : 0;
                }
            }
            for (int b = 0; b < nBig; ++b) {
                const QNode nd = L[S.bigList[b]];
                int mx, my;
                node_mid(nd, mx, my);
                const uint32_t ch = ((nd.count + nw - 1) / nw + 31u) & ~31u;
                const uint32_t c0 = min((uint32_t)wid * ch, nd.count), c1 = min(c0 + ch, nd.count);
                const uint4 cc = warp_count_range(node_keys(nd, bufs, candCap) + c0, c1 - c0, mx, my, lane);
                if (lane == 0) S.bigCnt[b * nw + wid] = cc;
            }
            __syncthreads();
            for (int b = tid; b < nBig; b += blockDim.x) {
                uint4 t = make_uint4(0, 0, 0, 0);
                for (int w = 0; w < nw; ++w) {
                    const uint4 c = S.bigCnt[b * nw + w];
                    t.x += c.x; t.y += c.y; t.z += c.z; t.w += c.w;
                }
                const int i = S.bigList[b];
                S.cc[i] = t;
                S.sA[i] = nonzero4(t) - 1;
            }
            __syncthreads();
            block_exscan(S.sA, n, warpSums);
            for (int i = tid; i < n; i += blockDim.x)
                if (prev + S.sA[i] >= N) atomicMin(&misc[1], i);
            __syncthreads();
            const int cut = misc[1];
            for (int i = tid; i < n; i += blockDim.x) {
                const bool split = i < cut && L[i].count > 1;
                S.sA[i] = split ? nonzero4(S.cc[i]) : 0;
                S.sB[i] = split ? 0 : 1;
            }
            __syncthreads();
            const int pushed = block_exscan(S.sA, n, warpSums);
            const int surv = block_exscan(S.sB, n, warpSums);
            for (int i = wid; i < n; i += nw) {
                const QNode nd = L[i];
                if (i < cut && nd.count > 1) {
                    if (!(nBig && S.sC[i] >= 0)) warp_split(nd, S.cc[i], bufs, candCap, Lnew, pushed - 1 - S.sA[i], lane);
                } else if (lane == 0) {
                    Lnew[pushed + S.sB[i]] = nd;
                }
            }
            for (int b = 0; b < nBig; ++b) {
                const int i = S.bigList[b];
                if (i >= cut) continue;  // the walk stopped before this node: it survives unsplit (copied above)
                const QNode nd = L[i];
                const uint4 cc = S.cc[i];
                int mx, my;
                node_mid(nd, mx, my);
                if (wid == 0) write_children(nd, cc, Lnew, pushed - 1 - S.sA[i], lane);
                uint32_t r0 = 0, r1 = cc.x, r2 = cc.x + cc.y, r3 = cc.x + cc.y + cc.z;
                for (int w = 0; w < wid; ++w) {  // keys of the chunks before mine
                    const uint4 c = S.bigCnt[b * nw + w];
                    r0 += c.x; r1 += c.y; r2 += c.z; r3 += c.w;
                }
                const uint32_t ch = ((nd.count + nw - 1) / nw + 31u) & ~31u;
                const uint32_t c0 = min((uint32_t)wid * ch, nd.count), c1 = min(c0 + ch, nd.count);
                const uint32_t srcBuf = nd.begin >> 31, first = nd.begin & 0x7FFFFFFFu;
                warp_split_range(bufs + (srcBuf ? candCap : 0) + first + c0, c1 - c0, bufs + (srcBuf ? 0 : candCap) + first, r0, r1, r2, r3,
                                 mx, my, lane);
            }
            __syncthreads();
            n = pushed + surv;
            P = pushed;
        } else {
            // ---- final phase pass over the expandable children of the previous pass (positions < P) ----
            for (int i = wid; i < P; i += nw) {
                const QNode nd = L[i];
                uint4 cc = make_uint4(0, 0, 0, 0);
                if (nd.count > 1) cc = warp_count_children(nd, bufs, candCap, lane);
                if (lane == 0) S.cc[i] = cc;
            }
            if (tid == 0) misc[1] = 0x7fffffff;  // rank of the node whose split reaches N
            __syncthreads();
            // processing order: descending (count, creation order); creation order = P-1-position
            int E = 0;
            const bool packed = candCap < (1 << 18) && P <= 8192;  // (count, creation order) fits one positive 32-bit sort key
            if (packed) {
                for (int i = tid; i < P; i += blockDim.x) {
                    const uint32_t ci = L[i].count;
                    S.sA[i] = ci > 1 ? (int)((ci << 13) | (uint32_t)(8191 - i)) : 0;
                }
                __syncthreads();
            }
            for (int i = tid; i < P; i += blockDim.x) {
                const uint32_t ci = L[i].count;
                if (ci > 1) {
                    int rank = 0;
                    if (packed) {
                        const int ki = S.sA[i];
#pragma unroll 4
                        for (int j = 0; j < P; ++j) rank += S.sA[j] > ki ? 1 : 0;
                    } else {
                        for (int j = 0; j < P; ++j) {
                            const uint32_t cj = L[j].count;
                            rank += (cj > 1 && (cj > ci || (cj == ci && j < i))) ? 1 : 0;
                        }
                    }
                    S.sC[rank] = i;
                }
                S.sB[i] = ci > 1 ? 1 : 0;
            }
            __syncthreads();
            E = block_exscan(S.sB, P, warpSums);  // only the total is needed
            for (int r = tid; r < E; r += blockDim.x) S.sA[r] = nonzero4(S.cc[S.sC[r]]);
            __syncthreads();
            block_exscan(S.sA, E, warpSums);  // sA[r] = children pushed before rank r
            for (int r = tid; r < E; r += blockDim.x) {
                // size after processing ranks 0..r = prev + (pushed through r) - (r + 1)
                const int after = prev + S.sA[r] + nonzero4(S.cc[S.sC[r]]) - (r + 1);
                if (after >= N) atomicMin(&misc[1], r);
            }
            __syncthreads();
            const int last = min(misc[1], E - 1);  // ranks 0..last are processed
            const int pushed = (last >= 0) ? S.sA[last] + nonzero4(S.cc[S.sC[last]]) : 0;
            // survivors: every list node that is not a processed parent
            for (int i = tid; i < n; i += blockDim.x) S.sB[i] = 1;
            __syncthreads();
            for (int r = tid; r <= last; r += blockDim.x) S.sB[S.sC[r]] = 0;
            __syncthreads();
            const int surv = block_exscan(S.sB, n, warpSums);
            for (int r = wid; r <= last; r += nw) {
                const int i = S.sC[r];
                warp_split(L[i], S.cc[i], bufs, candCap, Lnew, pushed - 1 - S.sA[r], lane);
            }
            for (int i = tid; i < n; i += blockDim.x) {
                const bool gone = (i + 1 < n) ? (S.sB[i + 1] == S.sB[i]) : (S.sB[i] == surv);
                if (!gone) Lnew[pushed + S.sB[i]] = L[i];
            }
            __syncthreads();
            n = pushed + surv;
            P = pushed;
        }
        cur ^= 1;
        // ---- termination (:1026-1030, :1092-1093) ----
        if (n >= N || n == prev) break;
        if (!finalPhase) {
            // nToExpand = pushed children holding more than one key
            if (tid == 0) misc[2] = 0;
            __syncthreads();
            int mine = 0;
            for (int i = tid; i < P; i += blockDim.x) mine += (cur ? S.list[1] : S.list[0])[i].count > 1 ? 1 : 0;
            if (mine) atomicAdd(&misc[2], mine);
            __syncthreads();
            if (n + 3 * misc[2] > N) finalPhase = true;
            __syncthreads();
        }
    }

    // ---- keep the best key of every node (:1098-1117): first key of maximal response, list order ----
    const QNode* L = cur ? S.list[1] : S.list[0];
    const int nOut = min(n, outCap);
    // Final nodes hold few keys (about candidates / N), so four lanes share a node: 8 nodes per warp step
    for (int i0 = wid * 8; i0 < nOut; i0 += nw * 8) {
        const int i = i0 + (lane >> 2), sub = lane & 3;
        uint32_t best = 0;
        const uint32_t* k = bufs;
        if (i < nOut) {
            const QNode nd = L[i];
            k = bufs + ((nd.begin >> 31) ? candCap : 0) + (nd.begin & 0x7FFFFFFFu);
            for (uint32_t j = sub; j < nd.count; j += 4) {  // (response, first index) packed: counts stay below 2^24
                const uint32_t v = ((uint32_t)key_resp(k[j]) << 24) | (0xFFFFFFu - j);
                best = max(best, v);
            }
        }
        best = max(best, __shfl_xor_sync(0xffffffffu, best, 1));
        best = max(best, __shfl_xor_sync(0xffffffffu, best, 2));
        if (i < nOut && sub == 0) {
            const uint32_t key = k[0xFFFFFFu - (best & 0xFFFFFFu)];
            outKeys[i] = pack_key(key_x(key) + addX, key_y(key) + addY, key_resp(key));
        }
    }
    return n;
}

__device__ __forceinline__ void carve(uint8_t* smem, int maxNodes, int maxCells, QtShared& S) {
    const int warps = blockDim.x >> 5;
    size_t o = 0;
    S.list[0] = reinterpret_cast<QNode*>(smem + o); o += (size_t)maxNodes * sizeof(QNode);
    S.list[1] = reinterpret_cast<QNode*>(smem + o); o += (size_t)maxNodes * sizeof(QNode);
    S.cc = reinterpret_cast<uint4*>(smem + o); o += (size_t)maxNodes * sizeof(uint4);
    S.bigCnt = reinterpret_cast<uint4*>(smem + o); o += (size_t)QT_MAXBIG * warps * sizeof(uint4);
    S.sA = reinterpret_cast<int*>(smem + o); o += (size_t)maxNodes * 4;
    S.sB = reinterpret_cast<int*>(smem + o); o += (size_t)maxNodes * 4;
    S.sC = reinterpret_cast<int*>(smem + o); o += (size_t)maxNodes * 4;
    S.misc = reinterpret_cast<int*>(smem + o); o += (48 + MAX_ROOTS + 4) * 4;
    S.rootW = reinterpret_cast<int*>(smem + o); o += (size_t)warps * MAX_ROOTS * 4;
    S.bigList = reinterpret_cast<int*>(smem + o); o += (size_t)QT_MAXBIG * 4;
    S.cellOff = reinterpret_cast<int*>(smem + o); o += (size_t)(maxCells + 1) * 4;
}

static size_t qt_smem(int maxNodes, int maxCells, int threads = QT_THREADS) {
    const int warps = threads / 32;
    return (size_t)maxNodes * (2 * sizeof(QNode) + sizeof(uint4) + 12) + (48 + MAX_ROOTS + 4) * 4 + (size_t)(maxCells + 1) * 4 +
           (size_t)QT_MAXBIG * warps * sizeof(uint4) + (size_t)warps * MAX_ROOTS * 4 + (size_t)QT_MAXBIG * 4;
}

// grid (nlevels, batch)
template <int THREADS>
__global__ void __launch_bounds__(THREADS) k_quadtree(const __grid_constant__ FrameGeom g, int maxNodes, int maxCells,
                                                          const int* __restrict__ cellCnt,
                                                          const uint32_t* __restrict__ cellKeys, uint32_t* qtScratch,
                                                          int* __restrict__ lvlCnt, uint32_t* __restrict__ lvlKeys,
                                                          int* dStatus, int levelBase) {
    extern __shared__ __align__(16) uint8_t qsmem[];
    QtShared S;
    carve(qsmem, maxNodes, maxCells, S);
    if (g.pdl) {
        griddep_launch_dependents();
        griddep_wait();
    }
    const int level = blockIdx.x + levelBase, f = blockIdx.y, tid = threadIdx.x;
    const LevelGeom& L = g.lv[level];
    const int nCells = L.nCols * L.nRows;
    const int* cnt = cellCnt + (long long)f * g.nCellsTotal + L.cellBase;
    const uint32_t* keys = cellKeys + (long long)f * g.keysPerFrame + L.keyBase;
    const int cellCap = L.cellCap;
    uint32_t* bufs = qtScratch + (long long)f * g.qtFrameStride + L.qtBase;
    // gather the per-cell slots into one list in cell order = vToDistributeKeys order (:1256-1299)
    for (int i = tid; i < nCells; i += blockDim.x) S.cellOff[i] = cnt[i];
    __syncthreads();
    int K = block_exscan(S.cellOff, nCells, S.misc + 16);
    if (K > L.candCap) {
        if (tid == 0) atomicExch(dStatus, ORBX_E_OVERFLOW);
        K = L.candCap;
    }
    // eight lanes copy one cell's slot (a cell of a natural image holds a handful of corners): four cells per warp trip
    for (int c = tid >> 3; c < nCells; c += blockDim.x >> 3) {
        const int nc = min(cnt[c], cellCap), o = S.cellOff[c];
        for (int k = tid & 7; k < nc; k += 8)
            if (o + k < K) bufs[o + k] = keys[(long long)c * cellCap + k];
    }
    __syncthreads();
    uint32_t* out = lvlKeys + (long long)f * g.slotsTotal + L.slotBase;
    const int n = quadtree_core(bufs, L.candCap, K, L.maxBX - L.minBX, L.maxBY - L.minBY, L.nIni, L.hX, L.quota, L.minBX,
                                L.minBY, out, L.quota + 4, S);
    if (tid == 0) lvlCnt[(long long)f * g.nlevels + level] = min(n, L.quota + 4);
}

__global__ void __launch_bounds__(QT_THREADS) k_quadtree_single(const uint32_t* keysIn, int K, int W, int H, int nIni,
                                                                 float hX, int N, int maxNodes, uint32_t* bufs, int candCap,
                                                                 uint32_t* out, int* outCnt) {
    extern __shared__ __align__(16) uint8_t qsmem[];
    QtShared S;
    carve(qsmem, maxNodes, 0, S);
    for (int i = threadIdx.x; i < K; i += blockDim.x) bufs[i] = keysIn[i];
    __syncthreads();
    const int n = quadtree_core(bufs, candCap, K, W, H, nIni, hX, N, 0, 0, out, N + 4, S);
    if (threadIdx.x == 0) *outCnt = min(n, N + 4);
}

int quadtree_smem_bytes(int maxQuota) { return (int)qt_smem(maxQuota + MAX_ROOTS + 8, 0); }

int launch_quadtree(const FrameGeom& g, int batch, const int* cellCnt, const uint32_t* cellKeys, uint32_t* qtScratch,
                    int* lvlCnt, uint32_t* lvlKeys, int* dStatus, cudaStream_t s, int onlyLevel) {
    int maxQuota = 0, maxCells = 0;
    for (int l = 0; l < g.nlevels; ++l) {
        maxQuota = max(maxQuota, g.lv[l].quota);
        maxCells = max(maxCells, g.lv[l].nCols * g.lv[l].nRows);
    }
    const int maxNodes = maxQuota + MAX_ROOTS + 8;
    const size_t smem = qt_smem(maxNodes, maxCells);
    if (smem > 220 * 1024) {
        set_error("quadtree: %zu bytes of shared memory needed (nfeatures per level %d, cells %d)", smem, maxQuota, maxCells);
        return ORBX_E_INVALID;
    }
    // onlyLevel >= 0: that pyramid level only (single frames: a level is distributed as soon as its corners exist)
    dim3 grid(onlyLevel >= 0 ? 1 : g.nlevels, batch);
    const int levelBase = onlyLevel >= 0 ? onlyLevel : 0;
    const bool attr = g.pdl && !g.plainLaunch;
    // Small batches (single-frame latency; at most one CTA per SM): a CTA gets 512 threads instead of 256 — the walk is a chain
    // of short block-wide steps, measured on one KITTI frame (level 0, cycles): 108 k at 256 threads, 88 k at 512, 103 k at
    // 1024; keeping the key buffers in shared memory changed nothing (the chain is instruction latency, not memory latency).
    // ORBX_QT_THREADS = 256 / 512 / 1024 overrides.
    static const int threadsPref = []() { const char* e = getenv("ORBX_QT_THREADS"); return e ? atoi(e) : QT_THREADS_SMALL; }();
    const bool small = g.nlevels * batch <= 148;
    const int threads = small && (threadsPref == 512 || threadsPref == 1024) ? threadsPref : QT_THREADS;
    const size_t smem2 = qt_smem(maxNodes, maxCells, threads);
    static size_t configured[3][ORBX_MAX_DEVICES] = {};
    int rcAttr;
    if (threads == 1024) {
        if ((rcAttr = ensure_dynamic_smem((const void*)k_quadtree<1024>, smem2, configured[2]))) return rcAttr;
        ORBX_CUDA(launch_kernel(k_quadtree<1024>, grid, dim3(1024), smem2, s, attr, g, maxNodes, maxCells, cellCnt, cellKeys, qtScratch, lvlCnt,
                                lvlKeys, dStatus, levelBase));
    } else if (threads == 512) {
        if ((rcAttr = ensure_dynamic_smem((const void*)k_quadtree<512>, smem2, configured[1]))) return rcAttr;
        ORBX_CUDA(launch_kernel(k_quadtree<512>, grid, dim3(512), smem2, s, attr, g, maxNodes, maxCells, cellCnt, cellKeys, qtScratch, lvlCnt,
                                lvlKeys, dStatus, levelBase));
    } else {
        if ((rcAttr = ensure_dynamic_smem((const void*)k_quadtree<QT_THREADS>, smem2, configured[0]))) return rcAttr;
        ORBX_CUDA(launch_kernel(k_quadtree<QT_THREADS>, grid, dim3(QT_THREADS), smem2, s, attr, g, maxNodes, maxCells, cellCnt, cellKeys,
                                qtScratch, lvlCnt, lvlKeys, dStatus, levelBase));
    }
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

int launch_quadtree_single(const uint32_t* dKeysIn, int n, int minX, int maxX, int minY, int maxY, int N,
                           uint32_t* dScratch, int candCap, uint32_t* dOut, int* dCount, int* dStatus, cudaStream_t s) {
    (void)dStatus;
    const int W = maxX - minX, H = maxY - minY;
    const int nIni = (int)roundf((float)W / (float)H);
    const float hX = nIni > 0 ? (float)W / (float)nIni : 1.f;
    if (nIni > MAX_ROOTS) {
        set_error("quadtree: aspect ratio %d:1 needs more than %d root nodes", nIni, MAX_ROOTS);
        return ORBX_E_INVALID;
    }
    const int maxNodes = N + MAX_ROOTS + 8;
    const size_t smem = qt_smem(maxNodes, 0);
    if (smem > 220 * 1024) {
        set_error("quadtree: N=%d does not fit shared memory", N);
        return ORBX_E_INVALID;
    }
    static size_t configured[ORBX_MAX_DEVICES] = {};
    if (int rcAttr = ensure_dynamic_smem((const void*)k_quadtree_single, smem, configured)) return rcAttr;
    k_quadtree_single<<<1, QT_THREADS, smem, s>>>(dKeysIn, n, W, H, nIni, hX, N, maxNodes, dScratch, candCap, dOut, dCount);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx

