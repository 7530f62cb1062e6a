// csrc/k_fast.cu — per-30-px-cell FAST-9/16 with the iniThFAST -> minThFAST fallback and 3x3 score non-max
// suppression (reference src/ORBextractor.cc:1122-1149 grid set-up, :1256-1299 per-cell cv::FAST calls).
//
// The reference calls cv::FAST once or twice per 30-px cell on the window [ini, ini + cell + 6) (:1258-1275).
// cv::FAST ignores a 3-px rim, so the DETECTABLE regions [ini + 3, ini + cell + 3) tile a level exactly, and a pixel's
// corner score does not depend on the cell it is in.  What is per cell is (a) non-max suppression — rim pixels are
// never scored, so NMS must not look across a cell seam — and (b) the fallback to minThFAST when the cell produced no
// keypoint.  One CTA therefore handles a STRIP of several horizontally adjacent cells of one cell row:
//   tile   the strip's window is staged into shared memory by one TMA box load (origin rounded down to 16 bytes);
//   A      quick reject, four pixels per lane as packed bytes: saturating v +- t once per word, then carry-less byte
//          compares of the 8 EVEN ring points (rows y+-3, y+-2, y realigned with funnel shifts).  A run of 9 ring points
//          holds 4 consecutive even ones, so a pixel survives iff 4 consecutive even points are all brighter (or all
//          darker).  Warps walk rows, lanes walk 32-bit words, so there is no index division;
//   B      survivors (dense list) get OpenCV's cornerScore<16>: max over the 16 arcs of 9 of the min signed difference,
//          both polarities at once as packed 16-bit pairs (VIMNMX3.S16x2), minus 1.  score >= t is exactly the
//          16-point segment test, so no separate run-of-9 detection is needed; corners store the score in a shared
//          score tile (0 = not a corner) and join the corner list;
//   C      corners strictly greater than their 8 neighbours OF THE SAME CELL survive and are counted per cell;
//   pass 2 cells that are still empty repeat A-C at minThFAST (:1281-1285); pixels of other cells are masked out;
//   D      survivors are ranked by position inside their cell (cv::FAST's row-major order) and written to the cell's slot.
//
// Roofline: reads every level pixel once (strip windows overlap by 6 px and hit L2); the kernel is bound by the
// integer pipe, not by HBM.
#include <vector>

#include "orbx_internal.h"
#include "tma.cuh"

namespace orbx {

#ifndef ORBX_FS_THREADS
#define ORBX_FS_THREADS 128
#endif
#ifndef ORBX_FS_MINBLOCKS
#define ORBX_FS_MINBLOCKS 8  // measured on B200: 128 x 8 beats 256 x {1, 5, 6} and 128 x 12 (tools/variants.py)
#endif
constexpr int FS_THREADS = ORBX_FS_THREADS;
#ifndef ORBX_FS_LIST1
#define ORBX_FS_LIST1 2048
#endif
#ifndef ORBX_FS_LIST2
#define ORBX_FS_LIST2 1024
#endif
#ifndef ORBX_FS_MAXCELLS
#define ORBX_FS_MAXCELLS 4  // measured on B200 (tools/variants.py): 2 / 3 / 4 / 5 / 7 cells -> 2.24 / 2.05 / 1.96 / 2.02 / 2.09 ms per 512 KITTI frames
#endif
constexpr int FS_LIST1 = ORBX_FS_LIST1;  // quick-test survivors per row chunk
constexpr int FS_LIST2 = ORBX_FS_LIST2;  // corners per strip and pass (beyond that NMS scans the score tile)
constexpr int FS_KP = ORBX_FS_MAXCELLS * CELL_CAP;     // keypoints per strip

__constant__ int c_ringDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
__constant__ int c_ringDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// OpenCV's cornerScore<16> for both polarities at once.  e[k] holds (256 + d_k) in its low and (256 - d_k) in its high
// 16-bit half, d_k = v - ring_k, so one packed min / max (VIMNMX3.S16x2) serves the "darker ring" and the "brighter
// ring" case together.  Returns max over the 16 circular arcs of 9 of the arc minimum, per half (still biased by 256).
__device__ __forceinline__ uint32_t best_arc_min_x2(const uint32_t (&e)[16]) {
    uint32_t m3[16], m9[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) m3[k] = __vimin3_s16x2(e[k], e[(k + 1) & 15], e[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m9[k] = __vimin3_s16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
    const uint32_t a = __vimax3_s16x2(m9[0], m9[1], m9[2]), b = __vimax3_s16x2(m9[3], m9[4], m9[5]);
    const uint32_t c = __vimax3_s16x2(m9[6], m9[7], m9[8]), d = __vimax3_s16x2(m9[9], m9[10], m9[11]);
    const uint32_t f = __vimax3_s16x2(m9[12], m9[13], m9[14]);
    return __vimax3_s16x2(__vimax3_s16x2(a, b, c), __vimax3_s16x2(d, f, m9[15]), a);
}

__device__ __forceinline__ uint32_t lop3_and3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0x80;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t b, uint32_t c) {  // a & (b | c)
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xe0;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t lop3_or3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xfe;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

struct FastShared {
    uint8_t* tile;
    uint8_t* score;
    uint8_t* act;        // per tile column: 0x80 where the quick test is enabled in this pass
    uint8_t* cellOfCol;  // per tile column: cell index inside the strip, 255 = not detectable
    uint16_t* list1;
    uint16_t* list2;
    uint32_t* kp;        // cell << 24 | pos << 8 | score
    int* cnt;            // [0],[1] n1 (ping-pong per row chunk), [2] corners of the pass, [4] nkp, [5] pass-2 needed, [6..] per-cell counts
    uint64_t* bar;
};

__global__ void __launch_bounds__(FS_THREADS, ORBX_FS_MINBLOCKS) k_fast_strips(const __grid_constant__ FrameGeom g,
                                                             const __grid_constant__ TmaMaps maps,
                                                             const FastStrip* __restrict__ strips,
                                                             int* __restrict__ cellCnt, uint32_t* __restrict__ cellKeys,
                                                             int* __restrict__ dStatus) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int TW = g.fastBoxW, TH = g.fastBoxH, TN = TW * TH, TWq = TW >> 2;
    FastShared S;
    S.tile = smem;
    S.score = smem + TN;
    S.act = smem + 2 * TN;
    S.cellOfCol = S.act + TW;
    S.list1 = reinterpret_cast<uint16_t*>(S.cellOfCol + TW);
    S.list2 = S.list1 + FS_LIST1;
    S.kp = reinterpret_cast<uint32_t*>(S.list2 + FS_LIST2);
    S.cnt = reinterpret_cast<int*>(S.kp + FS_KP);
    S.bar = reinterpret_cast<uint64_t*>(S.cnt + 16);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int f = blockIdx.y;
    const FastStrip st = strips[blockIdx.x];
    const LevelGeom& L = g.lv[st.level];
    const int ci = st.ci, cj0 = st.cj0, nc = st.nc;
    const int cellGlobal0 = L.cellBase + ci * L.nCols + cj0;
    int* outCnt = cellCnt + (long long)f * g.nCellsTotal + cellGlobal0;
    uint32_t* outKeys = cellKeys + ((long long)f * g.nCellsTotal + cellGlobal0) * CELL_CAP;

    // window of the strip (src/ORBextractor.cc:1258-1275): rows of cell row ci, columns of cells cj0 .. cj0 + nc - 1
    const int iniY = L.minBY + ci * L.hCell, iniX = L.minBX + cj0 * L.wCell;
    const int maxY = min(iniY + L.hCell + 6, L.maxBY);
    const int dH = (iniY >= L.maxBY - 3) ? 0 : maxY - iniY - 6;  // detectable rows
    // detectable columns: cell j is valid unless iniX_j >= maxBX - 6 or its window is narrower than 7 px
    int dW = 0;
    for (int j = 0; j < nc; ++j) {
        const int ix = iniX + j * L.wCell;
        const int mx = min(ix + L.wCell + 6, L.maxBX);
        if (ix < L.maxBX - 6 && mx - ix >= 7) dW = j * L.wCell + (mx - ix - 6);
    }
    if (dH < 1 || dW < 1) {
        if (tid < nc) outCnt[tid] = 0;
        return;
    }
    if (tid == 0) {
        mbar_init(S.bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(S.bar, (uint32_t)TN);
        tma_load_3d(S.tile, &maps.m[st.level], (XOFF + iniX) & ~15, EDGE + iniY, f + g.f0, S.bar);  // box origin 16-byte aligned
    }
    const int xo = (XOFF + iniX) & 15;  // window column 0 inside the box
    const int X0 = xo + 3;              // first detectable tile column
    const int wq0 = X0 >> 2, nW = ((X0 + dW + 3) >> 2) - wq0;
    // while the tile is in flight: clear the score tile, set up the column tables and counters
    for (int i = tid; i < TN / 16; i += FS_THREADS) reinterpret_cast<uint4*>(S.score)[i] = make_uint4(0u, 0u, 0u, 0u);  // TW % 16 == 0
    for (int x = tid; x < TW; x += FS_THREADS) {
        const int dx = x - X0;
        const bool det = dx >= 0 && dx < dW;
        int cell = 0;  // dx / wCell for the few cells of a strip, without a division
        for (int j = 1; j < nc; ++j) cell += dx >= j * L.wCell ? 1 : 0;
        S.cellOfCol[x] = det ? (uint8_t)cell : (uint8_t)255;
        S.act[x] = det ? 0x80 : 0;
    }
    if (tid < 16) S.cnt[tid] = 0;
    __syncthreads();
    mbar_wait(S.bar, 0);

    const uint32_t* tile32 = reinterpret_cast<const uint32_t*>(S.tile);
    const uint32_t* act32 = reinterpret_cast<const uint32_t*>(S.act);
    const int rowsPerChunk = max(1, FS_LIST1 / (nW * 4));
    int rpc = dH;  // optimistic: the whole strip is one chunk (survivors are a small fraction of the pixels)
    const int stepY = FS_THREADS / nW, stepW = FS_THREADS - stepY * nW;
    int off[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) off[k] = c_ringDy[k] * TW + c_ringDx[k];

    // C for one corner: keep it iff its score beats the 8 neighbours that lie in the same cell
    auto nms_one = [&](int pos, int s) {
        const int y = pos / TW, x = pos - y * TW;
        const int cell = S.cellOfCol[x];
        const bool lft = S.cellOfCol[x - 1] == cell, rgt = S.cellOfCol[x + 1] == cell;
        const uint8_t* c = S.score + pos;
        bool keep = s > c[-TW] && s > c[TW];
        if (lft) keep = keep && s > c[-1] && s > c[-TW - 1] && s > c[TW - 1];
        if (rgt) keep = keep && s > c[1] && s > c[-TW + 1] && s > c[TW + 1];
        if (keep) {
            const int o = atomicAdd(&S.cnt[4], 1);
            if (o < FS_KP) S.kp[o] = ((uint32_t)cell << 24) | ((uint32_t)pos << 8) | (uint32_t)s;
            atomicAdd(&S.cnt[6 + cell], 1);
        }
    };

    int par = 0;  // which n1 counter the current chunk appends to
    for (int pass = 0; pass < 2; ++pass) {
        const int t = min(max(pass == 0 ? g.iniTh : g.minTh, 0), 255);
        const uint32_t T4 = (uint32_t)t * 0x01010101u;
        for (int y0 = 0; y0 < dH; y0 += rpc) {
            const int y1 = min(y0 + rpc, dH);
            // ---- A: packed quick test on the 8 even ring points: a run of 9 holds 4 consecutive even points ----
            // the (row, word) pairs of the chunk are walked as one flat index, so every lane has work whatever the strip
            // width; the row / word of a thread advance incrementally (no division inside the loop)
            const int total = (y1 - y0) * nW;
            int yy = tid / nW, w = tid - yy * nW;
            for (int idx = tid; idx < total; idx += FS_THREADS, w += stepW, yy += stepY) {
                if (w >= nW) { w -= nW; ++yy; }
                {
                    const int y = y0 + yy;
                    const uint32_t* rowc = tile32 + (y + 3) * TWq + wq0;
                    const uint32_t vm = act32[wq0 + w];
                    if (vm == 0) continue;
                    const uint32_t c = rowc[w];
                    const uint32_t hi = __vaddus4(c, T4), lo = __vsubus4(c, T4);  // saturating v + t, v - t
                    // per-byte unsigned compares without cross-byte carries: with r7 = r & 0x7f.., bit 7 of (r7 + (~hi & 0x7f..))
                    // is (r & 127) > (hi & 127) and bit 7 of (r7 + (~lo & 0x7f..) + 0x01..) is (r & 127) >= (lo & 127); one
                    // LOP3 per compare merges it with the top bits (0xb2: r > hi; 0x4d: !(r >= lo), i.e. v - t > r)
                    const uint32_t HN7 = ~hi & 0x7f7f7f7fu, LN7 = (~lo & 0x7f7f7f7fu) + 0x01010101u;
                    uint32_t ring[8];
                    ring[0] = rowc[w + 3 * TWq];                                                       // ( 0, +3)
                    ring[4] = rowc[w - 3 * TWq];                                                       // ( 0, -3)
                    {
                        const uint32_t* r = rowc + 2 * TWq;
                        const uint32_t m = r[w];
                        ring[1] = __funnelshift_r(m, r[w + 1], 16);                                    // (+2, +2)
                        ring[7] = __funnelshift_r(r[w - 1], m, 16);                                    // (-2, +2)
                    }
                    ring[2] = __funnelshift_r(c, rowc[w + 1], 24);                                     // (+3,  0)
                    ring[6] = __funnelshift_r(rowc[w - 1], c, 8);                                      // (-3,  0)
                    {
                        const uint32_t* r = rowc - 2 * TWq;
                        const uint32_t m = r[w];
                        ring[3] = __funnelshift_r(m, r[w + 1], 16);                                    // (+2, -2)
                        ring[5] = __funnelshift_r(r[w - 1], m, 16);                                    // (-2, -2)
                    }
                    uint32_t br[8], dk[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const uint32_t r7 = ring[k] & 0x7f7f7f7fu;
                        const uint32_t sb = r7 + HN7, sd = r7 + LN7;
                        asm("lop3.b32 %0, %1, %2, %3, 0xb2;" : "=r"(br[k]) : "r"(ring[k]), "r"(hi), "r"(sb));
                        asm("lop3.b32 %0, %1, %2, %3, 0x4d;" : "=r"(dk[k]) : "r"(ring[k]), "r"(lo), "r"(sd));
                    }
                    // a pixel survives iff 4 consecutive even ring points agree: with pair flags p[k] = f[k] & f[k+1] that
                    // is OR_k (p[k] & p[k+2]), and p[k+2] & (p[k] | p[k+4]) covers two of the eight products per LOP3
                    uint32_t pb[8], pd[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        pb[k] = br[k] & br[(k + 1) & 7];
                        pd[k] = dk[k] & dk[(k + 1) & 7];
                    }
                    const uint32_t fl = (lop3_or3(lop3_and_or(pb[2], pb[0], pb[4]), lop3_and_or(pb[6], pb[4], pb[0]),
                                                  lop3_and_or(pb[3], pb[1], pb[5])) |
                                         lop3_or3(lop3_and_or(pb[7], pb[5], pb[1]), lop3_and_or(pd[2], pd[0], pd[4]),
                                                  lop3_and_or(pd[6], pd[4], pd[0])) |
                                         lop3_and_or(pd[3], pd[1], pd[5]) | lop3_and_or(pd[7], pd[5], pd[1])) & vm;
                    if (fl) {
                        const int np = __popc(fl);
                        int o = atomicAdd(&S.cnt[par], np);
                        if (o + np <= FS_LIST1) {  // (an overflowing optimistic pass is repeated in row chunks, see below)
                            const int pos0 = (y + 3) * TW + ((wq0 + w) << 2);
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (fl & (0x80u << (8 * j))) S.list1[o++] = (uint16_t)(pos0 + j);
                        }
                    }
                }
            }
            __syncthreads();
            const int n1 = S.cnt[par];
            if (tid == 0) S.cnt[par ^ 1] = 0;  // nobody touches the other counter before the next barrier
            par ^= 1;
            if (n1 > FS_LIST1) {  // more survivors than the list holds (only possible while rpc == dH): redo the pass in
                                  // chunks of rows whose pixels all fit.  Nothing but the survivor count was written.
                rpc = rowsPerChunk;
                y0 = -rpc;
                __syncthreads();
                continue;
            }
            // ---- B: full 16-point segment test; corners get cornerScore<16> (both polarities: the one without a run
            //         never wins the max) and join the corner list of this pass ----
            for (int i = tid; i < n1; i += FS_THREADS) {
                const int pos = S.list1[i];
                const uint32_t v = S.tile[pos];
                const uint32_t cv = (v + 256u) + ((256u - v) << 16);  // (256 + v, 256 - v)
                uint32_t e[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) e[k] = (uint32_t)S.tile[pos + off[k]] * 0xFFFFu + cv;  // (256 + d, 256 - d), d = v - ring
                const uint32_t best = best_arc_min_x2(e);
                const int s = (int)max(best & 0xFFFFu, best >> 16) - 257;
                if (s >= t) {  // some arc of 9 has all |differences| > t with one sign: a corner, and s is its score
                    S.score[pos] = (uint8_t)s;
                    const int o = atomicAdd(&S.cnt[2], 1);
                    if (o < FS_LIST2) S.list2[o] = (uint16_t)pos;
                }
            }
            __syncthreads();
        }
        // ---- C: 3x3 non-max suppression among pixels of the same cell ----
        const int n2 = S.cnt[2];
        if (n2 <= FS_LIST2) {
            for (int i = tid; i < n2; i += FS_THREADS) {
                const int pos = S.list2[i];
                nms_one(pos, S.score[pos]);
            }
        } else {  // more corners than the list holds: scan the score tile of the active columns instead
            for (int y = warp; y < dH; y += FS_THREADS / 32) {
                const uint32_t* srow = reinterpret_cast<const uint32_t*>(S.score) + (y + 3) * TWq + wq0;
                for (int w = lane; w < nW; w += 32) {
                    const uint32_t sc = srow[w] & ((act32[wq0 + w] >> 7) * 0xFFu);
                    if (sc == 0) continue;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int s = (sc >> (8 * j)) & 0xFF;
                        if (s != 0) nms_one((y + 3) * TW + ((wq0 + w) << 2) + j, s);
                    }
                }
            }
        }
        __syncthreads();
        if (pass == 1) break;
        // ---- cells without a keypoint go again at minThFAST ----
        if (tid < nc) {
            const int ix = iniX + tid * L.wCell;
            const bool valid = ix < L.maxBX - 6 && min(ix + L.wCell + 6, L.maxBX) - ix >= 7;
            if (valid && S.cnt[6 + tid] == 0) S.cnt[5] = 1;
        }
        __syncthreads();
        if (S.cnt[5] == 0) break;
        for (int x = tid; x < TW; x += FS_THREADS) {
            const int cell = S.cellOfCol[x];
            S.act[x] = (cell != 255 && S.cnt[6 + cell] == 0) ? 0x80 : 0;
        }
        if (tid == 0) S.cnt[2] = 0;
        __syncthreads();
    }
    // ---- D: row-major order inside every cell (positions are unique; rank = number of smaller keys of the cell) ----
    int nkp = S.cnt[4];
    if (nkp > FS_KP) {
        if (tid == 0) atomicExch(dStatus, ORBX_E_OVERFLOW);
        nkp = FS_KP;
    }
    for (int i = tid; i < nkp; i += FS_THREADS) {
        const uint32_t e = S.kp[i];
        const uint32_t cell = e >> 24;
        int rank = 0;
        for (int j = 0; j < nkp; ++j) {
            const uint32_t o = S.kp[j];
            rank += ((o >> 24) == cell && o < e) ? 1 : 0;
        }
        if (rank < CELL_CAP) {
            const int pos = (int)((e >> 8) & 0xFFFFu);
            const int y = pos / TW, x = pos - y * TW;
            outKeys[(long long)cell * CELL_CAP + rank] = pack_key(x - xo + cj0 * L.wCell, y + ci * L.hCell, (int)(e & 0xFFu));
        }
    }
    if (tid < nc) {
        int c = S.cnt[6 + tid];
        if (c > CELL_CAP) {
            atomicExch(dStatus, ORBX_E_OVERFLOW);
            c = CELL_CAP;
        }
        outCnt[tid] = c;
    }
}

size_t fast_smem_bytes(const FrameGeom& g) {
    return (size_t)2 * g.fastBoxW * g.fastBoxH + 2 * g.fastBoxW + FS_LIST1 * 2 + FS_LIST2 * 2 + FS_KP * 4 +
           16 * 4 + 16;
}

// Cells per strip of a level: as many as ORBX_FS_MAXCELLS allows while the TMA box (strip window + 15 bytes of origin
// slack) stays within 256 columns.  Phase A walks (row, word) pairs as a flat index, so lane utilisation does not depend
// on the strip width; the width trades shared memory per CTA (occupancy) against halo re-reads and per-CTA set-up.
int fast_cells_per_strip(int wCell, int nCols) {
    int best = 1;
    for (int c = 1; c <= ORBX_FS_MAXCELLS && c <= FAST_MAX_STRIP_CELLS && c <= nCols; ++c) {
        if (c * wCell + 6 + 15 > 256) break;
        best = c;
    }
    return best;
}

int launch_fast(const FrameGeom& g, const TmaMaps& maps, const FastStrip* dStrips, int batch, int* cellCnt,
                uint32_t* cellKeys, int* dStatus, cudaStream_t s) {
    if (g.nStrips == 0) return ORBX_OK;
    const size_t smem = fast_smem_bytes(g);
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        ORBX_CUDA(cudaFuncSetAttribute(k_fast_strips, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    dim3 grid(g.nStrips, batch);
    k_fast_strips<<<grid, FS_THREADS, smem, s>>>(g, maps, dStrips, cellCnt, cellKeys, dStatus);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
