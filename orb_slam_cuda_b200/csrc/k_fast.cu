// csrc/k_fast.cu — per-30-px-cell FAST-9/16 with the iniThFAST -> minThFAST fallback and 3x3 score non-max
// suppression (reference src/ORBextractor.cc:1122-1149 grid set-up, :1256-1299 per-cell cv::FAST calls).
//
// The reference calls cv::FAST once or twice per 30-px cell on the window [ini, ini + cell + 6) (:1258-1275).
// cv::FAST ignores a 3-px rim, so the DETECTABLE regions [ini + 3, ini + cell + 3) tile a level exactly, and a pixel's
// corner score does not depend on the cell it is in.  What is per cell is (a) non-max suppression — rim pixels are
// never scored, so NMS must not look across a cell seam — and (b) the fallback to minThFAST when the cell produced no
// keypoint.  One CTA owns a STRIP of several horizontally adjacent cells of one cell row and is PERSISTENT OVER FRAMES:
// it walks a slice of the batch with a two-stage TMA / mbarrier ring (the window of frame f + 1 is in flight while
// frame f is scored), so the strip geometry, the column tables and the thread -> (row, word) mapping are built once per
// CTA instead of once per frame.  Per frame:
//   A1     opposite-pair reject, four pixels per lane as packed bytes: an arc of 9 ring points holds one point of every
//          diametral pair, so a pixel can only be a corner if ring point (0,+3) or (0,-3) differs from v by more than t
//          (OpenCV's own first test, d = tab[ptr[0]] | tab[ptr[8]], without the sign).  A lane walks a word column three
//          rows at a time, so every row difference serves two rows: one row load, one VABSDIFF4 and a carry test per
//          word; words with a live pixel are compacted into a word list (one warp scan and one atomic per warp round);
//   A2     the words of the list get the test on all 8 EVEN ring points (rows y+-3, y+-2, y realigned with funnel
//          shifts): a run of 9 holds 4 consecutive even points, so a pixel survives iff 4 consecutive even points are
//          all brighter (or all darker); survivors join a pixel list;
//   B      survivors get OpenCV's cornerScore<16>: max over the 16 arcs of 9 of the min signed difference, both
//          polarities at once as packed 16-bit pairs (VIMNMX3.S16x2), minus 1.  score >= t is exactly the 16-point
//          segment test; corners store the score in a shared score tile (0 = not a corner) and join the corner list;
//   C      corners strictly greater than their 8 neighbours OF THE SAME CELL set their bit in a keep bitmap;
//   pass 2 cells that are still empty repeat A-C at minThFAST (:1281-1285); pixels of other cells are masked out;
//   D      one warp per cell walks the cell's bitmap rows (lane = row), a warp prefix sum of the row counts gives every
//          keypoint its rank in cv::FAST's row-major order, and the keys go straight to the cell's slot.
// The byte compares never saturate: with s = (v & 127) + t and u = (v & 127) + 128 - t (no carries between bytes for
// t < 128) the top bits of v + t / v - t are (v ^ s) / (v & u), an overflowing v + t is bit 7 of (v & s) and an
// underflowing v - t is bit 7 of ~(v | u); both only mask the final flags.
//
// Roofline: reads every level pixel once (strip windows overlap by 6 px and hit L2); the kernel is bound by the
// integer pipe, not by HBM.
#include <mutex>
#include <vector>

#include "orbx_internal.h"
#include "tma.cuh"

namespace orbx {

#ifndef ORBX_FS_THREADS
#define ORBX_FS_THREADS 128
#endif
#ifndef ORBX_FS_MINBLOCKS
#define ORBX_FS_MINBLOCKS 7
#endif
constexpr int FS_THREADS = ORBX_FS_THREADS;
#ifndef ORBX_FS_WL
#define ORBX_FS_WL 1280
#endif
#ifndef ORBX_FS_LIST1
#define ORBX_FS_LIST1 1536  // (with LIST2 768: 30.9 KB of shared memory per CTA, 7 CTAs per SM instead of 6; smaller lists overflow into the redo path: profiles/r02_variants.txt)
#endif
#ifndef ORBX_FS_LIST2
#define ORBX_FS_LIST2 768
#endif
#ifndef ORBX_FS_MAXCELLS
#define ORBX_FS_MAXCELLS 4
#endif
#ifndef ORBX_FS_FPC
#define ORBX_FS_FPC 8   // frames one CTA walks (per-CTA set-up is paid once per slice)
#endif
constexpr int FS_WL = ORBX_FS_WL;        // words with a live pixel after A1, per row chunk
constexpr int FS_LIST1 = ORBX_FS_LIST1;  // even-point survivors per row chunk
constexpr int FS_LIST2 = ORBX_FS_LIST2;  // corners per strip and frame (beyond that NMS scans the score tile)

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

template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(LUT));
    return r;
}
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t b, uint32_t c) { return lop3<0xe0>(a, b, c); }  // a & (b | c)
__device__ __forceinline__ uint32_t lop3_or3(uint32_t a, uint32_t b, uint32_t c) { return lop3<0xfe>(a, b, c); }

// Thresholds of four centre pixels (one word) for the carry-less byte compares.
struct Thr {
    uint32_t hx, HN7;  // brighter: bit 7 of hx = top bit of v + t; HN7 = 127 - ((v + t) & 127)
    uint32_t lx, LN7;  // darker:   bit 7 of lx = top bit of v - t; LN7 = 128 - ((v - t) & 127)
    uint32_t okB, okD; // bit 7: v + t <= 255 / v - t >= 0 (a saturated threshold can never be passed)
};
// t < 128: no saturating arithmetic needed (see the header); t >= 128: the generic saturating form
__device__ __forceinline__ Thr make_thr(uint32_t c, uint32_t T4, uint32_t K4, bool small) {
    Thr r;
    if (small) {
        const uint32_t c7 = c & 0x7f7f7f7fu;
        const uint32_t s = c7 + T4;  // (v & 127) + t            <= 254
        const uint32_t u = c7 + K4;  // (v & 127) + 128 - t      in [1, 255]
        r.hx = c ^ s;
        r.HN7 = ~s & 0x7f7f7f7fu;
        r.lx = c & u;
        r.LN7 = (~u & 0x7f7f7f7fu) + 0x01010101u;
        r.okB = ~(c & s);
        r.okD = c | u;
    } else {
        const uint32_t hi = __vaddus4(c, T4), lo = __vsubus4(c, T4);
        r.hx = hi;
        r.HN7 = ~hi & 0x7f7f7f7fu;
        r.lx = lo;
        r.LN7 = (~lo & 0x7f7f7f7fu) + 0x01010101u;
        r.okB = 0xFFFFFFFFu;  // ring > 255 and ring < 0 are false by themselves
        r.okD = 0xFFFFFFFFu;
    }
    return r;
}
// bit 7 of every byte: ring > v + t (br) and ring < v - t (dk).  With r7 = ring & 127, bit 7 of r7 + HN7 is
// (ring & 127) > (hi & 127) and bit 7 of r7 + LN7 is (ring & 127) >= (lo & 127); one LOP3 per compare merges it with the
// top bits (0xb2: ring > hi; 0x4d: !(ring >= lo)).
__device__ __forceinline__ void cmp_ring(uint32_t ring, const Thr& th, uint32_t& br, uint32_t& dk) {
    const uint32_t r7 = ring & 0x7f7f7f7fu;
    br = lop3<0xb2>(ring, th.hx, r7 + th.HN7);
    dk = lop3<0x4d>(ring, th.lx, r7 + th.LN7);
}

// A1 building block: the bytes in which two rows three apart differ by more than t.  |a - b| is one VABSDIFF4 (a full-rate ALU
// instruction on sm_100a, tools/pipe_rates.cu); a byte d exceeds t iff d >= 128 or (d & 127) + 127 - t carries into bit 7
// (t < 128), resp. d >= 128 and (d & 127) + 127 - (t - 128) carries (t >= 128).  K7 holds 127 - (t & 127) in every byte.
// Bit 7 of every byte that passes is set; the other bits are not meaningful.
__device__ __forceinline__ uint32_t a1_flag(uint32_t a, uint32_t b, uint32_t K7, bool small) {
    const uint32_t d = __vabsdiffu4(a, b);
    const uint32_t g = (d & 0x7f7f7f7fu) + K7;
    return small ? (g | d) : (g & d);
}

__global__ void __launch_bounds__(FS_THREADS, ORBX_FS_MINBLOCKS) k_fast_strips(const __grid_constant__ FrameGeom g,
                                                             const __grid_constant__ TmaMaps maps,
                                                             const FastStrip* __restrict__ strips,
                                                             int* __restrict__ cellCnt, uint32_t* __restrict__ cellKeys,
                                                             int* __restrict__ dStatus, int batch, int fpc) {
    extern __shared__ __align__(128) uint8_t smem[];
    if (g.pdl) griddep_launch_dependents();
    const int TW = g.fastBoxW, TH = g.fastBoxH, TN = TW * TH, TWq = TW >> 2;
    const int TNa = (TN + 127) & ~127;  // a ring stage (TMA destinations are 128-byte aligned)
    const int BW = (TW + 31) >> 5;      // keep-bitmap words per tile row
    uint8_t* const score = smem + 2 * TNa;
    uint64_t* const bar = reinterpret_cast<uint64_t*>(score + TNa);
    int* const cntAll = reinterpret_cast<int*>(bar + 2);  // two sets (frame parity) of [0] live words, [1] survivors, [2] corners, [6..] per-cell keypoints
    uint32_t* const keep = reinterpret_cast<uint32_t*>(cntAll + 32);
    uint8_t* const act = reinterpret_cast<uint8_t*>(keep + TH * BW);  // per tile column: 0x80 where the tests are enabled in this pass
    uint8_t* const cellOfCol = act + TW;                               // per tile column: cell index inside the strip, 255 = not detectable
    uint8_t* const colList = cellOfCol + TW;                           // pass 2: the word columns of the empty cells
    uint16_t* const wlist = reinterpret_cast<uint16_t*>(colList + 64);  // y << 6 | word
    uint16_t* const list1 = wlist + FS_WL;                                // tile row << 8 | tile column
    uint16_t* const list2 = list1 + FS_LIST1;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int fBeg = blockIdx.y * fpc, fEnd = min(fBeg + fpc, batch);
    const FastStrip st = strips[blockIdx.x];
    const LevelGeom& L = g.lv[st.level];
    const int ci = st.ci, cj0 = st.cj0, nc = st.nc, wCell = L.wCell;
    const int cellGlobal0 = L.cellBase + ci * L.nCols + cj0;

    // window of the strip (src/ORBextractor.cc:1258-1275): rows of cell row ci, columns of cells cj0 .. cj0 + nc - 1
    const int iniY = L.minBY + ci * L.hCell, iniX = L.minBX + cj0 * wCell;
    const int maxY = min(iniY + L.hCell + 6, L.maxBY);
    const int dH = (iniY >= L.maxBY - 3) ? 0 : maxY - iniY - 6;  // detectable rows
    // detectable columns: cell j is valid unless iniX_j >= maxBX - 6 or its window is narrower than 7 px
    int dW = 0;
    for (int j = 0; j < nc; ++j) {
        const int ix = iniX + j * wCell;
        const int mx = min(ix + wCell + 6, L.maxBX);
        if (ix < L.maxBX - 6 && mx - ix >= 7) dW = j * wCell + (mx - ix - 6);
    }
    if (dH < 1 || dW < 1) {
        if (tid < nc)
            for (int f = fBeg; f < fEnd; ++f) cellCnt[(long long)f * g.nCellsTotal + cellGlobal0 + tid] = 0;
        return;
    }
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_mbar_init();
        tma_prefetch_desc(&maps.m[st.level]);
    }
    const int boxX = (XOFF + iniX) & ~15, boxY = EDGE + iniY;  // box origin 16-byte aligned
    const int xo = (XOFF + iniX) & 15;                         // window column 0 inside the box
    const int X0 = xo + 3;                                     // first detectable tile column
    const int wq0 = X0 >> 2, nW = ((X0 + dW + 3) >> 2) - wq0;
    // once per CTA: clear the score tile and the bitmap, set up the column tables and counters
    for (int i = tid; i < TNa / 16; i += FS_THREADS) reinterpret_cast<uint4*>(score)[i] = make_uint4(0u, 0u, 0u, 0u);
    for (int i = tid; i < TH * BW; i += FS_THREADS) keep[i] = 0u;
    for (int x = tid; x < TW; x += FS_THREADS) {
        const int dx = x - X0;
        const bool det = dx >= 0 && dx < dW;
        int cell = 0;  // dx / wCell for the few cells of a strip, without a division
        for (int j = 1; j < nc; ++j) cell += dx >= j * wCell ? 1 : 0;
        cellOfCol[x] = det ? (uint8_t)cell : (uint8_t)255;
        act[x] = det ? 0x80 : 0;
    }
    if (tid < 32) cntAll[tid] = 0;
    __syncthreads();
    if (g.pdl) griddep_wait();  // strip geometry and column tables are built; the pyramid is complete once the predecessor is
    if (tid == 0) {
        mbar_expect_tx(&bar[0], (uint32_t)TN);
        tma_load_3d(smem, &maps.m[st.level], boxX, boxY, fBeg + g.f0, &bar[0]);
    }

    const uint32_t* act32 = reinterpret_cast<const uint32_t*>(act);
    // row chunks: the word list always holds a chunk; the survivor list holds it in the common case (a denser chunk is
    // redone with rpcSafe rows, whose pixels all fit)
    const int rpcOpt = max(1, min(dH, FS_WL / nW)), rpcSafe = max(1, FS_LIST1 / (nW * 4));
    const int nWG = (nW + 31) >> 5;  // 32-column groups of the strip (nW <= 64)
    const uint32_t laneLt = (1u << lane) - 1u;
    // emptiness test of the strip's cells after pass 1 (src/ORBextractor.cc:1281): which cells exist at all
    uint32_t validCells = 0;
    for (int j = 0; j < nc; ++j) {
        const int ix = iniX + j * wCell;
        if (ix < L.maxBX - 6 && min(ix + wCell + 6, L.maxBX) - ix >= 7) validCells |= 1u << j;
    }

    for (int f = fBeg, it = 0; f < fEnd; ++f, ++it) {
        const int stg = it & 1;
        int* const cnt = cntAll + stg * 16;  // this frame's counters; the other set is cleared below for frame f + 1
        if (tid == 0 && f + 1 < fEnd) {  // the other stage was released by the barrier that ended frame f - 1
            mbar_expect_tx(&bar[stg ^ 1], (uint32_t)TN);
            tma_load_3d(smem + (stg ^ 1) * TNa, &maps.m[st.level], boxX, boxY, f + 1 + g.f0, &bar[stg ^ 1]);
        }
        if (tid < 16) cntAll[(stg ^ 1) * 16 + tid] = 0;  // last used by frame f - 1, next used after this frame's last barrier
        const uint8_t* tile = smem + stg * TNa;
        const uint32_t* tile32 = reinterpret_cast<const uint32_t*>(tile);
        mbar_wait(&bar[stg], (uint32_t)(it >> 1) & 1u);

        // C for one corner: keep it iff its score beats the 8 neighbours that lie in the same cell
        auto nms_one = [&](int e, int s) {
            const int y = e >> 8, x = e & 255;
            const int cell = cellOfCol[x];
            const bool lft = cellOfCol[x - 1] == cell, rgt = cellOfCol[x + 1] == cell;
            const uint8_t* c = score + y * TW + x;
            bool kp = s > c[-TW] && s > c[TW];
            if (lft) kp = kp && s > c[-1] && s > c[-TW - 1] && s > c[TW - 1];
            if (rgt) kp = kp && s > c[1] && s > c[-TW + 1] && s > c[TW + 1];
            if (kp) {
                atomicOr(&keep[y * BW + (x >> 5)], 1u << (x & 31));
                atomicAdd(&cnt[6 + cell], 1);
            }
        };

        int n2done = 0;     // corners of list2 that C has already seen (pass 1)
        int nCols = nW;     // word columns A1 walks in this pass
        bool pass2 = false;
        for (int pass = 0; pass < 2; ++pass) {
            const int t = min(max(pass == 0 ? g.iniTh : g.minTh, 0), 255);
            const bool small = t < 128;
            const uint32_t T4 = (uint32_t)t * 0x01010101u, K4 = 0x80808080u - T4;
            const uint32_t K7 = (uint32_t)(127 - (t & 127)) * 0x01010101u;
            int rpc = rpcOpt;
            for (int y0 = 0; y0 < dH;) {
                const int y1 = min(y0 + rpc, dH);
                // ---- A1: opposite-pair reject on (0, +-3).  |T[k + 3] - T[k]| > t of tile rows k, k + 3 is the "down" test of
                //      detectable row k and the "up" test of row k + 3, so a lane walks one word column in steps of THREE rows
                //      (task = residue class x word column; all columns in pass 1, the empty cells' columns in pass 2): one
                //      row load and one difference test per word.  A lane collects its live rows in a bit mask; a warp prefix
                //      sum and ONE shared atomic per warp round place them in the word list ----
                {
                    const int rows = y1 - y0, nTasks = 3 * nCols;
                    const int per = min(32, (nTasks + FS_THREADS / 32 - 1) / (FS_THREADS / 32));  // tasks per warp and round
                    for (int tb = warp * per; tb < nTasks; tb += (FS_THREADS / 32) * per) {
                        const int task = tb + lane;
                        uint32_t live = 0;
                        int e0 = 0;
                        if (lane < per && task < nTasks) {
                            const int res = (task >= nCols ? 1 : 0) + (task >= 2 * nCols ? 1 : 0);
                            const int ci = task - res * nCols;
                            const int w = pass == 0 ? ci : colList[ci];
                            const uint32_t vm = act32[wq0 + w];
                            e0 = ((y0 + res) << 6) | w;
                            if (vm != 0 && res < rows) {
                                const uint32_t* p = tile32 + (y0 + res) * TWq + wq0 + w;
                                uint32_t cur = p[3 * TWq];
                                uint32_t fPrev = a1_flag(p[0], cur, K7, small);
                                p += 6 * TWq;
                                const int nu = (rows - res + 2) / 3;
#pragma unroll 4
                                for (int j = 0; j < nu; ++j) {
                                    const uint32_t next = *p;
                                    p += 3 * TWq;
                                    const uint32_t fNext = a1_flag(cur, next, K7, small);
                                    if (((fPrev | fNext) & vm) != 0) live |= 1u << j;
                                    cur = next;
                                    fPrev = fNext;
                                }
                            }
                        }
                        const int c = __popc(live);
                        int incl = c;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const int v = __shfl_up_sync(0xffffffffu, incl, o);
                            if (lane >= o) incl += v;
                        }
                        const int total = __shfl_sync(0xffffffffu, incl, 31);
                        if (total != 0) {
                            int base = 0;
                            if (lane == 31) base = atomicAdd(&cnt[0], total);
                            int o = __shfl_sync(0xffffffffu, base, 31) + incl - c;
                            while (live) {
                                const int j = __ffs(live) - 1;
                                live &= live - 1;
                                wlist[o++] = (uint16_t)(e0 + ((3 * j) << 6));
                            }
                        }
                    }
                }
                __syncthreads();
                // ---- A2: the 8 even ring points of the live words: a run of 9 holds 4 consecutive even points ----
                const int n0 = cnt[0];
                for (int i = tid; i < n0; i += FS_THREADS) {
                    const int e = wlist[i];
                    const int y = e >> 6, w = e & 63;
                    const uint32_t* rowc = tile32 + (y + 3) * TWq + wq0 + w;
                    const uint32_t vm = act32[wq0 + w];
                    const uint32_t c = rowc[0];
                    const Thr th = make_thr(c, T4, K4, small);
                    uint32_t ring[8];
                    ring[0] = rowc[3 * TWq];                                                        // ( 0, +3)
                    ring[4] = rowc[-3 * TWq];                                                       // ( 0, -3)
                    {
                        const uint32_t* r = rowc + 2 * TWq;
                        const uint32_t m = r[0];
                        ring[1] = __funnelshift_r(m, r[1], 16);                                     // (+2, +2)
                        ring[7] = __funnelshift_r(r[-1], m, 16);                                    // (-2, +2)
                    }
                    ring[2] = __funnelshift_r(c, rowc[1], 24);                                      // (+3,  0)
                    ring[6] = __funnelshift_r(rowc[-1], c, 8);                                      // (-3,  0)
                    {
                        const uint32_t* r = rowc - 2 * TWq;
                        const uint32_t m = r[0];
                        ring[3] = __funnelshift_r(m, r[1], 16);                                     // (+2, -2)
                        ring[5] = __funnelshift_r(r[-1], m, 16);                                    // (-2, -2)
                    }
                    uint32_t br[8], dk[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) cmp_ring(ring[k], th, br[k], dk[k]);
                    // a pixel survives iff 4 consecutive even ring points agree: with pair flags p[k] = f[k] & f[k+1] that
                    // is OR_k (p[k] & p[k+2]), and p[k+2] & (p[k] | p[k+4]) covers two of the eight products per LOP3
                    uint32_t pb[8], pd[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        pb[k] = br[k] & br[(k + 1) & 7];
                        pd[k] = dk[k] & dk[(k + 1) & 7];
                    }
                    const uint32_t fb = lop3_or3(lop3_and_or(pb[2], pb[0], pb[4]), lop3_and_or(pb[6], pb[4], pb[0]),
                                                 lop3_and_or(pb[3], pb[1], pb[5])) | lop3_and_or(pb[7], pb[5], pb[1]);
                    const uint32_t fd = lop3_or3(lop3_and_or(pd[2], pd[0], pd[4]), lop3_and_or(pd[6], pd[4], pd[0]),
                                                 lop3_and_or(pd[3], pd[1], pd[5])) | lop3_and_or(pd[7], pd[5], pd[1]);
                    const uint32_t fl = ((fb & th.okB) | (fd & th.okD)) & vm;
                    if (fl) {
                        const int np = __popc(fl);
                        int o = atomicAdd(&cnt[1], np);
                        if (o + np <= FS_LIST1) {  // (an overflowing chunk is redone with fewer rows, see below)
                            const int e0 = ((y + 3) << 8) | ((wq0 + w) << 2);
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (fl & (0x80u << (8 * j))) list1[o++] = (uint16_t)(e0 + j);
                        }
                    }
                }
                __syncthreads();
                const int n1 = cnt[1];
                if (tid == 0) cnt[0] = 0;  // every thread has read n0
                if (n1 > FS_LIST1) {       // more survivors than the list holds: redo this chunk with rows that always fit
                    __syncthreads();
                    if (tid == 0) cnt[1] = 0;
                    rpc = rpcSafe;
                    __syncthreads();
                    continue;
                }
                // ---- B: full 16-point segment test; corners get cornerScore<16> (both polarities: the one without a run
                //         never wins the max) and join the corner list ----
                for (int i = tid; i < n1; i += FS_THREADS) {
                    const int e = list1[i];
                    const uint8_t* p = tile + (e >> 8) * TW + (e & 255);
                    const uint32_t v = p[0];
                    const uint32_t cv = (v + 256u) + ((256u - v) << 16);  // (256 + v, 256 - v)
                    const uint8_t *pu3 = p + 3 * TW, *pu2 = p + 2 * TW, *pu1 = p + TW, *pd1 = p - TW, *pd2 = p - 2 * TW, *pd3 = p - 3 * TW;
                    uint32_t ev[16];  // (256 + d, 256 - d), d = v - ring
                    ev[0] = (uint32_t)pu3[0] * 0xFFFFu + cv;
                    ev[1] = (uint32_t)pu3[1] * 0xFFFFu + cv;
                    ev[2] = (uint32_t)pu2[2] * 0xFFFFu + cv;
                    ev[3] = (uint32_t)pu1[3] * 0xFFFFu + cv;
                    ev[4] = (uint32_t)p[3] * 0xFFFFu + cv;
                    ev[5] = (uint32_t)pd1[3] * 0xFFFFu + cv;
                    ev[6] = (uint32_t)pd2[2] * 0xFFFFu + cv;
                    ev[7] = (uint32_t)pd3[1] * 0xFFFFu + cv;
                    ev[8] = (uint32_t)pd3[0] * 0xFFFFu + cv;
                    ev[9] = (uint32_t)pd3[-1] * 0xFFFFu + cv;
                    ev[10] = (uint32_t)pd2[-2] * 0xFFFFu + cv;
                    ev[11] = (uint32_t)pd1[-3] * 0xFFFFu + cv;
                    ev[12] = (uint32_t)p[-3] * 0xFFFFu + cv;
                    ev[13] = (uint32_t)pu1[-3] * 0xFFFFu + cv;
                    ev[14] = (uint32_t)pu2[-2] * 0xFFFFu + cv;
                    ev[15] = (uint32_t)pu3[-1] * 0xFFFFu + cv;
                    const uint32_t best = best_arc_min_x2(ev);
                    const int s = (int)max(best & 0xFFFFu, best >> 16) - 257;
                    if (s >= t) {  // some arc of 9 has all |differences| > t with one sign: a corner, and s is its score
                        score[(e >> 8) * TW + (e & 255)] = (uint8_t)s;
                        const int o = atomicAdd(&cnt[2], 1);
                        if (o < FS_LIST2) list2[o] = (uint16_t)e;
                    }
                }
                __syncthreads();
                if (tid == 0) cnt[1] = 0;  // ordered before the next chunk's A2 by the barrier after its A1
                y0 = y1;
            }
            // ---- C: 3x3 non-max suppression among pixels of the same cell ----
            const int n2 = cnt[2];
            if (n2 <= FS_LIST2) {
                for (int i = n2done + tid; i < n2; i += FS_THREADS) {
                    const int e = list2[i];
                    nms_one(e, score[(e >> 8) * TW + (e & 255)]);
                }
                n2done = n2;
            } else {  // more corners than the list holds: scan the score tile of the active columns instead
                for (int y = warp; y < dH; y += FS_THREADS / 32) {
                    const uint32_t* srow = reinterpret_cast<const uint32_t*>(score) + (y + 3) * TWq + wq0;
                    for (int w = lane; w < nW; w += 32) {
                        const uint32_t sc = srow[w] & ((act32[wq0 + w] >> 7) * 0xFFu);
                        if (sc == 0) continue;
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const int s = (sc >> (8 * j)) & 0xFF;
                            if (s != 0) nms_one(((y + 3) << 8) | (((wq0 + w) << 2) + j), s);
                        }
                    }
                }
            }
            __syncthreads();
            if (pass == 1) break;
            // ---- cells without a keypoint go again at minThFAST ----
            uint32_t empty = 0;
            for (int j = 0; j < nc; ++j)
                if (cnt[6 + j] == 0) empty |= 1u << j;
            empty &= validCells;
            if (empty == 0) break;
            pass2 = true;
            for (int x = tid; x < TW; x += FS_THREADS) {
                const int cell = cellOfCol[x];
                act[x] = (cell != 255 && ((empty >> cell) & 1u)) ? 0x80 : 0;
            }
            __syncthreads();
            // the word columns pass 2 walks (every warp writes the same list, so no further barrier is needed)
            nCols = 0;
            for (int q = 0; q < nWG; ++q) {
                const int w = (q << 5) + lane;
                const bool a = w < nW && act32[wq0 + w] != 0;
                const uint32_t bal = __ballot_sync(0xffffffffu, a);
                if (a) colList[nCols + __popc(bal & laneLt)] = (uint8_t)w;
                nCols += __popc(bal);
            }
            __syncwarp();
        }
        // ---- D: cv::FAST's row-major order inside every cell, straight from the bitmap: one warp per cell, lane = row of
        //      the cell; a warp prefix sum of the row counts ranks every keypoint and the keys go to the cell's slot ----
        const int n2all = cnt[2];
        {
            int* outCnt = cellCnt + (long long)f * g.nCellsTotal + cellGlobal0;
            const int cellCap = L.cellCap;
            uint32_t* outKeys = cellKeys + (long long)f * g.keysPerFrame + L.keyBase + (long long)(ci * L.nCols + cj0) * cellCap;
            for (int cj = warp; cj < nc; cj += FS_THREADS / 32) {
                const int cx0 = X0 + min(cj * wCell, dW), cx1 = X0 + min(cj * wCell + wCell, dW);  // tile columns of the cell
                int run = 0;
                if (cx1 > cx0 && cnt[6 + cj] > 0) {
                    // the cell's columns as one 64-bit row mask (a cell is < 64 px wide): bit i = column cx0 + i
                    const int k0 = cx0 >> 5, cw = cx1 - cx0;
                    const uint32_t sh = (uint32_t)(cx0 & 31);
                    const uint32_t wLo = cw >= 32 ? 0xFFFFFFFFu : (1u << cw) - 1u, wHi = cw > 32 ? (1u << (cw - 32)) - 1u : 0u;
                    for (int r0 = 0; r0 < dH; r0 += 32) {
                        const int yT = r0 + lane + 3;
                        uint32_t lo = 0, hi = 0;
                        if (r0 + lane < dH) {
                            const uint32_t* kr = keep + yT * BW + k0;
                            const uint32_t a0 = kr[0], a1 = k0 + 1 < BW ? kr[1] : 0u, a2 = k0 + 2 < BW ? kr[2] : 0u;
                            lo = __funnelshift_r(a0, a1, sh) & wLo;
                            hi = __funnelshift_r(a1, a2, sh) & wHi;
                        }
                        const int c = __popc(lo) + __popc(hi);
                        int incl = c;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const int v = __shfl_up_sync(0xffffffffu, incl, o);
                            if (lane >= o) incl += v;
                        }
                        int rank = run + incl - c;
                        run += __shfl_sync(0xffffffffu, incl, 31);
                        const uint8_t* srow = score + yT * TW + cx0;
                        const uint32_t keyRow = pack_key(cx0 - xo + cj0 * wCell, yT + ci * L.hCell, 0);
                        while (lo) {
                            const int i = __ffs(lo) - 1;
                            lo &= lo - 1;
                            if (rank < cellCap) outKeys[cj * cellCap + rank] = keyRow + (uint32_t)i + ((uint32_t)srow[i] << 24);
                            ++rank;
                        }
                        while (hi) {
                            const int i = __ffs(hi) + 31;
                            hi &= hi - 1;
                            if (rank < cellCap) outKeys[cj * cellCap + rank] = keyRow + (uint32_t)i + ((uint32_t)srow[i] << 24);
                            ++rank;
                        }
                    }
                }
                if (lane == 0) {
                    if (run > cellCap) {  // cannot happen: cellCap is the NMS packing bound of the cell (cell_capacity)
                        atomicExch(dStatus, ORBX_E_OVERFLOW);
                        run = cellCap;
                    }
                    outCnt[cj] = run;
                }
            }
        }
        // ---- leave the score tile, the bitmap and the column mask clean for the next frame (D reads the scores of the
        //      kept corners, so the clean-up waits for it) ----
        if (f + 1 < fEnd) {
            if (pass2)
                for (int x = tid; x < TW; x += FS_THREADS) act[x] = cellOfCol[x] != 255 ? 0x80 : 0;
            __syncthreads();
            if (n2all <= FS_LIST2) {
                for (int i = tid; i < n2all; i += FS_THREADS) {
                    const int e = list2[i];
                    score[(e >> 8) * TW + (e & 255)] = 0;
                }
            } else {
                for (int i = tid; i < TNa / 16; i += FS_THREADS) reinterpret_cast<uint4*>(score)[i] = make_uint4(0u, 0u, 0u, 0u);
            }
            for (int i = tid; i < TH * BW; i += FS_THREADS) keep[i] = 0u;
            __syncthreads();
        }
    }
}

size_t fast_smem_bytes(const FrameGeom& g) {
    const size_t TN = (size_t)g.fastBoxW * g.fastBoxH, TNa = (TN + 127) & ~(size_t)127;
    const size_t BW = (g.fastBoxW + 31) >> 5;
    return 3 * TNa + 2 * 8 + 32 * 4 + (size_t)g.fastBoxH * BW * 4 + 2 * g.fastBoxW + 64 + (FS_WL + FS_LIST1 + FS_LIST2) * 2;
}

// Cells per strip of a level: as many as ORBX_FS_MAXCELLS allows while the TMA box (strip window + 15 bytes of origin
// slack) stays within 256 columns.  A1 walks (row, word) pairs as a flat index, so lane utilisation does not depend
// on the strip width; the width trades shared memory per CTA (occupancy) against halo re-reads.
int fast_cells_per_strip(int wCell, int nCols) {
    int best = 1;
    for (int c = 1; c <= ORBX_FS_MAXCELLS && c <= FAST_MAX_STRIP_CELLS && c <= nCols; ++c) {
        if (c * wCell + 6 + 15 > 256) break;
        best = c;
    }
    return best;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is per device: remember what each device has been given
int ensure_dynamic_smem(const void* func, size_t bytes, size_t* perDevice /* [ORBX_MAX_DEVICES] */) {
    static std::mutex mu;
    if (bytes <= 48 * 1024) return ORBX_OK;
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < ORBX_MAX_DEVICES && perDevice[dev] >= bytes) return ORBX_OK;
    ORBX_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    if (dev >= 0 && dev < ORBX_MAX_DEVICES) perDevice[dev] = bytes;
    return ORBX_OK;
}

int launch_fast(const FrameGeom& g, const TmaMaps& maps, const FastStrip* dStrips, int batch, int* cellCnt,
                uint32_t* cellKeys, int* dStatus, cudaStream_t s, int onlyLevel) {
    // onlyLevel >= 0: the strips of that pyramid level only (single frames: a level is searched as soon as it exists)
    const int strip0 = onlyLevel >= 0 ? g.lv[onlyLevel].stripBase : 0, nStrips = onlyLevel >= 0 ? g.lv[onlyLevel].stripCount : g.nStrips;
    if (nStrips == 0) return ORBX_OK;
    const size_t smem = fast_smem_bytes(g);
    static size_t configured[ORBX_MAX_DEVICES] = {};
    int rc = ensure_dynamic_smem((const void*)k_fast_strips, smem, configured);
    if (rc) return rc;
    // frames per CTA: ORBX_FS_FPC while the grid keeps >= 4 waves of 148 SMs x 7 CTAs
    int fpc = ORBX_FS_FPC;
    while (fpc > 1 && (long long)nStrips * ((batch + fpc - 1) / fpc) < 4LL * 148 * ORBX_FS_MINBLOCKS) fpc >>= 1;
    dim3 grid(nStrips, (batch + fpc - 1) / fpc);
    ORBX_CUDA(launch_kernel(k_fast_strips, grid, dim3(FS_THREADS), smem, s, g.pdl && !g.plainLaunch, g, maps, dStrips + strip0, cellCnt, cellKeys,
                            dStatus, batch, fpc));
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
