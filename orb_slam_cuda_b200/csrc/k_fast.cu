// csrc/k_fast.cu — per-30-px-cell FAST-9/16 with the iniThFAST -> minThFAST fallback and 3x3 score non-max
// suppression (reference src/ORBextractor.cc:1122-1149 grid set-up, :1256-1299 per-cell cv::FAST calls).
//
// One CTA per (cell, frame).  The cell window [ini, ini + cell + 6) (clipped to maxBorder, :1258-1275) is staged
// into shared memory by one TMA box load from the padded plane (box origin rounded down to 16 bytes); then, at threshold t = iniThFAST:
//   A. every detectable pixel (window minus cv::FAST's 3-px rim) takes the 4-point antipodal quick test; survivors
//      are appended to a shared list (warp-aggregated atomics),
//   B. survivors run the full 16-point segment test (bit-mask run detection) and, for corners, the OpenCV
//      cornerScore<16> value = max over the 16 arcs of 9 of the minimum |difference|, minus 1 (doubling minima),
//      written to a shared score tile (0 = not a corner),
//   C. corners strictly greater than their 8 neighbours in the score tile survive (rim pixels score 0, exactly
//      like the zero-initialised score rows of cv::FAST, so NMS never looks across a cell seam),
//   D. survivors are ranked by position, giving cv::FAST's row-major output order, and stored into the cell's slot.
// If no keypoint survives, A-D run again at minThFAST (:1281-1285).  Scores do not depend on the threshold.
//
// Roofline: each level pixel is read once from HBM/L2 (windows of neighbouring cells overlap by 6 px and hit L2);
// the kernel itself is integer-ALU / shared-memory bound, not HBM bound.
#include "orbx_internal.h"
#include "tma.cuh"

namespace orbx {

__constant__ int c_ringDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
__constant__ int c_ringDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// true iff the 16-bit circular mask holds a run of >= 9 set bits
__device__ __forceinline__ bool has_run9(uint32_t m16) {
    uint32_t x = m16 | (m16 << 16);
    uint32_t a = x & (x >> 1);
    a &= a >> 2;
    a &= a >> 4;
    a &= x >> 8;
    return (a & 0xFFFFu) != 0;
}

// max over the 16 circular arcs of length 9 of the minimum of e[]
__device__ __forceinline__ int best_arc_min(const int (&e)[16]) {
    int m2[16], m4[16], best = -1000;
#pragma unroll
    for (int k = 0; k < 16; ++k) m2[k] = min(e[k], e[(k + 1) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m4[k] = min(m2[k], m2[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int m8 = min(m4[k], m4[(k + 4) & 15]);
        best = max(best, min(m8, e[(k + 8) & 15]));
    }
    return best;
}

__global__ void __launch_bounds__(FAST_THREADS) k_fast_cells(const __grid_constant__ FrameGeom g,
                                                              const __grid_constant__ TmaMaps maps,
                                                              int* __restrict__ cellCnt, uint32_t* __restrict__ cellKeys,
                                                              int* __restrict__ dStatus) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int TW = g.fastBoxW, TH = g.fastBoxH, TN = TW * TH;
    uint8_t* tile = smem;
    uint8_t* score = smem + TN;
    uint16_t* list1 = reinterpret_cast<uint16_t*>(smem + 2 * TN);
    uint32_t* list2 = reinterpret_cast<uint32_t*>(smem + 4 * TN);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 4 * TN + CELL_CAP * 4);
    int* cnt = reinterpret_cast<int*>(bar + 1);  // cnt[0] = n1, cnt[1] = n2

    const int tid = threadIdx.x, lane = tid & 31;
    const int f = blockIdx.y;
    int cell = blockIdx.x, level = 0;
    while (level + 1 < g.nlevels && cell >= g.lv[level + 1].cellBase) ++level;
    const LevelGeom& L = g.lv[level];
    cell -= L.cellBase;
    const int ci = cell / L.nCols, cj = cell - ci * L.nCols;
    int* outCnt = cellCnt + (long long)f * g.nCellsTotal + blockIdx.x;
    uint32_t* outKeys = cellKeys + ((long long)f * g.nCellsTotal + blockIdx.x) * CELL_CAP;

    // window of this cell (src/ORBextractor.cc:1258-1275)
    const int iniY = L.minBY + ci * L.hCell, iniX = L.minBX + cj * L.wCell;
    const int maxY = min(iniY + L.hCell + 6, L.maxBY), maxX = min(iniX + L.wCell + 6, L.maxBX);
    const int ww = maxX - iniX, wh = maxY - iniY;
    if (iniY >= L.maxBY - 3 || iniX >= L.maxBX - 6 || ww < 7 || wh < 7) {
        if (tid == 0) *outCnt = 0;
        return;
    }
    if (tid == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(bar, (uint32_t)TN);
        tma_load_3d(tile, &maps.m[level], (XOFF + iniX) & ~15, EDGE + iniY, f, bar);  // box origin must be 16-byte aligned
    }
    const int dW = ww - 6, dH = wh - 6, npx = dW * dH;
    const int xo = (XOFF + iniX) & 15;  // window column 0 inside the box
    mbar_wait(bar, 0);

    int n2 = 0;
    for (int pass = 0; pass < 2; ++pass) {
        const int t = min(max(pass == 0 ? g.iniTh : g.minTh, 0), 255);
        if (tid < 2) cnt[tid] = 0;
        for (int i = tid; i < TN / 4; i += FAST_THREADS) reinterpret_cast<uint32_t*>(score)[i] = 0;
        __syncthreads();
        // ---- A: antipodal quick test on ring points 0, 4, 8, 12 ----
        for (int base = 0; base < npx; base += FAST_THREADS) {
            const int p = base + tid;
            bool pass4 = false;
            int pos = 0;
            if (p < npx) {
                const int y = p / dW, x = p - y * dW;
                pos = (y + 3) * TW + x + 3 + xo;
                const int v = tile[pos];
                const int r0 = tile[pos + 3 * TW], r8 = tile[pos - 3 * TW], r4 = tile[pos + 3], r12 = tile[pos - 3];
                const int hi = v + t, lo = v - t;
                const bool br = (r0 > hi || r8 > hi) && (r4 > hi || r12 > hi);
                const bool dk = (r0 < lo || r8 < lo) && (r4 < lo || r12 < lo);
                pass4 = br || dk;
            }
            const unsigned m = __ballot_sync(0xffffffffu, pass4);
            if (m) {
                const int lead = __ffs(m) - 1;
                int off = 0;
                if (lane == lead) off = atomicAdd(&cnt[0], __popc(m));
                off = __shfl_sync(0xffffffffu, off, lead);
                if (pass4) list1[off + __popc(m & ((1u << lane) - 1))] = (uint16_t)pos;
            }
        }
        __syncthreads();
        const int n1 = cnt[0];
        // ---- B: full segment test + corner score ----
        for (int i = tid; i < n1; i += FAST_THREADS) {
            const int pos = list1[i];
            const int v = tile[pos];
            int d[16];
            uint32_t dark = 0, bright = 0;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                d[k] = v - (int)tile[pos + c_ringDy[k] * TW + c_ringDx[k]];
                dark |= (d[k] > t ? 1u : 0u) << k;
                bright |= (d[k] < -t ? 1u : 0u) << k;
            }
            int s = -1;
            if (has_run9(dark)) s = best_arc_min(d) - 1;
            if (has_run9(bright)) {
#pragma unroll
                for (int k = 0; k < 16; ++k) d[k] = -d[k];
                s = max(s, best_arc_min(d) - 1);
            }
            if (s > 0) score[pos] = (uint8_t)s;
        }
        __syncthreads();
        // ---- C: 3x3 non-max suppression ----
        for (int i = tid; i < n1; i += FAST_THREADS) {
            const int pos = list1[i];
            const int s = score[pos];
            if (s == 0) continue;
            const uint8_t* c = score + pos;
            if (s > c[-1] && s > c[1] && s > c[-TW - 1] && s > c[-TW] && s > c[-TW + 1] && s > c[TW - 1] && s > c[TW] &&
                s > c[TW + 1]) {
                const int o = atomicAdd(&cnt[1], 1);
                if (o < CELL_CAP) list2[o] = ((uint32_t)pos << 8) | (uint32_t)s;
            }
        }
        __syncthreads();
        n2 = cnt[1];
        if (n2 > 0) break;
        __syncthreads();  // everyone has read cnt[] before the next pass clears it
    }
    if (n2 > CELL_CAP) {
        if (tid == 0) atomicExch(dStatus, ORBX_E_OVERFLOW);
        n2 = CELL_CAP;
    }
    // ---- D: row-major order (positions are unique, rank = number of smaller positions) ----
    for (int i = tid; i < n2; i += FAST_THREADS) {
        const uint32_t e = list2[i];
        int rank = 0;
        for (int j = 0; j < n2; ++j) rank += (list2[j] < e) ? 1 : 0;
        const int pos = (int)(e >> 8);
        const int y = pos / TW, x = pos - y * TW;
        outKeys[rank] = pack_key(x - xo + cj * L.wCell, y + ci * L.hCell, (int)(e & 0xFFu));
    }
    if (tid == 0) *outCnt = n2;
}

size_t fast_smem_bytes(const FrameGeom& g) { return (size_t)4 * g.fastBoxW * g.fastBoxH + CELL_CAP * 4 + 8 + 16; }

int launch_fast(const FrameGeom& g, const TmaMaps& maps, int batch, int* cellCnt, uint32_t* cellKeys, int* dStatus,
                cudaStream_t s) {
    const size_t smem = fast_smem_bytes(g);
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        ORBX_CUDA(cudaFuncSetAttribute(k_fast_cells, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    dim3 grid(g.nCellsTotal, batch);
    k_fast_cells<<<grid, FAST_THREADS, smem, s>>>(g, maps, cellCnt, cellKeys, dStatus);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
