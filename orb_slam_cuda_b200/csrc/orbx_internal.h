// csrc/orbx_internal.h — shared declarations of liborbx.so (not installed; the public surface is include/orbx.h).
//
// Device data layout (all per extractor handle, B = batch capacity, l = pyramid level):
//   plane[l]   : uint8  [B][h_l + 38][pitch_l]   padded pyramid plane; interior pixel (x, y) lives at byte
//                                               (y + 19) * pitch_l + (x + 32): the 19-px reflect-101 border of
//                                               the reference (EDGE_THRESHOLD, src/ORBextractor.cc:97) sits at
//                                               bytes 13..31, so interior rows start 32-byte aligned.
//   blur[l]    : uint8  [B][h_l][bpitch_l]        7x7 sigma-2 blurred level, interior only (descriptor taps stay
//                                               >= 1 px inside, see k_describe.cu).
//   cellCnt    : int32  [B][nCellsTotal]          FAST keypoints per 30-px cell (all levels concatenated)
//   cellKeys   : uint32 [B][level][cell][cellCap_l] packed (resp << 24 | y << 12 | x), x,y relative to minBorder,
//                                               row-major inside the cell (cv::FAST output order)
//   lvlCnt     : int32  [B][nlevels]              keypoints kept by the quadtree
//   lvlKeys    : uint32 [B][slotsTotal]           packed kept keypoints, level l at slotBase[l], quota_l + 4 slots
//   qtScratch  : uint32 [B][nlevels][2][candCap_l] key ping-pong buffers of the quadtree
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/orbx.h"

namespace orbx {

constexpr int EDGE = 19;         // EDGE_THRESHOLD
constexpr int XOFF = 32;         // byte column of interior x = 0 inside a plane row
constexpr int PADL = XOFF - EDGE;  // byte column of padded x = 0  (13)
constexpr int HALF_PATCH = 15;
constexpr int PATCH = 31;
// Key slots per FAST cell.  cv::FAST's non-maximum suppression keeps a corner only if its score is strictly greater than
// those of its 8 neighbours, so no two kept corners touch: a cell that evaluates w x h positions keeps at most
// ceil(w / 2) * ceil(h / 2).  Sizing the slot by that bound means a cell can never overflow, whatever the image (white
// noise included).  Rounded up to a multiple of 32.
inline int cell_capacity(int wCell, int hCell) { return (((wCell + 1) / 2) * ((hCell + 1) / 2) + 31) / 32 * 32; }
constexpr int FAST_MAX_STRIP_CELLS = 8;  // cells one FAST CTA handles (a strip of one cell row)
#ifndef ORBX_BLUR_TH
#define ORBX_BLUR_TH 32  // output rows of a blur tile (the TMA box is 160 x (ORBX_BLUR_TH + 6))
#endif
constexpr int BLUR_TILE_H = ORBX_BLUR_TH;
#ifndef ORBX_RZT_TH
#define ORBX_RZT_TH 24  // destination rows of a TMA-tiled resize tile (<= 32: one lane prepares one row's constants); 16 / 24 / 32
                       // measured 0.764 / 0.750 / 0.766 ms for the pyramid stage (profiles/r02_variants.txt)
#endif
constexpr int RZT_TH = ORBX_RZT_TH;
static_assert(RZT_TH >= 4 && RZT_TH <= 32, "one lane per tile row");
constexpr int RZT_BW = 192, RZT_BH = (RZT_TH * 5 + 3) / 4 + 8;  // source box: 128 x 1.2 + alignment slack; TH x 1.25 + slack (48 rows for TH 32)
constexpr int QT_THREADS = 256;
constexpr int MAX_ROOTS = 64;
constexpr int ORBX_MAX_DEVICES = 64;  // per-device attribute caches (ensure_dynamic_smem)

struct LevelGeom {
    int w, h;               // interior size
    int pitch, rows;        // plane row pitch (bytes), rows = h + 38
    long long planeStride;  // bytes per frame in plane[l]
    int bpitch;
    long long blurStride;   // bytes per frame in blur[l]
    // FAST grid (src/ORBextractor.cc:1133-1147)
    int minBX, minBY, maxBX, maxBY, nCols, nRows, wCell, hCell;
    int stripBase, stripCount;  // this level's FAST strips in the strip table
    int cellBase;           // first cell of this level in the per-frame cell arrays
    int cellCap;            // key slots per cell of this level = the most corners FAST + 3x3 NMS can leave in a cell (cell_capacity)
    long long keyBase;      // first key slot (uint32) of this level inside one frame's cellKeys
    int quota;              // mnFeaturesPerLevel[l]
    int slotBase;           // first slot of this level in lvlKeys
    int candCap;            // capacity of one quadtree key buffer
    long long qtBase;       // offset (in uint32) of this level's scratch inside one frame's qtScratch
    float scale;            // mvScaleFactor[l]
    float size;             // (float)(int)(PATCH_SIZE * mvScaleFactor[l])
    // quadtree roots (src/ORBextractor.cc:894-914)
    int nIni;
    float hX;
    int blurTilesX, blurTileBase;  // 128x32 output tiles of the blur kernel
    int kpBase;                    // first warp slot of this level in the describe kernel's grid
    uint8_t* plane;
    uint8_t* blur;
};

struct FrameGeom {
    int nlevels;
    int nCellsTotal;
    long long keysPerFrame;   // uint32 per frame in cellKeys (sum over levels of cells x cellCap)
    int slotsTotal;
    long long qtFrameStride;  // uint32 per frame in qtScratch
    int fastBoxW, fastBoxH;   // TMA box of the FAST kernel
    int nStrips;              // FAST strips per frame (all levels)
    int blurTilesTotal;
    int iniTh, minTh;
    int patternVariant;       // ORBX_PATTERN_REFERENCE (0) / ORBX_PATTERN_CANONICAL (1): which bit_pattern_31_ the descriptors use
    int f0;                   // first scratch frame of this launch (lanes of the host-batch pipeline own disjoint ranges)
    int plainLaunch;          // with pdl: launch WITHOUT the programmatic attribute (first kernel of a forked stream; the kernel's
                              // griddep instructions are then no-ops) — the single-frame kernel shapes stay selected
    int pdl;                  // 1: the kernels of this launch sequence are chained by programmatic dependent launch (single-frame
                              //    latency path): each lets its successor become resident at once and waits for its predecessor
                              //    itself (griddep_wait) before it touches anything the predecessor wrote
    LevelGeom lv[ORBX_MAX_LEVELS];
};

struct FastStrip {  // a run of horizontally adjacent cells of one cell row, processed by one CTA of the FAST kernel
    uint16_t level, ci, cj0, nc;
};

struct alignas(64) TmaMaps {  // one descriptor per level, passed to kernels as a __grid_constant__ parameter
    CUtensorMap m[ORBX_MAX_LEVELS];
};

struct ResizeTab {  // one entry per destination column / row: source index pair + 11-bit coefficients
    uint16_t s0, s1;
    int16_t c0, c1;
};

void count_launch();
void set_error(const char* fmt, ...);

// <<<grid, block, smem, s>>> with, when pdl is set, cudaLaunchAttributeProgrammaticStreamSerialization: the kernel may start while
// its predecessor in the stream still runs; such a kernel calls griddep_wait() (tma.cuh) before its first dependent access
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, bool pdl, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define ORBX_CUDA(call)                                                                          \
    do {                                                                                         \
        cudaError_t e__ = (call);                                                                \
        if (e__ != cudaSuccess) {                                                                \
            ::orbx::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e__)); \
            return ORBX_E_CUDA;                                                                  \
        }                                                                                        \
    } while (0)

struct DevBuf {  // owning device allocation (grow-only)
    void* p = nullptr;
    size_t bytes = 0;
    inline int ensure(size_t n) {
        if (n <= bytes) return ORBX_OK;
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
        cudaError_t e = cudaMalloc(&p, n);
        if (e != cudaSuccess) {
            set_error("cudaMalloc(%zu) failed: %s", n, cudaGetErrorString(e));
            p = nullptr;
            return e == cudaErrorMemoryAllocation ? ORBX_E_OOM : ORBX_E_CUDA;
        }
        bytes = n;
        return ORBX_OK;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct PinBuf {  // owning pinned host allocation
    void* p = nullptr;
    size_t bytes = 0;
    int ensure(size_t n) {
        if (n <= bytes) return ORBX_OK;
        if (p) cudaFreeHost(p);
        p = nullptr;
        bytes = 0;
        cudaError_t e = cudaMallocHost(&p, n);
        if (e != cudaSuccess) {
            set_error("cudaMallocHost(%zu) failed: %s", n, cudaGetErrorString(e));
            p = nullptr;
            return ORBX_E_OOM;
        }
        bytes = n;
        return ORBX_OK;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        bytes = 0;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};


// ---- what the matcher entry points need from an extractor handle (stereo reads both pyramids) -------------------
struct ExtractorView {
    const FrameGeom* g;
    cudaStream_t stream;
    int device;
    const float* scale;
    const float* invScale;
    long long serial;  // 0 = nothing extracted yet
    int lastBatch;
    // the keypoints / descriptors of the last single-frame host call as they still sit on the device (null after any other call)
    const orbx_keypoint* lastKps;
    const uint8_t* lastDesc;
    int lastCount;
};
int extractor_view(orbx_extractor* ex, ExtractorView* v);

struct StereoLevels {  // interior pointers (pixel 0,0) of frame 0 of both extractors, per level
    const uint8_t* left[ORBX_MAX_LEVELS];
    const uint8_t* right[ORBX_MAX_LEVELS];
    int pitchL[ORBX_MAX_LEVELS], pitchR[ORBX_MAX_LEVELS];
    int widthR[ORBX_MAX_LEVELS];
    float scale[ORBX_MAX_LEVELS], invScale[ORBX_MAX_LEVELS];
    int nlevels;
};

// ---- kernel launchers (each file k_*.cu holds the kernels of one stage) ---------------------------------------
// ComputePyramid: level 0 = frame + reflect-101 border, level l = resize(level l-1) + border
int launch_pyramid_base(const FrameGeom& g, const uint8_t* dFrames, int batch, size_t stride, size_t frameStride,
                        cudaStream_t s);
int launch_pyramid_level(const FrameGeom& g, int level, const ResizeTab* xtab, const ResizeTab* ytab, int areaFast,
                         int batch, cudaStream_t s);
int launch_pyramid_level_tma(const FrameGeom& g, const TmaMaps& maps, int level, const ResizeTab* xtab, const ResizeTab* ytab,
                             const int* tileX0, const int* tileY0, int batch, cudaStream_t s, bool pdl = false);
// per-cell FAST-9 + threshold fallback + NMS
int launch_fast(const FrameGeom& g, const TmaMaps& maps, const FastStrip* dStrips, int batch, int* cellCnt,
                uint32_t* cellKeys, int* dStatus, cudaStream_t s, int onlyLevel = -1);
int fast_cells_per_strip(int wCell, int nCols);
// opt-in to > 48 KB of dynamic shared memory, once per (kernel, device): perDevice is the kernel's own
// size_t[ORBX_MAX_DEVICES] cache (the attribute is per device; the cache is guarded by a mutex)
int ensure_dynamic_smem(const void* func, size_t bytes, size_t* perDevice);
// DistributeOctTree per (frame, level)
int launch_quadtree(const FrameGeom& g, int batch, const int* cellCnt, const uint32_t* cellKeys, uint32_t* qtScratch,
                    int* lvlCnt, uint32_t* lvlKeys, int* dStatus, cudaStream_t s, int onlyLevel = -1);
int quadtree_smem_bytes(int maxQuota);
// stand-alone quadtree on a caller-supplied candidate list (debug / parity entry)
int launch_quadtree_single(const uint32_t* dKeysIn, int n, int minX, int maxX, int minY, int maxY, int N,
                           uint32_t* dScratch, int candCap, uint32_t* dOut, int* dCount, int* dStatus, cudaStream_t s);
// 7x7 sigma=2 fixed-point GaussianBlur of every level
int launch_blur(const FrameGeom& g, const TmaMaps& maps, int batch, cudaStream_t s);
// IC_Angle + steered BRIEF + scale-back + level-major compaction into the caller's output
int launch_describe(const FrameGeom& g, int batch, const int* lvlCnt, const uint32_t* lvlKeys, orbx_keypoint* dKps,
                    uint8_t* dDesc, int cap, int* dCounts, float* dLvlAngles, int* dStatus, cudaStream_t s);
// matchers
int launch_hamming_pairs(const uint8_t* a, const uint8_t* b, int n, int* dist, cudaStream_t s);
int launch_bruteforce(const uint8_t* q, int nq, const uint8_t* d, int nd, int th, float ratio, int* idx, int* d1, int* d2,
                      int* match, unsigned long long* dPartial, unsigned int* dCounters, cudaStream_t s);
size_t bruteforce_scratch_bytes(int nq, int nd);
size_t bruteforce_counter_bytes(int nq);
// score[i * ldScore + (colMap ? colMap[j] : j)]; colMap[j] < 0 skips database keyframe j
int launch_allpairs(const uint8_t* q, int kq, const uint8_t* db, int kd, int n, int th, float ratio, int* score, int ldScore,
                    const int* colMap, cudaStream_t s);

// candCnt[n] followed (16-byte aligned) by key[n][8] and idx[n][8] (uint32): see resolve_common.cuh
inline size_t resolve_aux_bytes(int n) { return ((size_t)n + 3) / 4 * 16 + (size_t)n * 64; }
int launch_search_for_initialization(const orbx_keypoint* k1, int n1, const uint8_t* desc1, const orbx_keypoint* k2, int n2,
                                     const uint8_t* desc2, const float bounds[4], float* prev, int* matches12, int window,
                                     float nnratio, int checkOri, int* cellOf, int* sorted, int* cellStart, uint32_t* cand,
                                     int candCap, int* candCnt, int* nmatches, int* dStatus, cudaStream_t s);
int launch_search_by_projection(int n1, const float* x3Dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mpDesc1,
                                const int* octave1, const float* angle1, const orbx_keypoint* k2, int n2, const uint8_t* desc2,
                                const float* uRight2, const uint8_t* occupied2, const float* scaleFactors, const float cam[4],
                                float mbf, const float bounds[4], float th, int direction, int checkOri, int thDist,
                                int noDepthTest, int* cellOf, int* sorted, int* cellStart, uint32_t* cand, int candCap,
                                int* candCnt, int* match21, int* nmatches, int* dStatus, cudaStream_t s);
int launch_search_by_projection_map(int n1, const float* proj, const float* viewCos, const int* level1, const uint8_t* valid1,
                                    const uint8_t* obs1, const uint8_t* mpDesc1, const orbx_keypoint* k2, int n2,
                                    const uint8_t* desc2, const float* uRight2, const uint8_t* occupied2, const float* scaleFactors,
                                    const float bounds[4], float th, float nnratio, int* cellOf, int* sorted, int* cellStart,
                                    uint32_t* cand, int candCap, int* candCnt, int* match21, int* nmatches, int* dStatus,
                                    cudaStream_t s);
int launch_search_by_projection_sim3(int n1, const float* uvr, const int* level1, const uint8_t* valid1, const uint8_t* ones1,
                                     const uint8_t* mpDesc1, const orbx_keypoint* k2, int n2, const uint8_t* desc2,
                                     const uint8_t* occupied2, const float bounds[4], int thDist, int* cellOf, int* sorted,
                                     int* cellStart, uint32_t* cand, int candCap, int* candCnt, int* match21, int* nmatches,
                                     int* dStatus, cudaStream_t s);
int launch_frame_grid(const orbx_keypoint* k, int n, const float bounds[4], int* cellOf, int* sorted, int* cellStart, cudaStream_t s);
int launch_search_by_bow(int nItems, const int* nodeOf, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                         const uint8_t* valid1, const uint8_t* desc1, const float* angle1, int n2, const uint8_t* valid2,
                         const uint8_t* desc2, const float* angle2, const int* candOff, uint32_t* cand, int* candCnt, int thDist,
                         float nnratio, int checkOri, int* match12, int* match21, int* nmatches, cudaStream_t s);
int launch_search_for_triangulation(int nItems, const int* nodeOf, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                                    const orbx_keypoint* k1, const uint8_t* desc1, const uint8_t* valid1, const uint8_t* stereo1, int n2,
                                    const orbx_keypoint* k2, const uint8_t* desc2, const uint8_t* valid2, const uint8_t* stereo2,
                                    const float* scaleFactors2, const float* levelSigma2, const float F12[9], float ex, float ey,
                                    int checkOri, int* match12, signed char* evBin, int* hist, int* nmatches, cudaStream_t s);
int launch_fuse_match(int n1, const float* uvr, const float* ur1, const int* level1, const uint8_t* valid1, const uint8_t* mpDesc1,
                      const orbx_keypoint* k2, int n2, const uint8_t* desc2, const float* uRight2, const float* invLevelSigma2,
                      const float bounds[4], int* cellOf, int* sorted, int* cellStart, int* bestIdx, int* bestDist, cudaStream_t s);
int launch_gray(const uint8_t* src, long long srcStride, int w, int h, int channels, int bgr, int shift, uint8_t* dst, long long dstStride,
                cudaStream_t s);
// Frame::UndistortKeyPoints: K = fx, fy, cx, cy; dist = k1 k2 p1 p2 [k3]
int launch_undistort(const orbx_keypoint* in, int n, const float K[4], const float* dist, int ndist, orbx_keypoint* out, cudaStream_t s);
int launch_distinctive(int nPoints, const int* start, const uint8_t* desc, int* bestIdx, int* bestMedian, cudaStream_t s);
int launch_stereo(const StereoLevels& lv, const orbx_keypoint* kL, int nL, const uint8_t* dL, const orbx_keypoint* kR, int nR,
                  const uint8_t* dR, float mbf, float maxD, float* uRight, float* depth, int* sad, cudaStream_t s);

// ---- packed keypoint helpers ---------------------------------------------------------------------------------
__host__ __device__ inline uint32_t pack_key(int x, int y, int resp) {
    return ((uint32_t)resp << 24) | ((uint32_t)y << 12) | (uint32_t)x;
}
__host__ __device__ inline int key_x(uint32_t k) { return (int)(k & 0xFFFu); }
__host__ __device__ inline int key_y(uint32_t k) { return (int)((k >> 12) & 0xFFFu); }
__host__ __device__ inline int key_resp(uint32_t k) { return (int)(k >> 24); }

}  // namespace orbx
