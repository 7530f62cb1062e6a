// csrc/extractor.cu — host side of the extractor handle behind include/orbx.h: constructor tables, per-geometry
// device layout, TMA descriptors, the launch sequence of operator() and the C entry points.
//
// Reference lines restated here (host arithmetic only; pixel work is in k_*.cu):
//   ORBextractor::ORBextractor   src/ORBextractor.cc:496-557   scale tables, per-level quotas, umax
//   ComputePyramid               src/ORBextractor.cc:1843-1844 level sizes = cvRound((float)cols * mvInvScaleFactor)
//   ComputeKeyPointsOctTree      src/ORBextractor.cc:1133-1147 borders and the 30-px cell grid
//   DistributeOctTree            src/ORBextractor.cc:894-896   number and width of the root nodes
//   cv::resize INTER_LINEAR 8U   (OpenCV; call site :1852)     per-column / per-row index + 11-bit coefficient tables
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "orbx_internal.h"

namespace orbx {

static thread_local std::string g_err;
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
}
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static inline int cv_round(float v) { return (int)lrintf(v); }   // cvRound: round half to even
static inline int cv_round(double v) { return (int)lrint(v); }
static inline short sat_short(float v) {
    const int i = cv_round(v);
    return (short)std::min(std::max(i, (int)SHRT_MIN), (int)SHRT_MAX);
}
static inline long long round_up(long long v, long long a) { return (v + a - 1) / a * a; }

}  // namespace orbx

using namespace orbx;

#ifndef ORBX_BLUR_OVERLAP_DEFAULT
#define ORBX_BLUR_OVERLAP_DEFAULT 0  // 1: GaussianBlur on a side stream beside FAST + quadtree; ORBX_BLUR_OVERLAP overrides
#endif
#ifndef ORBX_DESC_SUB_DEFAULT
#define ORBX_DESC_SUB_DEFAULT 0  // frames per blur + describe sub-batch (0 = whole batch); ORBX_DESC_SUB overrides
#endif

struct orbx_extractor {
    orbx_params prm{};
    int device = 0;
    cudaStream_t stream = nullptr;
    // constructor tables (src/ORBextractor.cc:501-555)
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> quota;
    int umax[16];
    // geometry-dependent state
    int w = 0, h = 0;        // frame size the layout was built for
    int batchCap = 0;        // frames the scratch holds
    int outCap = 0;          // keypoints per frame the output staging holds
    FrameGeom g{};
    TmaMaps fastMaps{}, blurMaps{}, resizeMaps{};  // resizeMaps.m[l]: SOURCE level l of the TMA-tiled resize
    std::vector<int> areaFast;
    DevBuf planes[ORBX_MAX_LEVELS], blurs[ORBX_MAX_LEVELS], xtab[ORBX_MAX_LEVELS], ytab[ORBX_MAX_LEVELS], rzTiles[ORBX_MAX_LEVELS];
    int rzTilesX[ORBX_MAX_LEVELS] = {};  // rzTiles[l] = tileX0[tilesX] | tileY0[tilesY]
    DevBuf strips, cellCnt, cellKeys, qtScratch, lvlCnt, lvlKeys, lvlAngles, status, dDesc, dCounts;
    DevBuf dFrames[2], dKps[2];  // device input / output staging of host-buffer calls, one set per in-flight call (slot)
    PinBuf hFrames, hKps, hDesc, hCounts, hStatus;
    // orbx_extract_batch_async: up to two calls in flight; slot k of a call = its sequence number & 1
    struct Pending {
        cudaEvent_t done = nullptr;  // recorded after the call's last D2H
        const int* hst = nullptr;    // pinned status word of the call
        int cap = 0;
    } pend[2];
    int pendHead = 0, pendCount = 0;  // oldest in-flight slot, calls in flight
    // host copies of pyramid levels for mvImagePyramid readers
    std::vector<uint8_t> hostLevel[ORBX_MAX_LEVELS];
    int hostLevelFrame[ORBX_MAX_LEVELS];
    long long hostLevelSerial[ORBX_MAX_LEVELS];
    long long serial = 0;    // bumped by every extract call
    // host-buffer batches are cut into chunks pipelined over three streams: H2D | kernels (stream) | D2H
    cudaStream_t h2dStream = nullptr, d2hStream = nullptr, stream2 = nullptr;  // stream2: kernels of odd chunks
    std::vector<cudaEvent_t> pipeEv;  // 2 per chunk: frames landed, kernels done
    // single-frame host calls replay a captured CUDA graph (memset + H2D + 13 kernels + one D2H): one launch per frame
    struct GraphKey {
        int w = 0, h = 0, cap = 0;
        const void *hIn = nullptr, *hOut = nullptr, *dIn = nullptr, *dOut = nullptr, *plane0 = nullptr;
        bool operator==(const GraphKey& o) const {
            return w == o.w && h == o.h && cap == o.cap && hIn == o.hIn && hOut == o.hOut && dIn == o.dIn && dOut == o.dOut && plane0 == o.plane0;
        }
    } graphKey;
    cudaGraphExec_t graphExec = nullptr;
    long long graphKernels = 0;  // kernel nodes of graphExec
    int lastKernels = 0;         // kernels the last run_pipeline call launched
    bool graphSeen = false, graphOk = true;
    bool timing = false;
    bool timed = false;      // the last call recorded events
    cudaEvent_t ev[ORBX_NUM_STAGES + 1] = {};
    std::vector<cudaEvent_t> subEv;  // timing of the blur / describe sub-batches: (blur done, describe done) per sub-batch
    int nSub = 0;
    // GaussianBlur needs the pyramid only, so it runs on a side stream beside FAST + quadtree and joins before the descriptors
    struct Side {
        cudaStream_t main = nullptr, side = nullptr;
        cudaEvent_t fork = nullptr, join = nullptr, t0 = nullptr, t1 = nullptr;
    };
    std::vector<Side> sides;  // one per stream the pipeline has been launched on
    // single frames: every pyramid level is searched (FAST + quadtree) on a stream of its own as soon as it exists, the blur
    // runs on another one; the main stream joins them before the descriptors
    struct Dag {
        cudaStream_t main = nullptr, lvl[ORBX_MAX_LEVELS] = {}, blur = nullptr;
        cudaEvent_t ready[ORBX_MAX_LEVELS] = {}, done[ORBX_MAX_LEVELS] = {}, blurDone = nullptr;
    };
    std::vector<Dag> dags;
    bool overlapped = false;  // the last timed call ran the blur on a side stream (stage_ms reads t0 / t1)
    int timedSide = -1;
    int lastBatch = 0;
    int lastScratchBase = 0; // scratch frame holding frame `lastFrameBase` of the caller's batch
    int lastFrameBase = 0;   // index (within the caller's batch) of frame 0 of the scratch: > 0 after a chunked host batch
    int patternVariant = ORBX_PATTERN_REFERENCE;
    // output of the last synchronous single-frame host call, still on the device (orbx_stereo_match_last reads it)
    const orbx_keypoint* lastDevKps = nullptr;
    const uint8_t* lastDevDesc = nullptr;
    int lastDevCount = 0;
};

// ---- constructor tables (src/ORBextractor.cc:501-555) -----------------------------------------------------------
static void build_tables(orbx_extractor* ex) {
    const orbx_params& p = ex->prm;
    const int nl = p.nlevels;
    const double scaleFactor = p.scale_factor;  // the member is a double initialised from the float argument
    ex->scale.assign(nl, 1.f);
    ex->sigma2.assign(nl, 1.f);
    for (int i = 1; i < nl; ++i) {
        ex->scale[i] = (float)(ex->scale[i - 1] * scaleFactor);
        ex->sigma2[i] = ex->scale[i] * ex->scale[i];
    }
    ex->invScale.resize(nl);
    ex->invSigma2.resize(nl);
    for (int i = 0; i < nl; ++i) {
        ex->invScale[i] = 1.0f / ex->scale[i];
        ex->invSigma2[i] = 1.0f / ex->sigma2[i];
    }
    ex->quota.assign(nl, 0);
    const float factor = (float)(1.0f / scaleFactor);
    float nDesired = p.nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        ex->quota[l] = cv_round(nDesired);
        sum += ex->quota[l];
        nDesired *= factor;
    }
    ex->quota[nl - 1] = std::max(p.nfeatures - sum, 0);
    // umax (:540-555)
    int* umax = ex->umax;
    int v, v0;
    const int vmax = (int)std::floor(HALF_PATCH * std::sqrt(2.f) / 2 + 1);
    const int vmin = (int)std::ceil(HALF_PATCH * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH * HALF_PATCH;
    for (v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
    for (v = HALF_PATCH, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
}

// cv::resize(INTER_LINEAR, CV_8U) coefficient tables for one axis: n destination samples from `sn` source samples
static void build_resize_axis(int sn, int dn, bool clampFrac, std::vector<ResizeTab>& tab) {
    const double inv = (double)dn / sn, scale = 1. / inv;
    tab.resize(dn);
    for (int d = 0; d < dn; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        ResizeTab t;
        if (clampFrac) {  // x axis: OpenCV zeroes the fraction when the index is clamped
            if (s < 0) { f = 0; s = 0; }
            if (s >= sn - 1) { f = 0; s = sn - 1; }
            t.s0 = (uint16_t)s;
            t.s1 = (uint16_t)std::min(s + 1, sn - 1);
        } else {          // y axis: rows are clamped, the fraction is kept
            t.s0 = (uint16_t)std::min(std::max(s, 0), sn - 1);
            t.s1 = (uint16_t)std::min(std::max(s + 1, 0), sn - 1);
        }
        t.c0 = sat_short((1.f - f) * 2048);
        t.c1 = sat_short(f * 2048);
        tab[d] = t;
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

// 3-D (byte column, row, frame) view of a padded plane with a (boxW x boxH x 1) box
static int make_plane_map(CUtensorMap* m, void* base, int pitch, int rows, long long planeStride, int batch, int boxW, int boxH) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
        return ORBX_E_CUDA;
    }
    cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)rows, (cuuint64_t)batch};
    cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)planeStride};
    cuuint32_t box[3] = {(cuuint32_t)boxW, (cuuint32_t)boxH, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (pitch %d rows %d stride %lld batch %d box %dx%d)", (int)r,
                  pitch, rows, planeStride, batch, boxW, boxH);
        return ORBX_E_CUDA;
    }
    return ORBX_OK;
}

// Builds the level geometry for a w x h frame and allocates scratch for `batch` frames.
static int configure(orbx_extractor* ex, int w, int h, int batch) {
    const int nl = ex->prm.nlevels;
    if (w == ex->w && h == ex->h && batch <= ex->batchCap) return ORBX_OK;
    if (w > 4000 || h > 4000) {
        set_error("frames larger than 4000 x 4000 are not supported (got %d x %d)", w, h);
        return ORBX_E_INVALID;
    }
    ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    batch = std::max(batch, ex->w == w && ex->h == h ? ex->batchCap : 1);
    // the layout is rebuilt in place: until it is complete the handle has NO geometry, so a failure below (OOM while growing,
    // an empty level, a TMA encode error) leaves a handle that reconfigures on the next call instead of one that launches on
    // half-built state
    ex->w = ex->h = 0;
    ex->batchCap = 0;
    ex->lastBatch = 0;
    if (ex->graphExec) {
        cudaGraphExecDestroy(ex->graphExec);
        ex->graphExec = nullptr;
    }
    ex->graphSeen = false;
    FrameGeom& g = ex->g;
    memset(&g, 0, sizeof g);
    g.nlevels = nl;
    g.iniTh = ex->prm.ini_th_fast;
    g.minTh = ex->prm.min_th_fast;
    long long keyBase = 0;
    int cellBase = 0, slotBase = 0, tileBase = 0, maxStripW = 0, maxHCell = 0;
    std::vector<FastStrip> strips;
    long long qtBase = 0;
    ex->areaFast.assign(nl, 0);
    for (int l = 0; l < nl; ++l) {
        LevelGeom& L = g.lv[l];
        L.w = cv_round((float)w * ex->invScale[l]);
        L.h = cv_round((float)h * ex->invScale[l]);
        if (L.w < 1 || L.h < 1) {
            set_error("pyramid level %d of a %d x %d frame is empty", l, w, h);
            return ORBX_E_INVALID;
        }
        L.pitch = (int)round_up(XOFF + L.w + EDGE + 16, 128);
        L.rows = L.h + 2 * EDGE;
        L.planeStride = round_up((long long)L.pitch * L.rows, 256);
        L.bpitch = (int)round_up(L.w, 128);
        L.blurStride = (long long)L.bpitch * L.h;
        L.minBX = L.minBY = EDGE - 3;
        L.maxBX = L.w - EDGE + 3;
        L.maxBY = L.h - EDGE + 3;
        const float width = (float)(L.maxBX - L.minBX), height = (float)(L.maxBY - L.minBY);
        const float W = 30;
        L.nCols = width > 0 ? (int)(width / W) : 0;
        L.nRows = height > 0 ? (int)(height / W) : 0;
        L.stripBase = (int)strips.size();
        if (L.nCols < 1 || L.nRows < 1) {  // the reference divides by zero here; such a level yields nothing
            L.nCols = L.nRows = 0;
            L.wCell = L.hCell = 1;
        } else {
            L.wCell = (int)std::ceil(width / L.nCols);
            L.hCell = (int)std::ceil(height / L.nRows);
            maxHCell = std::max(maxHCell, L.hCell);
            // FAST strips: every cell row is cut into runs of `per` cells, sizes balanced
            const int per = fast_cells_per_strip(L.wCell, L.nCols);
            const int nRuns = (L.nCols + per - 1) / per;
            for (int i = 0; i < L.nRows; ++i)
                for (int r = 0; r < nRuns; ++r) {
                    const int j0 = (int)((long long)L.nCols * r / nRuns), j1 = (int)((long long)L.nCols * (r + 1) / nRuns);
                    strips.push_back(FastStrip{(uint16_t)l, (uint16_t)i, (uint16_t)j0, (uint16_t)(j1 - j0)});
                    maxStripW = std::max(maxStripW, (j1 - j0) * L.wCell);
                }
        }
        L.stripCount = (int)strips.size() - L.stripBase;
        L.cellBase = cellBase;
        cellBase += L.nCols * L.nRows;
        L.cellCap = cell_capacity(L.wCell, L.hCell);
        L.keyBase = keyBase;
        keyBase += (long long)L.nCols * L.nRows * L.cellCap;
        L.quota = ex->quota[l];
        L.slotBase = slotBase;
        slotBase += L.quota + 4;
        L.candCap = (int)std::max<long long>(32, std::min<long long>((long long)L.nCols * L.nRows * L.cellCap, 1 << 20));
        L.qtBase = qtBase;
        qtBase += 2LL * L.candCap;
        L.scale = ex->scale[l];
        L.size = (float)(int)(PATCH * ex->scale[l]);  // const int scaledPatchSize = PATCH_SIZE*mvScaleFactor[level]
        L.nIni = 0;
        L.hX = 1.f;
        if (L.nCols > 0) {
            L.nIni = (int)std::round((float)(L.maxBX - L.minBX) / (L.maxBY - L.minBY));
            if (L.nIni > MAX_ROOTS) {
                set_error("level %d aspect ratio needs %d quadtree roots (max %d)", l, L.nIni, MAX_ROOTS);
                return ORBX_E_INVALID;
            }
            if (L.nIni >= 1) L.hX = (float)(L.maxBX - L.minBX) / L.nIni;
        }
        L.blurTilesX = (L.w + 127) / 128;
        L.blurTileBase = tileBase;
        tileBase += L.blurTilesX * ((L.h + BLUR_TILE_H - 1) / BLUR_TILE_H);
        L.kpBase = L.slotBase;
    }
    g.nCellsTotal = std::max(cellBase, 1);
    g.keysPerFrame = std::max<long long>(keyBase, 1);
    g.slotsTotal = slotBase;
    g.qtFrameStride = qtBase;
    g.blurTilesTotal = tileBase;
    g.nStrips = (int)strips.size();
    g.fastBoxW = (int)round_up(std::max(maxStripW, 1) + 6 + 15, 16);  // + alignment slack of the box origin
    g.fastBoxH = std::max(maxHCell, 1) + 6;
    if (g.fastBoxW > 256 || g.fastBoxH > 256 || g.fastBoxW * g.fastBoxH > 60000) {
        set_error("FAST cell window %d x %d exceeds the TMA box limit", g.fastBoxW, g.fastBoxH);
        return ORBX_E_INVALID;
    }
    // allocations
    int rc;
    for (int l = 0; l < nl; ++l) {
        LevelGeom& L = g.lv[l];
        if ((rc = ex->planes[l].ensure((size_t)L.planeStride * batch))) return rc;
        if ((rc = ex->blurs[l].ensure((size_t)L.blurStride * batch + 256))) return rc;
        L.plane = ex->planes[l].as<uint8_t>();
        L.blur = ex->blurs[l].as<uint8_t>();
        if (l > 0) {
            std::vector<ResizeTab> xt, yt;
            const LevelGeom& S = g.lv[l - 1];
            build_resize_axis(S.w, L.w, true, xt);
            build_resize_axis(S.h, L.h, false, yt);
            const double sx = 1. / ((double)L.w / S.w), sy = 1. / ((double)L.h / S.h);
            const int ix = cv_round(sx), iy = cv_round(sy);
            const bool area = std::abs(sx - ix) < DBL_EPSILON && std::abs(sy - iy) < DBL_EPSILON && ix == 2 && iy == 2;
            ex->areaFast[l] = area ? 1 : 0;
            if (area) {  // exact 2x decimation is routed to INTER_AREA by cv::resize: mean of the 2x2 block, which
                         // is the bilinear formula with all coefficients 1024 (((1024 * (1024 (p + q) >> 4)) >> 16) = p + q)
                for (int d = 0; d < L.w; ++d) xt[d] = ResizeTab{(uint16_t)(2 * d), (uint16_t)(2 * d + 1), 1024, 1024};
                for (int d = 0; d < L.h; ++d) yt[d] = ResizeTab{(uint16_t)(2 * d), (uint16_t)(2 * d + 1), 1024, 1024};
            }
            // the word-based kernel needs every group of 4 padded destination columns to read <= 8 consecutive source bytes
            bool fits = true;
            for (int bx = 0; bx < L.pitch && fits; bx += 4) {
                int lo = INT_MAX, hi = -1;
                for (int j = 0; j < 4; ++j) {
                    const int px = bx + j - PADL;
                    if (px < 0 || px >= L.w + 2 * EDGE) continue;
                    int d = px - EDGE;
                    while (d < 0 || d >= L.w) d = L.w == 1 ? 0 : (d < 0 ? -d : 2 * L.w - 2 - d);
                    lo = std::min(lo, (int)xt[d].s0);
                    hi = std::max(hi, std::max((int)xt[d].s0, (int)xt[d].s1));
                }
                if (hi >= 0 && hi - lo > 7) fits = false;
            }
            if (fits) ex->areaFast[l] |= 2;
            // TMA-tiled kernel: the source window of every 128-byte x RZT_TH-row destination tile must fit the RZT_BW x RZT_BH box
            // whose origin is 16-byte aligned in x
            static const bool noTma = getenv("ORBX_RESIZE_NO_TMA") != nullptr;
            if (fits && !noTma) {
                const int tilesX = (L.pitch + 127) / 128, tilesY = (L.rows + RZT_TH - 1) / RZT_TH;
                std::vector<int> org(tilesX + tilesY, 0);
                bool ok = true;
                auto refl = [](int d, int n) { while (d < 0 || d >= n) d = n == 1 ? 0 : (d < 0 ? -d : 2 * n - 2 - d); return d; };
                for (int tx = 0; tx < tilesX && ok; ++tx) {
                    int lo = INT_MAX, hi = -1;
                    for (int b = tx * 128; b < std::min(tx * 128 + 128, L.pitch); ++b) {
                        const int px = b - PADL;
                        if (px < 0 || px >= L.w + 2 * EDGE) continue;
                        const int d = refl(px - EDGE, L.w);
                        lo = std::min(lo, (int)xt[d].s0);
                        hi = std::max(hi, std::max((int)xt[d].s0, (int)xt[d].s1));
                    }
                    if (hi < 0) { lo = 0; hi = 0; }
                    org[tx] = (XOFF + lo) & ~15;
                    if (XOFF + hi + 12 - org[tx] > RZT_BW) ok = false;
                }
                for (int ty = 0; ty < tilesY && ok; ++ty) {
                    int lo = INT_MAX, hi = -1;
                    for (int r = ty * RZT_TH; r < std::min(ty * RZT_TH + RZT_TH, L.rows); ++r) {
                        const int d = refl(r - EDGE, L.h);
                        lo = std::min(lo, (int)yt[d].s0);
                        hi = std::max(hi, std::max((int)yt[d].s0, (int)yt[d].s1));
                    }
                    org[tilesX + ty] = EDGE + lo;
                    if (hi - lo + 1 > RZT_BH) ok = false;
                }
                if (ok) {
                    ex->areaFast[l] |= 4;
                    ex->rzTilesX[l] = tilesX;
                    if ((rc = ex->rzTiles[l].ensure(org.size() * sizeof(int)))) return rc;
                    ORBX_CUDA(cudaMemcpy(ex->rzTiles[l].p, org.data(), org.size() * sizeof(int), cudaMemcpyHostToDevice));
                }
            }
            if ((rc = ex->xtab[l].ensure(xt.size() * sizeof(ResizeTab)))) return rc;
            if ((rc = ex->ytab[l].ensure(yt.size() * sizeof(ResizeTab)))) return rc;
            ORBX_CUDA(cudaMemcpy(ex->xtab[l].p, xt.data(), xt.size() * sizeof(ResizeTab), cudaMemcpyHostToDevice));
            ORBX_CUDA(cudaMemcpy(ex->ytab[l].p, yt.data(), yt.size() * sizeof(ResizeTab), cudaMemcpyHostToDevice));
        }
    }
    if ((rc = ex->strips.ensure(std::max<size_t>(strips.size(), 1) * sizeof(FastStrip)))) return rc;
    ORBX_CUDA(cudaMemcpy(ex->strips.p, strips.data(), strips.size() * sizeof(FastStrip), cudaMemcpyHostToDevice));
    if ((rc = ex->cellCnt.ensure((size_t)batch * g.nCellsTotal * 4))) return rc;
    if ((rc = ex->cellKeys.ensure((size_t)batch * (size_t)g.keysPerFrame * 4))) return rc;
    if ((rc = ex->qtScratch.ensure((size_t)batch * g.qtFrameStride * 4))) return rc;
    if ((rc = ex->lvlCnt.ensure((size_t)batch * nl * 4))) return rc;
    if ((rc = ex->lvlKeys.ensure((size_t)batch * g.slotsTotal * 4))) return rc;
    if ((rc = ex->lvlAngles.ensure((size_t)batch * g.slotsTotal * 4))) return rc;
    if ((rc = ex->status.ensure(256))) return rc;
    if ((rc = ex->hStatus.ensure(64))) return rc;
    // TMA descriptors of the FAST windows and the blur windows
    for (int l = 0; l < nl; ++l) {
        const LevelGeom& L = g.lv[l];
        if ((rc = make_plane_map(&ex->fastMaps.m[l], L.plane, L.pitch, L.rows, L.planeStride, batch, g.fastBoxW, g.fastBoxH))) return rc;
        if ((rc = make_plane_map(&ex->blurMaps.m[l], L.plane, L.pitch, L.rows, L.planeStride, batch, 160, BLUR_TILE_H + 6))) return rc;
        if (l + 1 < nl && (ex->areaFast[l + 1] & 4) &&
            (rc = make_plane_map(&ex->resizeMaps.m[l], L.plane, L.pitch, L.rows, L.planeStride, batch, RZT_BW, RZT_BH)))
            return rc;
    }
    ex->w = w;
    ex->h = h;
    ex->batchCap = batch;
    for (int l = 0; l < ORBX_MAX_LEVELS; ++l) ex->hostLevelSerial[l] = -1;
    return ORBX_OK;
}

// ---- the launch sequence of operator() (src/ORBextractor.cc:1701-1809) on device-resident buffers ---------------
static int stage_sync(const char* stage, cudaStream_t s) {  // ORBX_DEBUG_SYNC=1: attribute device faults to a stage
    static const bool on = getenv("ORBX_DEBUG_SYNC") != nullptr;
    if (!on) return ORBX_OK;
    cudaError_t e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) {
        set_error("stage '%s' failed: %s", stage, cudaGetErrorString(e));
        return ORBX_E_CUDA;
    }
    return ORBX_OK;
}

static int run_pipeline(orbx_extractor* ex, const uint8_t* dFrames, int batch, size_t stride, size_t frameStride,
                        orbx_keypoint* dKps, uint8_t* dDesc, int cap, int* dCounts, int* dStatus, cudaStream_t s,
                        bool resetStatus = true, int f0 = 0) {
    FrameGeom g = ex->g;
    g.f0 = f0;  // scratch frames [f0, f0 + batch)
    g.patternVariant = ex->patternVariant;
    // Single frames: the launches of the sequence overlap (programmatic dependent launch) — a kernel is resident, with its
    // tables and constants loaded, when its predecessor completes.  ORBX_PDL=0 switches it off.
    static const int pdlPref = []() { const char* e = getenv("ORBX_PDL"); return e ? atoi(e) : 2; }();  // 1: resize chain only, 2: all
    const bool pdlOk = batch <= 2 && !getenv("ORBX_SYNC_STAGES") && !getenv("ORBX_BLUR_OVERLAP");
    const bool pdl = pdlOk && pdlPref >= 1;
    g.pdl = pdlOk && pdlPref >= 2 ? 1 : 0;
    g.plainLaunch = 0;
    const int nl = g.nlevels;
    int* cellCnt = ex->cellCnt.as<int>() + (size_t)f0 * g.nCellsTotal;
    uint32_t* cellKeys = ex->cellKeys.as<uint32_t>() + (size_t)f0 * (size_t)g.keysPerFrame;
    uint32_t* qtScratch = ex->qtScratch.as<uint32_t>() + (size_t)f0 * g.qtFrameStride;
    int* lvlCnt = ex->lvlCnt.as<int>() + (size_t)f0 * nl;
    uint32_t* lvlKeys = ex->lvlKeys.as<uint32_t>() + (size_t)f0 * g.slotsTotal;
    float* lvlAngles = ex->lvlAngles.as<float>() + (size_t)f0 * g.slotsTotal;
    int rc;
    if (resetStatus) ORBX_CUDA(cudaMemsetAsync(dStatus, 0, sizeof(int), s));
    if ((rc = stage_sync("input copy", s))) return rc;
    const bool tm = ex->timing && f0 == 0;
    ex->timed = tm;
    if (tm) ORBX_CUDA(cudaEventRecord(ex->ev[0], s));
    // Single frames, untimed: a DAG instead of a chain.  Level l is searched (FAST + quadtree, on stream lvl[l]) as soon as it
    // exists, while the levels below it are still being resized, and the blur runs beside the searches; only the descriptor
    // kernel needs everything.  Critical path: level 0 + FAST(0) + quadtree(0) + describe instead of the sum of all 13 kernels.
    static const bool dagPref = []() { const char* e = getenv("ORBX_DAG"); return e ? atoi(e) != 0 : true; }();
    const bool dag = g.pdl && dagPref && !tm && !getenv("ORBX_DEBUG_SYNC");
    orbx_extractor::Dag* D = nullptr;
    if (dag) {
        for (auto& d : ex->dags)
            if (d.main == s) D = &d;
        if (!D) {
            orbx_extractor::Dag d;
            d.main = s;
            for (int l = 0; l < nl; ++l) {
                ORBX_CUDA(cudaStreamCreateWithFlags(&d.lvl[l], cudaStreamNonBlocking));
                ORBX_CUDA(cudaEventCreateWithFlags(&d.ready[l], cudaEventDisableTiming));
                ORBX_CUDA(cudaEventCreateWithFlags(&d.done[l], cudaEventDisableTiming));
            }
            ORBX_CUDA(cudaStreamCreateWithFlags(&d.blur, cudaStreamNonBlocking));
            ORBX_CUDA(cudaEventCreateWithFlags(&d.blurDone, cudaEventDisableTiming));
            ex->dags.push_back(d);
            D = &ex->dags.back();
        }
    }
    auto search_level = [&](int l) -> int {  // FAST + quadtree of level l on its own stream, after ready[l]
        FrameGeom gl = g;
        gl.plainLaunch = 1;  // first kernel of the forked stream: an ordinary launch behind the event
        ORBX_CUDA(cudaEventRecord(D->ready[l], s));
        ORBX_CUDA(cudaStreamWaitEvent(D->lvl[l], D->ready[l], 0));
        int r = launch_fast(gl, ex->fastMaps, ex->strips.as<FastStrip>(), batch, cellCnt, cellKeys, dStatus, D->lvl[l], l);
        if (r) return r;
        gl.plainLaunch = g.lv[l].stripCount == 0;  // the quadtree follows its FAST kernel by programmatic dependent launch
        if ((r = launch_quadtree(gl, batch, cellCnt, cellKeys, qtScratch, lvlCnt, lvlKeys, dStatus, D->lvl[l], l))) return r;
        ORBX_CUDA(cudaEventRecord(D->done[l], D->lvl[l]));
        return ORBX_OK;
    };
    if ((rc = launch_pyramid_base(g, dFrames, batch, stride, frameStride, s))) return rc;                 // ComputePyramid
    if ((rc = stage_sync("Pyramid/Resize level 0", s))) return rc;
    if (dag && (rc = search_level(0))) return rc;
    for (int l = 1; l < nl; ++l) {
        if (ex->areaFast[l] & 4)
            rc = launch_pyramid_level_tma(g, ex->resizeMaps, l, ex->xtab[l].as<ResizeTab>(), ex->ytab[l].as<ResizeTab>(), ex->rzTiles[l].as<int>(),
                                          ex->rzTiles[l].as<int>() + ex->rzTilesX[l], batch, s, pdl);
        else
            rc = launch_pyramid_level(g, l, ex->xtab[l].as<ResizeTab>(), ex->ytab[l].as<ResizeTab>(), ex->areaFast[l], batch, s);
        if (rc) return rc;
        if (dag && (rc = search_level(l))) return rc;
    }
    if (dag) {
        FrameGeom gb = g;
        gb.plainLaunch = 1;
        ORBX_CUDA(cudaStreamWaitEvent(D->blur, D->ready[nl - 1], 0));  // the chain is sequential: the last level is the last to exist
        if ((rc = launch_blur(gb, ex->blurMaps, batch, D->blur))) return rc;                                // GaussianBlur
        ORBX_CUDA(cudaEventRecord(D->blurDone, D->blur));
        for (int l = 0; l < nl; ++l) ORBX_CUDA(cudaStreamWaitEvent(s, D->done[l], 0));
        ORBX_CUDA(cudaStreamWaitEvent(s, D->blurDone, 0));
        if ((rc = launch_describe(gb, batch, lvlCnt, lvlKeys, dKps, dDesc, cap, dCounts, lvlAngles, dStatus, s))) return rc;
        ex->lastKernels = 2 + (nl - 1) + nl + 2;  // level 0 (two kernels), resizes, quadtrees, blur, describe
        for (int l = 0; l < nl; ++l) ex->lastKernels += g.lv[l].stripCount > 0;  // + a FAST launch per level that has cells
        ex->timed = false;
        ex->serial++;
        ex->lastBatch = batch;
        ex->lastDevKps = nullptr;
        ex->lastDevDesc = nullptr;
        ex->lastScratchBase = f0;
        return ORBX_OK;
    }
    if ((rc = stage_sync("Pyramid/Resize", s))) return rc;
    if (tm) ORBX_CUDA(cudaEventRecord(ex->ev[1], s));
    static const int subPref = []() { const char* e = getenv("ORBX_DESC_SUB"); return e ? atoi(e) : ORBX_DESC_SUB_DEFAULT; }();
    static const bool overlapPref = []() { const char* e = getenv("ORBX_BLUR_OVERLAP"); return e ? atoi(e) != 0 : ORBX_BLUR_OVERLAP_DEFAULT != 0; }();
    const bool overlap = overlapPref && !(subPref > 0 && batch > subPref) && !getenv("ORBX_SYNC_STAGES");
    int sideIdx = -1;
    if (overlap) {
        for (size_t i = 0; i < ex->sides.size(); ++i)
            if (ex->sides[i].main == s) sideIdx = (int)i;
        if (sideIdx < 0) {
            orbx_extractor::Side sd;
            sd.main = s;
            ORBX_CUDA(cudaStreamCreateWithFlags(&sd.side, cudaStreamNonBlocking));
            ORBX_CUDA(cudaEventCreateWithFlags(&sd.fork, cudaEventDisableTiming));
            ORBX_CUDA(cudaEventCreateWithFlags(&sd.join, cudaEventDisableTiming));
            ORBX_CUDA(cudaEventCreate(&sd.t0));
            ORBX_CUDA(cudaEventCreate(&sd.t1));
            ex->sides.push_back(sd);
            sideIdx = (int)ex->sides.size() - 1;
        }
        const orbx_extractor::Side& sd = ex->sides[sideIdx];
        ORBX_CUDA(cudaEventRecord(sd.fork, s));
        ORBX_CUDA(cudaStreamWaitEvent(sd.side, sd.fork, 0));
        if (tm) ORBX_CUDA(cudaEventRecord(sd.t0, sd.side));
        if ((rc = launch_blur(g, ex->blurMaps, batch, sd.side))) return rc;                                  // GaussianBlur
        if (tm) ORBX_CUDA(cudaEventRecord(sd.t1, sd.side));
        ORBX_CUDA(cudaEventRecord(sd.join, sd.side));
    }
    if (tm) {
        ex->overlapped = overlap;
        ex->timedSide = sideIdx;
    }
    if ((rc = launch_fast(g, ex->fastMaps, ex->strips.as<FastStrip>(), batch, cellCnt, cellKeys, dStatus, s))) return rc;
    if ((rc = stage_sync("FAST+Grid", s))) return rc;
    if (tm) ORBX_CUDA(cudaEventRecord(ex->ev[2], s));
    if ((rc = launch_quadtree(g, batch, cellCnt, cellKeys, qtScratch, lvlCnt, lvlKeys, dStatus, s))) return rc;
    if ((rc = stage_sync("Make quadtree", s))) return rc;
    if (tm) ORBX_CUDA(cudaEventRecord(ex->ev[3], s));
    // GaussianBlur + descriptors run over SUB-BATCHES: the blurred levels of `sub` frames (1.47 MB per KITTI frame) are
    // still in the 126 MB L2 when the descriptor kernel gathers its 37 x 37 patches from them, instead of coming back from
    // HBM as they do when all frames are blurred first (ncu: 1.5 GB of DRAM reads for 62 MB of output on 512 frames).
    const int sub = subPref > 0 && batch > subPref ? (batch + (batch + subPref - 1) / subPref - 1) / ((batch + subPref - 1) / subPref) : batch;
    ex->nSub = 0;
    for (int b0 = 0; b0 < batch; b0 += sub) {
        const int nb = std::min(sub, batch - b0);
        FrameGeom gs = g;
        gs.f0 = g.f0 + b0;
        if (overlap) {
            ORBX_CUDA(cudaStreamWaitEvent(s, ex->sides[sideIdx].join, 0));
        } else {
            if ((rc = launch_blur(gs, ex->blurMaps, nb, s))) return rc;                                      // GaussianBlur
            if ((rc = stage_sync("Gaussian Blur", s))) return rc;
        }
        if (tm) {
            while ((int)ex->subEv.size() < 2 * (ex->nSub + 1)) {
                cudaEvent_t e;
                ORBX_CUDA(cudaEventCreate(&e));
                ex->subEv.push_back(e);
            }
            ORBX_CUDA(cudaEventRecord(ex->subEv[2 * ex->nSub], s));
        }
        if ((rc = launch_describe(gs, nb, lvlCnt + (size_t)b0 * nl, lvlKeys + (size_t)b0 * g.slotsTotal, dKps + (size_t)b0 * cap,
                                  dDesc + (size_t)b0 * cap * 32, cap, dCounts + b0, lvlAngles + (size_t)b0 * g.slotsTotal, dStatus, s)))
            return rc;
        if ((rc = stage_sync("Compute angle + ORB descriptor", s))) return rc;
        if (tm) ORBX_CUDA(cudaEventRecord(ex->subEv[2 * ex->nSub + 1], s));
        ex->nSub++;
    }
    if (tm) ORBX_CUDA(cudaEventRecord(ex->ev[5], s));
    ex->lastKernels = nl + 5 + (ex->nSub > 1 ? 2 * (ex->nSub - 1) : 0);
    ex->serial++;
    ex->lastBatch = batch;
    ex->lastDevKps = nullptr;  // set again by the single-frame host paths once their output is complete
    ex->lastDevDesc = nullptr;
    ex->lastScratchBase = f0;
    return ORBX_OK;
}

static int check_device(const orbx_extractor* ex) {
    if (!ex) {
        set_error("null extractor handle");
        return ORBX_E_INVALID;
    }
    ORBX_CUDA(cudaSetDevice(ex->device));
    return ORBX_OK;
}

namespace orbx {
int extractor_view(orbx_extractor* ex, ExtractorView* v) {
    if (!ex || !v) {
        set_error("null extractor handle");
        return ORBX_E_INVALID;
    }
    v->g = &ex->g;
    v->stream = ex->stream;
    v->device = ex->device;
    v->scale = ex->scale.data();
    v->invScale = ex->invScale.data();
    v->serial = ex->serial;
    v->lastBatch = ex->lastBatch;
    v->lastKps = ex->lastDevKps;
    v->lastDesc = ex->lastDevDesc;
    v->lastCount = ex->lastDevCount;
    return ORBX_OK;
}
}  // namespace orbx

extern "C" {

const char* orbx_last_error(void) { return g_err.c_str(); }
int orbx_version(void) { return 100; }
long long orbx_kernel_launches(void) { return g_launches.load(); }

int orbx_extractor_create(const orbx_params* p, orbx_extractor** out) {
    if (!p || !out) {
        set_error("orbx_extractor_create: null argument");
        return ORBX_E_INVALID;
    }
    *out = nullptr;
    if (p->nlevels < 1 || p->nlevels > ORBX_MAX_LEVELS || p->nfeatures < 1 || !(p->scale_factor > 1.0f) || p->nfeatures > 8000) {
        set_error("orbx_extractor_create: need 1 <= nlevels <= %d, 1 <= nfeatures <= 8000, scaleFactor > 1", ORBX_MAX_LEVELS);
        return ORBX_E_INVALID;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        set_error("no CUDA device: %s (this library has no CPU path)", e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        return ORBX_E_CUDA;
    }
    orbx_extractor* ex = new orbx_extractor();
    ex->prm = *p;
    if (p->device >= 0) ex->device = p->device;
    else ORBX_CUDA(cudaGetDevice(&ex->device));
    if (ex->device >= ndev) {
        set_error("device %d requested, %d present", ex->device, ndev);
        delete ex;
        return ORBX_E_INVALID;
    }
    ORBX_CUDA(cudaSetDevice(ex->device));
    cudaDeviceProp prop;
    ORBX_CUDA(cudaGetDeviceProperties(&prop, ex->device));
    if (prop.major < 10) {
        set_error("device %d is sm_%d%d; this library carries sm_100a code only", ex->device, prop.major, prop.minor);
        delete ex;
        return ORBX_E_CUDA;
    }
    ORBX_CUDA(cudaStreamCreateWithFlags(&ex->stream, cudaStreamNonBlocking));
    build_tables(ex);
    for (int l = 0; l < ORBX_MAX_LEVELS; ++l) ex->hostLevelSerial[l] = -1;
    if (p->width > 0 && p->height > 0) {
        int rc = configure(ex, p->width, p->height, std::max(1, p->max_batch));
        if (rc) {
            orbx_extractor_destroy(ex);
            return rc;
        }
    }
    *out = ex;
    return ORBX_OK;
}

void orbx_extractor_destroy(orbx_extractor* ex) {
    if (!ex) return;
    cudaSetDevice(ex->device);
    if (ex->stream) {
        cudaStreamSynchronize(ex->stream);
        cudaStreamDestroy(ex->stream);
    }
    for (cudaStream_t st : {ex->h2dStream, ex->d2hStream, ex->stream2})
        if (st) {
            cudaStreamSynchronize(st);
            cudaStreamDestroy(st);
        }
    for (cudaEvent_t e : ex->pipeEv) cudaEventDestroy(e);
    for (auto& pd : ex->pend)
        if (pd.done) cudaEventDestroy(pd.done);
    if (ex->graphExec) cudaGraphExecDestroy(ex->graphExec);
    for (cudaEvent_t e : ex->ev)
        if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : ex->subEv) cudaEventDestroy(e);
    for (auto& d : ex->dags) {
        for (int l = 0; l < ORBX_MAX_LEVELS; ++l) {
            if (d.lvl[l]) cudaStreamDestroy(d.lvl[l]);
            if (d.ready[l]) cudaEventDestroy(d.ready[l]);
            if (d.done[l]) cudaEventDestroy(d.done[l]);
        }
        if (d.blur) cudaStreamDestroy(d.blur);
        if (d.blurDone) cudaEventDestroy(d.blurDone);
    }
    for (auto& sd : ex->sides) {
        cudaStreamDestroy(sd.side);
        for (cudaEvent_t e : {sd.fork, sd.join, sd.t0, sd.t1}) cudaEventDestroy(e);
    }
    for (int l = 0; l < ORBX_MAX_LEVELS; ++l) {
        ex->planes[l].release(); ex->blurs[l].release(); ex->xtab[l].release(); ex->ytab[l].release(); ex->rzTiles[l].release();
    }
    for (DevBuf* b : {&ex->strips, &ex->cellCnt, &ex->cellKeys, &ex->qtScratch, &ex->lvlCnt, &ex->lvlKeys, &ex->lvlAngles, &ex->status,
                      &ex->dFrames[0], &ex->dFrames[1], &ex->dKps[0], &ex->dKps[1], &ex->dDesc, &ex->dCounts})
        b->release();
    for (PinBuf* b : {&ex->hFrames, &ex->hKps, &ex->hDesc, &ex->hCounts, &ex->hStatus}) b->release();
    delete ex;
}

int orbx_get_scale_tables(const orbx_extractor* ex, float* scale, float* inv, float* s2, float* inv2, int* quota) {
    if (!ex) { set_error("null extractor handle"); return ORBX_E_INVALID; }
    for (int l = 0; l < ex->prm.nlevels; ++l) {
        if (scale) scale[l] = ex->scale[l];
        if (inv) inv[l] = ex->invScale[l];
        if (s2) s2[l] = ex->sigma2[l];
        if (inv2) inv2[l] = ex->invSigma2[l];
        if (quota) quota[l] = ex->quota[l];
    }
    return ORBX_OK;
}

int orbx_extractor_enable_timing(orbx_extractor* ex, int on) {
    int rc = check_device(ex);
    if (rc) return rc;
    if (on && !ex->ev[0])
        for (int i = 0; i <= ORBX_NUM_STAGES; ++i) ORBX_CUDA(cudaEventCreate(&ex->ev[i]));
    ex->timing = on != 0;
    ex->timed = false;
    return ORBX_OK;
}

int orbx_extractor_stage_ms(orbx_extractor* ex, float* ms) {
    int rc = check_device(ex);
    if (rc) return rc;
    if (!ms || !ex->timed) { set_error("orbx_extractor_stage_ms: the last extract call was not timed"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaEventSynchronize(ex->ev[ORBX_NUM_STAGES]));
    for (int i = 0; i < 3; ++i) ORBX_CUDA(cudaEventElapsedTime(&ms[i], ex->ev[i], ex->ev[i + 1]));
    ms[3] = ms[4] = 0.f;  // blur and describe alternate over sub-batches: their times are summed per kind
    for (int k = 0; k < ex->nSub; ++k) {
        float a = 0.f, b = 0.f;
        ORBX_CUDA(cudaEventElapsedTime(&a, k == 0 ? ex->ev[3] : ex->subEv[2 * k - 1], ex->subEv[2 * k]));
        ORBX_CUDA(cudaEventElapsedTime(&b, ex->subEv[2 * k], ex->subEv[2 * k + 1]));
        ms[3] += a;
        ms[4] += b;
    }
    if (ex->overlapped && ex->timedSide >= 0)  // the blur ran beside FAST + quadtree: its own span on the side stream
        ORBX_CUDA(cudaEventElapsedTime(&ms[3], ex->sides[ex->timedSide].t0, ex->sides[ex->timedSide].t1));
    return ORBX_OK;
}

const char* orbx_stage_name(int stage) {
    static const char* names[ORBX_NUM_STAGES] = {"Pyramid/Resize", "FAST+Grid", "Make quadtree", "Gaussian Blur",
                                                 "Compute angle + ORB descriptor + Compute scale"};
    return stage >= 0 && stage < ORBX_NUM_STAGES ? names[stage] : "";
}

int orbx_extractor_set_pattern(orbx_extractor* ex, int variant) {
    if (!ex || (variant != ORBX_PATTERN_REFERENCE && variant != ORBX_PATTERN_CANONICAL)) {
        set_error("orbx_extractor_set_pattern: variant must be ORBX_PATTERN_REFERENCE or ORBX_PATTERN_CANONICAL");
        return ORBX_E_INVALID;
    }
    if (ex->patternVariant != variant && ex->graphExec) {  // the captured single-frame graph holds the old kernel parameters
        cudaStreamSynchronize(ex->stream);
        cudaGraphExecDestroy(ex->graphExec);
        ex->graphExec = nullptr;
        ex->graphSeen = false;
    }
    ex->patternVariant = variant;
    return ORBX_OK;
}

int orbx_get_umax(const orbx_extractor* ex, int* umax16) {
    if (!ex || !umax16) { set_error("null argument"); return ORBX_E_INVALID; }
    memcpy(umax16, ex->umax, sizeof ex->umax);
    return ORBX_OK;
}

void* orbx_extractor_stream(const orbx_extractor* ex) { return ex ? (void*)ex->stream : nullptr; }

int orbx_level_size(const orbx_extractor* ex, int level, int* w, int* h) {
    if (!ex || level < 0 || level >= ex->prm.nlevels || ex->w == 0) { set_error("orbx_level_size: no geometry yet"); return ORBX_E_INVALID; }
    if (w) *w = ex->g.lv[level].w;
    if (h) *h = ex->g.lv[level].h;
    return ORBX_OK;
}

int orbx_extract_batch_device(orbx_extractor* ex, const uint8_t* dFrames, int batch, int w, int h, size_t stride,
                              size_t frameStride, orbx_keypoint* dKps, uint8_t* dDesc, int cap, int* dCounts, int* dStatus,
                              void* stream) {
    int rc = check_device(ex);
    if (rc) return rc;
    if (!dFrames || batch < 1 || w < 1 || h < 1 || stride < (size_t)w || !dKps || !dDesc || cap < 1 || !dCounts) {
        set_error("orbx_extract_batch_device: bad argument");
        return ORBX_E_INVALID;
    }
    if ((rc = configure(ex, w, h, batch))) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : ex->stream;
    ex->lastFrameBase = 0;
    int* dSt = dStatus ? dStatus : ex->status.as<int>();
    // Large batches run as SUB-BATCHES whose intermediates (planes, blurred levels, cell lists: ~3.8 MB per KITTI frame) stay
    // in the 126 MB L2 between the stages, alternating over two streams so one sub-batch's small tail kernels overlap the
    // next one's wide ones.  Stage timing (orbx_extractor_enable_timing) measures the unsplit sequence.
    static const int subPref = []() { const char* e = getenv("ORBX_SUB_BATCH"); return e ? atoi(e) : 0; }();
    if (subPref <= 0 || batch <= subPref || ex->timing)
        return run_pipeline(ex, dFrames, batch, stride, frameStride, dKps, dDesc, cap, dCounts, dSt, s);
    if (!ex->stream2) ORBX_CUDA(cudaStreamCreateWithFlags(&ex->stream2, cudaStreamNonBlocking));
    while (ex->pipeEv.size() < 2) {
        cudaEvent_t e;
        ORBX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ex->pipeEv.push_back(e);
    }
    ORBX_CUDA(cudaMemsetAsync(dSt, 0, sizeof(int), s));
    ORBX_CUDA(cudaEventRecord(ex->pipeEv[0], s));
    ORBX_CUDA(cudaStreamWaitEvent(ex->stream2, ex->pipeEv[0], 0));
    for (int f0 = 0, k = 0; f0 < batch; f0 += subPref, ++k) {
        const int nb = std::min(subPref, batch - f0);
        if ((rc = run_pipeline(ex, dFrames + (size_t)f0 * frameStride, nb, stride, frameStride, dKps + (size_t)f0 * cap,
                               dDesc + (size_t)f0 * cap * 32, cap, dCounts + f0, dSt, (k & 1) ? ex->stream2 : s, false, f0)))
            return rc;
    }
    ORBX_CUDA(cudaEventRecord(ex->pipeEv[1], ex->stream2));
    ORBX_CUDA(cudaStreamWaitEvent(s, ex->pipeEv[1], 0));
    ex->lastBatch = batch;
    ex->lastScratchBase = 0;
    return ORBX_OK;
}

static bool is_pinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

static int overflow_error(int cap) {
    set_error("orbx_extract: a device buffer overflowed (more than cap=%d keypoints in a frame, or more than 2^20 FAST corners in one level)", cap);
    return ORBX_E_OVERFLOW;
}

// async = true: everything is enqueued (pinned, tightly packed caller buffers only) and the call returns; the caller collects
// it with orbx_extract_batch_wait.  Two calls may be in flight: they use disjoint device staging (slot) and share the
// streams, so call k + 1's H2D copies run while call k's kernels and D2H copies are still going.
static int extract_batch_impl(orbx_extractor* ex, const uint8_t* frames, int batch, int w, int h, size_t stride, size_t frameStride,
                              orbx_keypoint* kps, uint8_t* desc, int cap, int* counts, bool async) {
    int rc = check_device(ex);
    if (rc) return rc;
    if (batch < 1 || !counts || cap < 0) { set_error("orbx_extract_batch: bad argument"); return ORBX_E_INVALID; }
    if (!async && ex->pendCount > 0) {
        set_error("orbx_extract_batch: %d asynchronous call(s) in flight; collect them with orbx_extract_batch_wait first", ex->pendCount);
        return ORBX_E_INVALID;
    }
    if (async && ex->pendCount >= 2) {
        set_error("orbx_extract_batch_async: two calls are in flight already; orbx_extract_batch_wait one first");
        return ORBX_E_INVALID;
    }
    if (!frames || w <= 0 || h <= 0) {  // empty image: the reference returns without touching the outputs (:1542-1543)
        if (async) { set_error("orbx_extract_batch_async: empty frames"); return ORBX_E_INVALID; }
        for (int f = 0; f < batch; ++f) counts[f] = 0;
        return ORBX_OK;
    }
    if (stride < (size_t)w || !kps || !desc || cap < 1) { set_error("orbx_extract_batch: bad argument"); return ORBX_E_INVALID; }
    const int slot = async ? ((ex->pendHead + ex->pendCount) & 1) : 0;
    // The batch is cut into chunks that flow through three streams — H2D | kernels | D2H — so the copies of chunk c+1 and
    // c-1 overlap the kernels of chunk c.  The scratch (pyramids, cell lists ...) holds TWO chunks: even and odd chunks
    // run their kernels on two streams ("lanes") over disjoint halves of it; small batches are a single chunk.
    static const int chunkPref = []() { const char* e = getenv("ORBX_PIPE_CHUNK"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 128; }();
    static const bool taper = []() { const char* e = getenv("ORBX_PIPE_TAPER"); return !e || atoi(e) != 0; }();
    const int nFull = std::max(1, batch / chunkPref), C = (batch + nFull - 1) / nFull;
    // chunk sizes: C each; with >= 4 chunks the first and the last are halved, so the kernels start after a short first
    // copy and the drain after the last copy is short
    std::vector<int> sizes;
    for (int left = batch; left > 0;) {
        const int n = std::min(C, left);
        sizes.push_back(n);
        left -= n;
    }
    if (taper && sizes.size() >= 4 && C >= 16) {
        const int lastN = sizes.back();
        sizes.pop_back();
        sizes.erase(sizes.begin());
        sizes.insert(sizes.begin(), {C / 2, C - C / 2});
        sizes.push_back(lastN - lastN / 2);
        sizes.push_back(lastN / 2);
        if (sizes.back() == 0) sizes.pop_back();
    }
    const int nChunks = (int)sizes.size();
    const int lanes = nChunks > 1 ? 2 : 1;  // kernels of even / odd chunks run on two streams over disjoint scratch halves, so
                                            // the tail of one chunk's kernels overlaps the next chunk's
    if (ex->pendCount > 0 && !(w == ex->w && h == ex->h && lanes * C <= ex->batchCap)) {
        set_error("orbx_extract_batch_async: the geometry or the chunk size changes while a call is in flight; wait for it first");
        return ORBX_E_INVALID;
    }
    if ((rc = configure(ex, w, h, lanes * C))) return rc;
    cudaStream_t s = ex->stream;
    const size_t rowBytes = (size_t)w, frameBytes = rowBytes * h;
    if ((rc = ex->dFrames[slot].ensure(frameBytes * batch + 16))) return rc;
    // device outputs live in ONE allocation [keypoints | descriptors | counts | status] mirrored by one pinned block, so a
    // small call returns everything with a single D2H copy
    const size_t kpBytes = (size_t)round_up((long long)batch * cap * sizeof(orbx_keypoint), 256);
    const size_t dBytes = (size_t)round_up((long long)batch * cap * 32, 256), cBytes = (size_t)round_up((long long)batch * 4, 256);
    const size_t outBytes = kpBytes + dBytes + cBytes + 256;
    if ((rc = ex->dKps[slot].ensure(outBytes))) return rc;
    cudaStream_t sIn = s, sOut = s;
    if (nChunks > 1) {
        if (!ex->h2dStream) ORBX_CUDA(cudaStreamCreateWithFlags(&ex->h2dStream, cudaStreamNonBlocking));
        if (!ex->d2hStream) ORBX_CUDA(cudaStreamCreateWithFlags(&ex->d2hStream, cudaStreamNonBlocking));
        if (!ex->stream2) ORBX_CUDA(cudaStreamCreateWithFlags(&ex->stream2, cudaStreamNonBlocking));
        while ((int)ex->pipeEv.size() < 2 * nChunks + 1) {
            cudaEvent_t e;
            ORBX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            ex->pipeEv.push_back(e);
        }
        sIn = ex->h2dStream;
        sOut = ex->d2hStream;
    }
    // H2D straight from the caller's buffer when it is pinned and tightly packed, else through pinned staging;
    // D2H straight into the caller's buffers when they are pinned
    const bool tight = stride == rowBytes && (frameStride == frameBytes || batch == 1);
    const bool directIn = tight && is_pinned(frames);
    const bool direct = is_pinned(kps) && is_pinned(desc) && is_pinned(counts);
    if (async && !(directIn && direct)) {
        set_error("orbx_extract_batch_async needs pinned (cudaHostAlloc / cudaHostRegister) tightly packed frames and pinned outputs");
        return ORBX_E_INVALID;
    }
    if (!directIn && (rc = ex->hFrames.ensure(frameBytes * batch))) return rc;
    orbx_keypoint* hk = kps;
    uint8_t* hd = desc;
    int* hc = counts;
    int* hst = ex->hStatus.as<int>() + 4 * slot;
    if (!direct) {
        if ((rc = ex->hKps.ensure(outBytes))) return rc;
        uint8_t* hb = ex->hKps.as<uint8_t>();
        hk = reinterpret_cast<orbx_keypoint*>(hb);
        hd = hb + kpBytes;
        hc = reinterpret_cast<int*>(hb + kpBytes + dBytes);
        hst = reinterpret_cast<int*>(hb + kpBytes + dBytes + cBytes);
    }
    uint8_t* dOut = ex->dKps[slot].as<uint8_t>();
    orbx_keypoint* dK = reinterpret_cast<orbx_keypoint*>(dOut);
    uint8_t* dD = dOut + kpBytes;
    int* dC = reinterpret_cast<int*>(dOut + kpBytes + dBytes);
    int* dSt = reinterpret_cast<int*>(dOut + kpBytes + dBytes + cBytes);
    const bool oneCopy = !direct && nChunks == 1;
    uint8_t* dF = ex->dFrames[slot].as<uint8_t>();
    static const bool noGraph = getenv("ORBX_NO_GRAPH") != nullptr || getenv("ORBX_DEBUG_SYNC") != nullptr;
    if (batch == 1 && oneCopy && !directIn && !ex->timing && ex->graphOk && !noGraph && !async) {
        // ---- single frame, pageable buffers (what ORBextractor::operator() does per frame): the whole sequence is captured
        //      into a CUDA graph the second time a geometry is seen and replayed from then on ----
        uint8_t* st = ex->hFrames.as<uint8_t>();
        for (int y = 0; y < h; ++y) memcpy(st + (size_t)y * rowBytes, frames + (size_t)y * stride, rowBytes);
        orbx_extractor::GraphKey key;
        key.w = w; key.h = h; key.cap = cap;
        key.hIn = st; key.hOut = hk; key.dIn = dF; key.dOut = dOut; key.plane0 = ex->g.lv[0].plane;
        bool launched = false;
        if (ex->graphExec && key == ex->graphKey) {
            ORBX_CUDA(cudaGraphLaunch(ex->graphExec, s));
            ex->serial++;
            for (long long i = 0; i < ex->graphKernels; ++i) count_launch();  // the kernels of the replayed graph
            launched = true;
        } else {
            if (ex->graphExec) {
                cudaGraphExecDestroy(ex->graphExec);
                ex->graphExec = nullptr;
            }
            const bool capture = ex->graphSeen && key == ex->graphKey;
            ex->graphKey = key;
            ex->graphSeen = true;
            if (capture && cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
                cudaGetLastError();
                ex->graphOk = false;
            } else {
                rc = ORBX_OK;
                cudaError_t ce = cudaMemsetAsync(dSt, 0, sizeof(int), s);
                if (ce == cudaSuccess) ce = cudaMemcpyAsync(dF, st, frameBytes, cudaMemcpyHostToDevice, s);
                if (ce == cudaSuccess) rc = run_pipeline(ex, dF, 1, rowBytes, frameBytes, dK, dD, cap, dC, dSt, s, false, 0);
                if (capture) ex->graphKernels = ex->lastKernels;  // what one replay of the graph launches
                if (ce == cudaSuccess && !rc) ce = cudaMemcpyAsync(hk, dOut, outBytes, cudaMemcpyDeviceToHost, s);
                if (capture) {
                    cudaGraph_t graph = nullptr;
                    cudaError_t ee = cudaStreamEndCapture(s, &graph);
                    if (ce == cudaSuccess && !rc && ee == cudaSuccess && graph &&
                        cudaGraphInstantiate(&ex->graphExec, graph, 0) == cudaSuccess) {
                        ORBX_CUDA(cudaGraphLaunch(ex->graphExec, s));
                        launched = true;
                    } else {  // capture is not possible here: run the plain sequence from now on
                        cudaGetLastError();
                        ex->graphExec = nullptr;
                        ex->graphOk = false;
                    }
                    if (graph) cudaGraphDestroy(graph);
                } else {
                    if (ce != cudaSuccess) {
                        set_error("orbx_extract: %s", cudaGetErrorString(ce));
                        return ORBX_E_CUDA;
                    }
                    if (rc) return rc;
                    launched = true;
                }
            }
        }
        if (launched) {
            ex->lastFrameBase = ex->lastScratchBase = 0;
            ex->lastBatch = 1;
            ORBX_CUDA(cudaStreamSynchronize(s));
            counts[0] = hc[0];
            ex->lastDevKps = dK;
            ex->lastDevDesc = dD;
            ex->lastDevCount = std::min(hc[0], cap);
            memcpy(kps, hk, (size_t)hc[0] * sizeof(orbx_keypoint));
            memcpy(desc, hd, (size_t)hc[0] * 32);
            if (*hst != 0) return overflow_error(cap);
            return ORBX_OK;
        }
    }
    ORBX_CUDA(cudaMemsetAsync(dSt, 0, sizeof(int), s));
    if (nChunks > 1) {  // the second lane starts after the status word is cleared
        ORBX_CUDA(cudaEventRecord(ex->pipeEv[2 * nChunks], s));
        ORBX_CUDA(cudaStreamWaitEvent(ex->stream2, ex->pipeEv[2 * nChunks], 0));
    }
    for (int c = 0, f0 = 0; c < nChunks; f0 += sizes[c], ++c) {
        const int nb = sizes[c];
        const int lane = c & (lanes - 1);
        cudaStream_t sK = lane ? ex->stream2 : s;
        const uint8_t* src = frames + (size_t)f0 * frameStride;
        if (!directIn) {
            uint8_t* st = ex->hFrames.as<uint8_t>() + (size_t)f0 * frameBytes;
            for (int f = 0; f < nb; ++f)
                for (int y = 0; y < h; ++y)
                    memcpy(st + (size_t)f * frameBytes + (size_t)y * rowBytes, src + (size_t)f * frameStride + (size_t)y * stride, rowBytes);
            src = st;
        }
        ORBX_CUDA(cudaMemcpyAsync(dF + (size_t)f0 * frameBytes, src, frameBytes * nb, cudaMemcpyHostToDevice, sIn));
        if (nChunks > 1) {
            ORBX_CUDA(cudaEventRecord(ex->pipeEv[2 * c], sIn));
            ORBX_CUDA(cudaStreamWaitEvent(sK, ex->pipeEv[2 * c], 0));
        }
        if ((rc = run_pipeline(ex, dF + (size_t)f0 * frameBytes, nb, rowBytes, frameBytes, dK + (size_t)f0 * cap,
                               dD + (size_t)f0 * cap * 32, cap, dC + f0, dSt, sK, false, lane * C)))
            return rc;
        if (nChunks > 1) {
            ORBX_CUDA(cudaEventRecord(ex->pipeEv[2 * c + 1], sK));
            ORBX_CUDA(cudaStreamWaitEvent(sOut, ex->pipeEv[2 * c + 1], 0));
        }
        ex->lastFrameBase = f0;
        if (oneCopy) {
            ORBX_CUDA(cudaMemcpyAsync(hk, dOut, outBytes, cudaMemcpyDeviceToHost, sOut));
            break;
        }
        ORBX_CUDA(cudaMemcpyAsync(hc + f0, dC + f0, (size_t)nb * 4, cudaMemcpyDeviceToHost, sOut));
        ORBX_CUDA(cudaMemcpyAsync(hk + (size_t)f0 * cap, dK + (size_t)f0 * cap, (size_t)nb * cap * sizeof(orbx_keypoint),
                                  cudaMemcpyDeviceToHost, sOut));
        ORBX_CUDA(cudaMemcpyAsync(hd + (size_t)f0 * cap * 32, dD + (size_t)f0 * cap * 32, (size_t)nb * cap * 32,
                                  cudaMemcpyDeviceToHost, sOut));
    }
    if (!oneCopy) ORBX_CUDA(cudaMemcpyAsync(hst, dSt, 4, cudaMemcpyDeviceToHost, sOut));
    if (async) {  // every chunk's D2H waits for its kernels, which wait for its H2D: this event closes the whole call
        orbx_extractor::Pending& pd = ex->pend[slot];
        if (!pd.done) ORBX_CUDA(cudaEventCreateWithFlags(&pd.done, cudaEventDisableTiming));
        ORBX_CUDA(cudaEventRecord(pd.done, sOut));
        pd.hst = hst;
        pd.cap = cap;
        ex->pendCount++;
        return ORBX_OK;
    }
    ORBX_CUDA(cudaStreamSynchronize(sOut));
    if (batch == 1) {
        ex->lastDevKps = dK;
        ex->lastDevDesc = dD;
        ex->lastDevCount = std::min(hc[0], cap);
    }
    if (nChunks > 1) {
        ORBX_CUDA(cudaStreamSynchronize(sIn));
        ORBX_CUDA(cudaStreamSynchronize(ex->stream2));
        ORBX_CUDA(cudaStreamSynchronize(s));
    }
    if (!direct) {
        memcpy(counts, hc, (size_t)batch * 4);
        for (int f = 0; f < batch; ++f) {
            memcpy(kps + (size_t)f * cap, hk + (size_t)f * cap, (size_t)hc[f] * sizeof(orbx_keypoint));
            memcpy(desc + (size_t)f * cap * 32, hd + (size_t)f * cap * 32, (size_t)hc[f] * 32);
        }
    }
    return *hst != 0 ? overflow_error(cap) : ORBX_OK;
}

int orbx_extract_batch(orbx_extractor* ex, const uint8_t* frames, int batch, int w, int h, size_t stride, size_t frameStride,
                       orbx_keypoint* kps, uint8_t* desc, int cap, int* counts) {
    return extract_batch_impl(ex, frames, batch, w, h, stride, frameStride, kps, desc, cap, counts, false);
}

int orbx_extract_batch_async(orbx_extractor* ex, const uint8_t* frames, int batch, int w, int h, size_t stride, size_t frameStride,
                             orbx_keypoint* kps, uint8_t* desc, int cap, int* counts) {
    return extract_batch_impl(ex, frames, batch, w, h, stride, frameStride, kps, desc, cap, counts, true);
}

int orbx_extract_batch_wait(orbx_extractor* ex) {
    int rc = check_device(ex);
    if (rc) return rc;
    if (ex->pendCount == 0) {
        set_error("orbx_extract_batch_wait: no asynchronous call in flight");
        return ORBX_E_INVALID;
    }
    orbx_extractor::Pending& pd = ex->pend[ex->pendHead];
    ex->pendHead ^= 1;
    ex->pendCount--;
    ORBX_CUDA(cudaEventSynchronize(pd.done));
    return *pd.hst != 0 ? overflow_error(pd.cap) : ORBX_OK;
}

int orbx_extract(orbx_extractor* ex, const uint8_t* img, int w, int h, size_t stride, orbx_keypoint* kps, uint8_t* desc, int cap,
                 int* n) {
    if (!n) { set_error("orbx_extract: null count pointer"); return ORBX_E_INVALID; }
    return orbx_extract_batch(ex, img, 1, w, h, stride, (size_t)h * stride, kps, desc, cap, n);
}

}  // extern "C"

// ---- mvImagePyramid access and stage taps -----------------------------------------------------------------------
extern "C" {

int orbx_pyramid_level_host(orbx_extractor* ex, int frame, int level, const uint8_t** interior, int* w, int* h, size_t* pitch) {
    int rc = check_device(ex);
    if (rc) return rc;
    if (ex->serial == 0 || level < 0 || level >= ex->prm.nlevels || (frame -= ex->lastFrameBase) < 0 || frame >= ex->lastBatch) {
        set_error("orbx_pyramid_level_host: no such frame/level (frame %d level %d, last batch %d)", frame, level, ex->lastBatch);
        return ORBX_E_INVALID;
    }
    frame += ex->lastScratchBase;  // index into the scratch
    const LevelGeom& L = ex->g.lv[level];
    if (ex->hostLevelSerial[level] != ex->serial || ex->hostLevelFrame[level] != frame) {  // lazy D2H
        ex->hostLevel[level].resize((size_t)L.planeStride);
        ORBX_CUDA(cudaMemcpyAsync(ex->hostLevel[level].data(), L.plane + (size_t)frame * L.planeStride, (size_t)L.pitch * L.rows,
                                  cudaMemcpyDeviceToHost, ex->stream));
        ORBX_CUDA(cudaStreamSynchronize(ex->stream));
        ex->hostLevelSerial[level] = ex->serial;
        ex->hostLevelFrame[level] = frame;
    }
    if (interior) *interior = ex->hostLevel[level].data() + (size_t)EDGE * L.pitch + XOFF;
    if (w) *w = L.w;
    if (h) *h = L.h;
    if (pitch) *pitch = (size_t)L.pitch;
    return ORBX_OK;
}

int orbx_debug_blurred_level(orbx_extractor* ex, int frame, int level, uint8_t* out) {
    int rc = check_device(ex);
    if (rc) return rc;
    if (ex->serial == 0 || level < 0 || level >= ex->prm.nlevels || (frame -= ex->lastFrameBase) < 0 || frame >= ex->lastBatch || !out) {
        set_error("orbx_debug_blurred_level: bad argument");
        return ORBX_E_INVALID;
    }
    frame += ex->lastScratchBase;  // index into the scratch
    const LevelGeom& L = ex->g.lv[level];
    ORBX_CUDA(cudaMemcpy2DAsync(out, (size_t)L.w, L.blur + (size_t)frame * L.blurStride, (size_t)L.bpitch, (size_t)L.w, (size_t)L.h,
                                cudaMemcpyDeviceToHost, ex->stream));
    ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    return ORBX_OK;
}

int orbx_debug_candidates(orbx_extractor* ex, int frame, int level, orbx_keypoint* out, int cap, int* n) {
    int rc = check_device(ex);
    if (rc) return rc;
    if (ex->serial == 0 || level < 0 || level >= ex->prm.nlevels || (frame -= ex->lastFrameBase) < 0 || frame >= ex->lastBatch || !n) {
        set_error("orbx_debug_candidates: bad argument");
        return ORBX_E_INVALID;
    }
    frame += ex->lastScratchBase;  // index into the scratch
    const FrameGeom& g = ex->g;
    const LevelGeom& L = g.lv[level];
    const int nCells = L.nCols * L.nRows;
    std::vector<int> cnt(std::max(nCells, 1));
    std::vector<uint32_t> keys((size_t)std::max(nCells, 1) * L.cellCap);
    ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    if (nCells > 0) {
        ORBX_CUDA(cudaMemcpy(cnt.data(), ex->cellCnt.as<int>() + (size_t)frame * g.nCellsTotal + L.cellBase, (size_t)nCells * 4,
                             cudaMemcpyDeviceToHost));
        ORBX_CUDA(cudaMemcpy(keys.data(), ex->cellKeys.as<uint32_t>() + (size_t)frame * (size_t)g.keysPerFrame + (size_t)L.keyBase,
                             (size_t)nCells * L.cellCap * 4, cudaMemcpyDeviceToHost));
    }
    int m = 0;
    for (int c = 0; c < nCells; ++c)
        for (int k = 0; k < std::min(cnt[c], L.cellCap); ++k, ++m)
            if (out && m < cap) {
                const uint32_t key = keys[(size_t)c * L.cellCap + k];
                out[m] = orbx_keypoint{(float)key_x(key), (float)key_y(key), 7.f, -1.f, (float)key_resp(key), 0, -1};
            }
    *n = m;
    return ORBX_OK;
}

int orbx_debug_level_keypoints(orbx_extractor* ex, int frame, int level, orbx_keypoint* out, int cap, int* n) {
    int rc = check_device(ex);
    if (rc) return rc;
    if (ex->serial == 0 || level < 0 || level >= ex->prm.nlevels || (frame -= ex->lastFrameBase) < 0 || frame >= ex->lastBatch || !n) {
        set_error("orbx_debug_level_keypoints: bad argument");
        return ORBX_E_INVALID;
    }
    frame += ex->lastScratchBase;  // index into the scratch
    const FrameGeom& g = ex->g;
    const LevelGeom& L = g.lv[level];
    int cnt = 0;
    std::vector<uint32_t> keys(L.quota + 4);
    std::vector<float> ang(L.quota + 4);
    ORBX_CUDA(cudaStreamSynchronize(ex->stream));
    ORBX_CUDA(cudaMemcpy(&cnt, ex->lvlCnt.as<int>() + (size_t)frame * g.nlevels + level, 4, cudaMemcpyDeviceToHost));
    ORBX_CUDA(cudaMemcpy(keys.data(), ex->lvlKeys.as<uint32_t>() + (size_t)frame * g.slotsTotal + L.slotBase, keys.size() * 4,
                         cudaMemcpyDeviceToHost));
    ORBX_CUDA(cudaMemcpy(ang.data(), ex->lvlAngles.as<float>() + (size_t)frame * g.slotsTotal + L.slotBase, ang.size() * 4,
                         cudaMemcpyDeviceToHost));
    for (int i = 0; i < cnt && out && i < cap; ++i)
        out[i] = orbx_keypoint{(float)key_x(keys[i]), (float)key_y(keys[i]), L.size, ang[i], (float)key_resp(keys[i]), level, -1};
    *n = cnt;
    return ORBX_OK;
}

int orbx_debug_distribute(const orbx_keypoint* in, int n, int minX, int maxX, int minY, int maxY, int N, orbx_keypoint* out,
                          int cap, int* nout) {
    if (!nout || n < 0 || N < 1 || N > 8000 || maxX <= minX || maxY <= minY || maxX - minX > 4095 || maxY - minY > 4095) {
        set_error("orbx_debug_distribute: bad argument");
        return ORBX_E_INVALID;
    }
    std::vector<uint32_t> keys(std::max(n, 1));
    for (int i = 0; i < n; ++i) keys[i] = pack_key((int)in[i].x, (int)in[i].y, (int)in[i].response);
    DevBuf dIn, dScratch, dOut, dCnt;
    const int candCap = std::max(n, 32);
    int rc;
    if ((rc = dIn.ensure((size_t)candCap * 4)) || (rc = dScratch.ensure((size_t)candCap * 8)) || (rc = dOut.ensure((size_t)(N + 4) * 4)) ||
        (rc = dCnt.ensure(8)))
        return rc;
    ORBX_CUDA(cudaMemcpy(dIn.p, keys.data(), (size_t)n * 4, cudaMemcpyHostToDevice));
    rc = launch_quadtree_single(dIn.as<uint32_t>(), n, minX, maxX, minY, maxY, N, dScratch.as<uint32_t>(), candCap, dOut.as<uint32_t>(),
                                dCnt.as<int>(), dCnt.as<int>() + 1, 0);
    if (!rc) {
        int cnt = 0;
        std::vector<uint32_t> res(N + 4);
        cudaError_t e = cudaMemcpy(&cnt, dCnt.p, 4, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(res.data(), dOut.p, res.size() * 4, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) {
            set_error("orbx_debug_distribute: %s", cudaGetErrorString(e));
            rc = ORBX_E_CUDA;
        } else {
            for (int i = 0; i < cnt && out && i < cap; ++i)
                out[i] = orbx_keypoint{(float)key_x(res[i]), (float)key_y(res[i]), 7.f, -1.f, (float)key_resp(res[i]), 0, -1};
            *nout = cnt;
        }
    }
    dIn.release(); dScratch.release(); dOut.release(); dCnt.release();
    return rc;
}

}  // extern "C"
