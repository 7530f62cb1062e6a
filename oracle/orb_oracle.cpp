// oracle/orb_oracle.cpp — TEST INFRASTRUCTURE ONLY.  See orb_oracle.hpp for scope and pinning notes.
// Build: g++ -std=c++17 -O2 -ffp-contract=off -fno-fast-math (x86-64 baseline, no -march=native), so that
// float expressions evaluate as separately rounded mul/add like the reference's -O3 x86 build would.
#include "orb_oracle.hpp"

#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstring>
#include <list>
#include <utility>

namespace orc {

// ================================================================================================
// OpenCV primitive models (the arithmetic lives in OpenCV, which is not under /root/reference; call
// sites: src/ORBextractor.cc:1852 resize, :1854-1860 copyMakeBorder, :1278-1284 FAST, :1747 GaussianBlur,
// :190 fastAtan2, cvRound at :168,202,206-207,528,546,1844).
// ================================================================================================

// cvRound: SSE cvtss2si / cvtsd2si under the default rounding mode = round half to even.
int cv_round(float v) { return (int)lrintf(v); }
int cv_round(double v) { return (int)lrint(v); }
static inline int cv_floor(float v) { return (int)std::floor(v); }
static inline int cv_floor(double v) { return (int)std::floor(v); }
static inline int cv_ceil(float v) { return (int)std::ceil(v); }
static inline short sat_short(float v) {
    int i = cv_round(v);
    return (short)std::min(std::max(i, (int)SHRT_MIN), (int)SHRT_MAX);
}

// cv::resize(..., INTER_LINEAR) for CV_8UC1: 11-bit fixed-point separable bilinear (INTER_RESIZE_COEF_SCALE=2048),
// horizontal pass to int, vertical pass ((b0*(r0>>4))>>16) + ((b1*(r1>>4))>>16) + 2) >> 2.  An exact 2x2
// decimation is routed by OpenCV to the INTER_AREA fast path ((a+b+c+d+2)>>2).
void resize_linear_8u(const uint8_t* src, int sw, int sh, size_t sstep, uint8_t* dst, int dw, int dh, size_t dstep) {
    const double inv_x = (double)dw / sw, inv_y = (double)dh / sh;
    const double scale_x = 1. / inv_x, scale_y = 1. / inv_y;
    const int iscale_x = cv_round(scale_x) , iscale_y = cv_round(scale_y);
    const bool area_fast = std::abs(scale_x - iscale_x) < DBL_EPSILON && std::abs(scale_y - iscale_y) < DBL_EPSILON;
    if (area_fast && iscale_x == 2 && iscale_y == 2) {
        for (int dy = 0; dy < dh; ++dy) {
            const uint8_t* s0 = src + (size_t)(2 * dy) * sstep;
            const uint8_t* s1 = s0 + sstep;
            uint8_t* d = dst + (size_t)dy * dstep;
            for (int dx = 0; dx < dw; ++dx) d[dx] = (uint8_t)((s0[2 * dx] + s0[2 * dx + 1] + s1[2 * dx] + s1[2 * dx + 1] + 2) >> 2);
        }
        return;
    }
    std::vector<int> xofs(dw);
    std::vector<short> alpha(2 * (size_t)dw);
    for (int dx = 0; dx < dw; ++dx) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cv_floor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        alpha[2 * dx] = sat_short((1.f - fx) * 2048);
        alpha[2 * dx + 1] = sat_short(fx * 2048);
    }
    std::vector<int> r0(dw), r1(dw);
    for (int dy = 0; dy < dh; ++dy) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cv_floor(fy);
        fy -= sy;
        const int b0 = sat_short((1.f - fy) * 2048), b1 = sat_short(fy * 2048);
        const int y0 = std::min(std::max(sy, 0), sh - 1), y1 = std::min(std::max(sy + 1, 0), sh - 1);
        const uint8_t* s0 = src + (size_t)y0 * sstep;
        const uint8_t* s1 = src + (size_t)y1 * sstep;
        for (int dx = 0; dx < dw; ++dx) {
            const int sx = xofs[dx], sx1 = std::min(sx + 1, sw - 1);
            const int a0 = alpha[2 * dx], a1 = alpha[2 * dx + 1];
            r0[dx] = s0[sx] * a0 + s0[sx1] * a1;
            r1[dx] = s1[sx] * a0 + s1[sx1] * a1;
        }
        uint8_t* d = dst + (size_t)dy * dstep;
        for (int dx = 0; dx < dw; ++dx) {
            int v = (((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2;
            d[dx] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

// cv::borderInterpolate(p, n, BORDER_REFLECT_101)
int reflect101(int p, int n) {
    if ((unsigned)p < (unsigned)n) return p;
    if (n == 1) return 0;
    do {
        if (p < 0) p = -p;
        else p = 2 * n - 2 - p;
    } while ((unsigned)p >= (unsigned)n);
    return p;
}

// cv::copyMakeBorder(src, dst, pad, pad, pad, pad, BORDER_REFLECT_101 | BORDER_ISOLATED); dst is (w+2pad) x (h+2pad).
// src may alias the centre of dst (as at src/ORBextractor.cc:1854, where the ROI is a view of temp).
void border_reflect101(const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep, int pad) {
    for (int y = 0; y < h; ++y) {
        const uint8_t* s = src + (size_t)y * sstep;
        uint8_t* d = dst + (size_t)(y + pad) * dstep;
        if (d + pad != s) std::memmove(d + pad, s, (size_t)w);
        for (int x = 0; x < pad; ++x) {
            d[x] = s[reflect101(x - pad, w)];
            d[pad + w + x] = s[reflect101(w + x, w)];
        }
    }
    for (int y = 0; y < pad; ++y) {
        std::memcpy(dst + (size_t)y * dstep, dst + (size_t)(reflect101(y - pad, h) + pad) * dstep, (size_t)w + 2 * pad);
        std::memcpy(dst + (size_t)(pad + h + y) * dstep, dst + (size_t)(reflect101(h + y, h) + pad) * dstep, (size_t)w + 2 * pad);
    }
}

// cv::FAST(img, kps, threshold, nonmaxSuppression, TYPE_9_16).  Corner: >= 9 contiguous ring pixels all
// brighter than v+t or all darker than v-t.  Score (cornerScore<16>): the largest t' for which the pixel is
// still a corner = max over the 16 arcs of 9 and both polarities of min signed difference, minus 1.  With
// NMS a corner is kept iff its score is strictly greater than the scores of all 8 neighbours, non-corners
// (and the 3-px rim) counting as 0.  Output is row-major.  KeyPoint(x, y, size 7, angle -1, response score).
static const int kRing[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},   {3, 0},  {3, -1}, {2, -2}, {1, -3},
                                 {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

static int fast_score_at(const uint8_t* p, const int* off, int threshold) {
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 25; ++k) d[k] = v - p[off[k & 15]];
    // is it a corner at `threshold`, and what is the best arc-min?
    int best = INT_MIN;
    for (int s = 0; s < 16; ++s) {
        int mn = INT_MAX, mx = INT_MIN;
        for (int k = 0; k < 9; ++k) {
            mn = std::min(mn, d[s + k]);
            mx = std::max(mx, d[s + k]);
        }
        best = std::max(best, std::max(mn, -mx));
    }
    if (best <= threshold) return -1;  // not a corner
    return best - 1;
}

void fast9_16(const uint8_t* img, int w, int h, size_t step, int threshold, bool nms, std::vector<KeyPoint>& out) {
    out.clear();
    threshold = std::min(std::max(threshold, 0), 255);
    if (w < 7 || h < 7) return;
    int off[16];
    for (int k = 0; k < 16; ++k) off[k] = kRing[k][0] + kRing[k][1] * (int)step;
    std::vector<int> score((size_t)w * h, 0);
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            const uint8_t* p = img + (size_t)y * step + x;
            // quick rejection (same as OpenCV's first tests): an arc of 9 contains one pixel of every antipodal pair
            const int v = p[0], hi = v + threshold, lo = v - threshold;
            const int r0 = p[off[0]], r8 = p[off[8]], r4 = p[off[4]], r12 = p[off[12]];
            const bool br = (r0 > hi || r8 > hi) && (r4 > hi || r12 > hi);
            const bool dk = (r0 < lo || r8 < lo) && (r4 < lo || r12 < lo);
            int s = (br || dk) ? fast_score_at(p, off, threshold) : -1;
            score[(size_t)y * w + x] = s < 0 ? 0 : (nms ? s : 1);
            if (!nms && s >= 0) out.push_back({(float)x, (float)y, 7.f, -1.f, 0.f, 0, -1});
        }
    if (!nms) return;
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            const int s = score[(size_t)y * w + x];
            if (s == 0) {
                // A corner whose score is exactly 0 can only occur for threshold 0; OpenCV then still requires
                // score > neighbours, which fails against the zero rim, so it is dropped.
                continue;
            }
            const int* c = &score[(size_t)y * w + x];
            if (s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] && s > c[w - 1] && s > c[w] && s > c[w + 1])
                out.push_back({(float)x, (float)y, 7.f, -1.f, (float)s, 0, -1});
        }
}

// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8U: OpenCV's bit-exact 8.8 fixed-point
// path.  getGaussianKernelBitExact(7, 2.0) quantises to [18,34,48,56,48,34,18]/256; rows then columns;
// result = (sum + 32768) >> 16.
void gaussian7x7_s2(const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep) {
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<uint16_t> tmp((size_t)w * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* s = src + (size_t)y * sstep;
        for (int x = 0; x < w; ++x) {
            int acc = 0;
            if (x >= 3 && x < w - 3)
                for (int k = 0; k < 7; ++k) acc += K[k] * s[x + k - 3];
            else
                for (int k = 0; k < 7; ++k) acc += K[k] * s[reflect101(x + k - 3, w)];
            tmp[(size_t)y * w + x] = (uint16_t)acc;
        }
    }
    for (int y = 0; y < h; ++y) {
        uint8_t* d = dst + (size_t)y * dstep;
        const uint16_t* r[7];
        for (int k = 0; k < 7; ++k) r[k] = tmp.data() + (size_t)reflect101(y + k - 3, h) * w;
        for (int x = 0; x < w; ++x) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; ++k) acc += (uint32_t)K[k] * r[k][x];
            d[x] = (uint8_t)((acc + 32768u) >> 16);
        }
    }
}

// cv::fastAtan2 (degrees, [0,360)): 7th-order odd polynomial on min/max, float32, separately rounded mul/add.
float fast_atan2(float y, float x) {
    static const float p1 = 0.9997878412794807f * (float)(180 / M_PI);
    static const float p3 = -0.3258083974640975f * (float)(180 / M_PI);
    static const float p5 = 0.1555786518463281f * (float)(180 / M_PI);
    static const float p7 = -0.04432655554792128f * (float)(180 / M_PI);
    const float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// ================================================================================================
// ORBextractor — src/ORBextractor.cc (CPU branches only)
// ================================================================================================
// bit_pattern_31_ (:236-494).  The table below is the CANONICAL ORB pattern (OpenCV, upstream ORB-SLAM2).  The reference's
// working copy has the token VX_FAILURE between two rows (:261): `... -6,-2/*...*/,VX_FAILURE <newline> -2,11, ...` is the
// expression VX_FAILURE - 2, and OpenVX defines VX_FAILURE = -1, so the reference AS IT COMPILES holds -3 where the canonical
// table holds -2 (element 96: x of the first point of pair 24).  variant 0 = the reference as compiled (the default: parity is
// with the reference's own CPU path), 1 = canonical.  tests/test_oracle_golden.py compiles the reference's lines to show it.
static const int kPatternCanonical[256 * 4] = {
#include "../orb_slam_cuda_b200/csrc/orb_pattern_31.inc"
};
static int g_patternVariant = 0;
static const int* pattern_table() {
    static int ref[256 * 4];
    static bool init = false;
    if (!init) {
        for (int i = 0; i < 1024; ++i) ref[i] = kPatternCanonical[i];
        ref[96] = -1 /* VX_FAILURE */ - 2;
        init = true;
    }
    return g_patternVariant == 0 ? ref : kPatternCanonical;
}
void set_pattern_variant(int v) { g_patternVariant = v ? 1 : 0; }
int pattern_variant() { return g_patternVariant; }
const int* pattern_values() { return pattern_table(); }

// Constructor tables — src/ORBextractor.cc:496-557 (buildGraph at :559 is the OpenVX path, not restated).
Extractor::Extractor(int nf, float sf, int nl, int ini, int mn)
    : nfeatures(nf), nlevels(nl), iniThFAST(ini), minThFAST(mn), scaleFactor(sf) {
    mvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvScaleFactor[0] = 1.0f;
    mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        mvScaleFactor[i] = mvScaleFactor[i - 1] * scaleFactor;  // float * double -> double -> float (:508)
        mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
    }
    mvInvScaleFactor.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
        mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
        mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
    }
    mnFeaturesPerLevel.resize(nlevels);
    float factor = 1.0f / scaleFactor;  // double division, rounded to float (:522)
    float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int level = 0; level < nlevels - 1; level++) {
        mnFeaturesPerLevel[level] = cv_round(nDesired);
        sum += mnFeaturesPerLevel[level];
        nDesired *= factor;
    }
    mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sum, 0);

    // umax: end of each row of the circular patch (:540-555)
    umax.resize(HALF_PATCH_SIZE + 1);
    int v, v0, vmax = cv_floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
    int vmin = cv_ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
}

// ComputePyramid — src/ORBextractor.cc:1837-1862 (CPU branch): level l is resized from level l-1's ROI,
// then the whole padded plane is rebuilt by reflect-101.
void Extractor::compute_pyramid(const uint8_t* img, int w, int h, size_t step) {
    padded.assign(nlevels, Image());
    lw.assign(nlevels, 0);
    lh.assign(nlevels, 0);
    for (int level = 0; level < nlevels; ++level) {
        const float scale = mvInvScaleFactor[level];
        const int sw = cv_round((float)w * scale), sh = cv_round((float)h * scale);
        lw[level] = sw;
        lh[level] = sh;
        padded[level] = Image(sw + 2 * EDGE_THRESHOLD, sh + 2 * EDGE_THRESHOLD);
        Image& P = padded[level];
        uint8_t* roi = P.row(EDGE_THRESHOLD) + EDGE_THRESHOLD;
        if (level != 0) {
            resize_linear_8u(level_ptr(level - 1), lw[level - 1], lh[level - 1], level_step(level - 1), roi, sw, sh, P.w);
            border_reflect101(roi, sw, sh, P.w, P.px.data(), P.w, EDGE_THRESHOLD);
        } else {
            border_reflect101(img, sw, sh, step, P.px.data(), P.w, EDGE_THRESHOLD);
        }
    }
}

namespace {
struct Node {  // ExtractorNode, include/ORBextractor.h:62-73
    int ULx, ULy, URx, URy, BLx, BLy, BRx, BRy;
    std::vector<KeyPoint> keys;
    std::list<Node>::iterator lit;
    bool noMore = false;
    long seq = 0;  // creation order: stands in for the heap address the reference's sort() ties on (:1041)
};

// ExtractorNode::DivideNode — src/ORBextractor.cc:831-887
void divide(const Node& p, Node& n1, Node& n2, Node& n3, Node& n4) {
    const int halfX = (int)std::ceil(static_cast<float>(p.URx - p.ULx) / 2);
    const int halfY = (int)std::ceil(static_cast<float>(p.BRy - p.ULy) / 2);
    n1.ULx = p.ULx; n1.ULy = p.ULy;
    n1.URx = p.ULx + halfX; n1.URy = p.ULy;
    n1.BLx = p.ULx; n1.BLy = p.ULy + halfY;
    n1.BRx = p.ULx + halfX; n1.BRy = p.ULy + halfY;
    n2.ULx = n1.URx; n2.ULy = n1.URy;
    n2.URx = p.URx; n2.URy = p.URy;
    n2.BLx = n1.BRx; n2.BLy = n1.BRy;
    n2.BRx = p.URx; n2.BRy = p.ULy + halfY;
    n3.ULx = n1.BLx; n3.ULy = n1.BLy;
    n3.URx = n1.BRx; n3.URy = n1.BRy;
    n3.BLx = p.BLx; n3.BLy = p.BLy;
    n3.BRx = n1.BRx; n3.BRy = p.BLy;
    n4.ULx = n3.URx; n4.ULy = n3.URy;
    n4.URx = n2.BRx; n4.URy = n2.BRy;
    n4.BLx = n3.BRx; n4.BLy = n3.BRy;
    n4.BRx = p.BRx; n4.BRy = p.BRy;
    for (const KeyPoint& kp : p.keys) {
        if (kp.x < n1.URx) {
            if (kp.y < n1.BRy) n1.keys.push_back(kp);
            else n3.keys.push_back(kp);
        } else if (kp.y < n1.BRy) n2.keys.push_back(kp);
        else n4.keys.push_back(kp);
    }
    for (Node* c : {&n1, &n2, &n3, &n4})
        if (c->keys.size() == 1) c->noMore = true;
}
}  // namespace

// ORBextractor::DistributeOctTree — src/ORBextractor.cc:889-1120, including the fork-specific early exit at
// :959-963.  The reference's final-phase sort (:1041) ties on heap pointer values, which is allocator-dependent;
// here ties on key count are broken by creation order (later-created node = larger "address").
std::vector<KeyPoint> distribute_octtree(const std::vector<KeyPoint>& keys, int minX, int maxX, int minY, int maxY, int N) {
    const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));
    std::vector<KeyPoint> result;
    if (nIni < 1) return result;  // the reference indexes an empty vector here (undefined); we return nothing
    const float hX = static_cast<float>(maxX - minX) / nIni;
    std::list<Node> L;
    std::vector<Node*> roots(nIni);
    long seq = 0;
    for (int i = 0; i < nIni; i++) {
        Node ni;
        ni.ULx = (int)(hX * static_cast<float>(i)); ni.ULy = 0;
        ni.URx = (int)(hX * static_cast<float>(i + 1)); ni.URy = 0;
        ni.BLx = ni.ULx; ni.BLy = maxY - minY;
        ni.BRx = ni.URx; ni.BRy = maxY - minY;
        ni.seq = seq++;
        L.push_back(ni);
        roots[i] = &L.back();
    }
    for (const KeyPoint& kp : keys) {
        size_t r = (size_t)(kp.x / hX);
        if (r >= roots.size()) r = roots.size() - 1;  // unreachable for in-range keys; guards the oracle only
        roots[r]->keys.push_back(kp);
    }
    for (auto it = L.begin(); it != L.end();) {
        if (it->keys.size() == 1) { it->noMore = true; ++it; }
        else if (it->keys.empty()) it = L.erase(it);
        else ++it;
    }
    bool finish = false;
    std::vector<std::pair<int, Node*>> expandable;
    auto add_children = [&](Node* kids[4], int* nToExpand) {
        for (int k = 0; k < 4; ++k) {
            Node& c = *kids[k];
            if (c.keys.empty()) continue;
            c.seq = seq++;
            L.push_front(c);
            if (c.keys.size() > 1) {
                if (nToExpand) ++*nToExpand;
                expandable.push_back(std::make_pair((int)c.keys.size(), &L.front()));
                L.front().lit = L.begin();
            }
        }
    };
    auto by_size_then_creation = [](const std::pair<int, Node*>& a, const std::pair<int, Node*>& b) {
        if (a.first != b.first) return a.first < b.first;
        return a.second->seq < b.second->seq;
    };
    while (!finish) {
        const int prevSize = (int)L.size();
        auto it = L.begin();
        int nToExpand = 0;
        expandable.clear();
        while (it != L.end()) {
            if ((int)L.size() >= N) { finish = true; break; }  // :959-963
            if (it->noMore) { ++it; continue; }
            Node n1, n2, n3, n4;
            divide(*it, n1, n2, n3, n4);
            Node* kids[4] = {&n1, &n2, &n3, &n4};
            add_children(kids, &nToExpand);
            it = L.erase(it);
        }
        if ((int)L.size() >= N || (int)L.size() == prevSize) {
            finish = true;
        } else if (((int)L.size() + nToExpand * 3) > N) {
            while (!finish) {
                const int prev = (int)L.size();
                std::vector<std::pair<int, Node*>> prevExp = expandable;
                expandable.clear();
                std::sort(prevExp.begin(), prevExp.end(), by_size_then_creation);
                for (int j = (int)prevExp.size() - 1; j >= 0; j--) {
                    Node n1, n2, n3, n4;
                    divide(*prevExp[j].second, n1, n2, n3, n4);
                    Node* kids[4] = {&n1, &n2, &n3, &n4};
                    add_children(kids, nullptr);
                    L.erase(prevExp[j].second->lit);
                    if ((int)L.size() >= N) break;
                }
                if ((int)L.size() >= N || (int)L.size() == prev) finish = true;
            }
        }
    }
    result.reserve(L.size());
    for (const Node& n : L) {
        const KeyPoint* best = &n.keys[0];
        float maxResponse = best->response;
        for (size_t k = 1; k < n.keys.size(); k++)
            if (n.keys[k].response > maxResponse) { best = &n.keys[k]; maxResponse = n.keys[k].response; }
        result.push_back(*best);
    }
    return result;
}

// IC_Angle — src/ORBextractor.cc:164-191
float ic_angle(const uint8_t* center, size_t stepz, const std::vector<int>& umax) {
    int m_01 = 0, m_10 = 0;
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    const int step = (int)stepz;
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
        int v_sum = 0;
        const int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            int val_plus = center[u + v * step], val_minus = center[u - v * step];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return fast_atan2((float)m_01, (float)m_10);
}

// computeOrbDescriptor — src/ORBextractor.cc:194-233.  cos/sin resolve to the float overloads (glibc cosf/sinf).
void orb_descriptor(float angle_deg, const uint8_t* center, size_t stepz, uint8_t* desc) {
    static const float factorPI = (float)(M_PI / 180.f);
    const float angle = angle_deg * factorPI;
    const float a = cosf(angle), b = sinf(angle);
    const int step = (int)stepz;
    const int* pat = pattern_table();
    auto tap = [&](int idx) -> int {
        const int px = pat[2 * idx], py = pat[2 * idx + 1];
        return center[cv_round(px * b + py * a) * step + cv_round(px * a - py * b)];
    };
    for (int i = 0; i < 32; ++i, pat += 32) {
        int val = 0;
        for (int k = 0; k < 8; ++k) {
            const int t0 = tap(2 * k), t1 = tap(2 * k + 1);
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

// ComputeKeyPointsOctTree — src/ORBextractor.cc:1122-1149 and the CPU branch :1256-1340.
void Extractor::compute_keypoints() {
    candidates.assign(nlevels, {});
    levelKeys.assign(nlevels, {});
    const float W = 30;
    std::vector<KeyPoint> cell;
    for (int level = 0; level < nlevels; ++level) {
        const int minBorderX = EDGE_THRESHOLD - 3, minBorderY = minBorderX;
        const int maxBorderX = lw[level] - EDGE_THRESHOLD + 3, maxBorderY = lh[level] - EDGE_THRESHOLD + 3;
        std::vector<KeyPoint>& toDistribute = candidates[level];
        const float width = (maxBorderX - minBorderX), height = (maxBorderY - minBorderY);
        const int nCols = width / W, nRows = height / W;
        if (nCols < 1 || nRows < 1) continue;  // the reference divides by zero here; such levels yield nothing
        const int wCell = std::ceil(width / nCols), hCell = std::ceil(height / nRows);
        const uint8_t* base = level_ptr(level);
        const size_t step = level_step(level);
        for (int i = 0; i < nRows; i++) {
            const float iniY = minBorderY + i * hCell;
            float maxY = iniY + hCell + 6;
            if (iniY >= maxBorderY - 3) continue;
            if (maxY > maxBorderY) maxY = maxBorderY;
            for (int j = 0; j < nCols; j++) {
                const float iniX = minBorderX + j * wCell;
                float maxX = iniX + wCell + 6;
                if (iniX >= maxBorderX - 6) continue;
                if (maxX > maxBorderX) maxX = maxBorderX;
                const int x0 = (int)iniX, x1 = (int)maxX, y0 = (int)iniY, y1 = (int)maxY;
                const uint8_t* sub = base + (size_t)y0 * step + x0;
                fast9_16(sub, x1 - x0, y1 - y0, step, iniThFAST, true, cell);
                if (cell.empty()) fast9_16(sub, x1 - x0, y1 - y0, step, minThFAST, true, cell);
                for (KeyPoint& kp : cell) {
                    kp.x += j * wCell;
                    kp.y += i * hCell;
                    toDistribute.push_back(kp);
                }
            }
        }
        std::vector<KeyPoint>& kps = levelKeys[level];
        kps = distribute_octtree(toDistribute, minBorderX, maxBorderX, minBorderY, maxBorderY, mnFeaturesPerLevel[level]);
        const int scaledPatchSize = PATCH_SIZE * mvScaleFactor[level];
        for (KeyPoint& kp : kps) {
            kp.x += minBorderX;
            kp.y += minBorderY;
            kp.octave = level;
            kp.size = scaledPatchSize;
        }
    }
    for (int level = 0; level < nlevels; ++level)  // computeOrientation :822-829
        for (KeyPoint& kp : levelKeys[level])
            kp.angle = ic_angle(level_ptr(level) + (size_t)cv_round(kp.y) * level_step(level) + cv_round(kp.x),
                                level_step(level), umax);
}

// operator() — src/ORBextractor.cc:1538-1548 and the CPU branch :1701-1809.
void Extractor::extract(const uint8_t* img, int w, int h, size_t step, std::vector<KeyPoint>& out, std::vector<uint8_t>& desc) {
    out.clear();
    desc.clear();
    if (!img || w <= 0 || h <= 0) return;
    compute_pyramid(img, w, h, step);
    compute_keypoints();
    int total = 0;
    for (int l = 0; l < nlevels; ++l) total += (int)levelKeys[l].size();
    desc.assign((size_t)total * 32, 0);
    out.reserve(total);
    blurred.assign(nlevels, Image());
    int offset = 0;
    for (int level = 0; level < nlevels; ++level) {
        std::vector<KeyPoint> kps = levelKeys[level];
        const int n = (int)kps.size();
        if (n == 0) continue;
        // workingMat = mvImagePyramid[level].clone(); GaussianBlur(7x7, 2, 2, REFLECT_101)  (:1746-1747)
        blurred[level] = Image(lw[level], lh[level]);
        Image& B = blurred[level];
        gaussian7x7_s2(level_ptr(level), lw[level], lh[level], level_step(level), B.px.data(), B.w);
        for (int i = 0; i < n; ++i)
            orb_descriptor(kps[i].angle, B.row(cv_round(kps[i].y)) + cv_round(kps[i].x), B.w, &desc[(size_t)(offset + i) * 32]);
        offset += n;
        if (level != 0) {
            const float scale = mvScaleFactor[level];
            for (KeyPoint& kp : kps) { kp.x *= scale; kp.y *= scale; }
        }
        out.insert(out.end(), kps.begin(), kps.end());
    }
}

// ================================================================================================
// ORBmatcher / Frame
// ================================================================================================

// ORBmatcher::DescriptorDistance — src/ORBmatcher.cc:1647-1663 (SWAR popcount over 8 x 32 bits)
int descriptor_distance(const uint8_t* a, const uint8_t* b) {
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t x, y;
        std::memcpy(&x, a + 4 * i, 4);
        std::memcpy(&y, b + 4 * i, 4);
        uint32_t v = x ^ y;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

// Frame grid — src/Frame.cc:229-244 (AssignFeaturesToGrid), :381-391 (PosInGrid), include/Frame.h:37-38.
GridFrame::GridFrame(const KeyPoint* k, int n_, float mnX, float mxX, float mnY, float mxY)
    : minX(mnX), minY(mnY), maxX(mxX), maxY(mxY), keys(k), n(n_) {
    invW = static_cast<float>(COLS) / static_cast<float>(maxX - minX);
    invH = static_cast<float>(ROWS) / static_cast<float>(maxY - minY);
    for (int i = 0; i < n; i++) {
        const int px = (int)std::round((keys[i].x - minX) * invW);
        const int py = (int)std::round((keys[i].y - minY) * invH);
        if (px < 0 || px >= COLS || py < 0 || py >= ROWS) continue;
        grid[px][py].push_back(i);
    }
}

// Frame::GetFeaturesInArea — src/Frame.cc:326-379
std::vector<size_t> GridFrame::features_in_area(float x, float y, float r, int minLevel, int maxLevel) const {
    std::vector<size_t> v;
    const int nMinCellX = std::max(0, (int)std::floor((x - minX - r) * invW));
    if (nMinCellX >= COLS) return v;
    const int nMaxCellX = std::min((int)COLS - 1, (int)std::ceil((x - minX + r) * invW));
    if (nMaxCellX < 0) return v;
    const int nMinCellY = std::max(0, (int)std::floor((y - minY - r) * invH));
    if (nMinCellY >= ROWS) return v;
    const int nMaxCellY = std::min((int)ROWS - 1, (int)std::ceil((y - minY + r) * invH));
    if (nMaxCellY < 0) return v;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
            for (size_t idx : grid[ix][iy]) {
                const KeyPoint& kp = keys[idx];
                if (bCheckLevels) {
                    if (kp.octave < minLevel) continue;
                    if (maxLevel >= 0 && kp.octave > maxLevel) continue;
                }
                const float dx = kp.x - x, dy = kp.y - y;
                if (std::fabs(dx) < r && std::fabs(dy) < r) v.push_back(idx);
            }
    return v;
}

// ORBmatcher::ComputeThreeMaxima — src/ORBmatcher.cc:1601-1642
static void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

// ORBmatcher::SearchForInitialization — src/ORBmatcher.cc:405-520 (TH_LOW=50, HISTO_LENGTH=30 at :37-39)
int search_for_initialization(const GridFrame& F1, const uint8_t* desc1, const GridFrame& F2, const uint8_t* desc2,
                              float* prev, int* matches12, int windowSize, float nnratio, bool checkOri) {
    const int TH_LOW = 50, HISTO_LENGTH = 30;
    int nmatches = 0;
    for (int i = 0; i < F1.n; ++i) matches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> matchedDist(F2.n, INT_MAX), matches21(F2.n, -1);
    for (int i1 = 0; i1 < F1.n; i1++) {
        const KeyPoint kp1 = F1.keys[i1];
        const int level1 = kp1.octave;
        if (level1 > 0) continue;
        std::vector<size_t> cand = F2.features_in_area(prev[2 * i1], prev[2 * i1 + 1], (float)windowSize, level1, level1);
        if (cand.empty()) continue;
        const uint8_t* d1 = desc1 + (size_t)i1 * 32;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (size_t i2 : cand) {
            const int dist = descriptor_distance(d1, desc2 + i2 * 32);
            if (matchedDist[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = (int)i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (matches21[bestIdx2] >= 0) { matches12[matches21[bestIdx2]] = -1; nmatches--; }
                matches12[i1] = bestIdx2;
                matches21[bestIdx2] = i1;
                matchedDist[bestIdx2] = bestDist;
                nmatches++;
                if (checkOri) {
                    float rot = F1.keys[i1].angle - F2.keys[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < F1.n; i1++)
        if (matches12[i1] >= 0) {
            prev[2 * i1] = F2.keys[matches12[i1]].x;
            prev[2 * i1 + 1] = F2.keys[matches12[i1]].y;
        }
    return nmatches;
}

// ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono) — src/ORBmatcher.cc:1328-1470
// (TH_HIGH=100, HISTO_LENGTH=30 at :37-39).  The boundary is drawn after the pose arithmetic: x3Dc[i] = Rcw*x3Dw+tcw
// The relocalisation variant SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (:1472-1599) is the same loop
// with orbDist in place of TH_HIGH, no depth-sign test, the predicted level as octave1 and every assigned keypoint taken.
// (:1362-1363, cv::Mat expressions evaluated by the caller exactly as the reference does) and `direction` (0: octave +-1,
// 1: bForward, 2: bBackward, :1348-1349) come in; everything from :1365 on is restated here.
//   valid1[i]  = LastFrame.mvpMapPoints[i] && !LastFrame.mvbOutlier[i]       (:1353-1357)
//   obs1[i]    = that map point's Observations() > 0                           (decides whether a keypoint it was assigned to
//                                                                                is skipped by later iterations, :1404-1406)
//   occupied2  = CurrentFrame.mvpMapPoints[i2] && ...->Observations() > 0 on entry
//   match21[i2] = index i of the last-frame point assigned to current keypoint i2 (CurrentFrame.mvpMapPoints[i2] = pMP),
//                 -1 = untouched, -2 = set to NULL by the rotation-consistency pass (:1451-1467)
int search_by_projection_frame(int n1, const float* x3Dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mpDesc1,
                               const int* octave1, const float* angle1, const GridFrame& F2, const uint8_t* desc2,
                               const float* uRight2, const uint8_t* occupied2, float fx, float fy, float cx, float cy, float mbf,
                               const float* scaleFactors, float th, int direction, bool checkOri, int* match21, int orbDist,
                               bool depthTest) {
    const int HISTO_LENGTH = 30, TH_HIGH = orbDist;  // 100 (:39) or the ORBdist argument of the relocalisation variant (:1552)
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> owner(F2.n, -1);         // CurrentFrame.mvpMapPoints[i2] as the index of the last-frame point
    std::vector<uint8_t> taken(F2.n);         // ... && Observations() > 0
    for (int i = 0; i < F2.n; ++i) taken[i] = occupied2[i];
    for (int i = 0; i < n1; i++) {
        if (!valid1[i]) continue;
        const float xc = x3Dc[3 * i], yc = x3Dc[3 * i + 1];
        const float invzc = (float)(1.0 / x3Dc[3 * i + 2]);
        if (depthTest && invzc < 0) continue;  // frame-to-frame only (:1369); the relocalisation variant has no such test
        float u = fx * xc * invzc + cx;
        float v = fy * yc * invzc + cy;
        if (u < F2.minX || u > F2.maxX) continue;
        if (v < F2.minY || v > F2.maxY) continue;
        const int nLastOctave = octave1[i];
        const float radius = th * scaleFactors[nLastOctave];
        std::vector<size_t> vIndices2;
        if (direction == 1) vIndices2 = F2.features_in_area(u, v, radius, nLastOctave, -1);
        else if (direction == 2) vIndices2 = F2.features_in_area(u, v, radius, 0, nLastOctave);
        else vIndices2 = F2.features_in_area(u, v, radius, nLastOctave - 1, nLastOctave + 1);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = mpDesc1 + (size_t)i * 32;
        int bestDist = 256, bestIdx2 = -1;
        for (size_t i2 : vIndices2) {
            if (taken[i2]) continue;
            if (uRight2 && uRight2[i2] > 0) {
                const float ur = u - mbf * invzc;
                const float er = std::fabs(ur - uRight2[i2]);
                if (er > radius) continue;
            }
            const int dist = descriptor_distance(dMP, desc2 + i2 * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = (int)i2; }
        }
        if (bestDist <= TH_HIGH) {
            owner[bestIdx2] = i;
            taken[bestIdx2] = obs1[i];
            nmatches++;
            if (checkOri) {
                float rot = angle1[i] - F2.keys[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { owner[idx] = -2; nmatches--; }  // = static_cast<MapPoint*>(NULL)
    }
    for (int i = 0; i < F2.n; ++i) match21[i] = owner[i];
    return nmatches;
}

// ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, th) — src/ORBmatcher.cc:45-129 with
// RadiusByViewingCos (:131-137); called by Tracking::SearchLocalPoints (src/Tracking.cc:1270-1277).
//   valid1[i] = pMP->mbTrackInView && !pMP->isBad();  proj[3i..] = mTrackProjX, mTrackProjY, mTrackProjXR;
//   viewCos = mTrackViewCos;  level1 = mnTrackScaleLevel;  obs1 / occupied2 / match21 as above (no rotation pass: never -2)
int search_by_projection_map(int n1, const float* proj, const float* viewCos, const int* level1, const uint8_t* valid1,
                             const uint8_t* obs1, const uint8_t* mpDesc1, const GridFrame& F, const uint8_t* desc2,
                             const float* uRight2, const uint8_t* occupied2, const float* scaleFactors, float th, float nnratio,
                             int* match21) {
    const int TH_HIGH = 100;
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<int> owner(F.n, -1);
    std::vector<uint8_t> taken(F.n);
    for (int i = 0; i < F.n; ++i) taken[i] = occupied2[i];
    for (int iMP = 0; iMP < n1; iMP++) {
        if (!valid1[iMP]) continue;
        const int nPredictedLevel = level1[iMP];
        float r = viewCos[iMP] > 0.998 ? 2.5f : 4.0f;  // RadiusByViewingCos
        if (bFactor) r *= th;
        const std::vector<size_t> vIndices = F.features_in_area(proj[3 * iMP], proj[3 * iMP + 1], r * scaleFactors[nPredictedLevel],
                                                                nPredictedLevel - 1, nPredictedLevel);
        if (vIndices.empty()) continue;
        const uint8_t* MPdescriptor = mpDesc1 + (size_t)iMP * 32;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (size_t idx : vIndices) {
            if (taken[idx]) continue;
            if (uRight2 && uRight2[idx] > 0) {
                const float er = std::fabs(proj[3 * iMP + 2] - uRight2[idx]);
                if (er > r * scaleFactors[nPredictedLevel]) continue;
            }
            const int dist = descriptor_distance(MPdescriptor, desc2 + idx * 32);
            if (dist < bestDist) {
                bestDist2 = bestDist;
                bestDist = dist;
                bestLevel2 = bestLevel;
                bestLevel = F.keys[idx].octave;
                bestIdx = (int)idx;
            } else if (dist < bestDist2) {
                bestLevel2 = F.keys[idx].octave;
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
            owner[bestIdx] = iMP;
            taken[bestIdx] = obs1[iMP];
            nmatches++;
        }
    }
    for (int i = 0; i < F.n; ++i) match21[i] = owner[i];
    return nmatches;
}

// ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, vpPoints, vpMatched, th) — src/ORBmatcher.cc:290-403 (loop
// detection, src/LoopClosing.cc:278) from the radius search on (:357-398); the Sim3 / visibility arithmetic (:298-355) stays the
// caller's.  uvr[3i..] = u, v, radius = th*mvScaleFactors[nPredictedLevel];  level1 = nPredictedLevel;  valid1 = survived every
// `continue` of :318-351;  occupied2[idx] = vpMatched[idx] != NULL.  KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:578-617) is the
// Frame routine without a level filter; the level gate follows at :377-380.  match21[idx] = iMP (vpMatched[idx] = pMP), -1 = untouched.
int search_by_projection_sim3(int n1, const float* uvr, const int* level1, const uint8_t* valid1, const uint8_t* mpDesc1,
                              const GridFrame& KF, const uint8_t* desc2, const uint8_t* occupied2, int thDist, int* match21) {
    int nmatches = 0;
    std::vector<int> owner(KF.n, -1);
    std::vector<uint8_t> taken(KF.n);
    for (int i = 0; i < KF.n; ++i) taken[i] = occupied2[i];
    for (int iMP = 0; iMP < n1; iMP++) {
        if (!valid1[iMP]) continue;
        const int nPredictedLevel = level1[iMP];
        const std::vector<size_t> vIndices = KF.features_in_area(uvr[3 * iMP], uvr[3 * iMP + 1], uvr[3 * iMP + 2], -1, -1);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = mpDesc1 + (size_t)iMP * 32;
        int bestDist = 256, bestIdx = -1;
        for (size_t idx : vIndices) {
            if (taken[idx]) continue;
            const int kpLevel = KF.keys[idx].octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            const int dist = descriptor_distance(dMP, desc2 + idx * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
        }
        if (bestDist <= thDist) {  // TH_LOW (:392)
            owner[bestIdx] = iMP;
            taken[bestIdx] = 1;
            nmatches++;
        }
    }
    for (int i = 0; i < KF.n; ++i) match21[i] = owner[i];
    return nmatches;
}

// The matching core of both ORBmatcher::Fuse overloads: (pKF, vpMapPoints, th) src/ORBmatcher.cc:825-975 (candidate loop
// :893-945, with the chi-square gate on the reprojection error) and (pKF, Scw, vpPoints, th, vpReplacePoint) :977-1100 (candidate
// loop :1055-1075, no gate).  The projection and visibility tests before it and the map surgery after it (Replace /
// AddObservation / AddMapPoint, :948-972) are the caller's.  uvr = (u, v, radius); ur1 = u - bf*invz or NULL for the Sim3 overload.
void fuse_match(int n1, const float* uvr, const float* ur1, const int* level1, const uint8_t* valid1, const uint8_t* mpDesc1,
                const GridFrame& KF, const uint8_t* desc2, const float* uRight2, const float* invLevelSigma2, int* bestIdxOut,
                int* bestDistOut) {
    for (int i = 0; i < n1; ++i) {
        bestIdxOut[i] = -1;
        bestDistOut[i] = 256;
        if (!valid1[i]) continue;
        const float u = uvr[3 * i], v = uvr[3 * i + 1], radius = uvr[3 * i + 2];
        const int nPredictedLevel = level1[i];
        const std::vector<size_t> vIndices = KF.features_in_area(u, v, radius, -1, -1);
        const uint8_t* dMP = mpDesc1 + (size_t)i * 32;
        int bestDist = 256, bestIdx = -1;
        for (size_t idx : vIndices) {
            const KeyPoint& kp = KF.keys[idx];
            const int kpLevel = kp.octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            if (ur1) {
                if (uRight2 && uRight2[idx] >= 0) {  // reprojection error in stereo (:904-917)
                    const float ex = u - kp.x, ey = v - kp.y, er = ur1[i] - uRight2[idx];
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * invLevelSigma2[kpLevel] > 7.8) continue;
                } else {
                    const float ex = u - kp.x, ey = v - kp.y;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * invLevelSigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = descriptor_distance(dMP, desc2 + idx * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
        }
        bestIdxOut[i] = bestIdx;
        bestDistOut[i] = bestDist;
    }
}

// ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) — src/ORBmatcher.cc:159-288, and
// SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) — :522-655.  The vocabulary walk (DBoW2 FeatureVector merge, :175-181,
// :264-272) stays the caller's: it hands in the common nodes in order, node k holding the feature indices
// idx1[start1[k]..start1[k+1]) of the first and idx2[start2[k]..) of the second argument.
//   valid1[i1] = the keyframe-1 feature has a map point that is not bad (:192-198, :560-564)
//   valid2[i2] = 1 for the Frame variant; pMP2 && !pMP2->isBad() for the keyframe variant (:577-583)
//   thDist     = best distance accepted: TH_LOW for `<= TH_LOW` (:230), TH_LOW - 1 for `< TH_LOW` (:600)
// match12[i1] = i2 / -1 none / -2 removed by the rotation check; match21[i2] likewise (every i2 is assigned at most once).
int search_by_bow(int nNodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                  const uint8_t* valid1, const uint8_t* desc1, const float* angle1, int n2, const uint8_t* valid2,
                  const uint8_t* desc2, const float* angle2, int thDist, float nnratio, bool checkOri, int* match12, int* match21) {
    const int HISTO_LENGTH = 30;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];  // holds i1 (the keyframe variant stores idx1, the Frame variant bestIdxF = match12[i1])
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<uint8_t> taken(n2, 0);
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    for (int i = 0; i < n2; ++i) match21[i] = -1;
    for (int k = 0; k < nNodes; ++k) {
        for (int a = start1[k]; a < start1[k + 1]; ++a) {
            const int i1 = (int)idx1[a];
            if (!valid1[i1]) continue;
            int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
            for (int b = start2[k]; b < start2[k + 1]; ++b) {
                const int i2 = (int)idx2[b];
                if (taken[i2] || !valid2[i2]) continue;
                const int dist = descriptor_distance(desc1 + (size_t)i1 * 32, desc2 + (size_t)i2 * 32);
                if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = i2; }
                else if (dist < bestDist2) bestDist2 = dist;
            }
            if (bestDist1 <= thDist && (float)bestDist1 < nnratio * (float)bestDist2) {
                match12[i1] = bestIdx2;
                match21[bestIdx2] = i1;
                taken[bestIdx2] = 1;
                if (checkOri) {
                    float rot = angle1[i1] - angle2[bestIdx2];
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(i1);
                }
                nmatches++;
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int i1 : rotHist[i]) {
                match21[match12[i1]] = -2;
                match12[i1] = -2;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// Brute-force best / second-best with the TH + ratio acceptance test.  There is no single reference function
// (SURVEY.md §8a16); this follows the inner loop of ORBmatcher::SearchByBoW, src/ORBmatcher.cc:199-232:
// bestDist1 = bestDist2 = 256, strict '<' updates, accept iff d1 <= th && (float)d1 < ratio*(float)d2.
void match_bruteforce(const uint8_t* q, int nq, const uint8_t* d, int nd, int th, float ratio, int* idx, int* d1,
                      int* d2, int* match) {
    for (int i = 0; i < nq; ++i) {
        int b1 = 256, b2 = 256, bi = -1;
        for (int j = 0; j < nd; ++j) {
            const int dist = descriptor_distance(q + (size_t)i * 32, d + (size_t)j * 32);
            if (dist < b1) { b2 = b1; b1 = dist; bi = j; }
            else if (dist < b2) b2 = dist;
        }
        idx[i] = bi; d1[i] = b1; d2[i] = b2;
        match[i] = (b1 <= th && static_cast<float>(b1) < ratio * static_cast<float>(b2)) ? bi : -1;
    }
}

// Frame::ComputeStereoMatches — src/Frame.cc:465-639.  cv::norm(IL, IR, NORM_L1) on centre-subtracted
// integer-valued float patches is an exact integer sum (<= 121*510), so it is evaluated in int here.
void stereo_matches(const Extractor& EL, const Extractor& ER, const KeyPoint* kL, int N, const uint8_t* dL,
                    const KeyPoint* kR, int Nr, const uint8_t* dR, float mbf, float fx, float* uRight, float* depth) {
    const int TH_HIGH = 100, TH_LOW = 50;
    for (int i = 0; i < N; ++i) { uRight[i] = -1.0f; depth[i] = -1.0f; }
    const float mb = mbf / fx;
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int nRows = EL.lh[0];
    std::vector<std::vector<size_t>> rowIdx(nRows);
    for (int iR = 0; iR < Nr; iR++) {
        const float kpY = kR[iR].y;
        const float r = 2.0f * EL.mvScaleFactor[kR[iR].octave];
        const int maxr = (int)std::ceil(kpY + r), minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) rowIdx[yi].push_back(iR);  // bounds guard only; in-range for real keypoints
    }
    const float minZ = mb, minD = 0, maxD = mbf / minZ;
    std::vector<std::pair<int, int>> distIdx;
    for (int iL = 0; iL < N; iL++) {
        const KeyPoint& kpL = kL[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        const std::vector<size_t>& cand = rowIdx[(size_t)vL];
        if (cand.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        size_t bestIdxR = 0;
        for (size_t iR : cand) {
            const KeyPoint& kpR = kR[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = descriptor_distance(dL + (size_t)iL * 32, dR + iR * 32);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {
            const float uR0 = kR[bestIdxR].x;
            const float scaleFactor = EL.mvInvScaleFactor[kpL.octave];
            const float scaleduL = std::round(kpL.x * scaleFactor);
            const float scaledvL = std::round(kpL.y * scaleFactor);
            const float scaleduR0 = std::round(uR0 * scaleFactor);
            const int w = 5, L = 5;
            const uint8_t* PL = EL.level_ptr(kpL.octave);
            const size_t sL = EL.level_step(kpL.octave);
            const uint8_t* PR = ER.level_ptr(kpL.octave);
            const size_t sR = ER.level_step(kpL.octave);
            const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w);
            const int cL = PL[(size_t)(y0 + w) * sL + xl0 + w];
            int best = INT_MAX, bestincR = 0;
            std::vector<float> vDists(2 * L + 1);
            const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= ER.lw[kpL.octave]) continue;
            for (int incR = -L; incR <= +L; incR++) {
                const int xr0 = (int)(scaleduR0 + incR - w);
                const int cR = PR[(size_t)(y0 + w) * sR + xr0 + w];
                int sad = 0;
                for (int yy = 0; yy < 2 * w + 1; ++yy)
                    for (int xx = 0; xx < 2 * w + 1; ++xx)
                        sad += std::abs((PL[(size_t)(y0 + yy) * sL + xl0 + xx] - cL) - (PR[(size_t)(y0 + yy) * sR + xr0 + xx] - cR));
                const float dist = (float)sad;
                if (dist < best) { best = (int)dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = EL.mvScaleFactor[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                depth[iL] = mbf / disparity;
                uRight[iL] = bestuR;
                distIdx.push_back(std::pair<int, int>(best, iL));
            }
        }
    }
    if (distIdx.empty()) return;  // the reference reads element 0 of an empty vector here (undefined)
    std::sort(distIdx.begin(), distIdx.end());
    const float median = distIdx[distIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    for (int i = (int)distIdx.size() - 1; i >= 0; i--) {
        if (distIdx[i].first < thDist) break;
        uRight[distIdx[i].second] = -1;
        depth[distIdx[i].second] = -1;
    }
}

// ---- DBoW2 tree descent ---------------------------------------------------------------------------------------------
// Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1261 (transform of one feature) over the node table of
// loadFromTextFile (:1338-1424): children in file order (push_back :1388), word ids counted over the leaf-flagged lines
// (:1406-1413), isLeaf() = children.empty() (:111).  F::distance = FORB::distance, the 256-bit Hamming distance
// (Thirdparty/DBoW2/DBoW2/FORB.cpp:98-121), compared as double with a strict '<' so the first child wins ties.
void vocabulary_transform(int nNodes, const int* parent, const uint8_t* isLeaf, const uint8_t* desc, const double* weight, int L,
                          int n, const uint8_t* feat, int levelsup, uint32_t* wordId, double* w, uint32_t* nodeId) {
    std::vector<std::vector<int>> children(nNodes);
    std::vector<uint32_t> word(nNodes, 0);
    uint32_t nWords = 0;
    for (int i = 1; i < nNodes; ++i) {
        children[parent[i]].push_back(i);
        if (isLeaf[i]) word[i] = nWords++;
    }
    const int nidLevel = L - levelsup;
    for (int f = 0; f < n; ++f) {
        const uint8_t* a = feat + (size_t)32 * f;
        uint32_t nid = nidLevel <= 0 ? 0u : 0xFFFFFFFFu;
        int finalId = 0, level = 0;
        do {
            ++level;
            const std::vector<int>& nodes = children[finalId];
            finalId = nodes[0];
            double best = (double)descriptor_distance(a, desc + (size_t)32 * finalId);
            for (size_t c = 1; c < nodes.size(); ++c) {
                const double d = (double)descriptor_distance(a, desc + (size_t)32 * nodes[c]);
                if (d < best) { best = d; finalId = nodes[c]; }
            }
            if (level == nidLevel) nid = (uint32_t)finalId;
        } while (!children[finalId].empty());
        wordId[f] = word[finalId];
        w[f] = weight[finalId];
        nodeId[f] = nid;
    }
}

// ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-823) over the common vocabulary nodes, CheckDistEpipolarLine
// (:140-157) inlined.  vbMatched2 (:683) is never set by the reference, so it is not modelled.
int search_for_triangulation(int nNodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                             const KeyPoint* k1, const uint8_t* desc1, const uint8_t* valid1, const uint8_t* stereo1, int n2,
                             const KeyPoint* k2, const uint8_t* desc2, const uint8_t* valid2, const uint8_t* stereo2, const float* F12,
                             float ex, float ey, const float* scaleFactors2, const float* levelSigma2, bool checkOri, int* match12) {
    (void)n2;
    int nmatches = 0;
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    std::vector<int> rotHist[30];
    const float factor = 1.0f / 30;
    for (int node = 0; node < nNodes; ++node) {
        for (int p1 = start1[node]; p1 < start1[node + 1]; ++p1) {
            const int i1 = (int)idx1[p1];
            if (!valid1[i1]) continue;  // :697-706
            const bool bStereo1 = stereo1[i1] != 0;
            const KeyPoint& kp1 = k1[i1];
            int bestDist = 50, bestIdx2 = -1;  // TH_LOW
            for (int p2 = start2[node]; p2 < start2[node + 1]; ++p2) {
                const int i2 = (int)idx2[p2];
                if (!valid2[i2]) continue;  // :724-733
                const bool bStereo2 = stereo2[i2] != 0;
                const int dist = descriptor_distance(desc1 + (size_t)32 * i1, desc2 + (size_t)32 * i2);
                if (dist > 50 || dist > bestDist) continue;
                const KeyPoint& kp2 = k2[i2];
                if (!bStereo1 && !bStereo2) {
                    const float distex = ex - kp2.x, distey = ey - kp2.y;
                    if (distex * distex + distey * distey < 100 * scaleFactors2[kp2.octave]) continue;
                }
                const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
                const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
                const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
                const float num = a * kp2.x + b * kp2.y + c;
                const float den = a * a + b * b;
                if (den == 0) continue;
                const float dsqr = num * num / den;
                if (dsqr < 3.84 * levelSigma2[kp2.octave]) {
                    bestIdx2 = i2;
                    bestDist = dist;
                }
            }
            if (bestIdx2 >= 0) {
                match12[i1] = bestIdx2;
                nmatches++;
                if (checkOri) {
                    float rot = kp1.angle - k2[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == 30) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (checkOri) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, 30, ind1, ind2, ind3);
        for (int i = 0; i < 30; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); ++j) {
                match12[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:246-312) on the N gathered descriptors: N x N distance table
// (:281-291, stored as float there; the values are small integers), per row sort + vDists[0.5 * (N - 1)] (:296-298), strict '<'.
int distinctive_descriptor(const uint8_t* desc, int n, int* medianOut) {
    int bestMedian = INT_MAX, bestIdx = 0;
    std::vector<int> row(n);
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) row[j] = i == j ? 0 : descriptor_distance(desc + (size_t)32 * i, desc + (size_t)32 * j);
        std::sort(row.begin(), row.end());
        const int median = row[(size_t)(0.5 * (n - 1))];
        if (median < bestMedian) { bestMedian = median; bestIdx = i; }
    }
    if (medianOut) *medianOut = n > 0 ? bestMedian : -1;
    return n > 0 ? bestIdx : -1;
}

}  // namespace orc
