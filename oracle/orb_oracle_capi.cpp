// oracle/orb_oracle_capi.cpp — TEST INFRASTRUCTURE ONLY: flat C entry points over orb_oracle.cpp so that the
// Python test-suite / bench cpu_baseline leg can drive the CPU oracle through ctypes.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <thread>

#include "orb_oracle.hpp"

using namespace orc;

extern "C" {

int orc_cv_round_f(float v) { return cv_round(v); }
float orc_fast_atan2(float y, float x) { return fast_atan2(y, x); }
void orc_fast_atan2_array(const float* y, const float* x, float* out, int n) {
    for (int i = 0; i < n; ++i) out[i] = fast_atan2(y[i], x[i]);
}
void orc_resize_linear_8u(const uint8_t* s, int sw, int sh, int sstep, uint8_t* d, int dw, int dh, int dstep) {
    resize_linear_8u(s, sw, sh, sstep, d, dw, dh, dstep);
}
void orc_border_reflect101(const uint8_t* s, int w, int h, int sstep, uint8_t* d, int dstep, int pad) {
    border_reflect101(s, w, h, sstep, d, dstep, pad);
}
int orc_fast9(const uint8_t* img, int w, int h, int step, int th, int nms, KeyPoint* out, int cap) {
    std::vector<KeyPoint> v;
    fast9_16(img, w, h, step, th, nms != 0, v);
    const int n = (int)v.size();
    if (out) std::memcpy(out, v.data(), sizeof(KeyPoint) * (size_t)std::min(n, cap));
    return n;
}
void orc_gaussian7(const uint8_t* s, int w, int h, int sstep, uint8_t* d, int dstep) { gaussian7x7_s2(s, w, h, sstep, d, dstep); }
void orc_sincosf(const float* ang, float* c, float* s, int n) {  // glibc float overloads, as the reference calls them
    for (int i = 0; i < n; ++i) { c[i] = cosf(ang[i]); s[i] = sinf(ang[i]); }
}

void* orc_extractor_create(int nf, float sf, int nl, int ini, int mn) { return new Extractor(nf, sf, nl, ini, mn); }
void orc_extractor_destroy(void* e) { delete (Extractor*)e; }
void orc_extractor_tables(void* e, float* sf, float* inv, float* s2, float* inv2, int* quota, int* umax) {
    Extractor& E = *(Extractor*)e;
    for (int i = 0; i < E.nlevels; ++i) {
        sf[i] = E.mvScaleFactor[i]; inv[i] = E.mvInvScaleFactor[i];
        s2[i] = E.mvLevelSigma2[i]; inv2[i] = E.mvInvLevelSigma2[i];
        quota[i] = E.mnFeaturesPerLevel[i];
    }
    for (int i = 0; i < 16; ++i) umax[i] = E.umax[i];
}
// returns the number of keypoints; fills at most cap
int orc_extract(void* e, const uint8_t* img, int w, int h, int step, KeyPoint* kps, uint8_t* desc, int cap) {
    Extractor& E = *(Extractor*)e;
    std::vector<KeyPoint> k;
    std::vector<uint8_t> d;
    E.extract(img, w, h, step, k, d);
    const int n = (int)k.size(), m = std::min(n, cap);
    if (kps) std::memcpy(kps, k.data(), sizeof(KeyPoint) * (size_t)m);
    if (desc) std::memcpy(desc, d.data(), 32 * (size_t)m);
    return n;
}
void orc_level_size(void* e, int l, int* w, int* h) { Extractor& E = *(Extractor*)e; *w = E.lw[l]; *h = E.lh[l]; }
// copies the padded (w+38)x(h+38) plane, tightly packed
void orc_level_padded(void* e, int l, uint8_t* out) { Extractor& E = *(Extractor*)e; std::memcpy(out, E.padded[l].px.data(), E.padded[l].px.size()); }
int orc_level_blurred(void* e, int l, uint8_t* out) {
    Extractor& E = *(Extractor*)e;
    if (E.blurred[l].px.empty()) return 0;
    std::memcpy(out, E.blurred[l].px.data(), E.blurred[l].px.size());
    return 1;
}
static int copy_kps(const std::vector<KeyPoint>& v, KeyPoint* out, int cap) {
    if (out) std::memcpy(out, v.data(), sizeof(KeyPoint) * (size_t)std::min((int)v.size(), cap));
    return (int)v.size();
}
int orc_level_candidates(void* e, int l, KeyPoint* out, int cap) { return copy_kps(((Extractor*)e)->candidates[l], out, cap); }
int orc_level_keypoints(void* e, int l, KeyPoint* out, int cap) { return copy_kps(((Extractor*)e)->levelKeys[l], out, cap); }

int orc_distribute(const KeyPoint* in, int n, int minX, int maxX, int minY, int maxY, int N, KeyPoint* out, int cap) {
    std::vector<KeyPoint> v(in, in + n);
    return copy_kps(distribute_octtree(v, minX, maxX, minY, maxY, N), out, cap);
}
float orc_ic_angle(const uint8_t* plane, int step, int x, int y) {
    static Extractor E(1000, 1.2f, 8, 20, 7);
    return ic_angle(plane + (size_t)y * step + x, step, E.umax);
}
void orc_descriptor(const uint8_t* plane, int step, int x, int y, float angle, uint8_t* desc) {
    orb_descriptor(angle, plane + (size_t)y * step + x, step, desc);
}
int orc_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }
void orc_set_pattern_variant(int v) { set_pattern_variant(v); }
int orc_pattern_variant() { return pattern_variant(); }
void orc_pattern_values(int* out) { std::memcpy(out, pattern_values(), 1024 * sizeof(int)); }

int orc_search_for_initialization(const KeyPoint* k1, int n1, const uint8_t* d1, const KeyPoint* k2, int n2, const uint8_t* d2,
                                  float minX, float maxX, float minY, float maxY, float* prevXY, int* matches12,
                                  int windowSize, float nnratio, int checkOri) {
    GridFrame F1(k1, n1, minX, maxX, minY, maxY), F2(k2, n2, minX, maxX, minY, maxY);
    return search_for_initialization(F1, d1, F2, d2, prevXY, matches12, windowSize, nnratio, checkOri != 0);
}
// Frame::GetFeaturesInArea on the 64 x 48 grid (src/Frame.cc:229-244, :326-379): indices in the reference's visiting order
int orc_features_in_area(const KeyPoint* k, int n, float minX, float maxX, float minY, float maxY, float x, float y, float r,
                         int minLevel, int maxLevel, int* out) {
    GridFrame F(k, n, minX, maxX, minY, maxY);
    const std::vector<size_t> v = F.features_in_area(x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size(); ++i) out[i] = (int)v[i];
    return (int)v.size();
}

int orc_search_by_projection_frame(int n1, const float* x3Dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mpDesc1,
                                   const int* octave1, const float* angle1, const KeyPoint* k2, int n2, const uint8_t* desc2,
                                   const float* uRight2, const uint8_t* occupied2, const float* cam, float mbf,
                                   const float* bounds, const float* scaleFactors, float th, int direction, int checkOri,
                                   int* match21, int orbDist, int depthTest) {
    GridFrame F2(k2, n2, bounds[0], bounds[1], bounds[2], bounds[3]);
    return search_by_projection_frame(n1, x3Dc, valid1, obs1, mpDesc1, octave1, angle1, F2, desc2, uRight2, occupied2, cam[0],
                                      cam[1], cam[2], cam[3], mbf, scaleFactors, th, direction, checkOri != 0, match21, orbDist, depthTest != 0);
}
int orc_search_by_projection_map(int n1, const float* proj, const float* viewCos, const int* level1, const uint8_t* valid1,
                                 const uint8_t* obs1, const uint8_t* mpDesc1, const KeyPoint* k2, int n2, const uint8_t* desc2,
                                 const float* uRight2, const uint8_t* occupied2, const float* bounds, const float* scaleFactors,
                                 float th, float nnratio, int* match21) {
    GridFrame F(k2, n2, bounds[0], bounds[1], bounds[2], bounds[3]);
    return search_by_projection_map(n1, proj, viewCos, level1, valid1, obs1, mpDesc1, F, desc2, uRight2, occupied2, scaleFactors,
                                    th, nnratio, match21);
}
int orc_search_by_projection_sim3(int n1, const float* uvr, const int* level1, const uint8_t* valid1, const uint8_t* mpDesc1,
                                  const KeyPoint* k2, int n2, const uint8_t* desc2, const uint8_t* occupied2, const float* bounds,
                                  int thDist, int* match21) {
    GridFrame KF(k2, n2, bounds[0], bounds[1], bounds[2], bounds[3]);
    return search_by_projection_sim3(n1, uvr, level1, valid1, mpDesc1, KF, desc2, occupied2, thDist, match21);
}
void orc_fuse_match(int n1, const float* uvr, const float* ur1, const int* level1, const uint8_t* valid1, const uint8_t* mpDesc1,
                    const KeyPoint* k2, int n2, const uint8_t* desc2, const float* uRight2, const float* invLevelSigma2,
                    const float* bounds, int* bestIdx, int* bestDist) {
    GridFrame KF(k2, n2, bounds[0], bounds[1], bounds[2], bounds[3]);
    fuse_match(n1, uvr, ur1, level1, valid1, mpDesc1, KF, desc2, uRight2, invLevelSigma2, bestIdx, bestDist);
}
int orc_search_by_bow(int nNodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                      const uint8_t* valid1, const uint8_t* desc1, const float* angle1, int n2, const uint8_t* valid2,
                      const uint8_t* desc2, const float* angle2, int thDist, float nnratio, int checkOri, int* match12, int* match21) {
    return search_by_bow(nNodes, start1, idx1, start2, idx2, n1, valid1, desc1, angle1, n2, valid2, desc2, angle2, thDist, nnratio,
                         checkOri != 0, match12, match21);
}
void orc_match_bruteforce(const uint8_t* q, int nq, const uint8_t* d, int nd, int th, float ratio, int* idx, int* d1, int* d2, int* match) {
    match_bruteforce(q, nq, d, nd, th, ratio, idx, d1, d2, match);
}
void orc_stereo_matches(void* eL, void* eR, const KeyPoint* kL, int nL, const uint8_t* dL, const KeyPoint* kR, int nR,
                        const uint8_t* dR, float bf, float fx, float* uRight, float* depth) {
    stereo_matches(*(Extractor*)eL, *(Extractor*)eR, kL, nL, dL, kR, nR, dR, bf, fx, uRight, depth);
}

// ---- timing helpers for bench.py's cpu_baseline / --impl reference legs -----------------------------------
// Extract `nframes` frames (tightly packed, same size) with `threads` host threads, frame-parallel, one
// Extractor per thread (the reference extractor is single-threaded; src/Frame.cc:77-80 runs one per eye).
// Returns wall seconds; total keypoints in *nk.
double orc_bench_extract(const uint8_t* frames, int nframes, int w, int h, int nf, float sf, int nl, int ini, int mn,
                         int threads, long long* nk) {
    std::vector<long long> counts(threads, 0);
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() {
            Extractor E(nf, sf, nl, ini, mn);
            std::vector<KeyPoint> k;
            std::vector<uint8_t> d;
            for (int f = t; f < nframes; f += threads) {
                E.extract(frames + (size_t)f * w * h, w, h, w, k, d);
                counts[t] += (long long)k.size();
            }
        });
    for (auto& th : pool) th.join();
    auto t1 = std::chrono::steady_clock::now();
    long long s = 0;
    for (long long c : counts) s += c;
    if (nk) *nk = s;
    return std::chrono::duration<double>(t1 - t0).count();
}
void orc_vocabulary_transform(int nNodes, const int* parent, const uint8_t* isLeaf, const uint8_t* desc, const double* weight, int L,
                              int n, const uint8_t* feat, int levelsup, uint32_t* wordId, double* w, uint32_t* nodeId) {
    vocabulary_transform(nNodes, parent, isLeaf, desc, weight, L, n, feat, levelsup, wordId, w, nodeId);
}
int orc_search_for_triangulation(int nNodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                                 const KeyPoint* k1, const uint8_t* desc1, const uint8_t* valid1, const uint8_t* stereo1, int n2,
                                 const KeyPoint* k2, const uint8_t* desc2, const uint8_t* valid2, const uint8_t* stereo2, const float* F12,
                                 float ex, float ey, const float* scaleFactors2, const float* levelSigma2, int checkOri, int* match12) {
    return search_for_triangulation(nNodes, start1, idx1, start2, idx2, n1, k1, desc1, valid1, stereo1, n2, k2, desc2, valid2, stereo2,
                                    F12, ex, ey, scaleFactors2, levelSigma2, checkOri != 0, match12);
}
void orc_distinctive_descriptors(int nPoints, const int* start, const uint8_t* desc, int* bestIdx, int* bestMedian) {
    for (int p = 0; p < nPoints; ++p) bestIdx[p] = distinctive_descriptor(desc + (size_t)32 * start[p], start[p + 1] - start[p], bestMedian + p);
}
double orc_bench_bruteforce(const uint8_t* q, int nq, const uint8_t* d, int nd, int threads, int* idx, int* d1, int* d2, int* match) {
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() {
            const int b = (int)((long long)nq * t / threads), e = (int)((long long)nq * (t + 1) / threads);
            match_bruteforce(q + (size_t)b * 32, e - b, d, nd, 50, 0.6f, idx + b, d1 + b, d2 + b, match + b);
        });
    for (auto& th : pool) th.join();
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // extern "C"
