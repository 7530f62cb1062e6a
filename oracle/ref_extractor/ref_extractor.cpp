// oracle/ref_extractor/ref_extractor.cpp — TEST INFRASTRUCTURE.  C entry points around the reference's OWN ORBextractor, whose
// src/ORBextractor.cc is compiled where it lies (Makefile) with OPENVX undefined, so operator() takes its CPU branch
// (src/ORBextractor.cc:1696-1815: ComputePyramid, ComputeKeyPointsOctTree, GaussianBlur, computeDescriptors, rescale).
// The OpenVX calls in the constructor's buildGraph() land in the do-nothing stand-in of shim/NVX/nvx.h.
//
// One thing buildGraph() does matters on the CPU branch: with USE_PYRAMID defined (src/ORBextractor.cc:640) it overwrites
// mvScaleFactor / mvInvScaleFactor from the level widths of the VisionWorks pyramid (:674-680).  There is no VisionWorks here,
// so the wrapper puts the constructor's own geometric tables (:501-517) back before the first frame.  The product and the
// oracle use those same stock tables (DESIGN.md §2 says so).
//
// Tie-breaks.  DistributeOctTree sorts pair<int, ExtractorNode*> (src/ORBextractor.cc:1010-1015 and :1040): among nodes holding
// the same number of keypoints the order is the order of their HEAP ADDRESSES, i.e. whatever the allocator did — with glibc's
// malloc (freed list nodes are reused last-in-first-out) the reference's output is a function of its allocation history.  To
// make the compiled reference well defined, this library replaces operator new / delete FOR ITSELF ONLY (linked -Bsymbolic):
// blocks of exactly one std::list<ExtractorNode> node come from a bump arena that never reuses memory, everything else from
// malloc.  Node addresses then increase with creation time, so "higher address first" (the reference walks the sorted vector
// from the back) reads "the node created last goes first" — the rule the oracle and the CUDA quadtree implement.
#include <pthread.h>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <cstring>
#include <list>
#include <new>
#include <vector>
#include "ORBextractor.h"

namespace {
const size_t kNodeBytes = sizeof(std::_List_node<ORB_SLAM2::ExtractorNode>);
const size_t kArenaBytes = (size_t)64 << 20;  // one frame creates a few 10^4 nodes of ~100 B
thread_local char* g_arena = nullptr;         // per thread: the timed CPU arm runs one extractor per host thread
thread_local size_t g_arenaUsed = 0;
// every arena ever handed out (a block may be freed on another thread than the one that took it); arenas live until exit
char* g_arenas[1024];
std::atomic<int> g_nArenas{0};
inline bool in_arena(const void* p) {
    const int n = g_nArenas.load(std::memory_order_acquire);
    for (int i = 0; i < n; ++i)
        if ((const char*)p >= g_arenas[i] && (const char*)p < g_arenas[i] + kArenaBytes) return true;
    return false;
}
char* g_idle[1024];  // arenas given back by finished bench threads
int g_nIdle = 0;
char* new_arena() {
    if (g_nIdle > 0) return g_idle[--g_nIdle];
    char* a = (char*)malloc(kArenaBytes);
    const int i = g_nArenas.load(std::memory_order_relaxed);
    if (!a || i >= 1024) {
        fprintf(stderr, "ref_extractor: no memory for a node arena\n");
        abort();
    }
    g_arenas[i] = a;  // callers serialise through g_arenaLock
    g_nArenas.store(i + 1, std::memory_order_release);
    return a;
}
pthread_mutex_t g_arenaLock = PTHREAD_MUTEX_INITIALIZER;
void arena_give_back() {  // at the end of a worker thread
    if (!g_arena) return;
    pthread_mutex_lock(&g_arenaLock);
    g_idle[g_nIdle++] = g_arena;
    pthread_mutex_unlock(&g_arenaLock);
    g_arena = nullptr;
    g_arenaUsed = 0;
}
void arena_reset() { g_arenaUsed = 0; }  // called between frames: every list of the previous call is gone by then
void* refx_alloc(size_t n) {
    if (n == kNodeBytes) {
        if (!g_arena) {
            pthread_mutex_lock(&g_arenaLock);
            g_arena = new_arena();
            pthread_mutex_unlock(&g_arenaLock);
        }
        const size_t need = (n + 15) & ~(size_t)15;
        if (g_arena && g_arenaUsed + need <= kArenaBytes) {
            void* p = g_arena + g_arenaUsed;
            g_arenaUsed += need;
            return p;
        }
        fprintf(stderr, "ref_extractor: node arena exhausted\n");
        abort();
    }
    void* p = malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
inline void refx_free(void* p) {
    if (p && !in_arena(p)) free(p);
}
}  // namespace
void* operator new(size_t n) { return refx_alloc(n); }
void* operator new[](size_t n) { return refx_alloc(n); }
void operator delete(void* p) noexcept { refx_free(p); }
void operator delete[](void* p) noexcept { refx_free(p); }
void operator delete(void* p, size_t) noexcept { refx_free(p); }
void operator delete[](void* p, size_t) noexcept { refx_free(p); }

namespace {
struct Probe : public ORB_SLAM2::ORBextractor {
    Probe(int nf, float sf, int nl, int ini, int mn, int w, int h) : ORB_SLAM2::ORBextractor(nf, sf, nl, ini, mn, w, h) {
        // the constructor's loop, src/ORBextractor.cc:501-517
        mvScaleFactor[0] = 1.0f;
        for (int i = 1; i < nlevels; i++) mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * scaleFactor);
        for (int i = 0; i < nlevels; i++) mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
    }
    ~Probe() { times.clear(); }  // otherwise the base destructor appends a times.csv to the working directory
    void quiet() { times.clear(); }
    std::vector<int>& perLevel() { return mnFeaturesPerLevel; }
    std::vector<float>& scales() { return mvScaleFactor; }
    std::vector<int>& umaxTab() { return umax; }
    std::vector<cv::Point>& pat() { return pattern; }
    std::vector<cv::KeyPoint> distribute(const std::vector<cv::KeyPoint>& k, int minX, int maxX, int minY, int maxY, int n, int level) {
        return DistributeOctTree(k, minX, maxX, minY, maxY, n, level);
    }
};
}  // namespace

namespace {
// the reference's constructor prints "w,h" and its destructor "Avg computed frame ORB: ..." on std::cout; bench.py's contract
// is ONE line on stdout, so both run with cout parked
struct QuietCout {
    std::streambuf* old;
    QuietCout() : old(std::cout.rdbuf(nullptr)) {}
    ~QuietCout() {
        std::cout.rdbuf(old);
        std::cout.clear();
    }
};
struct BenchArg {
    const uint8_t* frames;
    int nframes, w, h, t, threads;
    long long count;
    Probe* probe;
};
void* bench_thread(void* v) {
    BenchArg* a = (BenchArg*)v;
    std::vector<cv::KeyPoint> k;
    for (int f = a->t; f < a->nframes; f += a->threads) {
        arena_reset();
        cv::Mat image(a->h, a->w, CV_8UC1, (void*)(a->frames + (size_t)f * a->w * a->h), (size_t)a->w), mask, d;
        (*a->probe)(image, mask, k, d);
        a->probe->quiet();
        a->count += (long long)k.size();
    }
    arena_give_back();
    return nullptr;
}
}  // namespace

extern "C" {

void* refx_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int width, int height) {
    QuietCout q;
    return new Probe(nfeatures, scaleFactor, nlevels, iniTh, minTh, width, height);
}
void refx_destroy(void* h) {
    QuietCout q;
    delete (Probe*)h;
}

// The timed CPU arm of bench.py (--impl reference, cpu_baseline kind "reference"): `threads` host threads, one reference
// ORBextractor each (the reference's extractor is single-threaded; src/Frame.cc:77-80 runs one per eye), frame f on thread
// f mod threads.  Extractors are built before the clock starts.  Returns wall seconds; total keypoints in *nk.
double refx_bench_extract(const uint8_t* frames, int nframes, int w, int h, int nf, float sf, int nl, int ini, int mn, int threads,
                          long long* nk) {
    std::vector<BenchArg> args((size_t)threads);
    std::vector<pthread_t> tid((size_t)threads);
    {
        QuietCout q;
        for (int t = 0; t < threads; ++t) {
            args[t] = BenchArg{frames, nframes, w, h, t, threads, 0, nullptr};
            args[t].probe = new Probe(nf, sf, nl, ini, mn, w, h);
        }
    }
    auto t0 = std::chrono::steady_clock::now();
    for (int t = 0; t < threads; ++t) pthread_create(&tid[t], nullptr, bench_thread, &args[t]);
    for (int t = 0; t < threads; ++t) pthread_join(tid[t], nullptr);
    auto t1 = std::chrono::steady_clock::now();
    long long s = 0;
    {
        QuietCout q;
        for (int t = 0; t < threads; ++t) {
            s += args[t].count;
            delete args[t].probe;
        }
    }
    if (nk) *nk = s;
    return std::chrono::duration<double>(t1 - t0).count();
}

// tables the constructor builds: features per level, scale factors, umax, the 512 pattern points (x, y interleaved)
void refx_tables(void* h, int* perLevel, float* scale, int* umax16, int* pattern1024) {
    Probe* p = (Probe*)h;
    for (size_t i = 0; i < p->perLevel().size(); ++i) perLevel[i] = p->perLevel()[i];
    for (size_t i = 0; i < p->scales().size(); ++i) scale[i] = p->scales()[i];
    for (size_t i = 0; i < p->umaxTab().size() && i < 16; ++i) umax16[i] = p->umaxTab()[i];
    for (size_t i = 0; i < p->pat().size() && i < 512; ++i) {
        pattern1024[2 * i] = p->pat()[i].x;
        pattern1024[2 * i + 1] = p->pat()[i].y;
    }
}

// ORBextractor::operator() on one 8-bit frame.  kps: 7 floats per keypoint (x, y, size, angle, response, octave, class_id);
// desc: 32 bytes per keypoint.  Returns the keypoint count (or -needed when cap is too small).
int refx_extract(void* h, const uint8_t* img, int w, int ht, size_t step, float* kps, uint8_t* desc, int cap) {
    Probe* p = (Probe*)h;
    arena_reset();
    cv::Mat image(ht, w, CV_8UC1, (void*)img, step), mask, d;
    std::vector<cv::KeyPoint> k;
    (*p)(image, mask, k, d);
    p->quiet();
    const int n = (int)k.size();
    if (n > cap) return -n;
    for (int i = 0; i < n; ++i) {
        float* o = kps + 7 * (size_t)i;
        o[0] = k[i].pt.x; o[1] = k[i].pt.y; o[2] = k[i].size; o[3] = k[i].angle; o[4] = k[i].response;
        o[5] = (float)k[i].octave; o[6] = (float)k[i].class_id;
        memcpy(desc + 32 * (size_t)i, d.ptr(i), 32);
    }
    return n;
}

// the reference's pyramid level `level` of the last frame (without its border), packed; returns w | h << 16
int refx_level(void* h, int level, uint8_t* out, int cap) {
    Probe* p = (Probe*)h;
    const cv::Mat& m = p->mvImagePyramid[level];
    if (m.rows * m.cols > cap) return -1;
    for (int y = 0; y < m.rows; ++y) memcpy(out + (size_t)y * m.cols, m.ptr(y), (size_t)m.cols);
    return m.cols | (m.rows << 16);
}

// ORBextractor::DistributeOctTree on caller keypoints (x, y, response triples); returns the kept count, out = same triples
int refx_distribute(void* h, const float* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int level, float* out, int cap) {
    Probe* p = (Probe*)h;
    arena_reset();
    std::vector<cv::KeyPoint> k((size_t)n);
    for (int i = 0; i < n; ++i) {
        k[i].pt.x = xyr[3 * i];
        k[i].pt.y = xyr[3 * i + 1];
        k[i].response = xyr[3 * i + 2];
        k[i].size = 7.f;
    }
    std::vector<cv::KeyPoint> r = p->distribute(k, minX, maxX, minY, maxY, N, level);
    const int m = (int)r.size();
    if (m > cap) return -m;
    for (int i = 0; i < m; ++i) {
        out[3 * i] = r[i].pt.x;
        out[3 * i + 1] = r[i].pt.y;
        out[3 * i + 2] = r[i].response;
    }
    return m;
}

}  // extern "C"
