// csrc/matcher.cu — C entry points of the matcher half of include/orbx.h: ORBmatcher::DescriptorDistance,
// brute-force best / second-best, all-vs-all keyframe scoring, SearchForInitialization and
// Frame::ComputeStereoMatches.  The kernels live in k_match.cu / k_search.cu / k_stereo.cu; this file owns the
// per-thread device workspace (a stream and grow-only scratch per host thread: ORBmatcher objects are stack-local
// on three threads in the reference, src/Tracking.cc:671, src/LocalMapping.cc:248, src/LoopClosing.cc:278) and the
// host <-> device copies of the "host buffer" entry points.  There is no CPU path.
#include <dlfcn.h>

#include <algorithm>
#include <cstring>
#include <vector>

#include "orbx_internal.h"

using namespace orbx;

namespace {

struct Workspace {  // one per host thread and device
    int device = -1;
    cudaStream_t stream = nullptr;
    DevBuf a, b, c, d, e, f, g, h, out0, out1, out2, out3, status, partial, counters;
    PinBuf pin;
    // brute force: the split-K partials and arrival counters are shared by every call of this thread; a call on another
    // stream waits for the previous one (bfDone) instead of racing it
    cudaEvent_t bfDone = nullptr;
    cudaStream_t bfStream = nullptr;
    // sharded all-vs-all: communication stream, gathered database, column maps, events (one per chunk + "previous call done")
    cudaStream_t commStream = nullptr;
    DevBuf gathered, pad, colMaps;
    std::vector<cudaEvent_t> apEv;
    cudaEvent_t apDone = nullptr;
    ~Workspace() {
        // device memory is reclaimed at process exit; freeing here could run after the CUDA runtime is torn down
    }
};

thread_local Workspace tls[8];


int workspace(Workspace** out) {
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        set_error("no CUDA device: %s (this library has no CPU path)", e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        return ORBX_E_CUDA;
    }
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 8) {
        set_error("device ordinal %d not supported by the matcher workspace", dev);
        return ORBX_E_INVALID;
    }
    Workspace& w = tls[dev];
    if (!w.stream) {
        ORBX_CUDA(cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking));
        w.device = dev;
    }
    *out = &w;
    return ORBX_OK;
}

int upload(DevBuf& buf, const void* src, size_t bytes, cudaStream_t s) {
    int rc = buf.ensure(bytes ? bytes : 16);
    if (rc) return rc;
    if (bytes) ORBX_CUDA(cudaMemcpyAsync(buf.p, src, bytes, cudaMemcpyHostToDevice, s));
    return ORBX_OK;
}

}  // namespace

// A frame whose undistorted keypoints, descriptors, right coordinates and 64 x 48 grid stay on the device between matcher calls
struct orbx_frame {
    int device = 0, n = 0;
    bool hasRight = false;
    float bounds[4] = {0, 0, 0, 0};
    DevBuf kps, desc, uright, grid;  // grid: cellOf[n] | sorted[n] | cellStart[64 * 48 + 1]
    int* sorted() const { return grid.as<int>() + n; }
    int* cellStart() const { return grid.as<int>() + 2 * (size_t)n; }
};

extern "C" {

int orbx_frame_create(const orbx_keypoint* kps_un, int n, const uint8_t* desc, const float* u_right, const float bounds[4], orbx_frame** out) {
    if (!out || n < 0 || n > 65535 || !bounds || !(bounds[1] > bounds[0]) || !(bounds[3] > bounds[2]) || (n > 0 && (!kps_un || !desc))) {
        set_error("orbx_frame_create: bad argument (n <= 65535, bounds = {minX, maxX, minY, maxY})");
        return ORBX_E_INVALID;
    }
    *out = nullptr;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    orbx_frame* f = new orbx_frame();
    f->device = w->device;
    f->n = n;
    f->hasRight = u_right != nullptr;
    memcpy(f->bounds, bounds, sizeof f->bounds);
    cudaStream_t s = w->stream;
    const size_t nn = (size_t)std::max(n, 1);
    if ((rc = upload(f->kps, kps_un, (size_t)n * sizeof(orbx_keypoint), s)) || (rc = upload(f->desc, desc, (size_t)n * 32, s)) ||
        (rc = upload(f->uright, u_right, u_right ? (size_t)n * 4 : 0, s)) || (rc = f->grid.ensure((2 * nn + 64 * 48 + 1) * 4)) ||
        (rc = launch_frame_grid(f->kps.as<orbx_keypoint>(), n, bounds, f->grid.as<int>(), f->sorted(), f->cellStart(), s))) {
        orbx_frame_destroy(f);
        return rc;
    }
    cudaError_t e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) {
        set_error("orbx_frame_create: %s", cudaGetErrorString(e));
        orbx_frame_destroy(f);
        return ORBX_E_CUDA;
    }
    *out = f;
    return ORBX_OK;
}

void orbx_frame_destroy(orbx_frame* f) {
    if (!f) return;
    cudaSetDevice(f->device);
    f->kps.release(); f->desc.release(); f->uright.release(); f->grid.release();
    delete f;
}

int orbx_frame_size(const orbx_frame* f) { return f ? f->n : 0; }

int orbx_frame_grid_host(const orbx_frame* f, int* cell_start, int* sorted) {
    if (!f || !cell_start || (f->n > 0 && !sorted)) { set_error("orbx_frame_grid_host: bad argument"); return ORBX_E_INVALID; }
    ORBX_CUDA(cudaSetDevice(f->device));
    ORBX_CUDA(cudaMemcpy(cell_start, f->cellStart(), (64 * 48 + 1) * 4, cudaMemcpyDeviceToHost));
    if (f->n > 0) ORBX_CUDA(cudaMemcpy(sorted, f->sorted(), (size_t)f->n * 4, cudaMemcpyDeviceToHost));
    return ORBX_OK;
}

int orbx_hamming_distance(const uint8_t* a, const uint8_t* b, int n, int* dist) {
    if (n < 0 || (n > 0 && (!a || !b || !dist))) { set_error("orbx_hamming_distance: bad argument"); return ORBX_E_INVALID; }
    if (n == 0) return ORBX_OK;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = w->stream;
    if ((rc = upload(w->a, a, (size_t)n * 32, s)) || (rc = upload(w->b, b, (size_t)n * 32, s)) || (rc = w->out0.ensure((size_t)n * 4)))
        return rc;
    if ((rc = launch_hamming_pairs(w->a.as<uint8_t>(), w->b.as<uint8_t>(), n, w->out0.as<int>(), s))) return rc;
    ORBX_CUDA(cudaMemcpyAsync(dist, w->out0.p, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_match_bruteforce_device(const uint8_t* dq, int nq, const uint8_t* dd, int nd, int th, float ratio, int* idx, int* d1,
                                 int* d2, int* match, void* stream) {
    if (nq < 0 || nd < 0 || nd > 0x3FFFFE || (nq > 0 && !dq) || (nd > 0 && !dd)) {
        set_error("orbx_match_bruteforce_device: bad argument (nd must be < 4194302)");
        return ORBX_E_INVALID;
    }
    if (nq == 0) return ORBX_OK;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    if ((rc = w->partial.ensure(bruteforce_scratch_bytes(nq, nd)))) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : w->stream;
    if (!w->bfDone) ORBX_CUDA(cudaEventCreateWithFlags(&w->bfDone, cudaEventDisableTiming));
    if (w->bfStream && w->bfStream != s) ORBX_CUDA(cudaStreamWaitEvent(s, w->bfDone, 0));  // the scratch is still the previous call's
    if (bruteforce_counter_bytes(nq) > w->counters.bytes) {  // per-query-block arrival counters: zero once, the kernel leaves them zero
        ORBX_CUDA(cudaStreamSynchronize(s));                 // (the old, smaller array is freed below)
        if ((rc = w->counters.ensure(std::max<size_t>(bruteforce_counter_bytes(nq), 4096)))) return rc;
        ORBX_CUDA(cudaMemsetAsync(w->counters.p, 0, w->counters.bytes, s));  // on the launch stream: ordered before the kernel
    }
    rc = launch_bruteforce(dq, nq, dd, nd, th, ratio, idx, d1, d2, match, w->partial.as<unsigned long long>(),
                           w->counters.as<unsigned int>(), s);
    if (rc) return rc;
    ORBX_CUDA(cudaEventRecord(w->bfDone, s));
    w->bfStream = s;
    return ORBX_OK;
}

int orbx_match_bruteforce(const uint8_t* q, int nq, const uint8_t* d, int nd, int th, float ratio, int* idx, int* d1, int* d2,
                          int* match) {
    if (nq < 0 || nd < 0 || (nq > 0 && !q) || (nd > 0 && !d)) { set_error("orbx_match_bruteforce: bad argument"); return ORBX_E_INVALID; }
    if (nq == 0) return ORBX_OK;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = w->stream;
    const size_t ob = (size_t)nq * 4;
    if ((rc = upload(w->a, q, (size_t)nq * 32, s)) || (rc = upload(w->b, d, (size_t)nd * 32, s)) || (rc = w->out0.ensure(ob)) ||
        (rc = w->out1.ensure(ob)) || (rc = w->out2.ensure(ob)) || (rc = w->out3.ensure(ob)))
        return rc;
    if ((rc = orbx_match_bruteforce_device(w->a.as<uint8_t>(), nq, w->b.as<uint8_t>(), nd, th, ratio, w->out0.as<int>(),
                                           w->out1.as<int>(), w->out2.as<int>(), w->out3.as<int>(), s)))
        return rc;
    if (idx) ORBX_CUDA(cudaMemcpyAsync(idx, w->out0.p, ob, cudaMemcpyDeviceToHost, s));
    if (d1) ORBX_CUDA(cudaMemcpyAsync(d1, w->out1.p, ob, cudaMemcpyDeviceToHost, s));
    if (d2) ORBX_CUDA(cudaMemcpyAsync(d2, w->out2.p, ob, cudaMemcpyDeviceToHost, s));
    if (match) ORBX_CUDA(cudaMemcpyAsync(match, w->out3.p, ob, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_match_allpairs_device(const uint8_t* dq, int kq, const uint8_t* ddb, int kd, int n, int th, float ratio, int* dscore,
                               void* stream) {
    if (kq < 0 || kd < 0 || n < 1 || n > 0x3FFFFE || !dscore || (kq > 0 && !dq) || (kd > 0 && !ddb)) {
        set_error("orbx_match_allpairs_device: bad argument");
        return ORBX_E_INVALID;
    }
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    return launch_allpairs(dq, kq, ddb, kd, n, th, ratio, dscore, kd, nullptr, stream ? (cudaStream_t)stream : w->stream);
}

// ---- all-vs-all sharded by query block, database all-gathered over NCCL (BASELINE config 5) -----------------
// NCCL is reached through dlsym: liborbx.so has no link-time dependency on it, and the communicator handed in belongs to
// the NCCL the caller's process already holds (ORBX_NCCL_LIB names another one).
namespace {
typedef int (*NcclAllGatherFn)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef const char* (*NcclErrFn)(int);
struct NcclApi {
    NcclAllGatherFn allGather = nullptr;
    NcclErrFn errString = nullptr;
};
const NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* h = nullptr;
        if (const char* e = getenv("ORBX_NCCL_LIB")) h = dlopen(e, RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // the copy the process has loaded already (e.g. torch's)
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW);
        if (h) {
            api.allGather = (NcclAllGatherFn)dlsym(h, "ncclAllGather");
            api.errString = (NcclErrFn)dlsym(h, "ncclGetErrorString");
        }
    }
    return api.allGather ? &api : nullptr;
}
void block_range(int n, int rank, int world, int* b, int* e) {  // contiguous blocks, sizes differ by at most one, earlier ranks larger
    const int base = n / world, rem = n % world;
    *b = rank * base + std::min(rank, rem);
    *e = *b + base + (rank < rem ? 1 : 0);
}
}  // namespace

int orbx_allpairs_sharded(void* nccl_comm, int rank, int world, const uint8_t* d_local, int k_total, int n, int th, float ratio,
                          int* d_scores, int chunks, void* stream, unsigned long long* gathered_bytes) {
    if (world < 1 || rank < 0 || rank >= world || k_total < 0 || n < 1 || n > 0x3FFFFE) {
        set_error("orbx_allpairs_sharded: bad argument");
        return ORBX_E_INVALID;
    }
    if (gathered_bytes) *gathered_bytes = 0;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : w->stream;
    int begin, end;
    block_range(k_total, rank, world, &begin, &end);
    const int kLocal = end - begin;
    if (kLocal > 0 && (!d_local || !d_scores)) {  // a rank with an empty block still takes part in the gather
        set_error("orbx_allpairs_sharded: this rank owns %d keyframes but d_local / d_scores is null", kLocal);
        return ORBX_E_INVALID;
    }
    // the rank's own block needs no communication: it is scored first, while the first chunk of the gather is in flight
    if (kLocal > 0 && (rc = launch_allpairs(d_local, kLocal, d_local, kLocal, n, th, ratio, d_scores + begin, k_total, nullptr, s))) return rc;
    if (world == 1) return ORBX_OK;
    const NcclApi* nccl = nccl_api();
    if (!nccl || !nccl_comm) {
        set_error("orbx_allpairs_sharded: %s", nccl ? "null NCCL communicator" : "libnccl.so.2 could not be loaded (set ORBX_NCCL_LIB)");
        return ORBX_E_INVALID;
    }
    int b0, kMax;
    block_range(k_total, 0, world, &b0, &kMax);  // rank 0 owns the largest block
    chunks = std::max(1, std::min(chunks > 0 ? chunks : 4, std::max(kMax, 1)));
    const int step = (kMax + chunks - 1) / chunks;
    const size_t kfBytes = (size_t)n * 32;
    if (!w->commStream) ORBX_CUDA(cudaStreamCreateWithFlags(&w->commStream, cudaStreamNonBlocking));
    if (!w->apDone) ORBX_CUDA(cudaEventCreateWithFlags(&w->apDone, cudaEventDisableTiming));
    while ((int)w->apEv.size() < chunks + 1) {
        cudaEvent_t e;
        ORBX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        w->apEv.push_back(e);
    }
    // gathered database: chunk c holds [world][width_c] keyframes; every chunk is padded to the same width on every rank
    if ((size_t)world * kMax * kfBytes > w->gathered.bytes || (size_t)step * kfBytes > w->pad.bytes ||
        (size_t)world * kMax * sizeof(int) > w->colMaps.bytes) {
        ORBX_CUDA(cudaStreamSynchronize(w->commStream));
        ORBX_CUDA(cudaStreamSynchronize(s));
        if ((rc = w->gathered.ensure((size_t)world * kMax * kfBytes)) || (rc = w->pad.ensure((size_t)step * kfBytes)) ||
            (rc = w->colMaps.ensure((size_t)world * kMax * sizeof(int))))
            return rc;
    }
    // column maps of every chunk: gathered keyframe (src, i) of chunk [c0, c1) is global keyframe begin(src) + c0 + i; the
    // rank's own block (scored above) and the padding are skipped
    std::vector<int> maps((size_t)world * kMax, -1);
    size_t mapOff = 0;
    for (int c0 = 0; c0 < kMax; c0 += step) {
        const int width = std::min(step, kMax - c0);
        for (int src = 0; src < world; ++src) {
            int sb, se;
            block_range(k_total, src, world, &sb, &se);
            for (int i = 0; i < width; ++i)
                maps[mapOff + (size_t)src * width + i] = (src != rank && sb + c0 + i < se) ? sb + c0 + i : -1;
        }
        mapOff += (size_t)world * width;
    }
    ORBX_CUDA(cudaMemcpyAsync(w->colMaps.p, maps.data(), maps.size() * sizeof(int), cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaStreamSynchronize(s));  // `maps` is pageable host memory of this call (a few KB; the scoring takes seconds)
    // the previous call's kernels may still read the gathered database; d_local must be complete before it is sent
    ORBX_CUDA(cudaEventRecord(w->apEv[chunks], s));
    ORBX_CUDA(cudaStreamWaitEvent(w->commStream, w->apEv[chunks], 0));
    ORBX_CUDA(cudaStreamWaitEvent(w->commStream, w->apDone, 0));
    uint8_t* recv = w->gathered.as<uint8_t>();
    size_t recvOff = 0;
    int ci = 0;
    for (int c0 = 0; c0 < kMax; c0 += step, ++ci) {  // every chunk's all-gather is issued up front; they complete in order
        const int width = std::min(step, kMax - c0);
        const uint8_t* send = d_local + (size_t)c0 * kfBytes;
        if (c0 + width > kLocal) {  // this rank's block is shorter than the chunk: send a zero-padded copy
            ORBX_CUDA(cudaMemsetAsync(w->pad.p, 0, (size_t)width * kfBytes, w->commStream));
            if (kLocal > c0)
                ORBX_CUDA(cudaMemcpyAsync(w->pad.p, d_local + (size_t)c0 * kfBytes, (size_t)(kLocal - c0) * kfBytes, cudaMemcpyDeviceToDevice,
                                          w->commStream));
            send = w->pad.as<uint8_t>();
        }
        const int nr = nccl->allGather(send, recv + recvOff, (size_t)width * kfBytes, 1 /* ncclUint8 */, nccl_comm, w->commStream);
        if (nr != 0) {
            set_error("ncclAllGather failed: %s", nccl->errString ? nccl->errString(nr) : "?");
            return ORBX_E_CUDA;
        }
        ORBX_CUDA(cudaEventRecord(w->apEv[ci], w->commStream));
        recvOff += (size_t)world * width * kfBytes;
        if (gathered_bytes) *gathered_bytes += (unsigned long long)world * width * kfBytes;
    }
    recvOff = 0;
    mapOff = 0;
    ci = 0;
    for (int c0 = 0; c0 < kMax; c0 += step, ++ci) {  // one scoring launch per landed chunk over the keyframes of every source rank
        const int width = std::min(step, kMax - c0);
        ORBX_CUDA(cudaStreamWaitEvent(s, w->apEv[ci], 0));
        if (kLocal > 0 && (rc = launch_allpairs(d_local, kLocal, recv + recvOff, world * width, n, th, ratio, d_scores, k_total,
                                                w->colMaps.as<int>() + mapOff, s)))
            return rc;
        recvOff += (size_t)world * width * kfBytes;
        mapOff += (size_t)world * width;
    }
    ORBX_CUDA(cudaEventRecord(w->apDone, s));
    return ORBX_OK;
}

int orbx_search_for_initialization(const orbx_keypoint* k1, int n1, const uint8_t* desc1, const orbx_keypoint* k2, int n2,
                                   const uint8_t* desc2, const float bounds[4], float* prev_xy, int* matches12, int window,
                                   float nnratio, int check_ori, int* nmatches) {
    if (n1 < 0 || n2 < 0 || !bounds || !nmatches || (n1 > 0 && (!k1 || !desc1 || !prev_xy || !matches12)) ||
        (n2 > 0 && (!k2 || !desc2)) || n2 > 65535 || !(bounds[1] > bounds[0]) || !(bounds[3] > bounds[2])) {
        set_error("orbx_search_for_initialization: bad argument (n2 <= 65535, bounds = {minX, maxX, minY, maxY})");
        return ORBX_E_INVALID;
    }
    *nmatches = 0;
    if (n1 == 0) return ORBX_OK;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = w->stream;
    // candidate capacity per frame-1 keypoint: at most every octave-0 keypoint of frame 2
    int n2lvl0 = 0;
    for (int i = 0; i < n2; ++i) n2lvl0 += k2[i].octave == 0 ? 1 : 0;
    const int candCap = std::max(n2lvl0, 1);
    // one packed upload [k1 | k2 | desc1 | desc2] through pinned memory, and one copy back of [matches12 | overflow flag | nmatches |
    // prev_xy] (device layout: out0 = matches12 | status | prev_xy) — pageable cudaMemcpyAsync calls cost ~10 us each
    const size_t kb = sizeof(orbx_keypoint);
    const size_t oK1 = 0, oK2 = oK1 + (size_t)n1 * kb, oD1 = (oK2 + (size_t)n2 * kb + 15) / 16 * 16, oD2 = oD1 + (size_t)n1 * 32,
                 total = oD2 + (size_t)n2 * 32;
    const size_t rSt = ((size_t)n1 * 4 + 15) / 16 * 16, rXY = rSt + 16, rBytes = rXY + (size_t)n1 * 8;  // offsets inside out0 / the copy back
    const size_t oR = (total + 15) / 16 * 16;
    if ((rc = w->pin.ensure(oR + rBytes)) || (rc = w->a.ensure(std::max<size_t>(total, 16))) || (rc = w->out0.ensure(rBytes + 16)) ||
        (rc = w->f.ensure(((size_t)2 * std::max(n2, 1) + 64 * 48 + 1) * 4)) || (rc = w->g.ensure((size_t)n1 * candCap * 4)) ||
        (rc = w->h.ensure(resolve_aux_bytes(n1))))
        return rc;
    uint8_t* hp = w->pin.as<uint8_t>();
    memcpy(hp + oK1, k1, (size_t)n1 * kb);
    if (n2) memcpy(hp + oK2, k2, (size_t)n2 * kb);
    memcpy(hp + oD1, desc1, (size_t)n1 * 32);
    if (n2) memcpy(hp + oD2, desc2, (size_t)n2 * 32);
    memset(hp + oR + rSt, 0, 16);
    memcpy(hp + oR + rXY, prev_xy, (size_t)n1 * 8);
    ORBX_CUDA(cudaMemcpyAsync(w->a.p, hp, total, cudaMemcpyHostToDevice, s));
    uint8_t* dOut = w->out0.as<uint8_t>();
    ORBX_CUDA(cudaMemcpyAsync(dOut + rSt, hp + oR + rSt, 16 + (size_t)n1 * 8, cudaMemcpyHostToDevice, s));  // cleared status + vbPrevMatched
    const uint8_t* dp = w->a.as<uint8_t>();
    int* cellOf = w->f.as<int>();
    int* sorted = cellOf + std::max(n2, 1);
    int* cellStart = sorted + std::max(n2, 1);
    int* dSt = reinterpret_cast<int*>(dOut + rSt);
    if ((rc = launch_search_for_initialization(reinterpret_cast<const orbx_keypoint*>(dp + oK1), n1, dp + oD1,
                                               reinterpret_cast<const orbx_keypoint*>(dp + oK2), n2, dp + oD2, bounds,
                                               reinterpret_cast<float*>(dOut + rXY), reinterpret_cast<int*>(dOut), window, nnratio, check_ori,
                                               cellOf, sorted, cellStart, w->g.as<uint32_t>(), candCap, w->h.as<int>(), dSt + 1, dSt, s)))
        return rc;
    ORBX_CUDA(cudaMemcpyAsync(hp + oR, dOut, rBytes, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    memcpy(matches12, hp + oR, (size_t)n1 * 4);
    memcpy(prev_xy, hp + oR + rXY, (size_t)n1 * 8);
    int st[2];
    memcpy(st, hp + oR + rSt, 8);
    if (st[0] != 0) {
        set_error("orbx_search_for_initialization: candidate list overflow");
        return ORBX_E_OVERFLOW;
    }
    *nmatches = st[1];
    return ORBX_OK;
}

int orbx_search_by_projection(int n1, const float* x3dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mp_desc1,
                              const int* octave1, const float* angle1, const orbx_keypoint* k2, int n2, const uint8_t* desc2,
                              const float* u_right2, const uint8_t* occupied2, const float cam[4], float mbf,
                              const float bounds[4], const float* scale_factors, int nlevels, float th, int direction,
                              int check_ori, int* match21, int* nmatches) {
    return orbx_search_by_projection_ex(n1, x3dc, valid1, obs1, mp_desc1, octave1, angle1, k2, n2, desc2, u_right2, occupied2, cam, mbf,
                                        bounds, scale_factors, nlevels, th, direction, check_ori, 100 /* TH_HIGH */, 0, match21,
                                        nmatches);
}

} // extern "C"
static int sbp_impl(const orbx_frame* fr, int n1, const float* x3dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mp_desc1,
                    const int* octave1, const float* angle1, const orbx_keypoint* k2, int n2, const uint8_t* desc2,
                    const float* u_right2, const uint8_t* occupied2, const float cam[4], float mbf,
                    const float bounds[4], const float* scale_factors, int nlevels, float th, int direction,
                    int check_ori, int orb_dist, int flags, int* match21, int* nmatches) {
    if (fr) {  // the frame's keypoints, descriptors, right coordinates and grid are on the device already
        n2 = fr->n;
        bounds = fr->bounds;
    }
    if (n1 < 0 || n2 < 0 || !bounds || !cam || !nmatches || !scale_factors || nlevels < 1 || nlevels > ORBX_MAX_LEVELS ||
        direction < 0 || direction > 2 || (n1 > 0 && (!x3dc || !valid1 || !obs1 || !mp_desc1 || !octave1 || !angle1)) ||
        (n2 > 0 && ((!fr && (!k2 || !desc2)) || !occupied2 || !match21)) || n2 > 65535 || !(bounds[1] > bounds[0]) || !(bounds[3] > bounds[2])) {
        set_error("orbx_search_by_projection: bad argument (n2 <= 65535, direction 0..2, bounds = {minX, maxX, minY, maxY})");
        return ORBX_E_INVALID;
    }
    for (int i = 0; i < n1; ++i)
        if (valid1[i] && (octave1[i] < 0 || octave1[i] >= nlevels)) {
            set_error("orbx_search_by_projection: octave %d of point %d outside the %d scale factors", octave1[i], i, nlevels);
            return ORBX_E_INVALID;
        }
    *nmatches = 0;
    for (int i = 0; i < n2; ++i) match21[i] = -1;
    if (n1 == 0 || n2 == 0) return ORBX_OK;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = w->stream;
    const int candCap = n2;
    // one packed upload: [x3dc | angle1 | octave1 | scale factors | u_right2 | mp_desc1 | valid1 | obs1 | occupied2]
    const size_t oX = 0, oA = oX + (size_t)n1 * 12, oO = oA + (size_t)n1 * 4, oS = oO + (size_t)n1 * 4, oU = oS + ORBX_MAX_LEVELS * 4,
                 oD = (oU + (size_t)n2 * 4 + 15) / 16 * 16, oV = oD + (size_t)n1 * 32, oB = oV + n1, oC = oB + n1, total = oC + n2;
    // the pinned buffer also takes the results back: [match21 | overflow flag | nmatches] in ONE copy behind the upload area
    const size_t oR = (total + 15) / 16 * 16, rBytes = (size_t)n2 * 4 + 8;
    if ((rc = w->pin.ensure(oR + rBytes)) || (rc = w->a.ensure(total))) return rc;
    uint8_t* hp = w->pin.as<uint8_t>();
    memcpy(hp + oX, x3dc, (size_t)n1 * 12);
    memcpy(hp + oA, angle1, (size_t)n1 * 4);
    memcpy(hp + oO, octave1, (size_t)n1 * 4);
    memset(hp + oS, 0, ORBX_MAX_LEVELS * 4);
    memcpy(hp + oS, scale_factors, (size_t)nlevels * 4);
    if (u_right2 && !fr) memcpy(hp + oU, u_right2, (size_t)n2 * 4);
    memcpy(hp + oD, mp_desc1, (size_t)n1 * 32);
    memcpy(hp + oV, valid1, n1);
    memcpy(hp + oB, obs1, n1);
    memcpy(hp + oC, occupied2, n2);
    ORBX_CUDA(cudaMemcpyAsync(w->a.p, hp, total, cudaMemcpyHostToDevice, s));
    if ((!fr && ((rc = upload(w->c, k2, (size_t)n2 * sizeof(orbx_keypoint), s)) || (rc = upload(w->d, desc2, (size_t)n2 * 32, s)))) ||
        (rc = w->out0.ensure((size_t)n2 * 4 + 16)) || (rc = w->f.ensure(((size_t)2 * n2 + 64 * 48 + 1) * 4)) ||
        (rc = w->g.ensure((size_t)n1 * candCap * 4)) || (rc = w->h.ensure(resolve_aux_bytes(n1))) || (rc = w->status.ensure(16)))
        return rc;
    int* dSt = w->out0.as<int>() + n2;  // status words right behind match21
    ORBX_CUDA(cudaMemsetAsync(dSt, 0, 8, s));
    const uint8_t* dp = w->a.as<uint8_t>();
    int* cellOf = fr ? nullptr : w->f.as<int>();
    int* sorted = fr ? fr->sorted() : cellOf + n2;
    int* cellStart = fr ? fr->cellStart() : sorted + n2;
    int* dN = dSt + 1;
    const float* dUr = fr ? (fr->hasRight ? fr->uright.as<float>() : nullptr) : (u_right2 ? reinterpret_cast<const float*>(dp + oU) : nullptr);
    if ((rc = launch_search_by_projection(n1, reinterpret_cast<const float*>(dp + oX), dp + oV, dp + oB, dp + oD,
                                          reinterpret_cast<const int*>(dp + oO), reinterpret_cast<const float*>(dp + oA),
                                          fr ? fr->kps.as<orbx_keypoint>() : w->c.as<orbx_keypoint>(), n2,
                                          fr ? fr->desc.as<uint8_t>() : w->d.as<uint8_t>(), dUr, dp + oC,
                                          reinterpret_cast<const float*>(dp + oS), cam, mbf, bounds, th, direction, check_ori, orb_dist,
                                          flags & ORBX_SBP_NO_DEPTH_TEST ? 1 : 0, cellOf, sorted, cellStart, w->g.as<uint32_t>(), candCap, w->h.as<int>(), w->out0.as<int>(), dN,
                                          dSt, s)))
        return rc;
    ORBX_CUDA(cudaMemcpyAsync(hp + oR, w->out0.p, rBytes, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    memcpy(match21, hp + oR, (size_t)n2 * 4);
    int st[2];
    memcpy(st, hp + oR + (size_t)n2 * 4, 8);
    if (st[0] != 0) {
        set_error("orbx_search_by_projection: candidate list overflow");
        return ORBX_E_OVERFLOW;
    }
    *nmatches = st[1];
    return ORBX_OK;
}

extern "C" {
int orbx_search_by_projection_ex(int n1, const float* x3dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mp_desc1,
                                 const int* octave1, const float* angle1, const orbx_keypoint* k2, int n2, const uint8_t* desc2,
                                 const float* u_right2, const uint8_t* occupied2, const float cam[4], float mbf,
                                 const float bounds[4], const float* scale_factors, int nlevels, float th, int direction,
                                 int check_ori, int orb_dist, int flags, int* match21, int* nmatches) {
    return sbp_impl(nullptr, n1, x3dc, valid1, obs1, mp_desc1, octave1, angle1, k2, n2, desc2, u_right2, occupied2, cam, mbf, bounds,
                    scale_factors, nlevels, th, direction, check_ori, orb_dist, flags, match21, nmatches);
}

int orbx_search_by_projection_frame(int n1, const float* x3dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mp_desc1,
                                    const int* octave1, const float* angle1, const orbx_frame* f2, const uint8_t* occupied2,
                                    const float cam[4], float mbf, const float* scale_factors, int nlevels, float th, int direction,
                                    int check_ori, int orb_dist, int flags, int* match21, int* nmatches) {
    if (!f2) { set_error("orbx_search_by_projection_frame: null frame"); return ORBX_E_INVALID; }
    return sbp_impl(f2, n1, x3dc, valid1, obs1, mp_desc1, octave1, angle1, nullptr, 0, nullptr, nullptr, occupied2, cam, mbf, nullptr,
                    scale_factors, nlevels, th, direction, check_ori, orb_dist, flags, match21, nmatches);
}

} // extern "C"
static int sbpm_impl(const orbx_frame* fr, int n1, const float* proj, const float* view_cos, const int* level1, const uint8_t* valid1,
                     const uint8_t* obs1, const uint8_t* mp_desc1, const orbx_keypoint* k2, int n2,
                     const uint8_t* desc2, const float* u_right2, const uint8_t* occupied2, const float bounds[4],
                     const float* scale_factors, int nlevels, float th, float nnratio, int* match21, int* nmatches) {
    if (fr) {
        n2 = fr->n;
        bounds = fr->bounds;
    }
    if (n1 < 0 || n2 < 0 || !bounds || !nmatches || !scale_factors || nlevels < 1 || nlevels > ORBX_MAX_LEVELS ||
        (n1 > 0 && (!proj || !view_cos || !level1 || !valid1 || !obs1 || !mp_desc1)) ||
        (n2 > 0 && ((!fr && (!k2 || !desc2)) || !occupied2 || !match21)) || n2 > 65535 || !(bounds[1] > bounds[0]) || !(bounds[3] > bounds[2])) {
        set_error("orbx_search_by_projection_map: bad argument (n2 <= 65535, bounds = {minX, maxX, minY, maxY})");
        return ORBX_E_INVALID;
    }
    for (int i = 0; i < n1; ++i)
        if (valid1[i] && (level1[i] < 0 || level1[i] >= nlevels)) {
            set_error("orbx_search_by_projection_map: level %d of point %d outside the %d scale factors", level1[i], i, nlevels);
            return ORBX_E_INVALID;
        }
    *nmatches = 0;
    for (int i = 0; i < n2; ++i) match21[i] = -1;
    if (n1 == 0 || n2 == 0) return ORBX_OK;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = w->stream;
    const int candCap = n2;
    // one packed upload: [proj | view_cos | level1 | scale factors | u_right2 | mp_desc1 | valid1 | obs1 | occupied2]
    const size_t oX = 0, oA = oX + (size_t)n1 * 12, oO = oA + (size_t)n1 * 4, oS = oO + (size_t)n1 * 4, oU = oS + ORBX_MAX_LEVELS * 4,
                 oD = (oU + (size_t)n2 * 4 + 15) / 16 * 16, oV = oD + (size_t)n1 * 32, oB = oV + n1, oC = oB + n1, total = oC + n2;
    // the pinned buffer also takes the results back: [match21 | overflow flag | nmatches] in ONE copy behind the upload area
    const size_t oR = (total + 15) / 16 * 16, rBytes = (size_t)n2 * 4 + 8;
    if ((rc = w->pin.ensure(oR + rBytes)) || (rc = w->a.ensure(total))) return rc;
    uint8_t* hp = w->pin.as<uint8_t>();
    memcpy(hp + oX, proj, (size_t)n1 * 12);
    memcpy(hp + oA, view_cos, (size_t)n1 * 4);
    memcpy(hp + oO, level1, (size_t)n1 * 4);
    memset(hp + oS, 0, ORBX_MAX_LEVELS * 4);
    memcpy(hp + oS, scale_factors, (size_t)nlevels * 4);
    if (u_right2 && !fr) memcpy(hp + oU, u_right2, (size_t)n2 * 4);
    memcpy(hp + oD, mp_desc1, (size_t)n1 * 32);
    memcpy(hp + oV, valid1, n1);
    memcpy(hp + oB, obs1, n1);
    memcpy(hp + oC, occupied2, n2);
    ORBX_CUDA(cudaMemcpyAsync(w->a.p, hp, total, cudaMemcpyHostToDevice, s));
    if ((!fr && ((rc = upload(w->c, k2, (size_t)n2 * sizeof(orbx_keypoint), s)) || (rc = upload(w->d, desc2, (size_t)n2 * 32, s)))) ||
        (rc = w->out0.ensure((size_t)n2 * 4 + 16)) || (rc = w->f.ensure(((size_t)2 * n2 + 64 * 48 + 1) * 4)) ||
        (rc = w->g.ensure((size_t)n1 * candCap * 4)) || (rc = w->h.ensure(resolve_aux_bytes(n1))) || (rc = w->status.ensure(16)))
        return rc;
    int* dSt = w->out0.as<int>() + n2;  // status words right behind match21
    ORBX_CUDA(cudaMemsetAsync(dSt, 0, 8, s));
    const uint8_t* dp = w->a.as<uint8_t>();
    int* cellOf = fr ? nullptr : w->f.as<int>();
    int* sorted = fr ? fr->sorted() : cellOf + n2;
    int* cellStart = fr ? fr->cellStart() : sorted + n2;
    int* dN = dSt + 1;
    const float* dUr = fr ? (fr->hasRight ? fr->uright.as<float>() : nullptr) : (u_right2 ? reinterpret_cast<const float*>(dp + oU) : nullptr);
    if ((rc = launch_search_by_projection_map(n1, reinterpret_cast<const float*>(dp + oX), reinterpret_cast<const float*>(dp + oA),
                                              reinterpret_cast<const int*>(dp + oO), dp + oV, dp + oB, dp + oD,
                                              fr ? fr->kps.as<orbx_keypoint>() : w->c.as<orbx_keypoint>(), n2,
                                              fr ? fr->desc.as<uint8_t>() : w->d.as<uint8_t>(), dUr, dp + oC,
                                              reinterpret_cast<const float*>(dp + oS), bounds, th, nnratio, cellOf, sorted, cellStart,
                                              w->g.as<uint32_t>(), candCap, w->h.as<int>(), w->out0.as<int>(), dN,
                                              dSt, s)))
        return rc;
    ORBX_CUDA(cudaMemcpyAsync(hp + oR, w->out0.p, rBytes, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    memcpy(match21, hp + oR, (size_t)n2 * 4);
    int st[2];
    memcpy(st, hp + oR + (size_t)n2 * 4, 8);
    if (st[0] != 0) {
        set_error("orbx_search_by_projection_map: candidate list overflow");
        return ORBX_E_OVERFLOW;
    }
    *nmatches = st[1];
    return ORBX_OK;
}

extern "C" {
int orbx_search_by_projection_map(int n1, const float* proj, const float* view_cos, const int* level1, const uint8_t* valid1,
                                  const uint8_t* obs1, const uint8_t* mp_desc1, const orbx_keypoint* k2, int n2,
                                  const uint8_t* desc2, const float* u_right2, const uint8_t* occupied2, const float bounds[4],
                                  const float* scale_factors, int nlevels, float th, float nnratio, int* match21, int* nmatches) {
    return sbpm_impl(nullptr, n1, proj, view_cos, level1, valid1, obs1, mp_desc1, k2, n2, desc2, u_right2, occupied2, bounds, scale_factors,
                     nlevels, th, nnratio, match21, nmatches);
}

int orbx_search_by_projection_map_frame(int n1, const float* proj, const float* view_cos, const int* level1, const uint8_t* valid1,
                                        const uint8_t* obs1, const uint8_t* mp_desc1, const orbx_frame* f2, const uint8_t* occupied2,
                                        const float* scale_factors, int nlevels, float th, float nnratio, int* match21, int* nmatches) {
    if (!f2) { set_error("orbx_search_by_projection_map_frame: null frame"); return ORBX_E_INVALID; }
    return sbpm_impl(f2, n1, proj, view_cos, level1, valid1, obs1, mp_desc1, nullptr, 0, nullptr, nullptr, occupied2, nullptr, scale_factors,
                     nlevels, th, nnratio, match21, nmatches);
}

int orbx_search_by_projection_sim3(int n1, const float* uvr, const int* level1, const uint8_t* valid1, const uint8_t* mp_desc1,
                                   const orbx_keypoint* k2, int n2, const uint8_t* desc2, const uint8_t* occupied2,
                                   const float bounds[4], int th_dist, int* match21, int* nmatches) {
    if (n1 < 0 || n2 < 0 || !bounds || !nmatches || (n1 > 0 && (!uvr || !level1 || !valid1 || !mp_desc1)) ||
        (n2 > 0 && (!k2 || !desc2 || !occupied2 || !match21)) || n2 > 65535 || !(bounds[1] > bounds[0]) || !(bounds[3] > bounds[2])) {
        set_error("orbx_search_by_projection_sim3: bad argument (n2 <= 65535, bounds = {minX, maxX, minY, maxY})");
        return ORBX_E_INVALID;
    }
    *nmatches = 0;
    for (int i = 0; i < n2; ++i) match21[i] = -1;
    if (n1 == 0 || n2 == 0) return ORBX_OK;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = w->stream;
    const int candCap = n2;
    // one packed upload: [uvr | level1 | mp_desc1 | valid1 | ones | occupied2]
    const size_t oX = 0, oO = oX + (size_t)n1 * 12, oD = (oO + (size_t)n1 * 4 + 15) / 16 * 16, oV = oD + (size_t)n1 * 32, oB = oV + n1,
                 oC = oB + n1, total = oC + n2;
    // the pinned buffer also takes the results back: [match21 | overflow flag | nmatches] in ONE copy behind the upload area
    const size_t oR = (total + 15) / 16 * 16, rBytes = (size_t)n2 * 4 + 8;
    if ((rc = w->pin.ensure(oR + rBytes)) || (rc = w->a.ensure(total))) return rc;
    uint8_t* hp = w->pin.as<uint8_t>();
    memcpy(hp + oX, uvr, (size_t)n1 * 12);
    memcpy(hp + oO, level1, (size_t)n1 * 4);
    memcpy(hp + oD, mp_desc1, (size_t)n1 * 32);
    memcpy(hp + oV, valid1, n1);
    memset(hp + oB, 1, n1);
    memcpy(hp + oC, occupied2, n2);
    ORBX_CUDA(cudaMemcpyAsync(w->a.p, hp, total, cudaMemcpyHostToDevice, s));
    if ((rc = upload(w->c, k2, (size_t)n2 * sizeof(orbx_keypoint), s)) || (rc = upload(w->d, desc2, (size_t)n2 * 32, s)) ||
        (rc = w->out0.ensure((size_t)n2 * 4 + 16)) || (rc = w->f.ensure(((size_t)2 * n2 + 64 * 48 + 1) * 4)) ||
        (rc = w->g.ensure((size_t)n1 * candCap * 4)) || (rc = w->h.ensure(resolve_aux_bytes(n1))) || (rc = w->status.ensure(16)))
        return rc;
    int* dSt = w->out0.as<int>() + n2;  // status words right behind match21
    ORBX_CUDA(cudaMemsetAsync(dSt, 0, 8, s));
    const uint8_t* dp = w->a.as<uint8_t>();
    int* cellOf = w->f.as<int>();
    int* sorted = cellOf + n2;
    int* cellStart = sorted + n2;
    int* dN = dSt + 1;
    if ((rc = launch_search_by_projection_sim3(n1, reinterpret_cast<const float*>(dp + oX), reinterpret_cast<const int*>(dp + oO),
                                               dp + oV, dp + oB, dp + oD, w->c.as<orbx_keypoint>(), n2, w->d.as<uint8_t>(), dp + oC,
                                               bounds, th_dist, cellOf, sorted, cellStart, w->g.as<uint32_t>(), candCap,
                                               w->h.as<int>(), w->out0.as<int>(), dN, dSt, s)))
        return rc;
    ORBX_CUDA(cudaMemcpyAsync(hp + oR, w->out0.p, rBytes, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    memcpy(match21, hp + oR, (size_t)n2 * 4);
    int st[2];
    memcpy(st, hp + oR + (size_t)n2 * 4, 8);
    if (st[0] != 0) {
        set_error("orbx_search_by_projection_sim3: candidate list overflow");
        return ORBX_E_OVERFLOW;
    }
    *nmatches = st[1];
    return ORBX_OK;
}

int orbx_search_by_bow(int n_nodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                       const uint8_t* valid1, const uint8_t* desc1, const float* angle1, int n2, const uint8_t* valid2,
                       const uint8_t* desc2, const float* angle2, int th_dist, float nnratio, int check_ori, int* match12, int* match21,
                       int* nmatches) {
    if (n_nodes < 0 || n1 < 0 || n2 < 0 || n2 > 65535 || !nmatches || (n_nodes > 0 && (!start1 || !start2)) ||
        (n1 > 0 && (!valid1 || !desc1 || !angle1 || !match12)) || (n2 > 0 && (!valid2 || !desc2 || !angle2 || !match21))) {
        set_error("orbx_search_by_bow: bad argument (n2 <= 65535)");
        return ORBX_E_INVALID;
    }
    *nmatches = 0;
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    for (int i = 0; i < n2; ++i) match21[i] = -1;
    if (n_nodes == 0 || n1 == 0 || n2 == 0) return ORBX_OK;
    if (start1[0] != 0 || start2[0] != 0) {
        set_error("orbx_search_by_bow: start1[0] and start2[0] must be 0");
        return ORBX_E_INVALID;
    }
    for (int k = 0; k < n_nodes; ++k)
        if (start1[k + 1] < start1[k] || start2[k + 1] < start2[k] || start2[k + 1] - start2[k] > 65535) {
            set_error("orbx_search_by_bow: node %d has a malformed feature range", k);
            return ORBX_E_INVALID;
        }
    const int nItems = start1[n_nodes], nIdx2 = start2[n_nodes];
    if ((nItems > 0 && !idx1) || (nIdx2 > 0 && !idx2)) {
        set_error("orbx_search_by_bow: null feature index list");
        return ORBX_E_INVALID;
    }
    for (int i = 0; i < nItems; ++i)
        if (idx1[i] >= (uint32_t)n1) { set_error("orbx_search_by_bow: idx1[%d] out of range", i); return ORBX_E_INVALID; }
    for (int i = 0; i < nIdx2; ++i)
        if (idx2[i] >= (uint32_t)n2) { set_error("orbx_search_by_bow: idx2[%d] out of range", i); return ORBX_E_INVALID; }
    if (nItems == 0 || nIdx2 == 0) return ORBX_OK;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = w->stream;
    // one packed upload: [idx1 | nodeOf | candOff | start2 | idx2 | angle1 | angle2 | desc1 | desc2 | valid1 | valid2]
    const size_t oI1 = 0, oNode = oI1 + (size_t)nItems * 4, oOff = oNode + (size_t)nItems * 4, oS2 = oOff + (size_t)nItems * 4,
                 oI2 = oS2 + (size_t)(n_nodes + 1) * 4, oA1 = oI2 + (size_t)nIdx2 * 4, oA2 = oA1 + (size_t)n1 * 4,
                 oD1 = (oA2 + (size_t)n2 * 4 + 15) / 16 * 16, oD2 = oD1 + (size_t)n1 * 32, oV1 = oD2 + (size_t)n2 * 32, oV2 = oV1 + n1,
                 total = oV2 + n2;
    // ... and one copy back of [match12 | match21 | nmatches] behind it
    const size_t oR = (total + 15) / 16 * 16, rBytes = ((size_t)n1 + n2) * 4 + 4;
    if ((rc = w->pin.ensure(oR + rBytes))) return rc;
    uint8_t* hp = w->pin.as<uint8_t>();
    int* hNode = reinterpret_cast<int*>(hp + oNode);
    int* hOff = reinterpret_cast<int*>(hp + oOff);
    size_t nCand = 0;
    for (int k = 0; k < n_nodes; ++k) {
        const size_t len2 = (size_t)(start2[k + 1] - start2[k]);
        for (int i = start1[k]; i < start1[k + 1]; ++i) {
            hNode[i] = k;
            hOff[i] = (int)nCand;
            nCand += len2;
            if (nCand > ((size_t)1 << 30)) {
                set_error("orbx_search_by_bow: more than 2^30 candidate pairs");
                return ORBX_E_INVALID;
            }
        }
    }
    memcpy(hp + oI1, idx1, (size_t)nItems * 4);
    memcpy(hp + oS2, start2, (size_t)(n_nodes + 1) * 4);
    memcpy(hp + oI2, idx2, (size_t)nIdx2 * 4);
    memcpy(hp + oA1, angle1, (size_t)n1 * 4);
    memcpy(hp + oA2, angle2, (size_t)n2 * 4);
    memcpy(hp + oD1, desc1, (size_t)n1 * 32);
    memcpy(hp + oD2, desc2, (size_t)n2 * 32);
    memcpy(hp + oV1, valid1, n1);
    memcpy(hp + oV2, valid2, n2);
    if ((rc = w->a.ensure(total)) || (rc = w->g.ensure((nCand + 1) * 4)) || (rc = w->h.ensure(resolve_aux_bytes(nItems))) ||
        (rc = w->out0.ensure(rBytes + 16)))
        return rc;
    ORBX_CUDA(cudaMemcpyAsync(w->a.p, hp, total, cudaMemcpyHostToDevice, s));
    const uint8_t* dp = w->a.as<uint8_t>();
    if ((rc = launch_search_by_bow(nItems, reinterpret_cast<const int*>(dp + oNode), reinterpret_cast<const uint32_t*>(dp + oI1),
                                   reinterpret_cast<const int*>(dp + oS2), reinterpret_cast<const uint32_t*>(dp + oI2), n1, dp + oV1,
                                   dp + oD1, reinterpret_cast<const float*>(dp + oA1), n2, dp + oV2, dp + oD2,
                                   reinterpret_cast<const float*>(dp + oA2), reinterpret_cast<const int*>(dp + oOff), w->g.as<uint32_t>(),
                                   w->h.as<int>(), th_dist, nnratio, check_ori, w->out0.as<int>(), w->out0.as<int>() + n1,
                                   w->out0.as<int>() + n1 + n2, s)))
        return rc;
    ORBX_CUDA(cudaMemcpyAsync(hp + oR, w->out0.p, rBytes, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    memcpy(match12, hp + oR, (size_t)n1 * 4);
    memcpy(match21, hp + oR + (size_t)n1 * 4, (size_t)n2 * 4);
    memcpy(nmatches, hp + oR + ((size_t)n1 + n2) * 4, 4);
    return ORBX_OK;
}

int orbx_cvt_color_gray_device(const uint8_t* d_src, long long src_stride, int w, int h, int channels, int flags, uint8_t* d_dst,
                               long long dst_stride, void* stream) {
    if (!d_src || !d_dst || w <= 0 || h <= 0 || (channels != 3 && channels != 4) || src_stride < (long long)w * channels || dst_stride < w) {
        set_error("orbx_cvt_color_gray: bad argument (channels 3 or 4, strides in bytes)");
        return ORBX_E_INVALID;
    }
    return launch_gray(d_src, src_stride, w, h, channels, (flags & ORBX_GRAY_BGR) ? 1 : 0, (flags & ORBX_GRAY_CV2_COEFFS) ? 14 : 15, d_dst,
                       dst_stride, static_cast<cudaStream_t>(stream));
}

int orbx_cvt_color_gray(const uint8_t* src, long long src_stride, int w, int h, int channels, int flags, uint8_t* dst, long long dst_stride) {
    if (!src || !dst || w <= 0 || h <= 0 || (channels != 3 && channels != 4) || src_stride < (long long)w * channels || dst_stride < w) {
        set_error("orbx_cvt_color_gray: bad argument (channels 3 or 4, strides in bytes)");
        return ORBX_E_INVALID;
    }
    Workspace* wk;
    int rc = workspace(&wk);
    if (rc) return rc;
    cudaStream_t s = wk->stream;
    const size_t inBytes = (size_t)w * channels * h, outBytes = (size_t)w * h;
    if ((rc = wk->a.ensure(inBytes)) || (rc = wk->out0.ensure(outBytes))) return rc;
    ORBX_CUDA(cudaMemcpy2DAsync(wk->a.p, (size_t)w * channels, src, (size_t)src_stride, (size_t)w * channels, h, cudaMemcpyHostToDevice, s));
    if ((rc = launch_gray(wk->a.as<uint8_t>(), (long long)w * channels, w, h, channels, (flags & ORBX_GRAY_BGR) ? 1 : 0,
                          (flags & ORBX_GRAY_CV2_COEFFS) ? 14 : 15, wk->out0.as<uint8_t>(), w, s)))
        return rc;
    ORBX_CUDA(cudaMemcpy2DAsync(dst, (size_t)dst_stride, wk->out0.p, (size_t)w, (size_t)w, h, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_undistort_keypoints_device(const orbx_keypoint* d_in, int n, const float K[4], const float* dist, int ndist, orbx_keypoint* d_out,
                                    void* stream) {
    if (n < 0 || !K || ndist < 0 || ndist > 5 || (ndist > 0 && !dist) || (n > 0 && (!d_in || !d_out)) || !(K[0] != 0.f) || !(K[1] != 0.f)) {
        set_error("orbx_undistort_keypoints: bad argument (K = fx, fy, cx, cy; up to 5 distortion coefficients k1 k2 p1 p2 k3)");
        return ORBX_E_INVALID;
    }
    Workspace* wk;
    int rc = workspace(&wk);
    if (rc) return rc;
    return launch_undistort(d_in, n, K, dist, ndist, d_out, stream ? static_cast<cudaStream_t>(stream) : wk->stream);
}

int orbx_undistort_keypoints(const orbx_keypoint* in, int n, const float K[4], const float* dist, int ndist, orbx_keypoint* out) {
    if (n < 0 || !K || ndist < 0 || ndist > 5 || (ndist > 0 && !dist) || (n > 0 && (!in || !out)) || !(K[0] != 0.f) || !(K[1] != 0.f)) {
        set_error("orbx_undistort_keypoints: bad argument (K = fx, fy, cx, cy; up to 5 distortion coefficients k1 k2 p1 p2 k3)");
        return ORBX_E_INVALID;
    }
    if (n == 0) return ORBX_OK;
    if (ndist == 0 || dist[0] == 0.0f) {  // src/Frame.cc:405-409: no distortion, mvKeysUn = mvKeys
        if (out != in) memcpy(out, in, (size_t)n * sizeof(orbx_keypoint));
        return ORBX_OK;
    }
    Workspace* wk;
    int rc = workspace(&wk);
    if (rc) return rc;
    cudaStream_t s = wk->stream;
    const size_t bytes = (size_t)n * sizeof(orbx_keypoint);
    if ((rc = upload(wk->a, in, bytes, s)) || (rc = wk->b.ensure(bytes))) return rc;
    if ((rc = launch_undistort(wk->a.as<orbx_keypoint>(), n, K, dist, ndist, wk->b.as<orbx_keypoint>(), s))) return rc;
    ORBX_CUDA(cudaMemcpyAsync(out, wk->b.p, bytes, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_fuse_match(int n1, const float* uvr, const float* ur1, const int* level1, const uint8_t* valid1, const uint8_t* mp_desc1,
                    const orbx_keypoint* k2, int n2, const uint8_t* desc2, const float* u_right2, const float* inv_level_sigma2,
                    int nlevels, const float bounds[4], int* best_idx, int* best_dist) {
    if (n1 < 0 || n2 < 0 || !bounds || nlevels <= 0 || nlevels > 64 || (n1 > 0 && (!uvr || !level1 || !valid1 || !mp_desc1 || !best_idx)) ||
        (n2 > 0 && (!k2 || !desc2)) || (ur1 && !inv_level_sigma2) || !(bounds[1] > bounds[0]) || !(bounds[3] > bounds[2])) {
        set_error("orbx_fuse_match: bad argument (bounds = {minX, maxX, minY, maxY})");
        return ORBX_E_INVALID;
    }
    for (int i = 0; i < n1; ++i) {
        best_idx[i] = -1;
        if (best_dist) best_dist[i] = 256;
    }
    if (n1 == 0 || n2 == 0) return ORBX_OK;
    for (int j = 0; j < n2; ++j)
        if (k2[j].octave < 0 || k2[j].octave >= nlevels) { set_error("orbx_fuse_match: k2[%d].octave out of range", j); return ORBX_E_INVALID; }
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = w->stream;
    // one packed upload: [uvr | ur1 | level1 | uRight2 | invSigma2 | mp_desc1 | valid1]
    const size_t oX = 0, oU = oX + (size_t)n1 * 12, oL = oU + (size_t)n1 * 4, oR = oL + (size_t)n1 * 4, oS = oR + (size_t)n2 * 4,
                 oD = (oS + (size_t)nlevels * 4 + 15) / 16 * 16, oV = oD + (size_t)n1 * 32, total = oV + n1;
    if ((rc = w->pin.ensure(total)) || (rc = w->a.ensure(total))) return rc;
    uint8_t* hp = w->pin.as<uint8_t>();
    memcpy(hp + oX, uvr, (size_t)n1 * 12);
    if (ur1) memcpy(hp + oU, ur1, (size_t)n1 * 4);
    memcpy(hp + oL, level1, (size_t)n1 * 4);
    if (u_right2) memcpy(hp + oR, u_right2, (size_t)n2 * 4);
    if (inv_level_sigma2) memcpy(hp + oS, inv_level_sigma2, (size_t)nlevels * 4);
    memcpy(hp + oD, mp_desc1, (size_t)n1 * 32);
    memcpy(hp + oV, valid1, n1);
    ORBX_CUDA(cudaMemcpyAsync(w->a.p, hp, total, cudaMemcpyHostToDevice, s));
    if ((rc = upload(w->c, k2, (size_t)n2 * sizeof(orbx_keypoint), s)) || (rc = upload(w->d, desc2, (size_t)n2 * 32, s)) ||
        (rc = w->out0.ensure((size_t)n1 * 4)) || (rc = w->out1.ensure((size_t)n1 * 4)) ||
        (rc = w->f.ensure(((size_t)2 * n2 + 64 * 48 + 1) * 4)))
        return rc;
    const uint8_t* dp = w->a.as<uint8_t>();
    int* cellOf = w->f.as<int>();
    int* sorted = cellOf + n2;
    int* cellStart = sorted + n2;
    if ((rc = launch_fuse_match(n1, reinterpret_cast<const float*>(dp + oX), ur1 ? reinterpret_cast<const float*>(dp + oU) : nullptr,
                                reinterpret_cast<const int*>(dp + oL), dp + oV, dp + oD, w->c.as<orbx_keypoint>(), n2, w->d.as<uint8_t>(),
                                u_right2 ? reinterpret_cast<const float*>(dp + oR) : nullptr, reinterpret_cast<const float*>(dp + oS),
                                bounds, cellOf, sorted, cellStart, w->out0.as<int>(), w->out1.as<int>(), s)))
        return rc;
    ORBX_CUDA(cudaMemcpyAsync(best_idx, w->out0.p, (size_t)n1 * 4, cudaMemcpyDeviceToHost, s));
    if (best_dist) ORBX_CUDA(cudaMemcpyAsync(best_dist, w->out1.p, (size_t)n1 * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_search_for_triangulation(int n_nodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                                  const orbx_keypoint* k1, const uint8_t* desc1, const uint8_t* valid1, const uint8_t* stereo1, int n2,
                                  const orbx_keypoint* k2, const uint8_t* desc2, const uint8_t* valid2, const uint8_t* stereo2,
                                  const float f12[9], float ex, float ey, const float* scale_factors2, const float* level_sigma2_2,
                                  int nlevels, int check_ori, int* match12, int* nmatches) {
    if (n_nodes < 0 || n1 < 0 || n2 < 0 || !nmatches || !f12 || nlevels <= 0 || nlevels > 64 || !scale_factors2 || !level_sigma2_2 ||
        (n_nodes > 0 && (!start1 || !start2)) || (n1 > 0 && (!k1 || !desc1 || !valid1 || !stereo1 || !match12)) ||
        (n2 > 0 && (!k2 || !desc2 || !valid2 || !stereo2))) {
        set_error("orbx_search_for_triangulation: bad argument");
        return ORBX_E_INVALID;
    }
    *nmatches = 0;
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    if (n_nodes == 0 || n1 == 0 || n2 == 0) return ORBX_OK;
    if (start1[0] != 0 || start2[0] != 0) {
        set_error("orbx_search_for_triangulation: start1[0] and start2[0] must be 0");
        return ORBX_E_INVALID;
    }
    for (int k = 0; k < n_nodes; ++k)
        if (start1[k + 1] < start1[k] || start2[k + 1] < start2[k] || start2[k + 1] - start2[k] > 65535) {
            set_error("orbx_search_for_triangulation: node %d has a malformed feature range", k);
            return ORBX_E_INVALID;
        }
    const int nItems = start1[n_nodes], nIdx2 = start2[n_nodes];
    if ((nItems > 0 && !idx1) || (nIdx2 > 0 && !idx2)) {
        set_error("orbx_search_for_triangulation: null feature index list");
        return ORBX_E_INVALID;
    }
    for (int i = 0; i < nItems; ++i)
        if (idx1[i] >= (uint32_t)n1) { set_error("orbx_search_for_triangulation: idx1[%d] out of range", i); return ORBX_E_INVALID; }
    for (int i = 0; i < nIdx2; ++i)
        if (idx2[i] >= (uint32_t)n2) { set_error("orbx_search_for_triangulation: idx2[%d] out of range", i); return ORBX_E_INVALID; }
    for (int i = 0; i < n1; ++i)
        if (k1[i].octave < 0 || k1[i].octave >= nlevels) { set_error("orbx_search_for_triangulation: k1[%d].octave out of range", i); return ORBX_E_INVALID; }
    for (int i = 0; i < n2; ++i)
        if (k2[i].octave < 0 || k2[i].octave >= nlevels) { set_error("orbx_search_for_triangulation: k2[%d].octave out of range", i); return ORBX_E_INVALID; }
    if (nItems == 0 || nIdx2 == 0) return ORBX_OK;
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = w->stream;
    // one packed upload: [idx1 | nodeOf | start2 | idx2 | scale | sigma2 | k1 | k2 | desc1 | desc2 | valid1 | stereo1 | valid2 | stereo2]
    const size_t kb = sizeof(orbx_keypoint);
    const size_t oI1 = 0, oNode = oI1 + (size_t)nItems * 4, oS2 = oNode + (size_t)nItems * 4, oI2 = oS2 + (size_t)(n_nodes + 1) * 4,
                 oSF = oI2 + (size_t)nIdx2 * 4, oSG = oSF + (size_t)nlevels * 4, oK1 = oSG + (size_t)nlevels * 4, oK2 = oK1 + (size_t)n1 * kb,
                 oD1 = (oK2 + (size_t)n2 * kb + 15) / 16 * 16, oD2 = oD1 + (size_t)n1 * 32, oV1 = oD2 + (size_t)n2 * 32, oT1 = oV1 + n1,
                 oV2 = oT1 + n1, oT2 = oV2 + n2, total = oT2 + n2;
    if ((rc = w->pin.ensure(total))) return rc;
    uint8_t* hp = w->pin.as<uint8_t>();
    int* hNode = reinterpret_cast<int*>(hp + oNode);
    for (int k = 0; k < n_nodes; ++k)
        for (int i = start1[k]; i < start1[k + 1]; ++i) hNode[i] = k;
    memcpy(hp + oI1, idx1, (size_t)nItems * 4);
    memcpy(hp + oS2, start2, (size_t)(n_nodes + 1) * 4);
    memcpy(hp + oI2, idx2, (size_t)nIdx2 * 4);
    memcpy(hp + oSF, scale_factors2, (size_t)nlevels * 4);
    memcpy(hp + oSG, level_sigma2_2, (size_t)nlevels * 4);
    memcpy(hp + oK1, k1, (size_t)n1 * kb);
    memcpy(hp + oK2, k2, (size_t)n2 * kb);
    memcpy(hp + oD1, desc1, (size_t)n1 * 32);
    memcpy(hp + oD2, desc2, (size_t)n2 * 32);
    memcpy(hp + oV1, valid1, n1);
    memcpy(hp + oT1, stereo1, n1);
    memcpy(hp + oV2, valid2, n2);
    memcpy(hp + oT2, stereo2, n2);
    if ((rc = w->a.ensure(total)) || (rc = w->out0.ensure((size_t)n1 * 4 + 16)) || (rc = w->h.ensure((size_t)nItems + 16)) ||
        (rc = w->status.ensure(32 * 4 + 16)))
        return rc;
    ORBX_CUDA(cudaMemcpyAsync(w->a.p, hp, total, cudaMemcpyHostToDevice, s));
    const uint8_t* dp = w->a.as<uint8_t>();
    int* dHist = w->status.as<int>();
    int* dN = dHist + 32;
    if ((rc = launch_search_for_triangulation(
             nItems, reinterpret_cast<const int*>(dp + oNode), reinterpret_cast<const uint32_t*>(dp + oI1),
             reinterpret_cast<const int*>(dp + oS2), reinterpret_cast<const uint32_t*>(dp + oI2), n1,
             reinterpret_cast<const orbx_keypoint*>(dp + oK1), dp + oD1, dp + oV1, dp + oT1, n2,
             reinterpret_cast<const orbx_keypoint*>(dp + oK2), dp + oD2, dp + oV2, dp + oT2, reinterpret_cast<const float*>(dp + oSF),
             reinterpret_cast<const float*>(dp + oSG), f12, ex, ey, check_ori, w->out0.as<int>(), w->h.as<signed char>(), dHist, dN, s)))
        return rc;
    int st = 0;
    ORBX_CUDA(cudaMemcpyAsync(match12, w->out0.p, (size_t)n1 * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(&st, dN, 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    *nmatches = st;
    return ORBX_OK;
}

int orbx_distinctive_descriptors(int n_points, const int* start, const uint8_t* desc, int* best_idx, int* best_median) {
    if (n_points < 0 || (n_points > 0 && (!start || !best_idx))) {
        set_error("orbx_distinctive_descriptors: bad argument");
        return ORBX_E_INVALID;
    }
    if (n_points == 0) return ORBX_OK;
    if (start[0] != 0) {
        set_error("orbx_distinctive_descriptors: start[0] must be 0");
        return ORBX_E_INVALID;
    }
    for (int i = 0; i < n_points; ++i)
        if (start[i + 1] < start[i] || start[i + 1] - start[i] > 65535) {
            set_error("orbx_distinctive_descriptors: point %d has a malformed observation range (at most 65535 observations)", i);
            return ORBX_E_INVALID;
        }
    const int total = start[n_points];
    if (total > 0 && !desc) {
        set_error("orbx_distinctive_descriptors: null descriptor table");
        return ORBX_E_INVALID;
    }
    Workspace* w;
    int rc = workspace(&w);
    if (rc) return rc;
    cudaStream_t s = w->stream;
    if ((rc = upload(w->a, start, (size_t)(n_points + 1) * 4, s)) || (rc = upload(w->b, desc, (size_t)std::max(total, 1) * 32, s)) ||
        (rc = w->out0.ensure((size_t)n_points * 4)) || (rc = w->out1.ensure((size_t)n_points * 4)))
        return rc;
    if ((rc = launch_distinctive(n_points, w->a.as<int>(), w->b.as<uint8_t>(), w->out0.as<int>(), w->out1.as<int>(), s))) return rc;
    std::vector<int> med;
    if (!best_median) med.resize(n_points);
    ORBX_CUDA(cudaMemcpyAsync(best_idx, w->out0.p, (size_t)n_points * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(best_median ? best_median : med.data(), w->out1.p, (size_t)n_points * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_stereo_match(orbx_extractor* exL, orbx_extractor* exR, const orbx_keypoint* kL, int nL, const uint8_t* descL,
                      const orbx_keypoint* kR, int nR, const uint8_t* descR, float bf, float mb, float* u_right, float* depth) {
    ExtractorView L, R;
    int rc;
    if ((rc = extractor_view(exL, &L)) || (rc = extractor_view(exR, &R))) return rc;
    if (nL < 0 || nR < 0 || nR > 65535 || (nL > 0 && (!kL || !descL || !u_right || !depth)) || (nR > 0 && (!kR || !descR))) {
        set_error("orbx_stereo_match: bad argument (nR <= 65535)");
        return ORBX_E_INVALID;
    }
    if (L.serial == 0 || R.serial == 0 || L.device != R.device || L.g->nlevels != R.g->nlevels) {
        set_error("orbx_stereo_match: both extractors must have extracted a frame on the same device with the same pyramid depth");
        return ORBX_E_INVALID;
    }
    if (L.lastBatch != 1 || R.lastBatch != 1) {  // the scratch then holds exactly the frame orbx_pyramid_level_host(0, ...) serves
        set_error("orbx_stereo_match: the last call of each extractor must have been a single-frame orbx_extract (left: %d frames, right: %d)",
                  L.lastBatch, R.lastBatch);
        return ORBX_E_INVALID;
    }
    // keypoints must lie where the extractor puts them (>= 16 px inside their level): the SAD windows read 5 + 5 px around the
    // scaled position, which then stays inside the padded plane; stale or foreign keypoints are refused instead of read out of bounds
    for (int pass = 0; pass < 2; ++pass) {
        const orbx_keypoint* k = pass ? kR : kL;
        const ExtractorView& V = pass ? R : L;
        for (int i = 0; i < (pass ? nR : nL); ++i) {
            const int o = k[i].octave;
            bool ok = o >= 0 && o < V.g->nlevels;
            if (ok) {
                const float sx = k[i].x * L.invScale[o], sy = k[i].y * L.invScale[o];
                ok = sx >= -1.f && sy >= -1.f && sx <= (float)V.g->lv[o].w && sy <= (float)V.g->lv[o].h;
            }
            if (!ok) {
                set_error("orbx_stereo_match: %s keypoint %d (%.1f, %.1f, octave %d) lies outside its pyramid level", pass ? "right" : "left", i,
                          k[i].x, k[i].y, o);
                return ORBX_E_INVALID;
            }
        }
    }
    if (nL == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(L.device));
    Workspace* w;
    if ((rc = workspace(&w))) return rc;
    cudaStream_t s = w->stream;
    ORBX_CUDA(cudaStreamSynchronize(L.stream));  // the pyramids of the last extract calls are complete
    ORBX_CUDA(cudaStreamSynchronize(R.stream));
    StereoLevels lv;
    memset(&lv, 0, sizeof lv);
    lv.nlevels = L.g->nlevels;
    for (int l = 0; l < lv.nlevels; ++l) {
        const LevelGeom &a = L.g->lv[l], &b = R.g->lv[l];
        lv.left[l] = a.plane + (size_t)EDGE * a.pitch + XOFF;
        lv.right[l] = b.plane + (size_t)EDGE * b.pitch + XOFF;
        lv.pitchL[l] = a.pitch;
        lv.pitchR[l] = b.pitch;
        lv.widthR[l] = b.w;
        lv.scale[l] = L.scale[l];       // Frame copies the LEFT extractor's tables (src/Frame.cc:68-74)
        lv.invScale[l] = L.invScale[l];
    }
    const float maxD = bf / mb;      // minZ = mb, maxD = mbf / minZ (:494-496)
    const size_t fb = (size_t)nL * 4;
    if ((rc = upload(w->a, kL, (size_t)nL * sizeof(orbx_keypoint), s)) || (rc = upload(w->b, descL, (size_t)nL * 32, s)) ||
        (rc = upload(w->c, kR, (size_t)nR * sizeof(orbx_keypoint), s)) || (rc = upload(w->d, descR, (size_t)nR * 32, s)) ||
        (rc = w->out0.ensure(fb)) || (rc = w->out1.ensure(fb)) || (rc = w->out2.ensure(fb)))
        return rc;
    if ((rc = launch_stereo(lv, w->a.as<orbx_keypoint>(), nL, w->b.as<uint8_t>(), w->c.as<orbx_keypoint>(), nR, w->d.as<uint8_t>(), bf,
                            maxD, w->out0.as<float>(), w->out1.as<float>(), w->out2.as<int>(), s)))
        return rc;
    ORBX_CUDA(cudaMemcpyAsync(u_right, w->out0.p, fb, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(depth, w->out1.p, fb, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}


int orbx_stereo_match_last(orbx_extractor* exL, orbx_extractor* exR, float bf, float mb, float* u_right, float* depth, int cap) {
    ExtractorView L, R;
    int rc;
    if ((rc = extractor_view(exL, &L)) || (rc = extractor_view(exR, &R))) return rc;
    if (L.serial == 0 || R.serial == 0 || L.device != R.device || L.g->nlevels != R.g->nlevels) {
        set_error("orbx_stereo_match_last: both extractors must have extracted a frame on the same device with the same pyramid depth");
        return ORBX_E_INVALID;
    }
    if (L.lastBatch != 1 || R.lastBatch != 1 || !L.lastKps || !R.lastKps) {
        set_error("orbx_stereo_match_last: the last call of each extractor must have been a single-frame orbx_extract");
        return ORBX_E_INVALID;
    }
    const int nL = L.lastCount, nR = R.lastCount;
    if (nR > 65535 || cap < nL || (nL > 0 && (!u_right || !depth))) {
        set_error("orbx_stereo_match_last: bad argument (%d left keypoints, cap %d; right keypoints %d <= 65535)", nL, cap, nR);
        return ORBX_E_INVALID;
    }
    if (nL == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(L.device));
    Workspace* w;
    if ((rc = workspace(&w))) return rc;
    cudaStream_t s = w->stream;
    ORBX_CUDA(cudaStreamSynchronize(L.stream));  // the last extract calls were synchronous: these return at once
    ORBX_CUDA(cudaStreamSynchronize(R.stream));
    StereoLevels lv;
    memset(&lv, 0, sizeof lv);
    lv.nlevels = L.g->nlevels;
    for (int l = 0; l < lv.nlevels; ++l) {
        const LevelGeom &a = L.g->lv[l], &b = R.g->lv[l];
        lv.left[l] = a.plane + (size_t)EDGE * a.pitch + XOFF;
        lv.right[l] = b.plane + (size_t)EDGE * b.pitch + XOFF;
        lv.pitchL[l] = a.pitch;
        lv.pitchR[l] = b.pitch;
        lv.widthR[l] = b.w;
        lv.scale[l] = L.scale[l];
        lv.invScale[l] = L.invScale[l];
    }
    const float maxD = bf / mb;
    const size_t fb = (size_t)nL * 4;
    if ((rc = w->out0.ensure(2 * fb)) || (rc = w->out2.ensure(fb))) return rc;  // u_right | depth in one buffer: one copy back
    if ((rc = w->pin.ensure(2 * fb))) return rc;
    if ((rc = launch_stereo(lv, L.lastKps, nL, L.lastDesc, R.lastKps, nR, R.lastDesc, bf, maxD, w->out0.as<float>(), w->out0.as<float>() + nL,
                            w->out2.as<int>(), s)))
        return rc;
    ORBX_CUDA(cudaMemcpyAsync(w->pin.p, w->out0.p, 2 * fb, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    memcpy(u_right, w->pin.p, fb);
    memcpy(depth, (const uint8_t*)w->pin.p + fb, fb);
    return ORBX_OK;
}

}  // extern "C"
