// orbx_match_bruteforce_noalloc — the brute-force matcher for callers that cannot allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The kernel is k_bruteforce (k_match.cu) unchanged; what differs is where its scratch lives: the
// per-slice partials and the per-query-block arrival counters are file-scope __device__ arrays of a fixed maximum size, so
// the call allocates nothing — no cudaMalloc / cudaFree, no Workspace, no stream or event creation.
//
// Every GPU that loads liborbx.so holds the two arrays (4 MiB + 32 KiB) until the process exits.  All calls on one device
// share them: calls that overlap on the GPU (two streams, two threads) race on the scratch and give wrong matches — the
// caller serialises them.  In host code the array names are host-side stand-ins, not device addresses; the device
// addresses are taken once per device with cudaGetSymbolAddress (below) and only those are handed to the kernel.
#include <mutex>
#include "orbx_internal.h"

namespace {
constexpr int kMaxQueries = ORBX_NOALLOC_MAX_QUERIES;
constexpr int kMaxSlices = 64;  // launch_bruteforce never splits the database into more slices (k_match.cu)
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned long long g_partial[(size_t)kMaxSlices * kMaxQueries];  // 4 MiB
static __device__ __align__(256) unsigned int g_counters[kMaxQueries];                           // 32 KiB (one per 128 queries used)

struct DeviceScratch {
    unsigned long long* partial = nullptr;
    unsigned int* counters = nullptr;
};
DeviceScratch g_scratch[kMaxDevices];
std::mutex g_lock;

// set-up, once per device, with that device current: resolve the symbols to THIS device's addresses
int scratch_for_current_device(DeviceScratch** out) {
    int dev = -1;
    ORBX_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        orbx::set_error("orbx_match_bruteforce_noalloc: device ordinal %d not supported (max %d)", dev, kMaxDevices - 1);
        return ORBX_E_INVALID;
    }
    std::lock_guard<std::mutex> guard(g_lock);
    DeviceScratch& sc = g_scratch[dev];
    if (!sc.partial || !sc.counters) {
        void *p = nullptr, *c = nullptr;
        ORBX_CUDA(cudaGetSymbolAddress(&p, g_partial));
        ORBX_CUDA(cudaGetSymbolAddress(&c, g_counters));
        if (!p || !c) {
            orbx::set_error("orbx_match_bruteforce_noalloc: cudaGetSymbolAddress returned a null address on device %d", dev);
            return ORBX_E_CUDA;
        }
        sc.partial = (unsigned long long*)p;
        sc.counters = (unsigned int*)c;
    }
    *out = &sc;
    return ORBX_OK;
}
}  // namespace

extern "C" int orbx_match_bruteforce_noalloc(const uint8_t* d_q, int nq, const uint8_t* d_d, int nd, int th, float ratio, int* d_idx,
                                             int* d_d1, int* d_d2, int* d_match, void* stream) {
    using namespace orbx;
    if (nq < 0 || nd < 0 || nd > 0x3FFFFE || (nq > 0 && !d_q) || (nd > 0 && !d_d)) {
        set_error("orbx_match_bruteforce_noalloc: bad argument (nd must be < 4194302)");
        return ORBX_E_INVALID;
    }
    if (nq > kMaxQueries) {
        set_error("orbx_match_bruteforce_noalloc: nq = %d exceeds ORBX_NOALLOC_MAX_QUERIES = %d (the static scratch is sized for that; "
                  "use orbx_match_bruteforce_device for larger query sets)", nq, kMaxQueries);
        return ORBX_E_INVALID;
    }
    if (nq == 0) return ORBX_OK;
    DeviceScratch* sc = nullptr;
    int rc = scratch_for_current_device(&sc);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;  // NULL = the default stream: no stream is created here
    // the arrival counters must start at zero; the kernel leaves them at zero, but an aborted earlier launch may not have.
    // The partials need no clearing: every slot read by the merge is written by this launch first.
    const size_t counterBytes = (size_t)((nq + 127) / 128) * sizeof(unsigned int);
    ORBX_CUDA(cudaMemsetAsync(sc->counters, 0, counterBytes, s));
    return launch_bruteforce(d_q, nq, d_d, nd, th, ratio, d_idx, d_d1, d_d2, d_match, sc->partial, sc->counters, s);
}
