// csrc/vocabulary.cu — the DBoW2 vocabulary tree descent behind Frame::ComputeBoW (reference src/Frame.cc:394-401 ->
// TemplatedVocabulary::transform, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1194 and :1218-1261) and the text loader
// (:1338-1424) that defines the node numbering.  SURVEY.md §8f rank 2.
//
// HBM layout: the children of a node are stored contiguously in the order DBoW2 pushed them (file order):
//   childStart[n_nodes + 1], childNode[n_nodes - 1] (node id per child slot), childDesc[(n_nodes - 1) x 32 B] (descriptor per
//   child slot, so one level of the descent is ONE coalesced read of k x 32 B), wordOf[n_nodes], weightOf[n_nodes] (double).
// Kernel: 16 lanes per feature; lane c takes child c of the current node (XOR + POPC over 8 words), the half-warp reduces
// (distance << 20 | child position) with a min — the smallest distance, first child among equals, exactly the strict '<' of
// :1241-1245 — and moves down until a node without children.  L dependent 320-byte reads per feature: latency bound, the tree
// (35 MB of descriptors for ORBvoc's 10^6 words) stays L2 resident across the features of a frame.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "orbx_internal.h"

using namespace orbx;

struct orbx_vocabulary {
    int device = 0, k = 0, L = 0, scoring = 0, weighting = 0, nNodes = 0, nWords = 0;
    cudaStream_t stream = nullptr;
    DevBuf childStart, childNode, childDesc, wordOf, weightOf, in, out;
    PinBuf pin;
    std::mutex lock;
};

namespace {

constexpr int VG = 16;  // lanes per feature

__global__ void __launch_bounds__(128) k_vocab_transform(const uint4* __restrict__ feat, int n, const int* __restrict__ childStart,
                                                          const int* __restrict__ childNode, const uint4* __restrict__ childDesc,
                                                          const uint32_t* __restrict__ wordOf, const double* __restrict__ weightOf,
                                                          int nidLevel, uint32_t* __restrict__ wordId, double* __restrict__ weight,
                                                          uint32_t* __restrict__ nodeId) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int f = t / VG, sub = threadIdx.x & (VG - 1);
    if (f >= n) return;  // n is padded to whole half-warps by the indexing: a half-warp is either all in or all out
    const uint32_t mask = 0xFFFFu << (threadIdx.x & 16);
    const uint4 a0 = feat[2 * f], a1 = feat[2 * f + 1];
    uint32_t nid = nidLevel <= 0 ? 0u : 0xFFFFFFFFu;
    int node = 0, level = 0;
    for (;;) {
        const int b = childStart[node], e = childStart[node + 1];
        if (b == e) break;
        ++level;
        uint32_t best = 0xFFFFFFFFu;
        for (int c = b + sub; c < e; c += VG) {
            const uint4 d0 = __ldg(&childDesc[2 * c]), d1 = __ldg(&childDesc[2 * c + 1]);
            const int dist = __popc(a0.x ^ d0.x) + __popc(a0.y ^ d0.y) + __popc(a0.z ^ d0.z) + __popc(a0.w ^ d0.w) + __popc(a1.x ^ d1.x) +
                             __popc(a1.y ^ d1.y) + __popc(a1.z ^ d1.z) + __popc(a1.w ^ d1.w);
            best = min(best, ((uint32_t)dist << 20) | (uint32_t)(c - b));
        }
#pragma unroll
        for (int o = VG / 2; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(mask, best, o));
        node = childNode[b + (int)(best & 0xFFFFFu)];
        if (level == nidLevel) nid = (uint32_t)node;
    }
    if (sub == 0) {
        wordId[f] = wordOf[node];
        weight[f] = weightOf[node];
        nodeId[f] = nid;
    }
}

int build(orbx_vocabulary* v, int nNodes, const int32_t* parent, const uint8_t* isLeaf, const uint8_t* desc, const double* weight) {
    // children in push order (:1388), word ids over the leaf-flagged nodes in node order (:1406-1413)
    std::vector<int> start(nNodes + 1, 0), fill(nNodes, 0), childNode(nNodes > 1 ? nNodes - 1 : 0);
    for (int i = 1; i < nNodes; ++i) {
        if (parent[i] < 0 || parent[i] >= i) {
            set_error("vocabulary: node %d has parent %d (a parent must precede its children)", i, parent[i]);
            return ORBX_E_INVALID;
        }
        ++start[parent[i] + 1];
    }
    for (int i = 0; i < nNodes; ++i) {
        if (start[i + 1] >= (1 << 20)) {
            set_error("vocabulary: node %d has %d children (limit 2^20 - 1)", i, start[i + 1]);
            return ORBX_E_INVALID;
        }
        start[i + 1] += start[i];
    }
    if (nNodes < 2 || start[1] == 0) {
        set_error("vocabulary: the root has no children (empty vocabulary)");
        return ORBX_E_INVALID;
    }
    std::vector<uint8_t> childDesc((size_t)childNode.size() * 32);
    std::vector<uint32_t> wordOf(nNodes, 0);
    uint32_t nWords = 0;
    for (int i = 1; i < nNodes; ++i) {
        const int slot = start[parent[i]] + fill[parent[i]]++;
        childNode[slot] = i;
        memcpy(&childDesc[(size_t)slot * 32], desc + (size_t)i * 32, 32);
        if (isLeaf[i]) wordOf[i] = nWords++;
    }
    v->nNodes = nNodes;
    v->nWords = (int)nWords;
    int rc;
    if ((rc = v->childStart.ensure(start.size() * 4)) || (rc = v->childNode.ensure(childNode.size() * 4)) ||
        (rc = v->childDesc.ensure(childDesc.size())) || (rc = v->wordOf.ensure((size_t)nNodes * 4)) ||
        (rc = v->weightOf.ensure((size_t)nNodes * 8)))
        return rc;
    ORBX_CUDA(cudaMemcpy(v->childStart.p, start.data(), start.size() * 4, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(v->childNode.p, childNode.data(), childNode.size() * 4, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(v->childDesc.p, childDesc.data(), childDesc.size(), cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(v->wordOf.p, wordOf.data(), (size_t)nNodes * 4, cudaMemcpyHostToDevice));
    ORBX_CUDA(cudaMemcpy(v->weightOf.p, weight, (size_t)nNodes * 8, cudaMemcpyHostToDevice));
    return ORBX_OK;
}

// One `ssnode >> value` after another; `failed` once a token is missing or does not parse.
struct Tokens {
    const char* p;
    const char* end;
    bool failed = false;
    bool next(const char** b, const char** e) {
        if (failed) return false;
        while (p < end && (*p == ' ' || *p == '\t' || *p == '\r' || *p == '\v' || *p == '\f')) ++p;
        if (p >= end) { failed = true; return false; }
        *b = p;
        while (p < end && !(*p == ' ' || *p == '\t' || *p == '\r' || *p == '\v' || *p == '\f')) ++p;
        *e = p;
        return true;
    }
    long integer() {
        const char *b, *e;
        if (!next(&b, &e)) return 0;
        char* stop = nullptr;
        const long v = strtol(b, &stop, 10);
        if (stop == b) { failed = true; return 0; }
        return v;
    }
    double real() {
        const char *b, *e;
        if (!next(&b, &e)) return 0.0;
        char* stop = nullptr;
        const double v = strtod(b, &stop);
        if (stop == b) { failed = true; return 0.0; }
        return v;
    }
};

}  // namespace

extern "C" {

int orbx_vocabulary_create(int k, int L, int scoring, int weighting, int n_nodes, const int32_t* parent, const uint8_t* is_leaf,
                           const uint8_t* desc, const double* weight, orbx_vocabulary** out) {
    if (!out || n_nodes < 2 || !parent || !is_leaf || !desc || !weight || L < 0 || scoring < 0 || scoring > 5 || weighting < 0 ||
        weighting > 3) {
        set_error("orbx_vocabulary_create: bad argument");
        return ORBX_E_INVALID;
    }
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        set_error("no CUDA device: %s (this library has no CPU path)", e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        return ORBX_E_CUDA;
    }
    orbx_vocabulary* v = new orbx_vocabulary();
    v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
    cudaGetDevice(&v->device);
    int rc = build(v, n_nodes, parent, is_leaf, desc, weight);
    if (rc == ORBX_OK && cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking) != cudaSuccess) {
        set_error("orbx_vocabulary_create: cudaStreamCreate failed");
        rc = ORBX_E_CUDA;
    }
    if (rc) {
        orbx_vocabulary_destroy(v);
        return rc;
    }
    *out = v;
    return ORBX_OK;
}

int orbx_vocabulary_load_text(const char* path, orbx_vocabulary** out) {
    if (!path || !out) {
        set_error("orbx_vocabulary_load_text: bad argument");
        return ORBX_E_INVALID;
    }
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) {
        set_error("orbx_vocabulary_load_text: cannot open %s", path);
        return ORBX_E_INVALID;
    }
    std::string text;
    char buf[1 << 16];
    size_t got;
    while ((got = fread(buf, 1, sizeof buf, f)) > 0) text.append(buf, got);
    fclose(f);
    const char* p = text.data();
    const char* end = p + text.size();
    const char* nl = (const char*)memchr(p, '\n', end - p);
    Tokens head{p, nl ? nl : end};
    const int k = (int)head.integer(), L = (int)head.integer(), n1 = (int)head.integer(), n2 = (int)head.integer();
    if (k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) {  // :1359-1363
        set_error("orbx_vocabulary_load_text: %s is not a DBoW2 text vocabulary (first line must be `k L scoring weighting`)", path);
        return ORBX_E_INVALID;
    }
    std::vector<int32_t> parent(1, -1);
    std::vector<uint8_t> leaf(1, 0), desc(32, 0);
    std::vector<double> weight(1, 0.0);
    // One node per line (:1377-1417).  `while(!f.eof()) getline(...)` also runs once more on the empty remainder after the last
    // newline that saveToTextFile writes; there `ssnode >> pid` fails in the stream's sentry (nothing but end-of-file to skip), so
    // neither pid nor nIsLeaf is ever written and the reference indexes m_nodes with an uninitialised int — undefined behaviour
    // (observed with gcc: a stale value of the previous line, or a crash).  A line without any token therefore makes no node
    // here; a line that starts but does not parse is an error.
    bool more = nl != nullptr;
    p = nl ? nl + 1 : end;
    int lineNo = 1;
    while (more) {
        nl = (const char*)memchr(p, '\n', end - p);
        Tokens t{p, nl ? nl : end};
        more = nl != nullptr;
        p = nl ? nl + 1 : end;
        ++lineNo;
        const char *tb, *te;
        Tokens probe = t;
        if (!probe.next(&tb, &te)) continue;  // blank line
        const long pid = t.integer();
        const long isLeaf = t.integer();
        uint8_t d[32];
        for (int i = 0; i < 32; ++i) d[i] = (uint8_t)t.integer();  // FORB::fromString (FORB.cpp:120-135): int -> unsigned char
        const double w = t.real();
        const int nid = (int)parent.size();
        if (t.failed) {
            set_error("orbx_vocabulary_load_text: line %d is not `parent is_leaf d0 .. d31 weight`", lineNo);
            return ORBX_E_INVALID;
        }
        if (pid < 0 || pid >= nid) {
            set_error("orbx_vocabulary_load_text: line %d names parent %ld before it exists", lineNo, pid);
            return ORBX_E_INVALID;
        }
        parent.push_back((int32_t)pid);
        leaf.push_back(isLeaf > 0);
        desc.insert(desc.end(), d, d + 32);
        weight.push_back(w);
    }
    return orbx_vocabulary_create(k, L, n1, n2, (int)parent.size(), parent.data(), leaf.data(), desc.data(), weight.data(), out);
}

void orbx_vocabulary_destroy(orbx_vocabulary* v) {
    if (!v) return;
    for (DevBuf* b : {&v->childStart, &v->childNode, &v->childDesc, &v->wordOf, &v->weightOf, &v->in, &v->out}) b->release();
    if (v->stream) cudaStreamDestroy(v->stream);
    delete v;
}

int orbx_vocabulary_info(const orbx_vocabulary* v, int* k, int* L, int* scoring, int* weighting, int* n_nodes, int* n_words) {
    if (!v) {
        set_error("orbx_vocabulary_info: null vocabulary");
        return ORBX_E_INVALID;
    }
    if (k) *k = v->k;
    if (L) *L = v->L;
    if (scoring) *scoring = v->scoring;
    if (weighting) *weighting = v->weighting;
    if (n_nodes) *n_nodes = v->nNodes;
    if (n_words) *n_words = v->nWords;
    return ORBX_OK;
}

int orbx_vocabulary_transform(orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup, uint32_t* word_id, double* weight,
                              uint32_t* node_id) {
    if (!v || n < 0 || (n > 0 && (!desc || !word_id || !weight || !node_id))) {
        set_error("orbx_vocabulary_transform: bad argument");
        return ORBX_E_INVALID;
    }
    if (n == 0) return ORBX_OK;
    std::lock_guard<std::mutex> guard(v->lock);
    ORBX_CUDA(cudaSetDevice(v->device));
    const size_t inBytes = (size_t)n * 32, oW = 0, oX = oW + (size_t)n * 8, oN = oX + (size_t)n * 4, outBytes = oN + (size_t)n * 4;
    int rc;
    if ((rc = v->in.ensure(inBytes)) || (rc = v->out.ensure(outBytes)) || (rc = v->pin.ensure(inBytes + outBytes))) return rc;
    uint8_t* hp = v->pin.as<uint8_t>();
    memcpy(hp, desc, inBytes);
    ORBX_CUDA(cudaMemcpyAsync(v->in.p, hp, inBytes, cudaMemcpyHostToDevice, v->stream));
    uint8_t* dOut = v->out.as<uint8_t>();
    const long long threads = (long long)n * VG;
    k_vocab_transform<<<(unsigned)((threads + 127) / 128), 128, 0, v->stream>>>(
        v->in.as<uint4>(), n, v->childStart.as<int>(), v->childNode.as<int>(), v->childDesc.as<uint4>(), v->wordOf.as<uint32_t>(),
        v->weightOf.as<double>(), v->L - levelsup, reinterpret_cast<uint32_t*>(dOut + oX), reinterpret_cast<double*>(dOut + oW),
        reinterpret_cast<uint32_t*>(dOut + oN));
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(hp + inBytes, dOut, outBytes, cudaMemcpyDeviceToHost, v->stream));
    ORBX_CUDA(cudaStreamSynchronize(v->stream));
    memcpy(weight, hp + inBytes + oW, (size_t)n * 8);
    memcpy(word_id, hp + inBytes + oX, (size_t)n * 4);
    memcpy(node_id, hp + inBytes + oN, (size_t)n * 4);
    return ORBX_OK;
}

}  // extern "C"
