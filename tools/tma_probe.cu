// Development probe for the TMA / mbarrier wrappers in csrc/tma.cuh (run on the GPU box, one variant per process).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../orb_slam_cuda_b200/csrc/tma.cuh"
using namespace orbx;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void k_bar(int* out) {
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    __syncthreads();
    if (threadIdx.x == 0) mbar_expect_tx(&bar, 0);
    mbar_wait(&bar, 0);
    if (threadIdx.x == 0) out[0] = 1;
}
__global__ void k_tma(const __grid_constant__ CUtensorMap map, const CUtensorMap* gmap, int useG, int c0, int c1, int c2, int bytes, uint8_t* out) {
    extern __shared__ __align__(128) uint8_t sm[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(sm + ((bytes + 127) / 128) * 128);
    if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
    __syncthreads();
    if (threadIdx.x == 0) { mbar_expect_tx(bar, bytes); tma_load_3d(sm, useG ? gmap : &map, c0, c1, c2, bar); }
    mbar_wait(bar, 0);
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = sm[i];
}
typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv) {
    int v = argc > 1 ? atoi(argv[1]) : 0;
    int* dout; CK(cudaMalloc(&dout, 4));
    if (v == 0) {
        k_bar<<<1, 128>>>(dout); CK(cudaDeviceSynchronize()); printf("variant 0 (mbarrier only): ok\n"); return 0;
    }
    const int pitch = 768, rows = 518, batch = 2;
    std::vector<uint8_t> h((size_t)pitch * rows * batch);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 7 + i / pitch) & 255);
    uint8_t* d; CK(cudaMalloc(&d, h.size())); CK(cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
    void* fp; cudaDriverEntryPointQueryResult q; CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    int bw = argc > 2 ? atoi(argv[2]) : 64, bh = argc > 3 ? atoi(argv[3]) : 32;
    cuuint64_t dims[3] = {pitch, rows, batch}; cuuint64_t str[2] = {pitch, (cuuint64_t)pitch * rows};
    cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}, es[3] = {1, 1, 1};
    CUtensorMap m;
    CUresult r = ((Enc)fp)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc %d\n", (int)r);
    CUtensorMap* gm; CK(cudaMalloc(&gm, sizeof m)); CK(cudaMemcpy(gm, &m, sizeof m, cudaMemcpyHostToDevice));
    int bytes = bw * bh; uint8_t* o; CK(cudaMalloc(&o, bytes));
    int c0 = argc > 4 ? atoi(argv[4]) : 64, c1 = argc > 5 ? atoi(argv[5]) : 32, c2 = 1;
    k_tma<<<1, 128, bytes + 256>>>(m, gm, v == 3, c0, c1, c2, bytes, o);
    CK(cudaDeviceSynchronize());
    std::vector<uint8_t> ho(bytes); CK(cudaMemcpy(ho.data(), o, bytes, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int y = 0; y < bh; ++y) for (int x = 0; x < bw; ++x)
        if (ho[y * bw + x] != h[(size_t)c2 * pitch * rows + (size_t)(c1 + y) * pitch + c0 + x]) ++bad;
    printf("box %dx%d at (%d,%d) gmem-desc %d: ok, mismatches %d\n", bw, bh, c0, c1, v == 3, bad);
    return 0;
}
