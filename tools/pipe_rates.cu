// tools/pipe_rates.cu — measured issue rates of the integer instructions the ORB front end leans on (sm_100a).
// Development / evidence tool: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates tools/pipe_rates.cu && ./pipe_rates
// Every kernel runs 8 independent dependency chains per thread (opaque inline PTX, so ptxas cannot fold them), 32 warps per
// SM on every SM; the rate is warp instructions per clock per SM, from the CTA's own clock64() span (max over CTAs), and is
// written as one JSON object (profiles/rNN_pipe_rates.json).  BASELINE.md's POPC bound and DESIGN.md's pipe model quote it.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <string>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int ITERS = 2048, CHAINS = 8, THREADS = 256, CTAS_PER_SM = 4;

#define OP_LOP3(a, b, c) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a) : "r"(b), "r"(c))
#define OP_IADD3(a, b, c) asm volatile("add.u32 %0, %0, %1;" : "+r"(a) : "r"(b))
#define OP_IMAD(a, b, c) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))
#define OP_SHF(a, b, c) asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))
#define OP_PRMT(a, b, c) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))
#define OP_POPC(a, b, c) asm volatile("popc.b32 %0, %0;" : "+r"(a))
#define OP_VABSDIFF4(a, b, c) asm volatile("vabsdiff4.u32.u32.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(0))
#define OP_VSAD4(a, b, c) asm volatile("vabsdiff4.u32.u32.u32.add %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))
#define OP_VIMNMX3(a, b, c) asm volatile("{.reg .b32 t; min.s16x2 t, %0, %1; min.s16x2 %0, t, %2;}" : "+r"(a) : "r"(b), "r"(c))
#define OP_VIMNMX2(a, b, c) asm volatile("max.u16x2 %0, %0, %1;" : "+r"(a) : "r"(b))
#define OP_DP4A(a, b, c) asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a) : "r"(b), "r"(c))
#define OP_DP2A(a, b, c) asm volatile("dp2a.lo.u32.u32 %0, %1, %2, %0;" : "+r"(a) : "r"(b), "r"(c))
#define OP_FLO(a, b, c) asm volatile("bfind.u32 %0, %0;" : "+r"(a))
#define OP_BREV(a, b, c) asm volatile("brev.b32 %0, %0;" : "+r"(a))
#define OP_IMNMX(a, b, c) asm volatile("max.u32 %0, %0, %1;" : "+r"(a) : "r"(b))
#define OP_SHL(a, b, c) asm volatile("shl.b32 %0, %0, 1;" : "+r"(a))
#define OP_SHFL(a, b, c) asm volatile("shfl.sync.bfly.b32 %0, %0, 1, 0x1f, 0xffffffff;" : "+r"(a))
#define OP_BALLOT(a, b, c) asm volatile("{.reg .pred p; setp.ne.u32 p, %0, 0; vote.sync.ballot.b32 %0, p, 0xffffffff;}" : "+r"(a))
#define OP_SETP_SEL(a, b, c) asm volatile("{.reg .pred p; setp.gt.u32 p, %0, %1; selp.u32 %0, %1, %2, p;}" : "+r"(a) : "r"(b), "r"(c))
// mixes: one instruction of each pipe per slot (do the fma pipe's IMAD and the alu pipe's LOP3 / POPC issue side by side?)
#define OP_LOP3_IMAD(a, b, c) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96; mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))
#define OP_POPC_LOP3(a, b, c) asm volatile("popc.b32 %0, %0; lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a) : "r"(b), "r"(c))
#define OP_POPC_IMAD(a, b, c) asm volatile("popc.b32 %0, %0; mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))
#define OP_VABS_LOP3(a, b, c) asm volatile("vabsdiff4.u32.u32.u32 %0, %0, %1, %2; lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a) : "r"(b), "r"(c))
#define OP_VABS_IMAD(a, b, c) asm volatile("vabsdiff4.u32.u32.u32 %0, %0, %1, %2; mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))

#define KERNEL(NAME, OP)                                                                                       \
    __global__ void __launch_bounds__(THREADS) k_##NAME(uint32_t* out, long long* cyc, uint32_t s1, uint32_t s2) { \
        uint32_t r[CHAINS];                                                                                    \
        for (int i = 0; i < CHAINS; ++i) r[i] = threadIdx.x * 2654435761u + i * 40503u + s1;                   \
        uint32_t b = s1 | 1u, c = s2;                                                                          \
        __syncthreads();                                                                                       \
        const long long t0 = clock64();                                                                        \
        for (int it = 0; it < ITERS; ++it) {                                                                   \
            _Pragma("unroll") for (int i = 0; i < CHAINS; ++i) { OP(r[i], b, c); }                              \
        }                                                                                                      \
        const long long t1 = clock64();                                                                        \
        uint32_t acc = 0;                                                                                      \
        for (int i = 0; i < CHAINS; ++i) acc ^= r[i];                                                          \
        if (acc == 0x12345u) out[0] = acc;                                                                     \
        if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;                                                       \
    }

KERNEL(lop3, OP_LOP3) KERNEL(iadd3, OP_IADD3) KERNEL(imad, OP_IMAD) KERNEL(shf, OP_SHF) KERNEL(prmt, OP_PRMT) KERNEL(popc, OP_POPC)
KERNEL(vabsdiff4, OP_VABSDIFF4) KERNEL(vsad4, OP_VSAD4) KERNEL(vimnmx3_s16x2, OP_VIMNMX3) KERNEL(vimnmx_u16x2, OP_VIMNMX2)
KERNEL(dp4a, OP_DP4A) KERNEL(dp2a, OP_DP2A) KERNEL(flo, OP_FLO) KERNEL(brev, OP_BREV) KERNEL(imnmx, OP_IMNMX) KERNEL(shl, OP_SHL)
KERNEL(shfl, OP_SHFL) KERNEL(ballot, OP_BALLOT) KERNEL(setp_sel, OP_SETP_SEL) KERNEL(lop3_imad, OP_LOP3_IMAD)
KERNEL(popc_lop3, OP_POPC_LOP3) KERNEL(popc_imad, OP_POPC_IMAD) KERNEL(vabs_lop3, OP_VABS_LOP3) KERNEL(vabs_imad, OP_VABS_IMAD)

struct Case { const char* name; void (*k)(uint32_t*, long long*, uint32_t, uint32_t); int perSlot; };

int main() {
    int dev = 0;
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, dev));
    const int sms = p.multiProcessorCount, ctas = sms * CTAS_PER_SM;
    uint32_t* out;
    long long* cyc;
    CK(cudaMalloc(&out, 64));
    CK(cudaMalloc(&cyc, ctas * sizeof(long long)));
#define C(n, per) {#n, k_##n, per}
    std::vector<Case> cases = {C(lop3, 1), C(iadd3, 1), C(imad, 1), C(shf, 1), C(prmt, 1), C(popc, 1), C(vabsdiff4, 1), C(vsad4, 1),
                               C(vimnmx3_s16x2, 1), C(vimnmx_u16x2, 1), C(dp4a, 1), C(dp2a, 1), C(flo, 1), C(brev, 1), C(imnmx, 1), C(shl, 1),
                               C(shfl, 1), C(ballot, 2), C(setp_sel, 2), C(lop3_imad, 2), C(popc_lop3, 2), C(popc_imad, 2), C(vabs_lop3, 2),
                               C(vabs_imad, 2)};
    std::vector<long long> h(ctas);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    printf("{\"device\": \"%s\", \"sms\": %d, \"warps_per_sm\": %d, \"chains_per_thread\": %d, \"unit\": \"warp instructions / clk / SM\", \"rates\": {",
           p.name, sms, CTAS_PER_SM * THREADS / 32, CHAINS);
    bool first = true;
    for (const Case& cs : cases) {
        cs.k<<<ctas, THREADS>>>(out, cyc, 12345u, 777u);  // warm-up
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        cs.k<<<ctas, THREADS>>>(out, cyc, 12345u, 777u);
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        CK(cudaMemcpy(h.data(), cyc, ctas * sizeof(long long), cudaMemcpyDeviceToHost));
        long long mx = 0;
        for (long long v : h) mx = v > mx ? v : mx;
        const double warpInstrPerSm = (double)ITERS * CHAINS * cs.perSlot * (THREADS / 32) * CTAS_PER_SM;
        printf("%s\"%s\": {\"per_clk_per_sm\": %.2f, \"ms\": %.4f, \"ptx_ops_per_slot\": %d}", first ? "" : ", ", cs.name, warpInstrPerSm / (double)mx, ms,
               cs.perSlot);
        first = false;
    }
    printf("}}\n");
    return 0;
}
