// csrc/k_describe.cu — IC_Angle (reference src/ORBextractor.cc:164-191), computeOrbDescriptor (:195-233), the
// scale-back of keypoint coordinates (:1794-1803) and the level-major output order of operator() (:1701-1809).
//
// One warp per kept keypoint (a warp walks DESC_PER_WARP slots).
//   angle: the 31 x 31 patch is read as 32-bit words (4 rows x 8 words per step, unaligned rows realigned with a
//          funnel shift) and dotted with two byte-coefficient tables by dp4a: u inside the circle (umax, :540-555) for
//          m10, v for m01.  The sums are exact int32, reduced by shuffles; cv::fastAtan2 is restated as its float32
//          polynomial with every multiply / add / divide individually rounded (no FMA).
//   BRIEF: the 37 x 37 neighbourhood of the BLURRED level (the rotated pattern reaches +-18) is staged into a per-warp
//          shared-memory patch with coalesced word loads; lane i then produces descriptor byte i from 16 taps at the
//          pattern points rotated by the keypoint angle: cvRound(x*b + y*a), cvRound(x*a - y*b) with separately
//          rounded products; cvRound (round half to even) is done with the 1.5 * 2^23 bias trick, whose result feeds
//          the address multiply-add directly; bit k = tap(2k) < tap(2k+1).
//   cos / sin: the reference calls the float overloads, i.e. glibc cosf / sinf.  Those are not correctly rounded
//          (about 1.3 % of arguments differ by 1 ulp from the rounded exact value), so they are restated here as
//          glibc does them (sysdeps/ieee754/flt-32/s_sincosf.h since 2.28): reduction by multiples of pi/2 and
//          a degree-7 / degree-8 polynomial pair evaluated in double.  tests/test_oracle_cv2.py checks the same
//          restatement against the host libm on millions of angles.
// Output position = (number of keypoints of lower levels of the same frame) + index within the level.
#include "orbx_internal.h"
#include "tma.cuh"

namespace orbx {

__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

// cv::fastAtan2 (degrees in [0, 360)), float32, no contraction
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float scale = 57.295780181884766f;  // (float)(180 / CV_PI)
    const float p1 = __fmul_rn(0.9997878412794807f, scale), p3 = __fmul_rn(-0.3258083974640975f, scale);
    const float p5 = __fmul_rn(0.1555786518463281f, scale), p7 = __fmul_rn(-0.04432655554792128f, scale);
    const float eps = 2.220446049250313e-16f;  // (float)DBL_EPSILON
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0.f) a = __fsub_rn(180.f, a);
    if (y < 0.f) a = __fsub_rn(360.f, a);
    return a;
}

// glibc sinf_poly: n even -> sine polynomial, n odd -> cosine polynomial; neg selects __sincosf_table[1]
__device__ __forceinline__ float glibc_sinf_poly(double x, double x2, int n, bool neg) {
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    const double sg = neg ? -1.0 : 1.0;
    const double C0 = sg * 0x1p0, C1 = sg * -0x1.ffffffd0c621cp-2, C2 = sg * 0x1.55553e1068f19p-5;
    const double C3 = sg * -0x1.6c087e89a359dp-10, C4 = sg * 0x1.99343027bf8c3p-16;
    if ((n & 1) == 0) {
        const double x3 = __dmul_rn(x, x2);
        const double s1 = __dadd_rn(S2, __dmul_rn(x2, S3));
        const double x7 = __dmul_rn(x3, x2);
        const double s = __dadd_rn(x, __dmul_rn(x3, S1));
        return (float)__dadd_rn(s, __dmul_rn(x7, s1));
    }
    const double x4 = __dmul_rn(x2, x2);
    const double c2 = __dadd_rn(C3, __dmul_rn(x2, C4));
    const double c1 = __dadd_rn(C0, __dmul_rn(x2, C1));
    const double x6 = __dmul_rn(x4, x2);
    const double c = __dadd_rn(c1, __dmul_rn(x4, C2));
    return (float)__dadd_rn(c, __dmul_rn(x6, c2));
}

// glibc sinf / cosf for |y| < 120 (the angle is in [0, 2 pi])
__device__ __forceinline__ void glibc_sincosf(float y, float* sn, float* cs) {
    const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ffu;
    const double x = (double)y;
    if (top < 0x3f4u) {  // abstop12(y) < abstop12(pi/4)
        if (top < 0x398u) {  // abstop12(0x1p-12f)
            *sn = y;
            *cs = 1.0f;
            return;
        }
        const double x2 = __dmul_rn(x, x);
        *sn = glibc_sinf_poly(x, x2, 0, false);
        *cs = glibc_sinf_poly(x, x2, 1, false);
        return;
    }
    const double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
    const int n = (__double2int_rz(r) + 0x800000) >> 24;
    const double xr = __dadd_rn(x, -__dmul_rn((double)n, 0x1.921FB54442D18p0));
    const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    const bool neg = (n & 2) != 0;
    const double xs = __dmul_rn(xr, sgn), x2 = __dmul_rn(xr, xr);
    *sn = glibc_sinf_poly(xs, x2, n, neg);
    *cs = glibc_sinf_poly(xs, x2, n ^ 1, neg);
}

constexpr int DESC_WARPS = 8;
#ifndef ORBX_DESC_PER_WARP
#define ORBX_DESC_PER_WARP 4
#endif
constexpr int DESC_PER_WARP = ORBX_DESC_PER_WARP;   // keypoint slots a warp walks (amortises the table set-up of the CTA)
constexpr int BP_HALF = 18;        // the rotated pattern reaches |coordinate| <= cvRound(18.38) = 18
constexpr int BP_ROWS = 2 * BP_HALF + 1;  // 37 rows; a row is 48 bytes read from an 8-byte aligned address (6 x 64-bit loads)
constexpr int BP_WORDS = 14, BP_STRIDE = BP_WORDS * 4;  // shared row stride: 56 bytes (8-byte aligned rows for cp.async)

// Tables the kernel copies into shared memory, filled once per process by init_describe_tables():
//   d_patF  : the 512 pattern points as floats, laid out [pair k of the lane][lane] as (x0, y0, x1, y1), once per variant of
//             bit_pattern_31_ (orbx_extractor_set_pattern)
//   d_coefU : [row r = v + 15][word j] 4 signed bytes  u = 4j + q - 15  inside the circular patch (|u| <= umax[|v|]), else 0
//   d_coefV : same mask, bytes = v
__device__ float4 d_patF[2][8 * 32];  // [pattern variant]: 0 = the reference as it compiles (element 96 = -3), 1 = canonical
__device__ uint32_t d_coefU[32 * 8];
__device__ uint32_t d_coefV[32 * 8];

__device__ __forceinline__ int dp4a_us(uint32_t pixels, uint32_t coef, int acc) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pixels), "r"(coef), "r"(acc));
    return d;
}

// grid (ceil(slotsTotal / (DESC_WARPS * DESC_PER_WARP)), batch)
template <int perWarp>
__global__ void __launch_bounds__(DESC_WARPS * 32) k_orient_describe(const __grid_constant__ FrameGeom g,
                                                                     const int* __restrict__ lvlCnt,
                                                                     const uint32_t* __restrict__ lvlKeys,
                                                                     orbx_keypoint* __restrict__ kps,
                                                                     uint8_t* __restrict__ desc, int cap,
                                                                     int* __restrict__ counts, float* __restrict__ lvlAngles,
                                                                     int* dStatus) {
    // perWarp: keypoint slots a warp walks — DESC_PER_WARP in batches, 1 for single frames, where the walk is the critical path
    __shared__ float4 patF[8 * 32];
    __shared__ uint32_t coefU[32 * 8], coefV[32 * 8];
    __shared__ __align__(16) uint32_t bpatch[DESC_WARPS][BP_ROWS * BP_WORDS + 2];
    __shared__ int sBefore[ORBX_MAX_LEVELS + 1];
    __shared__ unsigned char sLevelOf[DESC_WARPS * DESC_PER_WARP];  // pyramid level of each of the CTA's slots
    if (g.pdl) griddep_launch_dependents();
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        patF[i] = d_patF[g.patternVariant & 1][i];
        coefU[i] = d_coefU[i];
        coefV[i] = d_coefV[i];
    }
    if (g.pdl) griddep_wait();  // the tables above are constants; everything below reads what the chain produced
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, f = blockIdx.y;
    const int* cnt = lvlCnt + (long long)f * g.nlevels;
    if (threadIdx.x >= 32 && threadIdx.x < 32 + DESC_WARPS * perWarp) {
        const int slot = blockIdx.x * (DESC_WARPS * perWarp) + (threadIdx.x - 32);
        int level = 0;
        while (level + 1 < g.nlevels && slot >= g.lv[level + 1].slotBase) ++level;
        sLevelOf[threadIdx.x - 32] = (unsigned char)level;
    }
    if (threadIdx.x == 0) {  // exclusive prefix of the per-level counts = output position of each level's first keypoint
        int total = 0;
        for (int l = 0; l < g.nlevels; ++l) {
            sBefore[l] = total;
            total += cnt[l];
        }
        sBefore[g.nlevels] = total;
        if (blockIdx.x == 0) {
            counts[f] = min(total, cap);
            if (total > cap) atomicExch(dStatus, ORBX_E_OVERFLOW);
        }
    }
    __syncthreads();
    uint32_t* bp = bpatch[warp];
    const uint8_t* bp8 = reinterpret_cast<const uint8_t*>(bp);

    for (int it = 0; it < perWarp; ++it) {
        const int slot = (blockIdx.x * perWarp + it) * DESC_WARPS + warp;
        if (slot >= g.slotsTotal) break;
        const int level = sLevelOf[it * DESC_WARPS + warp];
        const LevelGeom& L = g.lv[level];
        const int j = slot - L.slotBase;
        const int before = sBefore[level];
        if (j >= sBefore[level + 1] - before) continue;
        const uint32_t key = lvlKeys[(long long)f * g.slotsTotal + slot];
        const int x = key_x(key), y = key_y(key);

        // ---- blurred 37 x 37 patch: rows of 48 bytes from an 8-byte aligned address, copied global -> shared with
        //      cp.async (no registers).  The copies are issued here, before IC_Angle, so their latency overlaps the moment
        //      sums and the arctangent (the patch address needs x, y only); lanes = 5 rows x 6 double words per step,
        //      8 steps cover rows 0..39 (rows >= 37 are skipped) ----
        const uint8_t* brow = L.blur + (long long)(f + g.f0) * L.blurStride + (long long)(y - BP_HALF) * L.bpitch + (x - BP_HALF);
        const int mis = (int)(reinterpret_cast<uintptr_t>(brow) & 7);
        __syncwarp();  // the previous keypoint's taps are done with the patch
        if (lane < 30) {
            const int rsub = lane / 6, q = lane - rsub * 6;
            const uint2* bw = reinterpret_cast<const uint2*>(reinterpret_cast<uintptr_t>(brow) & ~(uintptr_t)7);
            const int bpo = L.bpitch >> 3;
            const uint32_t dst0 = (uint32_t)__cvta_generic_to_shared(bp) + (uint32_t)(rsub * BP_STRIDE + q * 8);
#pragma unroll
            for (int r5 = 0; r5 < 8; ++r5) {
                const int r = r5 * 5 + rsub;
                if (r < BP_ROWS)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst0 + (uint32_t)(r5 * 5 * BP_STRIDE)),
                                 "l"(bw + r * bpo + q)
                                 : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");

        // ---- IC_Angle on the unblurred level: lanes = (4 rows) x (8 words of 4 pixels), dp4a against the u / v tables ----
        int m10 = 0, m01 = 0;
        {
            const uint8_t* p0 = L.plane + (long long)(f + g.f0) * L.planeStride + (long long)(y - HALF_PATCH + EDGE) * L.pitch +
                                (x - HALF_PATCH + XOFF) + 4 * (lane & 7);
            const uint32_t sh = (uint32_t)(reinterpret_cast<uintptr_t>(p0) & 3) * 8;
            const uint32_t* w0 = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(p0) & ~(uintptr_t)3);
            const int rsub = lane >> 3, pitchq = L.pitch >> 2;
#pragma unroll
            for (int r4 = 0; r4 < 8; ++r4) {
                const int r = r4 * 4 + rsub;  // row 31 exists in the padded plane; its coefficients are zero
                const uint32_t* w = w0 + r * pitchq;
                const uint32_t px = __funnelshift_r(__ldg(w), __ldg(w + 1), sh);
                m10 = dp4a_us(px, coefU[r * 8 + (lane & 7)], m10);
                m01 = dp4a_us(px, coefV[r * 8 + (lane & 7)], m01);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m10 += __shfl_xor_sync(0xffffffffu, m10, o);
            m01 += __shfl_xor_sync(0xffffffffu, m01, o);
        }
        const float angle = fast_atan2_deg((float)m01, (float)m10);
        if (lvlAngles && lane == 0) lvlAngles[(long long)f * g.slotsTotal + slot] = angle;
        const int o = before + j;
        if (o >= cap) {  // output overflow (reported through dStatus): drain the copies before the patch is reused
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            continue;
        }

        // ---- the staged patch has landed ----
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();

        // ---- steered BRIEF: lane i -> byte i, 16 rotated taps read from the staged patch ----
        const float factorPI = 0.01745329238474369f;  // (float)(CV_PI / 180.f)
        float a, b;
        glibc_sincosf(__fmul_rn(angle, factorPI), &b, &a);  // a = cos, b = sin
        // cvRound(v) = bits(v + 1.5 * 2^23) - bits(1.5 * 2^23) (round half to even, |v| < 2^22); the two biases and the
        // patch origin are folded into one constant of the byte address  row * 44 + col
        const float MAGIC = 12582912.0f;
        const uint32_t base = 0u - 0x4B400000u * (uint32_t)(BP_STRIDE + 1) + (uint32_t)(BP_HALF * BP_STRIDE + BP_HALF + mis);  // mod 2^32
        int val = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float4 p = patF[k * 32 + lane];
            const float r0 = __fadd_rn(__fadd_rn(__fmul_rn(p.x, b), __fmul_rn(p.y, a)), MAGIC);
            const float c0 = __fadd_rn(__fsub_rn(__fmul_rn(p.x, a), __fmul_rn(p.y, b)), MAGIC);
            const float r1 = __fadd_rn(__fadd_rn(__fmul_rn(p.z, b), __fmul_rn(p.w, a)), MAGIC);
            const float c1 = __fadd_rn(__fsub_rn(__fmul_rn(p.z, a), __fmul_rn(p.w, b)), MAGIC);
            const int t0 = bp8[__float_as_uint(r0) * (uint32_t)BP_STRIDE + __float_as_uint(c0) + base];
            const int t1 = bp8[__float_as_uint(r1) * (uint32_t)BP_STRIDE + __float_as_uint(c1) + base];
            val |= (t0 < t1 ? 1 : 0) << k;
        }
        desc[((long long)f * cap + o) * 32 + lane] = (uint8_t)val;

        // ---- keypoint record (cv::KeyPoint layout), coordinates scaled back to level 0 (:1794-1803) ----
        if (lane < 7) {
            float fx = (float)x, fy = (float)y;
            if (level != 0) {
                fx = __fmul_rn(fx, L.scale);
                fy = __fmul_rn(fy, L.scale);
            }
            uint32_t w;
            switch (lane) {
                case 0: w = __float_as_uint(fx); break;
                case 1: w = __float_as_uint(fy); break;
                case 2: w = __float_as_uint(L.size); break;
                case 3: w = __float_as_uint(angle); break;
                case 4: w = __float_as_uint((float)key_resp(key)); break;
                case 5: w = (uint32_t)level; break;
                default: w = 0xFFFFFFFFu; break;  // class_id = -1
            }
            reinterpret_cast<uint32_t*>(kps + (long long)f * cap + o)[lane] = w;
        }
    }
}

static int init_describe_tables() {
    static bool done[16] = {};
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    if (dev < 16 && done[dev]) return ORBX_OK;
    static const signed char pat[1024] = {
#include "orb_pattern_31.inc"
    };
    static const int umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
    // variant 0 (ORBX_PATTERN_REFERENCE): the reference as it compiles — `-6,-2,VX_FAILURE <newline> -2,11` (src/ORBextractor.cc:261-262) is
    // the expression VX_FAILURE - 2 = -3 in element 96; variant 1 (ORBX_PATTERN_CANONICAL): the table of OpenCV / upstream ORB-SLAM2
    float4 pf[2][256];
    for (int variant = 0; variant < 2; ++variant)
        for (int lane = 0; lane < 32; ++lane)
            for (int k = 0; k < 8; ++k) {
                const int e = lane * 32 + 4 * k;  // pair 8 * lane + k: (x0, y0, x1, y1)
                float v[4];
                for (int q = 0; q < 4; ++q) v[q] = (float)((variant == 0 && e + q == 96) ? -3 : (int)pat[e + q]);
                pf[variant][k * 32 + lane] = make_float4(v[0], v[1], v[2], v[3]);
            }
    uint32_t cu[256], cv[256];
    for (int r = 0; r < 32; ++r)
        for (int j = 0; j < 8; ++j) {
            uint32_t wu = 0, wv = 0;
            for (int q = 0; q < 4; ++q) {
                const int c = 4 * j + q, u = c - HALF_PATCH, v = r - HALF_PATCH;
                const bool in = r <= 30 && c <= 30 && std::abs(u) <= umax[std::abs(v)];
                wu |= (uint32_t)(uint8_t)(signed char)(in ? u : 0) << (8 * q);
                wv |= (uint32_t)(uint8_t)(signed char)(in ? v : 0) << (8 * q);
            }
            cu[r * 8 + j] = wu;
            cv[r * 8 + j] = wv;
        }
    ORBX_CUDA(cudaMemcpyToSymbol(d_patF, pf, sizeof pf));
    ORBX_CUDA(cudaMemcpyToSymbol(d_coefU, cu, sizeof cu));
    ORBX_CUDA(cudaMemcpyToSymbol(d_coefV, cv, sizeof cv));
    if (dev < 16) done[dev] = true;
    return ORBX_OK;
}

int launch_describe(const FrameGeom& g, int batch, const int* lvlCnt, const uint32_t* lvlKeys, orbx_keypoint* dKps,
                    uint8_t* dDesc, int cap, int* dCounts, float* dLvlAngles, int* dStatus, cudaStream_t s) {
    int rc = init_describe_tables();
    if (rc) return rc;
    const int perWarp = g.pdl ? 1 : DESC_PER_WARP;
    const int perCta = DESC_WARPS * perWarp;
    dim3 grid((g.slotsTotal + perCta - 1) / perCta, batch);
    if (perWarp == 1)
        ORBX_CUDA(launch_kernel(k_orient_describe<1>, grid, dim3(DESC_WARPS * 32), 0, s, g.pdl && !g.plainLaunch, g, lvlCnt, lvlKeys, dKps, dDesc, cap, dCounts,
                                dLvlAngles, dStatus));
    else
        ORBX_CUDA(launch_kernel(k_orient_describe<DESC_PER_WARP>, grid, dim3(DESC_WARPS * 32), 0, s, g.pdl && !g.plainLaunch, g, lvlCnt, lvlKeys, dKps, dDesc, cap,
                                dCounts, dLvlAngles, dStatus));
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
