// csrc/k_describe.cu — IC_Angle (reference src/ORBextractor.cc:164-191), computeOrbDescriptor (:195-233), the
// scale-back of keypoint coordinates (:1794-1803) and the level-major output order of operator() (:1701-1809).
//
// One warp per kept keypoint.
//   angle: lanes span the 31 columns of the circular patch, the warp walks its 31 rows (umax gives the row
//          half-width, :540-555); m10 / m01 are exact int32 sums, reduced by shuffles; cv::fastAtan2 is restated
//          as its float32 polynomial with every multiply / add / divide individually rounded (no FMA).
//   BRIEF: lane i produces descriptor byte i from 16 taps of the blurred level at the pattern points rotated by
//          the keypoint angle: cvRound(x*b + y*a), cvRound(x*a - y*b) with separately rounded products and
//          round-half-even (__float2int_rn), bit k = tap(2k) < tap(2k+1).
//   cos / sin: the reference calls the float overloads, i.e. glibc cosf / sinf.  Those are not correctly rounded
//          (about 1.3 % of arguments differ by 1 ulp from the rounded exact value), so they are restated here as
//          glibc does them (sysdeps/ieee754/flt-32/s_sincosf.h since 2.28): reduction by multiples of pi/2 and
//          a degree-7 / degree-8 polynomial pair evaluated in double.  tests/test_oracle_cv2.py checks the same
//          restatement against the host libm on millions of angles.
// Output position = (number of keypoints of lower levels of the same frame) + index within the level.
#include "orbx_internal.h"

namespace orbx {

__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
__constant__ signed char c_pattern[1024] = {
#include "orb_pattern_31.inc"
};

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

// grid (ceil(slotsTotal / DESC_WARPS), batch)
__global__ void __launch_bounds__(DESC_WARPS * 32) k_orient_describe(const __grid_constant__ FrameGeom g,
                                                                     const int* __restrict__ lvlCnt,
                                                                     const uint32_t* __restrict__ lvlKeys,
                                                                     orbx_keypoint* __restrict__ kps,
                                                                     uint8_t* __restrict__ desc, int cap,
                                                                     int* __restrict__ counts, float* __restrict__ lvlAngles,
                                                                     int* dStatus) {
    __shared__ signed char pat[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) pat[i] = c_pattern[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, f = blockIdx.y;
    const int slot = blockIdx.x * DESC_WARPS + (threadIdx.x >> 5);
    if (slot >= g.slotsTotal) return;
    int level = 0;
    while (level + 1 < g.nlevels && slot >= g.lv[level + 1].slotBase) ++level;
    const LevelGeom& L = g.lv[level];
    const int j = slot - L.slotBase;
    const int* cnt = lvlCnt + (long long)f * g.nlevels;
    int before = 0, total = 0;
    for (int l = 0; l < g.nlevels; ++l) {
        const int c = cnt[l];
        if (l < level) before += c;
        total += c;
    }
    if (slot == 0 && lane == 0) {
        counts[f] = min(total, cap);
        if (total > cap) atomicExch(dStatus, ORBX_E_OVERFLOW);
    }
    if (j >= cnt[level]) return;
    const uint32_t key = lvlKeys[(long long)f * g.slotsTotal + slot];
    const int x = key_x(key), y = key_y(key);

    // ---- IC_Angle on the unblurred level ----
    const uint8_t* center = L.plane + (long long)f * L.planeStride + (long long)(y + EDGE) * L.pitch + (x + XOFF);
    const int u = lane - HALF_PATCH;
    int m10 = 0, m01 = 0;
    for (int v = -HALF_PATCH; v <= HALF_PATCH; ++v) {
        const int d = c_umax[v < 0 ? -v : v];
        if (lane < PATCH && u >= -d && u <= d) {
            const int val = center[v * L.pitch + u];
            m10 += u * val;
            m01 += v * val;
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
    if (o >= cap) return;

    // ---- steered BRIEF on the blurred level ----
    const float factorPI = 0.01745329238474369f;  // (float)(CV_PI / 180.f)
    float a, b;
    glibc_sincosf(__fmul_rn(angle, factorPI), &b, &a);  // a = cos, b = sin
    const uint8_t* bc = L.blur + (long long)f * L.blurStride + (long long)y * L.bpitch + x;
    const signed char* p = pat + lane * 32;
    int val = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float x0 = (float)p[4 * k], y0 = (float)p[4 * k + 1], x1 = (float)p[4 * k + 2], y1 = (float)p[4 * k + 3];
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a)));
        const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a)));
        const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)));
        const int t0 = bc[r0 * L.bpitch + c0], t1 = bc[r1 * L.bpitch + c1];
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

int launch_describe(const FrameGeom& g, int batch, const int* lvlCnt, const uint32_t* lvlKeys, orbx_keypoint* dKps,
                    uint8_t* dDesc, int cap, int* dCounts, float* dLvlAngles, int* dStatus, cudaStream_t s) {
    dim3 grid((g.slotsTotal + DESC_WARPS - 1) / DESC_WARPS, batch);
    k_orient_describe<<<grid, DESC_WARPS * 32, 0, s>>>(g, lvlCnt, lvlKeys, dKps, dDesc, cap, dCounts, dLvlAngles, dStatus);
    count_launch();
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx
