// Scalar math shared by the two extractors' orientation / descriptor kernels (orb_keypoints.cu, orb_cv.cu):
// cv::fastAtan2 and glibc's sinf / cosf, both evaluated without FMA.
#pragma once
#include <cuda_runtime.h>

namespace rgbdfe {

// cvRound (round half to even) for |v| < 2^22 without the XU pipe: adding 1.5 * 2^23 leaves the integer in the low mantissa bits
__device__ __forceinline__ int cv_round_fast(float v) { return __float_as_int(__fadd_rn(v, 12582912.0f)) - 0x4B400000; }

__device__ __forceinline__ float fast_atan2_deg(float y, float x)
{
    // cv::fastAtan2 (scalar), evaluated without FMA
    const float p1 = 0.9997878412794807f * (float)(180 / 3.14159265358979323846);
    const float p3 = -0.3258083974640975f * (float)(180 / 3.14159265358979323846);
    const float p5 = 0.1555786518463281f * (float)(180 / 3.14159265358979323846);
    const float p7 = -0.04432655554792128f * (float)(180 / 3.14159265358979323846);
    const float eps = (float)2.2204460492503131e-16;
    float ax = fabsf(x), ay = fabsf(y);
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
    if (x < 0)
        a = __fsub_rn(180.f, a);
    if (y < 0)
        a = __fsub_rn(360.f, a);
    return a;
}

// glibc >= 2.28 sinf/cosf for |x| < 120 (double arithmetic, no FMA).
__device__ __forceinline__ float sincos_poly(double x, double x2, int n, int tab)
{
    const double sgn = tab ? -1.0 : 1.0;
    if ((n & 1) == 0) {
        const double s1 = -0x1.555545995a603p-3, s2 = 0x1.1107605230bc4p-7, s3 = -0x1.994eb3774cf24p-13;
        double x3 = __dmul_rn(x, x2);
        double t1 = __dadd_rn(s2, __dmul_rn(x2, s3));
        double x7 = __dmul_rn(x3, x2);
        double s = __dadd_rn(x, __dmul_rn(x3, s1));
        return (float)__dadd_rn(s, __dmul_rn(x7, t1));
    } else {
        const double c0 = sgn * 0x1p0, c1 = sgn * -0x1.ffffffd0c621cp-2, c2 = sgn * 0x1.55553e1068f19p-5,
                     c3 = sgn * -0x1.6c087e89a359dp-10, c4 = sgn * 0x1.99343027bf8c3p-16;
        double x4 = __dmul_rn(x2, x2);
        double t2 = __dadd_rn(c3, __dmul_rn(x2, c4));
        double t1 = __dadd_rn(c0, __dmul_rn(x2, c1));
        double x6 = __dmul_rn(x4, x2);
        double c = __dadd_rn(t1, __dmul_rn(x4, c2));
        return (float)__dadd_rn(c, __dmul_rn(x6, t2));
    }
}

__device__ __forceinline__ unsigned abstop12(float x) { return (__float_as_uint(x) >> 20) & 0x7ff; }

__device__ __forceinline__ void sincosf_glibc(float y, float* sn, float* cs)
{
    double x = (double)y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        double x2 = __dmul_rn(x, x);
        if (abstop12(y) < abstop12(0x1p-12f)) {
            *sn = y;
            *cs = 1.0f;
            return;
        }
        *sn = sincos_poly(x, x2, 0, 0);
        *cs = sincos_poly(x, x2, 1, 0);
        return;
    }
    // reduce_fast: |y| < 120 guaranteed (angle in [0, 2*pi))
    double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
    int n = ((int)r + 0x800000) >> 24;
    x = __dsub_rn(x, __dmul_rn((double)n, 0x1.921FB54442D18p0));
    const double sg = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    const int tab = (n & 2) ? 1 : 0;
    double xs = __dmul_rn(x, sg), x2 = __dmul_rn(x, x);
    *sn = sincos_poly(xs, x2, n, tab);
    *cs = sincos_poly(xs, x2, n ^ 1, tab);
}

} // namespace rgbdfe
