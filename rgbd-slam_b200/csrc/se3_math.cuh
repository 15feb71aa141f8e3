// Device arithmetic of RansacSE3 (Solver/SolverSE3.cpp:161-297):
//   tfc_fit     — pcl::TransformationFromCorrespondences (incremental weighted mean /
//                 covariance, weights 1/(z1*z2)) + Eigen::JacobiSVD<Matrix3f> (two-sided
//                 Jacobi, Eigen 3.3 sweep order and thresholds) -> rigid transform, FP32
//   mahal2      — errorFunction2: squared Mahalanobis distance with a 3x3 LLT, FP64
// Arithmetic contract (DESIGN.md "FP contract"): no FMA contraction (this TU is compiled
// with -fmad=false), IEEE div/sqrt, 3-term sums associate as a + (b + c), 4x4*vec rows
// left to right.  These orders are what the CPU oracle evaluates, so results are bit-equal.
#pragma once
#include "common.cuh"
#include <cfloat>

namespace rgbdfe {

template <class T>
__device__ __forceinline__ T smax(T a, T b) { return (a < b) ? b : a; } // std::max

struct Rot {
    float c, s;
};

__device__ __forceinline__ void make_jacobi(float x, float y, float z, Rot& r)
{
    float deno = 2.0f * fabsf(y);
    if (deno < FLT_MIN) {
        r.c = 1.0f;
        r.s = 0.0f;
        return;
    }
    float tau = (x - z) / deno;
    float w = sqrtf(tau * tau + 1.0f);
    float t;
    if (tau > 0.0f)
        t = 1.0f / (tau + w);
    else
        t = 1.0f / (tau - w);
    float sign_t = t > 0.0f ? 1.0f : -1.0f;
    float n = 1.0f / sqrtf(t * t + 1.0f);
    r.s = -sign_t * (y / fabsf(y)) * fabsf(t) * n;
    r.c = n;
}

__device__ __forceinline__ void rot_rows(float* M, int p, int q, Rot j)
{
#pragma unroll
    for (int i = 0; i < 3; i++) {
        float x = M[p * 3 + i], y = M[q * 3 + i];
        M[p * 3 + i] = j.c * x + j.s * y;
        M[q * 3 + i] = -j.s * x + j.c * y;
    }
}
__device__ __forceinline__ void rot_cols(float* M, int p, int q, Rot j)
{
#pragma unroll
    for (int i = 0; i < 3; i++) {
        float x = M[i * 3 + p], y = M[i * 3 + q];
        M[i * 3 + p] = j.c * x + j.s * y;
        M[i * 3 + q] = -j.s * x + j.c * y;
    }
}

// Eigen::JacobiSVD<Matrix3f>(ComputeFullU | ComputeFullV); row-major 3x3.
__device__ inline void jacobi_svd3(const float* A, float* U, float* S, float* V)
{
    const float precision = 2.0f * FLT_EPSILON;
    const float considerAsZero = FLT_MIN;
    float scale = 0.0f;
#pragma unroll
    for (int i = 0; i < 9; i++) {
        float a = fabsf(A[i]);
        if (a > scale)
            scale = a;
    }
    if (scale == 0.0f)
        scale = 1.0f;
    float W[9];
#pragma unroll
    for (int i = 0; i < 9; i++) {
        W[i] = A[i] / scale;
        U[i] = V[i] = (i % 4 == 0) ? 1.0f : 0.0f;
    }
    float maxDiag = smax(fabsf(W[0]), smax(fabsf(W[4]), fabsf(W[8])));
    bool finished = false;
    for (int sweep = 0; !finished && sweep < 64; sweep++) {
        finished = true;
#pragma unroll
        for (int p = 1; p < 3; ++p) {
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                if (q >= p)
                    continue;
                float threshold = smax(considerAsZero, precision * maxDiag);
                if (fabsf(W[p * 3 + q]) > threshold || fabsf(W[q * 3 + p]) > threshold) {
                    finished = false;
                    float m00 = W[p * 3 + p], m01 = W[p * 3 + q], m10 = W[q * 3 + p], m11 = W[q * 3 + q];
                    Rot rot1;
                    float t = m00 + m11;
                    float d = m10 - m01;
                    if (fabsf(d) < FLT_MIN) {
                        rot1.s = 0.0f;
                        rot1.c = 1.0f;
                    } else {
                        float u = t / d;
                        float tmp = sqrtf(1.0f + u * u);
                        rot1.s = 1.0f / tmp;
                        rot1.c = u / tmp;
                    }
                    float n00 = rot1.c * m00 + rot1.s * m10, n01 = rot1.c * m01 + rot1.s * m11;
                    float n11 = -rot1.s * m01 + rot1.c * m11;
                    Rot jr;
                    make_jacobi(n00, n01, n11, jr);
                    Rot jrt = { jr.c, -jr.s };
                    Rot jl = { rot1.c * jrt.c - rot1.s * jrt.s, rot1.c * jrt.s + rot1.s * jrt.c };
                    rot_rows(W, p, q, jl);
                    rot_cols(U, p, q, jl);
                    rot_cols(W, p, q, jrt);
                    rot_cols(V, p, q, jrt);
                    maxDiag = smax(maxDiag, smax(fabsf(W[p * 3 + p]), fabsf(W[q * 3 + q])));
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        float a = W[i * 3 + i];
        S[i] = fabsf(a);
        if (a < 0.0f) {
            U[0 * 3 + i] = -U[0 * 3 + i];
            U[1 * 3 + i] = -U[1 * 3 + i];
            U[2 * 3 + i] = -U[2 * 3 + i];
        }
    }
#pragma unroll
    for (int i = 0; i < 3; i++)
        S[i] *= scale;
    // sort the singular values in descending order (Eigen: for each i the first maximum of S[i..2] is swapped
    // into place, stop at a zero maximum) — written out for n = 3 so that U, V, S stay in registers
    auto swap_cols = [&](int a, int b) {
        float t = S[a];
        S[a] = S[b];
        S[b] = t;
#pragma unroll
        for (int r = 0; r < 3; r++) {
            t = U[r * 3 + a];
            U[r * 3 + a] = U[r * 3 + b];
            U[r * 3 + b] = t;
            t = V[r * 3 + a];
            V[r * 3 + a] = V[r * 3 + b];
            V[r * 3 + b] = t;
        }
    };
    {
        int pos = 0;
        float mx = S[0];
        if (S[1] > mx) {
            mx = S[1];
            pos = 1;
        }
        if (S[2] > mx) {
            mx = S[2];
            pos = 2;
        }
        if (mx == 0.0f)
            return;
        if (pos == 1)
            swap_cols(0, 1);
        else if (pos == 2)
            swap_cols(0, 2);
    }
    {
        if (S[2] > S[1]) {
            swap_cols(1, 2);
        } else if (S[1] == 0.0f)
            return;
    }
}

__device__ __forceinline__ float det3(const float* m)
{
    return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
}

// Running state of pcl::TransformationFromCorrespondences.
struct Tfc {
    float accW;
    float mean1[3], mean2[3];
    float cov[9];
    __device__ __forceinline__ void reset()
    {
        accW = 0.0f;
#pragma unroll
        for (int i = 0; i < 3; i++)
            mean1[i] = mean2[i] = 0.0f;
#pragma unroll
        for (int i = 0; i < 9; i++)
            cov[i] = 0.0f;
    }
    __device__ __forceinline__ void add(const rgbdfe_point3f& from, const rgbdfe_point3f& to)
    {
        if (isnan(from.z) || isnan(to.z))
            return;
        float weight = 1.0f / (from.z * to.z);
        if (weight == 0.0f)
            return;
        accW += weight;
        float alpha = weight / accW;
        float d1[3] = { from.x - mean1[0], from.y - mean1[1], from.z - mean1[2] };
        float d2[3] = { to.x - mean2[0], to.y - mean2[1], to.z - mean2[2] };
        float oma = 1.0f - alpha;
#pragma unroll
        for (int i = 0; i < 3; i++) {
            float ad2 = alpha * d2[i];
#pragma unroll
            for (int j = 0; j < 3; j++)
                cov[i * 3 + j] = oma * (cov[i * 3 + j] + ad2 * d1[j]);
        }
#pragma unroll
        for (int i = 0; i < 3; i++) {
            mean1[i] += alpha * d1[i];
            mean2[i] += alpha * d2[i];
        }
    }
    __device__ inline void transform(float* T) const
    {
        float U[9], S[3], V[9];
        jacobi_svd3(cov, U, S, V);
        float s2 = (det3(U) * det3(V) < 0.0f) ? -1.0f : 1.0f;
        float R[9];
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = 0; j < 3; j++)
                R[i * 3 + j] = U[i * 3 + 0] * V[j * 3 + 0] + (U[i * 3 + 1] * V[j * 3 + 1] + (U[i * 3 + 2] * s2) * V[j * 3 + 2]);
#pragma unroll
        for (int i = 0; i < 3; i++) {
            float rm = R[i * 3 + 0] * mean1[0] + (R[i * 3 + 1] * mean1[1] + R[i * 3 + 2] * mean1[2]);
            T[i * 4 + 0] = R[i * 3 + 0];
            T[i * 4 + 1] = R[i * 3 + 1];
            T[i * 4 + 2] = R[i * 3 + 2];
            T[i * 4 + 3] = mean2[i] - rm;
        }
        T[12] = T[13] = T[14] = 0.0f;
        T[15] = 1.0f;
    }
};

// errorFunction2.  T: row-major 4x4 double (cast from float).  C = latched depth covariance.
__device__ inline double mahal2(const rgbdfe_point3f& a, const rgbdfe_point3f& b, const double* T, double C, double rcx, double rcy)
{
    if (isnan(a.z) || isnan(b.z))
        return DBL_MAX;
    const double x = a.x, y = a.y, z = a.z;
    const double mu2[3] = { (double)b.x, (double)b.y, (double)b.z };
    double d[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        double m = ((T[i * 4 + 0] * x + T[i * 4 + 1] * y) + T[i * 4 + 2] * z) + T[i * 4 + 3] * 1.0;
        d[i] = m - mu2[i];
    }
    {
        double dsq = d[0] * d[0] + (d[1] * d[1] + d[2] * d[2]);
        double s1 = smax(rcx, C), s2 = smax(rcx, C);
        if (dsq > 2.0 * (s1 + s2))
            return DBL_MAX;
    }
    const double c1[3] = { rcx * z, rcy * z, C };
    const double c2[3] = { rcx * mu2[2], rcy * mu2[2], C };
    double S[9];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
            double t0 = (T[0 * 4 + i] * c1[0]) * T[0 * 4 + j];
            double t1 = (T[1 * 4 + i] * c1[1]) * T[1 * 4 + j];
            double t2 = (T[2 * 4 + i] * c1[2]) * T[2 * 4 + j];
            S[i * 3 + j] = (t0 + (t1 + t2)) + (i == j ? c2[i] : 0.0);
        }
    if (isnan(d[2]))
        return DBL_MAX;
    do {
        double xk = S[0];
        if (xk <= 0.0)
            break;
        S[0] = xk = sqrt(xk);
        S[3] /= xk;
        S[6] /= xk;
        xk = S[4] - S[3] * S[3];
        if (xk <= 0.0)
            break;
        S[4] = xk = sqrt(xk);
        S[7] = S[7] - S[6] * S[3];
        S[7] /= xk;
        xk = S[8] - (S[6] * S[6] + S[7] * S[7]);
        if (xk <= 0.0)
            break;
        S[8] = sqrt(xk);
    } while (0);
    double b0 = d[0], b1 = d[1], b2 = d[2];
    b0 /= S[0];
    b1 -= S[3] * b0;
    b1 /= S[4];
    b2 -= (S[6] * b0 + S[7] * b1);
    b2 /= S[8];
    b2 /= S[8];
    b1 -= S[7] * b2;
    b1 /= S[4];
    b0 -= (S[3] * b1 + S[6] * b2);
    b0 /= S[0];
    double r = d[0] * b0 + (d[1] * b1 + d[2] * b2);
    if (!(r >= 0.0))
        return DBL_MAX;
    return r;
}

} // namespace rgbdfe
