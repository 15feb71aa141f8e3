// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// CPU restatement of RansacSE3 (Solver/SolverSE3.cpp:23-297) plus the two
// third-party pieces it calls that are NOT under /root/reference:
//   * pcl::TransformationFromCorrespondences (PCL `common`, unpinned in
//     CMakeLists.txt:55; restated from PCL 1.8-1.12's
//     common/impl/transformation_from_correspondences.hpp, unchanged across
//     those releases), used at SolverSE3.cpp:161-179;
//   * Eigen::JacobiSVD<Matrix3f> and Eigen::LLT<Matrix3d> (Eigen unpinned,
//     CMakeLists.txt:37; restated from Eigen 3.3.7: two-sided Jacobi, sweep
//     order p=1..2,q<p, threshold max(FLT_MIN, 2*eps*maxDiag), unblocked
//     lower Cholesky + two triangular solves).
// PARITY UNPINNED bit-for-bit for those two: the reference has no tests or vectors
// and neither library is installed here.  Their VALUES are pinned against numpy
// float64 / LAPACK (tests/test_numeric_pins.py: SVD 1e-5, fit 1e-4, LLT 1e-10).
// Arithmetic contract (what the CUDA kernels must reproduce bit-for-bit):
//   FP32 for the fit, FP64 for scoring, NO fused multiply-add anywhere,
//   3-term sums associate as a + (b + c) (Eigen's unrolled redux tree),
//   4x4*vec4 rows associate left to right.
#include "oracle.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <set>

namespace orc {

// ---- Eigen::JacobiSVD<Matrix3f>(ComputeFullU|ComputeFullV), row-major 3x3 ----
namespace {
    struct Rot {
        float c, s;
    };
    inline bool make_jacobi(float x, float y, float z, Rot& r)
    {
        float deno = 2.0f * fabsf(y);
        if (deno < FLT_MIN) {
            r.c = 1.0f;
            r.s = 0.0f;
            return false;
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
        return true;
    }
    // rows p,q of M (applyOnTheLeft) : x' = c x + s y ; y' = -s x + c y
    inline void rot_rows(float* M, int p, int q, Rot j)
    {
        for (int i = 0; i < 3; i++) {
            float x = M[p * 3 + i], y = M[q * 3 + i];
            M[p * 3 + i] = j.c * x + j.s * y;
            M[q * 3 + i] = -j.s * x + j.c * y;
        }
    }
    // cols p,q of M with the rotation (c,s): x' = c x + s y ; y' = -s x + c y
    inline void rot_cols(float* M, int p, int q, Rot j)
    {
        for (int i = 0; i < 3; i++) {
            float x = M[i * 3 + p], y = M[i * 3 + q];
            M[i * 3 + p] = j.c * x + j.s * y;
            M[i * 3 + q] = -j.s * x + j.c * y;
        }
    }
}

void jacobi_svd3(const float A[9], float U[9], float S[3], float V[9])
{
    const float precision = 2.0f * FLT_EPSILON;
    const float considerAsZero = FLT_MIN;
    float scale = 0.0f;
    for (int i = 0; i < 9; i++) {
        float a = fabsf(A[i]);
        if (a > scale)
            scale = a;
    }
    if (scale == 0.0f)
        scale = 1.0f;
    float W[9];
    for (int i = 0; i < 9; i++) {
        W[i] = A[i] / scale;
        U[i] = V[i] = (i % 4 == 0) ? 1.0f : 0.0f;
    }
    float maxDiag = std::max(fabsf(W[0]), std::max(fabsf(W[4]), fabsf(W[8])));
    bool finished = false;
    for (int sweep = 0; !finished && sweep < 64; sweep++) {
        finished = true;
        for (int p = 1; p < 3; ++p) {
            for (int q = 0; q < p; ++q) {
                float threshold = std::max(considerAsZero, precision * maxDiag);
                if (fabsf(W[p * 3 + q]) > threshold || fabsf(W[q * 3 + p]) > threshold) {
                    finished = false;
                    // real_2x2_jacobi_svd
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
                    // m.applyOnTheLeft(0,1,rot1)
                    float n00 = rot1.c * m00 + rot1.s * m10, n01 = rot1.c * m01 + rot1.s * m11;
                    float n11 = -rot1.s * m01 + rot1.c * m11;
                    Rot jr;
                    make_jacobi(n00, n01, n11, jr);
                    // j_left = rot1 * j_right.transpose()
                    Rot jrt = { jr.c, -jr.s };
                    Rot jl = { rot1.c * jrt.c - rot1.s * jrt.s, rot1.c * jrt.s + rot1.s * jrt.c };

                    rot_rows(W, p, q, jl); // W.applyOnTheLeft(p,q,jl)
                    rot_cols(U, p, q, jl); // U.applyOnTheRight(p,q,jl.transpose())
                    Rot jrT = { jr.c, -jr.s };
                    rot_cols(W, p, q, jrT); // W.applyOnTheRight(p,q,jr)
                    rot_cols(V, p, q, jrT); // V.applyOnTheRight(p,q,jr)
                    maxDiag = std::max(maxDiag, std::max(fabsf(W[p * 3 + p]), fabsf(W[q * 3 + q])));
                }
            }
        }
    }
    for (int i = 0; i < 3; i++) {
        float a = W[i * 3 + i];
        S[i] = fabsf(a);
        if (a < 0.0f)
            for (int r = 0; r < 3; r++)
                U[r * 3 + i] = -U[r * 3 + i];
    }
    for (int i = 0; i < 3; i++)
        S[i] *= scale;
    for (int i = 0; i < 3; i++) {
        int pos = i;
        float mx = S[i];
        for (int k = i + 1; k < 3; k++)
            if (S[k] > mx) {
                mx = S[k];
                pos = k;
            }
        if (mx == 0.0f)
            break;
        if (pos != i) {
            std::swap(S[i], S[pos]);
            for (int r = 0; r < 3; r++) {
                std::swap(U[r * 3 + i], U[r * 3 + pos]);
                std::swap(V[r * 3 + i], V[r * 3 + pos]);
            }
        }
    }
}

static inline float det3(const float* m)
{
    // Eigen bruteforce_det3_helper
    return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
}

// RansacSE3::getTransformFromMatches (SolverSE3.cpp:161-179) + PCL TFC.
void tfc_fit(const Point3f* p1, const Point3f* p2, const DMatch* m, int n, float T[16])
{
    float accW = 0.0f;
    float mean1[3] = { 0, 0, 0 }, mean2[3] = { 0, 0, 0 };
    float cov[9] = { 0, 0, 0, 0, 0, 0, 0, 0, 0 };
    for (int k = 0; k < n; k++) {
        const Point3f& from = p1[m[k].queryIdx];
        const Point3f& to = p2[m[k].trainIdx];
        if (std::isnan(from.z) || std::isnan(to.z))
            continue;
        float weight = 1.0f / (from.z * to.z);
        if (weight == 0.0f)
            continue;
        accW += weight;
        float alpha = weight / accW;
        float d1[3] = { from.x - mean1[0], from.y - mean1[1], from.z - mean1[2] };
        float d2[3] = { to.x - mean2[0], to.y - mean2[1], to.z - mean2[2] };
        float oma = 1.0f - alpha;
        for (int i = 0; i < 3; i++) {
            float ad2 = alpha * d2[i];
            for (int j = 0; j < 3; j++)
                cov[i * 3 + j] = oma * (cov[i * 3 + j] + ad2 * d1[j]);
        }
        for (int i = 0; i < 3; i++) {
            mean1[i] += alpha * d1[i];
            mean2[i] += alpha * d2[i];
        }
    }
    float U[9], S[3], V[9];
    jacobi_svd3(cov, U, S, V);
    float s2 = (det3(U) * det3(V) < 0.0f) ? -1.0f : 1.0f;
    float R[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            R[i * 3 + j] = U[i * 3 + 0] * V[j * 3 + 0] + (U[i * 3 + 1] * V[j * 3 + 1] + (U[i * 3 + 2] * s2) * V[j * 3 + 2]);
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

// RansacSE3::errorFunction2 (SolverSE3.cpp:216-280), T row-major 4x4 double.
double mahal2(const Point3f& a, const Point3f& b, const double T[16], DepthCovLatch& latch, int* full)
{
    static const double cam_angle_x = 58.0 / 180.0 * M_PI;
    static const double cam_angle_y = 45.0 / 180.0 * M_PI;
    static const double raster_stddev_x = 3 * tan(cam_angle_x / 640);
    static const double raster_stddev_y = 3 * tan(cam_angle_y / 480);
    static const double rcx = raster_stddev_x * raster_stddev_x;
    static const double rcy = raster_stddev_y * raster_stddev_y;
    if (full)
        *full = 0;
    if (std::isnan(a.z) || std::isnan(b.z))
        return DBL_MAX;
    const double x = a.x, y = a.y, z = a.z;
    const double mu2[3] = { b.x, b.y, b.z };
    double m[3], d[3];
    for (int i = 0; i < 3; i++) {
        m[i] = ((T[i * 4 + 0] * x + T[i * 4 + 1] * y) + T[i * 4 + 2] * z) + T[i * 4 + 3] * 1.0;
        d[i] = m[i] - mu2[i];
    }
    // depthCovariance latch (SolverSE3.cpp:282-287): first depth ever seen.
    if (!latch.latched) {
        double sd = 0.01 * z * z;
        latch.cov = sd * sd;
        latch.latched = 1;
    }
    const double C = latch.cov;
    {
        double dsq = d[0] * d[0] + (d[1] * d[1] + d[2] * d[2]);
        double s1 = std::max(rcx, C), s2 = std::max(rcx, C);
        if (dsq > 2.0 * (s1 + s2))
            return DBL_MAX;
    }
    if (full)
        *full = 1;
    const double c1[3] = { rcx * z, rcy * z, C };
    const double c2[3] = { rcx * mu2[2], rcy * mu2[2], C };
    double S[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            // (R^T cov1 R)(i,j) = sum_k (R(k,i) c1_k) R(k,j)
            double t0 = (T[0 * 4 + i] * c1[0]) * T[0 * 4 + j];
            double t1 = (T[1 * 4 + i] * c1[1]) * T[1 * 4 + j];
            double t2 = (T[2 * 4 + i] * c1[2]) * T[2 * 4 + j];
            S[i * 3 + j] = (t0 + (t1 + t2)) + (i == j ? c2[i] : 0.0);
        }
    if (std::isnan(d[2]))
        return DBL_MAX;
    // LLT (lower, unblocked, in place); on a non-positive pivot Eigen stops
    // factoring but solve() still runs on the partly factored matrix.
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

namespace {
    double inliers_and_error(const Point3f* p1, const Point3f* p2, const std::vector<DMatch>& all, const float T4f[16],
        float maxMahal, DepthCovLatch& latch, std::vector<DMatch>& inl, Se3Stats* st)
    {
        inl.clear();
        double meanError = 0.0;
        double T[16];
        for (int i = 0; i < 16; i++)
            T[i] = (double)T4f[i];
        const double th = (double)(maxMahal * maxMahal);
        for (const DMatch& m : all) {
            const Point3f& o = p1[m.queryIdx];
            const Point3f& t = p2[m.trainIdx];
            if (o.z == 0.0f || t.x == 0.0f)
                continue;
            int full = 0;
            double md = mahal2(o, t, T, latch, &full);
            if (st) {
                if (full)
                    st->ef++;
                else
                    st->es++;
            }
            if (md > th)
                continue;
            if (!(md >= 0.0))
                continue;
            meanError += md;
            inl.push_back(m);
        }
        if (inl.size() < 3)
            meanError = 1e9;
        else {
            meanError /= inl.size();
            meanError = sqrt(meanError);
        }
        return meanError;
    }
}

void ransac_se3(const Point3f* p1, const Point3f* p2, const std::vector<DMatch>& m12, const Se3Params& prm,
    GlibcRand& rng, DepthCovLatch& latch, Se3Result& res, Se3Stats* st)
{
    res.inliers.clear();
    res.rmse = 1e6f;
    res.ok = 0;
    res.iters_executed = 0;
    for (int i = 0; i < 16; i++)
        res.T[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    if (m12.size() < prm.min_inlier_th)
        return;
    std::vector<DMatch> used(m12.begin(), m12.end());
    int validIters = 0;
    double inlierError;
    std::sort(used.begin(), used.end(), [](const DMatch& a, const DMatch& b) { return a.distance < b.distance; });

    for (int n = 0; (n < prm.iterations && used.size() >= prm.sample_size); n++) {
        double refinedError = 1e6;
        std::vector<DMatch> refined;
        // sampleMatches (SolverSE3.cpp:135-159)
        std::vector<DMatch> cur;
        {
            std::set<size_t> ids;
            int safetyNet = 0;
            while (ids.size() < prm.sample_size && used.size() >= prm.sample_size) {
                int id1 = rng.random_int(0, (int)used.size() - 1);
                int id2 = rng.random_int(0, (int)used.size() - 1);
                if (id1 > id2)
                    id1 = id2;
                ids.insert((size_t)id1);
                if (++safetyNet > 10000)
                    break;
            }
            for (size_t id : ids)
                cur.push_back(used[id]);
        }
        float refinedT[16];
        for (int i = 0; i < 16; i++)
            refinedT[i] = (i % 5 == 0) ? 1.0f : 0.0f;
        res.iters_executed++;
        if (st)
            st->iters++;

        for (int refinements = 1; refinements < 20; refinements++) {
            float T[16];
            tfc_fit(p1, p2, cur.data(), (int)cur.size(), T);
            if (st) {
                st->adds += (long)cur.size();
                st->svds++;
                st->rounds++;
            }
            inlierError = inliers_and_error(p1, p2, used, T, prm.max_mahal, latch, cur, st);
            if (cur.size() < prm.min_inlier_th || inlierError > prm.max_mahal)
                break;
            if (cur.size() >= refined.size() && inlierError <= refinedError) {
                size_t prevNum = refined.size();
                for (int i = 0; i < 16; i++)
                    refinedT[i] = T[i];
                refined = cur;
                refinedError = inlierError;
                if (cur.size() == prevNum)
                    break;
            } else
                break;
        }

        if (refined.size() > 0) {
            validIters++;
            if (refinedError <= res.rmse && refined.size() >= res.inliers.size() && refined.size() >= prm.min_inlier_th) {
                res.rmse = (float)refinedError;
                for (int i = 0; i < 16; i++)
                    res.T[i] = refinedT[i];
                res.inliers = refined;
                if (refined.size() > used.size() * 0.5)
                    n += 10;
                if (refined.size() > used.size() * 0.75)
                    n += 10;
                if (refined.size() > used.size() * 0.8)
                    break;
            }
        }
    }

    if (validIters == 0) {
        float I[16];
        for (int i = 0; i < 16; i++)
            I[i] = (i % 5 == 0) ? 1.0f : 0.0f;
        std::vector<DMatch> inl;
        inlierError = inliers_and_error(p1, p2, used, I, prm.max_mahal, latch, inl, st);
        if (inl.size() > prm.min_inlier_th && inlierError < prm.max_mahal) {
            for (int i = 0; i < 16; i++)
                res.T[i] = I[i];
            res.inliers = inl;
            res.rmse = (float)((double)res.rmse + inlierError); // `rmse += inlierError` (:114), float += double
            validIters++;
        }
    }
    res.ok = res.inliers.size() >= prm.min_inlier_th ? 1 : 0;
}

} // namespace orc
