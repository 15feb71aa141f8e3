// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// CPU restatement of class Ransac (Solver/Ransac.cpp:12-59,61-181,183-208,
// 210-245,247-280,429-454).  Only EUCLIDEAN_ERROR / ADAPTIVE_ERROR are
// executable in the reference (the others dereference a null camera or read an
// uninitialised covariance), so only those are restated.  Eigen::umeyama
// (Eigen unpinned; restated from Eigen 3.3.7 Geometry/Umeyama.h, no scaling)
// -> PARITY UNPINNED bit-for-bit; values pinned against numpy Kabsch to 1e-4
// (tests/test_numeric_pins.py).  Arithmetic contract: FP32, no FMA, sums over the n
// correspondences and over 3-term products run left to right.
#include "oracle.h"

#include <algorithm>
#include <cmath>

namespace orc {

static inline float det3(const float* m)
{
    return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
}

// src, dst: n x 3 (row per point).  T maps src -> dst.  NaN-filled when n == 0.
void umeyama3(const float* src, const float* dst, int n, float T[16])
{
    const float one_over_n = 1.0f / (float)n;
    float sm[3] = { 0, 0, 0 }, dm[3] = { 0, 0, 0 };
    for (int k = 0; k < n; k++)
        for (int i = 0; i < 3; i++) {
            sm[i] += src[k * 3 + i];
            dm[i] += dst[k * 3 + i];
        }
    for (int i = 0; i < 3; i++) {
        sm[i] *= one_over_n;
        dm[i] *= one_over_n;
    }
    float sigma[9] = { 0, 0, 0, 0, 0, 0, 0, 0, 0 };
    for (int k = 0; k < n; k++) {
        float sd[3], dd[3];
        for (int i = 0; i < 3; i++) {
            sd[i] = src[k * 3 + i] - sm[i];
            dd[i] = one_over_n * (dst[k * 3 + i] - dm[i]);
        }
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++)
                sigma[i * 3 + j] += dd[i] * sd[j];
    }
    if (n == 0)
        for (int i = 0; i < 9; i++)
            sigma[i] = NAN;
    float U[9], S[3], V[9];
    jacobi_svd3(sigma, U, S, V);
    float s2 = (det3(U) * det3(V) < 0.0f) ? -1.0f : 1.0f;
    float R[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            R[i * 3 + j] = (U[i * 3 + 0] * V[j * 3 + 0] + U[i * 3 + 1] * V[j * 3 + 1]) + (U[i * 3 + 2] * s2) * V[j * 3 + 2];
    for (int i = 0; i < 3; i++) {
        float rm = (R[i * 3 + 0] * sm[0] + R[i * 3 + 1] * sm[1]) + R[i * 3 + 2] * sm[2];
        T[i * 4 + 0] = R[i * 3 + 0];
        T[i * 4 + 1] = R[i * 3 + 1];
        T[i * 4 + 2] = R[i * 3 + 2];
        T[i * 4 + 3] = dm[i] - rm;
    }
    T[12] = T[13] = T[14] = 0.0f;
    T[15] = 1.0f;
}

namespace {
    int ransac_iters(double inlierRatio, double successProbability = 0.98, int numberOfPairs = 3)
    {
        return int((log(1 - successProbability) / log(1 - pow(inlierRatio, numberOfPairs))));
    }

    bool model(const Point3f* prev, const Point3f* cur, const std::vector<DMatch>& m, float T[16])
    {
        std::vector<float> pf(m.size() * 3), cf(m.size() * 3);
        for (size_t j = 0; j < m.size(); j++) {
            const Point3f& a = prev[m[j].queryIdx];
            const Point3f& b = cur[m[j].trainIdx];
            pf[j * 3] = a.x, pf[j * 3 + 1] = a.y, pf[j * 3 + 2] = a.z;
            cf[j * 3] = b.x, cf[j * 3 + 1] = b.y, cf[j * 3 + 2] = b.z;
        }
        umeyama3(cf.data(), pf.data(), (int)m.size(), T);
        if (std::isnan(T[0])) {
            for (int i = 0; i < 16; i++)
                T[i] = (i % 5 == 0) ? 1.0f : 0.0f;
            return false;
        }
        return true;
    }

    float inlier_ratio_euclid(const Point3f* prev, const Point3f* cur, const std::vector<DMatch>& m, const float T[16],
        const UmeyamaParams& prm, std::vector<DMatch>& consistent)
    {
        int inlierCount = 0;
        for (const DMatch& it : m) {
            const Point3f& c = cur[it.trainIdx];
            const Point3f& p = prev[it.queryIdx];
            float e[3];
            for (int i = 0; i < 3; i++) {
                float rc = T[i * 4 + 0] * c.x + (T[i * 4 + 1] * c.y + T[i * 4 + 2] * c.z);
                e[i] = rc + T[i * 4 + 3];
            }
            double threshold = prm.inlierThresholdEuclidean;
            if (prm.errorVersion == 4)
                threshold *= p.z;
            float dx = e[0] - p.x, dy = e[1] - p.y, dz = e[2] - p.z;
            float nrm = sqrtf(dx * dx + (dy * dy + dz * dz));
            if (nrm < threshold) {
                inlierCount++;
                consistent.push_back(it);
            }
        }
        return float(inlierCount) / float(m.size());
    }
}

void ransac_umeyama(const Point3f* p1, const Point3f* p2, const std::vector<DMatch>& m12, const UmeyamaParams& prm,
    GlibcRand& rng, UmeyamaResult& res)
{
    res.inliers.clear();
    res.outlier_marked.clear();
    res.iters_executed = 0;
    float best[16];
    for (int i = 0; i < 16; i++)
        best[i] = res.T[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    double bestRatio = 0.0;
    int iterationCount = ransac_iters(0.20); // ctor ignores params.iterationCount (Ransac.cpp:26)

    std::vector<DMatch> matches;
    for (const DMatch& m : m12) {
        res.outlier_marked.push_back(m.trainIdx);
        const float z1 = p1[m.queryIdx].z, z2 = p2[m.trainIdx].z;
        if (z1 < 0.1f || z1 > 6.0f || z2 < 0.1f || z2 > 6.0f)
            continue;
        matches.push_back(m);
    }
    res.ok = 0;
    if ((int)matches.size() < prm.minimalNumberOfMatches)
        return;

    std::vector<DMatch> bestInl;
    for (int i = 0; i < iterationCount; i++) {
        res.iters_executed++;
        // getRandomMatches (Ransac.cpp:183-208)
        std::vector<DMatch> chosen;
        std::vector<bool> valid(matches.size(), true);
        while ((int)chosen.size() < prm.usedPairs) {
            int s = rng.random_int(0, (int)matches.size() - 1);
            if (valid[s]) {
                chosen.push_back(matches[s]);
                valid[s] = false;
            }
        }
        float T[16];
        if (model(p1, p2, chosen, T)) {
            std::vector<DMatch> consistent;
            float ratio = inlier_ratio_euclid(p1, p2, matches, T, prm, consistent);
            if (ratio > bestRatio) { // saveBetterModel (Ransac.cpp:429-447)
                for (int k = 0; k < 16; k++)
                    best[k] = T[k];
                bestRatio = ratio;
                bestInl.swap(consistent);
                iterationCount = std::min(ransac_iters(prm.minimalInlierRatioThreshold), ransac_iters(bestRatio));
            }
        }
    }
    model(p1, p2, bestInl, best);
    std::vector<DMatch> newBest;
    if (!bestInl.empty())
        inlier_ratio_euclid(p1, p2, bestInl, best, prm, newBest);
    newBest.swap(bestInl);
    if (bestRatio < prm.minimalInlierRatioThreshold) {
        for (int i = 0; i < 16; i++)
            best[i] = (i % 5 == 0) ? 1.0f : 0.0f;
        bestInl.clear();
    }
    for (int i = 0; i < 16; i++)
        res.T[i] = best[i];
    res.inliers = bestInl;
    // !T.isIdentity(1e-5)
    bool ident = true;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float x = best[i * 4 + j];
            if (i == j) {
                if (!(fabsf(x - 1.0f) <= 1e-5f * std::min(fabsf(x), 1.0f)))
                    ident = false;
            } else if (!(fabsf(x) <= 1e-5f))
                ident = false;
        }
    res.ok = ident ? 0 : 1;
}

} // namespace orc
