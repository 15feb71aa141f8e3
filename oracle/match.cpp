// ORACLE — TEST INFRASTRUCTURE ONLY.
// Matcher::match (Features/Matcher.cpp:106-139): BFMatcher knn k=2 (Hamming),
// ratio test `m1.distance < ratio * m2.distance` in float, then in query order:
// skip if the train index was already claimed, skip if discardOutliers and the
// query keypoint is flagged outlier, skip unless both keypoints have z > 0
// (Frame::isValidObs, Core/Frame.cpp:415-418); claim the train index.
// Guard (documented deviation): the reference reads matchesKnn[i][1] unchecked
// when the train set has < 2 rows (Matcher.cpp:117, UB); the oracle returns no
// matches in that case.
#include "oracle.h"

#include <set>

namespace orc {

int matcher_match(const uint8_t* qdesc, int nq, const uint8_t* tdesc, int nt, const uint8_t* q_outlier,
    const Point3f* q3d, const Point3f* t3d, float ratio, int discard_outliers, std::vector<DMatch>& out)
{
    out.clear();
    if (nq <= 0 || nt < 2)
        return 0;
    std::vector<int> idx((size_t)nq * 2), dist((size_t)nq * 2);
    {
        StageTimer t(ST_KNN);
        knn2_hamming(qdesc, nq, tdesc, nt, idx.data(), dist.data());
    }
    StageTimer t(ST_FILTER);
    std::set<int> used;
    for (int i = 0; i < nq; i++) {
        float d1 = (float)dist[i * 2], d2 = (float)dist[i * 2 + 1];
        if (d1 < ratio * d2) {
            int i2 = idx[i * 2];
            if (used.count(i2))
                continue;
            if (discard_outliers && q_outlier && q_outlier[i])
                continue;
            if (!(q3d[i].z > 0) || !(t3d[i2].z > 0))
                continue;
            used.insert(i2);
            DMatch m;
            m.queryIdx = i;
            m.trainIdx = i2;
            m.imgIdx = 0;
            m.distance = d1;
            out.push_back(m);
        }
    }
    return (int)out.size();
}

} // namespace orc
