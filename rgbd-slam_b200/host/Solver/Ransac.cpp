// Ransac shim (reference: Solver/Ransac.cpp:12-59 ctor / compute, :61-181 estimateTransformation).
#include "Ransac.h"
#include "../System/Random.h"
#include "../rgbdfe_host.h"
#ifdef RGBDFE_WITH_OPENCV
#include "Core/Frame.h"
#else
#include "../Core/Frame.h"
#endif
#include <cmath>
#include <iostream>

using namespace std;

Ransac::Ransac(const Frame::Ptr F1, Frame::Ptr F2, const vector<cv::DMatch>& matches, Ransac::parameters params)
    : Solver(F1, F2, matches)
{
    mpCamera = nullptr;

    mRANSACParams.verbose = params.verbose;
    mRANSACParams.errorVersion = params.errorVersion;
    mRANSACParams.usedPairs = params.usedPairs;
    mRANSACParams.inlierThresholdEuclidean = params.inlierThresholdEuclidean;
    mRANSACParams.inlierThresholdReprojection = params.inlierThresholdReprojection;
    mRANSACParams.inlierThresholdMahalanobis = params.inlierThresholdMahalanobis;
    mRANSACParams.minimalInlierRatioThreshold = params.minimalInlierRatioThreshold;
    mRANSACParams.minimalNumberOfMatches = params.minimalNumberOfMatches;

    // computeRANSACIteration(0.20): the ctor ignores params.iterationCount (Ransac.cpp:26)
    mRANSACParams.iterationCount = int((log(1 - 0.98) / log(1 - pow(0.20, 3))));

    if (mRANSACParams.verbose > 0) {
        cout << "RANSACParams.verbose --> " << mRANSACParams.verbose << endl;
        cout << "RANSACParams.usedPairs --> " << mRANSACParams.usedPairs << endl;
        cout << "RANSACParams.errorVersion --> " << mRANSACParams.errorVersion << endl;
        cout << "RANSACParams.inlierThresholdEuclidean --> " << mRANSACParams.inlierThresholdEuclidean << endl;
        cout << "RANSACParams.minimalInlierRatioThreshold --> " << mRANSACParams.minimalInlierRatioThreshold << endl;
    }
}

bool Ransac::compute(vector<cv::DMatch>& inliers)
{
    Eigen::Matrix4f T = estimateTransformation(mF1->mvKeys3Dc, mF2->mvKeys3Dc, mMatches, inliers);

    for (const auto& m : inliers)
        mF2->setInlier(m.trainIdx);

    cv::Mat Tm(4, 4, CV_32F); // Converter::toMat<float, 4, 4>(T)
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++)
            Tm.at<float>(r, c) = T(r, c);
    mF2->setPose(Tm.inv() * mF1->getPose());

    if (T.isIdentity())
        return false;
    else
        return true;
}

Eigen::Matrix4f Ransac::estimateTransformation(vector<cv::Point3f> prevFeatures, vector<cv::Point3f> features, vector<cv::DMatch> matches,
    vector<cv::DMatch>& bestInlierMatches)
{
    if (mRANSACParams.verbose > 0)
        cout << "RANSAC: original matches.size() = " << matches.size() << endl;

    // every matched keypoint of F2 becomes an outlier first (inside the remove_if of Ransac.cpp:72-83)
    for (const auto& m : matches)
        mF2->setOutlier(m.trainIdx);

    Random::initSeed(); // getRandomMatches draws through Random::randomInt (:195)

    rgbdfe_umeyama_params prm;
    prm.errorVersion = mRANSACParams.errorVersion;
    prm.inlierThresholdEuclidean = mRANSACParams.inlierThresholdEuclidean;
    prm.minimalInlierRatioThreshold = mRANSACParams.minimalInlierRatioThreshold;
    prm.minimalNumberOfMatches = mRANSACParams.minimalNumberOfMatches;
    prm.usedPairs = mRANSACParams.usedPairs;
    rgbdfe_umeyama_result res;
    bestInlierMatches.assign(matches.size(), cv::DMatch());
    const int rc = rgbdfe_ransac_umeyama(rgbdfe_host::ctx(), reinterpret_cast<const rgbdfe_point3f*>(prevFeatures.data()), (int)prevFeatures.size(),
        reinterpret_cast<const rgbdfe_point3f*>(features.data()), (int)features.size(), reinterpret_cast<const rgbdfe_dmatch*>(matches.data()),
        (int)matches.size(), &prm, &res, reinterpret_cast<rgbdfe_dmatch*>(bestInlierMatches.data()), (int)bestInlierMatches.size());
    if (rgbdfe_host::report("rgbdfe_ransac_umeyama", rc) != RGBDFE_OK) {
        bestInlierMatches.clear();
        return Eigen::Matrix4f::Identity();
    }
    bestInlierMatches.resize(res.n_inliers);
    Eigen::Matrix4f T;
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++)
            T(r, c) = res.T[r * 4 + c];

    if (mRANSACParams.verbose > 0)
        cout << "RANSAC: iterations executed = " << res.iters_executed << ", inliers = " << res.n_inliers << endl;
    return T;
}
