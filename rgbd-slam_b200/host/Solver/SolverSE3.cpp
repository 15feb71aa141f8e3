// RansacSE3 shim (reference: Solver/SolverSE3.cpp:10-21,23-133,299-305).  The arithmetic is
// librgbdfe's; what stays here is the reference's bookkeeping on Frame, in its order.
#include "SolverSE3.h"
#include "../System/Random.h"
#include "../rgbdfe_host.h"
#ifdef RGBDFE_WITH_OPENCV
#include "Core/Frame.h"
#else
#include "../Core/Frame.h"
#endif

using namespace std;

RansacSE3::RansacSE3()
    : mIterations(200)
    , mMinInlierTh(20)
    , mMaxMahalanobisDistance(3.0f)
    , mSampleSize(4)
{
}

RansacSE3::RansacSE3(int iters, uint minInlierTh, float maxMahalanobisDist, uint sampleSize)
    : mIterations(iters)
    , mMinInlierTh(minInlierTh)
    , mMaxMahalanobisDistance(maxMahalanobisDist)
    , mSampleSize(sampleSize)
{
}

bool RansacSE3::compute(Frame::Ptr pF1, Frame::Ptr pF2, const vector<cv::DMatch>& m12, const bool& updateF2)
{
    mvInliers.clear();
    rmse = 1e6;
    mT21 = Eigen::Matrix4f::Identity();

    if (m12.size() < mMinInlierTh) // :29
        return false;

    mpSourceFrame = pF1;
    mpTargetFrame = pF2;

    if (updateF2) { // :38-42
        for (const auto& m : m12)
            pF2->setOutlier(m.trainIdx);
    }

    Random::initSeed(); // sampleMatches draws through Random::randomInt (:141-142), which seeds on first use

    rgbdfe_se3_result res;
    vector<cv::DMatch> inl(m12.size());
    // the 3-D point sets are normally resident already (Matcher::match just ran on these frames); the key is the content
    const int n1 = (int)pF1->mvKeys3Dc.size(), n2 = (int)pF2->mvKeys3Dc.size();
    int rc;
    rgbdfe_frame *f1 = nullptr, *f2 = nullptr;
    if (n1 > 0 && n2 > 0 && (int)pF1->mDescriptors.rows == n1 && (int)pF2->mDescriptors.rows == n2 && pF1->mDescriptors.isContinuous()
        && pF2->mDescriptors.isContinuous()) {
        f1 = rgbdfe_host::resident(pF1->mDescriptors.data, pF1->mvKeys3Dc.data(), n1);
        f2 = rgbdfe_host::resident(pF2->mDescriptors.data, pF2->mvKeys3Dc.data(), n2);
    }
    if (f1 && f2)
        rc = rgbdfe_frames_ransac_se3(rgbdfe_host::ctx(), f1, f2, n1, n2, reinterpret_cast<const rgbdfe_dmatch*>(m12.data()), (int)m12.size(), mIterations,
            mMinInlierTh, mMaxMahalanobisDistance, mSampleSize, &res, reinterpret_cast<rgbdfe_dmatch*>(inl.data()), (int)inl.size());
    else
        rc = rgbdfe_ransac_se3(rgbdfe_host::ctx(), reinterpret_cast<const rgbdfe_point3f*>(pF1->mvKeys3Dc.data()), n1,
            reinterpret_cast<const rgbdfe_point3f*>(pF2->mvKeys3Dc.data()), n2, reinterpret_cast<const rgbdfe_dmatch*>(m12.data()), (int)m12.size(),
            mIterations, mMinInlierTh, mMaxMahalanobisDistance, mSampleSize, &res, reinterpret_cast<rgbdfe_dmatch*>(inl.data()), (int)inl.size());
    if (rgbdfe_host::report("rgbdfe_ransac_se3", rc) != RGBDFE_OK)
        return false; // no CPU path: a failed call is a failed estimate

    rmse = res.rmse;
    inl.resize(res.n_inliers);
    mvInliers.swap(inl);
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++)
            mT21(r, c) = res.T21[r * 4 + c];

    if (mvInliers.size() >= mMinInlierTh) { // :119-131
        if (updateF2) {
            for (const auto& m : mvInliers)
                pF2->setInlier(m.trainIdx);

            cv::Mat T21(4, 4, CV_32F); // Converter::toMat<float, 4, 4>(mT21)
            for (int r = 0; r < 4; r++)
                for (int c = 0; c < 4; c++)
                    T21.at<float>(r, c) = mT21(r, c);
            T21 = T21 * pF1->getPose();
            pF2->setPose(T21);
        }

        return true;
    } else {
        return false;
    }
}

void RansacSE3::setIterations(int iters)
{
    mIterations = iters;
}

void RansacSE3::setMaxMahalanobisDistance(float dist)
{
    mMaxMahalanobisDistance = dist;
}

void RansacSE3::setSampleSize(uint sampleSize)
{
    mSampleSize = sampleSize;
}

void RansacSE3::setInlierThreshold(uint th)
{
    mMinInlierTh = th;
}
