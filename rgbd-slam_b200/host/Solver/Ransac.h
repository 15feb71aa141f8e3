// Ransac — same public interface as the reference's Solver/Ransac.h:13-93 (class Ransac : Solver,
// Umeyama RANSAC).  estimateTransformation and its helpers run on the GPU
// (rgbdfe_ransac_umeyama: kernel group K12).  Only EUCLIDEAN_ERROR and ADAPTIVE_ERROR are
// executable in the reference (the other versions dereference the null mpCamera, Ransac.cpp:15,350,
// or read an uninitialised covariance, :282-321); the same two are executable here.
#ifndef RANSAC_H
#define RANSAC_H

#include "Solver.h"
#include <memory>
#include <set>
#include <vector>

class RGBDcamera;

class Ransac : public Solver {
public:
    enum ERROR_VERSION { EUCLIDEAN_ERROR,
        REPROJECTION_ERROR,
        EUCLIDEAN_AND_REPROJECTION_ERROR,
        MAHALANOBIS_ERROR,
        ADAPTIVE_ERROR };

    struct parameters {
        int verbose;
        int errorVersion, errorVersionVO, errorVersionMap;
        double inlierThresholdEuclidean, inlierThresholdReprojection, inlierThresholdMahalanobis;
        double minimalInlierRatioThreshold;
        int minimalNumberOfMatches;
        int usedPairs;
        int iterationCount;
    };

    Ransac(const std::shared_ptr<Frame> F1, std::shared_ptr<Frame> F2, const std::vector<cv::DMatch>& matches, Ransac::parameters params);

    virtual ~Ransac() {}

    bool compute(std::vector<cv::DMatch>& inliers) override;

protected:
    // Robustly estimate transformation from given two sets of 3D features and
    // potentially correct matches between those sets
    Eigen::Matrix4f estimateTransformation(std::vector<cv::Point3f> prevFeatures,
        std::vector<cv::Point3f> features, std::vector<cv::DMatch> matches,
        std::vector<cv::DMatch>& inlierMatches);

    // Compute inlier ratio w.r.t. points in the second image
    static double pointInlierRatio(std::vector<cv::DMatch>& inlierMatches,
        std::vector<cv::DMatch>& allMatches)
    {
        std::set<int> inlier, all;
        for (auto& m : allMatches)
            all.insert(m.trainIdx);

        for (auto& in : inlierMatches)
            inlier.insert(in.trainIdx);

        return double(inlier.size()) / double(all.size());
    }

    std::shared_ptr<RGBDcamera> mpCamera;
    parameters mRANSACParams;
};

#endif // RANSAC_H
