// ORBextractor — same public interface as the reference's Features/ORBextractor.h:9-64; the work
// is done by librgbdfe (CUDA, sm_100a) through rgbdfe_orb_configure / rgbdfe_orb_extract.
// Protected helpers of the reference (ComputePyramid, ComputeKeyPointsOctTree, DistributeOctTree,
// pattern, umax) have no host counterpart: they are kernels K1-K6 inside the library.
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include "../cvcompat.h"
#include <list>
#include <vector>

class ORBextractor : public cv::Feature2D {
public:
    enum {
        HARRIS_SCORE = 0,
        FAST_SCORE = 1
    };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);

    ~ORBextractor() {}

    // Compute the ORB features and descriptors on an image. ORB are dispersed on the image using an octree.
    // Mask is ignored in the current implementation.
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors);

    CV_WRAP virtual void detectAndCompute(cv::InputArray image, cv::InputArray mask, CV_OUT std::vector<cv::KeyPoint>& keypoints,
        cv::OutputArray descriptors, bool useProvidedKeypoints = false);

    virtual int defaultNorm() const { return cv::NORM_HAMMING; }

    int inline GetLevels() { return nlevels; }

    float inline GetScaleFactor() { return scaleFactor; }

    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }

    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }

    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }

    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // Filled on demand from the device after each extraction (the reference keeps the host pyramid
    // as a by-product of ComputePyramid, ORBextractor.cpp:773-797).  Level images are ROIs of
    // buffers padded by the 19 px border, as in the reference.
    std::vector<cv::Mat> mvImagePyramid;
    void fetchImagePyramid();

protected:
    void configureDevice();

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;

    std::vector<int> mnFeaturesPerLevel;

    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;
};

#endif // ORBEXTRACTOR_H
