// Frame stand-in for builds WITHOUT the reference tree (RGBDFE_WITH_OPENCV undefined): exactly the
// members of the reference's Core/Frame.h:23-216 that the hot-path classes touch
// (Matcher::match, RansacSE3::compute, Ransac::compute) plus the Frame-constructor steps that feed
// them (Core/Frame.cpp:34-73: extractFeatures :119-122, uprojectCamera :91-117 without lens
// distortion).  In a build inside the reference tree the real Core/Frame.h is used instead — the
// shims only rely on these names.
#ifndef FRAME_H
#define FRAME_H

#include "../System/Macros.h"
#include "../cvcompat.h"
#include <vector>

class Extractor;

// pinhole intrinsics the way RGBDcamera exposes them (Core/RGBDcamera.h: cx(), cy(), invfx(), invfy())
struct RGBDcamera {
    float mfx, mfy, mcx, mcy;
    float mDepthMapFactor; // multiplies the raw depth (Core/Frame.cpp:48)
    float fx() const { return mfx; }
    float fy() const { return mfy; }
    float cx() const { return mcx; }
    float cy() const { return mcy; }
    float invfx() const { return 1.0f / mfx; }
    float invfy() const { return 1.0f / mfy; }
};

class Frame {
public:
    SMART_POINTER_TYPEDEFS(Frame);

    Frame() : mpCamera(nullptr), mTimeStamp(0), N(0) {}

    // imGray: CV_8UC1 (the reference converts BGR->gray first, Frame.cpp:47); imDepth: CV_32F metres
    Frame(const cv::Mat& imGray, const cv::Mat& imDepth, const double& timeStamp, std::shared_ptr<Extractor> pExtractor, RGBDcamera* pRGBDcamera);

    void extractFeatures();

    void setPose(cv::Mat Tcw) { mTcw = Tcw.clone(); }
    cv::Mat getPose() const { return mTcw.clone(); }
    // pose blocks of Frame::updatePoseMatrices (Core/Frame.cpp:137-151), read by Matcher::projectionMatch
    cv::Mat getRotation() const;        // mRcw
    cv::Mat getTranslation() const;     // mtcw
    cv::Mat getRotationInverse() const; // mRwc = mRcw.t()
    cv::Mat getCameraCenter() const;    // mOw = -mRcw.t() * mtcw

    bool isInlier(const size_t& idx) const { return mvbOutlier[idx] == false; }
    bool isOutlier(const size_t& idx) const { return mvbOutlier[idx] == true; }
    void setInlier(const size_t& idx) { mvbOutlier[idx] = false; }
    void setOutlier(const size_t& idx) { mvbOutlier[idx] = true; }

    bool isValidObs(const size_t& idx) { return mvKeys3Dc[idx].z > 0; }

public:
    cv::Mat mImGray;
    cv::Mat mImDepth;

    std::shared_ptr<Extractor> mpExtractor;
    RGBDcamera* mpCamera;
    double mTimeStamp;

    // Number of KeyPoints.
    size_t N;

    std::vector<cv::KeyPoint> mvKeys;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<cv::Point3f> mvKeys3Dc;

    // ORB descriptor, each row associated to a keypoint.
    cv::Mat mDescriptors;

    // Flag to identify outlier associations.
    std::vector<bool> mvbOutlier;

    // Undistorted image bounds and grid cell sizes (static in the reference, Core/Frame.h:158-166; set by the first
    // frame, Core/Frame.cpp:60-70)
    static float mfGridElementWidthInv, mfGridElementHeightInv;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
    static bool mbInitialComputations;

private:
    void uprojectCamera();

    cv::Mat mTcw;
};

#endif
