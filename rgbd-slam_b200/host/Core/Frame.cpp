// Frame stand-in (see Frame.h): the Frame-constructor steps around the extractor,
// restated from Core/Frame.cpp:34-73,91-122 for a camera without lens distortion.
#include "Frame.h"
#include "../Features/Extractor.h"

Frame::Frame(const cv::Mat& imGray, const cv::Mat& imDepth, const double& timeStamp, Extractor::Ptr pExtractor, RGBDcamera* pRGBDcamera)
    : mImGray(imGray)
    , mImDepth(imDepth)
    , mpExtractor(pExtractor)
    , mpCamera(pRGBDcamera)
    , mTimeStamp(timeStamp)
    , N(0)
{
    mTcw = cv::Mat::eye(4, 4, CV_32F);

    // Feature extraction
    extractFeatures();

    N = mvKeys.size();
    if (mvKeys.empty())
        return;

    mvKeysUn = mvKeys; // undistortKeyPoints with distCoef[0] == 0 (Frame.cpp:251-255)
    mvbOutlier = std::vector<bool>(N, false);

    uprojectCamera();
}

void Frame::extractFeatures()
{
    mpExtractor->detectAndCompute(mImGray, cv::Mat(), mvKeys, mDescriptors); // Frame.cpp:119-122
}

void Frame::uprojectCamera()
{
    mvKeys3Dc = std::vector<cv::Point3f>(N, cv::Point3f(0, 0, 0));

    for (size_t i = 0; i < N; i++) {
        const cv::KeyPoint& kp = mvKeys[i];
        const cv::KeyPoint& kpU = mvKeysUn[i];

        const float& v = kp.pt.y;
        const float& u = kp.pt.x;

        const float z = mImDepth.at<float>((int)v, (int)u); // Mat::at<float>(float, float) truncates (Frame.cpp:103)

        if (z > 0) {
            // 3D KeyPoint position in Camera coordinates
            const float u = kpU.pt.x;
            const float v = kpU.pt.y;
            const float x = (u - mpCamera->cx()) * z * mpCamera->invfx();
            const float y = (v - mpCamera->cy()) * z * mpCamera->invfy();
            mvKeys3Dc[i] = cv::Point3f(x, y, z);
        }
    }
}
