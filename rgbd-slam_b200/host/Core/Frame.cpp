// Frame stand-in (see Frame.h): the Frame-constructor steps around the extractor,
// restated from Core/Frame.cpp:34-73,91-122 for a camera without lens distortion.
#include "Frame.h"
#include "../Features/Extractor.h"

float Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv;
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;
bool Frame::mbInitialComputations = true;

cv::Mat Frame::getRotation() const
{
    cv::Mat R(3, 3, CV_32F);
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++)
            R.at<float>(r, c) = mTcw.at<float>(r, c);
    return R;
}

cv::Mat Frame::getTranslation() const
{
    cv::Mat t(3, 1, CV_32F);
    for (int r = 0; r < 3; r++)
        t.at<float>(r, 0) = mTcw.at<float>(r, 3);
    return t;
}

cv::Mat Frame::getRotationInverse() const
{
    cv::Mat R(3, 3, CV_32F);
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++)
            R.at<float>(r, c) = mTcw.at<float>(c, r);
    return R;
}

cv::Mat Frame::getCameraCenter() const
{
    // -mRcw.t() * mtcw: cv::gemm(GEMM_1_T, alpha = -1) accumulates each row in double and narrows once
    cv::Mat o(3, 1, CV_32F);
    for (int r = 0; r < 3; r++) {
        double s = 0.0;
        for (int k = 0; k < 3; k++)
            s += (double)mTcw.at<float>(k, r) * (double)mTcw.at<float>(k, 3);
        o.at<float>(r, 0) = (float)(-s);
    }
    return o;
}

Frame::Frame(const cv::Mat& imGray, const cv::Mat& imDepth, const double& timeStamp, Extractor::Ptr pExtractor, RGBDcamera* pRGBDcamera)
    : mImGray(imGray)
    , mImDepth(imDepth)
    , mpExtractor(pExtractor)
    , mpCamera(pRGBDcamera)
    , mTimeStamp(timeStamp)
    , N(0)
{
    mTcw = cv::Mat::eye(4, 4, CV_32F);

    if (mbInitialComputations) { // Core/Frame.cpp:60-70 for distCoef[0] == 0 (computeImageBounds :306-311)
        mnMinX = 0.0f, mnMaxX = (float)imGray.cols, mnMinY = 0.0f, mnMaxY = (float)imGray.rows;
        mfGridElementWidthInv = 64.0f / (mnMaxX - mnMinX);  // FRAME_GRID_COLS
        mfGridElementHeightInv = 48.0f / (mnMaxY - mnMinY); // FRAME_GRID_ROWS
        mbInitialComputations = false;
    }

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
