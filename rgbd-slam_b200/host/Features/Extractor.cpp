// Extractor shim (reference: Features/Extractor.cpp:15-61,111-158,161-252).
#include "Extractor.h"
#include "ORBextractor.h"
#include "../rgbdfe_host.h"
#include <cstring>
#include <iostream>

using namespace std;

int Extractor::mNorm;

namespace {
// The ORB detector / ORB descriptor pair of the reference (cv::ORB::create(nfeatures, scaleFactor, nlevels) for
// detection, Extractor.cpp:163-166; a default cv::ORB::create() for description, :216-218): one object stands for
// both, because Extractor::detectAndCompute's detect -> retainBest(nfeatures) -> compute sequence (:55-59) runs as
// one device pass (rgbdfe_cvorb_configure + rgbdfe_orb_extract).
class CvOrbOnDevice : public cv::Feature2D {
public:
    CvOrbOnDevice(int nfeatures_, float scaleFactor_, int nlevels_)
        : nfeatures(nfeatures_)
        , scaleFactor(scaleFactor_)
        , nlevels(nlevels_)
    {
        configureDevice();
    }
    void configureDevice()
    {
        if (rgbdfe_host::report("rgbdfe_cvorb_configure", rgbdfe_cvorb_configure(rgbdfe_host::ctx(), nfeatures, scaleFactor, nlevels)) != RGBDFE_OK)
            std::terminate();
        rgbdfe_host::config_owner() = this;
    }
    void detectAndCompute(cv::InputArray _image, cv::InputArray, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray _descriptors, bool = false) override
    {
#ifdef RGBDFE_WITH_OPENCV
        cv::Mat image = _image.getMat();
#else
        const cv::Mat& image = _image;
#endif
        keypoints.clear();
        if (image.empty())
            return;
        if (rgbdfe_host::config_owner() != this)
            configureDevice();
        const int cap = nfeatures + 64; // ties at a retainBest boundary are all kept (keypoint.cpp)
        keypoints.resize(cap);
        cv::Mat desc(cap, 32, CV_8U);
        int n = 0;
        const int rc = rgbdfe_orb_extract(rgbdfe_host::ctx(), image.data, image.cols, image.rows, (int)image.step,
            reinterpret_cast<rgbdfe_keypoint*>(keypoints.data()), desc.data, cap, &n);
        if (rgbdfe_host::report("rgbdfe_orb_extract", rc) != RGBDFE_OK)
            n = 0; // errors map to "no features", never to a CPU path
        keypoints.resize(n);
        if (n == 0) {
            _descriptors.release();
            return;
        }
        _descriptors.create(n, 32, CV_8U);
#ifdef RGBDFE_WITH_OPENCV
        cv::Mat out = _descriptors.getMat();
#else
        cv::Mat& out = _descriptors;
#endif
        for (int i = 0; i < n; i++)
            memcpy(out.ptr<uchar>(i), desc.ptr<uchar>(i), 32);
    }
    int defaultNorm() const override { return cv::NORM_HAMMING; }

private:
    int nfeatures;
    float scaleFactor;
    int nlevels;
};
} // namespace

Extractor::Extractor(eType detector, eType descriptor, eMode mode)
{
    mDetectorType = detector;
    mDescriptorType = descriptor;
    mMode = mode;

    setParameters(1000, 1.2f, 8, 20, 7); // Extractor.cpp:21
}

void Extractor::setParameters(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
{
    nfeatures = _nfeatures;
    scaleFactor = _scaleFactor;
    nlevels = _nlevels;
    iniThFAST = _iniThFAST;
    minThFAST = _minThFAST;

    if (mMode == ADAPTIVE) {
        // the adaptive wrappers (DetectorAdjuster & co, Extractor.cpp:63-109) exist for FAST/GFTT/SURF/SIFT/ORB only
        cerr << "Not supported adaptive detector" << endl;
        terminate();
    } else if (mMode == NORMAL) {
        createDetector();
        cout << "- Normal Feature Extraction" << endl;
    }

    createDescriptor();
    mNorm = mpDescriptor->defaultNorm(); // NORM_HAMMING for ORB2 (Extractor.cpp:47)
}

void Extractor::detectAndCompute(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors)
{
    // ORB2/ORB2 (Extractor.cpp:52-53) and ORB/ORB (:55-59, detect -> retainBest -> compute fused in one device pass)
    // are the pairs that can be constructed (see createDetector)
    mpDetector->detectAndCompute(image, mask, keypoints, descriptors);
}

void Extractor::init()
{
    // Extractor.cpp:63-80
    mvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvScaleFactor[0] = 1.0f;
    mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        mvScaleFactor[i] = mvScaleFactor[i - 1] * scaleFactor;
        mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
    }
    mvInvScaleFactor.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
        mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
        mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
    }
}

void Extractor::createAdaptiveDetector()
{
    cerr << "Not supported adaptive detector" << endl;
    terminate();
}

int Extractor::getLevels()
{
    if (mDetectorType == ORB2 && mDescriptorType == ORB2)
        return dynamic_cast<ORBextractor*>(mpDetector.get())->GetLevels();
    return nlevels;
}

float Extractor::getScaleFactor()
{
    if (mDetectorType == ORB2 && mDescriptorType == ORB2)
        return dynamic_cast<ORBextractor*>(mpDetector.get())->GetScaleFactor();
    return scaleFactor;
}

vector<float> Extractor::getScaleFactors()
{
    if (mDetectorType == ORB2 && mDescriptorType == ORB2)
        return dynamic_cast<ORBextractor*>(mpDetector.get())->GetScaleFactors();
    return mvScaleFactor;
}

vector<float> Extractor::getInverseScaleFactors()
{
    if (mDetectorType == ORB2 && mDescriptorType == ORB2)
        return dynamic_cast<ORBextractor*>(mpDetector.get())->GetInverseScaleFactors();
    return mvInvScaleFactor;
}

vector<float> Extractor::getScaleSigmaSquares()
{
    if (mDetectorType == ORB2 && mDescriptorType == ORB2)
        return dynamic_cast<ORBextractor*>(mpDetector.get())->GetScaleSigmaSquares();
    return mvLevelSigma2;
}

vector<float> Extractor::getInverseScaleSigmaSquares()
{
    if (mDetectorType == ORB2 && mDescriptorType == ORB2)
        return dynamic_cast<ORBextractor*>(mpDetector.get())->GetInverseScaleSigmaSquares();
    return mvInvLevelSigma2;
}

void Extractor::createDetector()
{
    switch (mDetectorType) {
    case ORB: // Extractor.cpp:163-166
        if (mDescriptorType != ORB) {
            cerr << "Invalid descriptor!" << endl; // the ORB detector is rebuilt together with the ORB descriptor only
            terminate();
        }
        init();
        mpDetector.reset(new CvOrbOnDevice(nfeatures, (float)scaleFactor, nlevels));
        break;
    case ORB2: // Extractor.cpp:203-205
        mpDetector.reset(new ORBextractor(nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST));
        break;
    default:
        cerr << "Invalid detector!" << endl; // :208-209 (types other than ORB2 and ORB are OpenCV wrappers, not rebuilt)
        terminate();
    }
}

void Extractor::createDescriptor()
{
    switch (mDescriptorType) {
    case ORB: // Extractor.cpp:216-218: a default cv::ORB; its compute() is part of the detector object's device pass
        if (mDetectorType != ORB) {
            cerr << "Invalid detector!" << endl;
            terminate();
        }
        mpDescriptor = mpDetector;
        break;
    case ORB2: // Extractor.cpp:244-247: the ORB2 descriptor is the detector object itself
        mpDescriptor = mpDetector;
        break;
    default:
        cerr << "Invalid descriptor!" << endl; // :250-251
        terminate();
    }
}
