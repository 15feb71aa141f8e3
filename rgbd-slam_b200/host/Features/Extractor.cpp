// Extractor shim (reference: Features/Extractor.cpp:15-61,111-158,161-252).
#include "Extractor.h"
#include "ORBextractor.h"
#include <iostream>

using namespace std;

int Extractor::mNorm;

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
    // only the ORB2/ORB2 branch of Extractor.cpp:52-53 can be constructed (see createDetector)
    mpDetector->detectAndCompute(image, mask, keypoints, descriptors);
}

void Extractor::init() {}

void Extractor::createAdaptiveDetector()
{
    cerr << "Not supported adaptive detector" << endl;
    terminate();
}

int Extractor::getLevels()
{
    return dynamic_cast<ORBextractor*>(mpDetector.get())->GetLevels();
}

float Extractor::getScaleFactor()
{
    return dynamic_cast<ORBextractor*>(mpDetector.get())->GetScaleFactor();
}

vector<float> Extractor::getScaleFactors()
{
    return dynamic_cast<ORBextractor*>(mpDetector.get())->GetScaleFactors();
}

vector<float> Extractor::getInverseScaleFactors()
{
    return dynamic_cast<ORBextractor*>(mpDetector.get())->GetInverseScaleFactors();
}

vector<float> Extractor::getScaleSigmaSquares()
{
    return dynamic_cast<ORBextractor*>(mpDetector.get())->GetScaleSigmaSquares();
}

vector<float> Extractor::getInverseScaleSigmaSquares()
{
    return dynamic_cast<ORBextractor*>(mpDetector.get())->GetInverseScaleSigmaSquares();
}

void Extractor::createDetector()
{
    switch (mDetectorType) {
    case ORB2: // Extractor.cpp:203-205
        mpDetector.reset(new ORBextractor(nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST));
        break;
    default:
        cerr << "Invalid detector!" << endl; // :208-209 (types other than ORB2 are OpenCV wrappers, not rebuilt)
        terminate();
    }
}

void Extractor::createDescriptor()
{
    switch (mDescriptorType) {
    case ORB2: // Extractor.cpp:244-247: the ORB2 descriptor is the detector object itself
        mpDescriptor = mpDetector;
        break;
    default:
        cerr << "Invalid descriptor!" << endl; // :250-251
        terminate();
    }
}
