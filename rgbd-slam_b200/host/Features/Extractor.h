// Extractor — same public interface as the reference's Features/Extractor.h:9-72.
// The ORB2/ORB2 combination (the north-star path, Features/Extractor.cpp:52-53,203-205,244-247)
// runs on the GPU through ORBextractor -> librgbdfe; the ORB/ORB combination (OpenCV's cv::ORB with
// Harris scoring, Features/Extractor.cpp:55-59,163-166,216-218) through rgbdfe_cvorb_configure.
// The other detector / descriptor types of
// the reference are thin wrappers over OpenCV (xfeatures2d) classes and are outside the hot
// path this library replaces: asking for them ends the process the way the reference does for
// an invalid type (std::terminate, Extractor.cpp:39,209,251).
#ifndef EXTRACTOR_H
#define EXTRACTOR_H

#include "../System/Macros.h"
#include "../cvcompat.h"
#include <mutex>
#include <string>

class Extractor {
public:
    SMART_POINTER_TYPEDEFS(Extractor);

    enum eType {
        ORB = 0,
        ORB2,
        SVO,
        FAST,
        GFTT,
        STAR,
        BRISK,
        FREAK,
        BRIEF,
        LATCH,
        SURF,
        SIFT
    };

    enum eMode {
        NORMAL = 0,
        ADAPTIVE
    };

    eType mDetectorType, mDescriptorType;
    eMode mMode;

    Extractor(eType detector, eType descriptor, eMode mode);

    void setParameters(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);

    void detectAndCompute(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors);

    void init();
    void createAdaptiveDetector();
    void createDetector();
    void createDescriptor();

    int getLevels();
    float getScaleFactor();
    std::vector<float> getScaleFactors();
    std::vector<float> getInverseScaleFactors();
    std::vector<float> getScaleSigmaSquares();
    std::vector<float> getInverseScaleSigmaSquares();

    // L2 or HAMMING
    static int mNorm;

private:
    std::mutex mMutexExtraction;
    cv::Ptr<cv::DescriptorExtractor> mpDescriptor;
    cv::Ptr<cv::FeatureDetector> mpDetector;

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;

    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;
};

#endif // EXTRACTOR_H
