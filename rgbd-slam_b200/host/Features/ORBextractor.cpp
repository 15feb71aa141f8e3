// ORBextractor shim (reference: Features/ORBextractor.cpp:348-406 ctor tables, :706-771
// operator() / detectAndCompute).  Every table and every pixel is computed by librgbdfe.
#include "ORBextractor.h"

#include "../rgbdfe_host.h"

#include <cassert>

using namespace std;

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures)
    , scaleFactor(_scaleFactor)
    , nlevels(_nlevels)
    , iniThFAST(_iniThFAST)
    , minThFAST(_minThFAST)
{
    configureDevice();
}

void ORBextractor::configureDevice()
{
    // the thread's context is configured for THIS extractor's parameters; the tables come back from
    // the library so that getters return exactly what the kernels use (ORBextractor.cpp:353-383)
    rgbdfe_ctx* c = rgbdfe_host::ctx();
    if (rgbdfe_host::report("rgbdfe_orb_configure", rgbdfe_orb_configure(c, nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST))
        != RGBDFE_OK)
        std::terminate();
    mvScaleFactor.resize(nlevels);
    mvInvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    rgbdfe_host::config_owner() = this;
    rgbdfe_host::report("rgbdfe_orb_tables",
        rgbdfe_orb_tables(c, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data(), mnFeaturesPerLevel.data()));
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*mask: ignored, ORBextractor.h:21-22*/, vector<cv::KeyPoint>& _keypoints,
    cv::OutputArray _descriptors)
{
#ifdef RGBDFE_WITH_OPENCV
    if (_image.empty())
        return;
    cv::Mat image = _image.getMat();
#else
    const cv::Mat& image = _image;
    if (image.empty()) // ORBextractor.cpp:708-709
        return;
#endif
    assert(image.type() == CV_8UC1); // :712

    static_assert(sizeof(cv::KeyPoint) == sizeof(rgbdfe_keypoint), "cv::KeyPoint layout");
    rgbdfe_ctx* c = rgbdfe_host::ctx();
    const int cap = nfeatures + 64; // the quadtree may overshoot the per-level quota by up to 3 (ORBextractor.cpp:530,581-582)
    _keypoints.resize(cap);
    cv::Mat desc(cap, 32, CV_8U);
    int n = 0;
    if (rgbdfe_host::config_owner() != this) // another extractor (or thread) configured this thread's context in between
        configureDevice();
    const int rc = rgbdfe_orb_extract(c, image.data, image.cols, image.rows, (int)image.step, reinterpret_cast<rgbdfe_keypoint*>(_keypoints.data()),
        desc.data, cap, &n);
    if (rgbdfe_host::report("rgbdfe_orb_extract", rc) != RGBDFE_OK)
        n = 0; // errors map to "no features", never to a CPU path
    _keypoints.resize(n);
    if (n == 0) {
#ifdef RGBDFE_WITH_OPENCV
        _descriptors.release(); // :726-727
#else
        _descriptors.release();
#endif
        return;
    }
#ifdef RGBDFE_WITH_OPENCV
    _descriptors.create(n, 32, CV_8U); // :729
    cv::Mat out = _descriptors.getMat();
#else
    _descriptors.create(n, 32, CV_8U);
    cv::Mat& out = _descriptors;
#endif
    for (int i = 0; i < n; i++)
        memcpy(out.ptr<uchar>(i), desc.ptr<uchar>(i), 32);
}

void ORBextractor::detectAndCompute(cv::InputArray image, cv::InputArray mask, vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors, bool)
{
    (*this)(image, mask, keypoints, descriptors); // :768-771
}

void ORBextractor::fetchImagePyramid()
{
    rgbdfe_ctx* c = rgbdfe_host::ctx();
    mvImagePyramid.resize(nlevels);
    for (int level = 0; level < nlevels; level++) {
        int w = 0, h = 0;
        if (rgbdfe_orb_pyramid_level(c, level, nullptr, 0, &w, &h) != RGBDFE_OK || w <= 0)
            break;
        cv::Mat padded(h + 38, w + 38, CV_8U);
        if (rgbdfe_host::report("rgbdfe_orb_pyramid_level", rgbdfe_orb_pyramid_level(c, level, padded.data, (int)padded.step, &w, &h)) != RGBDFE_OK)
            break;
#ifdef RGBDFE_WITH_OPENCV
        mvImagePyramid[level] = padded(cv::Rect(19, 19, w, h)); // ROI inside the padded buffer, ORBextractor.cpp:779-782
#else
        cv::Mat roi = padded; // shares ownership of the padded buffer
        roi.data = padded.data + 19 * padded.step + 19;
        roi.rows = h;
        roi.cols = w;
        mvImagePyramid[level] = roi;
#endif
    }
}
