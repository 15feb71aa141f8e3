// Matcher shim (reference: Features/Matcher.cpp:13-17,106-139,141-144).
#include "Matcher.h"
#include "../rgbdfe_host.h"
#include "Extractor.h"
#ifdef RGBDFE_WITH_OPENCV
#include "Core/Frame.h"
#else
#include "../Core/Frame.h"
#endif

using namespace std;

const int Matcher::TH_HIGH = 100;
const int Matcher::TH_LOW = 50;
const int Matcher::HISTO_LENGTH = 30;

Matcher::Matcher(float nnratio)
    : mfNNratio(nnratio)
{
    // the reference creates a cv::BFMatcher for Extractor::mNorm here (Matcher.cpp:16); the Hamming
    // brute-force matcher is kernel K7 of librgbdfe and needs no per-object state
}

int Matcher::match(Frame::Ptr ref, Frame::Ptr cur, vector<cv::DMatch>& vMatches12, const bool discardOutliers)
{
    static_assert(sizeof(cv::DMatch) == sizeof(rgbdfe_dmatch), "cv::DMatch layout");
    static_assert(sizeof(cv::Point3f) == sizeof(rgbdfe_point3f), "cv::Point3f layout");
    vMatches12.clear();
    const int nq = (int)ref->N, nt = (int)cur->N;
    if (nq == 0 || nt == 0)
        return 0;
    if (Extractor::mNorm != cv::NORM_HAMMING) {
        fprintf(stderr, "Matcher: only NORM_HAMMING descriptors (ORB) are matched on the device\n");
        return 0;
    }
    // vector<bool> -> bytes for the C ABI (Frame::isOutlier, Core/Frame.cpp:334-337)
    vector<uint8_t> outl(nq);
    for (int i = 0; i < nq; ++i)
        outl[i] = ref->isOutlier(i) ? 1 : 0;
    // descriptor rows must be contiguous 32-byte rows (they are: ORBextractor creates an N x 32 CV_8U matrix)
    cv::Mat qd = ref->mDescriptors.isContinuous() ? ref->mDescriptors : ref->mDescriptors.clone();
    cv::Mat td = cur->mDescriptors.isContinuous() ? cur->mDescriptors : cur->mDescriptors.clone();
    vMatches12.resize(nq);
    int n = 0;
    const int rc = rgbdfe_match_knn2(rgbdfe_host::ctx(), qd.data, nq, td.data, nt, outl.data(),
        reinterpret_cast<const rgbdfe_point3f*>(ref->mvKeys3Dc.data()), reinterpret_cast<const rgbdfe_point3f*>(cur->mvKeys3Dc.data()), mfNNratio,
        discardOutliers ? 1 : 0, reinterpret_cast<rgbdfe_dmatch*>(vMatches12.data()), (int)vMatches12.size(), &n);
    if (rgbdfe_host::report("rgbdfe_match_knn2", rc) != RGBDFE_OK)
        n = 0; // errors map to "no matches", never to a CPU path
    vMatches12.resize(n);
    for (auto& m : vMatches12)
        m.imgIdx = 0; // what BFMatcher::knnMatch reports for a single train image
    return int(vMatches12.size());
}

double Matcher::descriptorDistance(const cv::Mat& a, const cv::Mat& b)
{
    // cv::norm(a, b, NORM_HAMMING) of two 32-byte rows (Matcher.cpp:141-144); trivial, stays on the host
    const unsigned char* pa = a.ptr<unsigned char>();
    const unsigned char* pb = b.ptr<unsigned char>();
    int dist = 0;
    for (int i = 0; i < a.cols; i++)
        dist += __builtin_popcount((unsigned)(pa[i] ^ pb[i]));
    return (double)dist;
}
