// Matcher shim (reference: Features/Matcher.cpp:13-17,35-104,106-139,141-144).
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
    // both frames stay resident on the device keyed by content: `cur` of this call is `ref` of the next, and
    // RansacSE3::compute names the same two frames right after (Tracking.cpp:127-130)
    rgbdfe_frame* fq = rgbdfe_host::resident(qd.data, ref->mvKeys3Dc.data(), nq);
    rgbdfe_frame* ft = rgbdfe_host::resident(td.data, cur->mvKeys3Dc.data(), nt);
    if (!fq || !ft) {
        vMatches12.clear();
        return 0;
    }
    const int rc = rgbdfe_frames_match(rgbdfe_host::ctx(), fq, ft, nq, nt, outl.data(), mfNNratio, discardOutliers ? 1 : 0,
        reinterpret_cast<rgbdfe_dmatch*>(vMatches12.data()), (int)vMatches12.size(), &n);
    if (rgbdfe_host::report("rgbdfe_frames_match", rc) != RGBDFE_OK)
        n = 0; // errors map to "no matches", never to a CPU path
    vMatches12.resize(n);
    for (auto& m : vMatches12)
        m.imgIdx = 0; // what BFMatcher::knnMatch reports for a single train image
    return int(vMatches12.size());
}

int Matcher::projectionMatch(Frame::Ptr F1, Frame::Ptr F2, vector<cv::DMatch>& vMatches12, const float th)
{
    // the reference appends to vMatches12 and returns its size (Matcher.cpp:96-103); so does the shim
    const int nq = (int)F1->N, nt = (int)F2->N;
    if (nq == 0 || nt == 0)
        return int(vMatches12.size());
    rgbdfe_projection P;
    P.fx = F1->mpCamera->fx(), P.fy = F1->mpCamera->fy(), P.cx = F1->mpCamera->cx(), P.cy = F1->mpCamera->cy();
    const cv::Mat Rwc = F1->getRotationInverse(), Ow = F1->getCameraCenter(); // what Frame::unprojectWorld uses (Frame.cpp:317-327)
    const cv::Mat Rcw = F2->getRotation(), tcw = F2->getTranslation();
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) {
            P.Rwc1[r * 3 + c] = Rwc.at<float>(r, c);
            P.Rcw2[r * 3 + c] = Rcw.at<float>(r, c);
        }
        P.Ow1[r] = Ow.at<float>(r, 0);
        P.tcw2[r] = tcw.at<float>(r, 0);
    }
    P.min_x = Frame::mnMinX, P.max_x = Frame::mnMaxX, P.min_y = Frame::mnMinY, P.max_y = Frame::mnMaxY;
    P.grid_w_inv = Frame::mfGridElementWidthInv, P.grid_h_inv = Frame::mfGridElementHeightInv;
    P.th = th;
    vector<uint8_t> outl(nq);
    for (int i = 0; i < nq; ++i)
        outl[i] = F1->isOutlier(i) ? 1 : 0;
    vector<float> un((size_t)nt * 2);
    for (int i = 0; i < nt; ++i)
        un[2 * i] = F2->mvKeysUn[i].pt.x, un[2 * i + 1] = F2->mvKeysUn[i].pt.y;
    cv::Mat qd = F1->mDescriptors.isContinuous() ? F1->mDescriptors : F1->mDescriptors.clone();
    cv::Mat td = F2->mDescriptors.isContinuous() ? F2->mDescriptors : F2->mDescriptors.clone();
    vector<cv::DMatch> found(nq);
    int n = 0;
    const int rc = rgbdfe_projection_match(rgbdfe_host::ctx(), qd.data, nq, reinterpret_cast<const rgbdfe_point3f*>(F1->mvKeys3Dc.data()), outl.data(),
        td.data, nt, reinterpret_cast<const rgbdfe_point3f*>(F2->mvKeys3Dc.data()), un.data(), &P, reinterpret_cast<rgbdfe_dmatch*>(found.data()),
        nq, &n);
    if (rgbdfe_host::report("rgbdfe_projection_match", rc) != RGBDFE_OK)
        n = 0;
    vMatches12.insert(vMatches12.end(), found.begin(), found.begin() + n);
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
