// Caller-side test of the class shims: the hot part of Tracking::visualOdometry
// (System/Tracking.cpp:121-163) and of the Frame constructor (Core/Frame.cpp:34-73), written
// against the reference's class interfaces exactly as the reference's callers use them.
//   host_track <in.bin> <out.txt> [seed] [orb]      ("orb": Extractor(ORB, ORB) = cv::ORB instead of ORB2)
// in.bin : int32 n, w, h, nfeatures; float fx, fy, cx, cy; then n x (gray w*h u8, depth w*h f32 metres)
// out.txt: one block per frame (keypoint / descriptor checksums, match + RANSAC results, outlier flags)
#include "Core/Frame.h"
#include "Features/Extractor.h"
#include "Features/Matcher.h"
#include "Solver/Ransac.h"
#include "Solver/SolverSE3.h"
#include "System/Random.h"
#include "rgbdfe_host.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

using namespace std;

static unsigned long long fnv(const void* p, size_t n, unsigned long long h = 1469598103934665603ULL)
{
    const unsigned char* b = (const unsigned char*)p;
    for (size_t i = 0; i < n; i++) {
        h ^= b[i];
        h *= 1099511628211ULL;
    }
    return h;
}

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// host_track <in.bin> <out.json> <seed> <orb|orb2> bench <frames>: the drop-in caller's loop (one camera, blocking calls
// through the class interfaces, Tracking.cpp:121-163) timed per call over `frames` frames of the ping-ponged sequence.
static int run_bench(FILE* f, const char* out_path, int n, int w, int h, int nfeat, const float* intr, unsigned seed, bool cvorb, int total)
{
    vector<cv::Mat> grays, depths;
    for (int i = 0; i < n; i++) {
        cv::Mat gray(h, w, CV_8UC1), depth(h, w, CV_32F);
        if (fread(gray.data, 1, (size_t)w * h, f) != (size_t)w * h || fread(depth.data, 4, (size_t)w * h, f) != (size_t)w * h)
            return 2;
        grays.push_back(gray), depths.push_back(depth);
    }
    rgbdfe_host::set_capacity(w, h, nfeat);
    rgbdfe_host::seed(seed);
    RGBDcamera cam = { intr[0], intr[1], intr[2], intr[3], 1.0f };
    const Extractor::eType et = cvorb ? Extractor::ORB : Extractor::ORB2;
    Extractor::Ptr extractor(new Extractor(et, et, Extractor::NORMAL));
    extractor->setParameters(nfeat, 1.2f, 8, 20, 7);
    double t_frame = 0, t_match = 0, t_sac = 0;
    long matches = 0, inliers = 0, ok = 0;
    Frame::Ptr ref;
    const int warm = 8;
    double t_begin = 0;
    for (int i = 0; i < total + warm; i++) {
        if (i == warm) {
            t_frame = t_match = t_sac = 0, matches = inliers = ok = 0;
            t_begin = now_s();
        }
        int k = i % (2 * n - 2);
        if (k >= n)
            k = 2 * n - 2 - k;
        const double t0 = now_s();
        Frame::Ptr cur(new Frame(grays[k], depths[k], i / 30.0, extractor, &cam));
        const double t1 = now_s();
        t_frame += t1 - t0;
        if (ref) {
            vector<cv::DMatch> vMatches;
            Matcher matcher(0.9f);
            matcher.match(ref, cur, vMatches);
            const double t2 = now_s();
            RansacSE3 sac(200, 10, 3.0f, 4); // Tracking.cpp:129
            const bool b = sac.compute(ref, cur, vMatches);
            const double t3 = now_s();
            t_match += t2 - t1, t_sac += t3 - t2;
            matches += (long)vMatches.size(), inliers += (long)sac.mvInliers.size(), ok += b ? 1 : 0;
        }
        ref = cur;
    }
    const double dt = now_s() - t_begin;
    FILE* out = fopen(out_path, "w");
    fprintf(out, "{\"frames\": %d, \"frames_per_s\": %.2f, \"ms_per_frame\": {\"Frame_ctor_detectAndCompute\": %.4f, \"Matcher_match\": %.4f, "
                 "\"RansacSE3_compute\": %.4f}, \"mean_matches\": %.1f, \"mean_inliers\": %.1f, \"ok_frac\": %.3f, \"nfeatures\": %d}\n",
        total, total / dt, 1e3 * t_frame / total, 1e3 * t_match / total, 1e3 * t_sac / total, (double)matches / total, (double)inliers / total,
        (double)ok / total, nfeat);
    fclose(out);
    return 0;
}

int main(int argc, char** argv)
{
    if (argc < 3) {
        fprintf(stderr, "usage: host_track in.bin out.txt [seed]\n");
        return 2;
    }
    FILE* f = fopen(argv[1], "rb");
    if (!f)
        return 2;
    int hdr[4];
    float intr[4];
    if (fread(hdr, 4, 4, f) != 4 || fread(intr, 4, 4, f) != 4)
        return 2;
    const int n = hdr[0], w = hdr[1], h = hdr[2], nfeat = hdr[3];
    const unsigned seed = argc > 3 ? (unsigned)atoi(argv[3]) : 42u;
    if (argc > 6 && std::string(argv[5]) == "bench")
        return run_bench(f, argv[2], n, w, h, nfeat, intr, seed, std::string(argv[4]) == "orb", atoi(argv[6]));

    rgbdfe_host::set_capacity(w, h, nfeat);
    rgbdfe_host::seed(seed); // tests pin the stream; the reference seeds from time(NULL)

    RGBDcamera cam = { intr[0], intr[1], intr[2], intr[3], 1.0f };
    const bool cvorb = argc > 4 && std::string(argv[4]) == "orb";
    const Extractor::eType et = cvorb ? Extractor::ORB : Extractor::ORB2;
    Extractor::Ptr extractor(new Extractor(et, et, Extractor::NORMAL)); // main.cpp:31 with the ORB path
    extractor->setParameters(nfeat, 1.2f, 8, 20, 7);

    FILE* out = fopen(argv[2], "w");
    fprintf(out, "levels %d scale %.9g norm %d\n", extractor->getLevels(), extractor->getScaleFactor(), Extractor::mNorm);
    vector<float> sf = extractor->getScaleFactors(), s2 = extractor->getInverseScaleSigmaSquares();
    fprintf(out, "scale7 %.9g invsigma7 %.9g\n", sf.back(), s2.back());

    Frame::Ptr ref;
    for (int i = 0; i < n; i++) {
        cv::Mat gray(h, w, CV_8UC1), depth(h, w, CV_32F);
        if (fread(gray.data, 1, (size_t)w * h, f) != (size_t)w * h || fread(depth.data, 4, (size_t)w * h, f) != (size_t)w * h)
            return 2;
        Frame::Ptr cur(new Frame(gray, depth, i / 30.0, extractor, &cam));
        fprintf(out, "frame %d N %zu kp %016llx desc %016llx xyz %016llx\n", i, cur->N, fnv(cur->mvKeys.data(), cur->N * sizeof(cv::KeyPoint)),
            fnv(cur->mDescriptors.data, cur->N * 32), fnv(cur->mvKeys3Dc.data(), cur->N * sizeof(cv::Point3f)));
        if (ref && cur->N > 1 && ref->N > 1) {
            // Matcher::descriptorDistance (Matcher.cpp:141-144) on three descriptor pairs; the test recomputes them from the oracle's descriptors
            fprintf(out, "ddist %d %.1f %.1f %.1f\n", i, Matcher::descriptorDistance(ref->mDescriptors.row(0), cur->mDescriptors.row(0)),
                Matcher::descriptorDistance(ref->mDescriptors.row(1), cur->mDescriptors.row((int)cur->N - 1)),
                Matcher::descriptorDistance(cur->mDescriptors.row(0), cur->mDescriptors.row(0)));
        }
        if (ref) {
            // Tracking::visualOdometry (Tracking.cpp:125-130)
            vector<cv::DMatch> vMatches;
            Matcher matcher(0.9f);
            matcher.match(ref, cur, vMatches);
            RansacSE3 sac(200, 10, 3.0f, 4);
            bool b = sac.compute(ref, cur, vMatches);
            fprintf(out, "track %d matches %zu m %016llx ok %d inliers %zu inl %016llx rmse %.9g T", i, vMatches.size(),
                fnv(vMatches.data(), vMatches.size() * sizeof(cv::DMatch)), (int)b, sac.mvInliers.size(),
                fnv(sac.mvInliers.data(), sac.mvInliers.size() * sizeof(cv::DMatch)), sac.rmse);
            for (int r = 0; r < 4; r++)
                for (int c = 0; c < 4; c++)
                    fprintf(out, " %.9g", sac.mT21(r, c));
            fprintf(out, "\n");
            cv::Mat pose = cur->getPose();
            fprintf(out, "pose %d", i);
            for (int r = 0; r < 4; r++)
                for (int c = 0; c < 4; c++)
                    fprintf(out, " %.9g", pose.at<float>(r, c));
            fprintf(out, "\n");
            int nout = 0;
            for (size_t k = 0; k < cur->N; k++)
                nout += cur->isOutlier(k) ? 1 : 0;
            fprintf(out, "outliers %d %d\n", i, nout);
            // Matcher::projectionMatch with the poses the solver just set (dead code in the reference, scope row N2);
            // the pose blocks are printed so that the test can hand the oracle the very same floats
            vector<cv::DMatch> vProj;
            const int np = matcher.projectionMatch(ref, cur, vProj, 5.0f);
            fprintf(out, "proj %d n %d m %016llx P", i, np, fnv(vProj.data(), vProj.size() * sizeof(cv::DMatch)));
            const cv::Mat blocks[4] = { ref->getRotationInverse(), ref->getCameraCenter(), cur->getRotation(), cur->getTranslation() };
            for (const cv::Mat& b : blocks)
                for (int r = 0; r < b.rows; r++)
                    for (int c = 0; c < b.cols; c++)
                        fprintf(out, " %.9g", b.at<float>(r, c));
            fprintf(out, "\n");
            if (i == 1) {
                // class Ransac on the same matches (dead code in the reference, named by the north star);
                // a copy of the frame keeps the tracking state untouched
                Frame::Ptr cur2(new Frame(*cur));
                Ransac::parameters prm = {};
                prm.errorVersion = Ransac::EUCLIDEAN_ERROR;
                prm.inlierThresholdEuclidean = 0.02;
                prm.minimalInlierRatioThreshold = 0.2;
                prm.minimalNumberOfMatches = 10;
                prm.usedPairs = 3;
                vector<cv::DMatch> inl;
                Ransac r2(ref, cur2, vMatches, prm);
                bool ok2 = r2.compute(inl);
                fprintf(out, "umeyama ok %d inliers %zu inl %016llx\n", (int)ok2, inl.size(), fnv(inl.data(), inl.size() * sizeof(cv::DMatch)));
                fprintf(out, "randomInt %d\n", Random::randomInt(0, 999));
            }
        }
        ref = cur;
    }
    fclose(out);
    fclose(f);
    return 0;
}
