// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
// arm may load this library.  Parity status: OpenCV primitives pinned against
// cv2 4.13.0; glibc rand/sinf/cosf pinned against this image's glibc 2.39;
// PCL TransformationFromCorrespondences / Eigen JacobiSVD / LLT / umeyama are
// restated from their published algorithms -> "parity unpinned" for those.
#ifndef ORC_ORACLE_H
#define ORC_ORACLE_H

#include <cstdint>
#include <vector>

namespace orc {

// Layout-compatible with cv::KeyPoint (28 B), cv::DMatch (16 B), cv::Point3f.
struct KeyPoint {
    float x, y, size, angle, response;
    int octave, class_id;
};
struct DMatch {
    int queryIdx, trainIdx, imgIdx;
    float distance;
};
struct Point3f {
    float x, y, z;
};

// ---- cvprim.cpp ----
// per-thread stage timers of the CPU baseline (bench.py): seconds spent per stage since the last reset
enum { ST_PYRAMID = 0, ST_FAST, ST_QUADTREE, ST_ANGLE, ST_BLUR, ST_DESCRIBE, ST_KNN, ST_FILTER, ST_RANSAC, ST_COUNT };
extern thread_local double g_stage_s[ST_COUNT];
struct StageTimer {
    int id;
    double t0;
    static double now();
    explicit StageTimer(int i) : id(i), t0(now()) {}
    ~StageTimer() { g_stage_s[id] += now() - t0; }
};
int cv_round(float v);
int cv_round_d(double v);
int cv_floor(float v);
void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride);
void border_reflect101(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int b);
int fast9(const uint8_t* img, int w, int h, int stride, int threshold, int nonmax, int* out_xys, int cap);
void gauss7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);
float fast_atan2(float y, float x);
float sinf_glibc(float y);
float cosf_glibc(float y);
int hamming256(const uint8_t* a, const uint8_t* b);
void knn2_hamming(const uint8_t* q, int nq, const uint8_t* t, int nt, int* idx, int* dist);

// ---- rng.cpp : glibc rand()/srand() TYPE_3 + System/Random.cpp:15-20 ----
struct GlibcRand {
    int32_t r[34];
    int f, b; // front / rear indices into r[3..33] ring (31 words)
    void seed(unsigned s);
    int next(); // rand()
    int random_int(int mn, int mx); // Random::randomInt
};

// ---- orb.cpp : Features/ORBextractor.cpp + ExtractorNode.cpp ----
struct OrbLevel {
    int w, h; // un-padded level size
    int stride; // padded row stride (w + 38)
    std::vector<uint8_t> padded; // (w+38) x (h+38)
    std::vector<uint8_t> blurred; // w x h (only if the level has keypoints)
    std::vector<KeyPoint> candidates; // per-cell FAST output (coords rel. to minBorder)
    std::vector<KeyPoint> keypoints; // after quadtree + orientation (level coords)
    const uint8_t* image() const { return padded.data() + 19 * stride + 19; }
};

class OrbExtractor {
public:
    OrbExtractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
    // ORBextractor::operator()  (Features/ORBextractor.cpp:706-766)
    void extract(const uint8_t* img, int w, int h, int stride, std::vector<KeyPoint>& kps, std::vector<uint8_t>& desc);

    int nfeatures;
    double scaleFactor;
    int nlevels, iniThFAST, minThFAST;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
    std::vector<int> mnFeaturesPerLevel, umax;
    std::vector<OrbLevel> levels;

    void compute_pyramid(const uint8_t* img, int w, int h, int stride);
    void compute_keypoints();
    static std::vector<KeyPoint> distribute_quadtree(const std::vector<KeyPoint>& keys, int minX, int maxX, int minY, int maxY, int N);
    float ic_angle(const OrbLevel& L, float px, float py) const;
    static void describe(const uint8_t* img, int stride, const KeyPoint& kp, uint8_t* desc);
};

// ---- match.cpp : Features/Matcher.cpp:106-139 ----
int matcher_match(const uint8_t* qdesc, int nq, const uint8_t* tdesc, int nt, const uint8_t* q_outlier,
    const Point3f* q3d, const Point3f* t3d, float ratio, int discard_outliers, std::vector<DMatch>& out);

// ---- ransac_se3.cpp : Solver/SolverSE3.cpp ----
struct Se3Stats { // work counters for the roofline
    long es, ef, adds, svds, iters, rounds;
};
struct Se3Result {
    int ok;
    float rmse;
    float T[16]; // row-major 4x4 (mT21)
    std::vector<DMatch> inliers;
    int iters_executed;
};
struct Se3Params {
    int iterations;
    unsigned min_inlier_th;
    float max_mahal;
    unsigned sample_size;
};
// depth-covariance latch (SolverSE3.cpp:282-287): <0 = not latched yet.
struct DepthCovLatch {
    int latched;
    double cov;
};
void ransac_se3(const Point3f* p1, const Point3f* p2, const std::vector<DMatch>& m12, const Se3Params& prm,
    GlibcRand& rng, DepthCovLatch& latch, Se3Result& res, Se3Stats* stats);
void tfc_fit(const Point3f* p1, const Point3f* p2, const DMatch* m, int n, float T[16]);
double mahal2(const Point3f& a, const Point3f& b, const double T[16], DepthCovLatch& latch, int* full);
void jacobi_svd3(const float A[9], float U[9], float S[3], float V[9]);

// ---- ransac_umeyama.cpp : Solver/Ransac.cpp ----
struct UmeyamaParams {
    int errorVersion; // 0 EUCLIDEAN, 4 ADAPTIVE
    double inlierThresholdEuclidean;
    double minimalInlierRatioThreshold;
    int minimalNumberOfMatches;
    int usedPairs;
};
struct UmeyamaResult {
    int ok; // !T.isIdentity()
    float T[16];
    std::vector<DMatch> inliers;
    std::vector<int> outlier_marked; // train indices setOutlier'ed
    int iters_executed;
};
void ransac_umeyama(const Point3f* p1, const Point3f* p2, const std::vector<DMatch>& m12, const UmeyamaParams& prm,
    GlibcRand& rng, UmeyamaResult& res);
void umeyama3(const float* src, const float* dst, int n, float T[16]);

// ---- cvorb.cpp : the Extractor(ORB, ORB) branch = cv::ORB (Features/Extractor.cpp:55-59,163-166,216-218) ----
void resize_linear_exact_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride);
void gauss7_float_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);
struct CvOrbLevel {
    int w, h, stride; // stride = w + 2 * border
    float scale;
    int n_fast;       // FAST corners inside the border, before retainBest
    std::vector<uint8_t> padded, blurred;
    std::vector<KeyPoint> harris_in; // after retainBest(2 N_l) + Harris response, level coordinates
    const uint8_t* image() const { return padded.data() + 32 * stride + 32; }
};
class CvOrb {
public:
    CvOrb(int nfeatures_, float scaleFactor_, int nlevels_) : nfeatures(nfeatures_), scaleFactor((double)scaleFactor_), nlevels(nlevels_) {}
    // detect + compute as Extractor::detectAndCompute does for the ORB/ORB pair; the descriptor object is a default
    // cv::ORB (scale 1.2f), so the restatement shares one pyramid and requires scaleFactor == 1.2f
    void extract(const uint8_t* img, int w, int h, int stride, std::vector<KeyPoint>& kps, std::vector<uint8_t>& desc);
    int nfeatures;
    double scaleFactor;
    int nlevels;
    int edgeThreshold = 31, fastThreshold = 20, border = 32;
    std::vector<int> quota;
    std::vector<CvOrbLevel> levels;
};

} // namespace orc
#endif
