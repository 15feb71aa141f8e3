// ORACLE — TEST INFRASTRUCTURE ONLY.  C entry points for ctypes (tests/,
// smoke(), bench.py cpu_baseline / --impl reference).  Never linked into the
// product library.
#include "oracle.h"

#include <algorithm>
#include <chrono>
#include <cstring>

using namespace orc;

extern "C" {

// ---------------- cv::ORB branch (row N3) ----------------
void* orc_cvorb_create(int nfeatures, float scale, int nlevels) { return new CvOrb(nfeatures, scale, nlevels); }
void orc_cvorb_destroy(void* h) { delete (CvOrb*)h; }
int orc_cvorb_extract(void* h, const uint8_t* img, int w, int ht, int stride, KeyPoint* kps, uint8_t* desc, int cap)
{
    CvOrb* e = (CvOrb*)h;
    std::vector<KeyPoint> k;
    std::vector<uint8_t> d;
    e->extract(img, w, ht, stride, k, d);
    int n = std::min((int)k.size(), cap);
    if (n > 0) {
        memcpy(kps, k.data(), (size_t)n * sizeof(KeyPoint));
        memcpy(desc, d.data(), (size_t)n * 32);
    }
    return (int)k.size();
}
void orc_cvorb_level_size(void* h, int level, int* w, int* ht, int* quota, int* n_fast, int* n_harris)
{
    CvOrb* e = (CvOrb*)h;
    *w = e->levels[level].w;
    *ht = e->levels[level].h;
    *quota = e->quota.empty() ? 0 : e->quota[level];
    *n_fast = e->levels[level].n_fast;
    *n_harris = (int)e->levels[level].harris_in.size();
}
// un-padded level image / blurred level (w x h, tight) of the last extract()
void orc_cvorb_level_image(void* h, int level, uint8_t* dst, int blurred)
{
    CvOrb* e = (CvOrb*)h;
    const CvOrbLevel& L = e->levels[level];
    for (int y = 0; y < L.h; y++)
        memcpy(dst + (size_t)y * L.w, blurred ? L.blurred.data() + (size_t)y * L.w : L.image() + (size_t)y * L.stride, L.w);
}
int orc_cvorb_level_harris(void* h, int level, KeyPoint* out, int cap)
{
    CvOrb* e = (CvOrb*)h;
    const auto& v = e->levels[level].harris_in;
    int n = std::min((int)v.size(), cap);
    if (n > 0)
        memcpy(out, v.data(), (size_t)n * sizeof(KeyPoint));
    return (int)v.size();
}
void orc_gauss7_float(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) { gauss7_float_u8(src, w, h, sstride, dst, dstride); }
void orc_resize_linear_exact(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride)
{
    resize_linear_exact_u8(src, sw, sh, sstride, dst, dw, dh, dstride);
}

// ---------------- ORB ----------------
void* orc_orb_create(int nfeatures, float scale, int nlevels, int iniTh, int minTh)
{
    return new OrbExtractor(nfeatures, scale, nlevels, iniTh, minTh);
}
void orc_orb_destroy(void* h) { delete (OrbExtractor*)h; }

// returns the number of keypoints (may exceed cap; only cap are written)
int orc_orb_extract(void* h, const uint8_t* img, int w, int ht, int stride, KeyPoint* kps, uint8_t* desc, int cap)
{
    OrbExtractor* e = (OrbExtractor*)h;
    std::vector<KeyPoint> k;
    std::vector<uint8_t> d;
    e->extract(img, w, ht, stride, k, d);
    int n = std::min((int)k.size(), cap);
    if (n > 0) {
        memcpy(kps, k.data(), (size_t)n * sizeof(KeyPoint));
        memcpy(desc, d.data(), (size_t)n * 32);
    }
    return (int)k.size();
}
void orc_orb_tables(void* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* quota, int* umax)
{
    OrbExtractor* e = (OrbExtractor*)h;
    for (int i = 0; i < e->nlevels; i++) {
        scale[i] = e->mvScaleFactor[i];
        inv_scale[i] = e->mvInvScaleFactor[i];
        sigma2[i] = e->mvLevelSigma2[i];
        inv_sigma2[i] = e->mvInvLevelSigma2[i];
        quota[i] = e->mnFeaturesPerLevel[i];
    }
    for (int i = 0; i < 16; i++)
        umax[i] = e->umax[i];
}
void orc_orb_level_size(void* h, int level, int* w, int* ht)
{
    OrbExtractor* e = (OrbExtractor*)h;
    *w = e->levels[level].w;
    *ht = e->levels[level].h;
}
// padded level image ((w+38) x (h+38)) of the last extract()
void orc_orb_level_padded(void* h, int level, uint8_t* dst)
{
    OrbExtractor* e = (OrbExtractor*)h;
    memcpy(dst, e->levels[level].padded.data(), e->levels[level].padded.size());
}
int orc_orb_level_blurred(void* h, int level, uint8_t* dst)
{
    OrbExtractor* e = (OrbExtractor*)h;
    if (e->levels[level].blurred.empty())
        return 0;
    memcpy(dst, e->levels[level].blurred.data(), e->levels[level].blurred.size());
    return 1;
}
int orc_orb_level_candidates(void* h, int level, KeyPoint* out, int cap)
{
    OrbExtractor* e = (OrbExtractor*)h;
    const std::vector<KeyPoint>& c = e->levels[level].candidates;
    int n = std::min((int)c.size(), cap);
    if (n > 0)
        memcpy(out, c.data(), (size_t)n * sizeof(KeyPoint));
    return (int)c.size();
}
int orc_orb_level_keypoints(void* h, int level, KeyPoint* out, int cap)
{
    OrbExtractor* e = (OrbExtractor*)h;
    const std::vector<KeyPoint>& c = e->levels[level].keypoints;
    int n = std::min((int)c.size(), cap);
    if (n > 0)
        memcpy(out, c.data(), (size_t)n * sizeof(KeyPoint));
    return (int)c.size();
}
int orc_quadtree(const KeyPoint* in, int n, int minX, int maxX, int minY, int maxY, int N, KeyPoint* out, int cap)
{
    std::vector<KeyPoint> k(in, in + n);
    std::vector<KeyPoint> r = OrbExtractor::distribute_quadtree(k, minX, maxX, minY, maxY, N);
    int m = std::min((int)r.size(), cap);
    if (m > 0)
        memcpy(out, r.data(), (size_t)m * sizeof(KeyPoint));
    return (int)r.size();
}

// ---------------- Matcher ----------------
int orc_match(const uint8_t* qdesc, int nq, const uint8_t* tdesc, int nt, const uint8_t* q_outlier, const Point3f* q3d,
    const Point3f* t3d, float ratio, int discard_outliers, DMatch* out, int cap)
{
    std::vector<DMatch> m;
    matcher_match(qdesc, nq, tdesc, nt, q_outlier, q3d, t3d, ratio, discard_outliers, m);
    int n = std::min((int)m.size(), cap);
    if (n > 0)
        memcpy(out, m.data(), (size_t)n * sizeof(DMatch));
    return (int)m.size();
}

// ---------------- RansacSE3 ----------------
struct OrcSe3Out {
    int ok;
    float rmse;
    float T[16];
    int n_inliers;
    int iters_executed;
    long es, ef, adds, svds, rounds;
};
// rng: GlibcRand state (in/out). latch: {latched, cov} (in/out).
void orc_ransac_se3(const Point3f* p1, const Point3f* p2, const DMatch* m, int nm, int iters, unsigned min_inl, float max_mahal,
    unsigned sample, GlibcRand* rng, int* latched, double* latch_cov, OrcSe3Out* out, DMatch* inliers, int cap)
{
    std::vector<DMatch> mv(m, m + nm);
    Se3Params prm = { iters, min_inl, max_mahal, sample };
    DepthCovLatch latch = { *latched, *latch_cov };
    Se3Result r;
    Se3Stats st = { 0, 0, 0, 0, 0, 0 };
    ransac_se3(p1, p2, mv, prm, *rng, latch, r, &st);
    *latched = latch.latched;
    *latch_cov = latch.cov;
    out->ok = r.ok;
    out->rmse = r.rmse;
    memcpy(out->T, r.T, sizeof(r.T));
    out->n_inliers = (int)r.inliers.size();
    out->iters_executed = r.iters_executed;
    out->es = st.es, out->ef = st.ef, out->adds = st.adds, out->svds = st.svds, out->rounds = st.rounds;
    int n = std::min((int)r.inliers.size(), cap);
    if (n > 0)
        memcpy(inliers, r.inliers.data(), (size_t)n * sizeof(DMatch));
}
void orc_tfc_fit(const Point3f* p1, const Point3f* p2, const DMatch* m, int n, float* T) { tfc_fit(p1, p2, m, n, T); }
double orc_mahal2(const Point3f* a, const Point3f* b, const float* T4f, int* latched, double* cov)
{
    double T[16];
    for (int i = 0; i < 16; i++)
        T[i] = T4f[i];
    DepthCovLatch l = { *latched, *cov };
    double r = mahal2(*a, *b, T, l, nullptr);
    *latched = l.latched;
    *cov = l.cov;
    return r;
}
void orc_jacobi_svd3(const float* A, float* U, float* S, float* V) { jacobi_svd3(A, U, S, V); }
// std::sort with DMatch::operator< (distance only), as SolverSE3.cpp:52
void orc_sort_matches(DMatch* m, int n)
{
    std::sort(m, m + n, [](const DMatch& a, const DMatch& b) { return a.distance < b.distance; });
}

// ---------------- Ransac (Umeyama) ----------------
struct OrcUmeOut {
    int ok;
    float T[16];
    int n_inliers;
    int iters_executed;
};
void orc_ransac_umeyama(const Point3f* p1, const Point3f* p2, const DMatch* m, int nm, int errorVersion, double thEuclid,
    double minRatio, int minMatches, int usedPairs, GlibcRand* rng, OrcUmeOut* out, DMatch* inliers, int cap)
{
    std::vector<DMatch> mv(m, m + nm);
    UmeyamaParams prm = { errorVersion, thEuclid, minRatio, minMatches, usedPairs };
    UmeyamaResult r;
    ransac_umeyama(p1, p2, mv, prm, *rng, r);
    out->ok = r.ok;
    memcpy(out->T, r.T, sizeof(r.T));
    out->n_inliers = (int)r.inliers.size();
    out->iters_executed = r.iters_executed;
    int n = std::min((int)r.inliers.size(), cap);
    if (n > 0)
        memcpy(inliers, r.inliers.data(), (size_t)n * sizeof(DMatch));
}
void orc_umeyama3(const float* src, const float* dst, int n, float* T) { umeyama3(src, dst, n, T); }
}

namespace orc {
thread_local double g_stage_s[ST_COUNT];
double StageTimer::now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
}

// ---------------- timed CPU baseline: the whole tracked-frame path ----------------
// Frame ctor hot part (Core/Frame.cpp:51,91-117, no lens distortion) + Tracking::visualOdometry
// core (System/Tracking.cpp:125-130) over a sequence, single thread.  Returns frames tracked.
// stage_s (may be NULL): ST_COUNT seconds per stage of THIS call; iters_out (may be NULL): executed RANSAC iterations per pair.
extern "C" int orc_track_sequence2(int nframes, const uint8_t* const* gray, const float* const* depth, int w, int h, float fx, float fy, float cx,
    float cy, int nfeatures, float ratio, int iters, unsigned min_inl, float max_mahal, unsigned sample, unsigned seed, int* inliers_out,
    int* iters_out, double* stage_s, long long* se3_stats);
extern "C" int orc_track_sequence(int nframes, const uint8_t* const* gray, const float* const* depth, int w, int h, float fx, float fy, float cx,
    float cy, int nfeatures, float ratio, int iters, unsigned min_inl, float max_mahal, unsigned sample, unsigned seed, int* inliers_out)
{
    return orc_track_sequence2(nframes, gray, depth, w, h, fx, fy, cx, cy, nfeatures, ratio, iters, min_inl, max_mahal, sample, seed, inliers_out,
        nullptr, nullptr, nullptr);
}
extern "C" int orc_track_sequence2(int nframes, const uint8_t* const* gray, const float* const* depth, int w, int h, float fx, float fy, float cx,
    float cy, int nfeatures, float ratio, int iters, unsigned min_inl, float max_mahal, unsigned sample, unsigned seed, int* inliers_out,
    int* iters_out, double* stage_s, long long* se3_stats)
{
    // se3_stats (may be NULL): totals over the call of {E_s, E_f, A, V, rounds, pairs}: shortcut / full evaluations of
    // errorFunction2, TFC adds, SVDs, refinement rounds (the operation counts of SURVEY 8d's RANSAC flop model)
    long long tot[6] = { 0, 0, 0, 0, 0, 0 };
    for (int i = 0; i < ST_COUNT; i++)
        g_stage_s[i] = 0.0;
    OrbExtractor orb(nfeatures, 1.2f, 8, 20, 7);
    GlibcRand rng;
    rng.seed(seed);
    DepthCovLatch latch = { 0, 0.0 };
    std::vector<KeyPoint> pk, ck;
    std::vector<uint8_t> pd, cd, po, co;
    std::vector<Point3f> pp, cp;
    const float invfx = 1.0f / fx, invfy = 1.0f / fy;
    int tracked = 0;
    for (int f = 0; f < nframes; f++) {
        orb.extract(gray[f], w, h, w, ck, cd);
        cp.assign(ck.size(), Point3f { 0, 0, 0 });
        co.assign(ck.size(), 0);
        for (size_t i = 0; i < ck.size(); i++) {
            const float u = ck[i].x, v = ck[i].y;
            const float z = depth[f][(size_t)(int)v * w + (int)u];
            if (z > 0)
                cp[i] = Point3f { (u - cx) * z * invfx, (v - cy) * z * invfy, z };
        }
        if (f > 0) {
            std::vector<DMatch> m;
            matcher_match(pd.data(), (int)pk.size(), cd.data(), (int)ck.size(), po.data(), pp.data(), cp.data(), ratio, 1, m);
            Se3Params prm = { iters, min_inl, max_mahal, sample };
            Se3Result r;
            if (m.size() >= min_inl)
                for (const DMatch& mm : m)
                    co[mm.trainIdx] = 1;
            {
                StageTimer t(ST_RANSAC);
                Se3Stats st = {};
                ransac_se3(pp.data(), cp.data(), m, prm, rng, latch, r, &st);
                tot[0] += st.es, tot[1] += st.ef, tot[2] += st.adds, tot[3] += st.svds, tot[4] += st.rounds, tot[5]++;
            }
            if (r.ok)
                for (const DMatch& mm : r.inliers)
                    co[mm.trainIdx] = 0;
            if (inliers_out)
                inliers_out[f - 1] = (int)r.inliers.size();
            if (iters_out)
                iters_out[f - 1] = r.iters_executed;
            tracked++;
        }
        pk.swap(ck), pd.swap(cd), pp.swap(cp), po.swap(co);
    }
    if (stage_s)
        for (int i = 0; i < ST_COUNT; i++)
            stage_s[i] = g_stage_s[i];
    if (se3_stats)
        for (int i = 0; i < 6; i++)
            se3_stats[i] = tot[i];
    return tracked;
}

// ---------------- LoopDetector (row N4) ----------------
// `sort(possibleLoops.begin(), possibleLoops.end(), Frame::CompareBoWscore)` (PlaceRecognition/LoopDetector.cpp:79) with the
// REAL std::sort of this image's libstdc++: scores descending, ties in whatever order introsort leaves them.
extern "C" void orc_sort_bow_scores(double* score, int* idx, int n)
{
    struct E {
        double s;
        int k;
    };
    std::vector<E> v(n);
    for (int i = 0; i < n; i++)
        v[i] = E { score[i], idx[i] };
    std::sort(v.begin(), v.end(), [](const E& a, const E& b) { return a.s > b.s; });
    for (int i = 0; i < n; i++)
        score[i] = v[i].s, idx[i] = v[i].k;
}

// ---------------- numeric pins (tests/test_numeric_pins.py): the third-party restatements, batch form ----------------
extern "C" {
// JacobiSVD<Matrix3f> restatement on n row-major 3x3 matrices.
void orc_jacobi_svd3_batch(int n, const float* A, float* U, float* S, float* V)
{
    for (int i = 0; i < n; i++)
        jacobi_svd3(A + 9 * i, U + 9 * i, S + 3 * i, V + 9 * i);
}
// errorFunction2 restatement with the depth covariance pinned to `cov`; out[i] = squared Mahalanobis distance of
// pair i under T (row-major 4x4 double), full[i] = 1 when the LLT solve ran (0: NaN gate or the error shortcut).
void orc_mahal2_batch(int n, const Point3f* a, const Point3f* b, const double* T16, double cov, double* out, int* full)
{
    for (int i = 0; i < n; i++) {
        DepthCovLatch l;
        l.latched = 1;
        l.cov = cov;
        out[i] = mahal2(a[i], b[i], T16, l, full + i);
    }
}
}
