// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement of the third-party primitives the reference's hot path calls
// but does not vendor (OpenCV, glibc libm).  The reference links an unpinned
// OpenCV (CMakeLists.txt:31); this restatement is pinned against the only
// OpenCV build available offline: opencv-python-headless 4.13.0 (tests/
// test_oracle_cv2.py checks every function here against cv2 bit-for-bit).
//
//   orc_resize_linear_u8   <- cv::resize(INTER_LINEAR), 8-bit, 11-bit fixed-point
//                             coefficients       (call site Features/ORBextractor.cpp:786)
//   orc_border_reflect101  <- cv::copyMakeBorder(BORDER_REFLECT_101)   (:789-794)
//   orc_fast9              <- cv::FAST(roi, kps, th, true), TYPE_9_16  (:655-660)
//   orc_gauss7_u8          <- cv::GaussianBlur(7x7, 2, 2, REFLECT_101) (:746)
//   orc_fast_atan2         <- cv::fastAtan2                            (:40)
//   orc_cv_round           <- cvRound(float)                           (:19,51,56)
//   orc_sinf / orc_cosf    <- glibc 2.39 sinf/cosf as selected by
//                             `using namespace std; cos(float)`        (:48)
//   orc_knn2_hamming       <- cv::BFMatcher(NORM_HAMMING)::knnMatch(k=2)
//                             (Features/Matcher.cpp:16,113)
//
// Build: see oracle/Makefile (-O2 -ffp-contract=off -fno-fast-math).
#include "oracle.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace orc {

// cvRound(float): round-half-to-even (SSE cvtss2si / lrintf in the default
// rounding mode).
int cv_round(float v) { return (int)lrintf(v); }
int cv_round_d(double v) { return (int)lrint(v); }
int cv_floor(float v)
{
    int i = (int)v;
    return i - (i > v);
}

// ---------------------------------------------------------------------------
// cv::resize, INTER_LINEAR, CV_8UC1.  OpenCV's 8-bit path: horizontal pass in
// int with 11-bit coefficients (INTER_RESIZE_COEF_BITS), vertical pass
//   dst = (((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2.
// ---------------------------------------------------------------------------
void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride)
{
    const int COEF = 2048;
    double inv_scale_x = (double)dw / sw, inv_scale_y = (double)dh / sh;
    double scale_x = 1. / inv_scale_x, scale_y = 1. / inv_scale_y;

    std::vector<int> xofs(dw), yofs(dh);
    std::vector<short> ialpha(dw * 2), ibeta(dh * 2);
    int xmax = dw;
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cv_floor(fx);
        fx -= sx;
        if (sx < 0) {
            fx = 0;
            sx = 0;
        }
        if (sx + 1 >= sw) {
            xmax = std::min(xmax, dx);
            if (sx >= sw - 1) {
                fx = 0;
                sx = sw - 1;
            }
        }
        xofs[dx] = sx;
        ialpha[dx * 2] = (short)cv_round((1.f - fx) * COEF);
        ialpha[dx * 2 + 1] = (short)cv_round(fx * COEF);
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cv_floor(fy);
        fy -= sy;
        yofs[dy] = sy;
        ibeta[dy * 2] = (short)cv_round((1.f - fy) * COEF);
        ibeta[dy * 2 + 1] = (short)cv_round(fy * COEF);
    }

    std::vector<int> row0(dw), row1(dw);
    auto hresize = [&](int sy, std::vector<int>& D) {
        sy = std::min(std::max(sy, 0), sh - 1);
        const uint8_t* S = src + (size_t)sy * sstride;
        for (int dx = 0; dx < dw; dx++) {
            int sx = xofs[dx];
            if (dx < xmax)
                D[dx] = S[sx] * ialpha[dx * 2] + S[sx + 1] * ialpha[dx * 2 + 1];
            else
                D[dx] = S[sx] * COEF;
        }
    };
    for (int dy = 0; dy < dh; dy++) {
        hresize(yofs[dy], row0);
        hresize(yofs[dy] + 1, row1);
        int b0 = ibeta[dy * 2], b1 = ibeta[dy * 2 + 1];
        uint8_t* D = dst + (size_t)dy * dstride;
        for (int x = 0; x < dw; x++) {
            int v = (((b0 * (row0[x] >> 4)) >> 16) + ((b1 * (row1[x] >> 4)) >> 16) + 2) >> 2;
            D[x] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

static inline int reflect101(int p, int len)
{
    if (len == 1)
        return 0;
    while (p < 0 || p >= len) {
        if (p < 0)
            p = -p;
        else
            p = 2 * (len - 1) - p;
    }
    return p;
}

// cv::copyMakeBorder(src, dst, b, b, b, b, BORDER_REFLECT_101 [| ISOLATED]).
// dst is (w+2b) x (h+2b); the interior is copied, the ring mirrors it.
void border_reflect101(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int b)
{
    for (int y = -b; y < h + b; y++) {
        const uint8_t* S = src + (size_t)reflect101(y, h) * sstride;
        uint8_t* D = dst + (size_t)(y + b) * dstride;
        for (int x = -b; x < w + b; x++)
            D[x + b] = S[reflect101(x, w)];
    }
}

// ---------------------------------------------------------------------------
// cv::FAST, TYPE_9_16.  Bresenham circle of radius 3, 16 pixels; a pixel is a
// corner at threshold t when >= 9 contiguous circle pixels are all darker than
// v - t or all brighter than v + t.  The response is OpenCV's cornerScore<16>:
// the largest t' for which the pixel is still a corner (returned as t'; OpenCV
// computes max(t, arc strength) - 1).  With nonmax, a corner survives when its
// score is strictly greater than its 8 neighbours' scores (non-corners score
// 0); rows/cols within 3 px of the ROI edge are never tested.  Output is
// row-major.
// ---------------------------------------------------------------------------
static const int kCircle[16][2] = {
    { 0, 3 }, { 1, 3 }, { 2, 2 }, { 3, 1 }, { 3, 0 }, { 3, -1 }, { 2, -2 }, { 1, -3 },
    { 0, -3 }, { -1, -3 }, { -2, -2 }, { -3, -1 }, { -3, 0 }, { -3, 1 }, { -2, 2 }, { -1, 3 }
};

static int corner_score16(const uint8_t* ptr, const int* pixel, int threshold)
{
    const int K = 8, N = K * 3 + 1;
    int v = ptr[0];
    short d[N];
    for (int k = 0; k < N; k++)
        d[k] = (short)(v - ptr[pixel[k]]);

    int a0 = threshold;
    for (int k = 0; k < 16; k += 2) {
        int a = std::min((int)d[k + 1], (int)d[k + 2]);
        a = std::min(a, (int)d[k + 3]);
        if (a <= a0)
            continue;
        for (int j = 4; j <= 8; j++)
            a = std::min(a, (int)d[k + j]);
        a0 = std::max(a0, std::min(a, (int)d[k]));
        a0 = std::max(a0, std::min(a, (int)d[k + 9]));
    }
    int b0 = -a0;
    for (int k = 0; k < 16; k += 2) {
        int b = std::max((int)d[k + 1], (int)d[k + 2]);
        for (int j = 3; j <= 5; j++)
            b = std::max(b, (int)d[k + j]);
        if (b >= b0)
            continue;
        for (int j = 6; j <= 8; j++)
            b = std::max(b, (int)d[k + j]);
        b0 = std::min(b0, std::max(b, (int)d[k]));
        b0 = std::min(b0, std::max(b, (int)d[k + 9]));
    }
    return -b0 - 1;
}

int fast9(const uint8_t* img, int w, int h, int stride, int threshold, int nonmax, int* out_xys, int cap)
{
    const int K = 8, N = 16 + K + 1;
    int pixel[25];
    for (int k = 0; k < 16; k++)
        pixel[k] = kCircle[k][0] + kCircle[k][1] * stride;
    for (int k = 16; k < 25; k++)
        pixel[k] = pixel[k - 16];
    threshold = std::min(std::max(threshold, 0), 255);
    int n = 0;
    if (w < 7 || h < 7)
        return 0;

    std::vector<uint8_t> score((size_t)w * h, 0);
    std::vector<uint8_t> iscorner((size_t)w * h, 0);
    for (int i = 3; i < h - 3; i++) {
        const uint8_t* row = img + (size_t)i * stride;
        for (int j = 3; j < w - 3; j++) {
            const uint8_t* ptr = row + j;
            int v = ptr[0];
            bool corner = false;
            {
                int vt = v - threshold, count = 0;
                for (int k = 0; k < N; k++) {
                    if (ptr[pixel[k]] < vt) {
                        if (++count > K) {
                            corner = true;
                            break;
                        }
                    } else
                        count = 0;
                }
            }
            if (!corner) {
                int vt = v + threshold, count = 0;
                for (int k = 0; k < N; k++) {
                    if (ptr[pixel[k]] > vt) {
                        if (++count > K) {
                            corner = true;
                            break;
                        }
                    } else
                        count = 0;
                }
            }
            if (corner) {
                iscorner[(size_t)i * w + j] = 1;
                score[(size_t)i * w + j] = (uint8_t)corner_score16(ptr, pixel, threshold);
            }
        }
    }
    for (int i = 3; i < h - 3; i++) {
        for (int j = 3; j < w - 3; j++) {
            size_t o = (size_t)i * w + j;
            if (!iscorner[o])
                continue;
            int s = score[o];
            bool keep = true;
            if (nonmax) {
                const uint8_t* p = &score[o];
                keep = s > p[-1] && s > p[1] && s > p[-w - 1] && s > p[-w] && s > p[-w + 1]
                    && s > p[w - 1] && s > p[w] && s > p[w + 1];
            }
            if (keep) {
                if (n < cap) {
                    out_xys[n * 3] = j;
                    out_xys[n * 3 + 1] = i;
                    out_xys[n * 3 + 2] = nonmax ? s : 0;
                }
                n++;
            }
        }
    }
    return n;
}

// ---------------------------------------------------------------------------
// cv::GaussianBlur(7x7, sigma 2), CV_8U: OpenCV 4.x fixed-point path.  The
// kernel is quantised to 8 fractional bits summing to 256:
// {18,34,48,56,48,34,18}; rows then columns, exact integers, one rounding:
// (sum + 2^15) >> 16.  (Pinned empirically against cv2 4.13.0.)
// ---------------------------------------------------------------------------
static const int kGauss7[7] = { 18, 34, 48, 56, 48, 34, 18 };

void gauss7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride)
{
    std::vector<int> tmp((size_t)w * h);
    for (int y = 0; y < h; y++) {
        const uint8_t* S = src + (size_t)y * sstride;
        for (int x = 0; x < w; x++) {
            int a = 0;
            for (int k = -3; k <= 3; k++)
                a += kGauss7[k + 3] * S[reflect101(x + k, w)];
            tmp[(size_t)y * w + x] = a;
        }
    }
    for (int y = 0; y < h; y++) {
        uint8_t* D = dst + (size_t)y * dstride;
        for (int x = 0; x < w; x++) {
            int a = 0;
            for (int k = -3; k <= 3; k++)
                a += kGauss7[k + 3] * tmp[(size_t)reflect101(y + k, h) * w + x];
            D[x] = (uint8_t)std::min((a + 32768) >> 16, 255);
        }
    }
}

// ---------------------------------------------------------------------------
// cv::fastAtan2(y, x): degrees in [0,360), 7th-order odd polynomial.
// ---------------------------------------------------------------------------
float fast_atan2(float y, float x)
{
    const float p1 = 0.9997878412794807f * (float)(180 / M_PI);
    const float p3 = -0.3258083974640975f * (float)(180 / M_PI);
    const float p5 = 0.1555786518463281f * (float)(180 / M_PI);
    const float p7 = -0.04432655554792128f * (float)(180 / M_PI);
    float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0)
        a = 180.f - a;
    if (y < 0)
        a = 360.f - a;
    return a;
}

// ---------------------------------------------------------------------------
// glibc >= 2.28 sinf/cosf (the ARM optimized-routines algorithm): double
// arithmetic, one multiply-subtract range reduction by pi/2, degree-7/8
// minimax polynomials, result rounded to float once.  Only the |x| < 120
// branches are restated (the descriptor angle lies in [0, 2*pi)).
// Evaluated without FMA contraction.
// ---------------------------------------------------------------------------
struct SinCosTab {
    double sign[4];
    double hpi_inv, hpi;
    double c0, c1, c2, c3, c4;
    double s1, s2, s3;
};
static const SinCosTab kTab[2] = {
    { { 1.0, -1.0, -1.0, 1.0 }, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0,
        0x1p0, -0x1.ffffffd0c621cp-2, 0x1.55553e1068f19p-5, -0x1.6c087e89a359dp-10, 0x1.99343027bf8c3p-16,
        -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13 },
    { { 1.0, -1.0, -1.0, 1.0 }, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0,
        -0x1p0, 0x1.ffffffd0c621cp-2, -0x1.55553e1068f19p-5, 0x1.6c087e89a359dp-10, -0x1.99343027bf8c3p-16,
        -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13 }
};

static inline uint32_t abstop12(float x)
{
    uint32_t u;
    memcpy(&u, &x, 4);
    return (u >> 20) & 0x7ff;
}

static inline float sinf_poly(double x, double x2, const SinCosTab* p, int n)
{
    if ((n & 1) == 0) {
        double x3 = x * x2;
        double s1 = p->s2 + x2 * p->s3;
        double x7 = x3 * x2;
        double s = x + x3 * p->s1;
        return (float)(s + x7 * s1);
    } else {
        double x4 = x2 * x2;
        double c2 = p->c3 + x2 * p->c4;
        double c1 = p->c0 + x2 * p->c1;
        double x6 = x4 * x2;
        double c = c1 + x4 * p->c2;
        return (float)(c + x6 * c2);
    }
}

static inline double reduce_fast(double x, const SinCosTab* p, int* np)
{
    double r = x * p->hpi_inv;
    int n = ((int32_t)r + 0x800000) >> 24;
    *np = n;
    return x - n * p->hpi;
}

float sinf_glibc(float y)
{
    double x = y;
    const SinCosTab* p = &kTab[0];
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        double s = x * x;
        if (abstop12(y) < abstop12(0x1p-12f))
            return y;
        return sinf_poly(x, s, p, 0);
    } else if (abstop12(y) < abstop12(120.0f)) {
        int n;
        x = reduce_fast(x, p, &n);
        double s = p->sign[n & 3];
        if (n & 2)
            p = &kTab[1];
        return sinf_poly(x * s, x * x, p, n);
    }
    return sinf(y); // outside the restated domain
}

float cosf_glibc(float y)
{
    double x = y;
    const SinCosTab* p = &kTab[0];
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        double x2 = x * x;
        if (abstop12(y) < abstop12(0x1p-12f))
            return 1.0f;
        return sinf_poly(x, x2, p, 1);
    } else if (abstop12(y) < abstop12(120.0f)) {
        int n;
        x = reduce_fast(x, p, &n);
        double s = p->sign[n & 3];
        if (n & 2)
            p = &kTab[1];
        return sinf_poly(x * s, x * x, p, n ^ 1);
    }
    return cosf(y);
}

// ---------------------------------------------------------------------------
// cv::BFMatcher(NORM_HAMMING, crossCheck=false)::knnMatch(query, train, 2).
// Neighbours ordered by (distance, trainIdx) — lower train index wins ties
// (checked against cv2 in tests/test_oracle_cv2.py).
// idx/dist hold k=2 entries per query; missing neighbours are -1.
// ---------------------------------------------------------------------------
int hamming256(const uint8_t* a, const uint8_t* b)
{
    uint64_t x[4], y[4];
    memcpy(x, a, 32);
    memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) + __builtin_popcountll(x[2] ^ y[2])
        + __builtin_popcountll(x[3] ^ y[3]);
}

void knn2_hamming(const uint8_t* q, int nq, const uint8_t* t, int nt, int* idx, int* dist)
{
    for (int i = 0; i < nq; i++) {
        int b0 = -1, b1 = -1, d0 = 1 << 30, d1 = 1 << 30;
        for (int j = 0; j < nt; j++) {
            int d = hamming256(q + (size_t)i * 32, t + (size_t)j * 32);
            if (d < d0) {
                d1 = d0;
                b1 = b0;
                d0 = d;
                b0 = j;
            } else if (d < d1) {
                d1 = d;
                b1 = j;
            }
        }
        idx[i * 2] = b0;
        idx[i * 2 + 1] = b1;
        dist[i * 2] = b0 < 0 ? -1 : d0;
        dist[i * 2 + 1] = b1 < 0 ? -1 : d1;
    }
}

} // namespace orc

extern "C" {
void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride)
{
    orc::resize_linear_u8(src, sw, sh, sstride, dst, dw, dh, dstride);
}
void orc_border_reflect101(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int b)
{
    orc::border_reflect101(src, w, h, sstride, dst, dstride, b);
}
int orc_fast9(const uint8_t* img, int w, int h, int stride, int th, int nonmax, int* out, int cap)
{
    return orc::fast9(img, w, h, stride, th, nonmax, out, cap);
}
void orc_gauss7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride)
{
    orc::gauss7_u8(src, w, h, sstride, dst, dstride);
}
float orc_fast_atan2(float y, float x) { return orc::fast_atan2(y, x); }
void orc_fast_atan2_v(const float* y, const float* x, float* out, int n)
{
    for (int i = 0; i < n; i++)
        out[i] = orc::fast_atan2(y[i], x[i]);
}
int orc_cv_round(float v) { return orc::cv_round(v); }
float orc_sinf(float v) { return orc::sinf_glibc(v); }
float orc_cosf(float v) { return orc::cosf_glibc(v); }
// Count of floats in [lo,hi] (stepping `stride` representable values) where
// the restatement differs from the C library's sinf/cosf on this machine.
long orc_sincos_mismatches(float lo, float hi, int stride, long* total)
{
    uint32_t a, b;
    memcpy(&a, &lo, 4);
    memcpy(&b, &hi, 4);
    long bad = 0, n = 0;
    for (uint64_t u = a; u <= b; u += (uint64_t)stride) {
        uint32_t uu = (uint32_t)u;
        float f;
        memcpy(&f, &uu, 4);
        volatile float vf = f;
        float s0 = sinf(vf), c0 = cosf(vf);
        if (s0 != orc::sinf_glibc(f) || c0 != orc::cosf_glibc(f))
            bad++;
        n++;
    }
    *total = n;
    return bad;
}
void orc_knn2_hamming(const uint8_t* q, int nq, const uint8_t* t, int nt, int* idx, int* dist)
{
    orc::knn2_hamming(q, nq, t, nt, idx, dist);
}
}
