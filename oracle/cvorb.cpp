// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// Row N3 of SURVEY.md 8(f): the `Extractor(ORB, ORB)` branch of the reference, i.e.
//   mpDetector   = cv::ORB::create(nfeatures, scaleFactor, nlevels)   (Features/Extractor.cpp:163-166)
//   mpDescriptor = cv::ORB::create()                                  (Features/Extractor.cpp:216-218)
//   detect -> retainBest(nfeatures) -> compute                        (Features/Extractor.cpp:55-59)
// The arithmetic lives in OpenCV (features2d/src/orb.cpp, imgproc/src/resize.cpp, features2d/src/keypoint.cpp),
// which is not under /root/reference; this file restates the published algorithm of cv::ORB (HARRIS_SCORE,
// k = 0.04, 7x7 block, FAST threshold 20, edge threshold 31, patch 31, WTA_K 2, firstLevel 0):
//   pyramid   : chain of resize(INTER_LINEAR_EXACT) (8.8 fixed point, coefficients from double arithmetic)
//   detection : whole-level FAST-9 with 3x3 non-max suppression, border filter, retainBest(2 N_l) on the FAST score
//               (std::nth_element + std::partition: the survivors' ORDER is libstdc++'s introselect permutation),
//               Harris response, retainBest(N_l), intensity-centroid angle, pt *= scale
//   describe  : GaussianBlur(7x7, sigma 2) of every level (the float sepFilter2D path, see gauss7_float_u8), steered
//               BRIEF at cvRound(pt / scale)
// Parity: pinned bit-for-bit against cv2 4.13.0 (cv2.ORB_create(...).detect + cv2.ORB_create().compute) by
// tests/test_cvorb.py live and through tests/golden/cvorb_cv2.npz.
#include "brief_pattern.h"
#include "oracle.h"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace orc {

// cv::resize(..., INTER_LINEAR_EXACT) for CV_8UC1 (resize.cpp: resize_bitExact<uchar, interpolationLinear<uchar>>)
namespace {
struct ExactTab {
    std::vector<int> ofs;
    std::vector<uint16_t> c0, c1; // ufixedpoint16, 8 fractional bits
    int mn, mx;
};
ExactTab exact_coeffs(int srcsize, int dstsize)
{
    ExactTab t;
    t.ofs.assign(dstsize, 0);
    t.c0.assign(dstsize, 0);
    t.c1.assign(dstsize, 0);
    const double inv_scale = (double)dstsize / srcsize;
    const double scale = 1.0 / inv_scale; // softdouble::one() / softdouble(inv_scale)
    int minofst = 0, maxofst = dstsize;
    for (int val = 0; val < dstsize; val++) {
        const double fval = scale * ((double)val + 0.5) - 0.5;
        const int ival = (int)std::floor(fval);
        if (ival >= 0 && srcsize > 1) {
            if (ival < srcsize - 1) {
                t.ofs[val] = ival;
                const double frac = fval - (double)ival;
                t.c1[val] = frac < 0 ? 0 : (uint16_t)cv_round_d(frac * 256.0);
                t.c0[val] = (uint16_t)(256 > t.c1[val] ? 256 - t.c1[val] : 0);
            } else {
                t.ofs[val] = srcsize - 1;
                maxofst = std::min(maxofst, val);
            }
        } else
            minofst = std::max(minofst, val + 1);
    }
    t.mn = minofst, t.mx = maxofst;
    return t;
}
} // namespace

void resize_linear_exact_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride)
{
    const ExactTab X = exact_coeffs(sw, dw), Y = exact_coeffs(sh, dh);
    auto hline = [&](const uint8_t* row, std::vector<uint16_t>& out) {
        int i = 0;
        for (; i < X.mn; i++)
            out[i] = (uint16_t)(row[0] << 8);
        for (; i < X.mx; i++) {
            const uint8_t* px = row + X.ofs[i];
            const unsigned a = std::min(65535u, (unsigned)X.c0[i] * px[0]), b = std::min(65535u, (unsigned)X.c1[i] * px[1]);
            out[i] = (uint16_t)std::min(65535u, a + b);
        }
        const uint8_t last = row[X.ofs[dw - 1]];
        for (; i < dw; i++)
            out[i] = (uint16_t)(last << 8);
    };
    std::vector<uint16_t> r0(dw), r1(dw);
    for (int dy = 0; dy < dh; dy++) {
        uint8_t* d = dst + (size_t)dy * dstride;
        if (dy < Y.mn || dy >= Y.mx) {
            hline(src + (size_t)(dy < Y.mn ? 0 : sh - 1) * sstride, r0);
            for (int i = 0; i < dw; i++)
                d[i] = (uint8_t)std::min(255u, ((unsigned)r0[i] + 128u) >> 8);
            continue;
        }
        const int iy = Y.ofs[dy];
        hline(src + (size_t)iy * sstride, r0);
        hline(src + (size_t)std::min(iy + 1, sh - 1) * sstride, r1);
        for (int i = 0; i < dw; i++) {
            const uint64_t v = (uint64_t)r0[i] * Y.c0[dy] + (uint64_t)r1[i] * Y.c1[dy];
            const uint32_t s = (uint32_t)std::min<uint64_t>(v, 0xffffffffull);
            d[i] = (uint8_t)std::min<uint32_t>(255u, (uint32_t)(((uint64_t)s + 32768u) >> 16));
        }
    }
}

// cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) as cv::ORB reaches it: the level is a SUBMATRIX of the pyramid
// buffer, so GaussianBlur skips its bit-exact fixed-point path (smooth.dispatch.cpp: "isSubmatrix()" test) and runs
// sepFilter2D with the CV_32F kernel getGaussianKernel(7, 2): row pass uchar -> float (RowFilter, taps left to
// right), column pass float -> uchar (SymmColumnFilter, centre tap then symmetric pairs, saturate_cast = cvRound).
// cv2 4.13 runs the AVX2 dispatch of filter.simd.hpp here: the vector body and the column pass fuse every
// `s += k * v` into one FMA, the row pass's 4-way unrolled scalar remainder does not (see the row loop below).
// Pinned against cv2.sepFilter2D on this image's CPU (AVX2, no AVX-512).
static void gaussian_kernel7_sigma2(float k[7])
{
    // getGaussianKernel(7, 2.0, CV_32F): exp() in double, normalised in double, narrowed to float
    double t[7], sum = 0;
    const double scale2X = -0.5 / (2.0 * 2.0);
    for (int i = 0; i < 7; i++) {
        const double x = i - 3.0;
        t[i] = std::exp(scale2X * x * x);
        sum += t[i];
    }
    sum = 1. / sum;
    for (int i = 0; i < 7; i++)
        k[i] = (float)(t[i] * sum);
}
void gauss7_float_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride)
{
    float k[7];
    gaussian_kernel7_sigma2(k);
    std::vector<float> rows((size_t)(h + 6) * w);
    auto r101 = [](int p, int len) {
        if (p < 0)
            p = -p;
        if (p >= len)
            p = 2 * (len - 1) - p;
        return p;
    };
    for (int y = -3; y < h + 3; y++) {
        const uint8_t* s = src + (size_t)r101(y, h) * sstride;
        float* d = &rows[(size_t)(y + 3) * w];
        // RowFilter<uchar, float, RowVec_8u32f>: 32-pixel vector blocks (v_muladd = FMA), then the scalar loop unrolled
        // by 4 (compiled WITHOUT contraction: separate multiply and add), then the last w % 4 pixels (contracted: FMA)
        const int vec_end = w / 32 * 32, unr_end = vec_end + (w - vec_end) / 4 * 4;
        for (int x = 0; x < w; x++) {
            float acc = k[0] * (float)s[r101(x - 3, w)];
            const bool fused = x < vec_end || x >= unr_end;
            for (int j = 1; j < 7; j++) {
                const float v = (float)s[r101(x - 3 + j, w)];
                acc = fused ? std::fmaf(k[j], v, acc) : acc + k[j] * v;
            }
            d[x] = acc;
        }
    }
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            const float* c = &rows[(size_t)(y + 3) * w + x];
            float acc = k[3] * c[0];
            for (int j = 1; j <= 3; j++)
                acc = std::fmaf(k[3 + j], c[(size_t)j * w] + c[-(ptrdiff_t)j * w], acc);
            int v = cv_round(acc);
            dst[(size_t)y * dstride + x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
}

// cv::KeyPointsFilter::retainBest (features2d/src/keypoint.cpp): the real std::nth_element / std::partition
static void retain_best(std::vector<KeyPoint>& kps, int n)
{
    if (n >= 0 && kps.size() > (size_t)n) {
        if (n == 0) {
            kps.clear();
            return;
        }
        std::nth_element(kps.begin(), kps.begin() + n - 1, kps.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
        const float ambiguous = kps[n - 1].response;
        auto new_end = std::partition(kps.begin() + n, kps.end(), [ambiguous](const KeyPoint& k) { return k.response >= ambiguous; });
        kps.resize(new_end - kps.begin());
    }
}

void CvOrb::extract(const uint8_t* img, int w, int h, int stride, std::vector<KeyPoint>& out, std::vector<uint8_t>& desc)
{
    out.clear();
    desc.clear();
    levels.clear();
    if (!img || w <= 0 || h <= 0)
        return;
    const int B = border;
    // --- pyramid (orb.cpp detectAndCompute: layerInfo / resize / copyMakeBorder) ---
    levels.resize(nlevels);
    for (int l = 0; l < nlevels; l++) {
        CvOrbLevel& L = levels[l];
        L.scale = (float)std::pow(scaleFactor, (double)l); // getScale(level, firstLevel = 0, scaleFactor)
        const float inv = 1.0f / L.scale;
        L.w = cv_round((float)w * inv), L.h = cv_round((float)h * inv);
        L.stride = L.w + 2 * B;
        L.padded.assign((size_t)L.stride * (L.h + 2 * B), 0);
        std::vector<uint8_t> cur((size_t)L.w * L.h);
        if (l == 0) {
            for (int y = 0; y < h; y++)
                memcpy(&cur[(size_t)y * w], img + (size_t)y * stride, w);
        } else {
            const CvOrbLevel& P = levels[l - 1];
            resize_linear_exact_u8(P.image(), P.w, P.h, P.stride, cur.data(), L.w, L.h, L.w);
        }
        border_reflect101(cur.data(), L.w, L.h, L.w, L.padded.data(), L.stride, B);
    }
    // --- computeKeyPoints ---
    std::vector<int> nfeaturesPerLevel(nlevels);
    {
        const float factor = (float)(1.0 / scaleFactor);
        float ndesired = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
        int sum = 0;
        for (int l = 0; l < nlevels - 1; l++) {
            nfeaturesPerLevel[l] = cv_round(ndesired);
            sum += nfeaturesPerLevel[l];
            ndesired *= factor;
        }
        nfeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sum, 0);
    }
    quota = nfeaturesPerLevel;
    const int half = 15;
    std::vector<int> umax(half + 2);
    {
        int v, v0, vmax = cv_floor(half * std::sqrt(2.f) / 2 + 1);
        int vmin = (int)std::ceil(half * std::sqrt(2.f) / 2);
        for (v = 0; v <= vmax; ++v)
            umax[v] = cv_round_d(std::sqrt((double)half * half - v * v));
        for (v = half, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1])
                ++v0;
            umax[v] = v0;
            ++v0;
        }
    }
    for (int l = 0; l < nlevels; l++) {
        CvOrbLevel& L = levels[l];
        const int N = nfeaturesPerLevel[l];
        std::vector<int> xys((size_t)L.w * L.h / 4 * 3 + 3);
        int n = fast9(L.image(), L.w, L.h, L.stride, fastThreshold, 1, xys.data(), (int)xys.size() / 3);
        std::vector<KeyPoint> kps;
        for (int i = 0; i < n; i++) {
            // FAST keypoint: size 7, angle -1, response = score; runByImageBorder(edgeThreshold)
            const int x = xys[3 * i], y = xys[3 * i + 1];
            if (x >= edgeThreshold && x < L.w - edgeThreshold && y >= edgeThreshold && y < L.h - edgeThreshold)
                kps.push_back(KeyPoint{ (float)x, (float)y, 7.f, -1.f, (float)xys[3 * i + 2], 0, -1 });
        }
        L.n_fast = (int)kps.size();
        retain_best(kps, 2 * N);
        for (auto& k : kps) {
            k.octave = l;
            k.size = 31 * L.scale;
        }
        // HarrisResponses(blockSize 7, k 0.04)
        const int step = L.stride, r = 3;
        const float scale = 1.f / ((1 << 2) * 7 * 255.f);
        const float scale_sq_sq = scale * scale * scale * scale;
        for (auto& k : kps) {
            const int x0 = cv_round(k.x), y0 = cv_round(k.y);
            const uint8_t* ptr0 = L.image() + (y0 - r) * step + x0 - r;
            int a = 0, b = 0, c = 0;
            for (int i = 0; i < 7; i++)
                for (int j = 0; j < 7; j++) {
                    const uint8_t* ptr = ptr0 + i * step + j;
                    int Ix = (ptr[1] - ptr[-1]) * 2 + (ptr[-step + 1] - ptr[-step - 1]) + (ptr[step + 1] - ptr[step - 1]);
                    int Iy = (ptr[step] - ptr[-step]) * 2 + (ptr[step - 1] - ptr[-step - 1]) + (ptr[step + 1] - ptr[-step + 1]);
                    a += Ix * Ix;
                    b += Iy * Iy;
                    c += Ix * Iy;
                }
            k.response = ((float)a * b - (float)c * c - 0.04f * ((float)a + b) * ((float)a + b)) * scale_sq_sq;
        }
        L.harris_in = kps;
        retain_best(kps, N);
        // ICAngles
        for (auto& k : kps) {
            const uint8_t* center = L.image() + cv_round(k.y) * step + cv_round(k.x);
            int m_01 = 0, m_10 = 0;
            for (int u = -half; u <= half; ++u)
                m_10 += u * center[u];
            for (int v = 1; v <= half; ++v) {
                int v_sum = 0, d = umax[v];
                for (int u = -d; u <= d; ++u) {
                    int val_plus = center[u + v * step], val_minus = center[u - v * step];
                    v_sum += (val_plus - val_minus);
                    m_10 += u * (val_plus + val_minus);
                }
                m_01 += v * v_sum;
            }
            k.angle = fast_atan2((float)m_01, (float)m_10);
        }
        for (auto& k : kps) {
            k.x *= L.scale;
            k.y *= L.scale;
            out.push_back(k);
        }
    }
    // Extractor.cpp:56-57: retainBest only if more than nfeatures came back (cv::ORB never returns more)
    if (out.size() > (size_t)nfeatures)
        retain_best(out, nfeatures);
    // --- mpDescriptor->compute(): a default cv::ORB (scaleFactor 1.2f) with useProvidedKeypoints ---
    // runByImageBorder(keypoints, image.size(), 31) at level-0 coordinates; keypoints stay sorted by level
    {
        std::vector<KeyPoint> kept;
        for (auto& k : out)
            if (k.x >= (float)edgeThreshold && k.x < (float)(w - edgeThreshold) && k.y >= (float)edgeThreshold && k.y < (float)(h - edgeThreshold))
                kept.push_back(k);
        out.swap(kept);
    }
    desc.assign(out.size() * 32, 0);
    for (int l = 0; l < nlevels; l++) {
        CvOrbLevel& L = levels[l];
        L.blurred.assign((size_t)L.w * L.h, 0);
        gauss7_float_u8(L.image(), L.w, L.h, L.stride, L.blurred.data(), L.w);
    }
    for (size_t j = 0; j < out.size(); j++) {
        const KeyPoint& kpt = out[j];
        const CvOrbLevel& L = levels[kpt.octave];
        // the descriptor object is cv::ORB::create(): its own scale factor 1.2f, same pyramid iff scaleFactor == 1.2f
        const float dscale = (float)std::pow((double)1.2f, (double)kpt.octave);
        const float scale = 1.f / dscale;
        float angle = kpt.angle;
        angle *= (float)(M_PI / 180.f);
        const float a = cosf_glibc(angle), b = sinf_glibc(angle);
        const int step = L.w;
        const uint8_t* center = L.blurred.data() + cv_round(kpt.y * scale) * step + cv_round(kpt.x * scale);
        const signed char* pat = kBriefPattern;
        uint8_t* d = desc.data() + j * 32;
        for (int i = 0; i < 32; ++i, pat += 32) {
            int val = 0;
            for (int k = 0; k < 8; k++) {
                const signed char* p0 = pat + k * 4;
                int t0 = center[cv_round(p0[0] * b + p0[1] * a) * step + cv_round(p0[0] * a - p0[1] * b)];
                int t1 = center[cv_round(p0[2] * b + p0[3] * a) * step + cv_round(p0[2] * a - p0[3] * b)];
                val |= (t0 < t1) << k;
            }
            d[i] = (uint8_t)val;
        }
    }
}

} // namespace orc
