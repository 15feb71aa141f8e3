// Stand-ins for the OpenCV / Eigen types that cross the hot-path class interfaces, used ONLY when
// the shims are built without RGBDFE_WITH_OPENCV (this image has neither OpenCV C++ headers nor
// Eigen).  Same names, same memory layout for the PODs the C ABI passes through
// (cv::KeyPoint 28 B, cv::DMatch 16 B, cv::Point3f 12 B), and just enough of cv::Mat /
// Eigen::Matrix4f for the shims and their tests.  With RGBDFE_WITH_OPENCV the real headers are
// included instead and this file contributes nothing.
#ifndef RGBDFE_CVCOMPAT_H
#define RGBDFE_CVCOMPAT_H

#ifdef RGBDFE_WITH_OPENCV
#include <Eigen/Core>
#include <opencv2/core.hpp>
#include <opencv2/features2d.hpp>
#else

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <utility>
#include <memory>
#include <vector>

typedef unsigned int uint;
typedef unsigned char uchar;

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_WRAP
#define CV_OUT

namespace cv {

enum { NORM_L2 = 4, NORM_HAMMING = 6 };

struct Point2f {
    float x, y;
    Point2f() : x(0), y(0) {}
    Point2f(float x_, float y_) : x(x_), y(y_) {}
};
struct Point3f {
    float x, y, z;
    Point3f() : x(0), y(0), z(0) {}
    Point3f(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};
struct Point {
    int x, y;
};

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};

struct DMatch {
    int queryIdx, trainIdx, imgIdx;
    float distance;
    DMatch() : queryIdx(-1), trainIdx(-1), imgIdx(-1), distance(3.402823466e+38f) {}
    DMatch(int q, int t, float d) : queryIdx(q), trainIdx(t), imgIdx(-1), distance(d) {}
    bool operator<(const DMatch& m) const { return distance < m.distance; }
};

// Dense 2-D matrix of CV_8U or CV_32F with shared ownership (what the hot path needs of cv::Mat).
class Mat {
public:
    int rows = 0, cols = 0;
    size_t step = 0; // bytes per row
    uchar* data = nullptr;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    // wraps caller memory (no ownership), like cv::Mat(rows, cols, type, ptr, step)
    Mat(int r, int c, int type, void* ptr, size_t step_ = 0) : rows(r), cols(c), data((uchar*)ptr), type_(type)
    {
        step = step_ ? step_ : (size_t)c * elemSize();
    }
    void create(int r, int c, int type)
    {
        if (r == rows && c == cols && type == type_ && data && own_)
            return;
        type_ = type;
        rows = r;
        cols = c;
        step = (size_t)c * elemSize();
        own_ = std::make_shared<std::vector<uchar>>((size_t)r * step);
        data = own_->data();
    }
    void release()
    {
        own_.reset();
        data = nullptr;
        rows = cols = 0;
    }
    int type() const { return type_; }
    size_t elemSize() const { return type_ == CV_32F ? 4 : 1; }
    bool empty() const { return rows == 0 || cols == 0 || !data; }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }
    template <class T>
    T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
    template <class T>
    const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
    template <class T>
    T& at(int r, int c) { return ptr<T>(r)[c]; }
    template <class T>
    const T& at(int r, int c) const { return ptr<T>(r)[c]; }
    Mat row(int r) const
    {
        Mat m = *this;
        m.rows = 1;
        m.data = data + (size_t)r * step;
        return m;
    }
    Mat clone() const
    {
        Mat m(rows, cols, type_);
        for (int r = 0; r < rows; r++)
            memcpy(m.data + (size_t)r * m.step, data + (size_t)r * step, (size_t)cols * elemSize());
        return m;
    }
    static Mat eye(int r, int c, int type)
    {
        Mat m(r, c, type);
        memset(m.data, 0, (size_t)r * m.step);
        for (int i = 0; i < r && i < c; i++)
            m.at<float>(i, i) = 1.0f;
        return m;
    }
    // CV_32F matrix product (poses are 4x4)
    Mat operator*(const Mat& b) const
    {
        Mat m(rows, b.cols, CV_32F);
        for (int i = 0; i < rows; i++)
            for (int j = 0; j < b.cols; j++) {
                float s = 0.0f;
                for (int k = 0; k < cols; k++)
                    s += at<float>(i, k) * b.at<float>(k, j);
                m.at<float>(i, j) = s;
            }
        return m;
    }
    // inverse of a small square CV_32F matrix (Gaussian elimination with partial pivoting, computed in
    // double); stands in for cv::Mat::inv(DECOMP_LU) in builds without OpenCV
    Mat inv() const
    {
        const int n = rows;
        std::vector<double> a((size_t)n * 2 * n, 0.0);
        for (int i = 0; i < n; i++) {
            for (int j = 0; j < n; j++)
                a[(size_t)i * 2 * n + j] = at<float>(i, j);
            a[(size_t)i * 2 * n + n + i] = 1.0;
        }
        for (int c = 0; c < n; c++) {
            int p = c;
            for (int r = c + 1; r < n; r++)
                if (std::abs(a[(size_t)r * 2 * n + c]) > std::abs(a[(size_t)p * 2 * n + c]))
                    p = r;
            if (a[(size_t)p * 2 * n + c] == 0.0)
                return Mat::eye(n, n, CV_32F);
            for (int j = 0; j < 2 * n; j++)
                std::swap(a[(size_t)c * 2 * n + j], a[(size_t)p * 2 * n + j]);
            const double d = a[(size_t)c * 2 * n + c];
            for (int j = 0; j < 2 * n; j++)
                a[(size_t)c * 2 * n + j] /= d;
            for (int r = 0; r < n; r++)
                if (r != c) {
                    const double f = a[(size_t)r * 2 * n + c];
                    for (int j = 0; j < 2 * n; j++)
                        a[(size_t)r * 2 * n + j] -= f * a[(size_t)c * 2 * n + j];
                }
        }
        Mat m(n, n, CV_32F);
        for (int i = 0; i < n; i++)
            for (int j = 0; j < n; j++)
                m.at<float>(i, j) = (float)a[(size_t)i * 2 * n + n + j];
        return m;
    }

private:
    int type_ = CV_8U;
    std::shared_ptr<std::vector<uchar>> own_;
};

typedef const Mat& InputArray;
typedef Mat& OutputArray;

template <class T>
using Ptr = std::shared_ptr<T>;

// the part of cv::Feature2D the reference's extractor hierarchy relies on
class Feature2D {
public:
    virtual ~Feature2D() {}
    virtual void detectAndCompute(InputArray, InputArray, std::vector<KeyPoint>&, OutputArray, bool = false) {}
    virtual int defaultNorm() const { return NORM_L2; }
};
typedef Feature2D FeatureDetector;
typedef Feature2D DescriptorExtractor;

} // namespace cv

namespace Eigen {
// Column-major 4x4 float, the storage order of Eigen::Matrix4f.
struct Matrix4f {
    float m[16];
    float& operator()(int r, int c) { return m[c * 4 + r]; }
    const float& operator()(int r, int c) const { return m[c * 4 + r]; }
    float* data() { return m; }
    const float* data() const { return m; }
    void setIdentity()
    {
        for (int i = 0; i < 16; i++)
            m[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    }
    static Matrix4f Identity()
    {
        Matrix4f r;
        r.setIdentity();
        return r;
    }
    // Eigen's isIdentity(prec): diagonal isApprox 1, off-diagonal isMuchSmallerThan 1
    bool isIdentity(float prec = 1e-5f) const
    {
        for (int c = 0; c < 4; c++)
            for (int r = 0; r < 4; r++) {
                const float x = (*this)(r, c);
                if (r == c) {
                    const float ax = std::fabs(x);
                    if (!(std::fabs(x - 1.0f) <= prec * (ax < 1.0f ? ax : 1.0f)))
                        return false;
                } else if (!(std::fabs(x) <= prec))
                    return false;
            }
        return true;
    }
};
} // namespace Eigen

#define EIGEN_MAKE_ALIGNED_OPERATOR_NEW

#endif // RGBDFE_WITH_OPENCV
#endif // RGBDFE_CVCOMPAT_H
