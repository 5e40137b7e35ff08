// oracle/cvshim/opencv2/cvshim.hpp — TEST INFRASTRUCTURE ONLY (see ../../orb_oracle.h).
//
// A minimal stand-in for the part of the OpenCV C++ API that /root/reference/src/ORBextractor.cc, src/Frame.cc and the
// vendored Thirdparty/DBoW2 use, so
// that the reference's OWN translation units can be compiled from where they lie (oracle/Makefile target _ref) although
// this image has no OpenCV C++ headers or libraries. Nothing here is OpenCV code: the containers (Mat, KeyPoint, Point_, Size, Rect,
// InputArray/OutputArray) are written from the documented interface, and the five primitives the reference calls
// (resize, copyMakeBorder, FAST, GaussianBlur, fastAtan2) forward to the oracle's restatements in oracle/prims.cpp, which
// tests/test_oracle_kat.py pins bit for bit against the real OpenCV kernels through cv2 4.13.
// Matrices are single-channel CV_8U or CV_32F. The float algebra (products, transposes, ...) exists so that ALL of
// Frame.cc compiles; what actually runs from that file is ComputeStereoMatches and ComputeStereoFromRGBD (convertTo,
// scalar * ones, difference, L1 norm on 11x11 patches of small integers: exact in float and double alike).
// Anything the reference does not call is absent.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <limits>
#include <numeric>
#include <sstream>
#include <string>
#include <vector>

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_Assert(expr) do { if (!(expr)) { std::fprintf(stderr, "CV_Assert failed: %s\n", #expr); std::abort(); } } while (0)

typedef unsigned char uchar;

// cvRound: round half to even (the SSE2 cvtss2si / cvtsd2si path of OpenCV on x86-64; default rounding mode)
static inline int cvRound(double v) { return (int)std::lrint(v); }
static inline int cvRound(float v) { return (int)std::lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return (int)std::floor(v); }
static inline int cvCeil(double v) { return (int)std::ceil(v); }

namespace cv {

template <typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    Point_& operator*=(float s)  // cv::Point_<float> *= float: each coordinate multiplied in T (saturate_cast is the identity for float)
    {
        x = (T)(x * s);
        y = (T)(y * s);
        return *this;
    }
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};

// Same 28-byte layout as the real cv::KeyPoint.
struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};

struct MatStep {
    size_t v;
    MatStep() : v(0) {}
    operator size_t() const { return v; }
};

// What Mat::zeros returns (cv::MatExpr): assigning it to a Mat of the same shape fills that Mat's EXISTING buffer with
// zeros instead of rebinding the header — computeDescriptors (src/ORBextractor.cc:1035) relies on this to write through
// a rowRange view of the output matrix.
struct MatZeros {
    int rows, cols, type;
};

// 8-bit single-channel matrix header over a reference-counted buffer. The buffer and its counter come from malloc/free,
// never from operator new: ref_wrap.cpp replaces operator new by a per-call arena, and pyramid levels outlive a call.
class Mat {
public:
    int rows, cols;
    uchar* data;
    MatStep step;

    Mat() : rows(0), cols(0), data(nullptr), buf(nullptr), type_(CV_8U) {}
    Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr), buf(nullptr), type_(CV_8U) { create(r, c, type); }
    Mat(Size sz, int type) : rows(0), cols(0), data(nullptr), buf(nullptr), type_(CV_8U) { create(sz.height, sz.width, type); }
    Mat(int r, int c, int type, void* user, size_t stp) : rows(r), cols(c), data((uchar*)user), buf(nullptr), type_(type)
    {
        assert(type == CV_8U || type == CV_32F);
        step.v = stp;
    }
    Mat(const Mat& o) : rows(o.rows), cols(o.cols), data(o.data), step(o.step), buf(o.buf), type_(o.type_)
    {
        if (buf) ++*buf;
    }
    Mat& operator=(const Mat& o)
    {
        if (o.buf) ++*o.buf;
        drop();
        rows = o.rows; cols = o.cols; data = o.data; step = o.step; buf = o.buf; type_ = o.type_;
        return *this;
    }
    ~Mat() { drop(); }

    void create(int r, int c, int type)
    {
        assert(type == CV_8U || type == CV_32F);
        if (data && r == rows && c == cols && type == type_) return;  // cv::Mat::create keeps a buffer of the right shape
        drop();
        rows = r;
        cols = c;
        type_ = type;
        step.v = (size_t)c * elemSize();
        buf = (int*)std::malloc(16 + (size_t)r * step.v);
        *buf = 1;
        data = (uchar*)buf + 16;
    }
    size_t elemSize() const { return type_ == CV_32F ? 4 : 1; }
    void create(Size sz, int type) { create(sz.height, sz.width, type); }
    void release()
    {
        drop();
        data = nullptr;
        rows = cols = 0;
    }
    static MatZeros zeros(int r, int c, int type) { return MatZeros{r, c, type}; }
    Mat(const MatZeros& z) : rows(0), cols(0), data(nullptr), buf(nullptr), type_(CV_8U) { *this = z; }
    Mat& operator=(const MatZeros& z)
    {
        create(z.rows, z.cols, z.type);  // keeps a buffer (or view) of the right shape
        for (int y = 0; y < rows; ++y) std::memset(data + (size_t)y * step.v, 0, (size_t)cols * elemSize());
        return *this;
    }

    int type() const { return type_; }
    size_t step1() const { return step.v / elemSize(); }  // step in elements
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    Size size() const { return Size(cols, rows); }

    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    // single-index form: element i of a row or column vector
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step.v; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step.v; }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step.v); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step.v); }

    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
    Mat operator()(const Rect& r) const
    {
        assert(r.x >= 0 && r.y >= 0 && r.x + r.width <= cols && r.y + r.height <= rows);
        Mat m(*this);  // shares the buffer
        m.rows = r.height;
        m.cols = r.width;
        m.data = data + (size_t)r.y * step.v + (size_t)r.x * elemSize();
        return m;
    }
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat col(int x) const { return colRange(x, x + 1); }
    Mat clone() const
    {
        Mat m;
        if (!data) return m;
        m.create(rows, cols, type_);
        for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step.v, data + (size_t)y * step.v, (size_t)cols * elemSize());
        return m;
    }
    inline void copyTo(const class _OutputArray& dst) const;  // cv::Mat::copyTo(OutputArray), defined below the proxies
    void copyTo(Mat& dst) const
    {
        dst.create(rows, cols, type_);
        for (int y = 0; y < rows; ++y) std::memcpy(dst.data + (size_t)y * dst.step.v, data + (size_t)y * step.v, (size_t)cols * elemSize());
    }
    // Mat::convertTo, CV_8U -> CV_32F (src/Frame.cc:623, :641; dst may be *this)
    void convertTo(Mat& dst, int rtype) const
    {
        assert(type_ == CV_8U && rtype == CV_32F);
        Mat out(rows, cols, CV_32F);
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) out.at<float>(y, x) = (float)at<uchar>(y, x);
        dst = out;
    }
    static Mat ones(int r, int c, int type)
    {
        assert(type == CV_32F);
        Mat m(r, c, type);
        for (int y = 0; y < r; ++y)
            for (int x = 0; x < c; ++x) m.at<float>(y, x) = 1.f;
        return m;
    }
    inline struct MatTExpr t() const;  // lazy, like cv::MatExpr: -A.t() * b becomes ONE gemm (see below)
    double dot(const Mat& o) const
    {
        double s = 0;
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) s += (double)at<float>(y, x) * o.at<float>(y, x);
        return s;
    }
    Mat reshape(int /*cn*/) const { return *this; }  // only around undistortPoints, which this build never reaches

private:
    int* buf;  // counter at the head of the owned allocation; nullptr for user data
    int type_;
    void drop()
    {
        if (buf && --*buf == 0) std::free(buf);
        buf = nullptr;
    }
};

// cv::InputArray / cv::OutputArray are const references to proxy objects that a Mat converts to.
class _InputArray {
public:
    _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
    bool empty() const { return m_->empty(); }
    Mat getMat() const { return *m_; }

protected:
    Mat* m_;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray(Mat& m) : _InputArray(m) {}
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void release() const { m_->release(); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
inline void Mat::copyTo(const _OutputArray& dst) const
{
    dst.create(rows, cols, type_);
    Mat d = dst.getMat();
    for (int y = 0; y < rows; ++y) std::memcpy(d.data + (size_t)y * d.step.v, data + (size_t)y * step.v, (size_t)cols * elemSize());
}

// ---- CV_32F algebra used by src/Frame.cc ----
inline Mat operator-(const Mat& a, const Mat& b)
{
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = a.at<float>(y, x) - b.at<float>(y, x);
    return m;
}
inline Mat operator+(const Mat& a, const Mat& b)
{
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = a.at<float>(y, x) + b.at<float>(y, x);
    return m;
}
inline Mat operator*(float s, const Mat& a)
{
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = s * a.at<float>(y, x);
    return m;
}
inline Mat operator-(const Mat& a) { return -1.f * a; }
inline Mat operator/(const Mat& a, float s)
{
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = a.at<float>(y, x) / s;
    return m;
}

// ---- matrix products are lazy, as cv::MatExpr makes them: `A * b + c` is ONE cv::gemm(A, b, 1, c, 1) and `-A.t() * b` is
// ONE cv::gemm(A, b, -1, noArray, 0, GEMM_1_T). The arithmetic of the two shapes the reference's SearchByProjection runs
// (src/ORBmatcher.cc:1358-1367, :1383) is the one tests/test_frame_oracle.py pins against cv2 4.13:
//   3x3 * 3x1, not transposed (OpenCV's small-matrix path): float products, float sums left to right, then
//     (float)((double)sum * alpha + (double)c * beta);
//   everything else (the general path): double accumulation over k, times alpha (plus c), one rounding to float.
struct MatTExpr {
    Mat a;
    double alpha;
    operator Mat() const
    {
        Mat m(a.cols, a.rows, CV_32F);
        for (int y = 0; y < a.rows; ++y)
            for (int x = 0; x < a.cols; ++x) m.at<float>(x, y) = (float)(alpha * a.at<float>(y, x));
        return m;
    }
};
inline MatTExpr Mat::t() const { return MatTExpr{*this, 1.0}; }
inline MatTExpr operator-(const MatTExpr& t) { return MatTExpr{t.a, -t.alpha}; }

inline Mat gemm_(const Mat& a, bool aT, const Mat& b, double alpha, const Mat* c)
{
    const int M = aT ? a.cols : a.rows, K = aT ? a.rows : a.cols, N = b.cols;
    assert(b.rows == K);
    Mat m(M, N, CV_32F);
    const bool small = !aT && M == 3 && K == 3 && N == 1;
    for (int y = 0; y < M; ++y)
        for (int x = 0; x < N; ++x) {
            const double cv = c ? (double)c->at<float>(y, x) : 0.0;
            if (small) {
                volatile float p0 = a.at<float>(y, 0) * b.at<float>(0, x), p1 = a.at<float>(y, 1) * b.at<float>(1, x),
                               p2 = a.at<float>(y, 2) * b.at<float>(2, x);
                volatile float sum = p0 + p1;
                sum = sum + p2;
                m.at<float>(y, x) = (float)((double)sum * alpha + cv * 1.0);
            } else {
                double acc = 0;
                for (int k = 0; k < K; ++k) acc += (double)(aT ? a.at<float>(k, y) : a.at<float>(y, k)) * (double)b.at<float>(k, x);
                m.at<float>(y, x) = (float)(acc * alpha + cv);
            }
        }
    return m;
}
struct MatMulExpr {
    Mat a, b;
    bool aT;
    double alpha;
    operator Mat() const { return gemm_(a, aT, b, alpha, nullptr); }
    template <typename T> T at(int i) const { return Mat(*this).at<T>(i); }
};
inline MatMulExpr operator*(const Mat& a, const Mat& b) { return MatMulExpr{a, b, false, 1.0}; }
inline MatMulExpr operator*(const MatTExpr& t, const Mat& b) { return MatMulExpr{t.a, b, true, t.alpha}; }
inline MatMulExpr operator*(const MatMulExpr& m, const Mat& b) { return MatMulExpr{Mat(m), b, false, 1.0}; }
inline MatMulExpr operator*(const Mat& a, const MatTExpr& t) { return MatMulExpr{a, Mat(t), false, 1.0}; }
inline MatMulExpr operator*(const MatMulExpr& m, const MatTExpr& t) { return MatMulExpr{Mat(m), Mat(t), false, 1.0}; }
inline MatMulExpr operator*(float s, const MatMulExpr& m) { return MatMulExpr{m.a, m.b, m.aT, m.alpha * s}; }
inline MatMulExpr operator-(const MatMulExpr& m) { return MatMulExpr{m.a, m.b, m.aT, -m.alpha}; }
inline Mat operator+(const MatMulExpr& m, const Mat& c) { return gemm_(m.a, m.aT, m.b, m.alpha, &c); }
inline Mat operator-(const MatMulExpr& m, const Mat& c) { return Mat(m) - c; }
inline Mat operator-(const Mat& c, const MatMulExpr& m) { return c - Mat(m); }
inline Mat operator+(const Mat& c, const MatMulExpr& m) { return gemm_(m.a, m.aT, m.b, m.alpha, &c); }
enum { NORM_L1 = 2, NORM_L2 = 4 };
inline double norm(const Mat& a)
{
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) s += (double)a.at<float>(y, x) * a.at<float>(y, x);
    return std::sqrt(s);
}
// cv::norm(a, b, NORM_L1) on CV_32F: sum of |a - b|, accumulated in double (src/Frame.cc:644)
inline double norm(const Mat& a, const Mat& b, int normType)
{
    assert(normType == NORM_L1);
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) s += std::fabs((double)a.at<float>(y, x) - (double)b.at<float>(y, x));
    return s;
}
void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& D, const Mat& R, const Mat& P);  // aborts: never reached

// cv::Mat_<float>(r, c) << a, b, c (src/Frame.cc:734)
template <typename T>
class Mat_ : public Mat {
public:
    Mat_(int r, int c) : Mat(r, c, CV_32F) {}
    struct Init {
        Mat_* m;
        int i;
        Init& operator,(T v)
        {
            m->template at<T>(i / m->cols, i % m->cols) = v;
            ++i;
            return *this;
        }
        operator Mat() const { return *m; }
    };
    Init operator<<(T v)
    {
        this->template at<T>(0, 0) = v;
        return Init{this, 1};
    }
};

// cv::FileStorage / cv::FileNode: the YAML save/load of DBoW2's TemplatedVocabulary are virtual, so they are instantiated
// and must compile; nothing here ever calls them (vocabularies come in through loadFromTextFile). Everything aborts.
class FileNode {
public:
    FileNode operator[](const char*) const { std::abort(); }
    FileNode operator[](const std::string&) const { std::abort(); }
    FileNode operator[](int) const { std::abort(); }
    size_t size() const { std::abort(); }
    operator int() const { std::abort(); }
    operator double() const { std::abort(); }
    operator std::string() const { std::abort(); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage() {}
    FileStorage(const std::string&, int) { std::abort(); }
    bool isOpened() const { return false; }
    void release() {}
    FileNode operator[](const std::string&) const { std::abort(); }
    FileNode operator[](const char*) const { std::abort(); }
};
template <typename T> inline FileStorage& operator<<(FileStorage& fs, const T&) { std::abort(); return fs; }

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };
enum { INTER_LINEAR = 1 };

// ---- the five OpenCV functions on the path; defined in oracle/ref_wrap.cpp over oracle/prims.cpp ----
void resize(const Mat& src, Mat& dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR);
void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int borderType);
void FAST(const Mat& image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true);
void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_REFLECT_101);
float fastAtan2(float y, float x);

// referenced only by ComputeKeyPointsOld, which the reference never calls (src/ORBextractor.cc:1055 is commented out)
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& keypoints, int npoints);
};

}  // namespace cv
