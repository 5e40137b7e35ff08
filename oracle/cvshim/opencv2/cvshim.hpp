// oracle/cvshim/opencv2/cvshim.hpp — TEST INFRASTRUCTURE ONLY (see ../../orb_oracle.h).
//
// A minimal stand-in for the part of the OpenCV C++ API that /root/reference/src/ORBextractor.cc uses, so that the
// reference's OWN translation unit can be compiled from where it lies (oracle/Makefile target _ref) although this image
// has no OpenCV C++ headers or libraries. Nothing here is OpenCV code: the containers (Mat, KeyPoint, Point_, Size, Rect,
// InputArray/OutputArray) are written from the documented interface, and the five primitives the reference calls
// (resize, copyMakeBorder, FAST, GaussianBlur, fastAtan2) forward to the oracle's restatements in oracle/prims.cpp, which
// tests/test_oracle_kat.py pins bit for bit against the real OpenCV kernels through cv2 4.13.
// Only CV_8UC1 matrices exist. Anything the reference does not call is absent.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0

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
    int rows, cols;
};

// 8-bit single-channel matrix header over a reference-counted buffer. The buffer and its counter come from malloc/free,
// never from operator new: ref_wrap.cpp replaces operator new by a per-call arena, and pyramid levels outlive a call.
class Mat {
public:
    int rows, cols;
    uchar* data;
    MatStep step;

    Mat() : rows(0), cols(0), data(nullptr), buf(nullptr) {}
    Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr), buf(nullptr) { create(r, c, type); }
    Mat(Size sz, int type) : rows(0), cols(0), data(nullptr), buf(nullptr) { create(sz.height, sz.width, type); }
    Mat(int r, int c, int type, void* user, size_t stp) : rows(r), cols(c), data((uchar*)user), buf(nullptr)
    {
        assert(type == CV_8UC1);
        step.v = stp;
    }
    Mat(const Mat& o) : rows(o.rows), cols(o.cols), data(o.data), step(o.step), buf(o.buf)
    {
        if (buf) ++*buf;
    }
    Mat& operator=(const Mat& o)
    {
        if (o.buf) ++*o.buf;
        drop();
        rows = o.rows; cols = o.cols; data = o.data; step = o.step; buf = o.buf;
        return *this;
    }
    ~Mat() { drop(); }

    void create(int r, int c, int type)
    {
        assert(type == CV_8UC1);
        if (data && r == rows && c == cols) return;  // cv::Mat::create keeps a buffer of the right shape
        drop();
        rows = r;
        cols = c;
        step.v = (size_t)c;
        buf = (int*)std::malloc(16 + (size_t)r * c);
        *buf = 1;
        data = (uchar*)buf + 16;
    }
    void create(Size sz, int type) { create(sz.height, sz.width, type); }
    void release()
    {
        drop();
        data = nullptr;
        rows = cols = 0;
    }
    static MatZeros zeros(int r, int c, int type)
    {
        assert(type == CV_8UC1);
        return MatZeros{r, c};
    }
    Mat(const MatZeros& z) : rows(0), cols(0), data(nullptr), buf(nullptr) { *this = z; }
    Mat& operator=(const MatZeros& z)
    {
        create(z.rows, z.cols, CV_8UC1);  // keeps a buffer (or view) of the right shape
        for (int y = 0; y < rows; ++y) std::memset(data + (size_t)y * step.v, 0, (size_t)cols);
        return *this;
    }

    int type() const { return CV_8UC1; }
    size_t step1() const { return step.v; }  // step in elements; element size is 1
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    Size size() const { return Size(cols, rows); }

    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
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
        m.data = data + (size_t)r.y * step.v + r.x;
        return m;
    }
    Mat clone() const
    {
        Mat m(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step.v, data + (size_t)y * step.v, (size_t)cols);
        return m;
    }

private:
    int* buf;  // counter at the head of the owned allocation; nullptr for user data
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
