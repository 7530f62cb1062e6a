// oracle/ref_matcher/shim/opencv2/core/core.hpp — TEST INFRASTRUCTURE.
// Lets the reference's own src/ORBmatcher.cc and src/Frame.cc (compiled where they lie by oracle/ref_matcher/Makefile) build
// in an image without OpenCV C++ headers.  cv::Mat here is a small dense 2-D matrix (CV_8U images / descriptor tables,
// CV_32F poses and points) with exactly the operations those two files use.  Arithmetic follows OpenCV's CV_32F paths:
// products, dot and norm accumulate in double and round once to float (GEMMSingleMul<float,double>, dotProd_32f,
// normL2_32f), sums and differences are float, a scalar factor is rounded to float first (cvtScale).  Nothing in the
// product or in oracle/orb_oracle.cpp includes this file.
#ifndef ORBX_REF_MATCHER_SHIM_CORE_HPP
#define ORBX_REF_MATCHER_SHIM_CORE_HPP
// the real header pulls these in; the reference relies on that
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <limits>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_32S 4
#define CV_32F 5
#define CV_32FC1 5
#define CV_PI 3.1415926535897932384626433832795

namespace cv {
typedef unsigned char uchar;
enum { NORM_L1 = 2, NORM_L2 = 4 };

template <class T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<int> Point;
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
struct Range {
    int start, end;
    Range(int s, int e) : start(s), end(e) {}
};

class KeyPoint {  // field order of OpenCV's cv::KeyPoint
public:
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};

class Mat {
public:
    int rows, cols;
    uchar* data;
    size_t step;  // bytes per row

    Mat() : rows(0), cols(0), data(nullptr), step(0), type_(CV_8U) {}
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    Mat(int r, int c, int type, void* ext, size_t step_ = 0) : rows(r), cols(c), data((uchar*)ext), type_(type) {  // user-owned data
        step = step_ ? step_ : (size_t)c * elemSize();
    }
    Mat operator()(const Range& rr, const Range& cr) const { return roi(rr.start, rr.end, cr.start, cr.end); }
    bool isContinuous() const { return step == (size_t)cols * elemSize() || rows <= 1; }
    int channels() const { return 1; }
    Size size() const { return Size(cols, rows); }
    void release() {
        own_.reset();
        data = nullptr;
        rows = cols = 0;
        step = 0;
    }
    void create(int r, int c, int type) {
        type_ = type;
        rows = r;
        cols = c;
        step = (size_t)c * elemSize();
        own_ = std::make_shared<std::vector<uchar>>((size_t)r * step + 16);
        data = own_->data();
    }
    static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }
    static Mat ones(int r, int c, int type) {
        Mat m(r, c, type);
        for (int i = 0; i < r; ++i)
            for (int j = 0; j < c; ++j) m.set(i, j, 1.0);
        return m;
    }
    static Mat eye(int r, int c, int type) {
        Mat m(r, c, type);
        for (int i = 0; i < std::min(r, c); ++i) m.set(i, i, 1.0);
        return m;
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    size_t elemSize() const { return (type_ == CV_32F || type_ == CV_32S) ? 4 : 1; }
    template <class T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
    template <class T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }
    template <class T> T& at(int r, int c) { return ptr<T>(r)[c]; }
    template <class T> const T& at(int r, int c) const { return ptr<T>(r)[c]; }
    template <class T> T& at(int i) { return rows == 1 ? ptr<T>(0)[i] : cols == 1 ? ptr<T>(i)[0] : ptr<T>(i / cols)[i % cols]; }
    template <class T> const T& at(int i) const { return rows == 1 ? ptr<T>(0)[i] : cols == 1 ? ptr<T>(i)[0] : ptr<T>(i / cols)[i % cols]; }
    Mat row(int r) const { return roi(r, r + 1, 0, cols); }
    Mat col(int c) const { return roi(0, rows, c, c + 1); }
    Mat rowRange(int a, int b) const { return roi(a, b, 0, cols); }
    Mat colRange(int a, int b) const { return roi(0, rows, a, b); }
    Mat clone() const {
        Mat m;
        if (empty()) return m;
        m.create(rows, cols, type_);
        for (int r = 0; r < rows; ++r) std::memcpy(m.ptr(r), ptr(r), (size_t)cols * elemSize());
        return m;
    }
    void copyTo(Mat& dst) const { dst = clone(); }
    Mat reshape(int) const { return *this; }  // channels are not modelled: only reached on the distorted-camera path
    void convertTo(Mat& dst, int type) const {
        Mat m(rows, cols, type);
        for (int r = 0; r < rows; ++r)
            for (int c = 0; c < cols; ++c) m.set(r, c, get(r, c));
        dst = m;
    }
    Mat t() const {
        Mat m(cols, rows, type_);
        for (int r = 0; r < rows; ++r)
            for (int c = 0; c < cols; ++c) m.set(c, r, get(r, c));
        return m;
    }
    double dot(const Mat& o) const {  // dotProd_32f: double accumulator
        double s = 0;
        for (int r = 0; r < rows; ++r)
            for (int c = 0; c < cols; ++c) s += (double)f(r, c) * (double)o.f(r, c);
        return s;
    }
    float f(int r, int c) const { return ptr<float>(r)[c]; }
    double get(int r, int c) const { return type_ == CV_32F ? (double)ptr<float>(r)[c] : type_ == CV_32S ? (double)ptr<int>(r)[c] : (double)ptr(r)[c]; }
    void set(int r, int c, double v) {
        if (type_ == CV_32F) ptr<float>(r)[c] = (float)v;
        else if (type_ == CV_32S) ptr<int>(r)[c] = (int)v;
        else ptr(r)[c] = (uchar)v;
    }

private:
    Mat roi(int r0, int r1, int c0, int c1) const {
        Mat m;
        m.rows = r1 - r0;
        m.cols = c1 - c0;
        m.step = step;
        m.type_ = type_;
        m.data = data + (size_t)r0 * step + (size_t)c0 * elemSize();
        m.own_ = own_;
        return m;
    }
    int type_;
    std::shared_ptr<std::vector<uchar>> own_;
};

// cv::InputArray / cv::OutputArray as ORBextractor::operator() uses them (a Mat converts implicitly, like in OpenCV)
class _InputArray {
public:
    _InputArray() : m_(nullptr) {}
    _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
    Mat getMat() const { return m_ ? *m_ : Mat(); }
    bool empty() const { return !m_ || m_->empty(); }

protected:
    Mat* m_;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat& m) { m_ = &m; }
    void create(int r, int c, int type) const {
        if (m_) m_->create(r, c, type);
    }
    void release() const {
        if (m_) m_->release();
    }
    Mat& getMatRef() const { return *m_; }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

template <class T> class Mat_ : public Mat {
public:
    Mat_(int r, int c) : Mat(r, c, CV_32F) {}
};
template <class T> struct MatCommaInitializer_ {
    Mat m;
    int i;
    template <class U> MatCommaInitializer_& operator,(U v) {
        m.at<T>(i++) = (T)v;
        return *this;
    }
    operator Mat() const { return m; }
};
template <class T, class U> inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, U v) {
    MatCommaInitializer_<T> c{m, 0};
    c, v;
    return c;
}

inline Mat operator*(const Mat& a, const Mat& b) {  // GEMMSingleMul<float, double>
    Mat m(a.rows, b.cols, CV_32F);
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < b.cols; ++c) {
            double s = 0;
            for (int k = 0; k < a.cols; ++k) s += (double)a.f(r, k) * (double)b.f(k, c);
            m.at<float>(r, c) = (float)s;
        }
    return m;
}
inline Mat scaled(const Mat& a, double s) {  // cvtScale: the factor is rounded to float, the product is float
    const float fs = (float)s;
    Mat m(a.rows, a.cols, CV_32F);
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) m.at<float>(r, c) = a.f(r, c) * fs;
    return m;
}
inline Mat operator*(double s, const Mat& a) { return scaled(a, s); }
inline Mat operator*(const Mat& a, double s) { return scaled(a, s); }
inline Mat operator/(const Mat& a, double s) { return scaled(a, 1. / s); }
inline Mat operator-(const Mat& a) { return scaled(a, -1.); }
inline Mat operator+(const Mat& a, const Mat& b) {
    Mat m(a.rows, a.cols, CV_32F);
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) m.at<float>(r, c) = a.f(r, c) + b.f(r, c);
    return m;
}
inline Mat operator-(const Mat& a, const Mat& b) {
    Mat m(a.rows, a.cols, CV_32F);
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) m.at<float>(r, c) = a.f(r, c) - b.f(r, c);
    return m;
}
inline double norm(const Mat& a) {  // normL2_32f: double accumulator
    double s = 0;
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) s += (double)a.f(r, c) * (double)a.f(r, c);
    return std::sqrt(s);
}
inline double norm(const Mat& a, const Mat& b, int type) {  // NORM_L1 of the difference (normDiffL1_32f: double accumulator)
    if (type != NORM_L1) std::abort();
    double s = 0;
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) s += std::fabs((double)a.f(r, c) - (double)b.f(r, c));
    return s;
}
// only reached with a distorted camera model, which the pinning tests do not use (the oracle's UndistortKeyPoints is
// pinned against cv2.undistortPoints instead)
inline void undistortPoints(const Mat&, Mat&, const Mat&, const Mat&, const Mat&, const Mat&) { std::abort(); }
}  // namespace cv
#endif
