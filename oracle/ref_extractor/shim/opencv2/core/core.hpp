// oracle/ref_extractor/shim/opencv2/core/core.hpp — TEST INFRASTRUCTURE.
// Lets the reference's own src/ORBextractor.cc (compiled where it lies by oracle/ref_extractor/Makefile) build in an image without
// OpenCV C++ headers.  cv::Mat here is a small 8-bit matrix with the operations that file uses; the five pixel primitives it calls
// (resize INTER_LINEAR, copyMakeBorder REFLECT_101, FAST, GaussianBlur 7x7 sigma 2, fastAtan2) are the oracle's models of them
// (oracle/orb_oracle.cpp), which tests/test_oracle_cv2.py pins bit-exactly against cv2 4.13.  So the compiled reference supplies
// the CONTROL FLOW (ComputePyramid, ComputeKeyPointsOctTree, DistributeOctTree, IC_Angle, computeOrbDescriptor, operator()) and
// cv2-pinned code supplies the pixels.  Nothing in the product includes this file.
#ifndef ORBX_REF_EXTRACTOR_SHIM_CORE_HPP
#define ORBX_REF_EXTRACTOR_SHIM_CORE_HPP
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <iterator>
#include <list>
#include <memory>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_PI 3.1415926535897932384626433832795

namespace cv {
typedef unsigned char uchar;
enum { INTER_LINEAR = 1, BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };

inline int cvRoundImpl(double v) { return (int)lrint(v); }  // round half to even, as cvRound does on SSE2 / with lrint
}  // namespace cv
inline int cvRound(double v) { return cv::cvRoundImpl(v); }
inline int cvRound(float v) { return (int)lrintf(v); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v) { return (int)floor(v); }
inline int cvCeil(double v) { return (int)ceil(v); }

namespace cv {
template <class T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <class U> Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
    Point_& operator*=(float s) {  // cv::Point2f *= float (the only use: keypoint->pt *= scale)
        x = (T)(x * s);
        y = (T)(y * s);
        return *this;
    }
};
typedef Point_<int> Point;
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;
struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
    int x, y, width, height;
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
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
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};

struct ZerosExpr {  // what Mat::zeros returns; assigned into a Mat below
    int rows, cols, type;
};

class Mat {  // 8-bit single channel only
public:
    int rows, cols;
    size_t step;
    uchar* data;
    std::shared_ptr<std::vector<uchar>> buf;

    Mat() : rows(0), cols(0), step(0), data(nullptr) {}
    Mat(int r, int c, int type) { alloc(r, c, type); }
    Mat(Size s, int type) { alloc(s.height, s.width, type); }
    Mat(int r, int c, int type, void* ext, size_t step_ = 0) : rows(r), cols(c), step(step_ ? step_ : (size_t)c), data((uchar*)ext) { (void)type; }
    Mat(const ZerosExpr& z) { alloc(z.rows, z.cols, z.type); }

    void alloc(int r, int c, int type) {
        assert(type == CV_8UC1);
        (void)type;
        rows = r;
        cols = c;
        step = (size_t)c;
        buf = std::make_shared<std::vector<uchar>>((size_t)r * c + 64, (uchar)0);
        data = buf->data();
    }
    void create(int r, int c, int type) {  // OpenCV: no reallocation when the shape already fits
        if (data && rows == r && cols == c) return;
        alloc(r, c, type);
    }
    void create(Size s, int type) { create(s.height, s.width, type); }
    void release() { *this = Mat(); }
    // OpenCV's MatExpr assignment: create() (a no-op on a header of the same shape, e.g. a rowRange view) then fill
    Mat& operator=(const ZerosExpr& z) {
        create(z.rows, z.cols, z.type);
        for (int y = 0; y < rows; ++y) memset(data + (size_t)y * step, 0, (size_t)cols);
        return *this;
    }
    static ZerosExpr zeros(int r, int c, int type) { return ZerosExpr{r, c, type}; }

    int type() const { return CV_8UC1; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    Size size() const { return Size(cols, rows); }
    size_t step1() const { return step; }
    bool isContinuous() const { return step == (size_t)cols; }
    template <class T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    template <class T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <class T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
    template <class T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }

    Mat view(int y0, int x0, int r, int c) const {
        Mat m;
        m.rows = r;
        m.cols = c;
        m.step = step;
        m.data = data + (size_t)y0 * step + x0;
        m.buf = buf;
        return m;
    }
    Mat rowRange(int a, int b) const { return view(a, 0, b - a, cols); }
    Mat colRange(int a, int b) const { return view(0, a, rows, b - a); }
    Mat rowRange(const Range& r) const { return rowRange(r.start, r.end); }
    Mat colRange(const Range& r) const { return colRange(r.start, r.end); }
    Mat operator()(const Rect& r) const { return view(r.y, r.x, r.height, r.width); }
    Mat clone() const {
        Mat m(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; ++y) memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols);
        return m;
    }
    void copyTo(Mat& d) const {
        d.create(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; ++y) memcpy(d.data + (size_t)y * d.step, data + (size_t)y * step, (size_t)cols);
    }
    std::vector<uchar> packed() const {  // contiguous copy handed to the oracle's primitives
        std::vector<uchar> v((size_t)rows * cols);
        for (int y = 0; y < rows; ++y) memcpy(v.data() + (size_t)y * cols, data + (size_t)y * step, (size_t)cols);
        return v;
    }
    void unpack(const uchar* src) {
        for (int y = 0; y < rows; ++y) memcpy(data + (size_t)y * step, src + (size_t)y * cols, (size_t)cols);
    }
};

class _InputArray {
public:
    const Mat* m;
    Mat own;
    _InputArray() : m(nullptr) {}
    _InputArray(const Mat& mm) : m(&mm) {}
    bool empty() const { return !m || m->empty(); }
    Mat getMat() const { return m ? *m : Mat(); }
};
class _OutputArray {
public:
    Mat* m;
    _OutputArray(Mat& mm) : m(&mm) {}
    void create(int r, int c, int type) const { m->create(r, c, type); }
    void release() const { m->release(); }
    Mat getMat() const { return *m; }
    bool empty() const { return m->empty(); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
}  // namespace cv

// ---- the five pixel primitives: the oracle's cv2-pinned models (oracle/orb_oracle.cpp) ---------------------------------------
#include "orb_oracle.hpp"

namespace cv {
inline float fastAtan2(float y, float x) { return orc::fast_atan2(y, x); }

// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR): dst.create(dsize) (kept when it already has that shape — the reference
// resizes into a ROI of its bordered plane), then the 8-bit fixed-point bilinear kernel
inline void resize(const Mat& src, Mat& dst, Size dsize, double fx = 0, double fy = 0, int interp = INTER_LINEAR) {
    assert(fx == 0 && fy == 0 && interp == INTER_LINEAR);
    (void)fx; (void)fy; (void)interp;
    std::vector<uchar> s = src.packed();
    dst.create(dsize.height, dsize.width, CV_8UC1);
    std::vector<uchar> d((size_t)dsize.width * dsize.height);
    orc::resize_linear_8u(s.data(), src.cols, src.rows, (size_t)src.cols, d.data(), dsize.width, dsize.height, (size_t)dsize.width);
    dst.unpack(d.data());
}

// cv::copyMakeBorder with BORDER_REFLECT_101 (+BORDER_ISOLATED: never look outside the source ROI; the un-flagged call in the
// reference has a whole matrix as source, so both behave the same).  The source may be a ROI of dst (the reference's case).
inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int borderType) {
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101 && top == bottom && top == left && top == right);
    (void)borderType; (void)bottom; (void)left; (void)right;
    std::vector<uchar> s = src.packed();
    const int W = src.cols + 2 * top, H = src.rows + 2 * top;
    dst.create(H, W, CV_8UC1);
    std::vector<uchar> d((size_t)W * H);
    orc::border_reflect101(s.data(), src.cols, src.rows, (size_t)src.cols, d.data(), (size_t)W, top);
    dst.unpack(d.data());
}

inline void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sx, double sy = 0, int borderType = BORDER_REFLECT_101) {
    assert(ksize.width == 7 && ksize.height == 7 && sx == 2 && sy == 2 && borderType == BORDER_REFLECT_101);
    (void)ksize; (void)sx; (void)sy; (void)borderType;
    std::vector<uchar> s = src.packed();  // taken before dst is touched: the reference blurs in place
    const int w = src.cols, h = src.rows;
    dst.create(h, w, CV_8UC1);
    std::vector<uchar> d((size_t)w * h);
    orc::gaussian7x7_s2(s.data(), w, h, (size_t)w, d.data(), (size_t)w);
    dst.unpack(d.data());
}

// cv::FAST(image, keypoints, threshold, nonmaxSuppression) — TYPE_9_16; size 7, angle -1, octave 0, class_id -1
inline void FAST(const Mat& img, std::vector<KeyPoint>& kps, int threshold, bool nms = true) {
    std::vector<orc::KeyPoint> o;
    orc::fast9_16(img.data, img.cols, img.rows, img.step, threshold, nms, o);
    kps.clear();
    kps.reserve(o.size());
    for (const orc::KeyPoint& k : o) kps.push_back(KeyPoint(k.x, k.y, k.size, k.angle, k.response, k.octave, k.class_id));
}

// KeyPointsFilter::retainBest is named by ComputeKeyPointsOld only (dead code: operator() calls ComputeKeyPointsOctTree)
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& k, int n) {
        if (n >= 0 && (int)k.size() > n) {
            std::stable_sort(k.begin(), k.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
            k.resize(n);
        }
    }
};
inline void imshow(const std::string&, const Mat&) {}
inline int waitKey(int = 0) { return 0; }
}  // namespace cv
#endif
