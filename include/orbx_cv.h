// include/orbx_cv.h — the handful of OpenCV core types the ORBextractor / ORBmatcher class surface mentions.
//
// The reference's headers include <opencv/cv.h> (include/ORBextractor.h:26) and pass cv::Mat / cv::KeyPoint /
// cv::InputArray / cv::OutputArray across the class boundary.  When OpenCV's headers are available (define
// ORBX_USE_OPENCV, or let __has_include find them) they are used as they are.  This image has no OpenCV C++
// headers (only the cv2 Python wheel), so otherwise a minimal, layout-compatible stand-in of exactly the members the
// front end and its callers (src/Frame.cc:246-252, :465-639; src/ORBmatcher.cc:405-520) touch is defined here.
// No pixel arithmetic lives in this file.
#ifndef ORBX_CV_H_
#define ORBX_CV_H_

#if !defined(ORBX_NO_OPENCV) && (defined(ORBX_USE_OPENCV) || (defined(__has_include) && __has_include(<opencv2/core/core.hpp>)))
#include <opencv2/core/core.hpp>
#define ORBX_HAVE_OPENCV 1
#else
#define ORBX_HAVE_OPENCV 0

#include <cassert>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#ifndef CV_8U
#define CV_8U 0
#define CV_8UC1 0
#define CV_32S 4
#define CV_32F 5
#endif
#ifndef CV_PI
#define CV_PI 3.1415926535897932384626433832795
#endif

namespace cv {

typedef unsigned char uchar;

template <class T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
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

// Same field order and size (28 bytes) as OpenCV's cv::KeyPoint.
class KeyPoint {
public:
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

// Reference-counted 2-D matrix header: just enough of cv::Mat for 8-bit images and N x 32 descriptor tables.
class Mat {
public:
    int rows, cols;
    uchar* data;
    size_t step;  // bytes per row

    Mat() : rows(0), cols(0), data(nullptr), step(0), type_(CV_8U) {}
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    Mat(int r, int c, int type, void* ext, size_t step_ = 0) : rows(r), cols(c), data((uchar*)ext), type_(type) {
        step = step_ ? step_ : (size_t)c * elemSize();
    }
    void create(int r, int c, int type) {
        if (r == rows && c == cols && type == type_ && data && own_ && data == own_->data()) return;
        type_ = type;
        rows = r;
        cols = c;
        step = (size_t)c * elemSize();
        own_ = std::make_shared<std::vector<uchar>>((size_t)r * step);
        data = own_->data();
    }
    void release() {
        own_.reset();
        data = nullptr;
        rows = cols = 0;
        step = 0;
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    int channels() const { return 1; }
    size_t elemSize() const { return (type_ == CV_32S || type_ == CV_32F) ? 4 : 1; }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }
    Size size() const { return Size(cols, rows); }
    template <class T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
    template <class T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }
    template <class T> T& at(int r, int c) { return ptr<T>(r)[c]; }
    template <class T> const T& at(int r, int c) const { return ptr<T>(r)[c]; }
    Mat row(int r) const { return roi(r, r + 1, 0, cols); }
    Mat rowRange(int a, int b) const { return roi(a, b, 0, cols); }
    Mat colRange(int a, int b) const { return roi(0, rows, a, b); }
    Mat operator()(const Range& rr, const Range& cr) const { return roi(rr.start, rr.end, cr.start, cr.end); }
    static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }  // create() value-initialises its storage
    Mat clone() const {
        Mat m(rows, cols, type_);
        for (int r = 0; r < rows; ++r) std::memcpy(m.ptr(r), ptr(r), (size_t)cols * elemSize());
        return m;
    }
    // keeps `holder` alive as long as this header (or a copy / ROI of it) lives
    void setHolder(std::shared_ptr<void> holder) { hold_ = std::move(holder); }

private:
    Mat roi(int r0, int r1, int c0, int c1) const {
        Mat m;
        m.rows = r1 - r0;
        m.cols = c1 - c0;
        m.step = step;
        m.type_ = type_;
        m.data = data + (size_t)r0 * step + (size_t)c0 * elemSize();
        m.own_ = own_;
        m.hold_ = hold_;
        return m;
    }
    int type_;
    std::shared_ptr<std::vector<uchar>> own_;
    std::shared_ptr<void> hold_;
};

// cv::InputArray / cv::OutputArray as used by operator(): image.getMat(), descriptors.create / getMat / release.
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

}  // namespace cv
#endif  // OpenCV or stand-in

#endif  // ORBX_CV_H_
