// oracle/ref_dbow2/shim/opencv2/core/core.hpp — TEST INFRASTRUCTURE.  Lets the reference's own DBoW2 sources
// (/root/reference/Thirdparty/DBoW2, compiled where they lie by oracle/ref_dbow2/Makefile) build in an image that has no
// OpenCV C++ headers: cv::Mat is the stand-in of include/orbx_cv.h; cv::FileStorage / cv::FileNode are declared only as far as
// TemplatedVocabulary.h's YAML save / load members need to parse — they are never called (the text loader is what ORB-SLAM2
// uses, src/System.cc:65) and abort if they ever are.
#ifndef ORBX_REF_DBOW2_SHIM_CORE_HPP
#define ORBX_REF_DBOW2_SHIM_CORE_HPP
#define ORBX_NO_OPENCV 1
// the real header pulls these in; the reference relies on that
#include <math.h>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <limits>
#include <numeric>
#include <sstream>
#include <string>

#include "orbx_cv.h"

namespace cv {
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
    FileStorage(const char*, int) { std::abort(); }
    bool isOpened() const { return false; }
    FileNode operator[](const std::string&) const { std::abort(); }
};
template <class T> inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }
}  // namespace cv
#endif
