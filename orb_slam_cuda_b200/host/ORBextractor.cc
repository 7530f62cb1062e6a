// host/ORBextractor.cc — ORB_SLAM2::ORBextractor (include/ORBextractor.h) over the C ABI of liborbx.so.
// Mirrors the behaviour of the reference's operator() CPU path (src/ORBextractor.cc:1538-1548, 1701-1815) at the class
// boundary: empty image -> silent return, CV_8UC1 asserted, keypoints cleared, descriptors created N x 32 CV_8U or
// released when N == 0.  All pixel work happens on the device; this file only moves containers.
#include "ORBextractor.h"

#include <cassert>
#include <cstring>
#include <stdexcept>
#include <string>

namespace ORB_SLAM2 {

static void check(int rc, const char* what) {
    if (rc != ORBX_OK) throw std::runtime_error(std::string(what) + ": " + orbx_last_error());
}

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST, int width, int height)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST), mpHandle(nullptr) {
    init(width, height);
}

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST), mpHandle(nullptr) {
    init(0, 0);
}

void ORBextractor::init(int width, int height) {
    orbx_params p;
    p.nfeatures = nfeatures;
    p.scale_factor = (float)scaleFactor;
    p.nlevels = nlevels;
    p.ini_th_fast = iniThFAST;
    p.min_th_fast = minThFAST;
    p.width = width > 0 ? width : 0;
    p.height = height > 0 ? height : 0;
    p.max_batch = 1;
    p.device = -1;
    check(orbx_extractor_create(&p, &mpHandle), "ORBextractor");
    mvScaleFactor.resize(nlevels);
    mvInvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    umax.resize(16);
    check(orbx_get_scale_tables(mpHandle, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                                mvInvLevelSigma2.data(), mnFeaturesPerLevel.data()),
          "ORBextractor");
    check(orbx_get_umax(mpHandle, umax.data()), "ORBextractor");
    mvImagePyramid.resize(nlevels);
    const int cap = nfeatures + 4 * nlevels;  // DistributeOctTree may overshoot a level's quota by up to 3
    mKeyBuf.resize(cap);
    mDescBuf.resize((size_t)cap * 32);
}

ORBextractor::~ORBextractor() { orbx_extractor_destroy(mpHandle); }

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                              cv::OutputArray _descriptors) {
    if (_image.empty()) return;
    cv::Mat image = _image.getMat();
    assert(image.type() == CV_8UC1);
    int n = 0;
    check(orbx_extract(mpHandle, image.data, image.cols, image.rows, (size_t)image.step, mKeyBuf.data(), mDescBuf.data(),
                       (int)mKeyBuf.size(), &n),
          "ORBextractor::operator()");
    _keypoints.clear();
    if (n == 0) {
        _descriptors.release();
    } else {
        _descriptors.create(n, 32, CV_8U);
        cv::Mat descriptors = _descriptors.getMat();
        for (int i = 0; i < n; ++i) std::memcpy(descriptors.ptr(i), mDescBuf.data() + (size_t)i * 32, 32);
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "orbx_keypoint must be bit-compatible with cv::KeyPoint");
        _keypoints.resize(n);
        std::memcpy(static_cast<void*>(_keypoints.data()), mKeyBuf.data(), (size_t)n * sizeof(orbx_keypoint));
    }
    if (keepImagePyramid) {
        const int E = 19;  // EDGE_THRESHOLD
        for (int l = 0; l < nlevels; ++l) {
            const uint8_t* interior = nullptr;
            int w = 0, h = 0;
            size_t pitch = 0;
            check(orbx_pyramid_level_host(mpHandle, 0, l, &interior, &w, &h, &pitch), "ORBextractor::mvImagePyramid");
            cv::Mat parent(h + 2 * E, w + 2 * E, CV_8UC1, const_cast<uint8_t*>(interior) - (size_t)E * pitch - E, pitch);
            mvImagePyramid[l] = parent(cv::Range(E, E + h), cv::Range(E, E + w));
        }
    }
}

}  // namespace ORB_SLAM2
