// oracle/ref_matcher/ref_frame_unchanged.cpp — TEST INFRASTRUCTURE: the north_star's "Frame.cc and ComputeStereoMatches run unchanged
// on top of it", literally.  The REFERENCE's own src/Frame.cc (compiled where it lies, /root/reference) is linked against THIS
// repository's class surface — include/ORBextractor.h + orb_slam_cuda_b200/host/ORBextractor.cc and include/ORBmatcher.h over
// liborbx.so — with the test-only <opencv2/...> shim standing in for OpenCV (which also compiles the ORBX_HAVE_OPENCV branches of
// the class layer).  The reference's stereo constructor (src/Frame.cc:61-116) then runs as it is: two std::threads call
// (*mpORBextractorLeft/Right)(im, cv::Mat(), mvKeys, mDescriptors) -> CUDA extraction; its UNCHANGED ComputeStereoMatches body
// (:465-639) reads mpORBextractorLeft->mvImagePyramid[l].rowRange().colRange() (host copies of the device pyramid) and calls
// ORBmatcher::DescriptorDistance per candidate pair (one device call each: slow, and correct).
//   ref_frame_unchanged in.raw w h nfeatures bf fx out.bin      (left frame, right frame)
// out.bin: keys + descriptors of both eyes, mvuRight, mvDepth, then the 64 x 48 grid of the reference's AssignFeaturesToGrid.
#include "ref_stubs.h"

#include "Frame.h"

#include <cstdio>

namespace ORB_SLAM2 {
std::vector<SurgeryOp> g_surgery;
int MapPoint::PredictScale(const float& currentDist, Frame* pF) {  // src/MapPoint.cc:407-422 (Frame::isInFrustum links against it)
    const float ratio = mfMaxDistance / currentDist;
    int nScale = ceil(log(ratio) / pF->mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= pF->mnScaleLevels) nScale = pF->mnScaleLevels - 1;
    return nScale;
}
}  // namespace ORB_SLAM2

using namespace ORB_SLAM2;

static void put(FILE* f, const void* p, size_t n) { if (n && fwrite(p, 1, n, f) != n) { perror("fwrite"); exit(3); } }
static void put_i(FILE* f, int v) { put(f, &v, 4); }
static void put_keys(FILE* f, const std::vector<cv::KeyPoint>& k, const cv::Mat& d) {
    put_i(f, (int)k.size());
    put(f, k.data(), k.size() * sizeof(cv::KeyPoint));
    for (int i = 0; i < d.rows; ++i) put(f, d.ptr(i), 32);
}

int main(int argc, char** argv) {
    if (argc < 8) { fprintf(stderr, "usage: ref_frame_unchanged in.raw w h nfeatures bf fx out.bin\n"); return 2; }
    const int w = atoi(argv[2]), h = atoi(argv[3]), nf = atoi(argv[4]);
    const float bf = (float)atof(argv[5]), fx = (float)atof(argv[6]);
    std::vector<unsigned char> raw((size_t)2 * w * h);
    FILE* in = fopen(argv[1], "rb");
    if (!in || fread(raw.data(), 1, raw.size(), in) != raw.size()) { fprintf(stderr, "cannot read two %dx%d frames\n", w, h); return 2; }
    fclose(in);
    cv::Mat imL(h, w, CV_8UC1, raw.data()), imR(h, w, CV_8UC1, raw.data() + (size_t)w * h);
    try {
        ORBextractor exL(nf, 1.2f, 8, 20, 7, w, h), exR(nf, 1.2f, 8, 20, 7, w, h);
        ORBVocabulary voc;
        cv::Mat K = cv::Mat::eye(3, 3, CV_32F), dist(4, 1, CV_32F);
        K.at<float>(0, 0) = fx; K.at<float>(1, 1) = fx; K.at<float>(0, 2) = 0.5f * w; K.at<float>(1, 2) = 0.5f * h;
        Frame::mbInitialComputations = true;
        Frame::fx = fx; Frame::fy = fx; Frame::cx = 0.5f * w; Frame::cy = 0.5f * h;
        // mb is read by ComputeStereoMatches before the constructor assigns it (src/Frame.cc:498 vs :112): seed the storage with
        // what a running system finds there, mbf / fx of the previous frame
        void* store = ::operator new(sizeof(Frame));
        std::memset(store, 0, sizeof(Frame));
        {
            const float seed = bf / fx;
            std::memcpy((char*)store + ((char*)&((Frame*)store)->mb - (char*)store), &seed, sizeof seed);
        }
        Frame* F = new (store) Frame(imL, imR, 0.0, &exL, &exR, &voc, K, dist, bf, 40.f);
        FILE* out = fopen(argv[7], "wb");
        put_keys(out, F->mvKeys, F->mDescriptors);
        put_keys(out, F->mvKeysRight, F->mDescriptorsRight);
        put(out, F->mvuRight.data(), F->mvuRight.size() * 4);
        put(out, F->mvDepth.data(), F->mvDepth.size() * 4);
        for (int ix = 0; ix < FRAME_GRID_COLS; ++ix)
            for (int iy = 0; iy < FRAME_GRID_ROWS; ++iy) {
                put_i(out, (int)F->mGrid[ix][iy].size());
                for (size_t v : F->mGrid[ix][iy]) put_i(out, (int)v);
            }
        fclose(out);
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    return 0;
}
