// tests/cpp/frame_harness.cpp — a Frame-shaped C++ caller of the class surface (include/ORBextractor.h,
// include/ORBmatcher.h), written the way the reference's src/Frame.cc and src/Tracking.cc use it:
//   Frame::Frame(mono)    : ExtractORB(0, im) -> (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors)   (:246-252)
//   Frame::Frame(stereo)  : two std::threads, one extractor per eye (:77-80), then ComputeStereoMatches (:89)
//   Tracking::MonocularInitialization : ORBmatcher(0.9, true).SearchForInitialization(F1, F2, prev, m12, 100) (:671)
// Input: a raw file of 8-bit frames; output: a raw dump that tests/test_cpp_boundary.py compares with the CPU oracle.
//   frame_harness mono   in.raw w h nfeatures out.bin          (two frames: extraction + SearchForInitialization)
//   frame_harness stereo in.raw w h nfeatures bf fx out.bin    (left, right: extraction + ComputeStereoMatches)
//   frame_harness bow    in.raw w h nfeatures voc.txt out.bin  (System.cc:64 loadFromTextFile, Frame::ComputeBoW, SearchByBoW)
//   frame_harness track  in.raw w h nfeatures th out.bin       (last, current: Tracking::TrackWithMotionModel's SearchByProjection, :962)
//   frame_harness fuse   in.raw w h nfeatures th out.bin       (last, current: both ORBmatcher::Fuse overloads incl. their map surgery)
//   frame_harness latency in.raw w h nfeatures iterations      (p50 / p95 of ORBextractor::operator() as one JSON line; no output file)
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <chrono>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "ORBextractor.h"
#include "ORBVocabulary.h"
#include "ORBmatcher.h"

using namespace ORB_SLAM2;

struct KeyFrame;
struct SurgeryOp { int op; long a, b; };  // 0 AddObservation(point a, idx b), 1 AddMapPoint(point a, idx b), 2 a->Replace(b)
static std::vector<SurgeryOp> g_surgery;

struct MapPoint {  // the members of ORB_SLAM2::MapPoint that SearchByProjection reads (include/MapPoint.h)
    // Fuse (src/ORBmatcher.cc:825-1100): the observation bookkeeping of src/MapPoint.cc:83-95 (AddObservation), :179-220 (Replace)
    long mnId = -1;
    std::map<KeyFrame*, size_t> mObservations;
    bool IsInKeyFrame(KeyFrame* k) const { return mObservations.count(k) != 0; }
    void addObservationQuiet(KeyFrame* k, size_t idx);
    void AddObservation(KeyFrame* k, size_t idx);
    void Replace(MapPoint* pMP);
    cv::Mat mWorldPos, mDescriptor;
    int nObs = 0;
    bool mbTrackInView = false, mbBad = false;  // Tracking::SearchLocalPoints state (include/MapPoint.h:111-117)
    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0, mTrackViewCos = 0;
    int mnTrackScaleLevel = 0;
    bool isBad() const { return mbBad; }
    int idxInOther = -1;  // SearchBySim3: GetIndexInKeyFrame(pKF2)
    template <class K> int GetIndexInKeyFrame(K*) const { return idxInOther; }
    float GetMinDistanceInvariance() const { return 3.5f; }
    float GetMaxDistanceInvariance() const { return 15.5f; }
    template <class F> int PredictScale(float, F*) const { return mnTrackScaleLevel; }
    cv::Mat mNormal;
    cv::Mat GetNormal() const { return mNormal; }
    cv::Mat GetWorldPos() const { return mWorldPos; }
    cv::Mat GetDescriptor() const { return mDescriptor; }
    int Observations() const { return nObs; }
};

struct KeyFrame {  // the members of ORB_SLAM2::KeyFrame that the relocalisation SearchByProjection reads (include/KeyFrame.h)
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<MapPoint*> GetMapPointMatches() const { return mvpMapPoints; }
    std::map<unsigned int, std::vector<unsigned int>> mFeatVec;  // DBoW2::FeatureVector (SearchByBoW, src/ORBmatcher.cc:159,522)
    // SearchForTriangulation (src/ORBmatcher.cc:657-823)
    int N = 0;
    std::vector<float> mvuRight, mvLevelSigma2;
    cv::Mat mRcw, mtcw, mOw;
    cv::Mat GetRotation() const { return mRcw; }
    cv::Mat GetTranslation() const { return mtcw; }
    cv::Mat GetCameraCenter() const { return mOw; }
    MapPoint* GetMapPoint(size_t i) const { return mvpMapPoints[i]; }
    // Fuse: src/KeyFrame.cc:209-213 (AddMapPoint), :215-225, :236-249 (GetMapPoints)
    float mbf = 0;
    std::vector<float> mvInvLevelSigma2;
    void AddMapPoint(MapPoint* pMP, size_t idx) { g_surgery.push_back({1, pMP->mnId, (long)idx}); mvpMapPoints[idx] = pMP; }
    void EraseMapPointMatch(size_t idx) { mvpMapPoints[idx] = nullptr; }
    void ReplaceMapPointMatch(size_t idx, MapPoint* pMP) { mvpMapPoints[idx] = pMP; }
    std::set<MapPoint*> GetMapPoints() const {
        std::set<MapPoint*> s;
        for (MapPoint* p : mvpMapPoints) if (p && !p->isBad()) s.insert(p);
        return s;
    }
    // loop-detection SearchByProjection (src/ORBmatcher.cc:290-403)
    float fx = 0, fy = 0, cx = 0, cy = 0, mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0;
    std::vector<float> mvScaleFactors;
    cv::Mat mDescriptors;
    bool IsInImage(const float& x, const float& y) const { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }
};

void MapPoint::addObservationQuiet(KeyFrame* k, size_t idx) {
    if (mObservations.count(k)) return;
    mObservations[k] = idx;
    nObs += (idx < k->mvuRight.size() && k->mvuRight[idx] >= 0) ? 2 : 1;
}
void MapPoint::AddObservation(KeyFrame* k, size_t idx) {
    g_surgery.push_back({0, mnId, (long)idx});
    addObservationQuiet(k, idx);
}
void MapPoint::Replace(MapPoint* pMP) {
    if (pMP->mnId == mnId) return;
    g_surgery.push_back({2, mnId, pMP->mnId});
    std::map<KeyFrame*, size_t> obs = mObservations;
    mObservations.clear();
    mbBad = true;
    for (auto& kv : obs) {
        if (!pMP->IsInKeyFrame(kv.first)) {
            kv.first->ReplaceMapPointMatch(kv.second, pMP);
            pMP->addObservationQuiet(kv.first, kv.second);
        } else {
            kv.first->EraseMapPointMatch(kv.second);
        }
    }
}

struct Frame {  // the members of ORB_SLAM2::Frame that the front end reads or writes (include/Frame.h)
    cv::Mat mTcw;
    std::map<unsigned int, std::vector<unsigned int>> mFeatVec;  // DBoW2::FeatureVector
    std::map<unsigned int, double> mBowVec;                      // DBoW2::BowVector
    ORBVocabulary* mpORBvocabulary = nullptr;
    float fx = 0, fy = 0, cx = 0, cy = 0;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    std::vector<float> mvScaleFactors;
    ORBextractor *mpORBextractorLeft = nullptr, *mpORBextractorRight = nullptr;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    cv::Mat mDescriptors, mDescriptorsRight;
    std::vector<float> mvuRight, mvDepth;
    float mbf = 0, mb = 0;
    int N = 0;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
    void ExtractORB(int flag, const cv::Mat& im) {  // src/Frame.cc:246-252
        if (flag == 0) (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors);
        else (*mpORBextractorRight)(im, cv::Mat(), mvKeysRight, mDescriptorsRight);
    }
};
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;

static void put(FILE* f, const void* p, size_t n) { if (n && fwrite(p, 1, n, f) != n) { perror("fwrite"); exit(3); } }
static void put_i(FILE* f, int v) { put(f, &v, 4); }
static void put_keys(FILE* f, const std::vector<cv::KeyPoint>& k, const cv::Mat& d) {
    put_i(f, (int)k.size());
    put(f, k.data(), k.size() * sizeof(cv::KeyPoint));
    for (int i = 0; i < d.rows; ++i) put(f, d.ptr(i), 32);
}

int main(int argc, char** argv) {
    if (argc < 7) { fprintf(stderr, "usage: see the header comment\n"); return 2; }
    const std::string mode = argv[1];
    const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
    std::vector<unsigned char> raw((size_t)2 * w * h);
    FILE* in = fopen(argv[2], "rb");
    if (!in || fread(raw.data(), 1, raw.size(), in) != raw.size()) { fprintf(stderr, "cannot read two %dx%d frames from %s\n", w, h, argv[2]); return 2; }
    fclose(in);
    cv::Mat im0(h, w, CV_8UC1, raw.data()), im1(h, w, CV_8UC1, raw.data() + (size_t)w * h);
    try {
        if (mode == "latency") {  // frame_harness latency in.raw w h nfeatures iterations: p50 of ORBextractor::operator() in microseconds
            const int iters = atoi(argv[6]);
            ORBextractor ex(nf, 1.2f, 8, 20, 7, w, h);
            ex.keepImagePyramid = false;  // monocular tracking reads keypoints and descriptors only
            std::vector<cv::KeyPoint> keys;
            cv::Mat desc;
            std::vector<double> us;
            for (int i = 0; i < iters + 10; ++i) {
                const cv::Mat& im = (i & 1) ? im1 : im0;
                const auto t0 = std::chrono::steady_clock::now();
                ex(im, cv::Mat(), keys, desc);
                const auto t1 = std::chrono::steady_clock::now();
                if (i >= 10) us.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
            }
            std::sort(us.begin(), us.end());
            printf("{\"p50_us\": %.1f, \"p95_us\": %.1f, \"calls\": %d, \"keypoints\": %d}\n", us[us.size() / 2], us[us.size() * 95 / 100], iters,
                   (int)keys.size());
        } else if (mode == "mono") {
            FILE* out = fopen(argv[6], "wb");
            ORBextractor ex(nf, 1.2f, 8, 20, 7, w, h);  // the reference's 7-argument constructor (src/Tracking.cc:127)
            Frame F1, F2;
            F1.mpORBextractorLeft = F2.mpORBextractorLeft = &ex;
            Frame::mnMinX = 0; Frame::mnMaxX = (float)w; Frame::mnMinY = 0; Frame::mnMaxY = (float)h;  // no distortion (:441-446)
            F1.ExtractORB(0, im0);
            // mvImagePyramid of the first frame: level 1, including one pixel of the border through the parent
            const cv::Mat& L1 = ex.mvImagePyramid[1];
            put_i(out, L1.cols); put_i(out, L1.rows);
            for (int r = 0; r < L1.rows; ++r) put(out, L1.ptr(r), L1.cols);
            put(out, L1.ptr(0) - L1.step - 1, 1);  // pixel (-1, -1): reflect-101 border, inside the parent allocation
            F2.ExtractORB(0, im1);
            F1.mvKeysUn = F1.mvKeys; F2.mvKeysUn = F2.mvKeys;
            put_keys(out, F1.mvKeys, F1.mDescriptors);
            put_keys(out, F2.mvKeys, F2.mDescriptors);
            std::vector<cv::Point2f> prev(F1.mvKeysUn.size());
            for (size_t i = 0; i < prev.size(); ++i) prev[i] = F1.mvKeysUn[i].pt;
            std::vector<int> m12;
            ORBmatcher matcher(0.9f, true);
            const int n = matcher.SearchForInitialization(F1, F2, prev, m12, 100);
            put_i(out, n);
            put(out, m12.data(), m12.size() * 4);
            put(out, prev.data(), prev.size() * 8);
            // accessors + DescriptorDistance + empty image behaviour
            put_i(out, ex.GetLevels());
            std::vector<float> sf = ex.GetScaleFactors(), is2 = ex.GetInverseScaleSigmaSquares();
            put(out, sf.data(), sf.size() * 4); put(out, is2.data(), is2.size() * 4);
            put_i(out, ORBmatcher::DescriptorDistance(F1.mDescriptors.row(0), F2.mDescriptors.row(0)));
            std::vector<cv::KeyPoint> keep(3);
            cv::Mat keepD(2, 32, CV_8U);
            ex(cv::Mat(), cv::Mat(), keep, keepD);  // silent return, outputs untouched (src/ORBextractor.cc:1542-1543)
            put_i(out, (int)keep.size() * 100 + keepD.rows);
            fclose(out);
        } else if (mode == "stereo" && argc >= 9) {
            FILE* out = fopen(argv[8], "wb");
            ORBextractor exL(nf, 1.2f, 8, 20, 7), exR(nf, 1.2f, 8, 20, 7);  // upstream's 5-argument form
            exL.keepImagePyramid = exR.keepImagePyramid = false;            // the stereo matcher reads device memory
            Frame F;
            F.mpORBextractorLeft = &exL; F.mpORBextractorRight = &exR;
            F.mbf = (float)atof(argv[6]);
            F.mb = F.mbf / (float)atof(argv[7]);
            std::thread tl([&] { F.ExtractORB(0, im0); }), tr([&] { F.ExtractORB(1, im1); });  // src/Frame.cc:77-80
            tl.join(); tr.join();
            F.N = (int)F.mvKeys.size();
            ComputeStereoMatches(F);
            {  // the constructor-time form (device-resident keypoints, no uploads) must give the same floats
                const std::vector<float> u = F.mvuRight, z = F.mvDepth;
                ComputeStereoMatchesOnDevice(F);
                if (u.size() != F.mvuRight.size() || memcmp(u.data(), F.mvuRight.data(), u.size() * 4) != 0 ||
                    memcmp(z.data(), F.mvDepth.data(), z.size() * 4) != 0) {
                    fprintf(stderr, "ComputeStereoMatchesOnDevice differs from ComputeStereoMatches\n");
                    return 3;
                }
            }
            put_keys(out, F.mvKeys, F.mDescriptors);
            put_keys(out, F.mvKeysRight, F.mDescriptorsRight);
            put(out, F.mvuRight.data(), F.mvuRight.size() * 4);
            put(out, F.mvDepth.data(), F.mvDepth.size() * 4);
            fclose(out);
        } else if (mode == "track") {
            FILE* out = fopen(argv[7], "wb");
            const float th = (float)atof(argv[6]);
            ORBextractor ex(nf, 1.2f, 8, 20, 7, w, h);
            Frame Last, Cur;
            Last.mpORBextractorLeft = Cur.mpORBextractorLeft = &ex;
            Frame::mnMinX = 0; Frame::mnMaxX = (float)w; Frame::mnMinY = 0; Frame::mnMaxY = (float)h;
            Last.ExtractORB(0, im0);
            Cur.ExtractORB(0, im1);
            for (Frame* F : {&Last, &Cur}) {
                F->mvKeysUn = F->mvKeys;
                F->N = (int)F->mvKeys.size();
                F->fx = F->fy = 500.f; F->cx = 0.5f * w; F->cy = 0.5f * h;
                F->mbf = 40.f; F->mb = 0.08f;
                F->mvScaleFactors = ex.GetScaleFactors();
                F->mvpMapPoints.assign(F->N, nullptr);
                F->mvbOutlier.assign(F->N, false);
                F->mTcw = cv::Mat(4, 4, CV_32F);
                for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) F->mTcw.at<float>(r, c) = r == c ? 1.f : 0.f;
            }
            // current pose: a small rotation about y and a translation (fixed constants, no libm involved)
            const float Rc[9] = {0.9998f, 0.f, 0.02f, 0.f, 1.f, 0.f, -0.02f, 0.f, 0.9998f}, tc[3] = {0.01f, -0.005f, 0.02f};
            for (int r = 0; r < 3; ++r) {
                for (int c = 0; c < 3; ++c) Cur.mTcw.at<float>(r, c) = Rc[3 * r + c];
                Cur.mTcw.at<float>(r, 3) = tc[r];
            }
            std::vector<MapPoint> points(Last.N);
            for (int i = 0; i < Last.N; ++i) {
                if (i % 5 == 0) continue;  // no map point
                MapPoint& P = points[i];
                const float z = 2.f + (float)(i % 17);
                P.mWorldPos = cv::Mat(3, 1, CV_32F);
                P.mWorldPos.at<float>(0, 0) = (Last.mvKeys[i].pt.x - Last.cx) * z / Last.fx;
                P.mWorldPos.at<float>(1, 0) = (Last.mvKeys[i].pt.y - Last.cy) * z / Last.fy;
                P.mWorldPos.at<float>(2, 0) = z;
                P.mDescriptor = Last.mDescriptors.row(i);
                P.nObs = i % 7 != 0 ? 2 : 0;
                Last.mvpMapPoints[i] = &P;
                Last.mvbOutlier[i] = i % 11 == 0;
            }
            ORBmatcher matcher(0.9f, true);
            const int n = matcher.SearchByProjection(Cur, Last, th, true);
            put_keys(out, Last.mvKeys, Last.mDescriptors);
            put_keys(out, Cur.mvKeys, Cur.mDescriptors);
            put_i(out, n);
            for (int j = 0; j < Cur.N; ++j) put_i(out, Cur.mvpMapPoints[j] ? (int)(Cur.mvpMapPoints[j] - points.data()) : -1);
            // Tracking::SearchLocalPoints (:1270-1277): the same map points "in view" of the current frame at the position of
            // their last-frame keypoint, searched with the local-map matcher on a frame without assignments
            std::vector<MapPoint*> local;
            for (int i = 0; i < Last.N; ++i) {
                MapPoint& P = points[i];
                P.mbTrackInView = i % 5 != 0 && i % 3 != 0;
                P.mbBad = i % 13 == 0;
                P.mTrackProjX = Last.mvKeys[i].pt.x + 7.f;   // synth.shifted_frame moves the image by (7, -4)
                P.mTrackProjY = Last.mvKeys[i].pt.y - 4.f;
                P.mTrackProjXR = P.mTrackProjX - 3.f;
                P.mTrackViewCos = i % 2 ? 0.9995f : 0.9f;
                P.mnTrackScaleLevel = Last.mvKeys[i].octave;
                if (i % 5 != 0) local.push_back(&P);
            }
            Cur.mvpMapPoints.assign(Cur.N, nullptr);
            ORBmatcher localMatcher(0.8f);
            const int nl = localMatcher.SearchByProjection(Cur, local, 3.f);
            put_i(out, nl);
            for (int j = 0; j < Cur.N; ++j) put_i(out, Cur.mvpMapPoints[j] ? (int)(Cur.mvpMapPoints[j] - points.data()) : -1);
            // Tracking::Relocalization (:1445): the last frame as a keyframe, some points already found
            KeyFrame KF;
            KF.mvpMapPoints = Last.mvpMapPoints;
            KF.mvKeysUn = Last.mvKeysUn;
            std::set<MapPoint*> found;
            for (int i = 0; i < Last.N; i += 9) if (Last.mvpMapPoints[i]) found.insert(Last.mvpMapPoints[i]);
            Cur.mvpMapPoints.assign(Cur.N, nullptr);
            for (int j = 0; j < Cur.N; j += 10) Cur.mvpMapPoints[j] = &points[0];  // keypoints that already hold a point
            const int nr = matcher.SearchByProjection(Cur, &KF, found, 10.f, 100);
            put_i(out, nr);
            for (int j = 0; j < Cur.N; ++j) put_i(out, Cur.mvpMapPoints[j] ? (int)(Cur.mvpMapPoints[j] - points.data()) : -1);
            // LoopClosing::ComputeSim3 (:278): the current frame as the keyframe, Scw = 2 x its pose (scale 2), every good point a
            // candidate, every 10th keypoint already matched
            KeyFrame KC;
            KC.mvKeysUn = Cur.mvKeysUn; KC.mDescriptors = Cur.mDescriptors; KC.mvScaleFactors = Cur.mvScaleFactors;
            KC.fx = Cur.fx; KC.fy = Cur.fy; KC.cx = Cur.cx; KC.cy = Cur.cy;
            KC.mnMinX = 0; KC.mnMaxX = (float)w; KC.mnMinY = 0; KC.mnMaxY = (float)h;
            cv::Mat Scw(4, 4, CV_32F);
            for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) Scw.at<float>(r, c) = (r < 3 ? 2.f : 1.f) * Cur.mTcw.at<float>(r, c);
            std::vector<MapPoint*> cands;
            for (int i = 0; i < Last.N; ++i) {
                if (i % 5 == 0) continue;
                points[i].mNormal = cv::Mat(3, 1, CV_32F);
                points[i].mNormal.at<float>(0, 0) = 0.f; points[i].mNormal.at<float>(1, 0) = 0.f; points[i].mNormal.at<float>(2, 0) = 1.f;
                cands.push_back(&points[i]);
            }
            std::vector<MapPoint*> matched(Cur.N, nullptr);
            for (int j = 0; j < Cur.N; j += 10) matched[j] = &points[0];
            ORBmatcher loopMatcher(0.75f, true);
            const int ns = loopMatcher.SearchByProjection(&KC, Scw, cands, matched, 10);
            put_i(out, ns);
            for (int j = 0; j < Cur.N; ++j) put_i(out, matched[j] ? (int)(matched[j] - points.data()) : -1);
            // LoopClosing::ComputeSim3 (:357): SearchBySim3 between the last frame (keyframe 1, its points at depth 5 in its camera)
            // and the current one (keyframe 2, own points likewise), both at the identity pose, Sim3 = translation (-0.07, 0.04, 0),
            // which at fx = 500 and z = 5 is the image shift (7, -4)
            {
                std::vector<MapPoint> p1(Last.N), p2(Cur.N);
                KeyFrame K1, K2;
                auto setup = [&](KeyFrame& K, Frame& F, std::vector<MapPoint>& pts, int skip) {
                    K.mvKeysUn = F.mvKeysUn; K.mDescriptors = F.mDescriptors; K.mvScaleFactors = F.mvScaleFactors;
                    K.fx = K.fy = 500.f; K.cx = 0.5f * w; K.cy = 0.5f * h;
                    K.mnMinX = 0; K.mnMaxX = (float)w; K.mnMinY = 0; K.mnMaxY = (float)h;
                    K.mRcw = cv::Mat(3, 3, CV_32F); K.mtcw = cv::Mat(3, 1, CV_32F);
                    for (int r = 0; r < 3; ++r) { K.mtcw.at<float>(r, 0) = 0.f; for (int c = 0; c < 3; ++c) K.mRcw.at<float>(r, c) = r == c ? 1.f : 0.f; }
                    K.mvpMapPoints.assign(F.N, nullptr);
                    for (int i = 0; i < F.N; ++i) {
                        if (i % skip == 0) continue;
                        MapPoint& P = pts[i];
                        P.mWorldPos = cv::Mat(3, 1, CV_32F);
                        P.mWorldPos.at<float>(0, 0) = (F.mvKeysUn[i].pt.x - K.cx) * 5.f / K.fx;
                        P.mWorldPos.at<float>(1, 0) = (F.mvKeysUn[i].pt.y - K.cy) * 5.f / K.fy;
                        P.mWorldPos.at<float>(2, 0) = 5.f;
                        P.mDescriptor = F.mDescriptors.row(i);
                        P.mnTrackScaleLevel = F.mvKeysUn[i].octave;
                        P.mbBad = i % 17 == 0;
                        K.mvpMapPoints[i] = &P;
                    }
                };
                setup(K1, Last, p1, 6);
                setup(K2, Cur, p2, 7);
                cv::Mat R12(3, 3, CV_32F), t12(3, 1, CV_32F);
                for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) R12.at<float>(r, c) = r == c ? 1.f : 0.f;
                t12.at<float>(0, 0) = -0.07f; t12.at<float>(1, 0) = 0.04f; t12.at<float>(2, 0) = 0.f;
                std::vector<MapPoint*> m12(Last.N, nullptr);
                for (int i = 4; i < Last.N; i += 25) if (K1.mvpMapPoints[i]) { m12[i] = &p2[1]; }   // already matched: skipped
                p2[1].idxInOther = 1;
                ORBmatcher simMatcher(0.75f, true);
                const float s12 = 1.f;
                const int nf2 = simMatcher.SearchBySim3(&K1, &K2, m12, s12, R12, t12, 7.5f);
                put_i(out, nf2);
                for (int i = 0; i < Last.N; ++i) put_i(out, m12[i] ? (int)(m12[i] - p2.data()) : -1);
            }
            // Tracking::TrackReferenceKeyFrame (:842): SearchByBoW(keyframe, frame) with a toy vocabulary — node id = 5 descriptor bits,
            // features appended in index order as DBoW2 does
            auto featvec = [](const cv::Mat& D, std::map<unsigned int, std::vector<unsigned int>>& fv) {
                for (int i = 0; i < D.rows; ++i) fv[(D.at<unsigned char>(i, 5) | (D.at<unsigned char>(i, 17) << 8)) & 31u].push_back(i);
            };
            KF.mDescriptors = Last.mDescriptors;
            featvec(Last.mDescriptors, KF.mFeatVec);
            featvec(Cur.mDescriptors, Cur.mFeatVec);
            std::vector<MapPoint*> bowMatches;
            ORBmatcher bowMatcher(0.7f, true);
            const int nb = bowMatcher.SearchByBoW(&KF, Cur, bowMatches);
            put_i(out, nb);
            for (int j = 0; j < Cur.N; ++j) put_i(out, bowMatches[j] ? (int)(bowMatches[j] - points.data()) : -1);
            // LoopClosing::ComputeSim3 (:304): SearchByBoW(keyframe, keyframe); the second keyframe holds point j % N1 at keypoint j
            KC.mFeatVec = Cur.mFeatVec;
            KC.mvpMapPoints.assign(Cur.N, nullptr);
            for (int j = 0; j < Cur.N; ++j) if (j % 4 != 0) KC.mvpMapPoints[j] = &points[j % Last.N];
            std::vector<MapPoint*> matches12;
            ORBmatcher kfMatcher(0.75f, true);
            const int nk = kfMatcher.SearchByBoW(&KF, &KC, matches12);
            put_i(out, nk);
            for (int i = 0; i < Last.N; ++i) put_i(out, matches12[i] ? (int)(matches12[i] - points.data()) : -1);
            fclose(out);
        } else if (mode == "fuse") {
            // LocalMapping::SearchInNeighbors (src/LocalMapping.cc:519ff) and LoopClosing::SearchAndFuse (src/LoopClosing.cc:644): the
            // current frame is the keyframe (identity pose), the last frame's map points sit at depth 5 where the image shift (7, -4)
            // puts them; some keyframe features already hold points of their own
            FILE* out = fopen(argv[7], "wb");
            const float th = (float)atof(argv[6]);
            ORBextractor ex(nf, 1.2f, 8, 20, 7, w, h);
            Frame Last, Cur;
            Last.mpORBextractorLeft = Cur.mpORBextractorLeft = &ex;
            Last.ExtractORB(0, im0);
            Cur.ExtractORB(0, im1);
            const int N1 = (int)Last.mvKeys.size(), N2 = (int)Cur.mvKeys.size();
            put_keys(out, Last.mvKeys, Last.mDescriptors);
            put_keys(out, Cur.mvKeys, Cur.mDescriptors);
            for (int pass = 0; pass < 2; ++pass) {
                KeyFrame KF;
                KF.mvKeysUn = Cur.mvKeys; KF.mDescriptors = Cur.mDescriptors; KF.N = N2;
                KF.mvScaleFactors = ex.GetScaleFactors(); KF.mvInvLevelSigma2 = ex.GetInverseScaleSigmaSquares();
                KF.fx = KF.fy = 500.f; KF.cx = 0.5f * w; KF.cy = 0.5f * h; KF.mbf = 40.f;
                KF.mnMinX = 0; KF.mnMaxX = (float)w; KF.mnMinY = 0; KF.mnMaxY = (float)h;
                KF.mRcw = cv::Mat(3, 3, CV_32F); KF.mtcw = cv::Mat(3, 1, CV_32F); KF.mOw = cv::Mat(3, 1, CV_32F);
                for (int r = 0; r < 3; ++r) {
                    KF.mtcw.at<float>(r, 0) = 0.f; KF.mOw.at<float>(r, 0) = 0.f;
                    for (int c = 0; c < 3; ++c) KF.mRcw.at<float>(r, c) = r == c ? 1.f : 0.f;
                }
                KF.mvuRight.assign(N2, -1.f);
                for (int j = 0; j < N2; j += 3) KF.mvuRight[j] = Cur.mvKeys[j].pt.x - 8.f;  // stereo keypoints: depth 5 at bf 40
                std::vector<MapPoint> own(N2), pts(N1);
                KF.mvpMapPoints.assign(N2, nullptr);
                for (int j = 0; j < N2; ++j) {
                    if (j % 4 != 1) continue;
                    own[j].mnId = N1 + j;
                    own[j].nObs = 1 + j % 5;
                    own[j].mbBad = j % 29 == 1;
                    own[j].mObservations[&KF] = j;
                    KF.mvpMapPoints[j] = &own[j];
                }
                std::vector<MapPoint*> cands;
                for (int i = 0; i < N1; ++i) {
                    MapPoint& P = pts[i];
                    P.mnId = i;
                    P.mWorldPos = cv::Mat(3, 1, CV_32F);
                    P.mWorldPos.at<float>(0, 0) = (Last.mvKeys[i].pt.x + 7.f - KF.cx) * 5.f / KF.fx;
                    P.mWorldPos.at<float>(1, 0) = (Last.mvKeys[i].pt.y - 4.f - KF.cy) * 5.f / KF.fy;
                    P.mWorldPos.at<float>(2, 0) = 5.f;
                    P.mNormal = cv::Mat(3, 1, CV_32F);
                    P.mNormal.at<float>(0, 0) = 0.f; P.mNormal.at<float>(1, 0) = 0.f; P.mNormal.at<float>(2, 0) = i % 19 == 0 ? -1.f : 1.f;
                    P.mDescriptor = Last.mDescriptors.row(i);
                    P.mnTrackScaleLevel = Last.mvKeys[i].octave;   // PredictScale
                    P.nObs = i % 6;
                    P.mbBad = i % 23 == 0;
                    cands.push_back(i % 5 == 0 ? nullptr : &P);
                }
                for (int i = 1; i < N1; i += 37) cands.push_back(&pts[i]);  // listed twice: in the keyframe by then, or still free
                g_surgery.clear();
                ORBmatcher matcher(0.6f, true);
                std::vector<MapPoint*> replaced(cands.size(), nullptr);
                int nFused;
                if (pass == 0) {
                    nFused = matcher.Fuse(&KF, cands, th);
                } else {
                    for (auto& c : cands) if (!c) c = &pts[0];  // the Sim3 overload has no null test (:1002): point 0 is bad (0 % 23 == 0)
                    cv::Mat Scw(4, 4, CV_32F);
                    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) Scw.at<float>(r, c) = r == c ? (r < 3 ? 2.f : 1.f) : 0.f;
                    nFused = matcher.Fuse(&KF, Scw, cands, th, replaced);
                }
                put_i(out, nFused);
                put_i(out, (int)g_surgery.size());
                for (const SurgeryOp& o : g_surgery) { put_i(out, o.op); put_i(out, (int)o.a); put_i(out, (int)o.b); }
                for (int j = 0; j < N2; ++j) put_i(out, KF.mvpMapPoints[j] ? (int)KF.mvpMapPoints[j]->mnId : -1);
                put_i(out, (int)cands.size());
                for (size_t e = 0; e < cands.size(); ++e) put_i(out, replaced[e] ? (int)replaced[e]->mnId : -1);
            }
            fclose(out);
        } else if (mode == "bow" && argc >= 8) {
            // System.cc:64-65 loads the vocabulary, Frame::ComputeBoW (src/Frame.cc:394-401) fills mBowVec / mFeatVec of two frames,
            // Tracking::TrackReferenceKeyFrame (:842) matches them with SearchByBoW
            FILE* out = fopen(argv[7], "wb");
            ORBVocabulary voc;
            if (!voc.loadFromTextFile(argv[6])) { fprintf(stderr, "cannot load %s: %s\n", argv[6], orbx_last_error()); return 1; }
            ORBVocabulary none;
            put_i(out, none.loadFromTextFile(std::string(argv[6]) + ".missing") ? 1 : 0);
            put_i(out, (int)voc.size());
            ORBextractor ex(nf, 1.2f, 8, 20, 7, w, h);
            Frame F1, F2;
            F1.mpORBextractorLeft = F2.mpORBextractorLeft = &ex;
            F1.mpORBvocabulary = F2.mpORBvocabulary = &voc;
            F1.ExtractORB(0, im0);
            F2.ExtractORB(0, im1);
            for (Frame* F : {&F1, &F2}) {
                F->mvKeysUn = F->mvKeys;
                F->N = (int)F->mvKeys.size();
                ComputeBoW(*F);
                put_keys(out, F->mvKeys, F->mDescriptors);
                put_i(out, (int)F->mBowVec.size());
                for (const auto& kv : F->mBowVec) { put_i(out, (int)kv.first); put(out, &kv.second, 8); }
                put_i(out, (int)F->mFeatVec.size());
                for (const auto& kv : F->mFeatVec) {
                    put_i(out, (int)kv.first); put_i(out, (int)kv.second.size());
                    put(out, kv.second.data(), kv.second.size() * 4);
                }
            }
            // the vector<cv::Mat> form Frame::ComputeBoW uses (Converter::toDescriptorVector) gives the same maps
            std::vector<cv::Mat> rows;
            for (int i = 0; i < F1.mDescriptors.rows; ++i) rows.push_back(F1.mDescriptors.row(i));
            std::map<unsigned int, double> bow2;
            std::map<unsigned int, std::vector<unsigned int>> fv2;
            voc.transform(rows, bow2, fv2, 4);
            put_i(out, bow2 == F1.mBowVec && fv2 == F1.mFeatVec ? 1 : 0);
            std::vector<MapPoint> points(F1.N);
            KeyFrame KF;
            KF.mvKeysUn = F1.mvKeysUn; KF.mDescriptors = F1.mDescriptors; KF.mFeatVec = F1.mFeatVec;
            KF.mvpMapPoints.assign(F1.N, nullptr);
            for (int i = 0; i < F1.N; ++i) {
                points[i].mbBad = i % 13 == 0;
                if (i % 5 != 0) KF.mvpMapPoints[i] = &points[i];
            }
            std::vector<MapPoint*> matches;
            ORBmatcher matcher(0.7f, true);
            const int n = matcher.SearchByBoW(&KF, F2, matches);
            put_i(out, n);
            for (int j = 0; j < F2.N; ++j) put_i(out, matches[j] ? (int)(matches[j] - points.data()) : -1);
            // LocalMapping::CreateNewMapPoints (:248ff): SearchForTriangulation between the two keyframes.  Camera 2 sits 0.5 m to the
            // right of camera 1 at the same orientation, so the epipole is at infinity on the x axis (here: 1e6-ish) and the
            // epipolar line of (x1, y1) is y = y1 + dy with the fundamental matrix the test passes in argv[8].
            if (argc >= 9) {
                KeyFrame K2;
                K2.mvKeysUn = F2.mvKeysUn; K2.mDescriptors = F2.mDescriptors; K2.mFeatVec = F2.mFeatVec;
                KF.N = F1.N; K2.N = F2.N;
                KF.mvuRight.assign(F1.N, -1.f); K2.mvuRight.assign(F2.N, -1.f);
                for (int i = 0; i < F1.N; i += 3) KF.mvuRight[i] = 1.f;
                for (int j = 0; j < F2.N; j += 4) K2.mvuRight[j] = 1.f;
                K2.mvpMapPoints.assign(F2.N, nullptr);
                for (int j = 0; j < F2.N; j += 6) K2.mvpMapPoints[j] = &points[0];
                K2.mvScaleFactors = ex.GetScaleFactors();
                K2.mvLevelSigma2 = ex.GetScaleSigmaSquares();
                K2.fx = K2.fy = 500.f; K2.cx = 0.5f * w; K2.cy = 0.5f * h;
                auto mat = [](int r, int c, std::initializer_list<float> v) {
                    cv::Mat m(r, c, CV_32F);
                    int i = 0;
                    for (float x : v) { m.at<float>(i / c, i % c) = x; ++i; }
                    return m;
                };
                KF.mOw = mat(3, 1, {0.f, 0.f, 0.f});
                K2.mRcw = mat(3, 3, {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f});
                K2.mtcw = mat(3, 1, {-0.5f, 0.01f, 1e-3f});
                const float dy = (float)atof(argv[8]);
                cv::Mat F12 = mat(3, 3, {0.f, 0.f, 0.f, 0.f, 0.f, -1.f, 0.f, 1.f, -dy});
                std::vector<std::pair<size_t, size_t> > pairs;
                ORBmatcher triMatcher(0.6f, false);
                const int nt = triMatcher.SearchForTriangulation(&KF, &K2, F12, pairs, false);
                put_i(out, nt);
                put_i(out, (int)pairs.size());
                for (const auto& pr : pairs) { put_i(out, (int)pr.first); put_i(out, (int)pr.second); }
            }
            fclose(out);
        } else {
            return 2;
        }
    } catch (const std::exception& e) {
        fprintf(stderr, "frame_harness: %s\n", e.what());
        return 1;
    }
    return 0;
}
