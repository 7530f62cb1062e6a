// oracle/ref_matcher/ref_matcher.cpp — TEST INFRASTRUCTURE.  C entry points that build reference objects from flat arrays and
// call the REFERENCE's own ORBmatcher / Frame member functions (object code of /root/reference/src/ORBmatcher.cc and
// src/Frame.cc, compiled where they lie by the Makefile next to this file).  tests/ compare oracle/orb_oracle.cpp with these,
// live in this container and through tests/golden/ref_matcher.npz elsewhere; nothing in the product links this.
//
// Frame objects are the reference's real class built by its real monocular / stereo constructors (so the 64 x 48 grid, the
// image bounds, GetFeaturesInArea and ComputeStereoMatches are reference code); KeyFrame / MapPoint are the stand-ins of
// ref_stubs.h, filled the way KeyFrame::KeyFrame(Frame&, ...) (src/KeyFrame.cc:41-68) copies a Frame.
#include "ref_stubs.h"

#include "Frame.h"
#include "ORBmatcher.h"

namespace ORB_SLAM2 {

std::vector<SurgeryOp> g_surgery;

std::set<MapPoint*> KeyFrame::GetMapPoints() {
    std::set<MapPoint*> s;
    for (MapPoint* p : mvpMapPoints)
        if (p && !p->isBad()) s.insert(p);
    return s;
}
void KeyFrame::AddMapPoint(MapPoint* pMP, const size_t& idx) {
    g_surgery.push_back({1, pMP->mnId, (long)idx});
    mvpMapPoints[idx] = pMP;
}
std::vector<size_t> KeyFrame::GetFeaturesInArea(const float& x, const float& y, const float& r) const {
    std::vector<size_t> vIndices;
    const int nMinCellX = std::max(0, (int)floor((x - mnMinX - r) * mfGridElementWidthInv));
    if (nMinCellX >= mnGridCols) return vIndices;
    const int nMaxCellX = std::min((int)mnGridCols - 1, (int)ceil((x - mnMinX + r) * mfGridElementWidthInv));
    if (nMaxCellX < 0) return vIndices;
    const int nMinCellY = std::max(0, (int)floor((y - mnMinY - r) * mfGridElementHeightInv));
    if (nMinCellY >= mnGridRows) return vIndices;
    const int nMaxCellY = std::min((int)mnGridRows - 1, (int)ceil((y - mnMinY + r) * mfGridElementHeightInv));
    if (nMaxCellY < 0) return vIndices;
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
            for (size_t j : mGrid[ix][iy]) {
                const cv::KeyPoint& kpUn = mvKeysUn[j];
                if (fabs(kpUn.pt.x - x) < r && fabs(kpUn.pt.y - y) < r) vIndices.push_back(j);
            }
    return vIndices;
}
void MapPoint::addObservationQuiet(KeyFrame* pKF, size_t idx) {
    if (mObservations.count(pKF)) return;
    mObservations[pKF] = idx;
    nObs += pKF->mvuRight[idx] >= 0 ? 2 : 1;
}
void MapPoint::AddObservation(KeyFrame* pKF, size_t idx) {
    g_surgery.push_back({0, mnId, (long)idx});
    addObservationQuiet(pKF, idx);
}
void MapPoint::Replace(MapPoint* pMP) {
    if (pMP->mnId == this->mnId) return;
    g_surgery.push_back({2, mnId, pMP->mnId});
    std::map<KeyFrame*, size_t> obs = mObservations;
    mObservations.clear();
    mbBad = true;
    for (auto& kv : obs) {
        KeyFrame* pKF = kv.first;
        if (!pMP->IsInKeyFrame(pKF)) {
            pKF->ReplaceMapPointMatch(kv.second, pMP);
            pMP->addObservationQuiet(pKF, kv.second);
        } else {
            pKF->EraseMapPointMatch(kv.second);
        }
    }
}
int MapPoint::PredictScale(const float& currentDist, KeyFrame* pKF) {
    if (presetLevel >= 0) return presetLevel;
    const float ratio = mfMaxDistance / currentDist;
    int nScale = ceil(log(ratio) / pKF->mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= pKF->mnScaleLevels) nScale = pKF->mnScaleLevels - 1;
    return nScale;
}
int MapPoint::PredictScale(const float& currentDist, Frame* pF) {
    if (presetLevel >= 0) return presetLevel;
    const float ratio = mfMaxDistance / currentDist;
    int nScale = ceil(log(ratio) / pF->mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= pF->mnScaleLevels) nScale = pF->mnScaleLevels - 1;
    return nScale;
}

}  // namespace ORB_SLAM2

using namespace ORB_SLAM2;

namespace {

struct Cam {  // w, h, fx, fy, cx, cy, bf
    int w, h;
    float fx, fy, cx, cy, bf;
};

cv::Mat desc_mat(const uint8_t* d, int n) {
    cv::Mat m(n, 32, CV_8U);
    if (n) std::memcpy(m.data, d, (size_t)n * 32);
    return m;
}
cv::Mat vec3(const float* p) {
    cv::Mat m(3, 1, CV_32F);
    for (int i = 0; i < 3; ++i) m.at<float>(i) = p[i];
    return m;
}
cv::Mat pose(const float* T) {  // 4 x 4 row-major, NULL = identity
    cv::Mat m = cv::Mat::eye(4, 4, CV_32F);
    if (T)
        for (int i = 0; i < 16; ++i) m.at<float>(i / 4, i % 4) = T[i];
    return m;
}

struct Built {  // a reference Frame with the stand-in extractor that fed its constructor
    ORBextractor ex, exR;
    ORBVocabulary voc;
    std::unique_ptr<Frame> F;
};

void fill_extractor(ORBextractor& ex, const cv::KeyPoint* keys, int n, const uint8_t* desc, const float* sf, int nl) {
    ex.keys.assign(keys, keys + n);
    ex.desc = desc_mat(desc, n);
    ex.sf.assign(sf, sf + nl);
    ex.isf.resize(nl);
    ex.s2.resize(nl);
    ex.is2.resize(nl);
    for (int l = 0; l < nl; ++l) {  // src/ORBextractor.cc:505-517
        ex.s2[l] = sf[l] * sf[l];
        ex.isf[l] = 1.0f / sf[l];
        ex.is2[l] = 1.0f / ex.s2[l];
    }
}

// the reference's monocular constructor (src/Frame.cc:173-227): undistorted camera, image bounds (0, w, 0, h)
std::unique_ptr<Built> make_frame(const cv::KeyPoint* keys, int n, const uint8_t* desc, const Cam& c, const float* sf, int nl,
                                  const float* uRight = nullptr) {
    std::unique_ptr<Built> b(new Built);
    fill_extractor(b->ex, keys, n, desc, sf, nl);
    cv::Mat im(c.h, c.w, CV_8U), K = cv::Mat::eye(3, 3, CV_32F), dist(4, 1, CV_32F);
    K.at<float>(0, 0) = c.fx;
    K.at<float>(1, 1) = c.fy;
    K.at<float>(0, 2) = c.cx;
    K.at<float>(1, 2) = c.cy;
    Frame::mbInitialComputations = true;
    b->F.reset(new Frame(im, 0.0, &b->ex, &b->voc, K, dist, c.bf, 40.f));
    if (n == 0) {  // the constructor returns early on an empty frame: give it the members the matchers read
        b->F->N = 0;
    }
    if (uRight) b->F->mvuRight.assign(uRight, uRight + n);
    return b;
}

// KeyFrame::KeyFrame(Frame&, ...) (src/KeyFrame.cc:41-68): what a keyframe copies from its frame
void fill_keyframe(KeyFrame& kf, Frame& F, long id) {
    kf.mnId = id;
    kf.N = F.N;
    kf.fx = F.fx; kf.fy = F.fy; kf.cx = F.cx; kf.cy = F.cy; kf.mbf = F.mbf;
    kf.mnMinX = F.mnMinX; kf.mnMinY = F.mnMinY; kf.mnMaxX = F.mnMaxX; kf.mnMaxY = F.mnMaxY;
    kf.mnGridCols = FRAME_GRID_COLS; kf.mnGridRows = FRAME_GRID_ROWS;
    kf.mfGridElementWidthInv = F.mfGridElementWidthInv; kf.mfGridElementHeightInv = F.mfGridElementHeightInv;
    kf.mnScaleLevels = F.mnScaleLevels; kf.mfLogScaleFactor = F.mfLogScaleFactor;
    kf.mvKeysUn = F.mvKeysUn; kf.mvuRight = F.mvuRight; kf.mDescriptors = F.mDescriptors.clone();
    kf.mFeatVec = F.mFeatVec;
    kf.mvScaleFactors = F.mvScaleFactors; kf.mvLevelSigma2 = F.mvLevelSigma2; kf.mvInvLevelSigma2 = F.mvInvLevelSigma2;
    kf.mvpMapPoints.assign(F.N, nullptr);
    kf.mGrid.resize(kf.mnGridCols);
    for (int i = 0; i < kf.mnGridCols; i++) {
        kf.mGrid[i].resize(kf.mnGridRows);
        for (int j = 0; j < kf.mnGridRows; j++) kf.mGrid[i][j] = F.mGrid[i][j];
    }
    kf.Rcw = cv::Mat::eye(3, 3, CV_32F);
    kf.tcw = cv::Mat::zeros(3, 1, CV_32F);
    kf.Ow = cv::Mat::zeros(3, 1, CV_32F);
}
void set_pose(KeyFrame& kf, const float* T) {  // KeyFrame::SetPose (src/KeyFrame.cc:79-96)
    cv::Mat Tcw = pose(T);
    kf.Rcw = Tcw.rowRange(0, 3).colRange(0, 3).clone();
    kf.tcw = Tcw.rowRange(0, 3).col(3).clone();
    kf.Ow = -kf.Rcw.t() * kf.tcw;
}
void feat_vec(DBoW2::FeatureVector& fv, int nNodes, const int* start, const uint32_t* idx) {
    fv.clear();
    for (int k = 0; k < nNodes; ++k)
        for (int j = start[k]; j < start[k + 1]; ++j) fv.addFeature((DBoW2::NodeId)k, idx[j]);
}

}  // namespace

extern "C" {

int refm_descriptor_distance(const uint8_t* a, const uint8_t* b) {  // ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663)
    cv::Mat ma = desc_mat(a, 1), mb = desc_mat(b, 1);
    return ORBmatcher::DescriptorDistance(ma, mb);
}

// Frame::GetFeaturesInArea on a reference Frame (src/Frame.cc:326-379, grid by AssignFeaturesToGrid :229-244)
int refm_features_in_area(const cv::KeyPoint* keys, int n, const int* wh, float x, float y, float r, int minLevel, int maxLevel,
                          int* out) {
    const float sf[8] = {1, 1.2f, 1.44f, 1.728f, 2.0736f, 2.48832f, 2.985984f, 3.5831808f};
    std::vector<uint8_t> d((size_t)std::max(n, 1) * 32);
    Cam c{wh[0], wh[1], 500.f, 500.f, wh[0] * 0.5f, wh[1] * 0.5f, 40.f};
    auto b = make_frame(keys, n, d.data(), c, sf, 8);
    const std::vector<size_t> v = b->F->GetFeaturesInArea(x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size(); ++i) out[i] = (int)v[i];
    return (int)v.size();
}

// ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520)
int refm_search_for_initialization(const cv::KeyPoint* k1, int n1, const uint8_t* d1, const cv::KeyPoint* k2, int n2, const uint8_t* d2,
                                   const int* wh, float* prevXY, int* matches12, int windowSize, float nnratio, int checkOri) {
    const float sf[8] = {1, 1.2f, 1.44f, 1.728f, 2.0736f, 2.48832f, 2.985984f, 3.5831808f};
    Cam c{wh[0], wh[1], 500.f, 500.f, wh[0] * 0.5f, wh[1] * 0.5f, 40.f};
    auto f1 = make_frame(k1, n1, d1, c, sf, 8), f2 = make_frame(k2, n2, d2, c, sf, 8);
    std::vector<cv::Point2f> prev(n1);
    for (int i = 0; i < n1; ++i) prev[i] = cv::Point2f(prevXY[2 * i], prevXY[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher m(nnratio, checkOri != 0);
    const int n = m.SearchForInitialization(*f1->F, *f2->F, prev, m12, windowSize);
    for (int i = 0; i < n1; ++i) {
        matches12[i] = m12[i];
        prevXY[2 * i] = prev[i].x;
        prevXY[2 * i + 1] = prev[i].y;
    }
    return n;
}

// ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono) (src/ORBmatcher.cc:1328-1470).
// CurrentFrame.mTcw = identity, so the world positions handed in ARE x3Dc; tlcz = z of LastFrame's translation, which (with
// twc = 0) is tlc(2) and decides forward / backward against CurrentFrame.mb.
int refm_search_by_projection_frame(int n1, const float* x3D, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mpDesc1,
                                    const cv::KeyPoint* k1, const cv::KeyPoint* k2, int n2, const uint8_t* d2, const float* uRight2,
                                    const uint8_t* occupied2, const int* wh, const float* cam, float mbf, const float* sf, int nl,
                                    float th, float tlcz, int bMono, int checkOri, int* match21) {
    Cam c{wh[0], wh[1], cam[0], cam[1], cam[2], cam[3], mbf};
    std::vector<uint8_t> zero((size_t)std::max(n1, 1) * 32);
    auto last = make_frame(k1, n1, zero.data(), c, sf, nl);
    auto cur = make_frame(k2, n2, d2, c, sf, nl, uRight2);
    std::vector<std::unique_ptr<MapPoint>> mps;
    for (int i = 0; i < n1; ++i) {
        if (!valid1[i]) continue;
        mps.emplace_back(new MapPoint);
        MapPoint* p = mps.back().get();
        p->mnId = i;
        p->mWorldPos = vec3(x3D + 3 * i);
        p->mDescriptor = desc_mat(mpDesc1 + (size_t)i * 32, 1);
        p->nObs = obs1[i] ? 1 : 0;
        last->F->mvpMapPoints[i] = p;
    }
    MapPoint occupied;
    occupied.mnId = -100;
    occupied.nObs = 1;
    for (int i = 0; i < n2; ++i)
        if (occupied2[i]) cur->F->mvpMapPoints[i] = &occupied;
    cur->F->SetPose(pose(nullptr));
    float T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, tlcz, 0, 0, 0, 1};
    last->F->SetPose(pose(T));
    ORBmatcher m(0.9f, checkOri != 0);
    const int n = m.SearchByProjection(*cur->F, *last->F, th, bMono != 0);
    for (int i = 0; i < n2; ++i) {
        MapPoint* p = cur->F->mvpMapPoints[i];
        match21[i] = !p || p == &occupied ? -1 : (int)p->mnId;
    }
    return n;
}

// ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:45-129); the per-point fields are
// Tracking's projection cache (mTrackProjX/Y/XR, mnTrackScaleLevel, mTrackViewCos, mbTrackInView)
int refm_search_by_projection_map(int n1, const float* proj, const float* viewCos, const int* level1, const uint8_t* valid1,
                                  const uint8_t* obs1, const uint8_t* mpDesc1, const cv::KeyPoint* k2, int n2, const uint8_t* d2,
                                  const float* uRight2, const uint8_t* occupied2, const int* wh, const float* sf, int nl, float th,
                                  float nnratio, int* match21) {
    Cam c{wh[0], wh[1], 500.f, 500.f, wh[0] * 0.5f, wh[1] * 0.5f, 40.f};
    auto cur = make_frame(k2, n2, d2, c, sf, nl, uRight2);
    std::vector<std::unique_ptr<MapPoint>> mps;
    std::vector<MapPoint*> vp;
    for (int i = 0; i < n1; ++i) {
        mps.emplace_back(new MapPoint);
        MapPoint* p = mps.back().get();
        p->mnId = i;
        p->mbTrackInView = valid1[i] != 0;
        p->mTrackProjX = proj[3 * i];
        p->mTrackProjY = proj[3 * i + 1];
        p->mTrackProjXR = proj[3 * i + 2];
        p->mnTrackScaleLevel = level1[i];
        p->mTrackViewCos = viewCos[i];
        p->mDescriptor = desc_mat(mpDesc1 + (size_t)i * 32, 1);
        p->nObs = obs1[i] ? 1 : 0;
        vp.push_back(p);
    }
    MapPoint occupied;
    occupied.mnId = -100;
    occupied.nObs = 1;
    for (int i = 0; i < n2; ++i)
        if (occupied2[i]) cur->F->mvpMapPoints[i] = &occupied;
    ORBmatcher m(nnratio, true);
    const int n = m.SearchByProjection(*cur->F, vp, th);
    for (int i = 0; i < n2; ++i) {
        MapPoint* p = cur->F->mvpMapPoints[i];
        match21[i] = !p || p == &occupied ? -1 : (int)p->mnId;
    }
    return n;
}

// ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:159-288) over `nNodes` common nodes
int refm_search_by_bow_frame(int nNodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                             const uint8_t* valid1, const cv::KeyPoint* k1, const uint8_t* d1, int n2, const cv::KeyPoint* k2,
                             const uint8_t* d2, const int* wh, float nnratio, int checkOri, int* match21) {
    const float sf[8] = {1, 1.2f, 1.44f, 1.728f, 2.0736f, 2.48832f, 2.985984f, 3.5831808f};
    Cam c{wh[0], wh[1], 500.f, 500.f, wh[0] * 0.5f, wh[1] * 0.5f, 40.f};
    auto fk = make_frame(k1, n1, d1, c, sf, 8), fr = make_frame(k2, n2, d2, c, sf, 8);
    KeyFrame kf;
    fill_keyframe(kf, *fk->F, 1);
    feat_vec(kf.mFeatVec, nNodes, start1, idx1);
    feat_vec(fr->F->mFeatVec, nNodes, start2, idx2);
    std::vector<std::unique_ptr<MapPoint>> mps;
    for (int i = 0; i < n1; ++i)
        if (valid1[i]) {
            mps.emplace_back(new MapPoint);
            mps.back()->mnId = i;
            kf.mvpMapPoints[i] = mps.back().get();
        }
    std::vector<MapPoint*> out;
    ORBmatcher m(nnratio, checkOri != 0);
    const int n = m.SearchByBoW(&kf, *fr->F, out);
    for (int i = 0; i < n2; ++i) match21[i] = out[i] ? (int)out[i]->mnId : -1;
    return n;
}

// ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (src/ORBmatcher.cc:522-655)
int refm_search_by_bow_keyframes(int nNodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                                 const uint8_t* valid1, const cv::KeyPoint* k1, const uint8_t* d1, int n2, const uint8_t* valid2,
                                 const cv::KeyPoint* k2, const uint8_t* d2, const int* wh, float nnratio, int checkOri, int* match12) {
    const float sf[8] = {1, 1.2f, 1.44f, 1.728f, 2.0736f, 2.48832f, 2.985984f, 3.5831808f};
    Cam c{wh[0], wh[1], 500.f, 500.f, wh[0] * 0.5f, wh[1] * 0.5f, 40.f};
    auto f1 = make_frame(k1, n1, d1, c, sf, 8), f2 = make_frame(k2, n2, d2, c, sf, 8);
    KeyFrame kf1, kf2;
    fill_keyframe(kf1, *f1->F, 1);
    fill_keyframe(kf2, *f2->F, 2);
    feat_vec(kf1.mFeatVec, nNodes, start1, idx1);
    feat_vec(kf2.mFeatVec, nNodes, start2, idx2);
    std::vector<std::unique_ptr<MapPoint>> mps;
    for (int i = 0; i < n1; ++i)
        if (valid1[i]) {
            mps.emplace_back(new MapPoint);
            mps.back()->mnId = i;
            kf1.mvpMapPoints[i] = mps.back().get();
        }
    for (int i = 0; i < n2; ++i)
        if (valid2[i]) {
            mps.emplace_back(new MapPoint);
            mps.back()->mnId = i;  // index in keyframe 2
            kf2.mvpMapPoints[i] = mps.back().get();
        }
    std::vector<MapPoint*> out;
    ORBmatcher m(nnratio, checkOri != 0);
    const int n = m.SearchByBoW(&kf1, &kf2, out);
    for (int i = 0; i < n1; ++i) match12[i] = out[i] ? (int)out[i]->mnId : -1;
    return n;
}

// ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-823).  Keyframe 2 has the identity pose and keyframe 1's camera
// centre is `Cw`, so the epipole the reference computes (:664-670) is fx * Cw.x / Cw.z + cx, fy * Cw.y / Cw.z + cy.
int refm_search_for_triangulation(int nNodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                                  const cv::KeyPoint* k1, const uint8_t* d1, const uint8_t* valid1, const float* uRight1, int n2,
                                  const cv::KeyPoint* k2, const uint8_t* d2, const uint8_t* valid2, const float* uRight2,
                                  const int* wh, const float* cam, const float* sf, int nl, const float* F12, const float* Cw,
                                  int onlyStereo, int checkOri, int* match12) {
    Cam c{wh[0], wh[1], cam[0], cam[1], cam[2], cam[3], 40.f};
    auto f1 = make_frame(k1, n1, d1, c, sf, nl, uRight1), f2 = make_frame(k2, n2, d2, c, sf, nl, uRight2);
    KeyFrame kf1, kf2;
    fill_keyframe(kf1, *f1->F, 1);
    fill_keyframe(kf2, *f2->F, 2);
    feat_vec(kf1.mFeatVec, nNodes, start1, idx1);
    feat_vec(kf2.mFeatVec, nNodes, start2, idx2);
    kf1.Ow = vec3(Cw);
    MapPoint tracked;  // valid = 0: the feature already has a map point and is skipped (:704-707, :726-729)
    for (int i = 0; i < n1; ++i)
        if (!valid1[i]) kf1.mvpMapPoints[i] = &tracked;
    for (int i = 0; i < n2; ++i)
        if (!valid2[i]) kf2.mvpMapPoints[i] = &tracked;
    cv::Mat F(3, 3, CV_32F);
    for (int i = 0; i < 9; ++i) F.at<float>(i / 3, i % 3) = F12[i];
    std::vector<std::pair<size_t, size_t>> pairs;
    ORBmatcher m(0.6f, checkOri != 0);
    const int n = m.SearchForTriangulation(&kf1, &kf2, F, pairs, onlyStereo != 0);
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    for (auto& pr : pairs) match12[pr.first] = (int)pr.second;
    return n;
}

// ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:290-403) with Scw = identity:
// points are given in the keyframe's camera frame; level = the level PredictScale returns; found1 = already in vpMatched.
int refm_search_by_projection_sim3(int n1, const float* x3D, const float* normal, const float* minmax, const int* level1,
                                   const uint8_t* bad1, const uint8_t* found1, const uint8_t* mpDesc1, const cv::KeyPoint* k2, int n2,
                                   const uint8_t* d2, const uint8_t* occupied2, const int* wh, const float* cam, const float* sf,
                                   int nl, int th, int* match21) {
    Cam c{wh[0], wh[1], cam[0], cam[1], cam[2], cam[3], 40.f};
    auto f = make_frame(k2, n2, d2, c, sf, nl);
    KeyFrame kf;
    fill_keyframe(kf, *f->F, 1);
    std::vector<std::unique_ptr<MapPoint>> mps;
    std::vector<MapPoint*> vp, matched(n2, nullptr);
    MapPoint occupied;
    occupied.mnId = -100;
    for (int i = 0; i < n2; ++i)
        if (occupied2[i]) matched[i] = &occupied;
    std::vector<MapPoint*> extraFound;
    for (int i = 0; i < n1; ++i) {
        mps.emplace_back(new MapPoint);
        MapPoint* p = mps.back().get();
        p->mnId = i;
        p->mWorldPos = vec3(x3D + 3 * i);
        p->mNormalVector = vec3(normal + 3 * i);
        p->mfMinDistance = minmax[2 * i];
        p->mfMaxDistance = minmax[2 * i + 1];
        p->presetLevel = level1[i];
        p->mbBad = bad1[i] != 0;
        p->mDescriptor = desc_mat(mpDesc1 + (size_t)i * 32, 1);
        vp.push_back(p);
        if (found1[i]) matched.push_back(p);  // spAlreadyFound is built from every entry of vpMatched (:306-307)
    }
    ORBmatcher m(0.75f, true);
    const int n = m.SearchByProjection(&kf, pose(nullptr), vp, matched, th);
    for (int i = 0; i < n2; ++i) match21[i] = !matched[i] || matched[i] == &occupied ? -1 : (int)matched[i]->mnId;
    return n;
}

// ORBmatcher::Fuse(KeyFrame*, vpMapPoints, th) (src/ORBmatcher.cc:825-975), keyframe pose = identity.  kfPoint[i2] >= 0: the
// keyframe's feature i2 already holds map point number nPoints + kfPoint[i2]... see tests/test_oracle_ref_matcher.py.
// ops receives the surgery log as (op, a, b) triples: 0 AddObservation(point a, idx b), 1 AddMapPoint(point a, idx b),
// 2 a->Replace(b).  Returns nFused; *nOps = number of triples.
int refm_fuse(int n1, const float* x3D, const float* normal, const float* minmax, const int* level1, const uint8_t* bad1,
              const int* obs1, const uint8_t* mpDesc1, const cv::KeyPoint* k2, int n2, const uint8_t* d2, const float* uRight2,
              const int* kfPointObs /* -1: none, else Observations() of the point the feature holds */, const int* wh,
              const float* cam, float mbf, const float* sf, int nl, float th, long* ops, int opsCap, int* nOps) {
    Cam c{wh[0], wh[1], cam[0], cam[1], cam[2], cam[3], mbf};
    auto f = make_frame(k2, n2, d2, c, sf, nl, uRight2);
    KeyFrame kf;
    fill_keyframe(kf, *f->F, 1);
    std::vector<std::unique_ptr<MapPoint>> mps;
    std::vector<MapPoint*> vp;
    for (int i = 0; i < n1; ++i) {
        mps.emplace_back(new MapPoint);
        MapPoint* p = mps.back().get();
        p->mnId = i;
        p->mWorldPos = vec3(x3D + 3 * i);
        p->mNormalVector = vec3(normal + 3 * i);
        p->mfMinDistance = minmax[2 * i];
        p->mfMaxDistance = minmax[2 * i + 1];
        p->presetLevel = level1[i];
        p->mbBad = bad1[i] != 0;
        p->nObs = obs1[i];
        p->mDescriptor = desc_mat(mpDesc1 + (size_t)i * 32, 1);
        vp.push_back(p);
    }
    for (int i = 0; i < n2; ++i)
        if (kfPointObs[i] >= 0) {  // the keyframe's own map points: ids n1 + i, observed by the keyframe at feature i
            mps.emplace_back(new MapPoint);
            MapPoint* p = mps.back().get();
            p->mnId = n1 + i;
            p->mObservations[&kf] = i;
            p->nObs = kfPointObs[i];
            kf.mvpMapPoints[i] = p;
        }
    g_surgery.clear();
    ORBmatcher m(0.6f, true);
    const int n = m.Fuse(&kf, vp, th);
    *nOps = (int)g_surgery.size();
    for (int i = 0; i < *nOps && i < opsCap; ++i) {
        ops[3 * i] = g_surgery[i].op;
        ops[3 * i + 1] = g_surgery[i].a;
        ops[3 * i + 2] = g_surgery[i].b;
    }
    return n;
}

// ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame*, const set<MapPoint*>& sAlreadyFound, th, ORBdist)
// (src/ORBmatcher.cc:1472-1599, Tracking::Relocalization).  CurrentFrame.mTcw = identity: the positions handed in are x3Dc.
// found1[i]: the keyframe's point i is in sAlreadyFound; level1 = what PredictScale returns; k1 = the keyframe's mvKeysUn (angles).
int refm_search_by_projection_reloc(int n1, const float* x3D, const uint8_t* has1, const uint8_t* bad1, const uint8_t* found1,
                                    const float* minmax, const int* level1, const uint8_t* mpDesc1, const cv::KeyPoint* k1,
                                    const cv::KeyPoint* k2, int n2, const uint8_t* d2, const uint8_t* occupied2, const int* wh,
                                    const float* cam, const float* sf, int nl, float th, int orbDist, int checkOri, int* match21) {
    Cam c{wh[0], wh[1], cam[0], cam[1], cam[2], cam[3], 40.f};
    std::vector<uint8_t> zero((size_t)std::max(n1, 1) * 32);
    auto fk = make_frame(k1, n1, zero.data(), c, sf, nl);
    auto cur = make_frame(k2, n2, d2, c, sf, nl);
    KeyFrame kf;
    fill_keyframe(kf, *fk->F, 1);
    std::vector<std::unique_ptr<MapPoint>> mps;
    std::set<MapPoint*> found;
    for (int i = 0; i < n1; ++i) {
        if (!has1[i]) continue;
        mps.emplace_back(new MapPoint);
        MapPoint* p = mps.back().get();
        p->mnId = i;
        p->mWorldPos = vec3(x3D + 3 * i);
        p->mDescriptor = desc_mat(mpDesc1 + (size_t)i * 32, 1);
        p->mbBad = bad1[i] != 0;
        p->mfMinDistance = minmax[2 * i];
        p->mfMaxDistance = minmax[2 * i + 1];
        p->presetLevel = level1[i];
        kf.mvpMapPoints[i] = p;
        if (found1[i]) found.insert(p);
    }
    MapPoint occupied;
    occupied.mnId = -100;
    for (int i = 0; i < n2; ++i)
        if (occupied2[i]) cur->F->mvpMapPoints[i] = &occupied;
    cur->F->SetPose(pose(nullptr));
    ORBmatcher m(0.9f, checkOri != 0);
    const int n = m.SearchByProjection(*cur->F, &kf, found, th, orbDist);
    for (int i = 0; i < n2; ++i) {
        MapPoint* p = cur->F->mvpMapPoints[i];
        match21[i] = !p || p == &occupied ? -1 : (int)p->mnId;
    }
    return n;
}

// ORBmatcher::Fuse(KeyFrame*, cv::Mat Scw, vpPoints, th, vpReplacePoint) (src/ORBmatcher.cc:977-1100); Scw = scale * [I | 0].
// replaced[e] = id of the keyframe point reported for list entry e (-1: none); keyframe points are numbered n1 + feature index.
int refm_fuse_sim3(int n1, const float* x3D, const float* normal, const float* minmax, const int* level1, const uint8_t* bad1,
                   const uint8_t* mpDesc1, int nEntries, const int* entryPoint, const cv::KeyPoint* k2, int n2, const uint8_t* d2,
                   const int* kfPointObs, const uint8_t* kfBad, const int* wh, const float* cam, const float* sf, int nl, float scale,
                   float th, long* ops, int opsCap, int* nOps, int* replaced) {
    Cam c{wh[0], wh[1], cam[0], cam[1], cam[2], cam[3], 40.f};
    auto f = make_frame(k2, n2, d2, c, sf, nl);
    KeyFrame kf;
    fill_keyframe(kf, *f->F, 1);
    std::vector<std::unique_ptr<MapPoint>> mps;
    std::vector<MapPoint*> pts(n1);
    for (int i = 0; i < n1; ++i) {
        mps.emplace_back(new MapPoint);
        MapPoint* p = mps.back().get();
        p->mnId = i;
        p->mWorldPos = vec3(x3D + 3 * i);
        p->mNormalVector = vec3(normal + 3 * i);
        p->mfMinDistance = minmax[2 * i];
        p->mfMaxDistance = minmax[2 * i + 1];
        p->presetLevel = level1[i];
        p->mbBad = bad1[i] != 0;
        p->mDescriptor = desc_mat(mpDesc1 + (size_t)i * 32, 1);
        pts[i] = p;
    }
    for (int i = 0; i < n2; ++i)
        if (kfPointObs[i] >= 0) {
            mps.emplace_back(new MapPoint);
            MapPoint* p = mps.back().get();
            p->mnId = n1 + i;
            p->mObservations[&kf] = i;
            p->nObs = kfPointObs[i];
            p->mbBad = kfBad[i] != 0;
            kf.mvpMapPoints[i] = p;
        }
    std::vector<MapPoint*> vp(nEntries), rep(nEntries, nullptr);
    for (int e = 0; e < nEntries; ++e) vp[e] = pts[entryPoint[e]];
    float T[16] = {scale, 0, 0, 0, 0, scale, 0, 0, 0, 0, scale, 0, 0, 0, 0, 1};
    g_surgery.clear();
    ORBmatcher m(0.6f, true);
    const int n = m.Fuse(&kf, pose(T), vp, th, rep);
    *nOps = (int)g_surgery.size();
    for (int i = 0; i < *nOps && i < opsCap; ++i) {
        ops[3 * i] = g_surgery[i].op;
        ops[3 * i + 1] = g_surgery[i].a;
        ops[3 * i + 2] = g_surgery[i].b;
    }
    for (int e = 0; e < nEntries; ++e) replaced[e] = rep[e] ? (int)rep[e]->mnId : -1;
    return n;
}

// ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1102-1326): both keyframes at the identity pose, Sim3 = (s12, I, t12); the points
// of keyframe k sit at x3D_k in its camera frame; matched12[i] >= 0: vpMatches12[i] is set on entry to keyframe-2 point that
// index (already matched).  out12[i] = index of the keyframe-2 point vpMatches12[i] holds afterwards, -1 = null.
int refm_search_by_sim3(int n1, const cv::KeyPoint* k1, const uint8_t* d1, const uint8_t* has1, const uint8_t* bad1, const float* x3D1,
                        const int* level1, int n2, const cv::KeyPoint* k2, const uint8_t* d2, const uint8_t* has2, const uint8_t* bad2,
                        const float* x3D2, const int* level2, const float* minmax, const int* matched12, const int* wh, const float* cam,
                        const float* sf, int nl, float s12, const float* t12, float th, int* out12) {
    Cam c{wh[0], wh[1], cam[0], cam[1], cam[2], cam[3], 40.f};
    auto f1 = make_frame(k1, n1, d1, c, sf, nl), f2 = make_frame(k2, n2, d2, c, sf, nl);
    KeyFrame kf1, kf2;
    fill_keyframe(kf1, *f1->F, 1);
    fill_keyframe(kf2, *f2->F, 2);
    std::vector<std::unique_ptr<MapPoint>> mps;
    auto fill = [&](KeyFrame& kf, int n, const uint8_t* has, const uint8_t* bad, const float* x3D, const int* level, const uint8_t* d) {
        for (int i = 0; i < n; ++i) {
            if (!has[i]) continue;
            mps.emplace_back(new MapPoint);
            MapPoint* p = mps.back().get();
            p->mnId = i;
            p->mWorldPos = vec3(x3D + 3 * i);
            p->mDescriptor = desc_mat(d + (size_t)i * 32, 1);
            p->mbBad = bad[i] != 0;
            p->mfMinDistance = minmax[0];
            p->mfMaxDistance = minmax[1];
            p->presetLevel = level[i];
            p->mObservations[&kf] = i;
            kf.mvpMapPoints[i] = p;
        }
    };
    fill(kf1, n1, has1, bad1, x3D1, level1, d1);
    fill(kf2, n2, has2, bad2, x3D2, level2, d2);
    std::vector<MapPoint*> m12(n1, nullptr);
    for (int i = 0; i < n1; ++i)
        if (matched12[i] >= 0) m12[i] = kf2.mvpMapPoints[matched12[i]];
    cv::Mat R12 = cv::Mat::eye(3, 3, CV_32F), t = vec3(t12);
    ORBmatcher m(0.75f, true);
    const int n = m.SearchBySim3(&kf1, &kf2, m12, s12, R12, t, th);
    for (int i = 0; i < n1; ++i) out12[i] = m12[i] ? (int)m12[i]->mnId : -1;
    return n;
}

// Frame::ComputeStereoMatches (src/Frame.cc:465-639) through the reference's stereo constructor (:61-116): the stand-in
// extractors return the given keypoints / descriptors and hold the given pyramid levels (interior, tightly packed).
int refm_stereo_matches(const cv::KeyPoint* kL, int nL, const uint8_t* dL, const cv::KeyPoint* kR, int nR, const uint8_t* dR,
                        const uint8_t* const* levelsL, const uint8_t* const* levelsR, const int* levelWH, const int* wh,
                        const float* cam, float bf, const float* sf, int nl, float* uRight, float* depth) {
    std::unique_ptr<Built> b(new Built);
    fill_extractor(b->ex, kL, nL, dL, sf, nl);
    fill_extractor(b->exR, kR, nR, dR, sf, nl);
    for (int l = 0; l < nl; ++l) {
        const int w = levelWH[2 * l], h = levelWH[2 * l + 1];
        cv::Mat a(h, w, CV_8U), r(h, w, CV_8U);
        std::memcpy(a.data, levelsL[l], (size_t)w * h);
        std::memcpy(r.data, levelsR[l], (size_t)w * h);
        b->ex.mvImagePyramid.push_back(a);
        b->exR.mvImagePyramid.push_back(r);
    }
    cv::Mat im(wh[1], wh[0], CV_8U), K = cv::Mat::eye(3, 3, CV_32F), dist(4, 1, CV_32F);
    K.at<float>(0, 0) = cam[0];
    K.at<float>(1, 1) = cam[1];
    K.at<float>(0, 2) = cam[2];
    K.at<float>(1, 2) = cam[3];
    Frame::mbInitialComputations = true;
    // fx is a static the constructor sets AFTER ComputeStereoMatches on the first frame (:89 vs :103); the reference always
    // has built frames before (System initialises with one), so give the statics their values first
    Frame::fx = cam[0];
    Frame::fy = cam[1];
    Frame::cx = cam[2];
    Frame::cy = cam[3];
    // ComputeStereoMatches reads mb (minZ, :498) BEFORE the constructor assigns it (:112): a real run sees what the previous
    // Frame left at that address, i.e. mbf / fx of the frame before (Tracking rebuilds mCurrentFrame in place).  The harness
    // seeds the object's storage with that value, which the constructor's initialiser list leaves alone.
    void* raw = ::operator new(sizeof(Frame));
    std::memset(raw, 0, sizeof(Frame));
    {
        const float seed = bf / cam[0];
        std::memcpy((char*)raw + ((char*)&((Frame*)raw)->mb - (char*)raw), &seed, sizeof seed);
    }
    b->F.reset(new (raw) Frame(im, im, 0.0, &b->ex, &b->exR, &b->voc, K, dist, bf, 40.f));
    for (int i = 0; i < nL; ++i) {
        uRight[i] = b->F->mvuRight[i];
        depth[i] = b->F->mvDepth[i];
    }
    return nL;
}

}  // extern "C"
