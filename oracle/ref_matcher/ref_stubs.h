// oracle/ref_matcher/ref_stubs.h — TEST INFRASTRUCTURE, force-included (-include) ahead of the reference sources.
//
// The reference's src/ORBmatcher.cc and src/Frame.cc are compiled where they lie (Makefile) against the reference's REAL
// include/ORBmatcher.h and include/Frame.h.  The headers those pull in for the rest of the SLAM system (MapPoint.h,
// KeyFrame.h, ORBextractor.h with its Jetson-only NVX includes, ORBVocabulary.h, Converter.h with g2o / Eigen) are switched
// off by pre-defining their include guards; this file supplies stand-ins with exactly the members the two sources touch.
// Stand-in bodies restate the few reference lines they replace (cited); they hold no matcher logic.
#ifndef ORBX_REF_STUBS_H
#define ORBX_REF_STUBS_H
#define MAPPOINT_H
#define KEYFRAME_H
#define ORBVOCABULARY_H
#define CONVERTER_H
#include <opencv2/core/core.hpp>
#ifdef ORBX_REF_USE_PRODUCT_CLASSES
// ref_frame_unchanged: the reference's Frame.cc on top of THIS repository's ORBextractor / ORBmatcher classes (their headers use
// the same include guards as the reference's, so the reference's own copies are skipped wherever they are included later)
#include "ORBextractor.h"
#include "ORBmatcher.h"
#else
#define ORBEXTRACTOR_H
#endif

#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"

namespace ORB_SLAM2 {
using namespace std;  // the reference headers rely on `using namespace std` leaking from the ones replaced here

class Frame;
class KeyFrame;
class MapPoint;

// what a test wants to know about the map surgery of Fuse (src/ORBmatcher.cc:947-970, :1077-1095)
struct SurgeryOp {
    int op;  // 0 = pMP->AddObservation(pKF, idx), 1 = pKF->AddMapPoint(pMP, idx), 2 = a->Replace(b)
    long a, b;
};
extern std::vector<SurgeryOp> g_surgery;

#ifndef ORBX_REF_USE_PRODUCT_CLASSES
class ORBextractor {  // include/ORBextractor.h: the accessors src/Frame.cc:68-74 reads, operator() returning preset features
public:
    std::vector<cv::KeyPoint> keys;
    cv::Mat desc;
    std::vector<cv::Mat> mvImagePyramid;
    std::vector<float> sf, isf, s2, is2;
    void operator()(const cv::Mat&, const cv::Mat&, std::vector<cv::KeyPoint>& k, cv::Mat& d) {
        k = keys;
        d = desc.clone();
    }
    int GetLevels() { return (int)sf.size(); }
    float GetScaleFactor() { return sf.size() > 1 ? sf[1] : 1.f; }
    std::vector<float> GetScaleFactors() { return sf; }
    std::vector<float> GetInverseScaleFactors() { return isf; }
    std::vector<float> GetScaleSigmaSquares() { return s2; }
    std::vector<float> GetInverseScaleSigmaSquares() { return is2; }
};
#endif

class ORBVocabulary {  // Frame::ComputeBoW is not exercised here (the DBoW2 descent is pinned by oracle/ref_dbow2)
public:
    void transform(const std::vector<cv::Mat>&, DBoW2::BowVector&, DBoW2::FeatureVector&, int) {}
};

class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& d) {  // src/Converter.cc:29-37
        std::vector<cv::Mat> v;
        for (int j = 0; j < d.rows; ++j) v.push_back(d.row(j));
        return v;
    }
};

class KeyFrame {  // include/KeyFrame.h: members read by src/ORBmatcher.cc
public:
    long mnId = 0;
    int N = 0;
    float fx = 0, fy = 0, cx = 0, cy = 0, mbf = 0;
    int mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;
    int mnGridCols = 64, mnGridRows = 48;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    int mnScaleLevels = 0;
    float mfLogScaleFactor = 0;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<float> mvuRight;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<std::vector<std::vector<size_t>>> mGrid;
    cv::Mat Rcw, tcw, Ow;

    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }  // src/KeyFrame.cc:291-295
    std::set<MapPoint*> GetMapPoints();                                     // src/KeyFrame.cc:236-249
    void AddMapPoint(MapPoint* pMP, const size_t& idx);                     // src/KeyFrame.cc:209-213
    void EraseMapPointMatch(const size_t& idx) { mvpMapPoints[idx] = nullptr; }
    void ReplaceMapPointMatch(const size_t& idx, MapPoint* pMP) { mvpMapPoints[idx] = pMP; }
    cv::Mat GetRotation() { return Rcw.clone(); }
    cv::Mat GetTranslation() { return tcw.clone(); }
    cv::Mat GetCameraCenter() { return Ow.clone(); }
    bool IsInImage(const float& x, const float& y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); }  // :619-622
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r) const;  // src/KeyFrame.cc:576-617
    void AssignFeaturesToGrid();  // the grid a KeyFrame copies from its Frame (src/Frame.cc:229-244 via src/KeyFrame.cc:49-55)
};

class MapPoint {  // include/MapPoint.h: members read by src/ORBmatcher.cc and Frame::isInFrustum
public:
    long mnId = 0;
    cv::Mat mWorldPos, mNormalVector, mDescriptor;
    float mfMinDistance = 0, mfMaxDistance = 0;
    int presetLevel = -1;  // >= 0: PredictScale returns it (the test dictates the level); else the reference formula
    bool mbBad = false;
    int nObs = 0;
    std::map<KeyFrame*, size_t> mObservations;
    // Tracking's per-frame projection cache (include/MapPoint.h:97-104)
    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0;
    bool mbTrackInView = false;
    int mnTrackScaleLevel = 0;
    float mTrackViewCos = 0;
    long unsigned int mnTrackReferenceForFrame = 0, mnLastFrameSeen = 0;

    bool isBad() { return mbBad; }
    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    cv::Mat GetNormal() { return mNormalVector.clone(); }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }  // src/MapPoint.cc:380-384
    float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }  // :386-390
    int Observations() { return nObs; }
    bool IsInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF); }
    int GetIndexInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) ? (int)mObservations[pKF] : -1; }
    void AddObservation(KeyFrame* pKF, size_t idx);  // src/MapPoint.cc:83-95
    void Replace(MapPoint* pMP);                     // src/MapPoint.cc:179-220 (observation transfer only)
    int PredictScale(const float& currentDist, KeyFrame* pKF);  // src/MapPoint.cc:390-405
    int PredictScale(const float& currentDist, Frame* pF);      // :407-422
    void addObservationQuiet(KeyFrame* pKF, size_t idx);
};

}  // namespace ORB_SLAM2
#endif
