// include/ORBmatcher.h — the hot-path part of ORB_SLAM2::ORBmatcher (reference include/ORBmatcher.h:37-102,
// src/ORBmatcher.cc) and Frame::ComputeStereoMatches (src/Frame.cc:465-639) over the C ABI of include/orbx.h.
//
// Kept from the reference surface: ORBmatcher(nnratio, checkOri), the static DescriptorDistance(const cv::Mat&,
// const cv::Mat&), TH_LOW / TH_HIGH / HISTO_LENGTH, and SearchForInitialization with its reference signature.
// SearchForInitialization and ComputeStereoMatches are templates over the frame type so that they bind to the
// reference's ORB_SLAM2::Frame (its public members mvKeysUn, mDescriptors, mnMinX ... are used exactly as
// src/ORBmatcher.cc does) without this repository carrying Frame.h.  SearchByProjection (all four overloads) and SearchByBoW
// (both overloads), SearchForTriangulation, SearchBySim3 and Fuse (both overloads) bind the same way to Frame / KeyFrame / MapPoint.
// In Fuse the projection tests and the candidate loops of ALL map points run first (one orbx_fuse_match call); the map surgery
// (Replace / AddObservation / AddMapPoint: pointer work whose order matters) then runs point by point in the reference's order.
//
// Every call here runs on the device; there is no CPU implementation behind it (errors throw std::runtime_error).
// DescriptorDistance on ONE pair is a kernel launch plus two small copies: callers with many pairs should use
// DescriptorDistances (row-wise, one launch) or BruteForceMatch.
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <cmath>
#include <cstdint>
#include <cstring>
#include <set>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "ORBextractor.h"
#include "orbx.h"
#include "orbx_cv.h"

namespace ORB_SLAM2 {

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

    // Hamming distance between two ORB descriptors (rows of a CV_8U N x 32 matrix); src/ORBmatcher.cc:1647-1663
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
        int d = 0;
        check(orbx_hamming_distance(a.ptr<unsigned char>(), b.ptr<unsigned char>(), 1, &d), "DescriptorDistance");
        return d;
    }

    // Row-wise distances of two continuous N x 32 descriptor tables in one launch.
    static std::vector<int> DescriptorDistances(const cv::Mat& A, const cv::Mat& B) {
        if (A.rows != B.rows || A.cols != 32 || B.cols != 32 || !A.isContinuous() || !B.isContinuous())
            throw std::runtime_error("DescriptorDistances: two continuous N x 32 CV_8U tables expected");
        std::vector<int> d(A.rows);
        check(orbx_hamming_distance(A.ptr<unsigned char>(), B.ptr<unsigned char>(), A.rows, d.data()), "DescriptorDistances");
        return d;
    }

    // Best / second-best of every row of Q in D with the acceptance test of the reference's inner loops
    // (src/ORBmatcher.cc:187-232): vnMatches[i] = index into D, or -1.  Returns the number of matches.
    int BruteForceMatch(const cv::Mat& Q, const cv::Mat& D, std::vector<int>& vnMatches, int th = 50 /* TH_LOW */,
                        std::vector<int>* bestDist = nullptr, std::vector<int>* secondDist = nullptr) const {
        if (Q.cols != 32 || (D.rows > 0 && D.cols != 32) || !Q.isContinuous() || !D.isContinuous())
            throw std::runtime_error("BruteForceMatch: continuous N x 32 CV_8U tables expected");
        vnMatches.assign(Q.rows, -1);
        std::vector<int> idx(Q.rows), d1(Q.rows), d2(Q.rows);
        check(orbx_match_bruteforce(Q.ptr<unsigned char>(), Q.rows, D.ptr<unsigned char>(), D.rows, th, mfNNratio, idx.data(),
                                    d1.data(), d2.data(), vnMatches.data()),
              "BruteForceMatch");
        int n = 0;
        for (int m : vnMatches) n += m >= 0;
        if (bestDist) bestDist->swap(d1);
        if (secondDist) secondDist->swap(d2);
        return n;
    }

    // Matching for the Map Initialization (only used in the monocular case); src/ORBmatcher.cc:405-520.
    // FrameT needs: mvKeysUn (std::vector<cv::KeyPoint>), mDescriptors (cv::Mat N x 32 CV_8U) and the static image
    // bounds mnMinX, mnMaxX, mnMinY, mnMaxY — the members of ORB_SLAM2::Frame the reference routine reads.
    template <class FrameT>
    int SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                int windowSize = 10) {
        const int n1 = (int)F1.mvKeysUn.size(), n2 = (int)F2.mvKeysUn.size();
        vnMatches12 = std::vector<int>(n1, -1);
        if (n1 == 0) return 0;
        if ((int)vbPrevMatched.size() != n1) throw std::runtime_error("SearchForInitialization: vbPrevMatched size");
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint layout");
        static_assert(sizeof(cv::Point2f) == 8, "cv::Point2f layout");
        const cv::Mat d1 = continuous(F1.mDescriptors), d2 = continuous(F2.mDescriptors);
        const float bounds[4] = {(float)FrameT::mnMinX, (float)FrameT::mnMaxX, (float)FrameT::mnMinY, (float)FrameT::mnMaxY};
        int nmatches = 0;
        check(orbx_search_for_initialization(reinterpret_cast<const orbx_keypoint*>(F1.mvKeysUn.data()), n1,
                                             d1.ptr<unsigned char>(), reinterpret_cast<const orbx_keypoint*>(F2.mvKeysUn.data()),
                                             n2, d2.ptr<unsigned char>(), bounds, reinterpret_cast<float*>(vbPrevMatched.data()),
                                             vnMatches12.data(), windowSize, mfNNratio, mbCheckOrientation ? 1 : 0, &nmatches),
              "SearchForInitialization");
        return nmatches;
    }

    // Project MapPoints tracked in last frame into the current frame and search matches; src/ORBmatcher.cc:1328-1470 (used to
    // track from previous frame, src/Tracking.cc:962,968).  FrameT / MapPoint members read: mTcw (4x4 CV_32F), mb, mbf, fx, fy, cx,
    // cy, N, mvpMapPoints, mvbOutlier, mvKeys, mvKeysUn, mDescriptors, mvuRight, mvScaleFactors, mnMinX..mnMaxY;
    // pMP->GetWorldPos(), GetDescriptor(), Observations() — exactly what the reference routine reads.  The pose arithmetic stays
    // on the host (with OpenCV: the reference's own cv::Mat expressions); window search, Hamming and the assignment run on the GPU.
    template <class FrameT>
    int SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool bMono) {
        const int n1 = LastFrame.N, n2 = (int)CurrentFrame.mvKeysUn.size();
        if (n1 == 0 || n2 == 0) return 0;
        float R[9], t[3], Rl[9], tl[3];
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) {
                R[3 * r + c] = CurrentFrame.mTcw.template at<float>(r, c);
                Rl[3 * r + c] = LastFrame.mTcw.template at<float>(r, c);
            }
            t[r] = CurrentFrame.mTcw.template at<float>(r, 3);
            tl[r] = LastFrame.mTcw.template at<float>(r, 3);
        }
#if ORBX_HAVE_OPENCV
        const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3), tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
        const cv::Mat twc = -Rcw.t() * tcw;
        const cv::Mat tlc = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3) * twc + LastFrame.mTcw.rowRange(0, 3).col(3);
        const float tlcz = tlc.at<float>(2);
#else  // stand-in Mat without arithmetic: the same expressions in plain float
        float twc[3];
        for (int c = 0; c < 3; ++c) twc[c] = -(R[c] * t[0] + R[3 + c] * t[1] + R[6 + c] * t[2]);
        const float tlcz = Rl[6] * twc[0] + Rl[7] * twc[1] + Rl[8] * twc[2] + tl[2];
#endif
        const bool bForward = tlcz > CurrentFrame.mb && !bMono, bBackward = -tlcz > CurrentFrame.mb && !bMono;  // :1348-1349
        std::vector<float> x3(3 * (size_t)n1, 0.f), angle1(n1, 0.f);
        std::vector<unsigned char> valid1(n1, 0), obs1(n1, 0), mpDesc((size_t)n1 * 32, 0), occupied2(n2, 0);
        std::vector<int> octave1(n1, 0);
        for (int i = 0; i < n1; ++i) {
            auto* pMP = LastFrame.mvpMapPoints[i];
            if (!pMP || LastFrame.mvbOutlier[i]) continue;
            valid1[i] = 1;
            obs1[i] = pMP->Observations() > 0;
            const cv::Mat x3Dw = pMP->GetWorldPos();
#if ORBX_HAVE_OPENCV
            const cv::Mat x3Dc = Rcw * x3Dw + tcw;  // :1362-1363
            for (int c = 0; c < 3; ++c) x3[3 * i + c] = x3Dc.at<float>(c);
#else
            for (int r = 0; r < 3; ++r)
                x3[3 * i + r] = R[3 * r] * x3Dw.template at<float>(0, 0) + R[3 * r + 1] * x3Dw.template at<float>(1, 0) +
                                R[3 * r + 2] * x3Dw.template at<float>(2, 0) + t[r];
#endif
            const cv::Mat dMP = pMP->GetDescriptor();
            std::memcpy(&mpDesc[(size_t)i * 32], dMP.template ptr<unsigned char>(), 32);
            octave1[i] = LastFrame.mvKeys[i].octave;
            angle1[i] = LastFrame.mvKeysUn[i].angle;
        }
        for (int j = 0; j < n2; ++j) occupied2[j] = CurrentFrame.mvpMapPoints[j] && CurrentFrame.mvpMapPoints[j]->Observations() > 0;
        const cv::Mat d2 = continuous(CurrentFrame.mDescriptors);
        const float cam[4] = {CurrentFrame.fx, CurrentFrame.fy, CurrentFrame.cx, CurrentFrame.cy};
        const float bounds[4] = {(float)FrameT::mnMinX, (float)FrameT::mnMaxX, (float)FrameT::mnMinY, (float)FrameT::mnMaxY};
        const bool hasRight = (int)CurrentFrame.mvuRight.size() == n2;
        std::vector<int> match21(n2, -1);
        int nmatches = 0;
        check(orbx_search_by_projection(n1, x3.data(), valid1.data(), obs1.data(), mpDesc.data(), octave1.data(), angle1.data(),
                                        reinterpret_cast<const orbx_keypoint*>(CurrentFrame.mvKeysUn.data()), n2,
                                        d2.template ptr<unsigned char>(), hasRight ? CurrentFrame.mvuRight.data() : nullptr,
                                        occupied2.data(), cam, CurrentFrame.mbf, bounds, CurrentFrame.mvScaleFactors.data(),
                                        (int)CurrentFrame.mvScaleFactors.size(), th, bForward ? 1 : (bBackward ? 2 : 0),
                                        mbCheckOrientation ? 1 : 0, match21.data(), &nmatches),
              "SearchByProjection");
        for (int j = 0; j < n2; ++j) {
            if (match21[j] >= 0) CurrentFrame.mvpMapPoints[j] = LastFrame.mvpMapPoints[match21[j]];
            else if (match21[j] == -2) CurrentFrame.mvpMapPoints[j] = nullptr;
        }
        return nmatches;
    }

    // Search matches between Frame keypoints and projected MapPoints (Tracking::SearchLocalPoints, src/Tracking.cc:1270-1277);
    // src/ORBmatcher.cc:45-129.  MapPoint members read: mbTrackInView, isBad(), mnTrackScaleLevel, mTrackViewCos, mTrackProjX,
    // mTrackProjY, mTrackProjXR, GetDescriptor(), Observations().
    template <class FrameT, class MapPointT>
    int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3) {
        const int n1 = (int)vpMapPoints.size(), n2 = (int)F.mvKeysUn.size();
        if (n1 == 0 || n2 == 0) return 0;
        std::vector<float> proj(3 * (size_t)n1, 0.f), viewCos(n1, 0.f);
        std::vector<int> level1(n1, 0);
        std::vector<unsigned char> valid1(n1, 0), obs1(n1, 0), mpDesc((size_t)n1 * 32, 0), occupied2(n2, 0);
        for (int i = 0; i < n1; ++i) {
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP->mbTrackInView || pMP->isBad()) continue;
            valid1[i] = 1;
            obs1[i] = pMP->Observations() > 0;
            proj[3 * i] = pMP->mTrackProjX;
            proj[3 * i + 1] = pMP->mTrackProjY;
            proj[3 * i + 2] = pMP->mTrackProjXR;
            viewCos[i] = pMP->mTrackViewCos;
            level1[i] = pMP->mnTrackScaleLevel;
            const cv::Mat d = pMP->GetDescriptor();
            std::memcpy(&mpDesc[(size_t)i * 32], d.template ptr<unsigned char>(), 32);
        }
        for (int j = 0; j < n2; ++j) occupied2[j] = F.mvpMapPoints[j] && F.mvpMapPoints[j]->Observations() > 0;
        const cv::Mat d2 = continuous(F.mDescriptors);
        const float bounds[4] = {(float)FrameT::mnMinX, (float)FrameT::mnMaxX, (float)FrameT::mnMinY, (float)FrameT::mnMaxY};
        const bool hasRight = (int)F.mvuRight.size() == n2;
        std::vector<int> match21(n2, -1);
        int nmatches = 0;
        check(orbx_search_by_projection_map(n1, proj.data(), viewCos.data(), level1.data(), valid1.data(), obs1.data(), mpDesc.data(),
                                            reinterpret_cast<const orbx_keypoint*>(F.mvKeysUn.data()), n2,
                                            d2.template ptr<unsigned char>(), hasRight ? F.mvuRight.data() : nullptr,
                                            occupied2.data(), bounds, F.mvScaleFactors.data(), (int)F.mvScaleFactors.size(), th,
                                            mfNNratio, match21.data(), &nmatches),
              "SearchByProjection");
        for (int j = 0; j < n2; ++j)
            if (match21[j] >= 0) F.mvpMapPoints[j] = vpMapPoints[match21[j]];
        return nmatches;
    }

    // Project MapPoints seen in KeyFrame into the Frame and search matches (relocalisation, src/Tracking.cc:1445,1484);
    // src/ORBmatcher.cc:1472-1599.  KeyFrame members read: GetMapPointMatches(), mvKeysUn; MapPoint: isBad(), GetWorldPos(),
    // GetMinDistanceInvariance(), GetMaxDistanceInvariance(), PredictScale(dist, &Frame), GetDescriptor().
    template <class FrameT, class KeyFrameT, class MapPointT>
    int SearchByProjection(FrameT& CurrentFrame, KeyFrameT* pKF, const std::set<MapPointT*>& sAlreadyFound, const float th,
                           const int ORBdist) {
        const std::vector<MapPointT*> vpMPs = pKF->GetMapPointMatches();
        const int n1 = (int)vpMPs.size(), n2 = (int)CurrentFrame.mvKeysUn.size();
        if (n1 == 0 || n2 == 0) return 0;
#if ORBX_HAVE_OPENCV
        const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3), tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
        const cv::Mat Ow = -Rcw.t() * tcw;
#else
        float R[9], t[3], Ow[3];
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) R[3 * r + c] = CurrentFrame.mTcw.template at<float>(r, c);
            t[r] = CurrentFrame.mTcw.template at<float>(r, 3);
        }
        for (int c = 0; c < 3; ++c) Ow[c] = -(R[c] * t[0] + R[3 + c] * t[1] + R[6 + c] * t[2]);
#endif
        std::vector<float> x3(3 * (size_t)n1, 0.f), angle1(n1, 0.f);
        std::vector<unsigned char> valid1(n1, 0), obs1(n1, 1), mpDesc((size_t)n1 * 32, 0), occupied2(n2, 0);
        std::vector<int> level1(n1, 0);
        for (int i = 0; i < n1; ++i) {
            MapPointT* pMP = vpMPs[i];
            if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
            const cv::Mat x3Dw = pMP->GetWorldPos();
#if ORBX_HAVE_OPENCV
            const cv::Mat x3Dc = Rcw * x3Dw + tcw;  // :1494-1495
            for (int c = 0; c < 3; ++c) x3[3 * i + c] = x3Dc.at<float>(c);
            const cv::Mat PO = x3Dw - Ow;
            const float dist3D = (float)cv::norm(PO);  // :1510-1511
#else
            double ss = 0;
            for (int r = 0; r < 3; ++r) {
                x3[3 * i + r] = R[3 * r] * x3Dw.template at<float>(0, 0) + R[3 * r + 1] * x3Dw.template at<float>(1, 0) +
                                R[3 * r + 2] * x3Dw.template at<float>(2, 0) + t[r];
                const float po = x3Dw.template at<float>(r, 0) - Ow[r];
                ss += (double)po * po;
            }
            const float dist3D = (float)std::sqrt(ss);
#endif
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;  // :1517-1518
            valid1[i] = 1;
            level1[i] = pMP->PredictScale(dist3D, &CurrentFrame);
            const cv::Mat dMP = pMP->GetDescriptor();
            std::memcpy(&mpDesc[(size_t)i * 32], dMP.template ptr<unsigned char>(), 32);
            angle1[i] = pKF->mvKeysUn[i].angle;
        }
        for (int j = 0; j < n2; ++j) occupied2[j] = CurrentFrame.mvpMapPoints[j] != nullptr;  // :1540-1541
        const cv::Mat d2 = continuous(CurrentFrame.mDescriptors);
        const float cam[4] = {CurrentFrame.fx, CurrentFrame.fy, CurrentFrame.cx, CurrentFrame.cy};
        const float bounds[4] = {(float)FrameT::mnMinX, (float)FrameT::mnMaxX, (float)FrameT::mnMinY, (float)FrameT::mnMaxY};
        std::vector<int> match21(n2, -1);
        int nmatches = 0;
        check(orbx_search_by_projection_ex(n1, x3.data(), valid1.data(), obs1.data(), mpDesc.data(), level1.data(), angle1.data(),
                                           reinterpret_cast<const orbx_keypoint*>(CurrentFrame.mvKeysUn.data()), n2,
                                           d2.template ptr<unsigned char>(), nullptr, occupied2.data(), cam, CurrentFrame.mbf, bounds,
                                           CurrentFrame.mvScaleFactors.data(), (int)CurrentFrame.mvScaleFactors.size(), th, 0,
                                           mbCheckOrientation ? 1 : 0, ORBdist, ORBX_SBP_NO_DEPTH_TEST, match21.data(), &nmatches),
              "SearchByProjection");
        for (int j = 0; j < n2; ++j) {
            if (match21[j] >= 0) CurrentFrame.mvpMapPoints[j] = vpMPs[match21[j]];
            else if (match21[j] == -2) CurrentFrame.mvpMapPoints[j] = nullptr;
        }
        return nmatches;
    }

    // Project MapPoints using a Similarity Transformation and search matches (loop detection, src/LoopClosing.cc:278);
    // src/ORBmatcher.cc:290-403.  KeyFrame members read: fx, fy, cx, cy, mvScaleFactors, mvKeysUn, mDescriptors, IsInImage(),
    // mnMinX..mnMaxY; MapPoint: isBad(), GetWorldPos(), GetMin/MaxDistanceInvariance(), GetNormal(), PredictScale(), GetDescriptor().
    template <class KeyFrameT, class MapPointT>
    int SearchByProjection(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, std::vector<MapPointT*>& vpMatched,
                           int th) {
        const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy;
        const int n1 = (int)vpPoints.size(), n2 = (int)pKF->mvKeysUn.size();
        if (n1 == 0 || n2 == 0) return 0;
#if ORBX_HAVE_OPENCV
        cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
        const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
        cv::Mat Rcw = sRcw / scw;
        cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
        cv::Mat Ow = -Rcw.t() * tcw;
#else
        float R[9], t[3], Ow[3];
        double dd = 0;
        for (int c = 0; c < 3; ++c) dd += (double)Scw.template at<float>(0, c) * Scw.template at<float>(0, c);
        const float scw = (float)std::sqrt(dd);
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) R[3 * r + c] = Scw.template at<float>(r, c) / scw;
            t[r] = Scw.template at<float>(r, 3) / scw;
        }
        for (int c = 0; c < 3; ++c) Ow[c] = -(R[c] * t[0] + R[3 + c] * t[1] + R[6 + c] * t[2]);
#endif
        std::set<MapPointT*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
        spAlreadyFound.erase(static_cast<MapPointT*>(nullptr));
        std::vector<float> uvr(3 * (size_t)n1, 0.f);
        std::vector<int> level1(n1, 0);
        std::vector<unsigned char> valid1(n1, 0), mpDesc((size_t)n1 * 32, 0), occupied2(n2, 0);
        for (int i = 0; i < n1; ++i) {
            MapPointT* pMP = vpPoints[i];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
            const cv::Mat p3Dw = pMP->GetWorldPos();
            const cv::Mat Pn = pMP->GetNormal();
            float pc[3], dist;
            double podotpn;  // the reference compares the double PO.dot(Pn) with 0.5*dist
#if ORBX_HAVE_OPENCV
            const cv::Mat p3Dc = Rcw * p3Dw + tcw;
            for (int c = 0; c < 3; ++c) pc[c] = p3Dc.at<float>(c);
            const cv::Mat PO = p3Dw - Ow;
            dist = (float)cv::norm(PO);
            podotpn = PO.dot(Pn);
#else
            double ss = 0, dp = 0;
            for (int r = 0; r < 3; ++r) {
                pc[r] = R[3 * r] * p3Dw.template at<float>(0, 0) + R[3 * r + 1] * p3Dw.template at<float>(1, 0) +
                        R[3 * r + 2] * p3Dw.template at<float>(2, 0) + t[r];
                const float po = p3Dw.template at<float>(r, 0) - Ow[r];
                ss += (double)po * po;
                dp += (double)po * Pn.template at<float>(r, 0);
            }
            dist = (float)std::sqrt(ss);
            podotpn = dp;
#endif
            if (pc[2] < 0.0) continue;                   // depth must be positive (:333-334)
            const float invz = 1 / pc[2];
            const float x = pc[0] * invz, y = pc[1] * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!pKF->IsInImage(u, v)) continue;
            if (dist < pMP->GetMinDistanceInvariance() || dist > pMP->GetMaxDistanceInvariance()) continue;
            if (podotpn < 0.5 * dist) continue;          // viewing angle must be less than 60 deg (:358-359)
            const int nPredictedLevel = pMP->PredictScale(dist, pKF);
            valid1[i] = 1;
            level1[i] = nPredictedLevel;
            uvr[3 * i] = u;
            uvr[3 * i + 1] = v;
            uvr[3 * i + 2] = th * pKF->mvScaleFactors[nPredictedLevel];
            const cv::Mat dMP = pMP->GetDescriptor();
            std::memcpy(&mpDesc[(size_t)i * 32], dMP.template ptr<unsigned char>(), 32);
        }
        for (int j = 0; j < n2; ++j) occupied2[j] = vpMatched[j] != nullptr;
        const cv::Mat d2 = continuous(pKF->mDescriptors);
        const float bounds[4] = {(float)pKF->mnMinX, (float)pKF->mnMaxX, (float)pKF->mnMinY, (float)pKF->mnMaxY};
        std::vector<int> match21(n2, -1);
        int nmatches = 0;
        check(orbx_search_by_projection_sim3(n1, uvr.data(), level1.data(), valid1.data(), mpDesc.data(),
                                             reinterpret_cast<const orbx_keypoint*>(pKF->mvKeysUn.data()), n2,
                                             d2.template ptr<unsigned char>(), occupied2.data(), bounds, TH_LOW, match21.data(),
                                             &nmatches),
              "SearchByProjection");
        for (int j = 0; j < n2; ++j)
            if (match21[j] >= 0) vpMatched[j] = vpPoints[match21[j]];
        return nmatches;
    }

    // Search matches between MapPoints in a KeyFrame and ORB in a Frame, brute force constrained to ORB that belong to the same
    // vocabulary node (TrackReferenceKeyFrame src/Tracking.cc:842, Relocalization :1465); src/ORBmatcher.cc:159-288.
    // KeyFrame members read: GetMapPointMatches(), mFeatVec, mDescriptors, mvKeysUn; Frame: N, mFeatVec, mDescriptors, mvKeys.
    template <class KeyFrameT, class FrameT, class MapPointT>
    int SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches) {
        const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();
        vpMapPointMatches = std::vector<MapPointT*>(F.N, static_cast<MapPointT*>(nullptr));
        const int n1 = (int)vpMapPointsKF.size(), n2 = (int)F.N;
        BowNodes nodes;
        common_nodes(pKF->mFeatVec, F.mFeatVec, nodes);
        if (n1 == 0 || n2 == 0 || nodes.idx1.empty() || nodes.idx2.empty()) return 0;
        std::vector<unsigned char> valid1(n1), valid2(n2, 1);
        std::vector<float> angle1(n1), angle2(n2);
        for (int i = 0; i < n1; ++i) {
            valid1[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();  // :192-198
            angle1[i] = pKF->mvKeysUn[i].angle;
        }
        for (int j = 0; j < n2; ++j) angle2[j] = F.mvKeys[j].angle;  // :238 reads the distorted keypoints of the frame
        const cv::Mat d1 = continuous(pKF->mDescriptors), d2 = continuous(F.mDescriptors);
        std::vector<int> match12(n1, -1), match21(n2, -1);
        int nmatches = 0;
        check(orbx_search_by_bow((int)nodes.start1.size() - 1, nodes.start1.data(), nodes.idx1.data(), nodes.start2.data(),
                                 nodes.idx2.data(), n1, valid1.data(), d1.template ptr<unsigned char>(), angle1.data(), n2, valid2.data(),
                                 d2.template ptr<unsigned char>(), angle2.data(), TH_LOW, mfNNratio, mbCheckOrientation ? 1 : 0,
                                 match12.data(), match21.data(), &nmatches),
              "SearchByBoW");
        for (int j = 0; j < n2; ++j)
            if (match21[j] >= 0) vpMapPointMatches[j] = vpMapPointsKF[match21[j]];
        return nmatches;
    }

    // Matching between two KeyFrames for loop detection (LoopClosing::ComputeSim3, src/LoopClosing.cc:304);
    // src/ORBmatcher.cc:522-655.  KeyFrame members read: mvKeysUn, mFeatVec, GetMapPointMatches(), mDescriptors.
    template <class KeyFrameT, class MapPointT>
    int SearchByBoW(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12) {
        const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
        vpMatches12 = std::vector<MapPointT*>(vpMapPoints1.size(), static_cast<MapPointT*>(nullptr));
        const int n1 = (int)vpMapPoints1.size(), n2 = (int)vpMapPoints2.size();
        BowNodes nodes;
        common_nodes(pKF1->mFeatVec, pKF2->mFeatVec, nodes);
        if (n1 == 0 || n2 == 0 || nodes.idx1.empty() || nodes.idx2.empty()) return 0;
        std::vector<unsigned char> valid1(n1), valid2(n2);
        std::vector<float> angle1(n1), angle2(n2);
        for (int i = 0; i < n1; ++i) {
            valid1[i] = vpMapPoints1[i] && !vpMapPoints1[i]->isBad();  // :556-562
            angle1[i] = pKF1->mvKeysUn[i].angle;
        }
        for (int j = 0; j < n2; ++j) {
            valid2[j] = vpMapPoints2[j] && !vpMapPoints2[j]->isBad();  // :582-587
            angle2[j] = pKF2->mvKeysUn[j].angle;
        }
        const cv::Mat d1 = continuous(pKF1->mDescriptors), d2 = continuous(pKF2->mDescriptors);
        std::vector<int> match12(n1, -1), match21(n2, -1);
        int nmatches = 0;
        check(orbx_search_by_bow((int)nodes.start1.size() - 1, nodes.start1.data(), nodes.idx1.data(), nodes.start2.data(),
                                 nodes.idx2.data(), n1, valid1.data(), d1.template ptr<unsigned char>(), angle1.data(), n2, valid2.data(),
                                 d2.template ptr<unsigned char>(), angle2.data(), TH_LOW - 1 /* bestDist1 < TH_LOW, :605 */, mfNNratio,
                                 mbCheckOrientation ? 1 : 0, match12.data(), match21.data(), &nmatches),
              "SearchByBoW");
        for (int i = 0; i < n1; ++i)
            if (match12[i] >= 0) vpMatches12[i] = vpMapPoints2[match12[i]];
        return nmatches;
    }

    // Matching to triangulate new MapPoints, with the epipolar constraint (LocalMapping::CreateNewMapPoints, src/LocalMapping.cc:248ff);
    // src/ORBmatcher.cc:657-823.  KeyFrame members read: GetCameraCenter(), GetRotation(), GetTranslation(), fx, fy, cx, cy, N,
    // mFeatVec, GetMapPoint(idx), mvuRight, mvKeysUn, mDescriptors, mvScaleFactors, mvLevelSigma2.
    template <class KeyFrameT>
    int SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                               const bool bOnlyStereo) {
        // epipole in the second image (:664-672)
        const cv::Mat Cw = pKF1->GetCameraCenter(), R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
#if ORBX_HAVE_OPENCV
        const cv::Mat C2 = R2w * Cw + t2w;
        const float c2[3] = {C2.at<float>(0), C2.at<float>(1), C2.at<float>(2)};
#else
        float c2[3];  // cv::gemm on CV_32F accumulates in double
        for (int r = 0; r < 3; ++r) {
            double acc = 0;
            for (int c = 0; c < 3; ++c) acc += (double)R2w.template at<float>(r, c) * (double)Cw.template at<float>(c, 0);
            c2[r] = (float)(acc + (double)t2w.template at<float>(r, 0));
        }
#endif
        const float invz = 1.0f / c2[2];
        const float ex = pKF2->fx * c2[0] * invz + pKF2->cx;
        const float ey = pKF2->fy * c2[1] * invz + pKF2->cy;
        vMatchedPairs.clear();
        const int n1 = (int)pKF1->N, n2 = (int)pKF2->N;
        BowNodes nodes;
        common_nodes(pKF1->mFeatVec, pKF2->mFeatVec, nodes);
        if (n1 == 0 || n2 == 0 || nodes.idx1.empty() || nodes.idx2.empty()) return 0;
        std::vector<unsigned char> valid1(n1), stereo1(n1), valid2(n2), stereo2(n2);
        for (int i = 0; i < n1; ++i) {
            stereo1[i] = pKF1->mvuRight[i] >= 0;
            valid1[i] = !pKF1->GetMapPoint(i) && (!bOnlyStereo || stereo1[i]);  // :697-706
        }
        for (int j = 0; j < n2; ++j) {
            stereo2[j] = pKF2->mvuRight[j] >= 0;
            valid2[j] = !pKF2->GetMapPoint(j) && (!bOnlyStereo || stereo2[j]);  // :724-733
        }
        float f12[9];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) f12[3 * r + c] = F12.template at<float>(r, c);
        const cv::Mat d1 = continuous(pKF1->mDescriptors), d2 = continuous(pKF2->mDescriptors);
        std::vector<int> match12(n1, -1);
        int nmatches = 0;
        check(orbx_search_for_triangulation((int)nodes.start1.size() - 1, nodes.start1.data(), nodes.idx1.data(), nodes.start2.data(),
                                            nodes.idx2.data(), n1, reinterpret_cast<const orbx_keypoint*>(pKF1->mvKeysUn.data()),
                                            d1.template ptr<unsigned char>(), valid1.data(), stereo1.data(), n2,
                                            reinterpret_cast<const orbx_keypoint*>(pKF2->mvKeysUn.data()), d2.template ptr<unsigned char>(),
                                            valid2.data(), stereo2.data(), f12, ex, ey, pKF2->mvScaleFactors.data(),
                                            pKF2->mvLevelSigma2.data(), (int)pKF2->mvScaleFactors.size(), mbCheckOrientation ? 1 : 0,
                                            match12.data(), &nmatches),
              "SearchForTriangulation");
        vMatchedPairs.reserve(nmatches);
        for (int i = 0; i < n1; ++i)
            if (match12[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)match12[i]));  // :808-818
        return nmatches;
    }

    // Search matches between MapPoints seen in KF1 and KF2 transforming by a Sim3 [s12*R12|t12] (LoopClosing::ComputeSim3,
    // src/LoopClosing.cc:357); src/ORBmatcher.cc:1102-1326.  The two projection passes (:1150-1225, :1228-1305) gather their
    // candidates' inputs here and run their candidate loops on the device (orbx_fuse_match without the chi-square gate); the
    // agreement check (:1307-1323) is host code.  KeyFrame members read: fx, fy, cx, cy, GetRotation(), GetTranslation(),
    // GetMapPointMatches(), IsInImage(), mvScaleFactors, mvKeysUn, mDescriptors, mnMinX..mnMaxY; MapPoint: isBad(), GetWorldPos(),
    // GetIndexInKeyFrame(), GetMin/MaxDistanceInvariance(), PredictScale(), GetDescriptor().
    template <class KeyFrameT, class MapPointT>
    int SearchBySim3(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12, const float& s12, const cv::Mat& R12,
                     const cv::Mat& t12, const float th) {
        const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
        const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
        std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
        for (int i = 0; i < N1; i++) {
            MapPointT* pMP = vpMatches12[i];
            if (pMP) {
                vbAlreadyMatched1[i] = true;
                const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
                if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
            }
        }
        // transformation between the cameras (:1119-1122)
        float sR12[9], sR21[9], t12f[3], t21[3];
#if ORBX_HAVE_OPENCV
        {
            const cv::Mat a = s12 * R12, b = (1.0 / s12) * R12.t();
            const cv::Mat c = -b * t12;
            mat3(a, sR12); mat3(b, sR21); vec3(t12, t12f); vec3(c, t21);
        }
#else
        {
            float r[9];
            mat3(R12, r);
            vec3(t12, t12f);
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) {
                    sR12[3 * i + j] = (float)((double)s12 * (double)r[3 * i + j]);
                    sR21[3 * i + j] = (float)((1.0 / (double)s12) * (double)r[3 * j + i]);
                }
            for (int i = 0; i < 3; ++i) {
                double acc = 0;
                for (int j = 0; j < 3; ++j) acc += (double)sR21[3 * i + j] * (double)t12f[j];
                t21[i] = (float)(-acc);
            }
        }
#endif
        std::vector<int> vnMatch1(N1, -1), vnMatch2(N2, -1);
        // transform from KF1 to KF2 and search (:1150-1225), then from KF2 to KF1 (:1228-1305); both project with pKF1's intrinsics
        sim3_pass(pKF1, pKF2, vpMapPoints1, vbAlreadyMatched1, sR21, t21, pKF1->fx, pKF1->fy, pKF1->cx, pKF1->cy, th, vnMatch1);
        sim3_pass(pKF2, pKF1, vpMapPoints2, vbAlreadyMatched2, sR12, t12f, pKF1->fx, pKF1->fy, pKF1->cx, pKF1->cy, th, vnMatch2);
        int nFound = 0;  // check agreement (:1307-1323)
        for (int i1 = 0; i1 < N1; i1++) {
            const int idx2 = vnMatch1[i1];
            if (idx2 >= 0 && vnMatch2[idx2] == i1) {
                vpMatches12[i1] = vpMapPoints2[idx2];
                nFound++;
            }
        }
        return nFound;
    }


    // Project MapPoints into KeyFrame and search for duplicated MapPoints (LocalMapping::SearchInNeighbors, src/LocalMapping.cc:519ff);
    // src/ORBmatcher.cc:825-975.  KeyFrame members read: GetRotation(), GetTranslation(), GetCameraCenter(), fx, fy, cx, cy, mbf,
    // IsInImage(), mvScaleFactors, mvInvLevelSigma2, mvKeysUn, mvuRight, mDescriptors, mnMinX..mnMaxY, GetMapPoint(), AddMapPoint();
    // MapPoint: isBad(), IsInKeyFrame(), GetWorldPos(), GetMin/MaxDistanceInvariance(), GetNormal(), PredictScale(), GetDescriptor(),
    // Observations(), Replace(), AddObservation().  The candidate loop of a point (:886-945) reads nothing the surgery of earlier
    // points changes, so all candidate loops run first; `isBad() || IsInKeyFrame(pKF)` (:849), which the surgery DOES change (a
    // point listed twice, a point replaced on the way), is evaluated in the surgery loop, in the reference's order.
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT* pKF, const std::vector<MapPointT*>& vpMapPoints, const float th = 3.0) {
        const int n1 = (int)vpMapPoints.size(), n2 = (int)pKF->mvKeysUn.size();
        if (n1 == 0 || n2 == 0) return 0;
        float R[9], t[3], Ow[3];
        mat3(pKF->GetRotation(), R);
        vec3(pKF->GetTranslation(), t);
        vec3(pKF->GetCameraCenter(), Ow);
        FuseCandidates c;
        project_for_fuse(pKF, R, t, Ow, vpMapPoints, th, pKF->mbf, true, c);
        fuse_candidates(pKF, c, true, "Fuse");
        int nFused = 0;
        for (int i = 0; i < n1; i++) {
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP) continue;
            if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
            if (!c.valid[i] || c.bestIdx[i] < 0) continue;
            if (c.bestDist[i] <= TH_LOW) {  // if there is already a MapPoint replace, otherwise add a new measurement (:947-970)
                const int bestIdx = c.bestIdx[i];
                MapPointT* pMPinKF = pKF->GetMapPoint(bestIdx);
                if (pMPinKF) {
                    if (!pMPinKF->isBad()) {
                        if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                        else pMPinKF->Replace(pMP);
                    }
                } else {
                    pMP->AddObservation(pKF, bestIdx);
                    pKF->AddMapPoint(pMP, bestIdx);
                }
                nFused++;
            }
        }
        return nFused;
    }

    // Project MapPoints into KeyFrame using a given Sim3 and search for duplicated MapPoints (LoopClosing::SearchAndFuse,
    // src/LoopClosing.cc:644); src/ORBmatcher.cc:977-1100.  Reads GetMapPoints() in addition; no chi-square gate, no Replace: a
    // keypoint that already holds a good point is reported through vpReplacePoint.
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT* pKF, cv::Mat Scw, const std::vector<MapPointT*>& vpPoints, float th, std::vector<MapPointT*>& vpReplacePoint) {
        const int n1 = (int)vpPoints.size(), n2 = (int)pKF->mvKeysUn.size();
        if (n1 == 0 || n2 == 0) return 0;
        float R[9], t[3], Ow[3];
        decompose_sim3(Scw, R, t, Ow);
        const std::set<MapPointT*> spAlreadyFound = pKF->GetMapPoints();  // fixed for the whole call (:993)
        std::vector<MapPointT*> pts(vpPoints);
        for (int i = 0; i < n1; ++i)
            if (pts[i] && (pts[i]->isBad() || spAlreadyFound.count(pts[i]))) pts[i] = nullptr;  // :1005-1006 (nothing in this overload changes it)
        FuseCandidates c;
        project_for_fuse(pKF, R, t, Ow, pts, th, 0.f, false, c);
        fuse_candidates(pKF, c, false, "Fuse");
        int nFused = 0;
        for (int iMP = 0; iMP < n1; iMP++) {
            MapPointT* pMP = pts[iMP];
            if (!pMP || !c.valid[iMP] || c.bestIdx[iMP] < 0) continue;
            if (c.bestDist[iMP] <= TH_LOW) {  // :1077-1095
                const int bestIdx = c.bestIdx[iMP];
                MapPointT* pMPinKF = pKF->GetMapPoint(bestIdx);
                if (pMPinKF) {
                    if (!pMPinKF->isBad()) vpReplacePoint[iMP] = pMPinKF;
                } else {
                    pMP->AddObservation(pKF, bestIdx);
                    pKF->AddMapPoint(pMP, bestIdx);
                }
                nFused++;
            }
        }
        return nFused;
    }

    static const int TH_LOW = 50;
    static const int TH_HIGH = 100;
    static const int HISTO_LENGTH = 30;

    static void check(int rc, const char* what) {
        if (rc != ORBX_OK) throw std::runtime_error(std::string(what) + ": " + orbx_last_error());
    }
    static cv::Mat continuous(const cv::Mat& m) { return m.isContinuous() ? m : m.clone(); }

    // SearchBySim3 helpers.  gemm3: out = M * p + t for a 3x3 float M the way cv::gemm evaluates it on CV_32F (double accumulator).
    static void gemm3(const float* M, const float* p, const float* t, float* out) {
        for (int r = 0; r < 3; ++r) {
            double acc = 0;
            for (int c = 0; c < 3; ++c) acc += (double)M[3 * r + c] * (double)p[c];
            out[r] = (float)(acc + (double)t[r]);
        }
    }
    static void mat3(const cv::Mat& m, float* out) {
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) out[3 * r + c] = m.template at<float>(r, c);
    }
    static void vec3(const cv::Mat& m, float* out) {
        for (int r = 0; r < 3; ++r) out[r] = m.template at<float>(r, 0);
    }
    // One projection pass of SearchBySim3: the map points seen in pSrc, moved into pSrc's camera (R p + t of pSrc) and across the
    // Sim3 ([M | t]), projected into pDst's image; the candidate loop runs on the device.
    template <class KeyFrameT, class MapPointT>
    void sim3_pass(KeyFrameT* pSrc, KeyFrameT* pDst, const std::vector<MapPointT*>& vpPoints, const std::vector<bool>& vbAlready,
                   const float* M, const float* t, float fx, float fy, float cx, float cy, float th, std::vector<int>& vnMatch) const {
        const int n1 = (int)vpPoints.size(), n2 = (int)pDst->mvKeysUn.size();
        if (n1 == 0 || n2 == 0) return;
        float Rs[9], ts[3];
        mat3(pSrc->GetRotation(), Rs);
        vec3(pSrc->GetTranslation(), ts);
        std::vector<float> uvr(3 * (size_t)n1, 0.f);
        std::vector<int> level(n1, 0);
        std::vector<unsigned char> valid(n1, 0), mpDesc((size_t)n1 * 32, 0);
        for (int i = 0; i < n1; ++i) {
            MapPointT* pMP = vpPoints[i];
            if (!pMP || vbAlready[i] || pMP->isBad()) continue;
            float pw[3], own[3], pc[3];
            vec3(pMP->GetWorldPos(), pw);
            gemm3(Rs, pw, ts, own);   // p3Dc1 = R1w*p3Dw + t1w
            gemm3(M, own, t, pc);     // p3Dc2 = sR21*p3Dc1 + t21
            if (pc[2] < 0.0) continue;  // depth must be positive
            const float invz = 1.0 / pc[2];
            const float x = pc[0] * invz, y = pc[1] * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!pDst->IsInImage(u, v)) continue;
            const float dist3D = (float)std::sqrt((double)pc[0] * pc[0] + (double)pc[1] * pc[1] + (double)pc[2] * pc[2]);  // cv::norm
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
            level[i] = pMP->PredictScale(dist3D, pDst);
            uvr[3 * i] = u;
            uvr[3 * i + 1] = v;
            uvr[3 * i + 2] = th * pDst->mvScaleFactors[level[i]];
            const cv::Mat d = pMP->GetDescriptor();
            std::memcpy(&mpDesc[(size_t)i * 32], d.template ptr<unsigned char>(), 32);
            valid[i] = 1;
        }
        const cv::Mat d2 = continuous(pDst->mDescriptors);
        const float bounds[4] = {(float)pDst->mnMinX, (float)pDst->mnMaxX, (float)pDst->mnMinY, (float)pDst->mnMaxY};
        std::vector<int> bestIdx(n1, -1), bestDist(n1, 256);
        check(orbx_fuse_match(n1, uvr.data(), nullptr, level.data(), valid.data(), mpDesc.data(),
                              reinterpret_cast<const orbx_keypoint*>(pDst->mvKeysUn.data()), n2, d2.template ptr<unsigned char>(), nullptr,
                              nullptr, 64, bounds, bestIdx.data(), bestDist.data()),
              "SearchBySim3");
        for (int i = 0; i < n1; ++i)
            if (bestIdx[i] >= 0 && bestDist[i] <= TH_HIGH) vnMatch[i] = bestIdx[i];  // :1221-1224
    }


    // ---- Fuse helpers ----
    struct FuseCandidates {  // per map point: the projection of :852-888 / :1009-1046 and the result of the candidate loop
        std::vector<float> uvr, ur;
        std::vector<int> level, bestIdx, bestDist;
        std::vector<unsigned char> valid, desc;
    };
    // Scw -> Rcw, tcw, Ow (:986-990)
    static void decompose_sim3(const cv::Mat& Scw, float* R, float* t, float* Ow) {
#if ORBX_HAVE_OPENCV
        cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
        const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
        cv::Mat Rcw = sRcw / scw;
        cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
        cv::Mat O = -Rcw.t() * tcw;
        mat3(Rcw, R); vec3(tcw, t); vec3(O, Ow);
#else
        double dd = 0;
        for (int c = 0; c < 3; ++c) dd += (double)Scw.template at<float>(0, c) * Scw.template at<float>(0, c);
        const float scw = (float)std::sqrt(dd);
        const float inv = (float)(1.0 / (double)scw);  // cv::Mat / scalar multiplies by the float-rounded reciprocal
        for (int r = 0; r < 3; ++r) {
            for (int c = 0; c < 3; ++c) R[3 * r + c] = Scw.template at<float>(r, c) * inv;
            t[r] = Scw.template at<float>(r, 3) * inv;
        }
        for (int c = 0; c < 3; ++c) Ow[c] = -(float)((double)R[c] * t[0] + (double)R[3 + c] * t[1] + (double)R[6 + c] * t[2]);
#endif
    }
    // The per-point projection and visibility tests shared by both Fuse overloads (null entries are skipped).  Products and the norm /
    // dot accumulate in double and round once, as cv::Mat's CV_32F paths do; comparisons against 0.5 * dist3D are made in double
    // like the reference's `PO.dot(Pn) < 0.5*dist3D`.
    template <class KeyFrameT, class MapPointT>
    static void project_for_fuse(KeyFrameT* pKF, const float* R, const float* t, const float* Ow, const std::vector<MapPointT*>& pts, float th,
                                 float bf, bool stereoGate, FuseCandidates& c) {
        const int n1 = (int)pts.size();
        const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy;
        c.uvr.assign(3 * (size_t)n1, 0.f);
        c.ur.assign(n1, 0.f);
        c.level.assign(n1, 0);
        c.valid.assign(n1, 0);
        c.desc.assign((size_t)n1 * 32, 0);
        (void)stereoGate;
#if ORBX_HAVE_OPENCV
        const cv::Mat Rcw(3, 3, CV_32F, (void*)R), tcw(3, 1, CV_32F, (void*)t), Owm(3, 1, CV_32F, (void*)Ow);
#endif
        for (int i = 0; i < n1; ++i) {
            MapPointT* pMP = pts[i];
            if (!pMP) continue;
            float pw[3], pc[3], pn[3];
            vec3(pMP->GetWorldPos(), pw);
#if ORBX_HAVE_OPENCV
            {  // the reference's own expression, evaluated by OpenCV
                const cv::Mat p3Dw = pMP->GetWorldPos();
                const cv::Mat p3Dc = Rcw * p3Dw + tcw;
                vec3(p3Dc, pc);
            }
#else
            for (int r = 0; r < 3; ++r) {  // p3Dc = Rcw*p3Dw + tcw
                double acc = 0;
                for (int k = 0; k < 3; ++k) acc += (double)R[3 * r + k] * (double)pw[k];
                pc[r] = (float)acc + t[r];
            }
#endif
            if (pc[2] < 0.0f) continue;  // depth must be positive
            const float invz = 1 / pc[2];
            const float x = pc[0] * invz, y = pc[1] * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!pKF->IsInImage(u, v)) continue;
            const float ur = u - bf * invz;
            const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
            vec3(pMP->GetNormal(), pn);
            double ss = 0, dp = 0;
#if ORBX_HAVE_OPENCV
            {
                const cv::Mat PO = pMP->GetWorldPos() - Owm;
                ss = cv::norm(PO);
                ss *= ss;
                dp = PO.dot(pMP->GetNormal());
            }
            const float dist3D = (float)cv::norm(pMP->GetWorldPos() - Owm);
#else
            for (int r = 0; r < 3; ++r) {
                const float po = pw[r] - Ow[r];
                ss += (double)po * (double)po;
                dp += (double)po * (double)pn[r];
            }
            const float dist3D = (float)std::sqrt(ss);
#endif
            if (dist3D < minDistance || dist3D > maxDistance) continue;  // depth must be inside the scale pyramid of the image
            if (dp < 0.5 * dist3D) continue;                              // viewing angle must be less than 60 deg
            const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
            c.valid[i] = 1;
            c.level[i] = nPredictedLevel;
            c.uvr[3 * i] = u;
            c.uvr[3 * i + 1] = v;
            c.uvr[3 * i + 2] = th * pKF->mvScaleFactors[nPredictedLevel];
            c.ur[i] = ur;
            const cv::Mat dMP = pMP->GetDescriptor();
            std::memcpy(&c.desc[(size_t)i * 32], dMP.template ptr<unsigned char>(), 32);
        }
    }
    template <class KeyFrameT>
    static void fuse_candidates(KeyFrameT* pKF, FuseCandidates& c, bool gate, const char* what) {
        const int n1 = (int)c.valid.size(), n2 = (int)pKF->mvKeysUn.size();
        const cv::Mat d2 = continuous(pKF->mDescriptors);
        const float bounds[4] = {(float)pKF->mnMinX, (float)pKF->mnMaxX, (float)pKF->mnMinY, (float)pKF->mnMaxY};
        c.bestIdx.assign(n1, -1);
        c.bestDist.assign(n1, 256);
        const bool hasRight = gate && (int)pKF->mvuRight.size() == n2;
        check(orbx_fuse_match(n1, c.uvr.data(), gate ? c.ur.data() : nullptr, c.level.data(), c.valid.data(), c.desc.data(),
                              reinterpret_cast<const orbx_keypoint*>(pKF->mvKeysUn.data()), n2, d2.template ptr<unsigned char>(),
                              hasRight ? pKF->mvuRight.data() : nullptr, gate ? pKF->mvInvLevelSigma2.data() : nullptr,
                              gate ? (int)pKF->mvInvLevelSigma2.size() : 64, bounds, c.bestIdx.data(), c.bestDist.data()),
              what);
    }

    // The common nodes of two DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned int>>) in map order — the walk of
    // src/ORBmatcher.cc:176-256 — flattened for orbx_search_by_bow.
    struct BowNodes {
        std::vector<int> start1{0}, start2{0};
        std::vector<uint32_t> idx1, idx2;
    };
    template <class FeatVecT> static void common_nodes(const FeatVecT& fv1, const FeatVecT& fv2, BowNodes& out) {
        typename FeatVecT::const_iterator it1 = fv1.begin(), it2 = fv2.begin();
        while (it1 != fv1.end() && it2 != fv2.end()) {
            if (it1->first == it2->first) {
                out.idx1.insert(out.idx1.end(), it1->second.begin(), it1->second.end());
                out.idx2.insert(out.idx2.end(), it2->second.begin(), it2->second.end());
                out.start1.push_back((int)out.idx1.size());
                out.start2.push_back((int)out.idx2.size());
                ++it1;
                ++it2;
            } else if (it1->first < it2->first) {
                it1 = fv1.lower_bound(it2->first);
            } else {
                it2 = fv2.lower_bound(it1->first);
            }
        }
    }

protected:
    float mfNNratio;
    bool mbCheckOrientation;
};

// Frame::ComputeStereoMatches (src/Frame.cc:465-639) for a frame whose two extractors are ORB_SLAM2::ORBextractor
// objects of this library: reads mvKeys / mvKeysRight / mDescriptors / mDescriptorsRight / mbf / mb / N and the two
// extractor pointers, fills mvuRight and mvDepth.  The pyramids are read where they already are — in device memory —
// so mvImagePyramid need not be copied to the host for it.  Body of the reference member function becomes
//     ORB_SLAM2::ComputeStereoMatches(*this);
template <class FrameT> void ComputeStereoMatches(FrameT& F) {
    const int N = (int)F.mvKeys.size(), Nr = (int)F.mvKeysRight.size();
    F.mvuRight = std::vector<float>(N, -1.0f);
    F.mvDepth = std::vector<float>(N, -1.0f);
    if (N == 0) return;
    const cv::Mat dL = ORBmatcher::continuous(F.mDescriptors), dR = ORBmatcher::continuous(F.mDescriptorsRight);
    ORBmatcher::check(orbx_stereo_match(F.mpORBextractorLeft->handle(), F.mpORBextractorRight->handle(),
                                        reinterpret_cast<const orbx_keypoint*>(F.mvKeys.data()), N, dL.template ptr<unsigned char>(),
                                        reinterpret_cast<const orbx_keypoint*>(F.mvKeysRight.data()), Nr,
                                        dR.template ptr<unsigned char>(), F.mbf, F.mb, F.mvuRight.data(), F.mvDepth.data()),
                      "ComputeStereoMatches");
}

// The same for the place the reference calls it from — the stereo Frame constructor (src/Frame.cc:77-94), where mvKeys /
// mvKeysRight and the descriptors are still exactly what the two operator() calls returned: they are read where the extractors
// left them on the device, nothing is uploaded (orbx_stereo_match_last).  Throws if either extractor has run anything else since.
template <class FrameT> void ComputeStereoMatchesOnDevice(FrameT& F) {
    const int N = (int)F.mvKeys.size();
    F.mvuRight = std::vector<float>(N, -1.0f);
    F.mvDepth = std::vector<float>(N, -1.0f);
    if (N == 0) return;
    ORBmatcher::check(orbx_stereo_match_last(F.mpORBextractorLeft->handle(), F.mpORBextractorRight->handle(), F.mbf, F.mb, F.mvuRight.data(),
                                             F.mvDepth.data(), N),
                      "ComputeStereoMatchesOnDevice");
}

}  // namespace ORB_SLAM2

#endif  // ORBMATCHER_H
