// include/ORBextractor.h — ORB_SLAM2::ORBextractor with the class surface of the reference
// (include/ORBextractor.h:76-116 there), implemented on the B200 through the C ABI of include/orbx.h.
//
// Drop-in for the reference header + src/ORBextractor.cc: src/Frame.cc (:68-74, :126-132, :181-187 accessors,
// :246-252 operator(), :472/:562/:574/:579 mvImagePyramid) and src/Tracking.cc (:127-133 construction) compile against
// it unchanged.  What is NOT carried over: the OpenVX members (vx_context, vx_graph, ... :139-160), buildGraph, the
// GetTime friend and ExtractorNode (the quadtree lives in a CUDA kernel) — Jetson / VisionWorks glue that the hot
// path drops.  There is no CPU implementation behind this class: construction throws std::runtime_error when the
// device library cannot run.
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <vector>

#include "orbx.h"
#include "orbx_cv.h"

namespace ORB_SLAM2 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // The reference's 7-argument form (width / height come from Camera.width / Camera.height of the YAML,
    // src/Tracking.cc:124-133, and may be 0 = unknown until the first frame) ...
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int width, int height);
    // ... and upstream ORB-SLAM2's 5-argument form.
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
    ~ORBextractor();
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // Compute the ORB features and descriptors on an image.  Mask is ignored, as in the reference.
    // Empty image: returns without touching the outputs (src/ORBextractor.cc:1542-1543).
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints,
                    cv::OutputArray descriptors);

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // Level l is the centre ROI of a (w + 38) x (h + 38) host copy of the device plane, so rowRange / colRange up to
    // 19 px outside the level stay inside allocated memory, like the reference's temp(Rect(EDGE, EDGE, w, h)) ROIs
    // (src/ORBextractor.cc:1845-1847).  Filled after every operator() call only when `keepImagePyramid` is true
    // (default): ComputeStereoMatches is the one reader (src/Frame.cc:472-579); monocular users can switch the
    // device-to-host copies off.
    std::vector<cv::Mat> mvImagePyramid;
    bool keepImagePyramid = true;

    // bit_pattern_31_ as the reference compiles it (default; element 96 = VX_FAILURE - 2 = -3, src/ORBextractor.cc:261-262) or the
    // canonical ORB table of OpenCV / upstream ORB-SLAM2 (see orbx_extractor_set_pattern in orbx.h)
    void UseCanonicalPattern(bool on) { orbx_extractor_set_pattern(mpHandle, on ? ORBX_PATTERN_CANONICAL : ORBX_PATTERN_REFERENCE); }

    // the C handle, for the matcher entry points that read the device-resident pyramid (orbx_stereo_match)
    orbx_extractor* handle() const { return mpHandle; }

protected:
    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<int> umax;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

private:
    void init(int width, int height);
    orbx_extractor* mpHandle;
    std::vector<orbx_keypoint> mKeyBuf;
    std::vector<unsigned char> mDescBuf;
};

}  // namespace ORB_SLAM2

#endif
