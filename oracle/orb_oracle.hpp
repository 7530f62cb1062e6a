// oracle/orb_oracle.hpp — TEST INFRASTRUCTURE ONLY (CPU oracle for the ORB front end).
//
// This is a plain C++17 restatement, with no OpenCV dependency, of the CPU path ("path B") of the
// reference's ORBextractor / ORBmatcher / Frame::ComputeStereoMatches.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may build, load or call it.  The product
// (orb_slam_cuda_b200/, include/) never links or calls anything in this directory.
//
// Parity pinning.  The reference ships no tests or golden vectors.
//   * Matchers, Frame grid, ComputeStereoMatches: PINNED TO REFERENCE OBJECT CODE.  /root/reference's own src/ORBmatcher.cc and
//     src/Frame.cc are compiled where they lie (oracle/ref_matcher/Makefile -> oracle/_ref/libref_matcher.so) and every matcher
//     below equals them on 25 seeded cases, live and through tests/golden/ref_matcher.npz (tests/test_oracle_ref_matcher.py).
//     The DBoW2 descent is pinned the same way against the reference's Thirdparty/DBoW2 (oracle/ref_dbow2).
//   * Extractor (ORBextractor.cc): CONTROL FLOW PINNED TO REFERENCE OBJECT CODE.  /root/reference's own src/ORBextractor.cc is
//     compiled where it lies (oracle/ref_extractor/Makefile -> oracle/_ref/libref_extractor.so; OPENVX undefined = its CPU
//     branch, a do-nothing OpenVX stand-in, a monotonic heap for quadtree nodes so its pointer tie-breaks are defined) and
//     Extractor::extract below equals it bit for bit on 18 whole frames + quadtree tie cases, live and through
//     tests/golden/ref_extractor.npz (tests/test_oracle_ref_extractor.py).  The five OpenCV pixel primitives it calls (resize,
//     copyMakeBorder, FAST, GaussianBlur, fastAtan2) are not under /root/reference: they are pinned bit-exactly against cv2
//     4.13 (tests/test_oracle_cv2.py and the fixtures under tests/golden/), the 1024 pattern integers against the reference
//     text, the compiled reference and cv2.ORB.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace orc {

// Same field order and size (28 B) as cv::KeyPoint: pt.x, pt.y, size, angle, response, octave, class_id.
struct KeyPoint {
    float x, y, size, angle, response;
    int octave, class_id;
};

struct Image {  // owning 8-bit plane
    int w = 0, h = 0;
    std::vector<uint8_t> px;
    Image() {}
    Image(int w_, int h_) : w(w_), h(h_), px((size_t)w_ * h_) {}
    uint8_t* row(int y) { return px.data() + (size_t)y * w; }
    const uint8_t* row(int y) const { return px.data() + (size_t)y * w; }
};

// ---- OpenCV primitive models -------------------------------------------------------------------
int cv_round(float v);
int cv_round(double v);
void resize_linear_8u(const uint8_t* src, int sw, int sh, size_t sstep, uint8_t* dst, int dw, int dh, size_t dstep);
int reflect101(int p, int n);
void border_reflect101(const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep, int pad);
void fast9_16(const uint8_t* img, int w, int h, size_t step, int threshold, bool nms, std::vector<KeyPoint>& out);
void gaussian7x7_s2(const uint8_t* src, int w, int h, size_t sstep, uint8_t* dst, size_t dstep);
float fast_atan2(float y, float x);

// ---- ORBextractor (src/ORBextractor.cc, CPU path) -------------------------------------------------
constexpr int PATCH_SIZE = 31, HALF_PATCH_SIZE = 15, EDGE_THRESHOLD = 19;

std::vector<KeyPoint> distribute_octtree(const std::vector<KeyPoint>& keys, int minX, int maxX, int minY, int maxY,
                                         int N);

struct Extractor {
    int nfeatures, nlevels, iniThFAST, minThFAST;
    double scaleFactor;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
    std::vector<int> mnFeaturesPerLevel, umax;
    // Per-frame state (kept for stage-by-stage parity tests)
    std::vector<Image> padded;               // (w+38) x (h+38) planes; mvImagePyramid[l] = centre ROI
    std::vector<int> lw, lh;                 // unpadded level sizes
    std::vector<std::vector<KeyPoint>> candidates;  // vToDistributeKeys per level
    std::vector<std::vector<KeyPoint>> levelKeys;   // after quadtree + border offset + angle (level coords)
    std::vector<Image> blurred;              // unpadded blurred planes (empty if level had no keypoints)

    Extractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
    void compute_pyramid(const uint8_t* img, int w, int h, size_t step);
    void compute_keypoints();
    // Full operator(): returns keypoints (level-0 coords) and N x 32 descriptors.
    void extract(const uint8_t* img, int w, int h, size_t step, std::vector<KeyPoint>& kps, std::vector<uint8_t>& desc);
    const uint8_t* level_ptr(int l) const { return padded[l].row(EDGE_THRESHOLD) + EDGE_THRESHOLD; }
    size_t level_step(int l) const { return (size_t)padded[l].w; }
};

// bit_pattern_31_ variant used by orb_descriptor / Extractor::extract: 0 = the reference as it compiles (element 96 = VX_FAILURE - 2
// = -3, src/ORBextractor.cc:261-262; the default), 1 = the canonical ORB pattern (-2 there)
void set_pattern_variant(int v);
int pattern_variant();
const int* pattern_values();  // the 1024 integers of the current variant
float ic_angle(const uint8_t* center, size_t step, const std::vector<int>& umax);
void orb_descriptor(float angle_deg, const uint8_t* center, size_t step, uint8_t* desc);

// ---- ORBmatcher / Frame pieces -------------------------------------------------------------------
int descriptor_distance(const uint8_t* a, const uint8_t* b);

struct GridFrame {  // the parts of Frame that SearchForInitialization touches
    static constexpr int COLS = 64, ROWS = 48;
    float minX, minY, maxX, maxY, invW, invH;
    const KeyPoint* keys;
    int n;
    std::vector<size_t> grid[COLS][ROWS];
    GridFrame(const KeyPoint* keys, int n, float minX, float maxX, float minY, float maxY);
    std::vector<size_t> features_in_area(float x, float y, float r, int minLevel, int maxLevel) const;
};

int search_for_initialization(const GridFrame& F1, const uint8_t* desc1, const GridFrame& F2, const uint8_t* desc2,
                              float* prevMatchedXY, int* matches12, int windowSize, float nnratio, bool checkOri);

int search_by_projection_frame(int n1, const float* x3Dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mpDesc1,
                               const int* octave1, const float* angle1, const GridFrame& F2, const uint8_t* desc2,
                               const float* uRight2, const uint8_t* occupied2, float fx, float fy, float cx, float cy, float mbf,
                               const float* scaleFactors, float th, int direction, bool checkOri, int* match21, int orbDist = 100,
                               bool depthTest = true);

int search_by_projection_map(int n1, const float* proj, const float* viewCos, const int* level1, const uint8_t* valid1,
                             const uint8_t* obs1, const uint8_t* mpDesc1, const GridFrame& F, const uint8_t* desc2,
                             const float* uRight2, const uint8_t* occupied2, const float* scaleFactors, float th, float nnratio,
                             int* match21);

int search_by_projection_sim3(int n1, const float* uvr, const int* level1, const uint8_t* valid1, const uint8_t* mpDesc1,
                              const GridFrame& KF, const uint8_t* desc2, const uint8_t* occupied2, int thDist, int* match21);

void fuse_match(int n1, const float* uvr, const float* ur1, const int* level1, const uint8_t* valid1, const uint8_t* mpDesc1,
                const GridFrame& KF, const uint8_t* desc2, const float* uRight2, const float* invLevelSigma2, int* bestIdxOut,
                int* bestDistOut);

int search_by_bow(int nNodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                  const uint8_t* valid1, const uint8_t* desc1, const float* angle1, int n2, const uint8_t* valid2,
                  const uint8_t* desc2, const float* angle2, int thDist, float nnratio, bool checkOri, int* match12, int* match21);

// DBoW2 vocabulary tree descent (Frame::ComputeBoW, src/Frame.cc:394-401 -> TemplatedVocabulary::transform).  Nodes as
// loadFromTextFile numbers them: node 0 = root, node i >= 1 in file order with parent[i], isLeaf[i], desc[32 i], weight[i].
// nodeId[f] = 0xFFFFFFFF when the descent ends above level L - levelsup (the reference leaves *nid unwritten there).
void vocabulary_transform(int nNodes, const int* parent, const uint8_t* isLeaf, const uint8_t* desc, const double* weight, int L,
                          int n, const uint8_t* feat, int levelsup, uint32_t* wordId, double* w, uint32_t* nodeId);

int search_for_triangulation(int nNodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                             const KeyPoint* k1, const uint8_t* desc1, const uint8_t* valid1, const uint8_t* stereo1, int n2,
                             const KeyPoint* k2, const uint8_t* desc2, const uint8_t* valid2, const uint8_t* stereo2, const float* F12,
                             float ex, float ey, const float* scaleFactors2, const float* levelSigma2, bool checkOri, int* match12);

int distinctive_descriptor(const uint8_t* desc, int n, int* medianOut);

void match_bruteforce(const uint8_t* q, int nq, const uint8_t* d, int nd, int th, float ratio, int* idx, int* d1,
                      int* d2, int* match);

void stereo_matches(const Extractor& L, const Extractor& R, const KeyPoint* kL, int nL, const uint8_t* dL,
                    const KeyPoint* kR, int nR, const uint8_t* dR, float bf, float fx, float* uRight, float* depth);

}  // namespace orc
