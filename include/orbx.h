/* include/orbx.h — C ABI of the B200-native ORB front end (liborbx.so).
 *
 * This is the drop-in boundary.  The reference has no FFI layer for this path: its boundary is
 * the C++ class surface of ORB_SLAM2::ORBextractor / ORBmatcher / Frame::ComputeStereoMatches.  The C++ classes in
 * include/ORBextractor.h and include/ORBmatcher.h keep that surface and forward to the entry points below; a
 * binding from any other host language (ctypes, cgo, JNI …) binds the same symbols.
 *
 * Conventions
 *   - every function returns an int status: 0 = ORBX_OK, negative = error; orbx_last_error() gives the
 *     thread-local message of the last failing call on this thread;
 *   - plain pointers and sizes only; "host" entry points take host memory and do their own H2D/D2H,
 *     "_device" entry points take device pointers + a cudaStream_t (passed as void*) and are asynchronous;
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with ORBX_E_CUDA;
 *   - a handle is used by one thread at a time; different handles are independent (own stream + scratch),
 *     like the two extractor objects of the reference's stereo path (src/Frame.cc:77-80).
 */
#ifndef ORBX_H_
#define ORBX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_OK 0
#define ORBX_E_INVALID (-1)  /* bad argument / unsupported geometry */
#define ORBX_E_CUDA (-2)     /* CUDA runtime or driver failure (includes "no device") */
#define ORBX_E_OOM (-3)      /* allocation failure */
#define ORBX_E_OVERFLOW (-4) /* the caller's output capacity (cap) was exceeded, or one level held more than 2^20 FAST corners; a FAST cell cannot overflow (slots sized by the NMS packing bound) */

#define ORBX_MAX_LEVELS 16

/* Bit-compatible with cv::KeyPoint (28 bytes): pt.x, pt.y, size, angle, response, octave, class_id. */
typedef struct orbx_keypoint {
    float x, y, size, angle, response;
    int octave, class_id;
} orbx_keypoint;

/* Constructor arguments of ORBextractor (reference include/ORBextractor.h:82-83, src/ORBextractor.cc:496-500).
 * width/height are hints and may be 0 (size taken from the first frame); max_batch = frames per launch the
 * scratch is sized for (grown on demand); device = CUDA ordinal, -1 = current device. */
typedef struct orbx_params {
    int nfeatures;
    float scale_factor;
    int nlevels;
    int ini_th_fast;
    int min_th_fast;
    int width, height;
    int max_batch;
    int device;
} orbx_params;

typedef struct orbx_extractor orbx_extractor;

const char* orbx_last_error(void);
int orbx_version(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
long long orbx_kernel_launches(void);

/* ---- ORBextractor: ctor / dtor / accessors  (src/ORBextractor.cc:496-560, include/ORBextractor.h:94-114) ---- */
int orbx_extractor_create(const orbx_params* p, orbx_extractor** out);
void orbx_extractor_destroy(orbx_extractor* ex);
/* each array has nlevels entries; any pointer may be NULL */
int orbx_get_scale_tables(const orbx_extractor* ex, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                          int* features_per_level);
int orbx_get_umax(const orbx_extractor* ex, int* umax16);
/* Which bit_pattern_31_ (src/ORBextractor.cc:236-494) the descriptors are computed with.  The reference's working copy has the
 * token VX_FAILURE between two rows of the table (:261): `-8,-7, -6,-2,VX_FAILURE <newline> -2,11, ...` is the expression
 * VX_FAILURE - 2, and OpenVX defines VX_FAILURE = -1 — so the reference AS IT COMPILES holds -3 in element 96 (x of the first
 * point of pair 24: descriptor byte 3, bit 0) where the canonical ORB pattern of OpenCV / upstream ORB-SLAM2 holds -2.
 *   ORBX_PATTERN_REFERENCE (default)  what the reference binary computes: parity with the reference's own CPU path
 *   ORBX_PATTERN_CANONICAL            the canonical table (the one ORBvoc.txt was trained with) */
#define ORBX_PATTERN_REFERENCE 0
#define ORBX_PATTERN_CANONICAL 1
int orbx_extractor_set_pattern(orbx_extractor* ex, int variant);
/* the stream the handle launches on (cudaStream_t as void*) */
void* orbx_extractor_stream(const orbx_extractor* ex);

/* ---- per-stage timing (the reference's GetTime rows, src/ORBextractor.cc:1876-1904; stage names as at :1841,
 * :1131, :892, :1737, :1331/:1753) ----  With timing on, CUDA events bracket every stage of the next extract call
 * on the stream it runs on; orbx_extractor_stage_ms synchronises and returns the milliseconds of the last call. */
#define ORBX_NUM_STAGES 5
int orbx_extractor_enable_timing(orbx_extractor* ex, int on);
int orbx_extractor_stage_ms(orbx_extractor* ex, float* ms /* ORBX_NUM_STAGES */);
const char* orbx_stage_name(int stage);

/* ---- ORBextractor::operator()  (src/ORBextractor.cc:1538-1548, 1701-1815, CPU path semantics) ----
 * One 8-bit frame in host memory -> keypoints (level-major, level-0 coordinates) + n x 32 descriptor bytes.
 * cap = capacity of kps/desc in keypoints; *n = number produced (<= cap, else ORBX_E_OVERFLOW).
 * An empty image (img NULL or w,h <= 0) returns ORBX_OK with *n = 0, like the reference's silent return. */
int orbx_extract(orbx_extractor* ex, const uint8_t* img, int w, int h, size_t stride, orbx_keypoint* kps, uint8_t* desc,
                 int cap, int* n);

/* Batched form: `batch` frames of identical size, frame f at frames + f*frame_stride, rows `stride` bytes apart.
 * Outputs: kps[f*cap .. ], desc[(f*cap + i)*32 .. ], counts[f].  Host memory; copies are inside the call.
 * Batches of >= 128 frames are cut into chunks of about 64 frames pipelined over H2D | kernels (two lanes) | D2H streams;
 * pinned caller buffers are used directly, pageable ones go through pinned staging.  The device scratch then holds two
 * chunks: orbx_pyramid_level_host() afterwards serves the frames of the LAST chunk only (by their index in the batch). */
int orbx_extract_batch(orbx_extractor* ex, const uint8_t* frames, int batch, int w, int h, size_t stride,
                       size_t frame_stride, orbx_keypoint* kps, uint8_t* desc, int cap, int* counts);

/* Offline sequences (BASELINE config 5): the same call split into "enqueue" and "collect", so that consecutive batches
 * overlap — the host->device copies of batch k + 1 run while the kernels and device->host copies of batch k are still
 * going.  All buffers must be pinned host memory (cudaHostAlloc / cudaHostRegister) and the frames tightly packed
 * (stride == w, frame_stride == w*h); they stay owned by the call until orbx_extract_batch_wait returns for it, so two
 * calls in flight need two sets of output buffers.  At most two calls may be in flight; orbx_extract_batch_wait collects
 * the OLDEST one and returns its status (ORBX_OK / ORBX_E_OVERFLOW / ORBX_E_CUDA).  The reference has no counterpart:
 * its operator() is one synchronous frame (src/ORBextractor.cc:1538). */
int orbx_extract_batch_async(orbx_extractor* ex, const uint8_t* frames, int batch, int w, int h, size_t stride,
                             size_t frame_stride, orbx_keypoint* kps, uint8_t* desc, int cap, int* counts);
int orbx_extract_batch_wait(orbx_extractor* ex);

/* Same with every buffer already resident in device memory; asynchronous on `stream` (NULL = the handle's own
 * stream).  d_status (device int, may be NULL) receives 0 or ORBX_E_OVERFLOW. */
int orbx_extract_batch_device(orbx_extractor* ex, const uint8_t* d_frames, int batch, int w, int h, size_t stride,
                              size_t frame_stride, orbx_keypoint* d_kps, uint8_t* d_desc, int cap, int* d_counts,
                              int* d_status, void* stream);

/* ---- mvImagePyramid (include/ORBextractor.h:116; read by src/Frame.cc:472,562,574,579) ----
 * Level `level` of frame `frame` of the last extract call, copied lazily to host memory owned by the handle.
 * *interior points at pixel (0,0) of the level; it is the centre ROI of a (w+38) x (h+38) parent with row pitch
 * *pitch, so rows -19..h+18 and columns -19..w+18 are readable (the EDGE_THRESHOLD reflect-101 border). */
int orbx_pyramid_level_host(orbx_extractor* ex, int frame, int level, const uint8_t** interior, int* w, int* h,
                            size_t* pitch);
/* level sizes of the current geometry */
int orbx_level_size(const orbx_extractor* ex, int level, int* w, int* h);

/* ---- stage taps for parity tests (not used by the product path) ---- */
/* blurred level (7x7 sigma 2), tightly packed w*h bytes into out */
int orbx_debug_blurred_level(orbx_extractor* ex, int frame, int level, uint8_t* out);
/* FAST candidates of one level in the reference's vToDistributeKeys order (x,y relative to minBorder) */
int orbx_debug_candidates(orbx_extractor* ex, int frame, int level, orbx_keypoint* out, int cap, int* n);
/* keypoints kept by the quadtree for one level, level coordinates, with angle */
int orbx_debug_level_keypoints(orbx_extractor* ex, int frame, int level, orbx_keypoint* out, int cap, int* n);
/* stand-alone DistributeOctTree (src/ORBextractor.cc:889-1120) on caller-supplied candidates */
int orbx_debug_distribute(const orbx_keypoint* in, int n, int minX, int maxX, int minY, int maxY, int N,
                          orbx_keypoint* out, int cap, int* nout);

/* ---- ORBmatcher (src/ORBmatcher.cc) ---- */
/* DescriptorDistance (:1647-1663) for n pairs a[i] vs b[i] on the device; host buffers */
int orbx_hamming_distance(const uint8_t* a, const uint8_t* b, int n, int* dist);

/* Brute-force best / second-best + acceptance test (inner loop of SearchByBoW, :187-232):
 * idx[i] = first database row at minimal distance, d1/d2 = best / second-best distance (256 if none),
 * match[i] = idx[i] if d1 <= th && (float)d1 < ratio*(float)d2 else -1.  Host buffers. */
int orbx_match_bruteforce(const uint8_t* q, int nq, const uint8_t* d, int nd, int th, float ratio, int* idx, int* d1,
                          int* d2, int* match);
int orbx_match_bruteforce_device(const uint8_t* d_q, int nq, const uint8_t* d_d, int nd, int th, float ratio,
                                 int* d_idx, int* d_d1, int* d_d2, int* d_match, void* stream);

/* The same matcher for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same kernel, same
 * results as orbx_match_bruteforce_device, but its scratch (per-slice partials + arrival counters) is a file-scope
 * __device__ array inside liborbx.so (4 MiB + 32 KiB per GPU, held from library load until process exit), so the call
 * performs no cudaMalloc / cudaFree and creates no stream, event or workspace.  Limits: nq <= ORBX_NOALLOC_MAX_QUERIES,
 * otherwise ORBX_E_INVALID with a message naming the limit (nd as above).  All calls on one device share the scratch:
 * calls that overlap on the GPU (different streams or threads) RACE on it — serialise them.  The first call on each device
 * resolves the array's device address (cudaGetSymbolAddress, once); stream NULL = the default stream.  Device buffers,
 * asynchronous on `stream`. */
#define ORBX_NOALLOC_MAX_QUERIES 8192
int orbx_match_bruteforce_noalloc(const uint8_t* d_q, int nq, const uint8_t* d_d, int nd, int th, float ratio,
                                  int* d_idx, int* d_d1, int* d_d2, int* d_match, void* stream);

/* All-vs-all keyframe matching (relocalisation / loop-candidate workload; N x N DescriptorDistance tables as in
 * src/MapPoint.cc:280-305 generalised to keyframe pairs).  Query keyframes [kq] and database keyframes [kd] hold
 * n descriptors each (n*32 bytes per keyframe).  score[i*kd + j] = number of descriptors of query keyframe i whose
 * best match in database keyframe j passes d1 <= th && d1 < ratio*d2.  Device buffers, asynchronous. */
int orbx_match_allpairs_device(const uint8_t* d_q, int kq, const uint8_t* d_db, int kd, int n, int th, float ratio,
                               int* d_score, void* stream);

/* The same workload over `world` GPUs, one process per GPU (BASELINE config 5; the reference is single-GPU):
 * the k_total keyframes are split into contiguous blocks (block sizes differ by at most one, earlier ranks larger; rank r
 * holds the descriptors of its block in d_local [k_local][n][32]) and every rank scores its block of QUERY keyframes against
 * all k_total keyframes: d_scores [k_local][k_total].  The database is all-gathered over NCCL in `chunks` pieces on an
 * internal stream; the rank's own block is scored while the first piece is in flight and every further piece while the
 * next one arrives.  nccl_comm = the caller's ncclComm_t for these `world` ranks (ignored when world == 1); NCCL is resolved
 * at run time from the library the process has loaded (libnccl.so.2, or ORBX_NCCL_LIB).  Asynchronous on `stream`.
 * gathered_bytes (may be NULL) receives the bytes this rank received. */
int orbx_allpairs_sharded(void* nccl_comm, int rank, int world, const uint8_t* d_local, int k_total, int n, int th, float ratio,
                          int* d_scores, int chunks, void* stream, unsigned long long* gathered_bytes);

/* SearchForInitialization (:405-520): frame-1 keypoints of octave 0 search a +-window box around prev_xy[i] in
 * frame 2 (Frame::GetFeaturesInArea over the 64x48 grid, src/Frame.cc:326-379), ratio test, one-to-one stealing,
 * rotation histogram.  bounds = {mnMinX, mnMaxX, mnMinY, mnMaxY}.  prev_xy (2*n1 floats) is updated in place like
 * vbPrevMatched; matches12[n1] = index into frame 2 or -1; *nmatches = return value.  Host buffers. */
int orbx_search_for_initialization(const orbx_keypoint* k1, int n1, const uint8_t* desc1, const orbx_keypoint* k2,
                                   int n2, const uint8_t* desc2, const float bounds[4], float* prev_xy, int* matches12,
                                   int window, float nnratio, int check_ori, int* nmatches);

/* SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1328-1470): the per-frame tracking matcher
 * (src/Tracking.cc:962,968).  The boundary sits after the pose arithmetic, which stays the caller's cv::Mat expressions:
 *   x3dc[3*i..]  = Rcw*x3Dw+tcw of LastFrame.mvpMapPoints[i] (:1362-1363);  direction = 0 / 1 (bForward) / 2 (bBackward), :1348-1349
 *   valid1[i]    = mvpMapPoints[i] != NULL && !mvbOutlier[i];  obs1[i] = that point's Observations() > 0
 *   mp_desc1     = pMP->GetDescriptor() (32 bytes per point);  octave1 = LastFrame.mvKeys[i].octave;  angle1 = mvKeysUn[i].angle
 *   k2, desc2    = CurrentFrame.mvKeysUn / mDescriptors;  u_right2 = mvuRight (NULL for monocular frames)
 *   occupied2[j] = CurrentFrame.mvpMapPoints[j] != NULL && ...->Observations() > 0 on entry
 *   cam = {fx, fy, cx, cy};  bounds = {mnMinX, mnMaxX, mnMinY, mnMaxY};  scale_factors = mvScaleFactors
 * match21[j] = i: CurrentFrame.mvpMapPoints[j] = LastFrame.mvpMapPoints[i];  -1: untouched;  -2: set to NULL by the rotation
 * check (:1451-1467).  *nmatches = return value.  Host buffers. */
int orbx_search_by_projection(int n1, const float* x3dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mp_desc1,
                              const int* octave1, const float* angle1, const orbx_keypoint* k2, int n2, const uint8_t* desc2,
                              const float* u_right2, const uint8_t* occupied2, const float cam[4], float mbf,
                              const float bounds[4], const float* scale_factors, int nlevels, float th, int direction,
                              int check_ori, int* match21, int* nmatches);

/* Same kernels with the knobs of the relocalisation variant SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist)
 * (src/ORBmatcher.cc:1472-1599, called at src/Tracking.cc:1445,1484): orb_dist replaces TH_HIGH (:1552); flag
 * ORBX_SBP_NO_DEPTH_TEST drops the `invzc < 0` test that only the frame-to-frame routine has (:1369); there
 *   valid1[i] = pMP && !pMP->isBad() && !sAlreadyFound.count(pMP) && minDistance <= dist3D <= maxDistance  (:1489-1521)
 *   octave1[i] = pMP->PredictScale(dist3D, &CurrentFrame) (:1523);  angle1[i] = pKF->mvKeysUn[i].angle;  obs1[i] = 1 and
 *   occupied2[j] = CurrentFrame.mvpMapPoints[j] != NULL (any assigned keypoint is skipped, :1540-1541);  direction = 0. */
#define ORBX_SBP_NO_DEPTH_TEST 1
int orbx_search_by_projection_ex(int n1, const float* x3dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mp_desc1,
                                 const int* octave1, const float* angle1, const orbx_keypoint* k2, int n2, const uint8_t* desc2,
                                 const float* u_right2, const uint8_t* occupied2, const float cam[4], float mbf,
                                 const float bounds[4], const float* scale_factors, int nlevels, float th, int direction,
                                 int check_ori, int orb_dist, int flags, int* match21, int* nmatches);

/* SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, th) (src/ORBmatcher.cc:45-129, RadiusByViewingCos
 * :131-137): the local-map matcher of Tracking::SearchLocalPoints (src/Tracking.cc:1270-1277).
 *   valid1[i] = pMP->mbTrackInView && !pMP->isBad();  proj[3*i..] = mTrackProjX, mTrackProjY, mTrackProjXR;
 *   view_cos = mTrackViewCos;  level1 = mnTrackScaleLevel;  obs1, mp_desc1, k2, desc2, u_right2, occupied2, bounds,
 *   scale_factors as for orbx_search_by_projection;  nnratio = mfNNratio.
 * match21[j] = i: F.mvpMapPoints[j] = vpMapPoints[i];  -1: untouched.  *nmatches = return value.  Host buffers. */
int orbx_search_by_projection_map(int n1, const float* proj, const float* view_cos, const int* level1, const uint8_t* valid1,
                                  const uint8_t* obs1, const uint8_t* mp_desc1, const orbx_keypoint* k2, int n2,
                                  const uint8_t* desc2, const float* u_right2, const uint8_t* occupied2, const float bounds[4],
                                  const float* scale_factors, int nlevels, float th, float nnratio, int* match21, int* nmatches);

/* SearchByProjection(KeyFrame* pKF, cv::Mat Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:290-403; loop detection,
 * src/LoopClosing.cc:278) from the radius search on (:357-398); the Sim3 decomposition and the visibility tests (:298-355) stay
 * the caller's.  uvr[3*i..] = u, v, th*pKF->mvScaleFactors[nPredictedLevel];  level1[i] = nPredictedLevel;  valid1[i] = the point
 * passed every `continue` of :318-351;  k2, desc2 = pKF->mvKeysUn / mDescriptors;  occupied2[j] = vpMatched[j] != NULL;
 * th_dist = TH_LOW.  match21[j] = i: vpMatched[j] = vpPoints[i];  -1: untouched.  Host buffers. */
int orbx_search_by_projection_sim3(int n1, const float* uvr, const int* level1, const uint8_t* valid1, const uint8_t* mp_desc1,
                                   const orbx_keypoint* k2, int n2, const uint8_t* desc2, const uint8_t* occupied2,
                                   const float bounds[4], int th_dist, int* match21, int* nmatches);

/* ORBmatcher::SearchByBoW, both overloads: (KeyFrame* pKF, Frame &F, vpMapPointMatches) (src/ORBmatcher.cc:159-288;
 * Tracking::TrackReferenceKeyFrame, src/Tracking.cc:842, and Relocalization, :1465) and (KeyFrame* pKF1, KeyFrame* pKF2,
 * vpMatches12) (:522-655; LoopClosing::ComputeSim3, src/LoopClosing.cc:304).  The caller walks the two DBoW2 FeatureVectors
 * (:176-180 / :541-545) and hands over the n_nodes COMMON nodes in map order: node k holds the feature indices
 * idx1[start1[k] .. start1[k+1]) of the first and idx2[start2[k] .. start2[k+1]) of the second argument, each in vector order.
 *   valid1[i] = first argument's map point i exists and !isBad() (:192-198 / :556-562)
 *   valid2[j] = 1 for the Frame overload;  pMP2 exists and !isBad() for the KeyFrame overload (:582-587)
 *   desc*, angle* = mDescriptors rows and mvKeysUn[.].angle of the two arguments
 *   th_dist = TH_LOW for the Frame overload (`bestDist1 <= TH_LOW`, :230), TH_LOW - 1 for the KeyFrame overload
 *             (`bestDist1 < TH_LOW`, :605);  nnratio = mfNNratio;  check_ori = mbCheckOrientation
 * match12[i] = j and match21[j] = i for every accepted pair (Frame overload: vpMapPointMatches[j] = vpMapPointsKF[i];
 * KeyFrame overload: vpMatches12[i] = vpMapPoints2[j]);  -1: untouched;  -2: assigned, then removed by the rotation
 * histogram (:262-285 / :631-652) — the slot still blocked later features while the loop ran.  *nmatches = return value.
 * Host buffers; n2 <= 65535 and at most 65535 second-argument features per node. */
int orbx_search_by_bow(int n_nodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                       const uint8_t* valid1, const uint8_t* desc1, const float* angle1, int n2, const uint8_t* valid2,
                       const uint8_t* desc2, const float* angle2, int th_dist, float nnratio, int check_ori, int* match12,
                       int* match21, int* nmatches);

/* cvtColor(im, im, CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY) of Tracking::GrabImageStereo / RGBD / Monocular
 * (src/Tracking.cc:221-246, :261-274, :294-307): channels = 3 or 4 (alpha ignored), 8 bit, strides in bytes.
 *   ORBX_GRAY_BGR          the first channel is blue (mbRGB == false)
 *   ORBX_GRAY_CV2_COEFFS   OpenCV 2.4 / 3.x fixed point (4899, 9617, 1868, >> 14); default is OpenCV 4.x (9798, 19235, 3735, >> 15)
 * The host form copies in and out; the device form converts device buffers on `stream` so that the gray frame feeds
 * orbx_extract_device without leaving the GPU. */
#define ORBX_GRAY_BGR 1
#define ORBX_GRAY_CV2_COEFFS 2
int orbx_cvt_color_gray(const uint8_t* src, long long src_stride, int w, int h, int channels, int flags, uint8_t* dst, long long dst_stride);
int orbx_cvt_color_gray_device(const uint8_t* d_src, long long src_stride, int w, int h, int channels, int flags, uint8_t* d_dst,
                               long long dst_stride, void* stream);

/* ---- a frame that stays on the device between matcher calls (Frame::AssignFeaturesToGrid, src/Frame.cc:229-244) ----
 * Tracking runs several matchers against the CURRENT frame (TrackWithMotionModel's SearchByProjection, src/Tracking.cc:962,968;
 * SearchLocalPoints' :1270-1277).  orbx_frame_create uploads its undistorted keypoints (mvKeysUn), descriptors and right
 * coordinates (mvuRight, may be NULL) once and bins the 64 x 48 grid (mGrid) on the device; the *_frame matcher entries take the
 * handle where the host-array entries take k2 / n2 / desc2 / u_right2 / bounds, and skip the uploads and the binning.  Results are
 * identical.  orbx_frame_grid_host copies the grid out for inspection: cell (ix, iy) = cell_start[ix*48 + iy] ..
 * cell_start[ix*48 + iy + 1] of `sorted` (keypoint indices in mGrid's push_back order). */
typedef struct orbx_frame orbx_frame;
int orbx_frame_create(const orbx_keypoint* kps_un, int n, const uint8_t* desc, const float* u_right, const float bounds[4],
                      orbx_frame** out);
void orbx_frame_destroy(orbx_frame* f);
int orbx_frame_size(const orbx_frame* f);
int orbx_frame_grid_host(const orbx_frame* f, int* cell_start /* 64*48 + 1 */, int* sorted /* n */);
int orbx_search_by_projection_frame(int n1, const float* x3dc, const uint8_t* valid1, const uint8_t* obs1, const uint8_t* mp_desc1,
                                    const int* octave1, const float* angle1, const orbx_frame* f2, const uint8_t* occupied2,
                                    const float cam[4], float mbf, const float* scale_factors, int nlevels, float th, int direction,
                                    int check_ori, int orb_dist, int flags, int* match21, int* nmatches);
int orbx_search_by_projection_map_frame(int n1, const float* proj, const float* view_cos, const int* level1, const uint8_t* valid1,
                                        const uint8_t* obs1, const uint8_t* mp_desc1, const orbx_frame* f2, const uint8_t* occupied2,
                                        const float* scale_factors, int nlevels, float th, float nnratio, int* match21, int* nmatches);

/* Frame::UndistortKeyPoints (src/Frame.cc:403-433): cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) on the keypoint
 * coordinates — OpenCV's five-iteration inverse of the k1 k2 p1 p2 [k3] model in double, re-projected with the same K — every
 * other field of the keypoint copied (mvKeysUn[i] = mvKeys[i] with the new pt).  K = {fx, fy, cx, cy}; dist = ndist (0..5)
 * coefficients in OpenCV's order.  dist[0] == 0 (or ndist == 0) copies the keypoints, like the reference's early return
 * (:405-409); the device form always runs the model.  in == out is allowed. */
int orbx_undistort_keypoints(const orbx_keypoint* in, int n, const float K[4], const float* dist, int ndist, orbx_keypoint* out);
int orbx_undistort_keypoints_device(const orbx_keypoint* d_in, int n, const float K[4], const float* dist, int ndist,
                                    orbx_keypoint* d_out, void* stream);

/* The matching core of both ORBmatcher::Fuse overloads: Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:825-975; candidate loop
 * :886-945; LocalMapping::SearchInNeighbors, src/LocalMapping.cc:600ff) and Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)
 * (:977-1100; candidate loop :1049-1075; LoopClosing::SearchAndFuse).  Per map point i that passed the caller's projection and
 * visibility tests (valid1[i]): uvr[3i..] = u, v, radius = th * mvScaleFactors[nPredictedLevel];  level1[i] = nPredictedLevel;
 * ur1[i] = u - bf * invz and inv_level_sigma2 = pKF->mvInvLevelSigma2 switch the chi-square gate of the first overload on
 * (:904-932; ur1 = NULL for the Sim3 overload, which has none);  k2, desc2, u_right2 = pKF->mvKeysUn / mDescriptors / mvuRight
 * (u_right2 may be NULL: no stereo).  best_idx[i] / best_dist[i] = bestIdx / bestDist of the loop (-1 / 256 when nothing
 * qualified); the caller applies `bestDist <= TH_LOW` and the map surgery (:948-972 / :1078-1092), which walks pointers and whose
 * order matters, in its own loop.  Host buffers. */
int orbx_fuse_match(int n1, const float* uvr, const float* ur1, const int* level1, const uint8_t* valid1, const uint8_t* mp_desc1,
                    const orbx_keypoint* k2, int n2, const uint8_t* desc2, const float* u_right2, const float* inv_level_sigma2,
                    int nlevels, const float bounds[4], int* best_idx, int* best_dist);

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:657-823 with
 * CheckDistEpipolarLine :140-157; LocalMapping::CreateNewMapPoints, src/LocalMapping.cc:248ff).  Common vocabulary nodes as for
 * orbx_search_by_bow.  k*, desc* = mvKeysUn / mDescriptors of the two keyframes;
 *   valid1[i] = !pKF1->GetMapPoint(i) && (!bOnlyStereo || mvuRight[i] >= 0)  (:697-706),  stereo1[i] = mvuRight[i] >= 0;
 *   valid2[j], stereo2[j] likewise for pKF2 (:724-733);  f12 = F12 row major;  (ex, ey) = the epipole in image 2 (:664-672);
 *   scale_factors2 / level_sigma2_2 = pKF2->mvScaleFactors / mvLevelSigma2.
 * match12[i] = j: the pair (i, j) of vMatchedPairs (which lists them in ascending i);  -1: none.  *nmatches = return value.
 * The reference never sets vbMatched2 (:683), so a second-keyframe feature may be paired with several first-keyframe ones; that is
 * reproduced.  Host buffers. */
int orbx_search_for_triangulation(int n_nodes, const int* start1, const uint32_t* idx1, const int* start2, const uint32_t* idx2, int n1,
                                  const orbx_keypoint* k1, const uint8_t* desc1, const uint8_t* valid1, const uint8_t* stereo1, int n2,
                                  const orbx_keypoint* k2, const uint8_t* desc2, const uint8_t* valid2, const uint8_t* stereo2,
                                  const float f12[9], float ex, float ey, const float* scale_factors2, const float* level_sigma2_2,
                                  int nlevels, int check_ori, int* match12, int* nmatches);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:246-312; called after every new observation, src/LocalMapping.cc:
 * 137,356,373,…, src/Tracking.cc:619,…) for a batch of map points: point p owns the descriptors desc[32 * start[p] ..
 * 32 * start[p+1]) = the rows pKF->mDescriptors.row(idx) of its good observations in map order (:266-273).
 * best_idx[p] = BestIdx within the point's rows (the row that becomes mDescriptor), -1 for a point without rows (the reference
 * returns early, :275-276); best_median[p] (may be NULL) = its median distance.  Host buffers. */
int orbx_distinctive_descriptors(int n_points, const int* start, const uint8_t* desc, int* best_idx, int* best_median);

/* ---- DBoW2 vocabulary: Frame::ComputeBoW (src/Frame.cc:394-401) -------------------------------------------------------
 * The tree of TemplatedVocabulary<FORB::TDescriptor, FORB> (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h) in device memory
 * and the descent of transform(feature, word_id, weight, nid, levelsup) (:1218-1261) for a whole frame in one launch.  The
 * std::map bookkeeping around it (BowVector::addWeight / normalize, FeatureVector::addFeature, :1146-1193) is host code in
 * include/ORBVocabulary.h. */
typedef struct orbx_vocabulary orbx_vocabulary;

/* Nodes as loadFromTextFile numbers them (:1338-1424): node 0 = root (its parent / is_leaf / desc / weight entries are ignored),
 * node i >= 1 = line i+1 of the text file: parent[i] < i, is_leaf[i] (gets the next word id), desc[32 i ..], weight[i].
 * k, L, scoring, weighting = the four integers of the first line (m_k, m_L, ScoringType, WeightingType). */
int orbx_vocabulary_create(int k, int L, int scoring, int weighting, int n_nodes, const int32_t* parent, const uint8_t* is_leaf,
                           const uint8_t* desc, const double* weight, orbx_vocabulary** out);
/* ORBvocabulary::loadFromTextFile (:1338-1424; called at src/System.cc:64-65).  Blank lines make no node: on the empty
 * remainder after the file's last newline the reference reads `pid` uninitialised (undefined behaviour), which is not
 * reproduced.  ORBX_E_INVALID for a file that is not a DBoW2 text vocabulary. */
int orbx_vocabulary_load_text(const char* path, orbx_vocabulary** out);
void orbx_vocabulary_destroy(orbx_vocabulary* v);
int orbx_vocabulary_info(const orbx_vocabulary* v, int* k, int* L, int* scoring, int* weighting, int* n_nodes, int* n_words);
/* For every row of desc[n x 32]: word_id / weight of the leaf the descent ends in and node_id = the node passed at level
 * L - levelsup (0 when that level is <= 0; ORBX_VOC_NO_NODE when the descent ends above it — the reference leaves *nid
 * unwritten there).  Host buffers. */
#define ORBX_VOC_NO_NODE 0xFFFFFFFFu
int orbx_vocabulary_transform(orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup, uint32_t* word_id, double* weight,
                              uint32_t* node_id);

/* Frame::ComputeStereoMatches (src/Frame.cc:465-639).  exL / exR hold the pyramids of the left / right frame of
 * their last orbx_extract call (device resident; nothing is copied to the host).  bf = mbf (baseline x fx),
 * mb = baseline in metres = mbf / fx (src/Frame.cc:113).  u_right[nL], depth[nL] as mvuRight / mvDepth
 * (-1 = no match).  Host buffers. */
int orbx_stereo_match(orbx_extractor* exL, orbx_extractor* exR, const orbx_keypoint* kL, int nL, const uint8_t* descL,
                      const orbx_keypoint* kR, int nR, const uint8_t* descR, float bf, float mb, float* u_right,
                      float* depth);
/* The same on the keypoints and descriptors the two extractors produced in their last orbx_extract call, as they still
 * sit on the device — what the Frame constructor does (src/Frame.cc:77-80 extracts both eyes, :94 calls
 * ComputeStereoMatches on exactly those mvKeys / mvKeysRight / descriptors): nothing is uploaded.  cap = capacity of
 * u_right / depth (>= the left extractor's last keypoint count, else ORBX_E_INVALID).  Any other extractor call in between
 * (a batch, a device-buffer call) makes it fail with ORBX_E_INVALID rather than read stale data. */
int orbx_stereo_match_last(orbx_extractor* exL, orbx_extractor* exR, float bf, float mb, float* u_right, float* depth, int cap);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
