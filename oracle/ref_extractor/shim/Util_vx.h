// TEST INFRASTRUCTURE (oracle/ref_extractor): stands in for the reference's include/Util_vx.h (OpenVX <-> cv::Mat interop and
// debug printers of its VisionWorks path — Jetson glue, out of scope).  src/ORBextractor.cc names nvx_cv::createVXImageFromCVMat and
// nvx_cv::VXImageToCVMatMapper in its OpenVX branches only; they must compile, they never run (all b_*_VX switches are false).
#ifndef ORBX_REF_SHIM_UTIL_VX_H
#define ORBX_REF_SHIM_UTIL_VX_H
#include <NVX/nvx.h>
#include <opencv2/core/core.hpp>
#include <string>
inline void printReferenceInfo(vx_reference, std::string = "") {}
inline void printNodeInfo(vx_node, std::string = "") {}
namespace nvx_cv {
inline vx_image createVXImageFromCVMat(vx_context, const cv::Mat&) { return ORBX_VX_OBJ(vx_image); }
class VXImageToCVMatMapper {
public:
    VXImageToCVMatMapper(vx_image, vx_uint32 = 0, void* = nullptr, vx_enum = 0, vx_enum = 0) {}
    cv::Mat getMat() const { return cv::Mat(); }
};
}  // namespace nvx_cv
#endif
