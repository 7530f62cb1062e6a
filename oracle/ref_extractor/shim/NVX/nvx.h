// oracle/ref_extractor/shim/NVX/nvx.h — TEST INFRASTRUCTURE.  A do-nothing OpenVX / VisionWorks so that the reference's
// src/ORBextractor.cc (compiled where it lies by oracle/ref_extractor/Makefile) builds in an image without VisionWorks: every
// type is an opaque pointer, every call returns a non-null dummy / VX_SUCCESS and touches nothing.  The reference's constructor
// runs buildGraph() unconditionally (src/ORBextractor.cc:559); with these stubs that is a sequence of no-ops, and operator()
// takes its CPU branches (all b_*_VX switches are false without -DOPENVX, :77-90).  Only the enum VALUES that reach arithmetic
// matter: VX_FAILURE = -1 (OpenVX 1.x vx_status_e), which the reference's bit_pattern_31_ picks up at :261.
#ifndef ORBX_REF_SHIM_NVX_H
#define ORBX_REF_SHIM_NVX_H
#include <stddef.h>
#include <stdint.h>

typedef struct _vx_ref* vx_reference;
typedef struct _vx_ctx* vx_context;
typedef struct _vx_graph* vx_graph;
typedef struct _vx_node* vx_node;
typedef struct _vx_image* vx_image;
typedef struct _vx_array* vx_array;
typedef struct _vx_scalar* vx_scalar;
typedef struct _vx_pyramid* vx_pyramid;
typedef struct _vx_conv* vx_convolution;
typedef struct _vx_matrix* vx_matrix;
typedef struct _vx_param* vx_parameter;
typedef struct _vx_kernel* vx_kernel;
typedef int vx_status, vx_enum, vx_bool;
typedef uint32_t vx_uint32, vx_df_image;
typedef int32_t vx_int32;
typedef uint64_t vx_uint64;
typedef int16_t vx_int16;
typedef uint8_t vx_uint8;
typedef float vx_float32;
typedef size_t vx_size, vx_map_id;
enum vx_status_e { VX_ERROR_NOT_IMPLEMENTED = -2, VX_FAILURE = -1, VX_SUCCESS = 0 };
enum { vx_false_e = 0, vx_true_e = 1 };
enum {
    VX_TYPE_INT32 = 1, VX_TYPE_UINT32, VX_TYPE_FLOAT32, VX_TYPE_SIZE, VX_TYPE_KEYPOINT, VX_TYPE_UINT8, VX_TYPE_INT16, VX_DF_IMAGE_U8 = 100,
    VX_READ_ONLY = 200, VX_WRITE_ONLY, VX_READ_AND_WRITE, VX_MEMORY_TYPE_HOST, VX_ARRAY_NUMITEMS, VX_ARRAY_CAPACITY, VX_IMAGE_WIDTH,
    VX_IMAGE_HEIGHT, VX_INTERPOLATION_BILINEAR, VX_NODE_ATTRIBUTE_PERFORMANCE, VX_NODE_PERFORMANCE, NVX_DIRECTIVE_ENABLE_PERFORMANCE
};
#define VX_SCALE_PYRAMID_ORB 0.8408964f
struct vx_keypoint_t { int32_t x, y; float strength, scale, orientation; int32_t tracking_status; float error; };
struct vx_perf_t { uint64_t tmp, beg, end, sum, avg, min, num, max; };
#define vxArrayItem(type, ptr, index, stride) (*(type*)((uint8_t*)(ptr) + (index) * (stride)))

namespace orbx_vx_stub {
template <class T> inline T dummy() { static char storage[64]; return (T)(void*)storage; }
}
#define ORBX_VX_OBJ(T) orbx_vx_stub::dummy<T>()
inline vx_context vxCreateContext() { return ORBX_VX_OBJ(vx_context); }
inline vx_graph vxCreateGraph(vx_context) { return ORBX_VX_OBJ(vx_graph); }
inline vx_status vxDirective(vx_reference, vx_enum) { return VX_SUCCESS; }
inline vx_scalar vxCreateScalar(vx_context, vx_enum, const void*) { return ORBX_VX_OBJ(vx_scalar); }
inline vx_status vxCopyScalar(vx_scalar, void*, vx_enum, vx_enum) { return VX_SUCCESS; }
inline vx_array vxCreateArray(vx_context, vx_enum, vx_size) { return ORBX_VX_OBJ(vx_array); }
inline vx_array vxCreateVirtualArray(vx_graph, vx_enum, vx_size) { return ORBX_VX_OBJ(vx_array); }
inline vx_status vxQueryArray(vx_array, vx_enum, void* p, vx_size n) { for (vx_size i = 0; i < n; ++i) ((char*)p)[i] = 0; return VX_SUCCESS; }
inline vx_status vxTruncateArray(vx_array, vx_size) { return VX_SUCCESS; }
inline vx_status vxAddArrayItems(vx_array, vx_size, const void*, vx_size) { return VX_SUCCESS; }
inline vx_status vxMapArrayRange(vx_array, vx_size, vx_size, vx_map_id*, vx_size*, void** p, vx_enum, vx_enum, vx_uint32) { *p = nullptr; return VX_SUCCESS; }
inline vx_status vxUnmapArrayRange(vx_array, vx_map_id) { return VX_SUCCESS; }
inline vx_status vxCopyArrayRange(vx_array, vx_size, vx_size, vx_size, void*, vx_enum, vx_enum) { return VX_SUCCESS; }
inline vx_image vxCreateImage(vx_context, vx_uint32, vx_uint32, vx_df_image) { return ORBX_VX_OBJ(vx_image); }
inline vx_image vxCreateVirtualImage(vx_graph, vx_uint32, vx_uint32, vx_df_image) { return ORBX_VX_OBJ(vx_image); }
inline vx_status vxReleaseImage(vx_image*) { return VX_SUCCESS; }
// the reference derives mvScaleFactor[level] = width / v_width from the VisionWorks pyramid (:674-680); the stub reports 1 x 1 images
// and the harness restores the constructor's own tables afterwards (ref_extractor.cpp)
inline vx_status vxQueryImage(vx_image, vx_enum, void* p, vx_size) { *(vx_uint32*)p = 1; return VX_SUCCESS; }
inline vx_pyramid vxCreatePyramid(vx_context, vx_size, vx_float32, vx_uint32, vx_uint32, vx_df_image) { return ORBX_VX_OBJ(vx_pyramid); }
inline vx_pyramid vxCreateVirtualPyramid(vx_graph, vx_size, vx_float32, vx_uint32, vx_uint32, vx_df_image) { return ORBX_VX_OBJ(vx_pyramid); }
inline vx_image vxGetPyramidLevel(vx_pyramid, vx_uint32) { return ORBX_VX_OBJ(vx_image); }
inline vx_node vxGaussianPyramidNode(vx_graph, vx_image, vx_pyramid) { return ORBX_VX_OBJ(vx_node); }
inline vx_node vxScaleImageNode(vx_graph, vx_image, vx_image, vx_enum) { return ORBX_VX_OBJ(vx_node); }
inline vx_node vxConvolveNode(vx_graph, vx_image, vx_convolution, vx_image) { return ORBX_VX_OBJ(vx_node); }
inline vx_node vxFastCornersNode(vx_graph, vx_image, vx_scalar, vx_bool, vx_array, vx_scalar) { return ORBX_VX_OBJ(vx_node); }
inline vx_node nvxCopyImageNode(vx_graph, vx_image, vx_image) { return ORBX_VX_OBJ(vx_node); }
inline vx_status nvxuCopyImage(vx_context, vx_image, vx_image) { return VX_SUCCESS; }
inline vx_status vxuScaleImage(vx_context, vx_image, vx_image, vx_enum) { return VX_SUCCESS; }
inline vx_status vxSetReferenceName(vx_reference, const char*) { return VX_SUCCESS; }
inline vx_status vxVerifyGraph(vx_graph) { return VX_SUCCESS; }
inline vx_status vxProcessGraph(vx_graph) { return VX_SUCCESS; }
inline vx_status vxScheduleGraph(vx_graph) { return VX_SUCCESS; }
inline vx_status vxWaitGraph(vx_graph) { return VX_SUCCESS; }
inline vx_status vxQueryNode(vx_node, vx_enum, void* p, vx_size n) { for (vx_size i = 0; i < n; ++i) ((char*)p)[i] = 0; return VX_SUCCESS; }
inline vx_status vxReleaseGraph(vx_graph*) { return VX_SUCCESS; }
inline vx_status vxReleaseContext(vx_context*) { return VX_SUCCESS; }
inline vx_status vxReleaseArray(vx_array*) { return VX_SUCCESS; }
inline vx_status vxReleaseScalar(vx_scalar*) { return VX_SUCCESS; }
inline vx_status vxReleaseNode(vx_node*) { return VX_SUCCESS; }
inline vx_status vxReleasePyramid(vx_pyramid*) { return VX_SUCCESS; }
inline vx_status vxReleaseConvolution(vx_convolution*) { return VX_SUCCESS; }
#define NVXIO_SAFE_CALL(x) (x)
#define NVXIO_CHECK_REFERENCE(x) ((void)(x))
#endif
