// TEST INFRASTRUCTURE (oracle/ref_extractor): stands in for the reference's include/CustomNodes.h (user kernels of its OpenVX graph —
// Jetson glue, out of scope).  src/ORBextractor.cc only registers the kernels and creates their nodes in buildGraph(); with the
// do-nothing OpenVX of NVX/nvx.h these are no-ops.
#ifndef ORBX_REF_SHIM_CUSTOMNODES_H
#define ORBX_REF_SHIM_CUSTOMNODES_H
#include <NVX/nvx.h>
#include <map>
#include <stdint.h>
typedef struct _descriptor_t { uint8_t descr[32]; } descriptor_t;  // 32-byte ORB descriptor item of the OpenVX branch (never filled here)
static vx_enum zvx_descriptor = 0;
inline vx_status registerMakeGrid(vx_context) { return VX_SUCCESS; }
inline vx_status registerMakeQuadtree(vx_context) { return VX_SUCCESS; }
inline vx_status registerComputeAngle(vx_context) { return VX_SUCCESS; }
inline vx_status registerORB(vx_context) { return VX_SUCCESS; }
inline vx_status registerScaleArray(vx_context) { return VX_SUCCESS; }
inline vx_status registerCustomKernels(vx_context) { return VX_SUCCESS; }
inline vx_node zvxMakeGridNode(vx_graph, vx_array, vx_scalar, vx_scalar, vx_scalar, vx_scalar, vx_scalar, vx_scalar, vx_scalar, vx_array) { return ORBX_VX_OBJ(vx_node); }
inline vx_node zvxMakeQuadtreeNode(vx_graph, vx_array, vx_scalar, vx_scalar, vx_scalar, vx_scalar, vx_scalar, vx_array) { return ORBX_VX_OBJ(vx_node); }
inline vx_node zvxComputeAngleNode(vx_graph, vx_image, vx_array, vx_array, vx_array) { return ORBX_VX_OBJ(vx_node); }
inline vx_node zvxORBNode(vx_graph, vx_image, vx_array, vx_array, vx_array) { return ORBX_VX_OBJ(vx_node); }
inline vx_node zvxScaleArrayNode(vx_graph, vx_array, vx_scalar, vx_array) { return ORBX_VX_OBJ(vx_node); }
inline vx_status zvxuMakeGrid(vx_context, vx_array, vx_uint32, vx_uint32, vx_uint32, vx_uint32, vx_uint32, vx_float32, vx_uint32, vx_array) { return VX_SUCCESS; }
inline vx_status zvxuMakeQuadtree(vx_context, vx_array, vx_uint32, vx_uint32, vx_uint32, vx_uint32, vx_size, vx_array) { return VX_SUCCESS; }
inline vx_status zvxuComputeAngle(vx_context, vx_image, vx_array, vx_array, vx_array) { return VX_SUCCESS; }
inline vx_status zvxuORB(vx_context, vx_image, vx_array, vx_array, vx_array) { return VX_SUCCESS; }
inline vx_status zvxuScaleArray(vx_context, vx_array, vx_float32, vx_array) { return VX_SUCCESS; }
inline vx_convolution createGaussianConvolution(vx_context, vx_size, vx_float32, vx_uint32, bool) { return ORBX_VX_OBJ(vx_convolution); }
#endif
