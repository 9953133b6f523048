/* Transform, C API (reference include/c/vkt/Transform.h:18-56). */
#ifndef VKT_TRANSFORM_H
#define VKT_TRANSFORM_H

#include "Voxel.h"
#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef void (*vktTransformUnaryOp)(int32_t x, int32_t y, int32_t z, vktVoxelView_t voxel);
typedef void (*vktTransformBinaryOp)(int32_t x1, int32_t y1, int32_t z1, vktVoxelView_t voxel1, vktVoxelView_t voxel2);

VKTAPI vktError vktTransformSV1(vktStructuredVolume volume, vktTransformUnaryOp unaryOp);
VKTAPI vktError vktTransformSV2(vktStructuredVolume volume1, vktStructuredVolume volume2, vktTransformBinaryOp binaryOp);
VKTAPI vktError vktTransformRangeSV1(vktStructuredVolume volume, int32_t firstX, int32_t firstY, int32_t firstZ,
                                     int32_t lastX, int32_t lastY, int32_t lastZ, vktTransformUnaryOp unaryOp);
VKTAPI vktError vktTransformRangeSV2(vktStructuredVolume volume1, vktStructuredVolume volume2,
                                     int32_t firstX, int32_t firstY, int32_t firstZ,
                                     int32_t lastX, int32_t lastY, int32_t lastZ, vktTransformBinaryOp binaryOp);
/* volkit-b200 extension: bind a host op to the built-in device op computing the same function (GPU policy) */
VKTAPI vktError vktBindDeviceOp1(vktTransformUnaryOp hostOp, char const* deviceOpName);
VKTAPI vktError vktBindDeviceOp2(vktTransformBinaryOp hostOp, char const* deviceOpName);

#ifdef __cplusplus
}
#endif
#endif
