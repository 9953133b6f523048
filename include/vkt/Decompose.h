/* BrickDecompose, C API (reference include/c/vkt/Decompose.h:22-44). */
#ifndef VKT_DECOMPOSE_H
#define VKT_DECOMPOSE_H

#include <stdint.h>

#include "Array3D.h"
#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

VKTAPI vktError vktBrickDecomposeSV(vktArray3D_vktStructuredVolume decomp, vktStructuredVolume source,
                                    int32_t brickSizeX, int32_t brickSizeY, int32_t brickSizeZ,
                                    int32_t haloSizeNegX, int32_t haloSizeNegY, int32_t haloSizeNegZ,
                                    int32_t haloSizePosX, int32_t haloSizePosY, int32_t haloSizePosZ);
/* resizes the array and creates one volume per slot (handles already in the slots are overwritten, as in the reference) */
VKTAPI vktError vktBrickDecomposeResizeSV(vktArray3D_vktStructuredVolume decomp, vktStructuredVolume source,
                                          int32_t brickSizeX, int32_t brickSizeY, int32_t brickSizeZ,
                                          int32_t haloSizeNegX, int32_t haloSizeNegY, int32_t haloSizeNegZ,
                                          int32_t haloSizePosX, int32_t haloSizePosY, int32_t haloSizePosZ);

#ifdef __cplusplus
}
#endif
#endif
