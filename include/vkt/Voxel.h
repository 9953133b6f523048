/* Voxel view + scalar codec, C API (reference include/c/vkt/Voxel.h:14-38). */
#ifndef VKT_VOXEL_H
#define VKT_VOXEL_H

#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct
{
    uint8_t* bytes;
    vktDataFormat dataFormat;
    float mappingLo;
    float mappingHi;
} vktVoxelView_t;

VKTAPI vktError vktMapVoxel(uint8_t* dst, float value, vktDataFormat dataFormat, float mappingLo, float mappingHi);
VKTAPI vktError vktUnmapVoxel(float* value, uint8_t const* src, vktDataFormat dataFormat, float mappingLo, float mappingHi);

#ifdef __cplusplus
}
#endif
#endif
