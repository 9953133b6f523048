/* Flip, C API (reference include/c/vkt/Flip.h:18-33). */
#ifndef VKT_FLIP_H
#define VKT_FLIP_H

#include "common.h"
#include "linalg.h"

#ifdef __cplusplus
extern "C" {
#endif

VKTAPI vktError vktFlipSV(vktStructuredVolume dest, vktStructuredVolume source, vktAxis axis);
VKTAPI vktError vktFlipRangeSV(vktStructuredVolume dest, vktStructuredVolume source,
                               int32_t firstX, int32_t firstY, int32_t firstZ,
                               int32_t lastX, int32_t lastY, int32_t lastZ,
                               int32_t dstOffsetX, int32_t dstOffsetY, int32_t dstOffsetZ, vktAxis axis);

#ifdef __cplusplus
}
#endif
#endif
