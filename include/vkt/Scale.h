/* Scale, C API (reference include/c/vkt/Scale.h:15-18). */
#ifndef VKT_SCALE_H
#define VKT_SCALE_H

#include "common.h"
#include "linalg.h"

#ifdef __cplusplus
extern "C" {
#endif

VKTAPI vktError vktScaleSV(vktStructuredVolume dest, vktStructuredVolume source, vktVec3f_t scalingFactor,
                           vktVec3f_t centerOfScaling);

#ifdef __cplusplus
}
#endif
#endif
