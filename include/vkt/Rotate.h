/* Rotate, C API (reference include/c/vkt/Rotate.h:15-19). */
#ifndef VKT_ROTATE_H
#define VKT_ROTATE_H

#include "common.h"
#include "linalg.h"

#ifdef __cplusplus
extern "C" {
#endif

VKTAPI vktError vktRotateSV(vktStructuredVolume dest, vktStructuredVolume source, vktVec3f_t axis, float angleInRadians,
                            vktVec3f_t centerOfRotation);

#ifdef __cplusplus
}
#endif
#endif
