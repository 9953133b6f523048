/* Fill, C API (reference include/c/vkt/Fill.h:16-32). */
#ifndef VKT_FILL_H
#define VKT_FILL_H

#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

VKTAPI vktError vktFillSV(vktStructuredVolume volume, float value);
VKTAPI vktError vktFillRangeSV(vktStructuredVolume volume, int32_t firstX, int32_t firstY, int32_t firstZ,
                               int32_t lastX, int32_t lastY, int32_t lastZ, float value);

#ifdef __cplusplus
}
#endif
#endif
