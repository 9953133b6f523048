/* Copy, C API (reference include/c/vkt/Copy.h:16-32). */
#ifndef VKT_COPY_H
#define VKT_COPY_H

#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

VKTAPI vktError vktCopySV(vktStructuredVolume dst, vktStructuredVolume src);
VKTAPI vktError vktCopyRangeSV(vktStructuredVolume dst, vktStructuredVolume src,
                               int32_t firstX, int32_t firstY, int32_t firstZ,
                               int32_t lastX, int32_t lastY, int32_t lastZ,
                               int32_t dstOffsetX, int32_t dstOffsetY, int32_t dstOffsetZ);

#ifdef __cplusplus
}
#endif
#endif
