/* Resample, C API.  The reference has no C binding for Resample; this one mirrors the C++
 * signature (include/vkt/Resample.hpp). */
#ifndef VKT_RESAMPLE_H
#define VKT_RESAMPLE_H

#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { vktFilterModeNearest, vktFilterModeLinear } vktFilterMode;

VKTAPI vktError vktResampleSV(vktStructuredVolume dst, vktStructuredVolume src, vktFilterMode fm);

#ifdef __cplusplus
}
#endif
#endif
