/* Policy-dispatched memory primitives, C API (reference include/c/vkt/Memory.h:13-28). */
#ifndef VKT_MEMORY_H
#define VKT_MEMORY_H

#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { vktCopyKindHostToHost, vktCopyKindHostToDevice, vktCopyKindDeviceToHost, vktCopyKindDeviceToDevice } vktCopyKind;

VKTAPI void vktAllocate(void** ptr, size_t size);
VKTAPI void vktFree(void* ptr);
VKTAPI void vktMemcpy(void* dst, void const* src, size_t size, vktCopyKind ck);
VKTAPI void vktMemsetRange(void* dst, void const* src, size_t dstSize, size_t srcSize);

#ifdef __cplusplus
}
#endif
#endif
