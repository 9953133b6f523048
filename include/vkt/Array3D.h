/* 3-D array of volume handles, C API (reference include/c/vkt/Array3D.h:18-231, the VKT_ARRAY3D_INIT(
 * vktStructuredVolume) instantiation -- the only one the reference makes).  Elements are vktStructuredVolume
 * handles, NULL until the caller (or vktBrickDecomposeResizeSV) creates a volume in the slot; as in the reference,
 * Destroy releases the array, not the volumes. */
#ifndef VKT_ARRAY3D_H
#define VKT_ARRAY3D_H

#include <stddef.h>

#include "common.h"
#include "linalg.h"

#ifdef __cplusplus
extern "C" {
#endif

struct vktArray3D_vktStructuredVolume_impl;
typedef struct vktArray3D_vktStructuredVolume_impl* vktArray3D_vktStructuredVolume;
typedef vktStructuredVolume  vktArray3D_vktStructuredVolume_ValueType;
typedef vktStructuredVolume* vktArray3D_vktStructuredVolume_Iterator;
typedef vktStructuredVolume const* vktArray3D_vktStructuredVolume_ConstIterator;

VKTAPI void vktArray3D_vktStructuredVolume_CreateEmpty(vktArray3D_vktStructuredVolume* arr);
VKTAPI void vktArray3D_vktStructuredVolume_Create(vktArray3D_vktStructuredVolume* arr, vktVec3i_t dims);
VKTAPI void vktArray3D_vktStructuredVolume_CreateCopy(vktArray3D_vktStructuredVolume* arr, vktArray3D_vktStructuredVolume rhs);
VKTAPI void vktArray3D_vktStructuredVolume_Destroy(vktArray3D_vktStructuredVolume arr);
VKTAPI void vktArray3D_vktStructuredVolume_Resize(vktArray3D_vktStructuredVolume arr, vktVec3i_t dims);
VKTAPI void vktArray3D_vktStructuredVolume_Fill(vktArray3D_vktStructuredVolume arr, vktStructuredVolume value);
VKTAPI vktArray3D_vktStructuredVolume_Iterator vktArray3D_vktStructuredVolume_Begin(vktArray3D_vktStructuredVolume arr);
VKTAPI vktArray3D_vktStructuredVolume_ConstIterator vktArray3D_vktStructuredVolume_CBegin(vktArray3D_vktStructuredVolume arr);
VKTAPI vktArray3D_vktStructuredVolume_Iterator vktArray3D_vktStructuredVolume_End(vktArray3D_vktStructuredVolume arr);
VKTAPI vktArray3D_vktStructuredVolume_ConstIterator vktArray3D_vktStructuredVolume_CEnd(vktArray3D_vktStructuredVolume arr);
VKTAPI vktStructuredVolume* vktArray3D_vktStructuredVolume_Access(vktArray3D_vktStructuredVolume arr, vktVec3i_t index);
VKTAPI vktStructuredVolume const* vktArray3D_vktStructuredVolume_CAccess(vktArray3D_vktStructuredVolume arr, vktVec3i_t index);
VKTAPI vktBool_t vktArray3D_vktStructuredVolume_Empty(vktArray3D_vktStructuredVolume arr);
VKTAPI vktStructuredVolume* vktArray3D_vktStructuredVolume_Data(vktArray3D_vktStructuredVolume arr);
VKTAPI vktStructuredVolume const* vktArray3D_vktStructuredVolume_CData(vktArray3D_vktStructuredVolume arr);
VKTAPI vktVec3i_t vktArray3D_vktStructuredVolume_Dims(vktArray3D_vktStructuredVolume arr);
VKTAPI size_t vktArray3D_vktStructuredVolume_NumElements(vktArray3D_vktStructuredVolume arr);

#ifdef __cplusplus
}
#endif
#endif
