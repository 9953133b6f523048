/* Arithmetic family, C API (reference include/c/vkt/Arithmetic.h:16-186): vkt<Op>SV and vkt<Op>RangeSV for
 * Op in Sum, Diff, Prod, Quot, AbsDiff and their Safe* twins. */
#ifndef VKT_ARITHMETIC_H
#define VKT_ARITHMETIC_H

#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

#define VKT_DECLARE_ARITHMETIC_OP_C__(NAME)                                                                      \
    VKTAPI vktError vkt##NAME##SV(vktStructuredVolume dest, vktStructuredVolume source1, vktStructuredVolume source2); \
    VKTAPI vktError vkt##NAME##RangeSV(vktStructuredVolume dest, vktStructuredVolume source1,                    \
                                       vktStructuredVolume source2, int32_t firstX, int32_t firstY,              \
                                       int32_t firstZ, int32_t lastX, int32_t lastY, int32_t lastZ,              \
                                       int32_t dstOffsetX, int32_t dstOffsetY, int32_t dstOffsetZ);

VKT_DECLARE_ARITHMETIC_OP_C__(Sum)
VKT_DECLARE_ARITHMETIC_OP_C__(Diff)
VKT_DECLARE_ARITHMETIC_OP_C__(Prod)
VKT_DECLARE_ARITHMETIC_OP_C__(Quot)
VKT_DECLARE_ARITHMETIC_OP_C__(AbsDiff)
VKT_DECLARE_ARITHMETIC_OP_C__(SafeSum)
VKT_DECLARE_ARITHMETIC_OP_C__(SafeDiff)
VKT_DECLARE_ARITHMETIC_OP_C__(SafeProd)
VKT_DECLARE_ARITHMETIC_OP_C__(SafeQuot)
VKT_DECLARE_ARITHMETIC_OP_C__(SafeAbsDiff)

#undef VKT_DECLARE_ARITHMETIC_OP_C__

#ifdef __cplusplus
}
#endif
#endif
