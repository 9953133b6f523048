// Arithmetic family (reference include/cpp/vkt/Arithmetic.hpp:29-300, src/vkt/Arithmetic_serial.hpp:15-260):
//     dest[p + dstOffset] = map_dest( f( unmap(source1[p]), unmap(source2[p]) ) ),  p absolute in [first,last)
// f = + - * / |-|; the Safe* twins clamp f to dest's voxel mapping [lo,hi].  All arithmetic is IEEE binary32;
// every volume may have its own format and mapping; in-place use (dest aliasing a source, offset 0) is allowed.
#pragma once

#include <cstdint>

#include "common.hpp"
#include "linalg.hpp"

#define VKT_DECLARE_ARITHMETIC_OP__(NAME)                                                                   \
    VKTAPI Error NAME(StructuredVolume& dest, StructuredVolume& source1, StructuredVolume& source2);        \
    VKTAPI Error NAME##Range(StructuredVolume& dest, StructuredVolume& source1, StructuredVolume& source2,  \
                             int32_t firstX, int32_t firstY, int32_t firstZ,                                \
                             int32_t lastX, int32_t lastY, int32_t lastZ,                                   \
                             int32_t dstOffsetX = 0, int32_t dstOffsetY = 0, int32_t dstOffsetZ = 0);       \
    VKTAPI Error NAME##Range(StructuredVolume& dest, StructuredVolume& source1, StructuredVolume& source2,  \
                             Vec3i first, Vec3i last, Vec3i dstOffset = { 0, 0, 0 });

namespace vkt
{
    VKT_DECLARE_ARITHMETIC_OP__(Sum)
    VKT_DECLARE_ARITHMETIC_OP__(Diff)
    VKT_DECLARE_ARITHMETIC_OP__(Prod)
    VKT_DECLARE_ARITHMETIC_OP__(Quot)
    VKT_DECLARE_ARITHMETIC_OP__(AbsDiff)
    VKT_DECLARE_ARITHMETIC_OP__(SafeSum)
    VKT_DECLARE_ARITHMETIC_OP__(SafeDiff)
    VKT_DECLARE_ARITHMETIC_OP__(SafeProd)
    VKT_DECLARE_ARITHMETIC_OP__(SafeQuot)
    VKT_DECLARE_ARITHMETIC_OP__(SafeAbsDiff)
} // vkt

#undef VKT_DECLARE_ARITHMETIC_OP__
