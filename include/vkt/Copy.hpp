// Copy / CopyRange (reference include/cpp/vkt/Copy.hpp:14-33, src/vkt/Copy.cpp:26-60, Copy_serial.hpp:13-82):
// dst[p - first + dstOffset] = src[clamp(p, 0, srcDims-1)] for p in [first,last); raw bytes when format AND mapping
// agree, otherwise by value.  Copy(dst, src) uses the DST dims as the range (Copy.cpp:32).
#pragma once

#include <cstdint>

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    VKTAPI Error Copy(StructuredVolume& dst, StructuredVolume& src);
    VKTAPI Error CopyRange(StructuredVolume& dst, StructuredVolume& src,
                           int32_t firstX, int32_t firstY, int32_t firstZ,
                           int32_t lastX, int32_t lastY, int32_t lastZ,
                           int32_t dstOffsetX = 0, int32_t dstOffsetY = 0, int32_t dstOffsetZ = 0);
    VKTAPI Error CopyRange(StructuredVolume& dst, StructuredVolume& src, Vec3i first, Vec3i last,
                           Vec3i dstOffset = { 0, 0, 0 });
} // vkt
