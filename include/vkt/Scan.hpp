// Scan / ScanRange: 3-D prefix sum volume (reference include/cpp/vkt/Scan.hpp:24-53, src/vkt/Scan.cpp:22-71,
// Scan_serial.hpp:18-152).  dst receives CopyRange(dst, src, first, last, dstOffset) unless dst and src are the same
// object; then every voxel of the range (moved to dst space) becomes the sum of all voxels between the `first`
// corner and itself, taken over the raw voxel integers modulo 2^(8 bytesPerVoxel).  first > last on an axis sums
// right to left (in place only: the reference's copy of such a range does not terminate; rejected here).
#pragma once

#include <cstdint>

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    VKTAPI Error Scan(StructuredVolume& dst, StructuredVolume& src);
    VKTAPI Error ScanRange(StructuredVolume& dst, StructuredVolume& src,
                           int32_t firstX, int32_t firstY, int32_t firstZ,
                           int32_t lastX, int32_t lastY, int32_t lastZ,
                           int32_t dstOffsetX = 0, int32_t dstOffsetY = 0, int32_t dstOffsetZ = 0);
    VKTAPI Error ScanRange(StructuredVolume& dst, StructuredVolume& src, Vec3i first, Vec3i last,
                           Vec3i dstOffset = { 0, 0, 0 });
} // vkt
