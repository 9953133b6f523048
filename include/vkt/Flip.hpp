// Flip / FlipRange (reference include/cpp/vkt/Flip.hpp:23-70, src/vkt/Flip.cpp:26-117, Flip_serial.hpp:11-63).
// The serial body is followed AS WRITTEN: the loops cover [0, last - first) halved on the axis (first is otherwise
// ignored), the partner index is last-1-i, and z is mirrored when the axis is Y -- Axis::Y mirrors y and z together,
// Axis::Z copies the lower half of the z range unchanged.  Raw bytes are exchanged crosswise at + dstOffset; dest may
// be source (in place, dstOffset 0).
#pragma once

#include <cstdint>

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    VKTAPI Error Flip(StructuredVolume& dest, StructuredVolume& source, Axis axis);
    VKTAPI Error FlipRange(StructuredVolume& dest, StructuredVolume& source,
                           int32_t firstX, int32_t firstY, int32_t firstZ,
                           int32_t lastX, int32_t lastY, int32_t lastZ, Axis axis);
    VKTAPI Error FlipRange(StructuredVolume& dest, StructuredVolume& source,
                           int32_t firstX, int32_t firstY, int32_t firstZ,
                           int32_t lastX, int32_t lastY, int32_t lastZ,
                           int32_t dstOffsetX, int32_t dstOffsetY, int32_t dstOffsetZ, Axis axis);
    VKTAPI Error FlipRange(StructuredVolume& dest, StructuredVolume& source, Vec3i first, Vec3i last, Axis axis);
    VKTAPI Error FlipRange(StructuredVolume& dest, StructuredVolume& source, Vec3i first, Vec3i last, Vec3i dstOffset,
                           Axis axis);
} // vkt
