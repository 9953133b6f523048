// Scale (reference include/cpp/vkt/Scale.hpp:12-15, src/vkt/Scale.cpp:21-40, Scale_serial.hpp:15-56): every voxel p of
// dest is mapped to p' = (p - center) / scalingFactor + center; where p' lies inside the source, dest receives the
// reference's sampleLinear there (integer parameters: the coordinates are truncated), elsewhere dest is unchanged.
#pragma once

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    VKTAPI Error Scale(StructuredVolume& dest, StructuredVolume& source, Vec3f scalingFactor, Vec3f centerOfScaling);
} // vkt
