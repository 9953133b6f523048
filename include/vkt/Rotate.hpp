// Rotate (reference include/cpp/vkt/Rotate.hpp:12-16, src/vkt/Rotate.cpp:21-42, Rotate_serial.hpp:17-94): every voxel p
// of dest is mapped through the inverse rotation about `axis` (normalised here) by angleInRadians around
// centerOfRotation; where the result lies inside the source, dest receives the reference's sampleLinear there,
// elsewhere dest is unchanged.  The matrix is built on the host exactly as the reference builds it (6.28319f - angle,
// quaternion, quaternionToRotationMatrix) under both policies.
#pragma once

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    VKTAPI Error Rotate(StructuredVolume& dest, StructuredVolume& source, Vec3f axis, float angleInRadians,
                        Vec3f centerOfRotation);
} // vkt
