// POD vector / box types of the public API (reference include/cpp/vkt/linalg.hpp:8-96, hot-path subset).
#pragma once

namespace vkt
{
    struct Vec2f { float x, y; };
    struct Vec3f { float x, y, z; };
    struct Vec2i { int x, y; };
    struct Vec3i { int x, y, z; };
    struct Box3f { Vec3f min, max; };
    struct Box3i { Vec3i min, max; };

    enum class Axis { X, Y, Z };
} // vkt
