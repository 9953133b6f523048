// Fill / FillRange (reference include/cpp/vkt/Fill.hpp:16-55, src/vkt/Fill.cpp:32-83): volume[p] = map(value) for
// p in [first,last); no bounds checks.  Under Device::GPU this runs vktb_fill_range (the reference's own dispatch
// has no GPU arm for Fill, src/vkt/Callable.cpp:53-66, so Fill is a no-op there; here it dispatches).
#pragma once

#include <cstdint>

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    VKTAPI Error Fill(StructuredVolume& volume, float value);
    VKTAPI Error FillRange(StructuredVolume& volume, int32_t firstX, int32_t firstY, int32_t firstZ,
                           int32_t lastX, int32_t lastY, int32_t lastZ, float value);
    VKTAPI Error FillRange(StructuredVolume& volume, Vec3i first, Vec3i last, float value);
} // vkt
