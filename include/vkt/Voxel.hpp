// Voxel view handed to Transform ops, and the scalar float <-> bytes codec
// (reference include/cpp/vkt/Voxel.hpp:15-40, src/vkt/VoxelMapping.hpp:15-180).
#pragma once

#include <cstdint>

#include "common.hpp"

namespace vkt
{
    struct VoxelView
    {
        uint8_t* bytes;
        DataFormat dataFormat;
        float mappingLo;
        float mappingHi;
    };

    VKTAPI Error MapVoxel(uint8_t* dst, float value, DataFormat dataFormat, float mappingLo, float mappingHi);
    VKTAPI Error UnmapVoxel(float& value, uint8_t const* src, DataFormat dataFormat, float mappingLo, float mappingHi);
} // vkt
