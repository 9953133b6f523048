// BrickDecompose / BrickDecomposeResize (reference include/cpp/vkt/Decompose.hpp:14-60, src/vkt/Decompose.cpp:22-150,
// Decompose_serial.hpp:15-48).  Resize lays out ceil(dims / brickSize) bricks, border bricks cropped to the volume,
// every brick grown by the halos, each a new volume with the source's format, dist and mapping.  Decompose fills brick
// `index` with CopyRange(dest[index], source, index*brickSize - haloNeg, min(index*brickSize + brickSize, dims) +
// haloPos): halo voxels beyond the volume repeat its border (Copy clamps).  Under the GPU policy all bricks are
// written by ONE kernel launch (vktb_copy_range_batch).
#pragma once

#include <cstdint>

#include "Array3D.hpp"
#include "StructuredVolume.hpp"
#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    VKTAPI Error BrickDecompose(Array3D<StructuredVolume>& dest, StructuredVolume& source,
                                int32_t brickSizeX, int32_t brickSizeY, int32_t brickSizeZ,
                                int32_t haloSizeNegX = 0, int32_t haloSizeNegY = 0, int32_t haloSizeNegZ = 0,
                                int32_t haloSizePosX = 0, int32_t haloSizePosY = 0, int32_t haloSizePosZ = 0);
    VKTAPI Error BrickDecompose(Array3D<StructuredVolume>& dest, StructuredVolume& source, Vec3i brickSize,
                                Vec3i haloSizeNeg = { 0, 0, 0 }, Vec3i haloSizePos = { 0, 0, 0 });
    VKTAPI Error BrickDecomposeResize(Array3D<StructuredVolume>& dest, StructuredVolume& source,
                                      int32_t brickSizeX, int32_t brickSizeY, int32_t brickSizeZ,
                                      int32_t haloSizeNegX = 0, int32_t haloSizeNegY = 0, int32_t haloSizeNegZ = 0,
                                      int32_t haloSizePosX = 0, int32_t haloSizePosY = 0, int32_t haloSizePosZ = 0);
    VKTAPI Error BrickDecomposeResize(Array3D<StructuredVolume>& dest, StructuredVolume& source, Vec3i brickSize,
                                      Vec3i haloSizeNeg = { 0, 0, 0 }, Vec3i haloSizePos = { 0, 0, 0 });
} // vkt
