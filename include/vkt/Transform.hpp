// Transform / TransformRange (reference include/cpp/vkt/Transform.hpp:16-62, src/vkt/Transform_serial.hpp:15-103):
// for every voxel of [first,last) the op is called with the ABSOLUTE coordinates and a VoxelView of an 8-byte
// scratch copy of the voxel's raw bytes; the (possibly modified) bytes are written back.
//
// Under Device::GPU a host function pointer cannot run: the op must have been bound to a device op
// with BindDeviceOp() -- built-in ops by name ("xor_coord", "invert", "checker8", "threshold"; binary "or_assign",
// "swap").  An unbound pointer makes Transform return InvalidValue; it never falls back to the CPU.
#pragma once

#include <cstddef>
#include <cstdint>

#include "Voxel.hpp"
#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    typedef void (*TransformUnaryOp)(int32_t x, int32_t y, int32_t z, VoxelView voxel);
    typedef void (*TransformBinaryOp)(int32_t x1, int32_t y1, int32_t z1, VoxelView voxel1, VoxelView voxel2);

    VKTAPI Error Transform(StructuredVolume& volume, TransformUnaryOp unaryOp);
    VKTAPI Error Transform(StructuredVolume& volume1, StructuredVolume& volume2, TransformBinaryOp binaryOp);
    VKTAPI Error TransformRange(StructuredVolume& volume, int32_t firstX, int32_t firstY, int32_t firstZ,
                                int32_t lastX, int32_t lastY, int32_t lastZ, TransformUnaryOp unaryOp);
    VKTAPI Error TransformRange(StructuredVolume& volume, Vec3i first, Vec3i last, TransformUnaryOp unaryOp);
    VKTAPI Error TransformRange(StructuredVolume& volume1, StructuredVolume& volume2,
                                int32_t firstX, int32_t firstY, int32_t firstZ,
                                int32_t lastX, int32_t lastY, int32_t lastZ, TransformBinaryOp binaryOp);
    VKTAPI Error TransformRange(StructuredVolume& volume1, StructuredVolume& volume2, Vec3i first, Vec3i last,
                                TransformBinaryOp binaryOp);

    // volkit-b200 extension: bind a host op to the device op that computes the same function
    VKTAPI Error BindDeviceOp(TransformUnaryOp hostOp, char const* deviceOpName);
    VKTAPI Error BindDeviceOp(TransformBinaryOp hostOp, char const* deviceOpName);
    // ... or to the user's OWN op: the device image (cubin / PTX / fatbin) or CUDA C source of the same function written
    // with VKTB_UNARY_OP / VKTB_BINARY_OP(name, ...) of <vktb/transform_op.cuh>.  Afterwards Transform(volume, hostOp)
    // runs under Device::GPU like any other algorithm.
    VKTAPI Error BindDeviceOpImage(TransformUnaryOp hostOp, void const* image, std::size_t size, char const* name);
    VKTAPI Error BindDeviceOpImage(TransformBinaryOp hostOp, void const* image, std::size_t size, char const* name);
    VKTAPI Error BindDeviceOpSource(TransformUnaryOp hostOp, char const* cudaSource, char const* name);
    VKTAPI Error BindDeviceOpSource(TransformBinaryOp hostOp, char const* cudaSource, char const* name);
} // vkt
