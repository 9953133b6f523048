// Device::CPU / HostAPI::Serial bodies of the hot-path algorithms.  These run only when the calling thread's
// policy SELECTS the CPU; nothing under Device::GPU ever reaches them (a failing GPU call returns InvalidValue).
// Semantics follow the reference's serial path (files cited per function in serial.cpp); unlike the reference they
// work on a raw host view instead of calling migrate() once per voxel.
#pragma once

#include <cstdint>

#include <vkt/Aggregates.hpp>
#include <vkt/Transform.hpp>
#include <vkt/common.hpp>
#include <vkt/linalg.hpp>

namespace vkt
{
namespace host
{
    struct View
    {
        uint8_t*   data;
        Vec3i      dims;
        DataFormat format;
        float      lo, hi;
        int        bpv;

        std::size_t offset(int32_t x, int32_t y, int32_t z) const
        {
            return ((std::size_t)z * dims.x * (std::size_t)dims.y + (std::size_t)y * dims.x + (std::size_t)x) * bpv;
        }
    };

    void fillRange(View v, Vec3i first, Vec3i last, float value);
    void copyRange(View dst, View src, Vec3i first, Vec3i last, Vec3i dstOffset);
    void arithRange(int op, View dst, View s1, View s2, Vec3i first, Vec3i last, Vec3i dstOffset);
    void transformRange(View v, Vec3i first, Vec3i last, TransformUnaryOp op);
    void transformRange(View v1, View v2, Vec3i first, Vec3i last, TransformBinaryOp op);
    void resample(View dst, View src, int filter);
    void flipRange(View dst, View src, Vec3i first, Vec3i last, Vec3i dstOffset, int axis);
    // in place on dst (already holding the copied range): first/last in dst space, first > last = downwards
    void scanRange(View dst, Vec3i first, Vec3i last);
    // Scale / Rotate: m = col0.xyz, col1.xyz, col2.xyz; dst.data != src.data
    void scaleMatrix(Vec3f factor, float m[9]);
    void rotateMatrix(Vec3f axis, float angleInRadians, float m[9]);
    void affineRange(View dst, View src, Vec3i first, Vec3i last, const float m[9], Vec3f center);
    // weights x fastest, fdims in 1..3; dst may alias src (a copy is filtered)
    void applyFilter(View dst, View src, const float* weights, Vec3i fdims);
    // out: the 13 words of vkt::Aggregates
    void aggregatesRange(View vol, Vec3i first, Vec3i last, Aggregates& out);
    void histogramRange(View vol, Vec3i first, Vec3i last, std::size_t* bins, std::size_t numBins);

    // BrickDecomposeResize's brick grid (reference src/vkt/Decompose.cpp:96-146): number of bricks and the dims of
    // brick `index` including its halos
    Vec3i brickCount(Vec3i sourceDims, Vec3i brickSize);
    Vec3i brickDims(Vec3i sourceDims, Vec3i brickSize, Vec3i haloNeg, Vec3i haloPos, Vec3i index);
} // host

    class StructuredVolume;
namespace host
{
    // body shared by vkt::BrickDecompose and vktBrickDecomposeSV: bricks[i] in x-fastest order (algorithms.cpp)
    Error brickDecompose(StructuredVolume* const* bricks, Vec3i numBricks, StructuredVolume& source, Vec3i brickSize,
                         Vec3i haloNeg, Vec3i haloPos);
} // host
} // vkt
