// Host twin of kernels/voxel_codec.cuh: float <-> stored bytes with the semantics of the reference's x86-64 build
// (reference src/vkt/VoxelMapping.hpp:15-96 map, :98-180 unmap; src/vkt/linalg.hpp:20-41 min/max/clamp/lerp).
// The reference's float -> integer casts are undefined behaviour that compiles to cvttss2si + truncation; here the
// same instructions are requested explicitly through SSE intrinsics, so the result does not depend on the optimiser.
// Built with -ffp-contract=off: every float operation rounds on its own, as on the device (--fmad=false).
#pragma once

#include <cstdint>
#include <cstring>
#include <xmmintrin.h>

#include <vkt/common.hpp>

namespace vkt
{
namespace host
{
    inline int bytesPerVoxel(DataFormat f)
    {
        switch (f) {
        case DataFormat::Int8: case DataFormat::UInt8: return 1;
        case DataFormat::Int16: case DataFormat::UInt16: return 2;
        case DataFormat::Int32: case DataFormat::UInt32: case DataFormat::Float32: return 4;
        default: return 0;
        }
    }

    inline float refMin(float a, float b) { return b < a ? b : a; }
    inline float refMax(float a, float b) { return a < b ? b : a; }
    inline float refClamp(float x, float a, float b) { return refMax(a, refMin(x, b)); }
    inline float refLerp(float a, float b, float x) { return (1.0f - x) * a + x * b; }

    inline int32_t cvtt32(float f) { return _mm_cvttss_si32(_mm_set_ss(f)); }          // 0x80000000 when out of range / NaN
    inline int64_t cvtt64(float f) { return _mm_cvttss_si64(_mm_set_ss(f)); }          // 0x8000000000000000 likewise

    // writes bytesPerVoxel(format) bytes; Int8 / Int32 have no case in the reference either (no-op)
    inline void mapVoxel(uint8_t* dst, float value, DataFormat format, float lo, float hi)
    {
        value -= lo;
        value /= hi - lo;
        switch (format) {
        case DataFormat::UInt8:  { uint8_t v = (uint8_t)cvtt32(value * 255.999f); dst[0] = v; break; }
        case DataFormat::UInt16: { uint16_t v = (uint16_t)cvtt32(value * 65535.999f); std::memcpy(dst, &v, 2); break; }
        case DataFormat::Int16:  { uint16_t v = (uint16_t)cvtt32(value * 65535.999f - 32767.f); std::memcpy(dst, &v, 2); break; }
        case DataFormat::UInt32: { uint32_t v = (uint32_t)(uint64_t)cvtt64(value * 4294967295.999f); std::memcpy(dst, &v, 4); break; }
        case DataFormat::Float32: std::memcpy(dst, &value, 4); break;
        default: break;
        }
    }

    inline float unmapVoxel(uint8_t const* src, DataFormat format, float lo, float hi)
    {
        float t;
        switch (format) {
        case DataFormat::UInt8:  t = (float)src[0] / 255.999f; break;
        case DataFormat::UInt16: { uint16_t v; std::memcpy(&v, src, 2); t = (float)v / 65535.999f; break; }
        case DataFormat::Int16:  { int16_t v; std::memcpy(&v, src, 2); t = ((float)v + 32767.f) / 65535.999f; break; }
        case DataFormat::UInt32: { uint32_t v; std::memcpy(&v, src, 4); t = (float)v / 4294967295.999f; break; }
        case DataFormat::Float32: { float v; std::memcpy(&v, src, 4); return v; }   // raw value, mapping not applied
        default: return 0.f;
        }
        return refLerp(lo, hi, t);
    }
} // host
} // vkt
