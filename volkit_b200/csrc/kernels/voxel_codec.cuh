// Device voxel codec: float <-> stored bytes, bit-exact with the reference's x86-64 CPU build.
//
// Restates MapVoxelImpl / UnmapVoxelImpl (reference src/vkt/VoxelMapping.hpp:15-96 / :98-180) and
// min/max/clamp/lerp (reference src/vkt/linalg.hpp:20-41) for sm_100a:
//   * every float operation is an explicit round-to-nearest intrinsic (__fadd_rn, __fmul_rn, __fdiv_rn)
//     so nvcc can never contract a multiply-add into an FMA (the CPU build has no FMA);
//   * float -> integer conversions emulate x86 `cvttss2si` followed by truncation to the low bits
//     (the reference's casts are UB that *wraps* on x86); PTX cvt saturates instead;
//   * comparisons keep the reference's operand order so NaN takes the same branch.
//
// Division shortcuts that are exact in binary32 (and therefore still bit-exact):
//   x / 65535.999f   == x * 2^-16   (65535.999f rounds to 65536.0f)
//   x / 4294967295.999f == x * 2^-32
//   x / (hi-lo)      == x * (1/(hi-lo)) when hi-lo is a power of two whose reciprocal is representable.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace vktb {

enum Format : int { F_INT8 = 1, F_INT16 = 2, F_INT32 = 3, F_U8 = 4, F_U16 = 5, F_U32 = 6, F_F32 = 7 };

__host__ __device__ constexpr int bytesPerVoxel(int f)
{
    return (f == F_INT8 || f == F_U8) ? 1 : (f == F_INT16 || f == F_U16) ? 2
         : (f == F_INT32 || f == F_U32 || f == F_F32) ? 4 : 0;
}

// Per-volume mapping constants, prepared on the host (codec_host.hpp).
struct Mapping
{
    float lo, hi;
    float range;      // hi - lo (computed in binary32 on the host, identical to the device result)
    float invRange;   // RN(1/(hi-lo)) when exactRecip != 0, else unused
    int   exactRecip; // 1: hi-lo is a power of two with a representable reciprocal: x/range == x*invRange for every x
                      // 2: x/range == divProven(x) for EVERY binary32 x, proven exhaustively on the device (arith.cu)
    int   identity;   // lo == +0.0f && hi == 1.0f (bit pattern): unmap/map reduce to pure scaling
};

// ---- reference linalg.hpp:20-41 ---------------------------------------------------------------------
__device__ __forceinline__ float refMin(float a, float b) { return b < a ? b : a; }
__device__ __forceinline__ float refMax(float a, float b) { return a < b ? b : a; }
__device__ __forceinline__ float refClamp(float x, float a, float b) { return refMax(a, refMin(x, b)); }
__device__ __forceinline__ float refLerp(float a, float b, float x)
{
    return __fadd_rn(__fmul_rn(__fsub_rn(1.0f, x), a), __fmul_rn(x, b));
}

// ---- x86 cvttss2si emulation -----------------------------------------------------------------------
// r32 form: NaN / out of [-2^31, 2^31) -> 0x80000000.  PTX cvt.rzi.s32.f32 saturates (NaN -> 0,
// +big -> 0x7fffffff, -big -> 0x80000000): only the positive overflow and NaN differ.
__device__ __forceinline__ uint32_t x86Cvtt32(float f)
{
    int32_t r = __float2int_rz(f);
    return (f < 2147483648.0f) ? (uint32_t)r : 0x80000000u;   // NaN compares false -> indefinite
}
// Low 8/16 bits only: NaN -> 0 either way, so one compare suffices.
__device__ __forceinline__ uint32_t x86Cvtt32Low16(float f)
{
    int32_t r = __float2int_rz(f);
    return (f >= 2147483648.0f) ? 0u : (uint32_t)r;
}
// r64 form, low 32 bits: >= 2^63 -> 0x8000000000000000 (low word 0); NaN -> low word 0 both ways.
__device__ __forceinline__ uint32_t x86Cvtt64Low32(float f)
{
    long long r = __float2ll_rz(f);
    return (f >= 9223372036854775808.0f) ? 0u : (uint32_t)(unsigned long long)r;
}

// ---- IEEE division by the mapping range ---------------------------------------------------------------
// x / d for a divisor that stays the same for a whole volume.  With y = RN(1/d): q0 = RN(x*y), the residual
// r = fma(-q0, d, x) is exact while nothing underflows or overflows, and q = fma(r, y, q0) is the correctly rounded
// quotient for all but a few exceptional divisors (the last step of the classic reciprocal-based division; it needs
// real FMAs, which --fmad=false does not take away from explicit fmaf calls).  "All but a few" is not bit-exact, so the
// form is only used for divisors for which it has been PROVEN: arith.cu compares it with __fdiv_rn for all 2^32
// dividends once per divisor.  Dividends outside [2^-63, 2^64) other than +-0 take the IEEE division (residual not
// exact there).  Three arithmetic instructions instead of the ~11 plus a guarded slow-path call of __fdiv_rn.
__device__ __forceinline__ float divProven(float x, float d, float y)
{
    const uint32_t mag = __float_as_uint(x) & 0x7fffffffu;
    const float q0 = __fmul_rn(x, y);
    if (mag - 0x20000000u < 0x3f800000u) {                     // exponent field in [64, 191)
        const float r = __fmaf_rn(-q0, d, x);
        return __fmaf_rn(r, y, q0);
    }
    // +-0 / d is the signed zero x * y already (the correction step would turn -0 into +0: caught by the proof)
    return mag == 0u ? q0 : __fdiv_rn(x, d);
}

// ---- map: float -> raw (returned in the low bits of a uint32) ---------------------------------------
// PROVEN: the launcher has checked m.exactRecip == 2 once for the whole launch (arithFlatKernel); every other kernel
// uses the default and treats a proven divisor like any other (IEEE division): still exact, just not faster.
template <int F, bool IDENT, bool PROVEN = false>
__device__ __forceinline__ uint32_t mapVoxel(float value, const Mapping& m)
{
    if (!IDENT) {
        value = __fsub_rn(value, m.lo);
        if (PROVEN) value = divProven(value, m.range, m.invRange);
        else value = m.exactRecip == 1 ? __fmul_rn(value, m.invRange) : __fdiv_rn(value, m.range);
    }
    // IDENT: (v - +0) == v and v / 1 == v for every v (incl. -0, NaN payloads aside)
    if (F == F_U8)    return x86Cvtt32Low16(__fmul_rn(value, 255.999f)) & 0xffu;
    if (F == F_U16)   return x86Cvtt32Low16(__fmul_rn(value, 65535.999f)) & 0xffffu;
    if (F == F_INT16) return x86Cvtt32Low16(__fsub_rn(__fmul_rn(value, 65535.999f), 32767.f)) & 0xffffu;
    if (F == F_U32)   return x86Cvtt64Low32(__fmul_rn(value, 4294967295.999f));
    if (F == F_F32)   return __float_as_uint(value);
    return 0u;
}

// Exact uint -> float for values below 2^23 without the conversion pipe: OR the value into the mantissa of 2^23
// and subtract 2^23 (both steps exact).  I2F runs at a fraction of the FP32 rate and the integer kernels convert
// two to eight voxels per output voxel.
__device__ __forceinline__ float smallUintToFloat(uint32_t v)
{
    return __fsub_rn(__uint_as_float(0x4B000000u | v), 8388608.0f);
}
// v * 2^-16 for v < 2^16 the same way: 128.0f has an ulp of 2^-16, so 128 + v*2^-16 is the bit pattern
// 0x43000000 | v and the subtraction is exact -- equal to float(v) / 65535.999f (= float(v) * 2^-16, also exact).
__device__ __forceinline__ float u16ToUnit(uint32_t v)
{
    return __fsub_rn(__uint_as_float(0x43000000u | v), 128.0f);
}

// ---- unmap: raw -> float ----------------------------------------------------------------------------
template <int F, bool IDENT>
__device__ __forceinline__ float unmapVoxel(uint32_t raw, const Mapping& m)
{
    if (F == F_F32) return __uint_as_float(raw);   // raw bits, mapping NOT applied (VoxelMapping.hpp:163-178)
    float t;
    if (F == F_U8)         t = __fdiv_rn(smallUintToFloat(raw & 0xffu), 255.999f);
    else if (F == F_U16)   t = u16ToUnit(raw & 0xffffu);
    else if (F == F_INT16) t = __fmul_rn(__fadd_rn(__int2float_rn((int)(short)(raw & 0xffffu)), 32767.f), 1.52587890625e-05f);
    else if (F == F_U32)   t = __fmul_rn(__uint2float_rn(raw), 2.3283064365386962890625e-10f);
    else return 0.f;
    // IDENT: (1-t)*(+0) + t*1 == t exactly for t >= 0 (t is never negative for the unsigned formats;
    // Int16 can give t < 0 only for raw = -32768 -> keep the general path there)
    if (IDENT && F != F_INT16) return t;
    return refLerp(m.lo, m.hi, t);
}

// UInt8 unmap through a 256-entry table (built once per launch with the exact ops above): replaces the
// IEEE division by one shared-memory load.
template <bool IDENT>
__device__ __forceinline__ void buildU8Table(float* lut /*[256] shared*/, const Mapping& m)
{
    for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = unmapVoxel<F_U8, IDENT>((uint32_t)i, m);
}

// ---- runtime-format variants (generic fallback paths) -------------------------------------------------
__device__ __forceinline__ uint32_t mapVoxelRt(int f, float v, const Mapping& m)
{
    switch (f) {
    case F_U8:    return mapVoxel<F_U8, false>(v, m);
    case F_U16:   return mapVoxel<F_U16, false>(v, m);
    case F_INT16: return mapVoxel<F_INT16, false>(v, m);
    case F_U32:   return mapVoxel<F_U32, false>(v, m);
    case F_F32:   return mapVoxel<F_F32, false>(v, m);
    }
    return 0u;
}

__device__ __forceinline__ float unmapVoxelRt(int f, uint32_t raw, const Mapping& m)
{
    switch (f) {
    case F_U8:    return unmapVoxel<F_U8, false>(raw, m);
    case F_U16:   return unmapVoxel<F_U16, false>(raw, m);
    case F_INT16: return unmapVoxel<F_INT16, false>(raw, m);
    case F_U32:   return unmapVoxel<F_U32, false>(raw, m);
    case F_F32:   return unmapVoxel<F_F32, false>(raw, m);
    }
    return 0.f;
}

// raw element load/store by runtime size (little endian, naturally aligned)
__device__ __forceinline__ uint32_t loadRaw(const uint8_t* p, int bpv)
{
    if (bpv == 1) return *p;
    if (bpv == 2) return *reinterpret_cast<const uint16_t*>(p);
    return *reinterpret_cast<const uint32_t*>(p);
}
__device__ __forceinline__ void storeRaw(uint8_t* p, int bpv, uint32_t v)
{
    if (bpv == 1) *p = (uint8_t)v;
    else if (bpv == 2) *reinterpret_cast<uint16_t*>(p) = (uint16_t)v;
    else *reinterpret_cast<uint32_t*>(p) = v;
}

} // namespace vktb
