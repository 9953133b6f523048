// A user program in the style of the reference's examples (src/examples/Arithmetic.cpp:9-40, ExecutionPolicy.cpp),
// built against the REFERENCE's public headers and its own front files (integration/Makefile): every algorithm of
// the hot path runs once under Device::CPU (the reference's serial bodies) and once under Device::GPU (the same
// fronts -> VKT_LEGACY_CALL__ / Call() -> B200Backend.cpp -> libvktb200.so), with the reference's header-inline
// ManagedBuffer<T>::migrate() driving vktb_alloc / vktb_memcpy / vktb_free.  The two results must agree byte for byte.
// Prints one line per op; exit code = number of mismatches.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include <vkt/Arithmetic.hpp>
#include <vkt/Copy.hpp>
#include <vkt/ExecutionPolicy.hpp>
#include <vkt/Fill.hpp>
#include <vkt/Resample.hpp>
#include <vkt/StructuredVolume.hpp>
#include <vkt/Transform.hpp>

#include <vktb/backend.h>

using namespace vkt;

static void setDevice(ExecutionPolicy::Device dev)
{
    ExecutionPolicy ep = GetThreadExecutionPolicy();
    ep.device = dev;
    SetThreadExecutionPolicy(ep);
}

static uint64_t splitmix(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

// fills on the CPU (getData() under the CPU policy is host memory)
static void synth(StructuredVolume& v, uint64_t seed)
{
    setDevice(ExecutionPolicy::Device::CPU);
    uint8_t* p = v.getData();
    for (std::size_t i = 0; i < v.getSizeInBytes(); ++i) p[i] = (uint8_t)splitmix(i ^ (seed * 0x9E3779B97F4A7C15ull));
}

static std::vector<uint8_t> bytes(StructuredVolume& v)
{
    setDevice(ExecutionPolicy::Device::CPU);          // deferred: the next accessor migrates the volume back
    uint8_t* p = v.getData();
    return std::vector<uint8_t>(p, p + v.getSizeInBytes());
}

// the ops of src/examples/CoreAlgorithms.c:12-24
static void xorCoord(int32_t x, int32_t y, int32_t z, VoxelView voxel) { voxel.bytes[0] ^= (uint8_t)(x + y + z); }

static int failures = 0;

template <typename F>
static void both(char const* name, Vec3i dims, DataFormat fmt, Vec3i dstDims, DataFormat dstFmt, F run)
{
    std::vector<uint8_t> result[2];
    for (int pass = 0; pass < 2; ++pass) {
        setDevice(ExecutionPolicy::Device::CPU);
        StructuredVolume a(dims.x, dims.y, dims.z, fmt), b(dims.x, dims.y, dims.z, fmt);
        StructuredVolume d(dstDims.x, dstDims.y, dstDims.z, dstFmt);
        synth(a, 1); synth(b, 2); synth(d, 3);
        setDevice(pass == 0 ? ExecutionPolicy::Device::CPU : ExecutionPolicy::Device::GPU);
        run(d, a, b);
        result[pass] = bytes(d);
    }
    const bool same = result[0] == result[1];
    std::printf("%-34s CPU vs GPU: %s (%zu bytes)\n", name, same ? "identical" : "MISMATCH", result[0].size());
    if (!same) ++failures;
}

int main()
{
    int n = 0;
    if (vktb_device_count(&n) != 0 || n == 0) {
        std::fprintf(stderr, "both_policies: no usable GPU: %s\n", vktb_last_error());
        return 99;
    }
    int op = -1;
    vktb_unary_op_by_name("xor_coord", &op);
    vktb_bind_unary_op((void const*)&xorCoord, op);   // Transform under the GPU policy needs a device twin

    const Vec3i S = { 48, 40, 36 };
    const DataFormat U16 = DataFormat::UInt16, U8 = DataFormat::UInt8, F32 = DataFormat::Float32;
    both("Fill", S, U16, S, U16, [](StructuredVolume& d, StructuredVolume&, StructuredVolume&) { Fill(d, 0.3f); });
    both("FillRange", S, U16, S, U16, [](StructuredVolume& d, StructuredVolume&, StructuredVolume&) {
        FillRange(d, Vec3i{ 3, 5, 7 }, Vec3i{ 40, 33, 30 }, 0.7f); });
    both("Copy", S, U16, S, U16, [](StructuredVolume& d, StructuredVolume& a, StructuredVolume&) { Copy(d, a); });
    both("CopyRange dstOffset, clamped", S, U16, S, U16, [](StructuredVolume& d, StructuredVolume& a, StructuredVolume&) {
        CopyRange(d, a, Vec3i{ -2, -1, -3 }, Vec3i{ 30, 25, 20 }, Vec3i{ 8, 8, 8 }); });
    both("Copy UInt16 -> UInt8 (converting)", S, U16, S, U8, [](StructuredVolume& d, StructuredVolume& a, StructuredVolume&) { Copy(d, a); });
    both("SafeSum", S, U16, S, U16, [](StructuredVolume& d, StructuredVolume& a, StructuredVolume& b) { SafeSum(d, a, b); });
    both("SafeDiff", S, U16, S, U16, [](StructuredVolume& d, StructuredVolume& a, StructuredVolume& b) { SafeDiff(d, a, b); });
    both("SafeProd", S, U8, S, U8, [](StructuredVolume& d, StructuredVolume& a, StructuredVolume& b) { SafeProd(d, a, b); });
    both("SafeQuot", S, U16, S, U16, [](StructuredVolume& d, StructuredVolume& a, StructuredVolume& b) { SafeQuot(d, a, b); });
    both("SumRange dstOffset", S, U16, S, U16, [](StructuredVolume& d, StructuredVolume& a, StructuredVolume& b) {
        SumRange(d, a, b, Vec3i{ 1, 2, 3 }, Vec3i{ 38, 30, 26 }, Vec3i{ 8, 8, 8 }); });
    both("Transform (xor_coord)", S, U16, S, U16, [](StructuredVolume& d, StructuredVolume&, StructuredVolume&) { Transform(d, xorCoord); });
    both("Resample Nearest 48x40x36 -> 24x20x18", S, U16, Vec3i{ 24, 20, 18 }, U16,
         [](StructuredVolume& d, StructuredVolume& a, StructuredVolume&) { Resample(d, a, FilterMode::Nearest); });
    both("Resample Linear -> 72x60x54", S, U16, Vec3i{ 72, 60, 54 }, U16,
         [](StructuredVolume& d, StructuredVolume& a, StructuredVolume&) { Resample(d, a, FilterMode::Linear); });
    both("Resample equal dims UInt16 -> UInt8", S, U16, S, U8,
         [](StructuredVolume& d, StructuredVolume& a, StructuredVolume&) { Resample(d, a, FilterMode::Nearest); });
    (void)F32;
    std::printf("%s\n", failures ? "FAILED" : "ALL IDENTICAL");
    return failures;
}
