// Host-side launcher interface between the C-ABI (backend/cabi.cpp) and the kernels (*.cu).
#pragma once

#include <atomic>
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "common.cuh"
#include "voxel_codec.cuh"

namespace vktb {

struct VolDesc
{
    uint8_t* data;
    int      dims[3];
    int      format;
    Mapping  map;
};

struct LaunchCtx
{
    cudaStream_t stream;
    int          smCount;
};

// Every kernel launch of the library goes through this counter (vktb_launch_count()).
extern std::atomic<uint64_t> g_launchCount;
inline void countLaunch() { g_launchCount.fetch_add(1, std::memory_order_relaxed); }

// Host twin of the Mapping preparation (binary32 arithmetic, identical on host and device).
Mapping makeMapping(float lo, float hi);

// first/last/off are the reference's Vec3i arguments, passed through unchanged.
cudaError_t launchFill(const LaunchCtx&, const VolDesc& vol, const int32_t first[3], const int32_t last[3], float value);
cudaError_t launchCopy(const LaunchCtx&, const VolDesc& dst, const VolDesc& src, const int32_t first[3],
                       const int32_t last[3], const int32_t off[3]);
cudaError_t launchCopyBatch(const LaunchCtx&, int n, const VolDesc* dsts, const VolDesc& src, const int32_t* firsts,
                            const int32_t* lasts, const int32_t* offs);
// the same batch with its box table in a file-scope device array: no allocation (at most copyBatchNoAllocMaxBoxes()
// boxes, bytewise boxes only: cudaErrorInvalidValue / cudaErrorNotSupported otherwise)
int copyBatchNoAllocMaxBoxes();
cudaError_t launchCopyBatchNoAlloc(const LaunchCtx&, int n, const VolDesc* dsts, const VolDesc& src, const int32_t* firsts,
                                   const int32_t* lasts, const int32_t* offs);
cudaError_t launchArith(const LaunchCtx&, int op, const VolDesc& dst, const VolDesc& s1, const VolDesc& s2,
                        const int32_t first[3], const int32_t last[3], const int32_t off[3]);
int divisorProven(const LaunchCtx&, float lo, float hi);
cudaError_t launchTransform1(const LaunchCtx&, const VolDesc& vol, const int32_t first[3], const int32_t last[3], int op);
cudaError_t launchTransform2(const LaunchCtx&, const VolDesc& v1, const VolDesc& v2, const int32_t first[3],
                             const int32_t last[3], int op);
cudaError_t launchResample(const LaunchCtx&, const VolDesc& dst, const VolDesc& src, int filter,
                           int dstZ0, const int32_t dstFull[3], int srcZ0, const int32_t srcFull[3]);
cudaError_t launchResampleRowRuns(const LaunchCtx&, const VolDesc& dst, const VolDesc& src,
                                  int dstZ0, const int32_t dstFull[3], int srcZ0, const int32_t srcFull[3], unsigned* tapFlag = nullptr);
cudaError_t launchScaleRowRuns(const LaunchCtx&, const VolDesc& dst, const VolDesc& src, const float m[3],
                               const float center[3], const int32_t first[3], const int32_t last[3]);
cudaError_t launchResampleMarch(const LaunchCtx&, const VolDesc& dst, const VolDesc& src, int mode /*1 Linear8, 2 trilinear*/,
                                int dstZ0, const int32_t dstFull[3], int srcZ0, const int32_t srcFull[3], const unsigned* gate = nullptr);
cudaError_t launchConvertFlat(const LaunchCtx&, const VolDesc& dst, uint8_t* d, const VolDesc& src, const uint8_t* s, size_t n);
cudaError_t launchConvertRows(const LaunchCtx&, const VolDesc& dst, const VolDesc& src, const int32_t first[3],
                              const int32_t last[3], const int32_t off[3]);
cudaError_t launchMemsetRange(const LaunchCtx&, void* dst, const void* patternDev, size_t dstBytes, size_t patBytes);
cudaError_t launchFlip(const LaunchCtx&, const VolDesc& dst, const VolDesc& src, const int32_t first[3], const int32_t last[3],
                       const int32_t off[3], int axis);
cudaError_t launchHistogram(const LaunchCtx&, const VolDesc& vol, const int32_t first[3], const int32_t last[3],
                            unsigned long long* bins, size_t numBins, unsigned long long* droppedDev);
cudaError_t launchHistCombine(const LaunchCtx& c, unsigned long long* bins, unsigned long long* dropped,
                              const unsigned long long* parts, int nParts, size_t numBins);
cudaError_t launchScan(const LaunchCtx&, const VolDesc& dst, const VolDesc& src, const int32_t first[3], const int32_t last[3],
                       const int32_t off[3]);
// Aggregates: combinable moments of a range (kernels/aggregates.cu).  minIdx / maxIdx are linear indices inside the
// range (x fastest), ~0 when no voxel was below FLT_MAX (above -FLT_MAX).
struct AggPartial
{
    double sumD, sumD2;          // sum (v - shift), sum (v - shift)^2
    double prodMant;             // prod = prodMant * 2^prodExp
    long long prodExp;
    unsigned long long minIdx, maxIdx;
    float min, max, shift;
    int pad;
};
size_t aggregatesScratchBytes(const LaunchCtx&);
// scratchDev: aggregatesScratchBytes() of device memory; *totalDev points at the range's total inside it
cudaError_t launchAggregates(const LaunchCtx&, const VolDesc& vol, const int32_t first[3], const int32_t last[3],
                             AggPartial* scratchDev, AggPartial** totalDev);
// Scale / Rotate: dst voxel p -> M (p - center) + center, sampled with the reference's sampleLinear when it falls
// inside [first, last); m = col0.xyz, col1.xyz, col2.xyz.  dst must not overlap src (the C-ABI copies src then).
cudaError_t launchAffine(const LaunchCtx&, const VolDesc& dst, const VolDesc& src, const int32_t first[3],
                         const int32_t last[3], const float m[9], const float center[3]);
// ApplyFilter over [0, src dims); weights: fdims[0] * fdims[1] * fdims[2] HOST floats, x fastest; fdims in 1..3
cudaError_t launchFilter(const LaunchCtx&, const VolDesc& dst, const VolDesc& src, const float* weights,
                         const int32_t fdims[3]);
// Headless ray-marcher (kernels/render.cu): one frame of pixel-centre pinhole rays, RGBA8 and / or RGBA32F out
bool renderFormatSupported(int format);
void renderDefaultCamera(const int32_t dims[3], const float dist[3], float eye[3], float center[3], float up[3]);
cudaError_t launchRayMarch(const LaunchCtx&, const VolDesc& vol, const float dist[3], const float eye[3], const float center[3],
                           const float up[3], float fovyDegrees, const float* lutDev, int lutN, float dt, int srgb, int width,
                           int height, uint8_t* rgba8, float* rgba32f);
cudaError_t launchDebugCodec(const LaunchCtx&, int format, Mapping m, const float* values, uint32_t* mapped,
                             float* unmapped, size_t n);

// number of built-in transform ops (transform.cu)
constexpr int kNumUnaryOps = 4;
constexpr int kNumBinaryOps = 2;

// 64-bit voxel offset of (x,y,z) in a volume, reference StructuredVolume.cpp:311-317
inline long long voxelIndex(const VolDesc& v, long long x, long long y, long long z)
{
    return (z * v.dims[1] + y) * (long long)v.dims[0] + x;
}

// ---- geometry helpers shared by the launchers ----
inline bool overlaps(const VolDesc& a, const VolDesc& b)
{
    size_t na = (size_t)a.dims[0] * a.dims[1] * a.dims[2] * bytesPerVoxel(a.format);
    size_t nb = (size_t)b.dims[0] * b.dims[1] * b.dims[2] * bytesPerVoxel(b.format);
    return a.data < b.data + nb && b.data < a.data + na;
}

inline Operand operandAt(const VolDesc& v, long long x, long long y, long long z)
{
    const int bpv = bytesPerVoxel(v.format);
    Operand o;
    o.base = v.data + voxelIndex(v, x, y, z) * bpv;
    o.pitchY = (long long)v.dims[0] * bpv;
    o.pitchZ = (long long)v.dims[0] * v.dims[1] * bpv;
    return o;
}

// the range is one contiguous run in volume v
inline bool contiguous(const VolDesc& v, const Range& g)
{
    if (g.ny == 1 && g.nz == 1) return true;
    if (g.nx != v.dims[0]) return false;
    return g.nz == 1 || g.ny == v.dims[1];
}


} // namespace vktb
