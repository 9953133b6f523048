// Box table of vktb_copy_range_batch (BrickDecompose: n CopyRange calls of one source in one launch), shared by the
// two kernels that consume it (stream_ops.cu: direct loads; copy_bricks.cu: TMA-staged).
#pragma once

#include <cstddef>
#include <cstdint>

#include <cuda.h>
#include <cuda_runtime.h>

#include "launch.hpp"

namespace vktb {

struct BoxRec
{
    uint8_t*  d;                // dst voxel that receives the box's first voxel
    long long pitchY, pitchZ;   // dst pitches in bytes
    int nx, ny, nz;             // extents of [first, last)
    int x0, y0, z0;             // first (may be negative: clamped)
    unsigned nvox;              // flat kernel: nx*ny*nz (< 2^31)
    unsigned mulX, mulY;        // e / nx == (e * mulX) >> shX for e < 2^31 (likewise ny): no division per vector
    int shX, shY;
};

// cuTensorMapEncodeTiled through the runtime (libcuda is never linked); nullptr when the driver has none
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encodeTiled();

// Boxes that are one contiguous run of dst bytes each (`dev[0 .. n)`, already in device memory), staged through
// shared memory by the TMA unit.  cudaErrorNotSupported: the operands do not qualify, use copyBoxesFlatKernel.
cudaError_t launchCopyBricksTma(const LaunchCtx& c, const BoxRec* dev, const BoxRec* host, size_t n, const uint8_t* src,
                                int bpv, const int32_t sdims[3]);

} // namespace vktb
