// Flip / FlipRange and ComputeHistogram / ComputeHistogramRange -- the first two "next" rows of SURVEY.md 8(f).
//
// Flip follows the reference's serial body AS WRITTEN (src/vkt/Flip_serial.hpp:11-63), quirks included: the loops run
// over [0, range) with range = last - first halved on the flip axis (first is otherwise ignored), the mirrored index is
// last-1-i, and the z index is mirrored when the axis is Y, not Z -- so Axis::Y mirrors y and z together and Axis::Z
// copies the lower half of the z range onto itself.  Every visited voxel is read together with its partner and the two
// are written crosswise at + dstOffset.  Visited voxels and partners are disjoint sets, so the pairs are independent:
// one thread per pair (or per 16-byte chunk of pairs when x is not mirrored), in place included.
// Algorithmic traffic: 2 * bpv per visited voxel and per partner.
//
// Histogram follows src/vkt/Histogram_serial.hpp:18-51: bin = (size_t)((value - lo) * (numBins / (hi - lo))) in
// binary32 with value = the unmapped voxel; counts are 64-bit.  Voxels whose index falls outside [0, numBins) make the
// reference write out of bounds; here they are dropped and counted.  Per-CTA shared-memory counters (32-bit) absorb the
// atomics when the bins fit in shared memory; they are flushed to the global 64-bit bins once per CTA.
// Algorithmic traffic: bpv per voxel.
#include <cstdlib>

#include "common.cuh"
#include "launch.hpp"
#include "voxel_codec.cuh"

namespace vktb {

namespace {

struct FlipParams
{
    uint8_t*       d;        // dst voxel (off.x, off.y, off.z)
    const uint8_t* s;        // src voxel (0, 0, 0)
    long long dPitchY, dPitchZ, sPitchY, sPitchZ;
    int rx, ry, rz;          // visited extents (halved on the axis)
    int lx, ly, lz;          // last - 1 per axis
};

// generic: one voxel pair per thread
template <int BPV, int AXIS>
__global__ void __launch_bounds__(kThreads)
flipPairsKernel(FlipParams p)
{
    const long long rows = (long long)p.ry * p.rz;
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const int z = (int)(r / p.ry), y = (int)(r - (long long)z * p.ry);
        const int yy = AXIS == 1 ? p.ly - y : y;
        const int zz = AXIS == 1 ? p.lz - z : z;           // the reference mirrors z on Axis::Y (Flip_serial.hpp:49)
        const uint8_t* s1 = p.s + y * p.sPitchY + z * p.sPitchZ;
        const uint8_t* s2 = p.s + yy * p.sPitchY + zz * p.sPitchZ;
        uint8_t* d1 = p.d + y * p.dPitchY + z * p.dPitchZ;
        uint8_t* d2 = p.d + yy * p.dPitchY + zz * p.dPitchZ;
        for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < p.rx; x += gridDim.x * blockDim.x) {
            const int xx = AXIS == 0 ? p.lx - x : x;
            const uint32_t v1 = loadRaw(s1 + (long long)x * BPV, BPV);
            const uint32_t v2 = loadRaw(s2 + (long long)xx * BPV, BPV);
            storeRaw(d1 + (long long)x * BPV, BPV, v2);
            storeRaw(d2 + (long long)xx * BPV, BPV, v1);
        }
    }
}

// x not mirrored, every row start and the row length multiples of 16 bytes: whole rows swap as 128-bit vectors
template <int AXIS>
__global__ void __launch_bounds__(kThreads)
flipRowsVecKernel(FlipParams p, int vecsPerRow)
{
    const long long rows = (long long)p.ry * p.rz;
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const int z = (int)(r / p.ry), y = (int)(r - (long long)z * p.ry);
        const int yy = AXIS == 1 ? p.ly - y : y;
        const int zz = AXIS == 1 ? p.lz - z : z;
        const uint8_t* s1 = p.s + y * p.sPitchY + z * p.sPitchZ;
        const uint8_t* s2 = p.s + yy * p.sPitchY + zz * p.sPitchZ;
        uint8_t* d1 = p.d + y * p.dPitchY + z * p.dPitchZ;
        uint8_t* d2 = p.d + yy * p.dPitchY + zz * p.dPitchZ;
        for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < vecsPerRow; v += gridDim.x * blockDim.x) {
            const uint4 a = ldPlain(s1 + v * 16LL), b = ldPlain(s2 + v * 16LL);   // in place: coherent loads
            stStream(d1 + v * 16LL, b);
            stStream(d2 + v * 16LL, a);
        }
    }
}

// 16 bytes with their BPV-sized elements in reverse order
template <int BPV>
__device__ __forceinline__ uint4 reverseElems(const uint4& v)
{
    if (BPV == 4) return make_uint4(v.w, v.z, v.y, v.x);
    const unsigned sel = BPV == 2 ? 0x1032u : 0x0123u;
    return make_uint4(__byte_perm(v.w, 0u, sel), __byte_perm(v.z, 0u, sel), __byte_perm(v.y, 0u, sel), __byte_perm(v.x, 0u, sel));
}

// x mirrored, row starts / pitches / last.x * bpv multiples of 16 bytes: a 16-byte chunk of visited voxels and the
// 16-byte chunk of their partners are exchanged with their elements reversed; the (rx mod V) voxels next to the
// middle go through the scalar path
template <int BPV>
__global__ void __launch_bounds__(kThreads)
flipXVecKernel(FlipParams p, int vecsPerRow)
{
    constexpr int V = 16 / BPV;
    const long long rows = (long long)p.ry * p.rz;
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const int z = (int)(r / p.ry), y = (int)(r - (long long)z * p.ry);
        const uint8_t* srow = p.s + y * p.sPitchY + z * p.sPitchZ;
        uint8_t* drow = p.d + y * p.dPitchY + z * p.dPitchZ;
        for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < vecsPerRow; v += gridDim.x * blockDim.x) {
            const long long lo = (long long)v * 16;                              // voxels x = v*V .. v*V + V-1
            const long long hi = ((long long)p.lx + 1 - (long long)(v + 1) * V) * BPV;   // partners lx - x, ascending
            const uint4 a = ldPlain(srow + lo), b = ldPlain(srow + hi);
            stStream(drow + lo, reverseElems<BPV>(b));
            stStream(drow + hi, reverseElems<BPV>(a));
        }
        for (int x = vecsPerRow * V + blockIdx.x * blockDim.x + threadIdx.x; x < p.rx; x += gridDim.x * blockDim.x) {
            const int xx = p.lx - x;
            const uint32_t v1 = loadRaw(srow + (long long)x * BPV, BPV), v2 = loadRaw(srow + (long long)xx * BPV, BPV);
            storeRaw(drow + (long long)x * BPV, BPV, v2);
            storeRaw(drow + (long long)xx * BPV, BPV, v1);
        }
    }
}

template <int BPV>
cudaError_t launchFlipBpv(const LaunchCtx& c, const FlipParams& p, int axis, bool vec)
{
    const long long rows = (long long)p.ry * p.rz;
    const unsigned gy = (unsigned)(rows > 65535 ? 65535 : rows);
    if (vec && axis == 0) {
        const int vecs = (int)((long long)p.rx * BPV / 16);
        dim3 grid((unsigned)((vecs + kThreads) / kThreads), gy, 1);
        flipXVecKernel<BPV><<<grid, kThreads, 0, c.stream>>>(p, vecs);
    } else if (vec) {
        const int vecs = (int)((long long)p.rx * BPV / 16);
        dim3 grid((unsigned)((vecs + kThreads - 1) / kThreads), gy, 1);
        if (axis == 1) flipRowsVecKernel<1><<<grid, kThreads, 0, c.stream>>>(p, vecs);
        else flipRowsVecKernel<2><<<grid, kThreads, 0, c.stream>>>(p, vecs);
    } else {
        dim3 grid((unsigned)((p.rx + kThreads - 1) / kThreads), gy, 1);
        if (grid.x > 1024) grid.x = 1024;
        if (axis == 0) flipPairsKernel<BPV, 0><<<grid, kThreads, 0, c.stream>>>(p);
        else if (axis == 1) flipPairsKernel<BPV, 1><<<grid, kThreads, 0, c.stream>>>(p);
        else flipPairsKernel<BPV, 2><<<grid, kThreads, 0, c.stream>>>(p);
    }
    countLaunch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------- histogram
struct HistParams
{
    const uint8_t* s;        // voxel (first.x, first.y, first.z)
    long long pitchY, pitchZ;
    int nx, ny, nz;
    int format;
    Mapping m;
    float scale;             // (float)numBins / (hi - lo), binary32 (host)
    unsigned long long* bins;
    unsigned long long* dropped;
    unsigned numBins;
    size_t n;                // voxels of the range
    int flat;                // the range is one contiguous run
};

// bin index of an unmapped value, or -1 where the reference would write out of bounds.  cvt.rzi saturates and
// truncates toward zero like the (size_t) cast does for in-range values: b in (-1, 0) -> 0, b <= -1 -> negative,
// b >= numBins -> >= numBins; NaN converts to 0 and is rejected separately.
template <int F>
__device__ __forceinline__ int histBin(float val, const HistParams& p)
{
    const float b = __fmul_rn(__fsub_rn(val, p.m.lo), p.scale);
    const int i = __float2int_rz(b);
    if (b != b) return -1;                                // Float32 NaN voxels, or a degenerate mapping (hi == lo)
    return (unsigned)i < p.numBins ? i : -1;
}

// SHARED: per-CTA 32-bit counters in shared memory, REPL copies of every bin laid out as sh[bin * REPL + lane % REPL].
// With REPL = 32 lane l only ever touches bank l: no bank conflicts and no same-address serialisation inside a warp
// whatever the data looks like (a warp of equal voxels would otherwise serialise 32-fold).
// What bounds this kernel is the BIN COMPUTATION, not the shared-memory reductions (round 2 measured it: with the
// reference's float expression per voxel 2048^3 UInt16 takes 7.5 ms; a kernel with lane-private 16-bit counters and no
// atomics at all, built on the opposite assumption, took the same 7.5 ms and 4.5 ms with an integer bin -- this kernel
// with the same integer bin takes 2.98 ms):
//   * identity-mapped UInt16, up to 256 bins: unmap = raw * 2^-16 exactly (65535.999f rounds to 65536), lo = 0, scale =
//     numBins exactly, and raw * numBins < 2^24 is exact in binary32, so the reference's bin IS (raw * numBins) >> 16;
//   * UInt8, any mapping, any bin count: the kernel counts the 256 byte VALUES (extract, one reduction, nothing else);
//     each value's total goes through the exact float bin expression once, at the end.
// FLAT: the range is one contiguous run of p.n voxels (grid-stride over 128-bit vectors, four in flight); otherwise
// one CTA walks rows.
template <int F, bool IDENT, bool SHARED, bool FLAT>
__global__ void __launch_bounds__(kThreads)
histogramKernel(HistParams p, int replShift)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    // UInt8 with shared counters: count the 256 byte VALUES (one shared-memory reduction per voxel, nothing else); each
    // value's count goes through the exact bin expression once, at the end
    constexpr bool kByValue = F == F_U8 && SHARED;
    extern __shared__ uint32_t sh[];
    const unsigned nRows = kByValue ? 256u : p.numBins;
    const unsigned nCounters = nRows << replShift;
    const unsigned myCopy = threadIdx.x & ((1u << replShift) - 1u);
    if (SHARED) {
        for (unsigned i = threadIdx.x; i < nCounters; i += blockDim.x) sh[i] = 0u;
        __syncthreads();
    }
    unsigned drops = 0;
    auto count = [&](uint32_t raw) {
        if (kByValue) { atomicAdd(&sh[(raw << replShift) + myCopy], 1u); return; }
        int b;
        if (F == F_U16 && IDENT && p.numBins <= 256u) b = (int)((raw * p.numBins) >> 16);   // exact, see above
        else b = histBin<F>(unmapVoxel<F, IDENT>(raw, p.m), p);
        if (b < 0) { ++drops; return; }
        if (SHARED) atomicAdd(&sh[((unsigned)b << replShift) + myCopy], 1u);
        else atomicAdd(&p.bins[b], 1ull);
    };
    auto countVec = [&](const uint4& q) {
#pragma unroll
        for (int i = 0; i < V; ++i) count(vecGet<BPV>(q, i));
    };
    if (FLAT) {
        const size_t n = p.n;
        size_t head = ((16u - (reinterpret_cast<uintptr_t>(p.s) & 15u)) & 15u) / BPV;
        if (head > n) head = n;
        const size_t nvec = (n - head) / V, tail0 = head + nvec * V;
        const uint8_t* body = p.s + head * BPV;
        const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
        size_t i = tid;
        for (; i + 3 * stride < nvec; i += 4 * stride) {
            const uint4 q0 = ldStream(body + i * 16), q1 = ldStream(body + (i + stride) * 16);
            const uint4 q2 = ldStream(body + (i + 2 * stride) * 16), q3 = ldStream(body + (i + 3 * stride) * 16);
            countVec(q0); countVec(q1); countVec(q2); countVec(q3);
        }
        for (; i < nvec; i += stride) countVec(ldStream(body + i * 16));
        const size_t nEdge = head + (n - tail0);
        for (size_t e = tid; e < nEdge; e += stride) {
            const size_t k = e < head ? e : tail0 + (e - head);
            count(loadRaw(p.s + k * BPV, BPV));
        }
    } else {
        const long long rows = (long long)p.ny * p.nz;
        for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
            const int z = (int)(r / p.ny), y = (int)(r - (long long)z * p.ny);
            const uint8_t* row = p.s + y * p.pitchY + z * p.pitchZ;
            // scalar head up to the first 16-byte boundary, 128-bit body (two vectors in flight), scalar tail
            int head = (int)(((16u - (reinterpret_cast<uintptr_t>(row) & 15u)) & 15u) / BPV);
            if (head > p.nx) head = p.nx;
            const int nvec = (p.nx - head) / V;
            const int tail0 = head + nvec * V;
            const uint8_t* body = row + (long long)head * BPV;
            int v = threadIdx.x;
            for (; v + (int)blockDim.x < nvec; v += 2 * blockDim.x) {
                const uint4 q0 = ldStream(body + v * 16LL), q1 = ldStream(body + (v + (int)blockDim.x) * 16LL);
                countVec(q0); countVec(q1);
            }
            for (; v < nvec; v += blockDim.x) countVec(ldStream(body + v * 16LL));
            const int nEdge = head + (p.nx - tail0);
            for (int e = threadIdx.x; e < nEdge; e += blockDim.x) {
                const int x = e < head ? e : tail0 + (e - head);
                count(loadRaw(row + (long long)x * BPV, BPV));
            }
        }
    }
    if (drops) atomicAdd(p.dropped, (unsigned long long)drops);
    if (SHARED) {
        __syncthreads();
        for (unsigned i = threadIdx.x; i < nRows; i += blockDim.x) {
            unsigned long long sum = 0;
            for (unsigned k = 0; k < (1u << replShift); ++k) sum += sh[(i << replShift) + ((k + i) & ((1u << replShift) - 1u))];
            if (!sum) continue;
            const int b = kByValue ? histBin<F>(unmapVoxel<F, IDENT>(i, p.m), p) : (int)i;
            if (b >= 0) atomicAdd(&p.bins[b], sum);
            else atomicAdd(p.dropped, sum);
        }
    }
}

template <int F, bool IDENT, bool SHARED>
cudaError_t launchHistK(const LaunchCtx& c, const HistParams& p, unsigned grid, size_t shBytes, int replShift)
{
    auto k = p.flat ? histogramKernel<F, IDENT, SHARED, true> : histogramKernel<F, IDENT, SHARED, false>;
    if (shBytes > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shBytes);
        if (e != cudaSuccess) return e;
    }
    k<<<grid, kThreads, shBytes, c.stream>>>(p, replShift);
    countLaunch();
    return cudaGetLastError();
}

template <int F>
cudaError_t launchHistF(const LaunchCtx& c, const HistParams& p)
{
    const long long rows = (long long)p.ny * p.nz;
    constexpr size_t kMaxShared = 96 * 1024;
    const size_t rowsSh = F == F_U8 ? 256 : p.numBins;            // UInt8 counts byte values, whatever numBins is
    const bool shared = rowsSh * 4 <= kMaxShared;
    // as many copies per bin as fit in 32 KiB (so several CTAs stay resident), at most one per lane
    int replShift = 0;
    while (shared && replShift < 5 && ((rowsSh * 4) << (replShift + 1)) <= 32 * 1024) ++replShift;
    const size_t shBytes = shared ? ((rowsSh * 4) << replShift) : 0;
    // persistent: a few CTAs per SM, every CTA flushes its shared counters once
    long long grid = (long long)c.smCount * 6;
    if (!p.flat && grid > rows) grid = rows;
    const bool ident = p.m.identity != 0;
    if (shared) return ident ? launchHistK<F, true, true>(c, p, (unsigned)grid, shBytes, replShift)
                             : launchHistK<F, false, true>(c, p, (unsigned)grid, shBytes, replShift);
    return ident ? launchHistK<F, true, false>(c, p, (unsigned)grid, 0, 0)
                 : launchHistK<F, false, false>(c, p, (unsigned)grid, 0, 0);
}

} // namespace

cudaError_t launchFlip(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src, const int32_t first[3],
                       const int32_t last[3], const int32_t off[3], int axis)
{
    const int bpv = bytesPerVoxel(src.format);
    if (axis < 0 || axis > 2 || bpv != bytesPerVoxel(dst.format)) return cudaErrorInvalidValue;
    FlipParams p;
    p.rx = last[0] - first[0]; p.ry = last[1] - first[1]; p.rz = last[2] - first[2];
    if (axis == 0) p.rx >>= 1; else if (axis == 1) p.ry >>= 1; else p.rz >>= 1;
    if (p.rx <= 0 || p.ry <= 0 || p.rz <= 0) return cudaSuccess;
    p.lx = last[0] - 1; p.ly = last[1] - 1; p.lz = last[2] - 1;
    p.sPitchY = (long long)src.dims[0] * bpv; p.sPitchZ = p.sPitchY * src.dims[1];
    p.dPitchY = (long long)dst.dims[0] * bpv; p.dPitchZ = p.dPitchY * dst.dims[1];
    p.s = src.data;
    p.d = dst.data + voxelIndex(dst, off[0], off[1], off[2]) * bpv;
    const bool aligned = ((reinterpret_cast<uintptr_t>(p.s) | reinterpret_cast<uintptr_t>(p.d)) & 15u) == 0
                      && ((p.sPitchY | p.sPitchZ | p.dPitchY | p.dPitchZ) & 15) == 0;
    // rows swap whole (x not mirrored) when the row length is a multiple of 16 bytes; mirrored x needs the partner
    // chunks [last.x - (v+1)V, last.x - vV) on 16-byte boundaries too
    const bool vec = aligned && (axis != 0 ? ((long long)p.rx * bpv) % 16 == 0
                                           : ((long long)last[0] * bpv) % 16 == 0 && (long long)p.rx * bpv >= 16);
    switch (bpv) {
    case 1: return launchFlipBpv<1>(c, p, axis, vec);
    case 2: return launchFlipBpv<2>(c, p, axis, vec);
    case 4: return launchFlipBpv<4>(c, p, axis, vec);
    }
    return cudaErrorInvalidValue;
}

cudaError_t launchHistogram(const LaunchCtx& c, const VolDesc& vol, const int32_t first[3], const int32_t last[3],
                            unsigned long long* bins, size_t numBins, unsigned long long* droppedDev)
{
    if (numBins == 0 || numBins > (1u << 24)) return cudaErrorInvalidValue;   // (float)numBins must be exact
    cudaError_t e = cudaMemsetAsync(bins, 0, numBins * sizeof(unsigned long long), c.stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(droppedDev, 0, sizeof(unsigned long long), c.stream);
    if (e != cudaSuccess) return e;
    HistParams p;
    p.nx = last[0] - first[0]; p.ny = last[1] - first[1]; p.nz = last[2] - first[2];
    if (p.nx <= 0 || p.ny <= 0 || p.nz <= 0) return cudaSuccess;
    const int bpv = bytesPerVoxel(vol.format);
    p.pitchY = (long long)vol.dims[0] * bpv; p.pitchZ = p.pitchY * vol.dims[1];
    p.s = vol.data + voxelIndex(vol, first[0], first[1], first[2]) * bpv;
    p.format = vol.format; p.m = vol.map;
    {   // binary32, one rounding per operation (the reference's size_t -> float conversion is exact below 2^24)
        volatile float range = vol.map.hi - vol.map.lo;
        volatile float scale = (float)numBins / range;
        p.scale = scale;
    }
    p.bins = bins; p.dropped = droppedDev; p.numBins = (unsigned)numBins;
    p.n = (size_t)p.nx * p.ny * p.nz;
    Range g; g.nx = p.nx; g.ny = p.ny; g.nz = p.nz; g.x0 = first[0]; g.y0 = first[1]; g.z0 = first[2];
    p.flat = contiguous(vol, g) ? 1 : 0;
    switch (vol.format) {
    case F_U8:    return launchHistF<F_U8>(c, p);
    case F_U16:   return launchHistF<F_U16>(c, p);
    case F_INT16: return launchHistF<F_INT16>(c, p);
    case F_U32:   return launchHistF<F_U32>(c, p);
    case F_F32:   return launchHistF<F_F32>(c, p);
    }
    return cudaErrorInvalidValue;
}

// ---- several devices: bins[i] = sum over the parts (vktb_histogram_range on a striped volume, cabi.cu) ---------------------
// parts: nParts records of numBins + 1 counters (the bins, then the dropped voxels), already on this device
__global__ void __launch_bounds__(kThreads)
histCombineKernel(unsigned long long* __restrict__ bins, unsigned long long* __restrict__ dropped,
                  const unsigned long long* __restrict__ parts, int nParts, size_t numBins)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i <= numBins; i += (size_t)gridDim.x * blockDim.x) {
        unsigned long long sum = 0;
        for (int k = 0; k < nParts; ++k) sum += parts[(size_t)k * (numBins + 1) + i];
        if (i < numBins) bins[i] = sum;
        else *dropped = sum;
    }
}

cudaError_t launchHistCombine(const LaunchCtx& c, unsigned long long* bins, unsigned long long* dropped,
                              const unsigned long long* parts, int nParts, size_t numBins)
{
    const size_t ctas = (numBins + 1 + kThreads - 1) / kThreads;
    histCombineKernel<<<(unsigned)(ctas < 1024 ? ctas : 1024), kThreads, 0, c.stream>>>(bins, dropped, parts, nParts, numBins);
    countLaunch();
    return cudaGetLastError();
}

} // namespace vktb
