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

template <int BPV>
cudaError_t launchFlipBpv(const LaunchCtx& c, const FlipParams& p, int axis, bool vec)
{
    const long long rows = (long long)p.ry * p.rz;
    const unsigned gy = (unsigned)(rows > 65535 ? 65535 : rows);
    if (vec) {
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
    float binsF;             // (float)numBins
    unsigned long long* bins;
    unsigned long long* dropped;
    unsigned numBins;
};

__device__ __forceinline__ int histBin(float val, const HistParams& p)
{
    const float b = __fmul_rn(__fsub_rn(val, p.m.lo), p.scale);
    if (b >= 0.0f && b < p.binsF) return (int)b;          // truncation toward zero, as the (size_t) cast
    if (b > -1.0f && b < 0.0f) return 0;
    return -1;                                            // the reference would write out of bounds (NaN included)
}

template <int F, bool SHARED>
__global__ void __launch_bounds__(kThreads)
histogramKernel(HistParams p)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    extern __shared__ uint32_t sh[];
    if (SHARED) {
        for (unsigned i = threadIdx.x; i < p.numBins; i += blockDim.x) sh[i] = 0u;
        __syncthreads();
    }
    unsigned drops = 0;
    auto count = [&](uint32_t raw) {
        const int b = histBin(unmapVoxel<F, false>(raw, p.m), p);
        if (b < 0) { ++drops; return; }
        if (SHARED) atomicAdd(&sh[b], 1u);
        else atomicAdd(&p.bins[b], 1ull);
    };
    const long long rows = (long long)p.ny * p.nz;
    for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
        const int z = (int)(r / p.ny), y = (int)(r - (long long)z * p.ny);
        const uint8_t* row = p.s + y * p.pitchY + z * p.pitchZ;
        // scalar head up to the first 16-byte boundary, 128-bit body, scalar tail
        int head = (int)(((16u - (reinterpret_cast<uintptr_t>(row) & 15u)) & 15u) / BPV);
        if (head > p.nx) head = p.nx;
        const int nvec = (p.nx - head) / V;
        const int tail0 = head + nvec * V;
        const uint8_t* body = row + (long long)head * BPV;
        for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
            const uint4 q = ldStream(body + v * 16LL);
#pragma unroll
            for (int i = 0; i < V; ++i) count(vecGet<BPV>(q, i));
        }
        const int nEdge = head + (p.nx - tail0);
        for (int e = threadIdx.x; e < nEdge; e += blockDim.x) {
            const int x = e < head ? e : tail0 + (e - head);
            count(loadRaw(row + (long long)x * BPV, BPV));
        }
    }
    if (drops) atomicAdd(p.dropped, (unsigned long long)drops);
    if (SHARED) {
        __syncthreads();
        for (unsigned i = threadIdx.x; i < p.numBins; i += blockDim.x)
            if (sh[i]) atomicAdd(&p.bins[i], (unsigned long long)sh[i]);
    }
}

template <int F>
cudaError_t launchHistF(const LaunchCtx& c, const HistParams& p)
{
    const long long rows = (long long)p.ny * p.nz;
    // persistent: a few CTAs per SM, every CTA flushes its shared counters once
    long long grid = (long long)c.smCount * 8;
    if (grid > rows) grid = rows;
    const size_t shBytes = (size_t)p.numBins * 4;
    const bool shared = shBytes <= 96 * 1024;
    if (shared) {
        auto k = histogramKernel<F, true>;
        if (shBytes > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shBytes);
            if (e != cudaSuccess) return e;
        }
        k<<<(unsigned)grid, kThreads, shBytes, c.stream>>>(p);
    } else {
        histogramKernel<F, false><<<(unsigned)grid, kThreads, 0, c.stream>>>(p);
    }
    countLaunch();
    return cudaGetLastError();
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
    const bool vec = axis != 0 && ((long long)p.rx * bpv) % 16 == 0
                  && ((reinterpret_cast<uintptr_t>(p.s) | reinterpret_cast<uintptr_t>(p.d)) & 15u) == 0
                  && ((p.sPitchY | p.sPitchZ | p.dPitchY | p.dPitchZ) & 15) == 0;
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
    p.binsF = (float)numBins;
    p.bins = bins; p.dropped = droppedDev; p.numBins = (unsigned)numBins;
    switch (vol.format) {
    case F_U8:    return launchHistF<F_U8>(c, p);
    case F_U16:   return launchHistF<F_U16>(c, p);
    case F_INT16: return launchHistF<F_INT16>(c, p);
    case F_U32:   return launchHistF<F_U32>(c, p);
    case F_F32:   return launchHistF<F_F32>(c, p);
    }
    return cudaErrorInvalidValue;
}

} // namespace vktb
