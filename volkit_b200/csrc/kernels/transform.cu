// Transform / TransformRange (unary and binary) with built-in device ops.
//
// Semantics: reference src/vkt/Transform_serial.hpp:15-50 (unary) and :53-103 (binary): for every voxel of
// [first,last) an 8-byte zeroed scratch receives the voxel's raw bytes, the op is called with the ABSOLUTE
// coordinates and a VoxelView{bytes, format, lo, hi}, and the bytes are written back.  A host function
// pointer cannot run on the device, so ops are device functors selected by handle; the
// catalogue below is mirrored for the CPU checkers in oracle/ref_shim.cpp and oracle/vkt_oracle.c.
//
// The voxel's bytes travel in the low bytes of a 32-bit register (little endian, bpv <= 4).
// Algorithmic traffic: unary 2*bpv, binary 2*(bpv1 + bpv2) bytes per voxel.
#include "common.cuh"
#include "launch.hpp"
#include "voxel_codec.cuh"

namespace vktb {

namespace {

struct DevVoxel { uint32_t raw; int format; float lo, hi; };

template <int OP>
__device__ __forceinline__ uint32_t unaryOp(int x, int y, int z, uint32_t raw, int fmt, uint32_t mask, const Mapping& m)
{
    if (OP == 0) return raw ^ (uint32_t)((x + y + z) & 0xff);                    // xor_coord: bytes[0] ^= x+y+z
    if (OP == 1) return ~raw & mask;                                             // invert
    if (OP == 2) return ((((x >> 3) + (y >> 3) + (z >> 3)) & 1) == 0) ? 0u : raw; // checker8
    // threshold through the voxel mapping: value < 0.5*(lo+hi) -> lo else hi
    float val = unmapVoxelRt(fmt, raw, m);
    float mid = __fmul_rn(0.5f, __fadd_rn(m.lo, m.hi));
    return mapVoxelRt(fmt, val < mid ? m.lo : m.hi, m);
}

struct T1Params { Operand v; Range g; Mapping m; int fmt; };

template <int BPV, int OP>
__global__ void __launch_bounds__(kThreads)
transform1Kernel(T1Params p)
{
    constexpr int V = 16 / BPV;
    constexpr uint32_t mask = BPV == 4 ? 0xffffffffu : (1u << (BPV * 8)) - 1u;
    forEachRowChunk<V, BPV>(p.g, p.v, [&](int y, int z, int xr, int n, bool full) {
        uint8_t* dp = p.v.base + y * p.v.pitchY + z * p.v.pitchZ + (long long)xr * BPV;
        const int ax = p.g.x0 + xr, ay = p.g.y0 + y, az = p.g.z0 + z;
        if (full) {
            uint4 in = ldPlain(dp);
            uint32_t w[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
            for (int i = 0; i < V; ++i)
                vecSet<BPV>(w, i, unaryOp<OP>(ax + i, ay, az, vecGet<BPV>(in, i), p.fmt, mask, p.m));
            stStream(dp, make_uint4(w[0], w[1], w[2], w[3]));
        } else {
            for (int i = 0; i < n; ++i)
                storeRaw(dp + i * BPV, BPV, unaryOp<OP>(ax + i, ay, az, loadRaw(dp + i * BPV, BPV), p.fmt, mask, p.m));
        }
    });
}

// binary ops act on bytesPerVoxel(volume1.format) bytes of both scratch buffers (mask1)
template <int OP>
__device__ __forceinline__ void binaryOp(uint32_t& a, uint32_t& b, uint32_t mask1)
{
    if (OP == 0) { a |= (b & mask1); b = (b & ~mask1) | (a & mask1); }                   // or_assign
    else { uint32_t t = a & mask1; a = (a & ~mask1) | (b & mask1); b = (b & ~mask1) | t; } // swap
}

struct T2Params { Operand v1, v2; Range g; int b1, b2; };

template <int BPV, int OP>
__global__ void __launch_bounds__(kThreads)
transform2Kernel(T2Params p)
{
    constexpr int V = 16 / BPV;
    constexpr uint32_t mask = BPV == 4 ? 0xffffffffu : (1u << (BPV * 8)) - 1u;
    forEachRowChunk<V, BPV>(p.g, p.v1, [&](int y, int z, int xr, int n, bool full) {
        uint8_t* p1 = p.v1.base + y * p.v1.pitchY + z * p.v1.pitchZ + (long long)xr * BPV;
        uint8_t* p2 = p.v2.base + y * p.v2.pitchY + z * p.v2.pitchZ + (long long)xr * BPV;
        if (full && (reinterpret_cast<uintptr_t>(p2) & 15u) == 0) {
            uint4 i1 = ldPlain(p1), i2 = ldPlain(p2);
            uint32_t w1[4] = { 0u, 0u, 0u, 0u }, w2[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
            for (int i = 0; i < V; ++i) {
                uint32_t a = vecGet<BPV>(i1, i), b = vecGet<BPV>(i2, i);
                binaryOp<OP>(a, b, mask);
                vecSet<BPV>(w1, i, a); vecSet<BPV>(w2, i, b);
            }
            stStream(p1, make_uint4(w1[0], w1[1], w1[2], w1[3]));
            stStream(p2, make_uint4(w2[0], w2[1], w2[2], w2[3]));
        } else {
            for (int i = 0; i < n; ++i) {
                uint32_t a = loadRaw(p1 + i * BPV, BPV), b = loadRaw(p2 + i * BPV, BPV);
                binaryOp<OP>(a, b, mask);
                storeRaw(p1 + i * BPV, BPV, a); storeRaw(p2 + i * BPV, BPV, b);
            }
        }
    });
}

// mixed bytes-per-voxel: one voxel per thread
template <int OP>
__global__ void __launch_bounds__(kThreads)
transform2GenericKernel(T2Params p)
{
    const uint32_t mask1 = p.b1 == 4 ? 0xffffffffu : (1u << (p.b1 * 8)) - 1u;
    const long long rows = (long long)p.g.ny * p.g.nz;
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const int z = (int)(r / p.g.ny), y = (int)(r - (long long)z * p.g.ny);
        uint8_t* r1 = p.v1.base + y * p.v1.pitchY + z * p.v1.pitchZ;
        uint8_t* r2 = p.v2.base + y * p.v2.pitchY + z * p.v2.pitchZ;
        for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < p.g.nx; x += gridDim.x * blockDim.x) {
            uint32_t a = loadRaw(r1 + (long long)x * p.b1, p.b1), b = loadRaw(r2 + (long long)x * p.b2, p.b2);
            binaryOp<OP>(a, b, mask1);
            storeRaw(r1 + (long long)x * p.b1, p.b1, a); storeRaw(r2 + (long long)x * p.b2, p.b2, b);
        }
    }
}

template <int BPV>
cudaError_t launchT1(const LaunchCtx& c, const T1Params& p, int op)
{
    RowsLaunch l = rowsLaunch(p.g.nx, (long long)p.g.ny * p.g.nz, 16 / BPV);
    switch (op) {
    case 0: transform1Kernel<BPV, 0><<<l.grid, l.block, 0, c.stream>>>(p); break;
    case 1: transform1Kernel<BPV, 1><<<l.grid, l.block, 0, c.stream>>>(p); break;
    case 2: transform1Kernel<BPV, 2><<<l.grid, l.block, 0, c.stream>>>(p); break;
    case 3: transform1Kernel<BPV, 3><<<l.grid, l.block, 0, c.stream>>>(p); break;
    default: return cudaErrorInvalidValue;
    }
    countLaunch();
    return cudaGetLastError();
}

template <int BPV>
cudaError_t launchT2(const LaunchCtx& c, const T2Params& p, int op)
{
    RowsLaunch l = rowsLaunch(p.g.nx, (long long)p.g.ny * p.g.nz, 16 / BPV);
    if (op == 0) transform2Kernel<BPV, 0><<<l.grid, l.block, 0, c.stream>>>(p);
    else if (op == 1) transform2Kernel<BPV, 1><<<l.grid, l.block, 0, c.stream>>>(p);
    else return cudaErrorInvalidValue;
    countLaunch();
    return cudaGetLastError();
}

bool makeRange(Range& g, const int32_t first[3], const int32_t last[3])
{
    g.nx = last[0] - first[0]; g.ny = last[1] - first[1]; g.nz = last[2] - first[2];
    g.x0 = first[0]; g.y0 = first[1]; g.z0 = first[2];
    return g.nx > 0 && g.ny > 0 && g.nz > 0;
}

} // namespace

cudaError_t launchTransform1(const LaunchCtx& c, const VolDesc& vol, const int32_t first[3], const int32_t last[3], int op)
{
    T1Params p;
    if (!makeRange(p.g, first, last)) return cudaSuccess;
    p.v = operandAt(vol, first[0], first[1], first[2]);
    p.m = vol.map; p.fmt = vol.format;
    switch (bytesPerVoxel(vol.format)) {
    case 1: return launchT1<1>(c, p, op);
    case 2: return launchT1<2>(c, p, op);
    case 4: return launchT1<4>(c, p, op);
    }
    return cudaErrorInvalidValue;
}

cudaError_t launchTransform2(const LaunchCtx& c, const VolDesc& v1, const VolDesc& v2, const int32_t first[3],
                             const int32_t last[3], int op)
{
    T2Params p;
    if (!makeRange(p.g, first, last)) return cudaSuccess;
    p.v1 = operandAt(v1, first[0], first[1], first[2]);
    p.v2 = operandAt(v2, first[0], first[1], first[2]);
    p.b1 = bytesPerVoxel(v1.format); p.b2 = bytesPerVoxel(v2.format);
    if (p.b1 == 0 || p.b2 == 0 || op < 0 || op >= kNumBinaryOps) return cudaErrorInvalidValue;
    if (p.b1 == p.b2) {
        switch (p.b1) {
        case 1: return launchT2<1>(c, p, op);
        case 2: return launchT2<2>(c, p, op);
        case 4: return launchT2<4>(c, p, op);
        }
    }
    long long rows = (long long)p.g.ny * p.g.nz;
    dim3 grid((unsigned)((p.g.nx + kThreads - 1) / kThreads), (unsigned)(rows > 65535 ? 65535 : rows), 1);
    if (grid.x > 1024) grid.x = 1024;
    if (op == 0) transform2GenericKernel<0><<<grid, kThreads, 0, c.stream>>>(p);
    else transform2GenericKernel<1><<<grid, kThreads, 0, c.stream>>>(p);
    countLaunch();
    return cudaGetLastError();
}

} // namespace vktb
