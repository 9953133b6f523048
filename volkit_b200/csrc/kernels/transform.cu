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

// Whole-vector form of an op for 16 bytes of ONE row starting at (x, y, z): ops that are byte-parallel get a
// few SIMD-in-word instructions instead of a per-voxel extract/compute/insert chain.
template <int BPV, int OP>
__device__ __forceinline__ uint4 unaryOpVec(int x, int y, int z, const uint4& in, int fmt, uint32_t mask, const Mapping& m)
{
    constexpr int V = 16 / BPV;
    const uint32_t iw[4] = { in.x, in.y, in.z, in.w };
    uint32_t w[4] = { 0u, 0u, 0u, 0u };
    if constexpr (OP == 0) {
        const uint32_t s = (uint32_t)(x + y + z);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if constexpr (BPV == 1) w[k] = iw[k] ^ __vadd4(((s + 4u * k) & 0xffu) * 0x01010101u, 0x03020100u);
            else if constexpr (BPV == 2) w[k] = iw[k] ^ (((s + 2u * k) & 0xffu) | (((s + 2u * k + 1u) & 0xffu) << 16));
            else w[k] = iw[k] ^ ((s + k) & 0xffu);
        }
    } else if constexpr (OP == 1) {
#pragma unroll
        for (int k = 0; k < 4; ++k) w[k] = ~iw[k];
    } else {
#pragma unroll
        for (int i = 0; i < V; ++i)
            vecSet<BPV>(w, i, unaryOp<OP>(x + i, y, z, vecGet<BPV>(in, i), fmt, mask, m));
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
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
            stStream(dp, unaryOpVec<BPV, OP>(ax, ay, az, ldPlain(dp), p.fmt, mask, p.m));
        } else {
            for (int i = 0; i < n; ++i)
                storeRaw(dp + i * BPV, BPV, unaryOp<OP>(ax + i, ay, az, loadRaw(dp + i * BPV, BPV), p.fmt, mask, p.m));
        }
    });
}

// boxed range whose rows share one 16-byte phase: four rows in flight per thread
template <int BPV, int OP>
__global__ void __launch_bounds__(kThreads)
transform1UniformKernel(T1Params p, int shift)
{
    constexpr int V = 16 / BPV;
    constexpr uint32_t mask = BPV == 4 ? 0xffffffffu : (1u << (BPV * 8)) - 1u;
    forEachRowChunkUniform<V, 4, uint4>(p.g, shift,
        [&](int y, int z, int xr) { return ldPlain(p.v.base + y * p.v.pitchY + z * p.v.pitchZ + (long long)xr * BPV); },
        [&](int y, int z, int xr, const uint4& q) {
            stStream(p.v.base + y * p.v.pitchY + z * p.v.pitchZ + (long long)xr * BPV,
                     unaryOpVec<BPV, OP>(p.g.x0 + xr, p.g.y0 + y, p.g.z0 + z, q, p.fmt, mask, p.m));
        },
        [&](int y, int z, int xr, int n) {
            uint8_t* dp = p.v.base + y * p.v.pitchY + z * p.v.pitchZ + (long long)xr * BPV;
            for (int i = 0; i < n; ++i)
                storeRaw(dp + i * BPV, BPV, unaryOp<OP>(p.g.x0 + xr + i, p.g.y0 + y, p.g.z0 + z, loadRaw(dp + i * BPV, BPV),
                                                       p.fmt, mask, p.m));
        });
}

// Contiguous ranges (whole volumes, whole planes): flat grid-stride stream, four 128-bit vectors in flight per thread.
// The op still sees absolute coordinates: they are rebuilt from the linear voxel index with 32-bit arithmetic.
template <int BPV, int OP>
__global__ void __launch_bounds__(kThreads)
transform1FlatKernel(uint8_t* __restrict__ d, size_t n, int nx, int ny, int x0, int y0, int z0, Mapping m, int fmt)
{
    constexpr int V = 16 / BPV;
    constexpr uint32_t mask = BPV == 4 ? 0xffffffffu : (1u << (BPV * 8)) - 1u;
    size_t head = ((16u - (reinterpret_cast<uintptr_t>(d) & 15u)) & 15u) / BPV;
    if (head > n) head = n;
    const size_t nvec = (n - head) / V;
    const size_t tail0 = head + nvec * V;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    uint8_t* dv = d + head * BPV;

    // coordinates advance with the grid stride by add-with-carry: one 64-bit division pair per thread, none per vector
    struct Coord { int x, y, z; };
    auto coordOf = [&](size_t v) {
        const size_t row = v / (unsigned)nx, zr = row / (unsigned)ny;
        return Coord{ (int)(v - row * (unsigned)nx), (int)(row - zr * (unsigned)ny), (int)zr };
    };
    const Coord step = coordOf(stride * V);
    auto advance = [&](Coord& c) {
        c.x += step.x; if (c.x >= nx) { c.x -= nx; ++c.y; }
        c.y += step.y; if (c.y >= ny) { c.y -= ny; ++c.z; }
        c.z += step.z;
    };
    auto apply = [&](size_t vec, Coord c, const uint4& in) {
        if (c.x + V <= nx) {            // the whole vector lies in one row
            stStream(dv + vec * 16, unaryOpVec<BPV, OP>(x0 + c.x, y0 + c.y, z0 + c.z, in, fmt, mask, m));
            return;
        }
        uint32_t w[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
        for (int i = 0; i < V; ++i) {   // the vector wraps into the next row(s)
            vecSet<BPV>(w, i, unaryOp<OP>(x0 + c.x, y0 + c.y, z0 + c.z, vecGet<BPV>(in, i), fmt, mask, m));
            if (++c.x == nx) { c.x = 0; if (++c.y == ny) { c.y = 0; ++c.z; } }
        }
        stStream(dv + vec * 16, make_uint4(w[0], w[1], w[2], w[3]));
    };

    size_t i = tid;
    Coord c = coordOf(head + tid * V);
    for (; i + 3 * stride < nvec; i += 4 * stride) {
        const uint4 v0 = ldPlain(dv + i * 16), v1 = ldPlain(dv + (i + stride) * 16);
        const uint4 v2 = ldPlain(dv + (i + 2 * stride) * 16), v3 = ldPlain(dv + (i + 3 * stride) * 16);
        apply(i, c, v0); advance(c);
        apply(i + stride, c, v1); advance(c);
        apply(i + 2 * stride, c, v2); advance(c);
        apply(i + 3 * stride, c, v3); advance(c);
    }
    for (; i < nvec; i += stride) { apply(i, c, ldPlain(dv + i * 16)); advance(c); }
    const size_t nEdge = head + (n - tail0);
    for (size_t e = tid; e < nEdge; e += stride) {
        const size_t k = e < head ? e : tail0 + (e - head);
        const size_t row = k / (unsigned)nx;
        const int x = (int)(k - row * (unsigned)nx);
        const size_t zr = row / (unsigned)ny;
        const int y = (int)(row - zr * (unsigned)ny), z = (int)zr;
        storeRaw(d + k * BPV, BPV, unaryOp<OP>(x0 + x, y0 + y, z0 + z, loadRaw(d + k * BPV, BPV), fmt, mask, m));
    }
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

struct VecPair { uint4 a, b; };

// both volumes' rows in one 16-byte phase: two rows (four loads) in flight per thread
template <int BPV, int OP>
__global__ void __launch_bounds__(kThreads)
transform2UniformKernel(T2Params p, int shift)
{
    constexpr int V = 16 / BPV;
    constexpr uint32_t mask = BPV == 4 ? 0xffffffffu : (1u << (BPV * 8)) - 1u;
    forEachRowChunkUniform<V, 2, VecPair>(p.g, shift,
        [&](int y, int z, int xr) {
            VecPair q;
            q.a = ldPlain(p.v1.base + y * p.v1.pitchY + z * p.v1.pitchZ + (long long)xr * BPV);
            q.b = ldPlain(p.v2.base + y * p.v2.pitchY + z * p.v2.pitchZ + (long long)xr * BPV);
            return q;
        },
        [&](int y, int z, int xr, const VecPair& q) {
            uint32_t w1[4] = { 0u, 0u, 0u, 0u }, w2[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
            for (int i = 0; i < V; ++i) {
                uint32_t a = vecGet<BPV>(q.a, i), b = vecGet<BPV>(q.b, i);
                binaryOp<OP>(a, b, mask);
                vecSet<BPV>(w1, i, a); vecSet<BPV>(w2, i, b);
            }
            stStream(p.v1.base + y * p.v1.pitchY + z * p.v1.pitchZ + (long long)xr * BPV, make_uint4(w1[0], w1[1], w1[2], w1[3]));
            stStream(p.v2.base + y * p.v2.pitchY + z * p.v2.pitchZ + (long long)xr * BPV, make_uint4(w2[0], w2[1], w2[2], w2[3]));
        },
        [&](int y, int z, int xr, int n) {
            uint8_t* p1 = p.v1.base + y * p.v1.pitchY + z * p.v1.pitchZ + (long long)xr * BPV;
            uint8_t* p2 = p.v2.base + y * p.v2.pitchY + z * p.v2.pitchZ + (long long)xr * BPV;
            for (int i = 0; i < n; ++i) {
                uint32_t a = loadRaw(p1 + i * BPV, BPV), b = loadRaw(p2 + i * BPV, BPV);
                binaryOp<OP>(a, b, mask);
                storeRaw(p1 + i * BPV, BPV, a); storeRaw(p2 + i * BPV, BPV, b);
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
cudaError_t launchT1(const LaunchCtx& c, const T1Params& p, int op, bool flat)
{
    if (flat) {
        const size_t n = (size_t)p.g.nx * p.g.ny * p.g.nz;
        const unsigned grid = flatGrid(n * BPV / 16 + 1, kThreads * 4, c.smCount, 8);
#define VKTB_T1_FLAT(OP) transform1FlatKernel<BPV, OP><<<grid, kThreads, 0, c.stream>>>(p.v.base, n, p.g.nx, p.g.ny, p.g.x0, p.g.y0, p.g.z0, p.m, p.fmt)
        switch (op) {
        case 0: VKTB_T1_FLAT(0); break;
        case 1: VKTB_T1_FLAT(1); break;
        case 2: VKTB_T1_FLAT(2); break;
        case 3: VKTB_T1_FLAT(3); break;
        default: return cudaErrorInvalidValue;
        }
#undef VKTB_T1_FLAT
        countLaunch();
        return cudaGetLastError();
    }
    RowsLaunch l = rowsLaunch(p.g.nx, (long long)p.g.ny * p.g.nz, 16 / BPV);
    if (((p.v.pitchY | p.v.pitchZ) & 15) == 0 && (long long)p.g.ny * p.g.nz < 0x7fffffffLL) {
        const int shift = (int)((reinterpret_cast<uintptr_t>(p.v.base) & 15u) / BPV);
        switch (op) {
        case 0: transform1UniformKernel<BPV, 0><<<l.grid, l.block, 0, c.stream>>>(p, shift); break;
        case 1: transform1UniformKernel<BPV, 1><<<l.grid, l.block, 0, c.stream>>>(p, shift); break;
        case 2: transform1UniformKernel<BPV, 2><<<l.grid, l.block, 0, c.stream>>>(p, shift); break;
        case 3: transform1UniformKernel<BPV, 3><<<l.grid, l.block, 0, c.stream>>>(p, shift); break;
        default: return cudaErrorInvalidValue;
        }
        countLaunch();
        return cudaGetLastError();
    }
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
    const uintptr_t phase = reinterpret_cast<uintptr_t>(p.v1.base) & 15u;
    const bool uniform = ((p.v1.pitchY | p.v1.pitchZ | p.v2.pitchY | p.v2.pitchZ) & 15) == 0
                      && phase == (reinterpret_cast<uintptr_t>(p.v2.base) & 15u) && p.v1.base != p.v2.base
                      && (long long)p.g.ny * p.g.nz < 0x7fffffffLL;
    if (op != 0 && op != 1) return cudaErrorInvalidValue;
    if (uniform) {
        const int shift = (int)(phase / BPV);
        if (op == 0) transform2UniformKernel<BPV, 0><<<l.grid, l.block, 0, c.stream>>>(p, shift);
        else transform2UniformKernel<BPV, 1><<<l.grid, l.block, 0, c.stream>>>(p, shift);
    }
    else if (op == 0) transform2Kernel<BPV, 0><<<l.grid, l.block, 0, c.stream>>>(p);
    else transform2Kernel<BPV, 1><<<l.grid, l.block, 0, c.stream>>>(p);
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
    // whole rows of whole planes (or a single plane): one contiguous run, 32-bit indexable
    const bool flat = contiguous(vol, p.g) && p.g.nx == vol.dims[0] && (p.g.nz == 1 || p.g.ny == vol.dims[1])
                   && (long long)p.g.nx * p.g.ny * p.g.nz < (1LL << 40);
    switch (bytesPerVoxel(vol.format)) {
    case 1: return launchT1<1>(c, p, op, flat);
    case 2: return launchT1<2>(c, p, op, flat);
    case 4: return launchT1<4>(c, p, op, flat);
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
