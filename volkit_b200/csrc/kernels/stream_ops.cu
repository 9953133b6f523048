// Fill / FillRange, Copy / CopyRange (byte-wise and converting), MemsetRange and the codec probe.
//
// Fill     reference src/vkt/Fill_serial.hpp:20-26      vol[p] = map(value), p in [first,last)       bpv B/voxel
// Copy     reference src/vkt/Copy_serial.hpp:13-82      dst[p-first+off] = src[clamp(p)]             2 bpv (or bpv_s+bpv_d)
// Memset   reference src/vkt/Memory_cuda.cu:15-49       dst[i*pat .. ] = pattern                     dstBytes
#include <cstring>

#include "common.cuh"
#include "launch.hpp"
#include "voxel_codec.cuh"

namespace vktb {

std::atomic<uint64_t> g_launchCount{0};

Mapping makeMapping(float lo, float hi)
{
    Mapping m;
    m.lo = lo; m.hi = hi;
    volatile float r = hi - lo;          // binary32 subtraction, same as the device's __fsub_rn
    m.range = r;
    uint32_t bits; float rr = r;
    memcpy(&bits, &rr, 4);
    const uint32_t expo = (bits >> 23) & 0xffu, mant = bits & 0x7fffffu;
    // power of two, normal, with a normal reciprocal (2^-126 .. 2^126); sign may be either
    m.exactRecip = (mant == 0 && expo >= 1 && expo <= 253) ? 1 : 0;
    m.invRange = m.exactRecip ? 1.0f / rr : 0.f;
    uint32_t lb, hb; memcpy(&lb, &lo, 4); memcpy(&hb, &hi, 4);
    m.identity = (lb == 0u && hb == 0x3f800000u) ? 1 : 0;
    return m;
}

namespace {

// ---------------------------------------------------------------- Fill ----------------------------------
// 16-byte pattern for a bpv-sized raw value, valid at 16-byte aligned addresses
__device__ __forceinline__ uint4 patternOf(uint32_t raw, int bpv)
{
    uint32_t w = bpv == 1 ? (raw & 0xffu) * 0x01010101u : bpv == 2 ? (raw & 0xffffu) * 0x00010001u : raw;
    return make_uint4(w, w, w, w);
}

// The value is mapped on the device, once per thread, with the same codec the arithmetic kernels use
// (the reference maps per voxel: same value -> same bytes every time).
template <int F>
__global__ void __launch_bounds__(kThreads)
fillFlatKernel(uint8_t* __restrict__ d, size_t nbytes, float value, Mapping m)
{
    const uint4 pat = patternOf(mapVoxel<F, false>(value, m), bytesPerVoxel(F));
    // d is bpv-aligned; the pattern register is rotated so that it is correct at 16-byte boundaries
    size_t head = (16u - (reinterpret_cast<uintptr_t>(d) & 15u)) & 15u;
    if (head > nbytes) head = nbytes;
    const size_t nvec = (nbytes - head) / 16;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    uint8_t* dv = d + head;
    for (size_t i = tid; i < nvec; i += stride) stStream(dv + i * 16, pat);
    // edges byte-wise: byte k of the volume-relative stream equals pattern byte ((addr) & 15)
    const uint8_t* pb = reinterpret_cast<const uint8_t*>(&pat);
    const size_t tail0 = head + nvec * 16;
    const size_t nEdge = head + (nbytes - tail0);
    for (size_t e = tid; e < nEdge; e += stride) {
        size_t k = e < head ? e : tail0 + (e - head);
        d[k] = pb[reinterpret_cast<uintptr_t>(d + k) & 15u];
    }
}

template <int F>
__global__ void __launch_bounds__(kThreads)
fillRowsKernel(Operand d, Range g, float value, Mapping m)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    const uint32_t raw = mapVoxel<F, false>(value, m);
    const uint4 pat = patternOf(raw, BPV);
    forEachRowChunk<V, BPV>(g, d, [&](int y, int z, int xr, int n, bool full) {
        uint8_t* dp = d.base + y * d.pitchY + z * d.pitchZ + (long long)xr * BPV;
        if (full) stStream(dp, pat);
        else for (int i = 0; i < n; ++i) storeRaw(dp + i * BPV, BPV, raw);
    });
}

// ---------------------------------------------------------------- Copy (bytes) --------------------------
__global__ void __launch_bounds__(kThreads)
copyFlatKernel(uint8_t* __restrict__ d, const uint8_t* __restrict__ s, size_t nbytes)
{
    size_t head = (16u - (reinterpret_cast<uintptr_t>(d) & 15u)) & 15u;
    if (head > nbytes) head = nbytes;
    const size_t nvec = (nbytes - head) / 16;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    uint8_t* dv = d + head; const uint8_t* sv = s + head;
    size_t i = tid;
    for (; i + 3 * stride < nvec; i += 4 * stride) {
        uint4 v0 = ldStream(sv + i * 16), v1 = ldStream(sv + (i + stride) * 16);
        uint4 v2 = ldStream(sv + (i + 2 * stride) * 16), v3 = ldStream(sv + (i + 3 * stride) * 16);
        stStream(dv + i * 16, v0); stStream(dv + (i + stride) * 16, v1);
        stStream(dv + (i + 2 * stride) * 16, v2); stStream(dv + (i + 3 * stride) * 16, v3);
    }
    for (; i < nvec; i += stride) stStream(dv + i * 16, ldStream(sv + i * 16));
    const size_t tail0 = head + nvec * 16;
    const size_t nEdge = head + (nbytes - tail0);
    for (size_t e = tid; e < nEdge; e += stride) {
        size_t k = e < head ? e : tail0 + (e - head);
        d[k] = s[k];
    }
}

struct CopyParams
{
    Operand d;            // dst at dstOffset
    const uint8_t* s;     // src volume base
    int sx, sy, sz;       // src dims
    Range g;              // x0,y0,z0 = first (may be negative: clamped)
    Mapping md, ms;
    int fd, fs;
};

__device__ __forceinline__ int clampi(int x, int a, int b) { int m = b < x ? b : x; return a < m ? m : a; }

template <int BPV>
__global__ void __launch_bounds__(kThreads)
copyRowsKernel(CopyParams p)
{
    constexpr int V = 16 / BPV;
    forEachRowChunk<V, BPV>(p.g, p.d, [&](int y, int z, int xr, int n, bool full) {
        const int cy = clampi(p.g.y0 + y, 0, p.sy - 1), cz = clampi(p.g.z0 + z, 0, p.sz - 1);
        const uint8_t* srow = p.s + ((long long)cz * p.sy + cy) * (long long)p.sx * BPV;
        uint8_t* dp = p.d.base + y * p.d.pitchY + z * p.d.pitchZ + (long long)xr * BPV;
        const int ax = p.g.x0 + xr;   // absolute src x of the chunk's first voxel
        const uint8_t* sp = srow + (long long)ax * BPV;
        if (full && ax >= 0 && ax + V <= p.sx && (reinterpret_cast<uintptr_t>(sp) & 15u) == 0) {
            stStream(dp, ldStream(sp));
        } else {
            for (int i = 0; i < n; ++i)
                storeRaw(dp + i * BPV, BPV, loadRaw(srow + (long long)clampi(ax + i, 0, p.sx - 1) * BPV, BPV));
        }
    });
}

// bytewise copy of a box whose rows share one 16-byte phase in dst and src and need no clamping
template <int BPV>
__global__ void __launch_bounds__(kThreads)
copyRowsUniformKernel(CopyParams p, int shift)
{
    constexpr int V = 16 / BPV;
    const uint8_t* s0 = p.s + (((long long)p.g.z0 * p.sy + p.g.y0) * (long long)p.sx + p.g.x0) * BPV;
    const long long spy = (long long)p.sx * BPV, spz = spy * p.sy;
    forEachRowChunkUniform<V, 4, uint4>(p.g, shift,
        [&](int y, int z, int xr) { return ldStream(s0 + y * spy + z * spz + (long long)xr * BPV); },
        [&](int y, int z, int xr, const uint4& q) { stStream(p.d.base + y * p.d.pitchY + z * p.d.pitchZ + (long long)xr * BPV, q); },
        [&](int y, int z, int xr, int n) {
            const uint8_t* sp = s0 + y * spy + z * spz + (long long)xr * BPV;
            uint8_t* dp = p.d.base + y * p.d.pitchY + z * p.d.pitchZ + (long long)xr * BPV;
            for (int i = 0; i < n; ++i) storeRaw(dp + i * BPV, BPV, loadRaw(sp + i * BPV, BPV));
        });
}

// ---------------------------------------------------------------- Copy (by value) ------------------------
// generic: runtime formats, one voxel per thread along x, clamped src
__global__ void __launch_bounds__(kThreads)
convertRowsKernel(CopyParams p)
{
    const int bd = bytesPerVoxel(p.fd), bs = bytesPerVoxel(p.fs);
    const long long rows = (long long)p.g.ny * p.g.nz;
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const int z = (int)(r / p.g.ny), y = (int)(r - (long long)z * p.g.ny);
        const int cy = clampi(p.g.y0 + y, 0, p.sy - 1), cz = clampi(p.g.z0 + z, 0, p.sz - 1);
        const uint8_t* srow = p.s + ((long long)cz * p.sy + cy) * (long long)p.sx * bs;
        uint8_t* drow = p.d.base + y * p.d.pitchY + z * p.d.pitchZ;
        for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < p.g.nx; x += gridDim.x * blockDim.x) {
            const int cx = clampi(p.g.x0 + x, 0, p.sx - 1);
            float v = unmapVoxelRt(p.fs, loadRaw(srow + (long long)cx * bs, bs), p.ms);
            storeRaw(drow + (long long)x * bd, bd, mapVoxelRt(p.fd, v, p.md));
        }
    }
}

template <int BYTES> struct VecOf;
template <> struct VecOf<16> { typedef uint4 type; };
template <> struct VecOf<8>  { typedef uint2 type; };
template <> struct VecOf<4>  { typedef uint32_t type; };
template <> struct VecOf<2>  { typedef uint16_t type; };
template <> struct VecOf<1>  { typedef uint8_t type; };

// flat converting copy: V voxels per thread, V*BS bytes in, V*BD bytes out, both naturally aligned
template <int FD, int FS>
__global__ void __launch_bounds__(kThreads)
convertFlatKernel(uint8_t* __restrict__ d, const uint8_t* __restrict__ s, size_t n, Mapping md, Mapping ms)
{
    constexpr int BD = bytesPerVoxel(FD), BS = bytesPerVoxel(FS);
    constexpr int V = 16 / (BD > BS ? BD : BS);
    typedef typename VecOf<V * BS>::type SV;
    typedef typename VecOf<V * BD>::type DV;
    __shared__ float lut[FS == F_U8 ? 256 : 1];
    if (FS == F_U8) {
        for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = unmapVoxel<F_U8, false>((uint32_t)i, ms);
        __syncthreads();
    }
    const size_t nvec = n / V;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = tid; i < nvec; i += stride) {
        union { SV v; uint32_t w[4]; uint8_t b[16]; } in;
        union { DV v; uint32_t w[4]; uint8_t b[16]; } out;
        in.v = reinterpret_cast<const SV*>(s)[i];
#pragma unroll
        for (int k = 0; k < 4; ++k) out.w[k] = 0u;
#pragma unroll
        for (int k = 0; k < V; ++k) {
            uint32_t raw;
            if (BS == 4) raw = in.w[k];
            else if (BS == 2) raw = (in.w[k >> 1] >> ((k & 1) * 16)) & 0xffffu;
            else raw = (in.w[k >> 2] >> ((k & 3) * 8)) & 0xffu;
            float f = (FS == F_U8) ? lut[raw] : unmapVoxel<FS, false>(raw, ms);
            uint32_t e = mapVoxel<FD, false>(f, md);
            if (BD == 4) out.w[k] = e;
            else if (BD == 2) out.w[k >> 1] |= (e & 0xffffu) << ((k & 1) * 16);
            else out.w[k >> 2] |= (e & 0xffu) << ((k & 3) * 8);
        }
        reinterpret_cast<DV*>(d)[i] = out.v;
    }
    for (size_t k = nvec * V + tid; k < n; k += stride) {
        float f = unmapVoxel<FS, false>(loadRaw(s + k * BS, BS), ms);
        storeRaw(d + k * BD, BD, mapVoxel<FD, false>(f, md));
    }
}

// ---------------------------------------------------------------- MemsetRange ----------------------------
__global__ void __launch_bounds__(kThreads)
memsetPatternKernel(uint8_t* __restrict__ d, size_t nbytes, size_t patBytes, const uint8_t* __restrict__ pat)
{
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = tid; i < nbytes; i += stride) d[i] = pat[i % patBytes];
}

__global__ void debugCodecKernel(int f, Mapping m, const float* values, uint32_t* mapped, float* unmapped, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t raw = mapVoxelRt(f, values[i], m);
        mapped[i] = raw;
        unmapped[i] = unmapVoxelRt(f, raw, m);
    }
}

} // namespace

// ================================================================ launchers ==============================
cudaError_t launchFill(const LaunchCtx& c, const VolDesc& vol, const int32_t first[3], const int32_t last[3], float value)
{
    Range g;
    g.nx = last[0] - first[0]; g.ny = last[1] - first[1]; g.nz = last[2] - first[2];
    g.x0 = first[0]; g.y0 = first[1]; g.z0 = first[2];
    if (g.nx <= 0 || g.ny <= 0 || g.nz <= 0) return cudaSuccess;
    const int bpv = bytesPerVoxel(vol.format);
    Operand d = operandAt(vol, first[0], first[1], first[2]);
    const size_t n = (size_t)g.nx * g.ny * g.nz;

#define VKTB_FILL(F)                                                                                          \
    case F:                                                                                                   \
        if (contiguous(vol, g)) {                                                                             \
            unsigned grid = flatGrid(n * bpv / 16 + 1, kThreads * 4, c.smCount, 8);                           \
            fillFlatKernel<F><<<grid, kThreads, 0, c.stream>>>(d.base, n * bpv, value, vol.map);              \
        } else {                                                                                              \
            RowsLaunch l = rowsLaunch(g.nx, (long long)g.ny * g.nz, 16 / bytesPerVoxel(F));                   \
            fillRowsKernel<F><<<l.grid, l.block, 0, c.stream>>>(d, g, value, vol.map);                        \
        }                                                                                                     \
        break;

    switch (vol.format) {
        VKTB_FILL(F_U8) VKTB_FILL(F_U16) VKTB_FILL(F_INT16) VKTB_FILL(F_U32) VKTB_FILL(F_F32)
    default: return cudaErrorInvalidValue;
    }
#undef VKTB_FILL
    countLaunch();
    return cudaGetLastError();
}

namespace {

template <int FD>
cudaError_t convertFlatDispatch(const LaunchCtx& c, int fs, uint8_t* d, const uint8_t* s, size_t n, const Mapping& md,
                                const Mapping& ms)
{
    unsigned grid = flatGrid(n / 4 + 1, kThreads * 2, c.smCount, 8);
    switch (fs) {
    case F_U8:  convertFlatKernel<FD, F_U8><<<grid, kThreads, 0, c.stream>>>(d, s, n, md, ms); break;
    case F_U16: convertFlatKernel<FD, F_U16><<<grid, kThreads, 0, c.stream>>>(d, s, n, md, ms); break;
    case F_U32: convertFlatKernel<FD, F_U32><<<grid, kThreads, 0, c.stream>>>(d, s, n, md, ms); break;
    case F_F32: convertFlatKernel<FD, F_F32><<<grid, kThreads, 0, c.stream>>>(d, s, n, md, ms); break;
    default: return cudaErrorInvalidValue;
    }
    countLaunch();
    return cudaGetLastError();
}

} // namespace

// Converting flat copy, also used by Resample's equal-dims branch.  Returns cudaErrorNotSupported when the
// operands do not qualify (caller falls back to the generic rows kernel).
cudaError_t launchConvertFlat(const LaunchCtx& c, const VolDesc& dst, uint8_t* d, const VolDesc& src, const uint8_t* s, size_t n)
{
    const int bd = bytesPerVoxel(dst.format), bs = bytesPerVoxel(src.format);
    const int v = 16 / (bd > bs ? bd : bs);
    if ((reinterpret_cast<uintptr_t>(d) % (size_t)(v * bd)) || (reinterpret_cast<uintptr_t>(s) % (size_t)(v * bs)))
        return cudaErrorNotSupported;
    switch (dst.format) {
    case F_U8:  return convertFlatDispatch<F_U8>(c, src.format, d, s, n, dst.map, src.map);
    case F_U16: return convertFlatDispatch<F_U16>(c, src.format, d, s, n, dst.map, src.map);
    case F_U32: return convertFlatDispatch<F_U32>(c, src.format, d, s, n, dst.map, src.map);
    case F_F32: return convertFlatDispatch<F_F32>(c, src.format, d, s, n, dst.map, src.map);
    }
    return cudaErrorNotSupported;
}

cudaError_t launchCopy(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src, const int32_t first[3],
                       const int32_t last[3], const int32_t off[3])
{
    CopyParams p;
    Range& g = p.g;
    g.nx = last[0] - first[0]; g.ny = last[1] - first[1]; g.nz = last[2] - first[2];
    g.x0 = first[0]; g.y0 = first[1]; g.z0 = first[2];
    if (g.nx <= 0 || g.ny <= 0 || g.nz <= 0) return cudaSuccess;
    if (src.dims[0] <= 0 || src.dims[1] <= 0 || src.dims[2] <= 0) return cudaErrorInvalidValue;

    p.d = operandAt(dst, off[0], off[1], off[2]);
    p.s = src.data; p.sx = src.dims[0]; p.sy = src.dims[1]; p.sz = src.dims[2];
    p.md = dst.map; p.ms = src.map; p.fd = dst.format; p.fs = src.format;
    const int bd = bytesPerVoxel(dst.format), bs = bytesPerVoxel(src.format);
    // reference Copy_serial.hpp:21-22: bytes iff format AND mapping (float ==) agree
    const bool bytewise = dst.format == src.format && dst.map.lo == src.map.lo && dst.map.hi == src.map.hi;
    const bool inside = first[0] >= 0 && first[1] >= 0 && first[2] >= 0
                     && last[0] <= src.dims[0] && last[1] <= src.dims[1] && last[2] <= src.dims[2];
    const size_t n = (size_t)g.nx * g.ny * g.nz;
    const bool flat = inside && contiguous(dst, g) && contiguous(src, g);
    const uint8_t* s0 = src.data + (inside ? voxelIndex(src, first[0], first[1], first[2]) * bs : 0);

    if (bytewise) {
        if (flat && ((reinterpret_cast<uintptr_t>(p.d.base) ^ reinterpret_cast<uintptr_t>(s0)) & 15u) == 0) {
            unsigned grid = flatGrid(n * bd / 16 + 1, kThreads * 4, c.smCount, 8);
            copyFlatKernel<<<grid, kThreads, 0, c.stream>>>(p.d.base, s0, n * bd);
        } else {
            RowsLaunch l = rowsLaunch(g.nx, (long long)g.ny * g.nz, 16 / bd);
            // all rows in one 16-byte phase, src == dst phase, nothing to clamp: several rows in flight per thread
            const long long spy = (long long)src.dims[0] * bs, spz = spy * src.dims[1];
            const bool uniform = inside && (p.d.pitchY % 16) == 0 && (p.d.pitchZ % 16) == 0 && (spy % 16) == 0 && (spz % 16) == 0
                              && ((reinterpret_cast<uintptr_t>(p.d.base) ^ reinterpret_cast<uintptr_t>(s0)) & 15u) == 0
                              && (long long)g.ny * g.nz < 0x7fffffffLL;
            if (uniform) {
                const int shift = (int)((reinterpret_cast<uintptr_t>(p.d.base) & 15u) / bd);
                if (bd == 1) copyRowsUniformKernel<1><<<l.grid, l.block, 0, c.stream>>>(p, shift);
                else if (bd == 2) copyRowsUniformKernel<2><<<l.grid, l.block, 0, c.stream>>>(p, shift);
                else copyRowsUniformKernel<4><<<l.grid, l.block, 0, c.stream>>>(p, shift);
            }
            else if (bd == 1) copyRowsKernel<1><<<l.grid, l.block, 0, c.stream>>>(p);
            else if (bd == 2) copyRowsKernel<2><<<l.grid, l.block, 0, c.stream>>>(p);
            else copyRowsKernel<4><<<l.grid, l.block, 0, c.stream>>>(p);
        }
        countLaunch();
        return cudaGetLastError();
    }
    if (flat) {
        cudaError_t e = launchConvertFlat(c, dst, p.d.base, src, s0, n);
        if (e != cudaErrorNotSupported) return e;
    }
    return launchConvertRows(c, dst, src, first, last, off);
}

// generic converting copy (runtime formats, clamped src), always by value
cudaError_t launchConvertRows(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src, const int32_t first[3],
                              const int32_t last[3], const int32_t off[3])
{
    CopyParams p;
    Range& g = p.g;
    g.nx = last[0] - first[0]; g.ny = last[1] - first[1]; g.nz = last[2] - first[2];
    g.x0 = first[0]; g.y0 = first[1]; g.z0 = first[2];
    if (g.nx <= 0 || g.ny <= 0 || g.nz <= 0) return cudaSuccess;
    p.d = operandAt(dst, off[0], off[1], off[2]);
    p.s = src.data; p.sx = src.dims[0]; p.sy = src.dims[1]; p.sz = src.dims[2];
    p.md = dst.map; p.ms = src.map; p.fd = dst.format; p.fs = src.format;
    long long rows = (long long)g.ny * g.nz;
    dim3 grid((unsigned)((g.nx + kThreads - 1) / kThreads), (unsigned)(rows > 65535 ? 65535 : rows), 1);
    if (grid.x > 1024) grid.x = 1024;
    convertRowsKernel<<<grid, kThreads, 0, c.stream>>>(p);
    countLaunch();
    return cudaGetLastError();
}

cudaError_t launchMemsetRange(const LaunchCtx& c, void* dst, const void* patternDev, size_t dstBytes, size_t patBytes)
{
    if (dstBytes == 0 || patBytes == 0) return cudaSuccess;
    unsigned grid = flatGrid(dstBytes, kThreads * 16, c.smCount, 8);
    memsetPatternKernel<<<grid, kThreads, 0, c.stream>>>((uint8_t*)dst, dstBytes, patBytes, (const uint8_t*)patternDev);
    countLaunch();
    return cudaGetLastError();
}

cudaError_t launchDebugCodec(const LaunchCtx& c, int format, Mapping m, const float* values, uint32_t* mapped,
                             float* unmapped, size_t n)
{
    debugCodecKernel<<<64, kThreads, 0, c.stream>>>(format, m, values, mapped, unmapped, n);
    countLaunch();
    return cudaGetLastError();
}

} // namespace vktb
