// Fill / FillRange, Copy / CopyRange (byte-wise and converting), MemsetRange and the codec probe.
//
// Fill     reference src/vkt/Fill_serial.hpp:20-26      vol[p] = map(value), p in [first,last)       bpv B/voxel
// Copy     reference src/vkt/Copy_serial.hpp:13-82      dst[p-first+off] = src[clamp(p)]             2 bpv (or bpv_s+bpv_d)
// Memset   reference src/vkt/Memory_cuda.cu:15-49       dst[i*pat .. ] = pattern                     dstBytes
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "copy_boxes.hpp"
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
    // RN(1 / range) (an IEEE division on the host); used as is for powers of two, and by divProven once the divisor has
    // been proven (proveDivisor, arith.cu)
    m.invRange = (expo >= 1 && expo <= 253) ? 1.0f / rr : 0.f;
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
    const uint8_t* sEnd;  // one past its last byte
    int sx, sy, sz;       // src dims
    int shifted;          // uniform kernel: src rows sit at another 16-byte phase than dst rows
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
        const bool whole = full && ax >= 0 && ax + V <= p.sx;   // nothing to clamp in this chunk
        if (whole && (reinterpret_cast<uintptr_t>(sp) & 15u) == 0) {
            stStream(dp, ldStream(sp));
        } else if (whole && unalignedReadable(sp, p.sEnd)) {
            stStream(dp, ldUnaligned(sp));
        } else {
            for (int i = 0; i < n; ++i)
                storeRaw(dp + i * BPV, BPV, loadRaw(srow + (long long)clampi(ax + i, 0, p.sx - 1) * BPV, BPV));
        }
    });
}

// bytewise copy of a box that needs no clamping and whose rows share one 16-byte phase per operand (p.shifted: the
// src phase differs from the dst phase -> funnel-shifted loads)
template <int BPV>
__global__ void __launch_bounds__(kThreads)
copyRowsUniformKernel(CopyParams p, int shift)
{
    constexpr int V = 16 / BPV;
    const uint8_t* s0 = p.s + (((long long)p.g.z0 * p.sy + p.g.y0) * (long long)p.sx + p.g.x0) * BPV;
    const long long spy = (long long)p.sx * BPV, spz = spy * p.sy;
    auto edge = [&](int y, int z, int xr, int n) {
        const uint8_t* sp = s0 + y * spy + z * spz + (long long)xr * BPV;
        uint8_t* dp = p.d.base + y * p.d.pitchY + z * p.d.pitchZ + (long long)xr * BPV;
        for (int i = 0; i < n; ++i) storeRaw(dp + i * BPV, BPV, loadRaw(sp + i * BPV, BPV));
    };
    if (!p.shifted) {
        forEachRowChunkUniform<V, 4, uint4>(p.g, shift,
            [&](int y, int z, int xr) { return ldStream(s0 + y * spy + z * spz + (long long)xr * BPV); },
            [&](int y, int z, int xr, const uint4& q) { stStream(p.d.base + y * p.d.pitchY + z * p.d.pitchZ + (long long)xr * BPV, q); },
            edge);
        return;
    }
    // shifted: the load step only issues the two aligned vectors around the chunk (through L1: a lane's second vector
    // is the next lane's first); the funnel shift waits for the store step, so the loads of all four rows in flight go
    // out back to back.  The one chunk whose second vector would leave the source buffer is copied per voxel.
    struct Pair { uint4 lo, hi; };
    const unsigned sh = (unsigned)((reinterpret_cast<uintptr_t>(s0) + (uintptr_t)(16 - shift * BPV)) & 15u);   // != 0
    forEachRowChunkUniform<V, 4, Pair>(p.g, shift,
        [&](int y, int z, int xr) {
            const uint8_t* sp = s0 + y * spy + z * spz + (long long)xr * BPV;
            Pair q;
            q.lo = q.hi = make_uint4(0u, 0u, 0u, 0u);
            if (unalignedReadable(sp, p.sEnd)) {
                const uint4* v = reinterpret_cast<const uint4*>(sp - sh);
                q.lo = __ldg(v); q.hi = __ldg(v + 1);
            }
            return q;
        },
        [&](int y, int z, int xr, const Pair& q) {
            const uint8_t* sp = s0 + y * spy + z * spz + (long long)xr * BPV;
            if (unalignedReadable(sp, p.sEnd))
                stStream(p.d.base + y * p.d.pitchY + z * p.d.pitchZ + (long long)xr * BPV, window16(q.lo, q.hi, sh));
            else edge(y, z, xr, V);
        },
        edge);
}

// ---------------------------------------------------------------- Copy, many boxes of one source -----------
// BrickDecompose (reference src/vkt/Decompose_serial.hpp:15-48) is a loop of CopyRange(dest[index], source, first,
// last) calls; at 64^3 bricks a 2048^3 volume makes 32768 of them, far too many launches.  One launch instead:
// blockIdx.y picks the box, blockIdx.x a group of its z planes; rows are walked two at a time (two 128-bit loads in
// flight per thread), chunks cut at 16-byte boundaries of the dst row, the src side clamped and funnel-shifted.
__device__ __forceinline__ unsigned magicDiv(unsigned e, unsigned mul, int sh)
{
    return (unsigned)(((unsigned long long)e * mul) >> sh);
}

// Boxes that fill their destination volume completely (every brick of a decomposition does) are ONE contiguous run
// of dst bytes: cut it into aligned 128-bit vectors regardless of the rows (a 66-voxel UInt16 row is 132 bytes: cut
// per row, two of its ten chunks would be partial).  A vector lies in one src row or straddles two; each piece is
// one (funnel-shifted) 128-bit load and the two are merged bytewise, so every lane of a warp runs the same code
// (ncu on the first version, which gathered the straddling vectors voxel by voxel: 12 of 32 lanes active on
// average, 25 % of the HBM roofline).  Only vectors that touch the clamped x border of the volume, or the first /
// last partial vector of a box, are gathered per voxel.  Four vectors in flight per thread.
template <int BPV>
__global__ void __launch_bounds__(kThreads)
copyBoxesFlatKernel(const BoxRec* __restrict__ boxes, const uint8_t* __restrict__ s, const uint8_t* sEnd,
                    int sx, int sy, int sz)
{
    constexpr int V = 16 / BPV;
    constexpr int U = 4;
    // blockIdx.x = box, blockIdx.y = which 1024 vectors of it: CTAs that run together work on the SAME rows of
    // consecutive boxes, i.e. on x-adjacent bricks whose src pieces are neighbours in one src row -- with the box as
    // the slow index each 128-byte piece of a src row was fetched at a different time (54 % of the roofline for
    // aligned 64^3 bricks)
    const BoxRec b = boxes[blockIdx.x];
    const int shift = (int)((reinterpret_cast<uintptr_t>(b.d) & 15u) / BPV);
    const unsigned nvec = (b.nvox + shift + V - 1) / V;
    const unsigned k0 = blockIdx.y * (kThreads * U) + threadIdx.x;
    if (blockIdx.y * (kThreads * U) >= nvec) return;
    const long long rowBytes = (long long)sx * BPV;
    auto srcRow = [&](int y, int z) {
        return s + ((long long)clampi(b.z0 + z, 0, sz - 1) * sy + clampi(b.y0 + y, 0, sy - 1)) * rowBytes;
    };
    // 16 bytes at any phase, or false when the funnel-shifted load would leave the src buffer
    auto load16 = [&](const uint8_t* sp, uint4& q) {
        if (sp < s) return false;
        if ((reinterpret_cast<uintptr_t>(sp) & 15u) == 0) { q = ldStream(sp); return true; }
        if (!unalignedReadable(sp, sEnd)) return false;
        q = ldUnaligned(sp);
        return true;
    };

    struct Item { uint8_t* dp; int x, y, z, n, mode; uint4 q; };
    Item it[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        Item& t = it[u];
        const unsigned k = k0 + u * kThreads;
        t.mode = 0;
        if (k >= nvec) continue;
        long long e = (long long)k * V - shift;
        int n = V;
        if (e < 0) { n += (int)e; e = 0; }
        if (e + n > (long long)b.nvox) n = (int)(b.nvox - e);
        const unsigned row = magicDiv((unsigned)e, b.mulX, b.shX);
        t.x = (int)((unsigned)e - row * (unsigned)b.nx);
        t.z = (int)magicDiv(row, b.mulY, b.shY);
        t.y = (int)(row - (unsigned)t.z * (unsigned)b.ny);
        t.n = n;
        t.dp = b.d + e * BPV;
        t.mode = 2;
        if (n != V || b.nx < V) continue;
        const int nA = min(V, b.nx - t.x);           // voxels of the vector that lie in its first row
        const int ax = b.x0 + t.x;
        if (ax < 0 || ax + nA > sx) continue;        // clamped border: gather
        uint4 qa;
        if (!load16(srcRow(t.y, t.z) + (long long)ax * BPV, qa)) continue;
        if (nA < V) {
            if (b.x0 < 0 || b.x0 + (V - nA) > sx) continue;
            int y1 = t.y + 1, z1 = t.z;
            if (y1 == b.ny) { y1 = 0; ++z1; }
            // the second row's voxels land at byte nA*BPV of the vector: load from nA voxels before its start
            uint4 qb;
            if (!load16(srcRow(y1, z1) + ((long long)b.x0 - nA) * BPV, qb)) continue;
            const int split = nA * BPV;              // bytes [0, split) from qa, the rest from qb
            const uint32_t a[4] = { qa.x, qa.y, qa.z, qa.w }, c[4] = { qb.x, qb.y, qb.z, qb.w };
            uint32_t w[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int kb = min(max(split - 4 * i, 0), 4);
                const uint32_t m = kb == 4 ? 0xffffffffu : ((1u << (8 * kb)) - 1u);
                w[i] = (a[i] & m) | (c[i] & ~m);
            }
            qa = make_uint4(w[0], w[1], w[2], w[3]);
        }
        t.q = qa;
        t.mode = 1;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const Item& t = it[u];
        if (t.mode == 1) stStream(t.dp, t.q);
        else if (t.mode == 2) {
            uint32_t w[4] = { 0u, 0u, 0u, 0u };
            int x = t.x, y = t.y, z = t.z;
            const uint8_t* srow = srcRow(y, z);
            for (int i = 0; i < t.n; ++i) {
                const uint32_t raw = loadRaw(srow + (long long)clampi(b.x0 + x, 0, sx - 1) * BPV, BPV);
                if (t.n == V) w[(i * BPV) >> 2] |= raw << (((i * BPV) & 3) * 8);
                else storeRaw(t.dp + i * BPV, BPV, raw);
                if (++x == b.nx) {
                    x = 0;
                    if (++y == b.ny) { y = 0; ++z; }
                    srow = srcRow(y, z);
                }
            }
            if (t.n == V) stStream(t.dp, make_uint4(w[0], w[1], w[2], w[3]));
        }
    }
}

template <int BPV>
__global__ void __launch_bounds__(kThreads)
copyBoxesKernel(const BoxRec* __restrict__ boxes, const uint8_t* __restrict__ s, const uint8_t* sEnd,
                int sx, int sy, int sz, int zPer)
{
    constexpr int V = 16 / BPV;
    const BoxRec b = boxes[blockIdx.y];
    const int zBeg = blockIdx.x * zPer;
    if (zBeg >= b.nz) return;
    const int rows = (min(zBeg + zPer, b.nz) - zBeg) * b.ny;
    const int chunks = (b.nx + V - 1) / V + 1;

    struct Item { uint8_t* dp; const uint8_t* srow; int ax, n, mode; uint4 q; };
    auto prepare = [&](int r, int chunk, Item& it) {
        it.mode = 0;
        if (r >= rows) return;
        const int zz = r / b.ny, y = r - zz * b.ny, z = zBeg + zz;
        uint8_t* drow = b.d + y * b.pitchY + z * b.pitchZ;
        const int shift = (int)((reinterpret_cast<uintptr_t>(drow) & 15u) / BPV);
        int xr = chunk * V - shift, n = V;
        if (xr < 0) { n += xr; xr = 0; }
        if (xr + n > b.nx) n = b.nx - xr;
        if (n <= 0) return;
        const int cy = clampi(b.y0 + y, 0, sy - 1), cz = clampi(b.z0 + z, 0, sz - 1);
        it.srow = s + ((long long)cz * sy + cy) * (long long)sx * BPV;
        it.dp = drow + (long long)xr * BPV;
        it.ax = b.x0 + xr;
        it.n = n;
        const uint8_t* sp = it.srow + (long long)it.ax * BPV;
        const bool whole = n == V && it.ax >= 0 && it.ax + V <= sx;
        if (whole && (reinterpret_cast<uintptr_t>(sp) & 15u) == 0) { it.mode = 1; it.q = ldStream(sp); }
        else if (whole && unalignedReadable(sp, sEnd))             { it.mode = 1; it.q = ldUnaligned(sp); }
        else it.mode = 2;
    };
    auto finish = [&](const Item& it) {
        if (it.mode == 1) stStream(it.dp, it.q);
        else if (it.mode == 2)
            for (int i = 0; i < it.n; ++i)
                storeRaw(it.dp + i * BPV, BPV, loadRaw(it.srow + (long long)clampi(it.ax + i, 0, sx - 1) * BPV, BPV));
    };
    for (int chunk = threadIdx.x; chunk < chunks; chunk += blockDim.x)
        for (int r = threadIdx.y; r < rows; r += 2 * blockDim.y) {
            Item i0, i1;
            prepare(r, chunk, i0);
            prepare(r + blockDim.y, chunk, i1);
            finish(i0);
            finish(i1);
        }
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
    p.sEnd = src.data + (size_t)src.dims[0] * src.dims[1] * src.dims[2] * bs;
    p.shifted = 0;
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
                              && (long long)g.ny * g.nz < 0x7fffffffLL;
            if (uniform) {
                p.shifted = ((reinterpret_cast<uintptr_t>(p.d.base) ^ reinterpret_cast<uintptr_t>(s0)) & 15u) != 0;
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

// n boxes of one source in one launch (bytewise case); converting boxes go through launchCopy one by one.
cudaError_t launchCopyBatch(const LaunchCtx& c, int n, const VolDesc* dsts, const VolDesc& src, const int32_t* firsts,
                            const int32_t* lasts, const int32_t* offs)
{
    if (n <= 0) return cudaSuccess;
    if (src.dims[0] <= 0 || src.dims[1] <= 0 || src.dims[2] <= 0) return cudaErrorInvalidValue;
    const int bs = bytesPerVoxel(src.format);
    std::vector<BoxRec> recs, flats;   // row-wise boxes / boxes that are one contiguous run of dst bytes
    recs.reserve((size_t)n);
    flats.reserve((size_t)n);
    int maxNx = 0, maxNz = 0;
    long long maxPlane = 0;
    unsigned maxVox = 0;
    for (int i = 0; i < n; ++i) {
        const VolDesc& dst = dsts[i];
        const int32_t* f = firsts + 3 * i; const int32_t* l = lasts + 3 * i; const int32_t* o = offs + 3 * i;
        const bool bytewise = dst.format == src.format && dst.map.lo == src.map.lo && dst.map.hi == src.map.hi;
        if (!bytewise) {
            cudaError_t e = launchCopy(c, dst, src, f, l, o);
            if (e != cudaSuccess) return e;
            continue;
        }
        BoxRec r;
        r.nx = l[0] - f[0]; r.ny = l[1] - f[1]; r.nz = l[2] - f[2];
        if (r.nx <= 0 || r.ny <= 0 || r.nz <= 0) continue;
        const Operand d = operandAt(dst, o[0], o[1], o[2]);
        r.d = d.base; r.pitchY = d.pitchY; r.pitchZ = d.pitchZ;
        r.x0 = f[0]; r.y0 = f[1]; r.z0 = f[2];
        r.nvox = 0; r.mulX = r.mulY = 0; r.shX = r.shY = 0;
        const long long vox = (long long)r.nx * r.ny * r.nz;
        Range g; g.nx = r.nx; g.ny = r.ny; g.nz = r.nz; g.x0 = g.y0 = g.z0 = 0;
        if (contiguous(dst, g) && vox < (long long)65535 * kThreads * 4 * (16 / bs)) {   // grid.y of the flat kernel
            r.nvox = (unsigned)vox;
            magicDivisor((unsigned)r.nx, r.mulX, r.shX);
            magicDivisor((unsigned)r.ny, r.mulY, r.shY);
            flats.push_back(r);
            maxVox = r.nvox > maxVox ? r.nvox : maxVox;
            continue;
        }
        recs.push_back(r);
        maxNx = r.nx > maxNx ? r.nx : maxNx;
        maxNz = r.nz > maxNz ? r.nz : maxNz;
        const long long plane = (long long)r.nx * r.ny * bs;
        maxPlane = plane > maxPlane ? plane : maxPlane;
    }
    if (recs.empty() && flats.empty()) return cudaSuccess;

    const size_t nFlat = flats.size();
    flats.insert(flats.end(), recs.begin(), recs.end());   // one table: flat boxes first
    BoxRec* dev = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&dev, flats.size() * sizeof(BoxRec), c.stream);
    if (e != cudaSuccess) return e;
    // pageable source: the runtime stages the table before returning, the vectors may die with this frame
    e = cudaMemcpyAsync(dev, flats.data(), flats.size() * sizeof(BoxRec), cudaMemcpyHostToDevice, c.stream);
    if (e != cudaSuccess) { cudaFreeAsync(dev, c.stream); return e; }
    const uint8_t* sEnd = src.data + (size_t)src.dims[0] * src.dims[1] * src.dims[2] * bs;

    bool staged = false;
    if (nFlat) {
        // bricks staged through shared memory by the TMA unit (copy_bricks.cu) where the operands qualify
        e = launchCopyBricksTma(c, dev, flats.data(), nFlat, src.data, bs, src.dims);
        staged = e == cudaSuccess;
        if (e == cudaErrorNotSupported) e = cudaSuccess;
    }
    if (nFlat && !staged && e == cudaSuccess) {
        const unsigned perCta = kThreads * 4 * (16 / bs);                 // voxels per CTA
        const unsigned gx = (maxVox + (16 / bs) + perCta - 1) / perCta;   // + one vector for a misaligned head
        const size_t perLaunch = 0x7fffffff;
        for (size_t b0 = 0; b0 < nFlat && e == cudaSuccess; b0 += perLaunch) {
            dim3 grid((unsigned)(nFlat - b0 < perLaunch ? nFlat - b0 : perLaunch), gx, 1);
            if (bs == 1) copyBoxesFlatKernel<1><<<grid, kThreads, 0, c.stream>>>(dev + b0, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2]);
            else if (bs == 2) copyBoxesFlatKernel<2><<<grid, kThreads, 0, c.stream>>>(dev + b0, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2]);
            else copyBoxesFlatKernel<4><<<grid, kThreads, 0, c.stream>>>(dev + b0, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2]);
            countLaunch();
            e = cudaGetLastError();
        }
    }
    if (recs.empty() || e != cudaSuccess) { cudaFreeAsync(dev, c.stream); return e; }
    BoxRec* devRows = dev + nFlat;

    // ~64 KiB of dst per CTA, but never fewer CTAs than two waves of the machine
    int zPer = (int)((65536 + maxPlane - 1) / maxPlane);
    if (zPer < 1) zPer = 1;
    while (zPer > 1 && (long long)recs.size() * ((maxNz + zPer - 1) / zPer) < 2LL * c.smCount * 8) zPer = (zPer + 1) / 2;
    const RowsLaunch l = rowsLaunch(maxNx, 1, 16 / bs);
    const unsigned gx = (unsigned)((maxNz + zPer - 1) / zPer);
    for (size_t b0 = 0; b0 < recs.size(); b0 += 65535) {
        const unsigned gy = (unsigned)(recs.size() - b0 < 65535 ? recs.size() - b0 : 65535);
        dim3 grid(gx, gy, 1);
        if (bs == 1) copyBoxesKernel<1><<<grid, l.block, 0, c.stream>>>(devRows + b0, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2], zPer);
        else if (bs == 2) copyBoxesKernel<2><<<grid, l.block, 0, c.stream>>>(devRows + b0, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2], zPer);
        else copyBoxesKernel<4><<<grid, l.block, 0, c.stream>>>(devRows + b0, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2], zPer);
        countLaunch();
        e = cudaGetLastError();
        if (e != cudaSuccess) break;
    }
    cudaFreeAsync(dev, c.stream);
    return e;
}

// ---- the same batch with NO allocation anywhere (vktb_copy_range_batch_noalloc) ---------------------------------------
// For callers that may not call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The
// box table -- the one piece of scratch memory launchCopyBatch allocates per call -- lives in a file-scope device
// array instead: kNoAllocMaxBoxes records, resident on every GPU that loads the library until the process exits.
// In host code `g_boxScratch` is only a stand-in for the symbol; the device address comes from cudaGetSymbolAddress,
// asked once per device with that device current.  The host-side table is a static array under a mutex, so the call
// makes no host heap allocation either.  Every call shares the device array: calls whose kernels overlap on the GPU
// (two streams, two host threads) race on it; nothing in it has to start at zero -- the n records a call uses are
// overwritten by that call's table upload, which is stream-ordered before its kernels and after the previous call's.
constexpr int kNoAllocMaxBoxes = 4096;
static __device__ __align__(256) unsigned char g_boxScratch[(size_t)kNoAllocMaxBoxes * sizeof(BoxRec)];

namespace {
std::mutex g_noAllocMutex;
BoxRec*    g_boxScratchAddr[64] = {};          // per device ordinal, from cudaGetSymbolAddress
BoxRec     g_hostBoxes[kNoAllocMaxBoxes];      // flat boxes from the front, row-wise boxes from the back
}

int copyBatchNoAllocMaxBoxes() { return kNoAllocMaxBoxes; }

cudaError_t launchCopyBatchNoAlloc(const LaunchCtx& c, int n, const VolDesc* dsts, const VolDesc& src, const int32_t* firsts,
                                   const int32_t* lasts, const int32_t* offs)
{
    if (n <= 0) return cudaSuccess;
    if (n > kNoAllocMaxBoxes) return cudaErrorInvalidValue;
    if (src.dims[0] <= 0 || src.dims[1] <= 0 || src.dims[2] <= 0) return cudaErrorInvalidValue;
    const int bs = bytesPerVoxel(src.format);
    std::lock_guard<std::mutex> lock(g_noAllocMutex);

    // set-up, once per device, with that device current (the C-ABI entry made it so)
    int ord = -1;
    cudaError_t e = cudaGetDevice(&ord);
    if (e != cudaSuccess) return e;
    if (ord < 0 || ord >= 64) return cudaErrorInvalidDevice;
    if (!g_boxScratchAddr[ord]) {
        void* addr = nullptr;
        e = cudaGetSymbolAddress(&addr, g_boxScratch);
        if (e != cudaSuccess) return e;
        g_boxScratchAddr[ord] = static_cast<BoxRec*>(addr);
    }
    BoxRec* dev = g_boxScratchAddr[ord];

    int nFlat = 0, nRows = 0, maxNx = 0, maxNz = 0;
    long long maxPlane = 0;
    unsigned maxVox = 0;
    for (int i = 0; i < n; ++i) {
        const VolDesc& dst = dsts[i];
        const int32_t* f = firsts + 3 * i; const int32_t* l = lasts + 3 * i; const int32_t* o = offs + 3 * i;
        // converting boxes would need one launchCopy each; this entry point is the bytewise batch only
        if (!(dst.format == src.format && dst.map.lo == src.map.lo && dst.map.hi == src.map.hi)) return cudaErrorNotSupported;
        BoxRec r;
        r.nx = l[0] - f[0]; r.ny = l[1] - f[1]; r.nz = l[2] - f[2];
        if (r.nx <= 0 || r.ny <= 0 || r.nz <= 0) continue;
        const Operand d = operandAt(dst, o[0], o[1], o[2]);
        r.d = d.base; r.pitchY = d.pitchY; r.pitchZ = d.pitchZ;
        r.x0 = f[0]; r.y0 = f[1]; r.z0 = f[2];
        r.nvox = 0; r.mulX = r.mulY = 0; r.shX = r.shY = 0;
        const long long vox = (long long)r.nx * r.ny * r.nz;
        Range g; g.nx = r.nx; g.ny = r.ny; g.nz = r.nz; g.x0 = g.y0 = g.z0 = 0;
        if (contiguous(dst, g) && vox < (long long)65535 * kThreads * 4 * (16 / bs)) {
            r.nvox = (unsigned)vox;
            magicDivisor((unsigned)r.nx, r.mulX, r.shX);
            magicDivisor((unsigned)r.ny, r.mulY, r.shY);
            g_hostBoxes[nFlat++] = r;
            maxVox = r.nvox > maxVox ? r.nvox : maxVox;
            continue;
        }
        g_hostBoxes[kNoAllocMaxBoxes - 1 - nRows++] = r;
        maxNx = r.nx > maxNx ? r.nx : maxNx;
        maxNz = r.nz > maxNz ? r.nz : maxNz;
        const long long plane = (long long)r.nx * r.ny * bs;
        maxPlane = plane > maxPlane ? plane : maxPlane;
    }
    if (nFlat == 0 && nRows == 0) return cudaSuccess;
    // pageable source: cudaMemcpyAsync returns once the bytes left the host array (staged), so the next call may
    // rebuild it; on the device both uploads and the kernels are ordered by the stream
    if (nFlat) {
        e = cudaMemcpyAsync(dev, g_hostBoxes, (size_t)nFlat * sizeof(BoxRec), cudaMemcpyHostToDevice, c.stream);
        if (e != cudaSuccess) return e;
    }
    BoxRec* devRows = dev + (kNoAllocMaxBoxes - nRows);
    if (nRows) {
        e = cudaMemcpyAsync(devRows, g_hostBoxes + (kNoAllocMaxBoxes - nRows), (size_t)nRows * sizeof(BoxRec),
                            cudaMemcpyHostToDevice, c.stream);
        if (e != cudaSuccess) return e;
    }
    const uint8_t* sEnd = src.data + (size_t)src.dims[0] * src.dims[1] * src.dims[2] * bs;
    if (nFlat) {
        const unsigned perCta = kThreads * 4 * (16 / bs);
        const unsigned gx = (maxVox + (16 / bs) + perCta - 1) / perCta;
        dim3 grid((unsigned)nFlat, gx, 1);
        if (bs == 1) copyBoxesFlatKernel<1><<<grid, kThreads, 0, c.stream>>>(dev, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2]);
        else if (bs == 2) copyBoxesFlatKernel<2><<<grid, kThreads, 0, c.stream>>>(dev, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2]);
        else copyBoxesFlatKernel<4><<<grid, kThreads, 0, c.stream>>>(dev, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2]);
        countLaunch();
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    if (nRows) {
        // row-wise boxes were stored back to front: the kernel indexes them by blockIdx.y in any order
        int zPer = (int)((65536 + maxPlane - 1) / maxPlane);
        if (zPer < 1) zPer = 1;
        while (zPer > 1 && (long long)nRows * ((maxNz + zPer - 1) / zPer) < 2LL * c.smCount * 8) zPer = (zPer + 1) / 2;
        const RowsLaunch l = rowsLaunch(maxNx, 1, 16 / bs);
        dim3 grid((unsigned)((maxNz + zPer - 1) / zPer), (unsigned)nRows, 1);
        if (bs == 1) copyBoxesKernel<1><<<grid, l.block, 0, c.stream>>>(devRows, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2], zPer);
        else if (bs == 2) copyBoxesKernel<2><<<grid, l.block, 0, c.stream>>>(devRows, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2], zPer);
        else copyBoxesKernel<4><<<grid, l.block, 0, c.stream>>>(devRows, src.data, sEnd, src.dims[0], src.dims[1], src.dims[2], zPer);
        countLaunch();
        e = cudaGetLastError();
    }
    return e;
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
    p.sEnd = src.data + (size_t)src.dims[0] * src.dims[1] * src.dims[2] * bytesPerVoxel(src.format);
    p.shifted = 0;
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
