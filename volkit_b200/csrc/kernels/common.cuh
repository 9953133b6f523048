// Shared device helpers: 128-bit streaming loads/stores, range geometry, the row-chunk driver.
//
// All in-scope algorithms are HBM-bound byte/float streams; the rules that matter are
// 128-bit coalesced x-fastest accesses, enough bytes in flight per SM and grids sized in multiples of
// the SM count.  No tensor cores anywhere.
#pragma once

#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

namespace vktb {

constexpr int kThreads = 256;   // CTA size of every streaming kernel

// ---- 128-bit global accesses with streaming hints (nothing is re-used: keep L1 clean) -----------------
__device__ __forceinline__ uint4 ldStream(const void* p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// plain (coherent) 128-bit load: for operands that may alias the destination (in-place arithmetic)
__device__ __forceinline__ uint4 ldPlain(const void* p)
{
    return *reinterpret_cast<const uint4*>(p);
}
__device__ __forceinline__ void stStream(void* p, const uint4& v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// 16 bytes from an address of any alignment: the two enclosing aligned vectors, word-rotated and funnel-shifted
// (a *Range call whose sources sit at another 16-byte phase than the destination would otherwise fall back to
// per-voxel loads).  Both vectors go through L1 (the neighbouring thread re-reads `hi` as its `lo`).  The caller
// guarantees that the second vector is readable: (p & ~15) + 32 <= end of the buffer.  `coherent`: the source may
// alias memory this kernel writes.
__device__ __forceinline__ uint4 ldUnaligned(const uint8_t* p, bool coherent = false)
{
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint4* v = reinterpret_cast<const uint4*>(a & ~(uintptr_t)15);
    const unsigned sh = (unsigned)a & 15u;
    const uint4 lo = coherent ? *v : __ldg(v);
    if (sh == 0) return lo;
    const uint4 hi = coherent ? v[1] : __ldg(v + 1);
    uint32_t w0 = lo.x, w1 = lo.y, w2 = lo.z, w3 = lo.w, w4 = hi.x, w5 = hi.y;
    if (sh & 8u) { w0 = w2; w1 = w3; w2 = w4; w3 = w5; w4 = hi.z; w5 = hi.w; }
    if (sh & 4u) { w0 = w1; w1 = w2; w2 = w3; w3 = w4; w4 = w5; }
    const unsigned bits = (sh & 3u) * 8u;
    return make_uint4(__funnelshift_r(w0, w1, bits), __funnelshift_r(w1, w2, bits),
                      __funnelshift_r(w2, w3, bits), __funnelshift_r(w3, w4, bits));
}
// whether ldUnaligned(p) stays inside a buffer that ends at `end`
__device__ __forceinline__ bool unalignedReadable(const uint8_t* p, const uint8_t* end)
{
    return (reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)15) + 32u <= reinterpret_cast<uintptr_t>(end)
        || (reinterpret_cast<uintptr_t>(p) & 15u) == 0;
}

// ldUnaligned with the (rare) byte-wise fallback for the one chunk whose second vector would leave the buffer
__device__ __forceinline__ uint4 ldUnalignedWithin(const uint8_t* p, const uint8_t* end, bool coherent = false)
{
    if (unalignedReadable(p, end)) return ldUnaligned(p, coherent);
    uint32_t w[4] = { 0u, 0u, 0u, 0u };
    for (int i = 0; i < 16; ++i) w[i >> 2] |= (uint32_t)p[i] << ((i & 3) * 8);
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// neighbour-lane exchange (arithRowsShiftedKernel): the next lane's vector, and the 16
// bytes at phase sh of two consecutive vectors
__device__ __forceinline__ uint4 shflDown1(const uint4& v)
{
    return make_uint4(__shfl_down_sync(0xffffffffu, v.x, 1), __shfl_down_sync(0xffffffffu, v.y, 1),
                      __shfl_down_sync(0xffffffffu, v.z, 1), __shfl_down_sync(0xffffffffu, v.w, 1));
}

// bytes [sh, sh + 16) of the 32 bytes lo | hi  (sh uniform, 1..15)
__device__ __forceinline__ uint4 window16(const uint4& lo, const uint4& hi, unsigned sh)
{
    uint32_t w0 = lo.x, w1 = lo.y, w2 = lo.z, w3 = lo.w, w4 = hi.x, w5 = hi.y;
    if (sh & 8u) { w0 = w2; w1 = w3; w2 = w4; w3 = w5; w4 = hi.z; w5 = hi.w; }
    if (sh & 4u) { w0 = w1; w1 = w2; w2 = w3; w3 = w4; w4 = w5; }
    const unsigned bits = (sh & 3u) * 8u;
    return make_uint4(__funnelshift_r(w0, w1, bits), __funnelshift_r(w1, w2, bits),
                      __funnelshift_r(w2, w3, bits), __funnelshift_r(w3, w4, bits));
}

__device__ __forceinline__ void stStream32(void* p, uint32_t v)
{
    asm volatile("st.global.L1::no_allocate.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// ---- element access inside a 16-byte register vector --------------------------------------------------
// BPV = 1, 2 or 4; i is a compile-time constant after unrolling.
template <int BPV>
__device__ __forceinline__ uint32_t vecGet(const uint4& v, int i)
{
    const uint32_t w[4] = { v.x, v.y, v.z, v.w };
    if (BPV == 4) return w[i];
    if (BPV == 2) return (w[i >> 1] >> ((i & 1) * 16)) & 0xffffu;
    return (w[i >> 2] >> ((i & 3) * 8)) & 0xffu;
}
template <int BPV>
__device__ __forceinline__ void vecSet(uint32_t (&w)[4], int i, uint32_t e)
{
    if (BPV == 4) w[i] = e;
    else if (BPV == 2) w[i >> 1] |= (e & 0xffffu) << ((i & 1) * 16);
    else w[i >> 2] |= (e & 0xffu) << ((i & 3) * 8);
}

// ---- 3-D range geometry ----------------------------------------------------------------------------------
// One operand of a range operation: `base` addresses the operand's voxel that corresponds to the first
// voxel of the range; pitches are in BYTES.
struct Operand
{
    uint8_t*  base;
    long long pitchY;
    long long pitchZ;
};

struct Range
{
    int nx, ny, nz;   // extents of [first,last)
    int x0, y0, z0;   // absolute coordinates of the first voxel (ops that see coordinates; copy clamps)
};

// Launch shape of a rows kernel: blockDim = (tx, 256/tx), grid = (chunk blocks, row blocks).
struct RowsLaunch { dim3 grid, block; };

inline RowsLaunch rowsLaunch(int nx, long long rows, int vecElems)
{
    // +1 chunk: the dst-aligned chunking may split the row's head and tail across two extra chunks
    int chunks = (nx + vecElems - 1) / vecElems + 1;
    int tx = 1;
    while (tx < chunks && tx < kThreads) tx <<= 1;
    int ty = kThreads / tx;
    RowsLaunch l;
    l.block = dim3(tx, ty, 1);
    long long gy = (rows + ty - 1) / ty;
    // VKTB_MAX_GRID_Y (tests): a small cap makes every thread walk many rows on volumes the oracle finishes quickly
    const char* e = getenv("VKTB_MAX_GRID_Y");
    const long long v = e ? atoll(e) : 0, cap = v >= 1 && v <= 65535 ? v : 65535LL;
    if (gy > cap) gy = cap;
    l.grid = dim3((chunks + tx - 1) / tx, (unsigned)gy, 1);
    return l;
}

// The row-chunk driver.  Rows are the (y,z) lines of the range; each thread owns chunks of V consecutive
// voxels cut at 16-byte boundaries of the DESTINATION row, so full chunks can be stored with one 128-bit
// store.  body(y, z, xr, n, full) is called with xr = first range-relative x of the chunk, n = valid voxels
// (n == V && full -> the dst address of xr is 16-byte aligned and the chunk lies fully inside the row).
template <int V, int BPVD, class Body>
__device__ __forceinline__ void forEachRowChunk(const Range& g, const Operand& dst, Body body)
{
    const long long rows = (long long)g.ny * g.nz;
    const int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    const bool small = rows < 0x7fffffffLL;   // 32-bit row arithmetic (the 64-bit division is a ~100-instruction call)
    for (long long r = (long long)blockIdx.y * blockDim.y + threadIdx.y; r < rows;
         r += (long long)gridDim.y * blockDim.y) {
        const int z = small ? (int)((unsigned)r / (unsigned)g.ny) : (int)(r / g.ny);
        const int y = (int)(r - (long long)z * g.ny);
        const uint8_t* drow = dst.base + y * dst.pitchY + z * dst.pitchZ;
        const int shift = (int)((reinterpret_cast<uintptr_t>(drow) & 15u) / BPVD);   // voxels into a 16B line
        int xr = chunk * V - shift;
        int n = V;
        if (xr < 0) { n += xr; xr = 0; }
        if (xr + n > g.nx) n = g.nx - xr;
        if (n <= 0) continue;
        body(y, z, xr, n, n == V);
    }
}

// The row-chunk driver for ranges whose rows all share one 16-byte phase (every pitch of every operand is a
// multiple of 16 and the operands are congruent mod 16; the launcher checks): the thread's chunk [xr, xr + n) is the
// same in every row, rows advance by add-with-carry instead of a division, and full chunks run U rows at a time --
// U independent 128-bit loads per operand in flight before the first store (a single row per iteration leaves a
// copy with 16 B in flight per thread, ~80 % of the HBM roofline; ncu: profiles/r01_notes.md).
//   load(y, z, xr) -> Regs        reads the chunk's vectors (no stores)
//   store(y, z, xr, regs)         computes and stores one full chunk
//   edge(y, z, xr, n)             scalar path for the row's partial head / tail chunk
// rows < 2^31 (launcher).
template <int V, int U, class Regs, class Load, class Store, class Edge>
__device__ __forceinline__ void forEachRowChunkUniform(const Range& g, int shift, Load load, Store store, Edge edge)
{
    const int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    int xr = chunk * V - shift, n = V;
    if (xr < 0) { n += xr; xr = 0; }
    if (xr + n > g.nx) n = g.nx - xr;
    if (n <= 0) return;
    const unsigned rows = (unsigned)g.ny * (unsigned)g.nz;
    const unsigned S = gridDim.y * blockDim.y;
    unsigned r = blockIdx.y * blockDim.y + threadIdx.y;
    if (r >= rows) return;
    int z = (int)(r / (unsigned)g.ny), y = (int)(r - (unsigned)z * (unsigned)g.ny);
    const int sz = (int)(S / (unsigned)g.ny), sy = (int)(S - (unsigned)sz * (unsigned)g.ny);
    auto next = [&] { y += sy; z += sz; if (y >= g.ny) { y -= g.ny; ++z; } };
    const bool full = n == V;
    for (; (unsigned long long)r + (unsigned long long)(U - 1) * S < rows; r += U * S) {
        int ys[U], zs[U];
        Regs q[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            ys[u] = y; zs[u] = z;
            if (full) q[u] = load(y, z, xr);
            next();
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (full) store(ys[u], zs[u], xr, q[u]);
            else edge(ys[u], zs[u], xr, n);
        }
    }
    for (; r < rows; r += S) {
        if (full) { Regs q = load(y, z, xr); store(y, z, xr, q); }
        else edge(y, z, xr, n);
        next();
    }
}

// floor(e / d) for e < 2^31 as one widening multiply and a shift: m = ceil(2^(31+l) / d), l = ceil(log2 d).
// The error e*r/(d*2^(31+l)) stays below 2^-l <= 1/d, so the floor never moves.
inline void magicDivisor(unsigned d, unsigned& mul, int& sh)
{
    int l = 0;
    while ((1ull << l) < d) ++l;
    const unsigned long long p = 1ull << (31 + l);
    mul = (unsigned)((p + d - 1) / d);
    sh = 31 + l;
}
// Grid for flat (1-D) streaming kernels: enough CTAs to fill every SM several times over, capped so the
// grid-stride loop amortises the launch; always a multiple of the SM count.
inline unsigned flatGrid(size_t work, int perCta, int smCount, int ctasPerSm)
{
    size_t need = (work + (size_t)perCta - 1) / (size_t)perCta;
    size_t cap = (size_t)smCount * (size_t)ctasPerSm;
    if (need >= cap) return (unsigned)cap;
    size_t g = (need + (size_t)smCount - 1) / (size_t)smCount * (size_t)smCount;   // round up to SM multiple
    if (g == 0) g = (size_t)smCount;
    return (unsigned)(g < cap ? g : cap);
}

} // namespace vktb
