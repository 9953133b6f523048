// Arithmetic family (Sum, Diff, Prod, Quot, AbsDiff and their Safe* twins), device side.
//
// Semantics: reference src/vkt/Arithmetic_serial.hpp:15-45 (ArithmeticOp loop) and :47-260 (the ten
// lambdas): dst[p + off] = map_dst( f( unmap_s1(s1[p]), unmap_s2(s2[p]) ) ), p absolute in [first,last),
// Safe* = clamp(f, dst.lo, dst.hi) with clamp = max(lo, min(x, hi)) in the reference's operand order.
//
// Algorithmic traffic: bpv_s1 + bpv_s2 + bpv_dst bytes per voxel, every byte touched exactly once.
#pragma once

#include <type_traits>

#include "common.cuh"
#include "voxel_codec.cuh"

namespace vktb {

template <int OP>
__device__ __forceinline__ float arithEval(float a, float b, float lo, float hi)
{
    float r;
    switch (OP % 5) {
    case 0:  r = __fadd_rn(a, b); break;
    case 1:  r = __fsub_rn(a, b); break;
    case 2:  r = __fmul_rn(a, b); break;
    case 3:  r = __fdiv_rn(a, b); break;
    default: r = fabsf(__fsub_rn(a, b)); break;
    }
    return OP >= 5 ? refClamp(r, lo, hi) : r;
}

struct ArithParams
{
    Operand d, a, b;      // dst at (first+off), sources at first
    Range   g;
    Mapping md, ma, mb;
    int     alias;        // a source overlaps dst (in-place use): no non-coherent loads
    const uint8_t* aEnd;  // one past the last byte of each source volume (bounds the funnel-shifted loads)
    const uint8_t* bEnd;
    int     shifted;      // uniform kernel: a source row sits at another 16-byte phase than the dst row
    // exchange kernel: byte steps of the d / a / b row pointers for one stride of gridDim.y * blockDim.y rows, without
    // and with a y overflow into the next plane (host-computed so that the loop adds constants)
    long long step[3], stepWrap[3];
    int     sy;
};

// Homogeneous formats: one 128-bit vector per operand per chunk.
template <int F, int OP, bool IDENT, bool PROVEN = false>
__device__ __forceinline__ uint4 arithVec(const uint4& va, const uint4& vb, const Mapping& md,
                                          const Mapping& ma, const Mapping& mb, const float* lutA, const float* lutB)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    uint32_t w[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
    for (int i = 0; i < V; ++i) {
        uint32_t ra = vecGet<BPV>(va, i), rb = vecGet<BPV>(vb, i);
        float fa, fb;
        if (F == F_U8) { fa = lutA[ra]; fb = lutB[rb]; }
        else { fa = unmapVoxel<F, IDENT>(ra, ma); fb = unmapVoxel<F, IDENT>(rb, mb); }
        float r = arithEval<OP>(fa, fb, md.lo, md.hi);
        vecSet<BPV>(w, i, mapVoxel<F, IDENT, PROVEN>(r, md));
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

template <int F, int OP, bool IDENT>
__device__ __forceinline__ uint32_t arithOne(uint32_t ra, uint32_t rb, const Mapping& md, const Mapping& ma,
                                             const Mapping& mb)
{
    float fa = unmapVoxel<F, IDENT>(ra, ma), fb = unmapVoxel<F, IDENT>(rb, mb);
    return mapVoxel<F, IDENT>(arithEval<OP>(fa, fb, md.lo, md.hi), md);
}

// ---- flat: the range is one contiguous run of n voxels in all three operands, bases congruent mod 16 ----
template <int F, int OP, bool IDENT>
__global__ void __launch_bounds__(kThreads)
arithFlatKernel(uint8_t* __restrict__ d, const uint8_t* a, const uint8_t* b, size_t n, Mapping md, Mapping ma,
                Mapping mb, int alias)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    __shared__ float lutA[F == F_U8 ? 256 : 1], lutB[F == F_U8 ? 256 : 1];
    if (F == F_U8) {
        for (int i = threadIdx.x; i < 256; i += blockDim.x) {
            lutA[i] = unmapVoxel<F_U8, IDENT>((uint32_t)i, ma);
            lutB[i] = unmapVoxel<F_U8, IDENT>((uint32_t)i, mb);
        }
        __syncthreads();
    }

    // scalar head up to the first 16-byte boundary, vector body, scalar tail
    size_t head = ((16u - (reinterpret_cast<uintptr_t>(d) & 15u)) & 15u) / BPV;
    if (head > n) head = n;
    const size_t nvec = (n - head) / V;
    const size_t tail0 = head + nvec * V;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;

    const uint8_t* av = a + head * BPV;
    const uint8_t* bv = b + head * BPV;
    uint8_t* dv = d + head * BPV;

    // two vectors per operand in flight per thread; the division mode of map_dst is fixed for the launch, so the
    // loops exist twice: with the proven reciprocal form (md.exactRecip == 2, see voxel_codec.cuh) and without
    auto body = [&](auto provenTag) {
        constexpr bool PROVEN = decltype(provenTag)::value;
        size_t i = tid;
        for (; i + stride < nvec; i += 2 * stride) {
            uint4 a0, b0, a1, b1;
            if (alias) { a0 = ldPlain(av + i * 16); b0 = ldPlain(bv + i * 16);
                         a1 = ldPlain(av + (i + stride) * 16); b1 = ldPlain(bv + (i + stride) * 16); }
            else       { a0 = ldStream(av + i * 16); b0 = ldStream(bv + i * 16);
                         a1 = ldStream(av + (i + stride) * 16); b1 = ldStream(bv + (i + stride) * 16); }
            stStream(dv + i * 16, arithVec<F, OP, IDENT, PROVEN>(a0, b0, md, ma, mb, lutA, lutB));
            stStream(dv + (i + stride) * 16, arithVec<F, OP, IDENT, PROVEN>(a1, b1, md, ma, mb, lutA, lutB));
        }
        for (; i < nvec; i += stride) {
            uint4 a0 = alias ? ldPlain(av + i * 16) : ldStream(av + i * 16);
            uint4 b0 = alias ? ldPlain(bv + i * 16) : ldStream(bv + i * 16);
            stStream(dv + i * 16, arithVec<F, OP, IDENT, PROVEN>(a0, b0, md, ma, mb, lutA, lutB));
        }
    };
    if (!IDENT && md.exactRecip == 2) body(std::true_type{});
    else body(std::false_type{});
    // head + tail voxels (at most 2*(V-1)), one per thread of the first CTA(s)
    const size_t nEdge = head + (n - tail0);
    for (size_t e = tid; e < nEdge; e += stride) {
        size_t k = e < head ? e : tail0 + (e - head);
        uint32_t r = arithOne<F, OP, IDENT>(loadRaw(a + k * BPV, BPV), loadRaw(b + k * BPV, BPV), md, ma, mb);
        storeRaw(d + k * BPV, BPV, r);
    }
}

// ---- flat, UInt8 x UInt8 -> UInt8 through a 64 KiB pair table ------------------------------------------------------
// result[a][b] = map(f(unmap(a), unmap(b))) is built once per CTA in shared memory with the exact codec above; the
// stream then costs one byte-wide table load per voxel instead of two float look-ups plus ~14 arithmetic
// instructions (the per-operand tables of arithFlatKernel made UInt8 shared-memory bound at ~57 % of the HBM
// roofline; ncu: profiles/r01_notes.md).  Persistent grid, so the table is built 148 * k times per launch.
constexpr int kPairTableBytes = 65536;

template <int OP, bool IDENT>
__global__ void __launch_bounds__(kThreads)
arithPairTableKernel(uint8_t* __restrict__ d, const uint8_t* a, const uint8_t* b, size_t n, Mapping md, Mapping ma,
                     Mapping mb, int alias)
{
    extern __shared__ __align__(16) uint8_t pairTab[];
    __shared__ float lutB[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) lutB[i] = unmapVoxel<F_U8, IDENT>((uint32_t)i, mb);
    __syncthreads();
    // thread t builds row a = t (blockDim.x == 256): 256 entries, written as 64 words
    {
        const float fa = unmapVoxel<F_U8, IDENT>((uint32_t)threadIdx.x, ma);
        uint32_t* row = reinterpret_cast<uint32_t*>(pairTab + threadIdx.x * 256);
        for (int w = 0; w < 64; ++w) {
            // rotate the word order per thread: conflict-free shared-memory stores
            const int ww = (w + threadIdx.x) & 63;
            uint32_t word = 0u;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                word |= (mapVoxel<F_U8, IDENT>(arithEval<OP>(fa, lutB[ww * 4 + k], md.lo, md.hi), md) & 0xffu) << (8 * k);
            row[ww] = word;
        }
    }
    __syncthreads();

    size_t head = ((16u - (reinterpret_cast<uintptr_t>(d) & 15u)) & 15u);
    if (head > n) head = n;
    const size_t nvec = (n - head) / 16;
    const size_t tail0 = head + nvec * 16;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const uint8_t* av = a + head;
    const uint8_t* bv = b + head;
    uint8_t* dv = d + head;

    auto lookup = [&](const uint4& va, const uint4& vb) -> uint4 {
        const uint32_t wa[4] = { va.x, va.y, va.z, va.w }, wb[4] = { vb.x, vb.y, vb.z, vb.w };
        uint32_t w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t r = 0u;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t ia = (wa[q] >> (8 * k)) & 0xffu, ib = (wb[q] >> (8 * k)) & 0xffu;
                r |= (uint32_t)pairTab[ia * 256u + ib] << (8 * k);
            }
            w[q] = r;
        }
        return make_uint4(w[0], w[1], w[2], w[3]);
    };

    size_t i = tid;
    for (; i + stride < nvec; i += 2 * stride) {
        uint4 a0, b0, a1, b1;
        if (alias) { a0 = ldPlain(av + i * 16); b0 = ldPlain(bv + i * 16);
                     a1 = ldPlain(av + (i + stride) * 16); b1 = ldPlain(bv + (i + stride) * 16); }
        else       { a0 = ldStream(av + i * 16); b0 = ldStream(bv + i * 16);
                     a1 = ldStream(av + (i + stride) * 16); b1 = ldStream(bv + (i + stride) * 16); }
        stStream(dv + i * 16, lookup(a0, b0));
        stStream(dv + (i + stride) * 16, lookup(a1, b1));
    }
    for (; i < nvec; i += stride) {
        uint4 a0 = alias ? ldPlain(av + i * 16) : ldStream(av + i * 16);
        uint4 b0 = alias ? ldPlain(bv + i * 16) : ldStream(bv + i * 16);
        stStream(dv + i * 16, lookup(a0, b0));
    }
    const size_t nEdge = head + (n - tail0);
    for (size_t e = tid; e < nEdge; e += stride) {
        size_t k = e < head ? e : tail0 + (e - head);
        d[k] = pairTab[(uint32_t)a[k] * 256u + b[k]];
    }
}

// ---- rows, UInt8 x UInt8 -> UInt8 through the same pair table: *Range calls on big UInt8 volumes ------------------
// Rows of every operand in one 16-byte phase (pitches multiples of 16 bytes).  Work items are (row, dst-aligned 16-byte
// chunk) pairs in row-major order over a persistent grid, two in flight per thread; row and chunk come from the item
// number by two multiply-shift divisions (host-computed magic numbers, items < 2^31).  Sources at another phase than
// dst use the funnel-shifted loads.  The arithmetic kernels spend ~12 instructions per voxel, which caps UInt8 at
// about half of the HBM roofline; one table byte per voxel does not.
struct PairRows
{
    unsigned items, cpr;          // rows * cpr work items, cpr chunks per row
    unsigned mulC, mulY;          // item / cpr and row / ny as (x * mul) >> sh
    int shC, shY;
};

template <int OP, bool IDENT>
__global__ void __launch_bounds__(kThreads)
arithRowsPairTableKernel(ArithParams p, int shift, PairRows w)
{
    extern __shared__ __align__(16) uint8_t pairTab[];
    __shared__ float lutB[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) lutB[i] = unmapVoxel<F_U8, IDENT>((uint32_t)i, p.mb);
    __syncthreads();
    {   // thread t builds row a = t (blockDim.x == 256), word order rotated per thread: conflict-free stores
        const float fa = unmapVoxel<F_U8, IDENT>((uint32_t)threadIdx.x, p.ma);
        uint32_t* row = reinterpret_cast<uint32_t*>(pairTab + threadIdx.x * 256);
        for (int q = 0; q < 64; ++q) {
            const int ww = (q + threadIdx.x) & 63;
            uint32_t word = 0u;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                word |= (mapVoxel<F_U8, IDENT>(arithEval<OP>(fa, lutB[ww * 4 + k], p.md.lo, p.md.hi), p.md) & 0xffu) << (8 * k);
            row[ww] = word;
        }
    }
    __syncthreads();

    // phase of a full chunk's first source byte inside its aligned vector: the same for every row and chunk
    const unsigned shA = (unsigned)((reinterpret_cast<uintptr_t>(p.a.base) + (uintptr_t)(16 - shift)) & 15u);
    const unsigned shB = (unsigned)((reinterpret_cast<uintptr_t>(p.b.base) + (uintptr_t)(16 - shift)) & 15u);
    // fetch() only issues loads (the aligned vectors around the chunk); the funnel shifts wait until finish(), so the
    // loads of both items in flight go out back to back.  bytewise: the one chunk whose second vector would leave
    // the source buffer.
    struct Item { uint8_t* d; const uint8_t* a; const uint8_t* b; int n; bool bytewise; uint4 la, ha, lb, hb; };
    // through L1: a lane's second vector is the next lane's first
    auto ld16 = [&](const uint8_t* q) { return p.alias ? ldPlain(q) : __ldg(reinterpret_cast<const uint4*>(q)); };
    auto fetch = [&](unsigned item) {
        Item it;
        const unsigned r = (unsigned)(((unsigned long long)item * w.mulC) >> w.shC), c = item - r * w.cpr;
        const unsigned z = (unsigned)(((unsigned long long)r * w.mulY) >> w.shY), y = r - z * (unsigned)p.g.ny;
        int xr = (int)c * 16 - shift, n = 16;
        if (xr < 0) { n += xr; xr = 0; }
        if (xr + n > p.g.nx) n = p.g.nx - xr;
        it.n = n;
        it.d = p.d.base + y * p.d.pitchY + z * p.d.pitchZ + xr;
        it.a = p.a.base + y * p.a.pitchY + z * p.a.pitchZ + xr;
        it.b = p.b.base + y * p.b.pitchY + z * p.b.pitchZ + xr;
        it.la = it.ha = it.lb = it.hb = make_uint4(0u, 0u, 0u, 0u);
        it.bytewise = n == 16 && !(unalignedReadable(it.a, p.aEnd) && unalignedReadable(it.b, p.bEnd));
        if (n == 16 && !it.bytewise) {
            it.la = ld16(it.a - shA); it.lb = ld16(it.b - shB);
            if (shA != 0u) it.ha = ld16(it.a - shA + 16);
            if (shB != 0u) it.hb = ld16(it.b - shB + 16);
        }
        return it;
    };
    auto finish = [&](const Item& it) {
        if (it.n == 16 && !it.bytewise) {
            const uint4 va = shA != 0u ? window16(it.la, it.ha, shA) : it.la;
            const uint4 vb = shB != 0u ? window16(it.lb, it.hb, shB) : it.lb;
            const uint32_t wa[4] = { va.x, va.y, va.z, va.w }, wb[4] = { vb.x, vb.y, vb.z, vb.w };
            uint32_t o[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uint32_t r = 0u;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t ia = (wa[q] >> (8 * k)) & 0xffu, ib = (wb[q] >> (8 * k)) & 0xffu;
                    r |= (uint32_t)pairTab[ia * 256u + ib] << (8 * k);
                }
                o[q] = r;
            }
            stStream(it.d, make_uint4(o[0], o[1], o[2], o[3]));
        } else {
            for (int i = 0; i < it.n; ++i) it.d[i] = pairTab[(uint32_t)it.a[i] * 256u + it.b[i]];
        }
    };
    const unsigned stride = gridDim.x * blockDim.x;
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    for (; (unsigned long long)i + stride < w.items; i += 2 * stride) {
        const Item i0 = fetch(i), i1 = fetch(i + stride);
        finish(i0); finish(i1);
    }
    if (i < w.items) finish(fetch(i));
}

// ---- rows: general 3-D range with dstOffset; chunks are cut at 16-byte boundaries of the dst row --------
template <int F, int OP, bool IDENT>
__global__ void __launch_bounds__(kThreads)
arithRowsKernel(ArithParams p)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    forEachRowChunk<V, BPV>(p.g, p.d, [&](int y, int z, int xr, int n, bool full) {
        uint8_t* dp = p.d.base + y * p.d.pitchY + z * p.d.pitchZ + (long long)xr * BPV;
        const uint8_t* ap = p.a.base + y * p.a.pitchY + z * p.a.pitchZ + (long long)xr * BPV;
        const uint8_t* bp = p.b.base + y * p.b.pitchY + z * p.b.pitchZ + (long long)xr * BPV;
        const bool aligned = ((reinterpret_cast<uintptr_t>(ap) | reinterpret_cast<uintptr_t>(bp)) & 15u) == 0;
        if (full && (aligned || (unalignedReadable(ap, p.aEnd) && unalignedReadable(bp, p.bEnd)))) {
            uint4 va, vb;
            if (!aligned)     { va = ldUnaligned(ap, p.alias); vb = ldUnaligned(bp, p.alias); }
            else if (p.alias) { va = ldPlain(ap); vb = ldPlain(bp); }
            else              { va = ldStream(ap); vb = ldStream(bp); }
            uint32_t w[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
            for (int i = 0; i < V; ++i)
                vecSet<BPV>(w, i, arithOne<F, OP, IDENT>(vecGet<BPV>(va, i), vecGet<BPV>(vb, i), p.md, p.ma, p.mb));
            stStream(dp, make_uint4(w[0], w[1], w[2], w[3]));
        } else {
            for (int i = 0; i < n; ++i) {
                uint32_t r = arithOne<F, OP, IDENT>(loadRaw(ap + i * BPV, BPV), loadRaw(bp + i * BPV, BPV),
                                                    p.md, p.ma, p.mb);
                storeRaw(dp + i * BPV, BPV, r);
            }
        }
    });
}

// ---- rows, every row of all three operands in the same 16-byte phase: two rows (four loads) in flight ----
struct ArithPair { uint4 a, b; };

template <int F, int OP, bool IDENT, bool SHIFTED>
__global__ void __launch_bounds__(kThreads)
arithRowsUniformKernel(ArithParams p, int shift)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    auto emit = [&](int y, int z, int xr, const uint4& va, const uint4& vb) {
        uint32_t w[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
        for (int i = 0; i < V; ++i)
            vecSet<BPV>(w, i, arithOne<F, OP, IDENT>(vecGet<BPV>(va, i), vecGet<BPV>(vb, i), p.md, p.ma, p.mb));
        stStream(p.d.base + y * p.d.pitchY + z * p.d.pitchZ + (long long)xr * BPV, make_uint4(w[0], w[1], w[2], w[3]));
    };
    auto edge = [&](int y, int z, int xr, int n) {
        uint8_t* dp = p.d.base + y * p.d.pitchY + z * p.d.pitchZ + (long long)xr * BPV;
        const uint8_t* ap = p.a.base + y * p.a.pitchY + z * p.a.pitchZ + (long long)xr * BPV;
        const uint8_t* bp = p.b.base + y * p.b.pitchY + z * p.b.pitchZ + (long long)xr * BPV;
        for (int i = 0; i < n; ++i)
            storeRaw(dp + i * BPV, BPV, arithOne<F, OP, IDENT>(loadRaw(ap + i * BPV, BPV), loadRaw(bp + i * BPV, BPV),
                                                              p.md, p.ma, p.mb));
    };
    if (!SHIFTED) {     // a template parameter: the aligned kernel keeps its 64 registers (4 CTAs per SM)
        forEachRowChunkUniform<V, 2, ArithPair>(p.g, shift,
            [&](int y, int z, int xr) {
                const uint8_t* ap = p.a.base + y * p.a.pitchY + z * p.a.pitchZ + (long long)xr * BPV;
                const uint8_t* bp = p.b.base + y * p.b.pitchY + z * p.b.pitchZ + (long long)xr * BPV;
                ArithPair q;
                if (p.alias) { q.a = ldPlain(ap); q.b = ldPlain(bp); }
                else         { q.a = ldStream(ap); q.b = ldStream(bp); }
                return q;
            },
            [&](int y, int z, int xr, const ArithPair& q) { emit(y, z, xr, q.a, q.b); },
            edge);
        return;
    }
    // a source at another phase than dst: the load step only issues the aligned vectors around the chunk (through L1:
    // a lane's second vector is the next lane's first), the funnel shifts wait for the store step, so the loads of both
    // rows in flight go out back to back.  The one chunk whose second vector would leave a source buffer runs per voxel.
    struct Quad { uint4 la, ha, lb, hb; };
    const unsigned shA = (unsigned)((reinterpret_cast<uintptr_t>(p.a.base) + (uintptr_t)(16 - shift * BPV)) & 15u);
    const unsigned shB = (unsigned)((reinterpret_cast<uintptr_t>(p.b.base) + (uintptr_t)(16 - shift * BPV)) & 15u);
    auto ld16 = [&](const uint8_t* q) { return p.alias ? ldPlain(q) : __ldg(reinterpret_cast<const uint4*>(q)); };
    forEachRowChunkUniform<V, 2, Quad>(p.g, shift,
        [&](int y, int z, int xr) {
            const uint8_t* ap = p.a.base + y * p.a.pitchY + z * p.a.pitchZ + (long long)xr * BPV;
            const uint8_t* bp = p.b.base + y * p.b.pitchY + z * p.b.pitchZ + (long long)xr * BPV;
            Quad q;
            q.la = q.ha = q.lb = q.hb = make_uint4(0u, 0u, 0u, 0u);
            if (unalignedReadable(ap, p.aEnd) && unalignedReadable(bp, p.bEnd)) {
                q.la = ld16(ap - shA); q.lb = ld16(bp - shB);
                if (shA != 0u) q.ha = ld16(ap - shA + 16);
                if (shB != 0u) q.hb = ld16(bp - shB + 16);
            }
            return q;
        },
        [&](int y, int z, int xr, const Quad& q) {
            const uint8_t* ap = p.a.base + y * p.a.pitchY + z * p.a.pitchZ + (long long)xr * BPV;
            const uint8_t* bp = p.b.base + y * p.b.pitchY + z * p.b.pitchZ + (long long)xr * BPV;
            if (unalignedReadable(ap, p.aEnd) && unalignedReadable(bp, p.bEnd))
                emit(y, z, xr, shA != 0u ? window16(q.la, q.ha, shA) : q.la, shB != 0u ? window16(q.lb, q.hb, shB) : q.lb);
            else edge(y, z, xr, V);
        },
        edge);
}

// ---- rows in one 16-byte phase, sources at another phase: every lane loads ONE aligned vector per source and takes
// the words it misses from the next lane (whose aligned vector they are); only lanes without such a neighbour
// (warp end, row end) load the second vector themselves.  The kernel above reads two aligned vectors per source
// instead and recomputes every row address; this one walks its three row pointers.  blockDim.x is a multiple of 32
// (a warp lies in one row) and no lane leaves before the last row, so the shuffles run with the full mask.  No
// operand aliases dst.
template <int F, int OP, bool IDENT>
__global__ void __launch_bounds__(kThreads)
arithRowsShiftedKernel(ArithParams p, int shift)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    const int lane = threadIdx.x & 31;
    const int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    int xr = chunk * V - shift, n = V;
    if (xr < 0) { n += xr; xr = 0; }
    if (xr + n > p.g.nx) n = p.g.nx - xr;
    const bool full = n == V;
    // the next lane holds the following full chunk of this row
    const bool nextFull = full && lane < 31 && xr + 2 * V <= p.g.nx;
    const bool selfHi = full && !nextFull;
    const long long xb = (long long)xr * BPV;
    // phase of a full chunk's first source byte inside its aligned vector: the same for every row and chunk
    const unsigned shA = (unsigned)((reinterpret_cast<uintptr_t>(p.a.base) + (uintptr_t)(16 - shift * BPV)) & 15u);
    const unsigned shB = (unsigned)((reinterpret_cast<uintptr_t>(p.b.base) + (uintptr_t)(16 - shift * BPV)) & 15u);
    const unsigned rows = (unsigned)p.g.ny * (unsigned)p.g.nz;
    const unsigned S = gridDim.y * blockDim.y;
    unsigned r = blockIdx.y * blockDim.y + threadIdx.y;
    if (r >= rows) return;                         // whole warps: threadIdx.y is warp-uniform
    int z = (int)(r / (unsigned)p.g.ny), y = (int)(r - (unsigned)z * (unsigned)p.g.ny);
    // row pointers walk by S rows; a y overflow moves them to the next plane
    const uint8_t* pa = p.a.base + y * p.a.pitchY + z * p.a.pitchZ + xb - (full ? (long long)shA : 0);
    const uint8_t* pb = p.b.base + y * p.b.pitchY + z * p.b.pitchZ + xb - (full ? (long long)shB : 0);
    uint8_t* pd = p.d.base + y * p.d.pitchY + z * p.d.pitchZ + xb;
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    for (; r < rows; r += S) {
        // all loads of the row up front: the lanes without a neighbour fetch their second vectors now, not after the
        // shuffles (a dependent load there would stall the whole warp for one more memory latency).  Issuing the next
        // row's loads before this row is worked on was measured too: UInt16 unchanged, UInt8 19.1 -> 23.3 ms.
        uint4 la = zero, lb = zero, sa = zero, sb = zero;
        // (through L1: with L1::no_allocate loads the second vectors of the warp-end lanes and the sectors two warps
        // share are fetched from L2 twice -- Float32 10.2 against 8.1 ms)
        if (full) { la = __ldg(reinterpret_cast<const uint4*>(pa)); lb = __ldg(reinterpret_cast<const uint4*>(pb)); }
        if (selfHi && shA != 0u) sa = __ldg(reinterpret_cast<const uint4*>(pa + 16));
        if (selfHi && shB != 0u) sb = __ldg(reinterpret_cast<const uint4*>(pb + 16));
        uint4 qa = la, qb = lb;
        if (shA != 0u) {
            const uint4 ha = shflDown1(la);
            qa = window16(la, nextFull ? ha : sa, shA);
        }
        if (shB != 0u) {
            const uint4 hb = shflDown1(lb);
            qb = window16(lb, nextFull ? hb : sb, shB);
        }
        if (full) {
            uint32_t w[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
            for (int i = 0; i < V; ++i)
                vecSet<BPV>(w, i, arithOne<F, OP, IDENT>(vecGet<BPV>(qa, i), vecGet<BPV>(qb, i), p.md, p.ma, p.mb));
            stStream(pd, make_uint4(w[0], w[1], w[2], w[3]));
        } else {
            for (int i = 0; i < n; ++i)
                storeRaw(pd + i * BPV, BPV, arithOne<F, OP, IDENT>(loadRaw(pa + i * BPV, BPV), loadRaw(pb + i * BPV, BPV),
                                                                  p.md, p.ma, p.mb));
        }
        y += p.sy;
        const bool wrap = y >= p.g.ny;
        if (wrap) y -= p.g.ny;
        pd += wrap ? p.stepWrap[0] : p.step[0];
        pa += wrap ? p.stepWrap[1] : p.step[1];
        pb += wrap ? p.stepWrap[2] : p.step[2];
    }
}

// ---- generic: any mix of formats (runtime codec), one voxel per thread along x ---------------------------
template <int OP>
__global__ void __launch_bounds__(kThreads)
arithGenericKernel(ArithParams p, int fd, int fa, int fb)
{
    const int bd = bytesPerVoxel(fd), ba = bytesPerVoxel(fa), bb = bytesPerVoxel(fb);
    const long long rows = (long long)p.g.ny * p.g.nz;
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const int z = (int)(r / p.g.ny), y = (int)(r - (long long)z * p.g.ny);
        uint8_t* drow = p.d.base + y * p.d.pitchY + z * p.d.pitchZ;
        const uint8_t* arow = p.a.base + y * p.a.pitchY + z * p.a.pitchZ;
        const uint8_t* brow = p.b.base + y * p.b.pitchY + z * p.b.pitchZ;
        for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < p.g.nx; x += gridDim.x * blockDim.x) {
            float va = unmapVoxelRt(fa, loadRaw(arow + (long long)x * ba, ba), p.ma);
            float vb = unmapVoxelRt(fb, loadRaw(brow + (long long)x * bb, bb), p.mb);
            float rr = arithEval<OP>(va, vb, p.md.lo, p.md.hi);
            storeRaw(drow + (long long)x * bd, bd, mapVoxelRt(fd, rr, p.md));
        }
    }
}

} // namespace vktb
