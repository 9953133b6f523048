// ComputeAggregates / ComputeAggregatesRange -- one HBM pass, warp-shuffle reductions.
//
// Reference: src/vkt/Aggregates_serial.hpp:20-83.  The reference walks the range twice (x fastest) with BINARY32
// running accumulators: min / max (strict compares, so the first occurrence wins and NaN never enters), sum, mean
// (= sum), prod, then var += (v - mean)^2.  min, max, argmin and argmax are order-independent facts and are
// reproduced exactly, bit pattern of the first occurrence included.  The four accumulated statistics are NOT
// reproducible by any parallel evaluation (binary32 addition does not associate, and the running sum of a large
// volume stops growing once it passes 2^24 times the voxel values); for them this kernel returns moments that are
// accurate to ~1e-7 of the real-number value -- per-vector binary32 partial sums of SHIFTED values d = v - c (c =
// the range's first voxel, which keeps near-constant volumes free of cancellation), accumulated in binary64; UInt8 and
// UInt16 volumes with the identity mapping, whose unmap is linear in the raw integer, use exact integer moments
// instead (sum raw by dp4a / dp2a, sum raw^2 by dp4a / 64-bit multiply-adds) -- and the contract (include/vktb/backend.h) is the reference's own a-priori rounding bound around that value.
//
// One pass instead of the reference's two: sum (v - mean)^2 = sum d^2 - 2 (mean - c) sum d + n (mean - c)^2 is
// evaluated in binary64 by the finisher (launchAggregates), `mean` being the binary32 mean the reference would
// subtract.  prod is carried as mantissa x 2^exponent (binary64 mantissa renormalised every four factors), so it
// neither overflows nor underflows before the final rounding to binary32; on the unsigned formats with the
// identity mapping every factor is below 1, and a thread whose partial product has fallen below 2^-400 stops
// multiplying (the total is then 0 in binary32 whatever follows).
//
// Algorithmic bytes: bpv per voxel (one read).  Loads: 4 x 128-bit streaming loads in flight per thread (flat) or
// 2 per row (ranges); min/max of UInt8/UInt16 volumes with the identity mapping run on packed integers
// (__vminu4 / __vminu2: the unmap is strictly increasing there), positions are resolved once per thread at the end
// by re-reading the one vector that last lowered (raised) its minimum (maximum).
#include "launch.hpp"

#include <cfloat>
#include <cmath>

namespace vktb {
namespace {

constexpr unsigned long long kNone = ~0ull;
constexpr unsigned long long kVecTag = 1ull << 63;

struct AggParams
{
    const uint8_t* s;
    long long pitchY, pitchZ;
    int nx, ny, nz;
    int flat;
    size_t n;
    Mapping m;
    AggPartial* partials;   // [gridDim.x] CTA partials, then [gridDim.x] = the total written by aggFinishKernel
    unsigned numPartials;
};

__device__ __forceinline__ void normProd(double& P, long long& E)
{
    int hi = __double2hiint(P);
    const int e = (hi >> 20) & 0x7ff;
    if (e != 0 && e != 0x7ff) {
        E += e - 1023;
        hi = (hi & 0x800fffff) | (1023 << 20);
        P = __hiloint2double(hi, __double2loint(P));
    }
}

// (value, index) candidates: the smaller value wins, equal values keep the smaller linear index (= the voxel the
// serial walk meets first); kNone loses every tie
__device__ __forceinline__ void takeMin(float& v, unsigned long long& i, float ov, unsigned long long oi)
{
    if (ov < v || (ov == v && oi < i)) { v = ov; i = oi; }
}
__device__ __forceinline__ void takeMax(float& v, unsigned long long& i, float ov, unsigned long long oi)
{
    if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
}

struct AggThread
{
    double sumD, sumD2, P;
    long long E;
    float mn, mx;
    unsigned long long mnIdx, mxIdx;
};

__device__ __forceinline__ void combine(AggThread& a, const AggThread& b)
{
    a.sumD += b.sumD; a.sumD2 += b.sumD2;
    a.P *= b.P; a.E += b.E; normProd(a.P, a.E);
    takeMin(a.mn, a.mnIdx, b.mn, b.mnIdx);
    takeMax(a.mx, a.mxIdx, b.mx, b.mxIdx);
}

__device__ __forceinline__ AggThread shuffleXor(const AggThread& a, int lane)
{
    AggThread b;
    b.sumD = __shfl_xor_sync(0xffffffffu, a.sumD, lane);
    b.sumD2 = __shfl_xor_sync(0xffffffffu, a.sumD2, lane);
    b.P = __shfl_xor_sync(0xffffffffu, a.P, lane);
    b.E = __shfl_xor_sync(0xffffffffu, a.E, lane);
    b.mn = __shfl_xor_sync(0xffffffffu, a.mn, lane);
    b.mx = __shfl_xor_sync(0xffffffffu, a.mx, lane);
    b.mnIdx = __shfl_xor_sync(0xffffffffu, a.mnIdx, lane);
    b.mxIdx = __shfl_xor_sync(0xffffffffu, a.mxIdx, lane);
    return b;
}

// CTA reduction in a fixed order (xor butterfly inside the warp -- every operation in it is commutative on the
// pair, so all lanes agree -- then warp 0..7 in sequence); the result is valid in thread 0
__device__ __forceinline__ void blockReduce(AggThread& t, AggThread* sh /*[kThreads / 32]*/)
{
#pragma unroll
    for (int l = 16; l >= 1; l >>= 1) { AggThread o = shuffleXor(t, l); combine(t, o); }
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0)
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) combine(t, sh[w]);
}

__device__ __forceinline__ void storePartial(AggPartial* out, const AggThread& t, float shift)
{
    out->sumD = t.sumD; out->sumD2 = t.sumD2; out->prodMant = t.P; out->prodExp = t.E;
    out->minIdx = t.mnIdx; out->maxIdx = t.mxIdx; out->min = t.mn; out->max = t.mx; out->shift = shift; out->pad = 0;
}

template <int F, bool IDENT>
__global__ void __launch_bounds__(kThreads) aggregatesKernel(AggParams p)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    constexpr bool PACKED = IDENT && (F == F_U8 || F == F_U16);         // min/max on raw integers
    constexpr bool CONTRACTING = IDENT && (F == F_U8 || F == F_U16 || F == F_U32);   // every factor of prod is < 1
    constexpr bool FINITE = F != F_F32;                                  // 0 stays 0: no inf / NaN factors
    __shared__ float lut[F == F_U8 ? 256 : 1];
    __shared__ AggThread red[kThreads / 32];
    if (F == F_U8) { buildU8Table<IDENT>(lut, p.m); __syncthreads(); }

    auto value = [&](uint32_t raw) -> float {
        if (F == F_U8) return lut[raw & 0xffu];
        return unmapVoxel<F, IDENT>(raw, p.m);
    };
    // the shift: the range's first voxel (every thread reads the same 1-4 bytes)
    float c = PACKED ? 0.f : value(loadRaw(p.s, BPV));
    if (!(fabsf(c) <= FLT_MAX)) c = 0.f;

    AggThread t;
    t.sumD = 0.0; t.sumD2 = 0.0; t.P = 1.0; t.E = 0;
    t.mn = FLT_MAX; t.mx = -FLT_MAX; t.mnIdx = kNone; t.mxIdx = kNone;
    uint32_t rmn = 0xffffffffu, rmx = 0u;      // PACKED: raw extrema; rmx is stored +1 so that 0 means "none"
    bool dead = false;

    auto mulProd4 = [&](float a, float b, float cc, float d) {
        t.P = t.P * (double)a * (double)b * (double)cc * (double)d;
        normProd(t.P, t.E);
        if ((FINITE && t.P == 0.0) || (CONTRACTING && t.E < -400)) dead = true;
    };

    // PACKED: exact integer moments S1 = sum raw, S2 = sum raw^2 (the unmap is raw * 2^-16 for UInt16, raw / 255.999f
    // for UInt8, so sum v = S1 * scale and sum v^2 = S2 * scale^2; the shift c is 0 on this path)
    unsigned long long S1 = 0, S2 = 0;

    // the linear index of a vector is base + i * V; it is only needed on the rare paths that move an extremum
    uint32_t mn2 = 0xffffffffu, mx2 = 0u;      // PACKED: running lane-wise extrema
    uint32_t s1acc = 0;                        // PACKED: 32-bit part of S1, flushed every vector batch
    auto accVec = [&](const uint4& q, unsigned long long base, unsigned long long i) {
        if (PACKED) {
            uint32_t lo, hi, s1 = 0;
            if (F == F_U16) {
                lo = __vminu2(__vminu2(q.x, q.y), __vminu2(q.z, q.w));
                hi = __vmaxu2(__vmaxu2(q.x, q.y), __vmaxu2(q.z, q.w));
                const uint32_t w[4] = { q.x, q.y, q.z, q.w };
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    s1 = __dp2a_lo(w[k], 0x0101u, s1);
                    const uint32_t b = __umulhi(w[k], 65536u), a = w[k] - (b << 16);
                    S2 += (unsigned long long)a * a;
                    S2 += (unsigned long long)b * b;
                }
            } else {
                lo = __vminu4(__vminu4(q.x, q.y), __vminu4(q.z, q.w));
                hi = __vmaxu4(__vmaxu4(q.x, q.y), __vmaxu4(q.z, q.w));
                uint32_t s2 = 0;
                const uint32_t w[4] = { q.x, q.y, q.z, q.w };
#pragma unroll
                for (int k = 0; k < 4; ++k) { s1 = __dp4a(w[k], 0x01010101u, s1); s2 = __dp4a(w[k], w[k], s2); }
                S2 += s2;
            }
            s1acc += s1;                       // <= 4 vectors (2 MiB of voxel sum) between flushes
            // lane-wise running extrema: one packed min / max and one compare per vector; the horizontal reduction
            // and the position bookkeeping run only when a lane moved
            const uint32_t nlo = F == F_U16 ? __vminu2(mn2, lo) : __vminu4(mn2, lo);
            const uint32_t nhi = F == F_U16 ? __vmaxu2(mx2, hi) : __vmaxu4(mx2, hi);
            if (nlo != mn2) {
                mn2 = nlo;
                uint32_t h;
                if (F == F_U16) h = min(nlo & 0xffffu, nlo >> 16);
                else { h = __vminu4(nlo, nlo >> 16); h = min(h & 0xffu, (h >> 8) & 0xffu); }
                if (h < rmn) { rmn = h; t.mnIdx = (base + i * V) | kVecTag; }
            }
            if (nhi != mx2) {
                mx2 = nhi;
                uint32_t h;
                if (F == F_U16) h = max(nhi & 0xffffu, nhi >> 16);
                else { h = __vmaxu4(nhi, nhi >> 16); h = max(h & 0xffu, (h >> 8) & 0xffu); }
                if (h + 1u > rmx) { rmx = h + 1u; t.mxIdx = (base + i * V) | kVecTag; }
            }
            if (!dead) {
#pragma unroll
                for (int k = 0; k < V; k += 4)
                    mulProd4(value(vecGet<BPV>(q, k)), value(vecGet<BPV>(q, k + 1)), value(vecGet<BPV>(q, k + 2)),
                             value(vecGet<BPV>(q, k + 3)));
            }
            return;
        }
        float v[V];
#pragma unroll
        for (int i = 0; i < V; ++i) v[i] = value(vecGet<BPV>(q, i));
        const float mn0 = t.mn, mx0 = t.mx;
#pragma unroll
        for (int i = 0; i < V; ++i) {
            t.mn = v[i] < t.mn ? v[i] : t.mn;
            t.mx = v[i] > t.mx ? v[i] : t.mx;
        }
        if (t.mn < mn0) t.mnIdx = (base + i * V) | kVecTag;
        if (t.mx > mx0) t.mxIdx = (base + i * V) | kVecTag;
        float s[V], sq[V];
#pragma unroll
        for (int i = 0; i < V; ++i) { s[i] = __fsub_rn(v[i], c); sq[i] = __fmul_rn(s[i], s[i]); }
#pragma unroll
        for (int w = V / 2; w >= 1; w >>= 1)
#pragma unroll
            for (int i = 0; i < w; ++i) { s[i] = __fadd_rn(s[i], s[i + w]); sq[i] = __fadd_rn(sq[i], sq[i + w]); }
        t.sumD += (double)s[0];
        t.sumD2 += (double)sq[0];
        if (!dead) {
#pragma unroll
            for (int i = 0; i < V; i += 4) mulProd4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        }
    };
    auto accOne = [&](uint32_t raw, unsigned long long idx) {
        if (PACKED) {
            if (raw < rmn) { rmn = raw; t.mnIdx = idx; }
            if (raw + 1u > rmx) { rmx = raw + 1u; t.mxIdx = idx; }
            S1 += raw;
            S2 += (unsigned long long)raw * raw;
            if (!dead) mulProd4(value(raw), 1.f, 1.f, 1.f);
            return;
        }
        const float v = value(raw);
        if (v < t.mn) { t.mn = v; t.mnIdx = idx; }
        if (v > t.mx) { t.mx = v; t.mxIdx = idx; }
        const float d = __fsub_rn(v, c);
        t.sumD += (double)d;
        t.sumD2 += (double)__fmul_rn(d, d);
        if (!dead) mulProd4(v, 1.f, 1.f, 1.f);
    };

    // address of the voxel with linear range index i
    auto addrOf = [&](unsigned long long i) -> const uint8_t* {
        if (p.flat) return p.s + i * BPV;
        const unsigned long long r = i / (unsigned)p.nx;
        const int x = (int)(i - r * (unsigned)p.nx);
        const long long z = (long long)(r / (unsigned)p.ny), y = (long long)(r - (unsigned long long)z * (unsigned)p.ny);
        return p.s + y * p.pitchY + z * p.pitchZ + (long long)x * BPV;
    };

    if (p.flat) {
        const size_t n = p.n;
        size_t head = ((16u - (reinterpret_cast<uintptr_t>(p.s) & 15u)) & 15u) / BPV;
        if (head > n) head = n;
        const size_t nvec = (n - head) / V, tail0 = head + nvec * V;
        const uint8_t* body = p.s + head * BPV;
        const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
        // head, body, tail in this order: a thread meets its voxels in increasing linear index, so the strict
        // compares keep the first occurrence
        for (size_t k = tid; k < head; k += stride) accOne(loadRaw(p.s + k * BPV, BPV), k);
        size_t i = tid;
        for (; i + 3 * stride < nvec; i += 4 * stride) {
            const uint4 q0 = ldStream(body + i * 16), q1 = ldStream(body + (i + stride) * 16);
            const uint4 q2 = ldStream(body + (i + 2 * stride) * 16), q3 = ldStream(body + (i + 3 * stride) * 16);
            accVec(q0, head, i); accVec(q1, head, i + stride);
            accVec(q2, head, i + 2 * stride); accVec(q3, head, i + 3 * stride);
            if (PACKED) { S1 += s1acc; s1acc = 0; }
        }
        for (; i < nvec; i += stride) { accVec(ldStream(body + i * 16), head, i); if (PACKED) { S1 += s1acc; s1acc = 0; } }
        for (size_t k = tail0 + tid; k < n; k += stride) accOne(loadRaw(p.s + k * BPV, BPV), k);
    } else {
        const long long rows = (long long)p.ny * p.nz;
        for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
            const int z = (int)(r / p.ny), y = (int)(r - (long long)z * p.ny);
            const uint8_t* row = p.s + y * p.pitchY + z * p.pitchZ;
            const unsigned long long rowIdx = (unsigned long long)r * (unsigned)p.nx;
            int head = (int)(((16u - (reinterpret_cast<uintptr_t>(row) & 15u)) & 15u) / BPV);
            if (head > p.nx) head = p.nx;
            const int nvec = (p.nx - head) / V;
            const int tail0 = head + nvec * V;
            const uint8_t* body = row + (long long)head * BPV;
            for (int x = threadIdx.x; x < head; x += blockDim.x)
                accOne(loadRaw(row + (long long)x * BPV, BPV), rowIdx + (unsigned long long)x);
            int v = threadIdx.x;
            for (; v + (int)blockDim.x < nvec; v += 2 * blockDim.x) {
                const uint4 q0 = ldStream(body + v * 16LL), q1 = ldStream(body + (v + (int)blockDim.x) * 16LL);
                accVec(q0, rowIdx + head, (unsigned long long)v);
                accVec(q1, rowIdx + head, (unsigned long long)(v + (int)blockDim.x));
                if (PACKED) { S1 += s1acc; s1acc = 0; }
            }
            for (; v < nvec; v += blockDim.x) {
                accVec(ldStream(body + v * 16LL), rowIdx + head, (unsigned long long)v);
                if (PACKED) { S1 += s1acc; s1acc = 0; }
            }
            for (int x = tail0 + threadIdx.x; x < p.nx; x += blockDim.x)
                accOne(loadRaw(row + (long long)x * BPV, BPV), rowIdx + (unsigned long long)x);
        }
    }

    // exact positions: the first voxel of the recorded vector that holds the thread's extremum
    if (PACKED) {
        if (t.mnIdx != kNone) t.mn = value(rmn);
        if (t.mxIdx != kNone) t.mx = value(rmx - 1u);
        const double scale = F == F_U16 ? 1.52587890625e-05 : 1.0 / (double)255.999f;
        t.sumD = (double)S1 * scale;
        t.sumD2 = (double)S2 * (scale * scale);
    }
    auto resolve = [&](unsigned long long tag, float target, uint32_t rawTarget) -> unsigned long long {
        if (tag == kNone || !(tag & kVecTag)) return tag;
        const unsigned long long idx = tag & ~kVecTag;
        const uint4 q = ldStream(addrOf(idx));
        int found = V - 1;
#pragma unroll
        for (int i = V - 1; i >= 0; --i) {
            const uint32_t raw = vecGet<BPV>(q, i);
            if (PACKED ? raw == rawTarget : value(raw) == target) found = i;
        }
        return idx + (unsigned)found;
    };
    t.mnIdx = resolve(t.mnIdx, t.mn, rmn);
    t.mxIdx = resolve(t.mxIdx, t.mx, rmx - 1u);

    blockReduce(t, red);
    if (threadIdx.x == 0) storePartial(&p.partials[blockIdx.x], t, c);
}

// One CTA folds the CTA partials (all share the shift c of the range's first voxel) in a fixed order.
__global__ void __launch_bounds__(kThreads) aggFinishKernel(AggPartial* partials, unsigned n)
{
    __shared__ AggThread red[kThreads / 32];
    AggThread t;
    t.sumD = 0.0; t.sumD2 = 0.0; t.P = 1.0; t.E = 0;
    t.mn = FLT_MAX; t.mx = -FLT_MAX; t.mnIdx = kNone; t.mxIdx = kNone;
    for (unsigned i = threadIdx.x; i < n; i += blockDim.x) {
        const AggPartial& q = partials[i];
        AggThread o;
        o.sumD = q.sumD; o.sumD2 = q.sumD2; o.P = q.prodMant; o.E = q.prodExp;
        o.mn = q.min; o.mx = q.max; o.mnIdx = q.minIdx; o.mxIdx = q.maxIdx;
        combine(t, o);
    }
    blockReduce(t, red);
    if (threadIdx.x == 0) storePartial(&partials[n], t, partials[0].shift);
}

template <int F>
cudaError_t launchAggF(const LaunchCtx& c, const AggParams& p, unsigned grid)
{
    if (p.m.identity && F != F_INT16) aggregatesKernel<F, true><<<grid, kThreads, 0, c.stream>>>(p);
    else aggregatesKernel<F, false><<<grid, kThreads, 0, c.stream>>>(p);
    countLaunch();
    return cudaGetLastError();
}

} // namespace

size_t aggregatesScratchBytes(const LaunchCtx& c)
{
    return ((size_t)c.smCount * 8 + 1) * sizeof(AggPartial);
}

cudaError_t launchAggregates(const LaunchCtx& c, const VolDesc& vol, const int32_t first[3], const int32_t last[3],
                             AggPartial* scratchDev, AggPartial** totalDev)
{
    AggParams p;
    p.nx = last[0] - first[0]; p.ny = last[1] - first[1]; p.nz = last[2] - first[2];
    if (p.nx <= 0 || p.ny <= 0 || p.nz <= 0) return cudaErrorInvalidValue;
    const int bpv = bytesPerVoxel(vol.format);
    p.pitchY = (long long)vol.dims[0] * bpv; p.pitchZ = p.pitchY * vol.dims[1];
    p.s = vol.data + voxelIndex(vol, first[0], first[1], first[2]) * bpv;
    p.m = vol.map;
    p.n = (size_t)p.nx * p.ny * p.nz;
    Range g; g.nx = p.nx; g.ny = p.ny; g.nz = p.nz; g.x0 = first[0]; g.y0 = first[1]; g.z0 = first[2];
    p.flat = contiguous(vol, g) ? 1 : 0;
    p.partials = scratchDev;
    // persistent grid: up to 8 CTAs per SM, every CTA leaves one partial
    long long grid = (long long)c.smCount * 8;
    const long long rows = (long long)p.ny * p.nz;
    if (p.flat) {
        const long long need = (long long)((p.n * bpv / 16 + kThreads - 1) / kThreads);
        if (grid > need) grid = need < 1 ? 1 : need;
    } else if (grid > rows) grid = rows;
    p.numPartials = (unsigned)grid;
    cudaError_t e;
    switch (vol.format) {
    case F_U8:    e = launchAggF<F_U8>(c, p, (unsigned)grid); break;
    case F_U16:   e = launchAggF<F_U16>(c, p, (unsigned)grid); break;
    case F_INT16: e = launchAggF<F_INT16>(c, p, (unsigned)grid); break;
    case F_U32:   e = launchAggF<F_U32>(c, p, (unsigned)grid); break;
    case F_F32:   e = launchAggF<F_F32>(c, p, (unsigned)grid); break;
    default:      return cudaErrorInvalidValue;
    }
    if (e != cudaSuccess) return e;
    aggFinishKernel<<<1, kThreads, 0, c.stream>>>(scratchDev, (unsigned)grid);
    countLaunch();
    *totalDev = scratchDev + grid;
    return cudaGetLastError();
}

} // namespace vktb
