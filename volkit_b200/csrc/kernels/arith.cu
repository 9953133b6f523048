// Arithmetic family launcher: picks flat / rows / generic kernels (see arith.cuh).
#include <cstdlib>
#include <mutex>
#include <unordered_map>
#include <cstring>
#include "arith.cuh"
#include "launch.hpp"

namespace vktb {

namespace {

// below this the table build (65536 entries per CTA) is not worth it
constexpr size_t kPairTableMinVoxels = (size_t)1 << 24;

struct ArithPlan
{
    ArithParams p;
    bool   flat;       // contiguous in all three operands and co-aligned
    size_t n;          // voxels
    int    fd, fa, fb;
    bool   homogeneous;
    bool   identity;
};

template <int F, int OP, bool IDENT>
cudaError_t launchHomogeneous(const LaunchCtx& c, const ArithPlan& pl)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    if constexpr (F == F_U8) {
        if (pl.flat && pl.n >= kPairTableMinVoxels) {
            // big UInt8 streams: 64 KiB pair table per CTA, 3 CTAs per SM (persistent)
            auto kern = arithPairTableKernel<OP, IDENT>;
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kPairTableBytes);
            if (e != cudaSuccess) return e;
            unsigned grid = flatGrid(pl.n / V + 1, kThreads * 2, c.smCount, 3);
            kern<<<grid, kThreads, kPairTableBytes, c.stream>>>(pl.p.d.base, pl.p.a.base, pl.p.b.base, pl.n, pl.p.md,
                                                                 pl.p.ma, pl.p.mb, pl.p.alias);
            countLaunch();
            return cudaGetLastError();
        }
    }
    if (pl.flat) {

        // 2 vectors/thread/iteration, up to 8 CTAs per SM resident
        unsigned grid = flatGrid(pl.n / V + 1, kThreads * 2, c.smCount, 8);
        arithFlatKernel<F, OP, IDENT><<<grid, kThreads, 0, c.stream>>>(
            pl.p.d.base, pl.p.a.base, pl.p.b.base, pl.n, pl.p.md, pl.p.ma, pl.p.mb, pl.p.alias);
    } else {
        RowsLaunch l = rowsLaunch(pl.p.g.nx, (long long)pl.p.g.ny * pl.p.g.nz, V);
        ArithParams q = pl.p;
        const uintptr_t phase = reinterpret_cast<uintptr_t>(q.d.base) & 15u;
        // every row of an operand in one 16-byte phase; a source phase other than dst's is absorbed by shifted loads
        const bool uniform = ((q.d.pitchY | q.d.pitchZ | q.a.pitchY | q.a.pitchZ | q.b.pitchY | q.b.pitchZ) & 15) == 0
                          && (long long)q.g.ny * q.g.nz < 0x7fffffffLL;
        q.shifted = phase != (reinterpret_cast<uintptr_t>(q.a.base) & 15u)
                 || phase != (reinterpret_cast<uintptr_t>(q.b.base) & 15u);
        if constexpr (F == F_U8) {
            // big UInt8 ranges: one pair-table byte per voxel instead of ~12 arithmetic instructions
            const unsigned long long cpr = (unsigned long long)((q.g.nx + 15) / 16 + 1);
            const unsigned long long items = (unsigned long long)q.g.ny * q.g.nz * cpr;
            if (uniform && pl.n >= kPairTableMinVoxels && items < 0x7fffffffULL) {
                PairRows w;
                w.items = (unsigned)items; w.cpr = (unsigned)cpr;
                magicDivisor((unsigned)cpr, w.mulC, w.shC);
                magicDivisor((unsigned)q.g.ny, w.mulY, w.shY);
                auto kern = arithRowsPairTableKernel<OP, IDENT>;
                cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kPairTableBytes);
                if (e != cudaSuccess) return e;
                kern<<<(unsigned)c.smCount * 3u, kThreads, kPairTableBytes, c.stream>>>(q, (int)phase, w);
                countLaunch();
                return cudaGetLastError();
            }
        }
        // neighbour-lane exchange needs whole warps inside one row and sources that dst does not overlap
        static const bool exchange = [] { const char* e = getenv("VKTB_ARITH_EXCHANGE"); return !(e && e[0] == '0'); }();
        if (uniform && q.shifted && !q.alias && l.block.x >= 32 && exchange) {
            const long long S = (long long)l.grid.y * l.block.y, sz = S / q.g.ny, sy = S - sz * q.g.ny;
            const Operand* ops[3] = { &q.d, &q.a, &q.b };
            for (int i = 0; i < 3; ++i) {
                q.step[i] = sy * ops[i]->pitchY + sz * ops[i]->pitchZ;
                q.stepWrap[i] = q.step[i] + ops[i]->pitchZ - q.g.ny * ops[i]->pitchY;
            }
            q.sy = (int)sy;
            arithRowsShiftedKernel<F, OP, IDENT><<<l.grid, l.block, 0, c.stream>>>(q, (int)(phase / bytesPerVoxel(F)));
        } else if (uniform)
            (q.shifted ? arithRowsUniformKernel<F, OP, IDENT, true> : arithRowsUniformKernel<F, OP, IDENT, false>)
                <<<l.grid, l.block, 0, c.stream>>>(q, (int)(phase / bytesPerVoxel(F)));
        else
            arithRowsKernel<F, OP, IDENT><<<l.grid, l.block, 0, c.stream>>>(pl.p);
    }
    countLaunch();
    return cudaGetLastError();
}

template <int OP>
cudaError_t launchGeneric(const LaunchCtx& c, const ArithPlan& pl)
{
    long long rows = (long long)pl.p.g.ny * pl.p.g.nz;
    dim3 grid((unsigned)((pl.p.g.nx + kThreads - 1) / kThreads), (unsigned)(rows > 65535 ? 65535 : rows), 1);
    if (grid.x > 1024) grid.x = 1024;
    arithGenericKernel<OP><<<grid, kThreads, 0, c.stream>>>(pl.p, pl.fd, pl.fa, pl.fb);
    countLaunch();
    return cudaGetLastError();
}

template <int OP>
cudaError_t dispatchFormat(const LaunchCtx& c, const ArithPlan& pl)
{
    if (pl.homogeneous) {
        switch (pl.fd) {
        case F_U8:  return pl.identity ? launchHomogeneous<F_U8, OP, true>(c, pl)  : launchHomogeneous<F_U8, OP, false>(c, pl);
        case F_U16: return pl.identity ? launchHomogeneous<F_U16, OP, true>(c, pl) : launchHomogeneous<F_U16, OP, false>(c, pl);
        case F_U32: return pl.identity ? launchHomogeneous<F_U32, OP, true>(c, pl) : launchHomogeneous<F_U32, OP, false>(c, pl);
        case F_F32: return pl.identity ? launchHomogeneous<F_F32, OP, true>(c, pl) : launchHomogeneous<F_F32, OP, false>(c, pl);
        default: break;
        }
    }
    return launchGeneric<OP>(c, pl);
}

} // namespace

// ---- proof of divProven for one divisor ------------------------------------------------------------------------------
// Every binary32 dividend (2^32 bit patterns): divProven(x, d, RN(1/d)) against __fdiv_rn(x, d), bit for bit (NaN
// results only have to be NaN on both sides, as everywhere in this library).  About 3 ms on a B200, once per divisor
// and process; the verdict is cached.  VKTB_NO_PROVEN_DIV=1 keeps the IEEE division (A/B runs).
__global__ void __launch_bounds__(256)
divProofKernel(float d, float y, unsigned long long* bad)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned mism = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < (1ull << 32); i += stride) {
        const float x = __uint_as_float((uint32_t)i);
        const float a = divProven(x, d, y), b = __fdiv_rn(x, d);
        const bool same = __float_as_uint(a) == __float_as_uint(b) || (a != a && b != b);
        mism += same ? 0u : 1u;
    }
    if (mism) atomicAdd(bad, (unsigned long long)mism);
}

static void proveDivisor(const LaunchCtx& c, Mapping& m)
{
    static std::mutex mu;
    static std::unordered_map<uint32_t, int> proven;     // range bits -> 1 proven, 0 refuted / not provable
    if (m.identity || m.exactRecip != 0 || m.invRange == 0.f) return;
    static const bool off = getenv("VKTB_NO_PROVEN_DIV") != nullptr;
    if (off) return;
    uint32_t bits;
    memcpy(&bits, &m.range, 4);
    {
        std::lock_guard<std::mutex> l(mu);
        auto it = proven.find(bits);
        if (it != proven.end()) { if (it->second) m.exactRecip = 2; return; }
    }
    int ok = 0;
    unsigned long long* bad = nullptr;
    if (cudaMallocAsync((void**)&bad, sizeof(*bad), c.stream) == cudaSuccess) {
        unsigned long long h = ~0ull;
        cudaMemsetAsync(bad, 0, sizeof(*bad), c.stream);
        divProofKernel<<<(unsigned)c.smCount * 8u, 256, 0, c.stream>>>(m.range, m.invRange, bad);
        countLaunch();
        if (cudaMemcpyAsync(&h, bad, sizeof(h), cudaMemcpyDeviceToHost, c.stream) == cudaSuccess
            && cudaStreamSynchronize(c.stream) == cudaSuccess)
            ok = h == 0ull ? 1 : 0;
        else
            cudaGetLastError();
        cudaFreeAsync(bad, c.stream);
    }
    std::lock_guard<std::mutex> l(mu);
    proven[bits] = ok;
    if (ok) m.exactRecip = 2;
}

// 1 / 0: whether x / (hi - lo) runs in the proven reciprocal form for this mapping (tests, notes)
int divisorProven(const LaunchCtx& c, float lo, float hi)
{
    Mapping m = makeMapping(lo, hi);
    proveDivisor(c, m);
    return m.exactRecip == 2 ? 1 : 0;
}

cudaError_t launchArith(const LaunchCtx& c, int op, const VolDesc& dst, const VolDesc& s1, const VolDesc& s2,
                        const int32_t first[3], const int32_t last[3], const int32_t off[3])
{
    ArithPlan pl;
    Range& g = pl.p.g;
    g.nx = last[0] - first[0]; g.ny = last[1] - first[1]; g.nz = last[2] - first[2];
    g.x0 = first[0]; g.y0 = first[1]; g.z0 = first[2];
    if (g.nx <= 0 || g.ny <= 0 || g.nz <= 0) return cudaSuccess;

    pl.p.d = operandAt(dst, (long long)first[0] + off[0], (long long)first[1] + off[1], (long long)first[2] + off[2]);
    pl.p.a = operandAt(s1, first[0], first[1], first[2]);
    pl.p.b = operandAt(s2, first[0], first[1], first[2]);
    pl.p.md = dst.map; pl.p.ma = s1.map; pl.p.mb = s2.map;
    pl.p.alias = (overlaps(dst, s1) || overlaps(dst, s2)) ? 1 : 0;
    pl.p.aEnd = s1.data + (size_t)s1.dims[0] * s1.dims[1] * s1.dims[2] * bytesPerVoxel(s1.format);
    pl.p.bEnd = s2.data + (size_t)s2.dims[0] * s2.dims[1] * s2.dims[2] * bytesPerVoxel(s2.format);
    pl.p.shifted = 0;
    pl.fd = dst.format; pl.fa = s1.format; pl.fb = s2.format;
    pl.homogeneous = pl.fd == pl.fa && pl.fd == pl.fb;
    pl.identity = dst.map.identity && s1.map.identity && s2.map.identity;
    pl.n = (size_t)g.nx * g.ny * g.nz;
    const uintptr_t ad = reinterpret_cast<uintptr_t>(pl.p.d.base) & 15u;
    pl.flat = pl.homogeneous && contiguous(dst, g) && contiguous(s1, g) && contiguous(s2, g)
           && ad == (reinterpret_cast<uintptr_t>(pl.p.a.base) & 15u)
           && ad == (reinterpret_cast<uintptr_t>(pl.p.b.base) & 15u);
    // the one division per voxel is map_dst's (v - lo) / (hi - lo): the flat kernel has a loop for proven divisors
    if (pl.flat && !pl.identity && pl.n >= ((size_t)1 << 22)) proveDivisor(c, pl.p.md);

    switch (op) {
    case 0: return dispatchFormat<0>(c, pl);
    case 1: return dispatchFormat<1>(c, pl);
    case 2: return dispatchFormat<2>(c, pl);
    case 3: return dispatchFormat<3>(c, pl);
    case 4: return dispatchFormat<4>(c, pl);
    case 5: return dispatchFormat<5>(c, pl);
    case 6: return dispatchFormat<6>(c, pl);
    case 7: return dispatchFormat<7>(c, pl);
    case 8: return dispatchFormat<8>(c, pl);
    case 9: return dispatchFormat<9>(c, pl);
    }
    return cudaErrorInvalidValue;
}

} // namespace vktb
