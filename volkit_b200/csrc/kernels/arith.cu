// Arithmetic family launcher: picks flat / rows / generic kernels (see arith.cuh).
#include <cstdlib>
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
