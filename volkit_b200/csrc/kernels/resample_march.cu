// Resample, 8-tap modes, as a z-marching brick kernel with TMA 3-D tile staging.
//
//   MODE_LINEAR8    reference-faithful FilterMode::Linear for sources whose taps cannot be proven finite (Float32):
//                   StructuredVolumeView::sampleLinear, reference src/vkt/StructuredVolumeView.hpp:80-119, called with
//                   truncated coordinates (src/vkt/Resample_serial.hpp:55-69): all 8 taps run through the lerp chain with
//                   fraction 0, hi.x is NOT clamped, lo.z is NOT clamped (the reference's typo'd clamps).
//   MODE_TRILINEAR  the opt-in mathematically correct filter (float coordinates, six clamps); not what the reference
//                   computes.
//
// One CTA owns a dst (x,y) tile of 64 32-bit words x BY rows and marches through the dst planes.  The src planes it
// needs (lo/hi plane of every dst plane, monotone in z) are streamed through a ring of kRing shared-memory slots; each
// slot holds the src brick [xOrigin, +bxS) x [yOrigin, +byS) of one plane, fetched with cp.async.bulk.tensor.3d
// (SASS UTMALDG) against a CUtensorMap of the src window (out-of-bounds elements arrive as zeros and are never tapped).
// Warp 0 issues the loads kRing-2 planes ahead; full[]/empty[] mbarriers hand slots between the TMA unit and the eight
// consumer warps, so there is no CTA-wide barrier in the march.  Every src plane tile is read from L2/HBM once per CTA
// and re-used for all dst planes that tap it (~1.5 planes at 2:3 up-sampling, both z taps at 2:1 down-sampling).
//
// Thread mapping: lane-consecutive dst words along x (conflict-free or 2-way shared-memory taps for ratios <= 2, one
// full 128-byte line per warp store), four rows per pass.  Per-thread tap offsets and x fractions live in registers,
// the tile's row offsets / y fractions in a small shared table.
//
// Algorithmic traffic: Sx*Sy*Sz*bpv_src (every src voxel is tapped when the ratio is <= 2) + Dx*Dy*Dz*bpv_dst.
#include <cstdlib>

#include <cuda.h>

#include "common.cuh"
#include "copy_boxes.hpp"
#include "launch.hpp"
#include "tma.cuh"
#include "voxel_codec.cuh"

namespace vktb {

namespace {

constexpr int kMarchThreads = 256;
constexpr int kLanesX = 64;     // threads along x: one 32-bit dst word each
constexpr int kRowsY = kMarchThreads / kLanesX;
constexpr int kMaxBY = 32;      // dst rows per tile (8 passes of kRowsY)
constexpr int kRing = 4;        // src plane slots
constexpr int kMaxBox = 256;    // TMA box extent limit per dimension
constexpr int kMaxBoxesX = 4;
constexpr int kHdr = 4;

enum { M_LINEAR8 = 1, M_TRILINEAR = 2 };

struct AxisEntry { int lo, hi; float fr; };

// lo / hi tap and fraction of dst index d on one axis (axis: 0 x, 1 y, 2 z)
__device__ __forceinline__ AxisEntry axisEntry(int d, int D, int S, int mode, int axis)
{
    const float s = __fmul_rn(__fdiv_rn(__int2float_rn(d), __int2float_rn(D)), __int2float_rn(S));
    const int l = (int)x86Cvtt32(s);
    AxisEntry e;
    if (mode == M_TRILINEAR) {
        e.lo = min(max(l, 0), S - 1);
        e.hi = min(max((int)x86Cvtt32(__fadd_rn(s, 1.f)), 0), S - 1);
        e.fr = __fsub_rn(s, __int2float_rn(e.lo));
    } else {
        const int cl = min(max(l, 0), S - 1);
        e.lo = axis == 2 ? l : cl;                                   // lo.z is never clamped
        e.hi = axis == 0 ? l + 1 : min(max(l + 1, 0), S - 1);        // hi.x is never clamped
        e.fr = __fsub_rn(__int2float_rn(l), __int2float_rn(axis == 2 ? l : cl));
    }
    return e;
}

// Table layout (ints; fractions bit-cast): hdr[4] = {nSeq}; x: lo[Dx] hi[Dx] fr[Dx]; y: lo[Dy] hi[Dy] fr[Dy];
// z: lo[dz] hi[dz] fr[dz] seqLo[dz] seqHi[dz]; planeOf[szCount]; mark[szCount]; rank[szCount];
// zpack[dz] = int4 {seqLo, seqHi, fr, -} at zpackOff (16-byte aligned)
__global__ void __launch_bounds__(1024)
marchPlanKernel(int* tab, int Dx, int Sx, int Dy, int Sy, int dzCount, int dz0, int Dz, int Sz, int srcZ0, int szCount, int mode,
                int zpackOff)
{
    __shared__ int scan[1024];
    const int t = threadIdx.x;
    int* X = tab + kHdr;
    int* Y = X + 3 * Dx;
    int* Z = Y + 3 * Dy;
    if (blockIdx.x < 2) {
        const int n = blockIdx.x == 0 ? Dx : Dy, S = blockIdx.x == 0 ? Sx : Sy;
        int* A = blockIdx.x == 0 ? X : Y;
        for (int i = t; i < n; i += 1024) {
            const AxisEntry e = axisEntry(i, n, S, mode, blockIdx.x);
            A[i] = e.lo; A[n + i] = e.hi; A[2 * n + i] = __float_as_int(e.fr);
        }
        return;
    }
    int* seqLo = Z + 3 * dzCount;
    int* seqHi = seqLo + dzCount;
    int* planeOf = seqHi + dzCount;
    int* mark = planeOf + szCount;
    int* rank = mark + szCount;
    for (int i = t; i < szCount; i += 1024) mark[i] = 0;
    __syncthreads();
    for (int i = t; i < dzCount; i += 1024) {
        const AxisEntry e = axisEntry(dz0 + i, Dz, Sz, mode, 2);
        Z[i] = e.lo; Z[dzCount + i] = e.hi; Z[2 * dzCount + i] = __float_as_int(e.fr);
        const int a = e.lo - srcZ0, b = e.hi - srcZ0;
        if (a >= 0 && a < szCount) mark[a] = 1;
        if (b >= 0 && b < szCount) mark[b] = 1;
    }
    __syncthreads();
    const int seg = (szCount + 1023) / 1024;
    const int i0 = min(szCount, t * seg), i1 = min(szCount, i0 + seg);
    int cnt = 0;
    for (int i = i0; i < i1; ++i) cnt += mark[i];
    scan[t] = cnt;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = t >= o ? scan[t - o] : 0;
        __syncthreads();
        scan[t] += v;
        __syncthreads();
    }
    int k = scan[t] - cnt;
    for (int i = i0; i < i1; ++i) {
        rank[i] = k;
        if (mark[i]) { planeOf[k] = i; ++k; }
    }
    if (t == 1023) tab[0] = scan[1023];
    __syncthreads();
    for (int i = t; i < dzCount; i += 1024) {
        const int a = min(max(Z[i] - srcZ0, 0), szCount - 1), b = min(max(Z[dzCount + i] - srcZ0, 0), szCount - 1);
        seqLo[i] = rank[a];
        seqHi[i] = rank[b];
        // everything the march needs per dst plane in one 16-byte record
        reinterpret_cast<int4*>(tab + zpackOff)[i] = make_int4(rank[a], rank[b], Z[2 * dzCount + i], 0);
    }
}

struct MarchParams
{
    uint8_t*       d;       // dst window base
    const uint8_t* s;       // src window base (spill tap only)
    int Dx, Dy, dzCount;
    int Sx, Sy, szCount;
    int BXv, BY, tilesX;    // dst tile: BXv voxels (= 64 words) x BY rows
    int bxS, byS, nBoxX;    // src brick: nBoxX boxes of bxS x byS elements per plane
    int slotBytes;
    int spillOff;           // byte offset of the spill column (byS rows x 16 bytes) inside a slot
    int zpackOff;           // int offset of the int4 per-plane records in tab
    const int* tab;
    const unsigned* gate;   // when set: the kernel runs only if *gate != 0 (see launchResample: Float32 Linear)
    Mapping md, ms;
};

template <int B> struct ElemT;
template <> struct ElemT<1> { typedef uint8_t type; };
template <> struct ElemT<2> { typedef uint16_t type; };
template <> struct ElemT<4> { typedef uint32_t type; };

// dst voxels per thread and row: one 64-bit store for 2- and 4-byte voxels, one 32-bit store for bytes
__host__ __device__ constexpr int marchVoxelsPerThread(int bd) { return bd == 1 ? 4 : 8 / bd; }
// row passes per thread (tile height = passes * kRowsY): bounded so that both register planes (2 * passes * voxels
// floats) leave room for two CTAs per SM
__host__ __device__ constexpr int marchPasses(int bd) { return bd == 4 ? kMaxBY / kRowsY : kMaxBY / kRowsY / 2; }

template <int BYTES> __device__ __forceinline__ void marchStore(uint8_t* p, uint32_t w0, uint32_t w1);
template <> __device__ __forceinline__ void marchStore<4>(uint8_t* p, uint32_t w0, uint32_t)
{
    asm volatile("st.global.L1::no_allocate.u32 [%0], %1;" :: "l"(p), "r"(w0));
}
template <> __device__ __forceinline__ void marchStore<8>(uint8_t* p, uint32_t w0, uint32_t w1)
{
    asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" :: "l"(p), "r"(w0), "r"(w1));
}

template <int FS, int FD, int MODE, bool IDENT>
__global__ void __launch_bounds__(kMarchThreads)
resampleMarchKernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmapSpill, MarchParams p)
{
    if (p.gate && *p.gate == 0u) return;     // uniform: the one-tap result already in dst is exact (launchResample)
    constexpr int BS = bytesPerVoxel(FS), BD = bytesPerVoxel(FD);
    constexpr int VPT = marchVoxelsPerThread(BD);
    constexpr int SB = VPT * BD;             // bytes per store
    constexpr int NPASS = marchPasses(BD);
    constexpr int NP = NPASS * VPT;
    typedef typename ElemT<BS>::type SrcT;
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[kRing], empty[kRing];
    __shared__ int4 rowInfo[kMaxBY];         // {lo row byte offset in the brick, hi row byte offset, y fraction, lo | hi << 16 brick rows}

    const int t = threadIdx.x, lx = t % kLanesX, ty = t / kLanesX;
    const int* X = p.tab + kHdr;
    const int* Y = X + 3 * p.Dx;
    const int* Z = Y + 3 * p.Dy;
    const int* seqLoT = Z + 3 * p.dzCount;
    const int* seqHiT = seqLoT + p.dzCount;
    const int* planeOf = seqHiT + p.dzCount;
    const int4* zpack = reinterpret_cast<const int4*>(p.tab + p.zpackOff);
    const int nSeq = p.tab[0];

    // the tiles holding the last columns run an extra spill pass in reference-faithful Linear: they are the slowest,
    // so they take the first block indices (all of them start in the first wave) instead of forming the launch's tail
    const int tilesY = gridDim.x / p.tilesX;
    int tileX, tileY;
    if ((int)blockIdx.x < tilesY) { tileX = p.tilesX - 1; tileY = blockIdx.x; }
    else { const int b = blockIdx.x - tilesY; tileX = b % (p.tilesX - 1); tileY = b / (p.tilesX - 1); }
    const int x0 = tileX * p.BXv, y0 = tileY * p.BY;
    const int h = min(p.BY, p.Dy - y0);
    // the innermost TMA coordinate must address a 16-byte aligned element (found the hard way: misaligned box
    // origins raise an illegal-instruction fault)
    const int xOrigin = X[x0] & ~(16 / BS - 1), yOrigin = Y[y0];
    const int boxElems = p.bxS * p.byS;

    // this thread's dst voxels: x0 + lx*VPT + i in rows y0 + ty + pass*kRowsY
    const int xw = x0 + lx * VPT;
    const bool active = xw < p.Dx;           // Dx is a multiple of VPT (launcher)
    const int nPass = active ? max(0, (h - ty + kRowsY - 1) / kRowsY) : 0;
    int oLo[VPT], oHi[VPT];                  // byte offsets of the x taps inside a brick row
    float fx[VPT], gx[VPT];                  // x fraction and 1 - fraction
    bool spill[VPT];
    bool anySpill = false;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        oLo[i] = oHi[i] = 0; fx[i] = 0.f; spill[i] = false;
        if (active) {
            const int x = xw + i;
            const int a = X[x] - xOrigin, b = X[p.Dx + x] - xOrigin;
            oLo[i] = ((a / p.bxS) * boxElems + a % p.bxS) * BS;
            spill[i] = MODE == M_LINEAR8 && X[p.Dx + x] >= p.Sx;   // the unclamped +1 tap leaves the row
            oHi[i] = spill[i] ? oLo[i] : ((b / p.bxS) * boxElems + b % p.bxS) * BS;
            fx[i] = __int_as_float(X[2 * p.Dx + x]);
            anySpill |= spill[i];
        }
        gx[i] = __fsub_rn(1.0f, fx[i]);
    }
    for (int r = t; r < h; r += kMarchThreads) {
        const int y = y0 + r;
        rowInfo[r] = make_int4((Y[y] - yOrigin) * p.bxS * BS, (Y[p.Dy + y] - yOrigin) * p.bxS * BS, Y[2 * p.Dy + y],
                               (Y[y] - yOrigin) | ((Y[p.Dy + y] - yOrigin) << 16));
    }
    if (t == 0) {
        for (int i = 0; i < kRing; ++i) { mbarInit(smemAddr(&full[i]), 1); mbarInit(smemAddr(&empty[i]), kMarchThreads / 32); }
        fenceBarrierInit();
    }
    __syncthreads();

    const long long dstRowPitch = (long long)p.Dx * BD;
    const long long dstPlanePitch = dstRowPitch * p.Dy;
    const long long rowStep = dstRowPitch * kRowsY;
    // the reference's unclamped hi.x tap of the last column reads linear index + 1: column 0 of the NEXT row (the next
    // plane's first row after the last row; nothing, i.e. a zero voxel, past the end of the window).  The tile that
    // holds those columns stages that column with a second, 2-D tensor map over the window flattened to rows
    // (16 bytes x byS rows starting one row below the brick; out-of-bounds rows arrive as zeros).
    const bool tileSpill = MODE == M_LINEAR8 && X[p.Dx + min(p.Dx, x0 + p.BXv) - 1] >= p.Sx;
    const uint32_t slotTx = (uint32_t)(p.nBoxX * boxElems * BS) + (tileSpill ? (uint32_t)p.byS * 16u : 0u);
    const Mapping ms = p.ms, md = p.md;

    auto tapF = [&](const uint8_t* row, int off) -> float {
        return unmapVoxel<FS, IDENT>((uint32_t)*reinterpret_cast<const SrcT*>(row + off), ms);
    };
    // the reference's hi.x tap at the last column reads linear index + 1, i.e. the first voxel of the next row /
    // plane, or nothing (zero voxel) past the end of the buffer
    // lerp with the complement precomputed: (1-x)*a + x*b, each operation rounded on its own (refLerp)
    auto lerp = [](float a, float b, float x, float g) { return __fadd_rn(__fmul_rn(g, a), __fmul_rn(x, b)); };

    // ---- the march ----------------------------------------------------------------------------------------------
    // The reference nests its lerps x -> y -> z (StructuredVolumeView.hpp:112-118), so the chain factors exactly:
    //     P(s, y, x) = lerp_y(lerp_x(v[lx,ly,s], v[hx,ly,s]), lerp_x(v[lx,hy,s], v[hx,hy,s]))     per src plane s
    //     out(z, y, x) = map(lerp_z(P(lz), P(hz)))
    // with every intermediate a binary32 value either way.  A thread owns the same (x, y) positions in every dst
    // plane, so the two P planes it needs stay in its REGISTERS: each src plane goes through the 4-tap x/y stage
    // once per tile (not once per dst plane that taps it), the slot is handed back to the TMA unit right after, and a
    // dst plane costs one lerp, one map and a share of one store per voxel.
    float Plo[NP], Phi[NP];
#pragma unroll
    for (int i = 0; i < NP; ++i) Plo[i] = Phi[i] = 0.f;

    auto stageXY = [&](float (&P)[NP], int k) {
        const int slotId = k % kRing;
        mbarWait(smemAddr(&full[slotId]), (uint32_t)((k / kRing) & 1));
        const uint8_t* slot = smem + slotId * p.slotBytes;
        // branch-free hot path: a spilled hi.x tap is read as its lo tap here and patched below
#pragma unroll
        for (int pass = 0; pass < NPASS; ++pass) {
            if (pass < nPass) {
                const int4 ri = rowInfo[ty + pass * kRowsY];
                const uint8_t* rowLo = slot + ri.x;
                const uint8_t* rowHi = slot + ri.y;
                const float fy = __int_as_float(ri.z), gy = __fsub_rn(1.0f, fy);
#pragma unroll
                for (int i = 0; i < VPT; ++i) {
                    const float v0 = tapF(rowLo, oLo[i]), v1 = tapF(rowLo, oHi[i]);
                    const float v2 = tapF(rowHi, oLo[i]), v3 = tapF(rowHi, oHi[i]);
                    P[pass * VPT + i] = lerp(lerp(v0, v1, fx[i], gx[i]), lerp(v2, v3, fx[i], gx[i]), fy, gy);
                }
            }
        }
        if (anySpill) {   // only the lanes holding the last columns, only reference-faithful Linear
            const uint8_t* col = slot + p.spillOff;
#pragma unroll
            for (int pass = 0; pass < NPASS; ++pass) {
                if (pass < nPass) {
                    const int4 ri = rowInfo[ty + pass * kRowsY];
                    const float fy = __int_as_float(ri.z), gy = __fsub_rn(1.0f, fy);
                    const float v1 = tapF(col, (ri.w & 0xffff) * 16), v3 = tapF(col, (ri.w >> 16) * 16);
#pragma unroll
                    for (int i = 0; i < VPT; ++i) {
                        if (spill[i]) {
                            const float v0 = tapF(slot + ri.x, oLo[i]), v2 = tapF(slot + ri.y, oLo[i]);
                            P[pass * VPT + i] = lerp(lerp(v0, v1, fx[i], gx[i]), lerp(v2, v3, fx[i], gx[i]), fy, gy);
                        }
                    }
                }
            }
        }
        // this warp is done with the slot
        __syncwarp();
        if ((t & 31) == 0) mbarArrive(smemAddr(&empty[slotId]));
    };

    int issued = 0, curLo = -1, curHi = -1, prevLo = -1, prevHi = -1;
    uint32_t outWord[NPASS * (SB / 4)];
#pragma unroll
    for (int i = 0; i < NPASS * (SB / 4); ++i) outWord[i] = 0u;
    uint8_t* dz = p.d + (long long)(y0 + ty) * dstRowPitch + (long long)xw * BD;   // this thread's first row, plane zi

    int nextPlane = (t == 0 && nSeq > 0) ? planeOf[0] : 0;
    int4 zNext = zpack[0];       // per-plane record, fetched one plane ahead (an L2 round trip otherwise stalls every plane)
    for (int zi = 0; zi < p.dzCount; ++zi, dz += dstPlanePitch) {
        const int4 zc = zNext;
        if (zi + 1 < p.dzCount) zNext = zpack[zi + 1];
        const int sLo = zc.x, sHi = zc.y;
        if (t < 32) {
            // warp 0 keeps the TMA unit kRing planes ahead of its own position; the slots it re-arms belong to planes
            // below sLo, which this warp has staged already, so the wait below only ever waits for OTHER warps
            const int want = min(nSeq, sLo + kRing);
            while (issued < want) {
                const int slot = issued % kRing;
                if (issued >= kRing) mbarWait(smemAddr(&empty[slot]), (uint32_t)(((issued / kRing) & 1) ^ 1));
                if (t == 0) {
                    const uint32_t bar = smemAddr(&full[slot]);
                    mbarArriveExpectTx(bar, slotTx);
                    const int plane = nextPlane;
                    if (issued + 1 < nSeq) nextPlane = planeOf[issued + 1];   // fetched before it is needed
                    for (int kb = 0; kb < p.nBoxX; ++kb)
                        tensorLoad3d(smemAddr(smem + slot * p.slotBytes + kb * boxElems * BS), &tmap,
                                     xOrigin + kb * p.bxS, yOrigin, plane, bar);
                    if (tileSpill)
                        tensorLoad2d(smemAddr(smem + slot * p.slotBytes + p.spillOff), &tmapSpill, 0,
                                     plane * p.Sy + yOrigin + 1, bar);
                }
                __syncwarp();
                ++issued;
            }
        }
        // bring the register planes to (sLo, sHi): the sequence is monotone, so a plane is staged exactly once and in
        // sequence order (lo first), or inherited from the other register set
        if (sLo != curLo) {
            if (sLo == curHi) {
#pragma unroll
                for (int i = 0; i < NP; ++i) Plo[i] = Phi[i];
            } else {
                stageXY(Plo, sLo);
            }
            curLo = sLo;
        }
        if (sHi != curHi) {
            if (sHi == curLo) {
#pragma unroll
                for (int i = 0; i < NP; ++i) Phi[i] = Plo[i];
            } else {
                stageXY(Phi, sHi);
            }
            curHi = sHi;
        }

        // reference-faithful Linear: fractions are all zero, so dst planes tapping the same src planes are identical
        const bool reuse = MODE == M_LINEAR8 && sLo == prevLo && sHi == prevHi;
        prevLo = sLo; prevHi = sHi;
        if (!reuse) {
            const float fz = __int_as_float(zc.z), gz = __fsub_rn(1.0f, fz);
#pragma unroll
            for (int pass = 0; pass < NPASS; ++pass) {
                uint32_t w[2] = { 0u, 0u };
#pragma unroll
                for (int i = 0; i < VPT; ++i) {
                    const uint32_t e = mapVoxel<FD, IDENT>(lerp(Plo[pass * VPT + i], Phi[pass * VPT + i], fz, gz), md);
                    if (BD == 4) w[i] = e;
                    else if (BD == 2) w[i >> 1] |= (e & 0xffffu) << (16 * (i & 1));
                    else w[0] |= (e & 0xffu) << (8 * i);
                }
                outWord[pass * (SB / 4)] = w[0];
                if (SB == 8) outWord[pass * (SB / 4) + SB / 8] = w[1];
            }
        }
        uint8_t* q = dz;
#pragma unroll
        for (int pass = 0; pass < NPASS; ++pass) {
            if (pass < nPass) marchStore<SB>(q, outWord[pass * (SB / 4)], outWord[pass * (SB / 4) + SB / 8]);
            q += rowStep;
        }
    }
}

template <int FS, int FD, int MODE, bool IDENT>
cudaError_t launchMarch(const LaunchCtx& c, const CUtensorMap* tm, const MarchParams& p, size_t smemBytes, unsigned grid)
{
    auto kern = resampleMarchKernel<FS, FD, MODE, IDENT>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
    if (e != cudaSuccess) return e;
    kern<<<grid, kMarchThreads, smemBytes, c.stream>>>(tm[0], tm[1], p);
    countLaunch();
    return cudaGetLastError();
}

template <int FS, int FD>
cudaError_t launchMarchMode(const LaunchCtx& c, const CUtensorMap* tm, const MarchParams& p, size_t smemBytes, unsigned grid, int mode)
{
    // default (0,1) mappings on both sides drop the mapping arithmetic from the codec
    if (p.ms.identity && p.md.identity)
        return mode == M_LINEAR8 ? launchMarch<FS, FD, M_LINEAR8, true>(c, tm, p, smemBytes, grid)
                                 : launchMarch<FS, FD, M_TRILINEAR, true>(c, tm, p, smemBytes, grid);
    return mode == M_LINEAR8 ? launchMarch<FS, FD, M_LINEAR8, false>(c, tm, p, smemBytes, grid)
                             : launchMarch<FS, FD, M_TRILINEAR, false>(c, tm, p, smemBytes, grid);
}

template <int FS>
cudaError_t launchMarchDst(const LaunchCtx& c, const CUtensorMap* tm, const MarchParams& p, size_t smemBytes, unsigned grid,
                           int fd, int mode)
{
    switch (fd) {
    case F_U8:  return launchMarchMode<FS, F_U8>(c, tm, p, smemBytes, grid, mode);
    case F_U16: return launchMarchMode<FS, F_U16>(c, tm, p, smemBytes, grid, mode);
    case F_U32: return launchMarchMode<FS, F_U32>(c, tm, p, smemBytes, grid, mode);
    case F_F32: return launchMarchMode<FS, F_F32>(c, tm, p, smemBytes, grid, mode);
    }
    return cudaErrorNotSupported;
}

} // namespace

// libcuda is never linked: the driver entry point is fetched through the runtime
EncodeTiledFn encodeTiled()
{
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        else
            cudaGetLastError();
    }
    return fn;
}

// 8-tap resample (mode 1 = reference-faithful Linear, 2 = true trilinear) through the TMA z-march.  Returns
// cudaErrorNotSupported when the operands do not qualify; the caller then uses the direct-gather kernel.
cudaError_t launchResampleMarch(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src, int mode,
                                int dstZ0, const int32_t D[3], int srcZ0, const int32_t S[3], const unsigned* gate)
{
    const int bs = bytesPerVoxel(src.format), bd = bytesPerVoxel(dst.format);
    const int dzCount = dst.dims[2], szCount = src.dims[2];
    auto plain = [](int f) { return f == F_U8 || f == F_U16 || f == F_U32 || f == F_F32; };
    if (!plain(src.format) || !plain(dst.format)) return cudaErrorNotSupported;
    if (getenv("VKTB_RS_NO_MARCH")) return cudaErrorNotSupported;
    // TMA: 16-byte aligned base and strides; dst rows are written as 32- or 64-bit words
    if ((reinterpret_cast<uintptr_t>(src.data) & 15u) || ((long long)S[0] * bs) % 16) return cudaErrorNotSupported;
    const int storeBytes = marchVoxelsPerThread(bd) * bd;
    if ((reinterpret_cast<uintptr_t>(dst.data) % storeBytes) || ((long long)D[0] * bd) % storeBytes) return cudaErrorNotSupported;
    for (int a = 0; a < 3; ++a) if (D[a] > (1 << 24) || S[a] > (1 << 24)) return cudaErrorNotSupported;
    EncodeTiledFn encode = encodeTiled();
    if (!encode) return cudaErrorNotSupported;

    MarchParams p;
    p.d = dst.data; p.s = src.data;
    p.Dx = D[0]; p.Dy = D[1]; p.dzCount = dzCount;
    p.Sx = S[0]; p.Sy = S[1]; p.szCount = szCount;
    p.md = dst.map; p.ms = src.map;
    p.gate = gate;
    p.BXv = kLanesX * marchVoxelsPerThread(bd);
    p.tilesX = (D[0] + p.BXv - 1) / p.BXv;
    // src brick extents: floor(tile * S / D) + 3 covers lo .. hi of the tile (hi may be lo + 2 after rounding)
    const int gran = 16 / bs;
    const long long spanX = (long long)p.BXv * S[0] / D[0] + 3 + (gran - 1);   // + the origin's round-down to 16 bytes
    if (spanX <= kMaxBox) { p.nBoxX = 1; p.bxS = (int)((spanX + gran - 1) / gran * gran); if (p.bxS > kMaxBox) p.bxS = kMaxBox; }
    else { p.nBoxX = (int)((spanX + kMaxBox - 1) / kMaxBox); p.bxS = kMaxBox; }   // whole boxes: 128-byte aligned smem dst
    // Measured (profiles/r01_notes.md): splitting the span into evenly sized boxes makes 2:1 down-sampling of Float32
    // fit the slot budget, but there the march runs at 2.3 ms against 1.06 ms for the direct gather (no plane is
    // re-used and the taps need no conversion, so the per-plane hand-off is pure latency); Float32 down-sampling
    // therefore stays on the gather kernel -- the slot budget below rejects it.
    if (p.nBoxX > kMaxBoxesX) return cudaErrorNotSupported;
    int BY = marchPasses(bd) * kRowsY;
    long long byS = 0, slot = 0, spillOff = 0;
    for (; BY >= kRowsY; BY /= 2) {
        byS = (long long)BY * S[1] / D[1] + 3;
        spillOff = ((long long)p.nBoxX * p.bxS * byS * bs + 127) / 128 * 128;
        slot = (spillOff + byS * 16 + 127) / 128 * 128;
        if (byS <= kMaxBox && slot <= 20 * 1024) break;
    }
    if (BY < kRowsY) return cudaErrorNotSupported;
    p.BY = BY; p.byS = (int)byS; p.slotBytes = (int)slot; p.spillOff = (int)spillOff;
    const int tilesY = (D[1] + BY - 1) / BY;
    const size_t smemBytes = (size_t)kRing * slot;

    CUtensorMap tm[2];
    const cuuint64_t gdim[3] = { (cuuint64_t)S[0], (cuuint64_t)S[1], (cuuint64_t)szCount };
    const cuuint64_t gstr[2] = { (cuuint64_t)S[0] * bs, (cuuint64_t)S[0] * S[1] * bs };
    const cuuint32_t box[3] = { (cuuint32_t)p.bxS, (cuuint32_t)p.byS, 1u };
    const cuuint32_t estr[3] = { 1u, 1u, 1u };
    const CUtensorMapDataType dt = bs == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : bs == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16
                                                                                      : CU_TENSOR_MAP_DATA_TYPE_UINT32;
    if (encode(&tm[0], dt, 3, const_cast<uint8_t*>(src.data), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return cudaErrorNotSupported;
    // the spill column: the window as Sy*szCount rows, boxes of 16 bytes x byS rows
    const cuuint64_t gdim2[2] = { (cuuint64_t)S[0], (cuuint64_t)S[1] * (cuuint64_t)szCount };
    const cuuint32_t box2[2] = { (cuuint32_t)(16 / bs), (cuuint32_t)p.byS };
    if ((cuuint64_t)S[1] * (cuuint64_t)szCount > 0x7fffffffull ||
        encode(&tm[1], dt, 2, const_cast<uint8_t*>(src.data), gdim2, gstr, box2, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return cudaErrorNotSupported;

    const size_t nBase = kHdr + 3 * (size_t)D[0] + 3 * (size_t)D[1] + 5 * (size_t)dzCount + 3 * (size_t)szCount;
    const size_t zpackOff = (nBase + 3) / 4 * 4;
    if (zpackOff > 0x7fffffff) return cudaErrorNotSupported;
    const size_t nTab = zpackOff + 4 * (size_t)dzCount;
    p.zpackOff = (int)zpackOff;
    int* tab = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&tab, nTab * sizeof(int), c.stream);
    if (e != cudaSuccess) return e;
    p.tab = tab;
    marchPlanKernel<<<3, 1024, 0, c.stream>>>(tab, D[0], S[0], D[1], S[1], dzCount, dstZ0, D[2], S[2], srcZ0, szCount, mode, p.zpackOff);
    countLaunch();

    const unsigned grid = (unsigned)(p.tilesX * tilesY);
    switch (src.format) {
    case F_U8:  e = launchMarchDst<F_U8>(c, tm, p, smemBytes, grid, dst.format, mode); break;
    case F_U16: e = launchMarchDst<F_U16>(c, tm, p, smemBytes, grid, dst.format, mode); break;
    case F_U32: e = launchMarchDst<F_U32>(c, tm, p, smemBytes, grid, dst.format, mode); break;
    case F_F32: e = launchMarchDst<F_F32>(c, tm, p, smemBytes, grid, dst.format, mode); break;
    default: e = cudaErrorNotSupported;
    }
    cudaFreeAsync(tab, c.stream);
    return e;
}

} // namespace vktb
