// Resample, one-tap modes (Nearest, and reference-faithful Linear on integer sources, which is bit-identical to
// Nearest -- see resample.cu), as a TMA row pipeline.
//
// The coordinate maps of reference src/vkt/Resample_serial.hpp:60-62 depend on one axis each and are monotone,
// so the dst rows (y,z) that read the same src row (sy,sz) form a box [y-run] x [z-run].  One work item is
// (x tile, group of R y-runs, one z-run):
//     bulk-load the R src row segments into shared memory   (cp.async.bulk global -> shared, mbarrier; SASS UBLKCP)
//     every thread gathers 16-byte dst chunks from the staged rows (x map held in registers as 16-bit tap offsets)
//     and streams each chunk to every dst row of the run box with 128-bit stores
// so every src byte that is needed is read from HBM exactly once in full lines, the x gather is paid once per src
// row (not once per dst voxel: 2.25x fewer gathers at 2048 -> 3072) and src rows that no dst row maps to
// (down-sampling) are never touched.  Loads run `stages - 1` items ahead of the consumers.
//
// Measured alternatives (profiles/r01_resample_notes.md): expanding into a shared dst row and writing it with
// cp.async.bulk shared -> global was 7 % faster at 2048^3 -> 3072^3 but 2x slower when each row is stored once
// (the per-item wait for the bulk stores to drain serialises the CTA); 32-bit-per-lane stores (conflict-free taps)
// lost to the 128-bit stores used here.
//
// Algorithmic traffic: bytes of the src rows that are hit + Dx*Dy*Dz*bpv_dst.
#include <cmath>
#include <cstdlib>

#include "common.cuh"
#include "launch.hpp"
#include "tma.cuh"
#include "voxel_codec.cuh"

namespace vktb {

namespace {

// tuning overrides for experiments (tools/rs_probe.py); defaults are what ships
inline int envInt(const char* name, int dflt, int lo, int hi)
{
    const char* e = getenv(name);
    if (!e || !*e) return dflt;
    int v = atoi(e);
    return v < lo ? lo : v > hi ? hi : v;
}

constexpr int kRowThreads = 128; // CTA size of the row pipeline
constexpr int kMaxChunks = 4;    // 16-byte dst chunks of one row segment per thread
constexpr int kMaxStages = 8;    // src stages (items) in flight per CTA
constexpr int kHdr = 4;           // table header ints: {#y-runs, #z-runs, 0, 0}

// reference Resample_serial.hpp:60-62 + the (int32_t) conversion at :67 (x86 cvttss2si)
__device__ __forceinline__ int srcCoord(int d, int D, int S)
{
    return (int)x86Cvtt32(__fmul_rn(__fdiv_rn(__int2float_rn(d), __int2float_rn(D)), __int2float_rn(S)));
}

// Per-axis coordinate map of the plan: mode 0 = Resample's d / float(D) * S, mode 1 = Scale's m * (d - c) + c (the
// diagonal case of the affine gather, affine_filter.cu; inside the range by the launcher's check, so the FP32-pipe
// truncation is not needed here -- the plan runs once per launch)
struct AxisMap { int mode, D, S; float m, c; };
__device__ __forceinline__ int coordOf(const AxisMap& a, int d)
{
    if (a.mode == 0) return srcCoord(d, a.D, a.S);
    return (int)x86Cvtt32(__fadd_rn(__fmul_rn(a.m, __fsub_rn(__int2float_rn(d), a.c)), a.c));
}

// Plan: block 0 -> x map; block 1 -> y runs; block 2 -> z runs.  1024 threads per block.
// runs: start[k] = first dst index of run k (start[nRuns] = count), src[k] = its src coordinate.
__global__ void __launch_bounds__(1024)
resamplePlanKernel(int* tab, int Dx, int Dy, int dzCount, int dz0, AxisMap mx, AxisMap my, AxisMap mz)
{
    __shared__ int scan[1024];
    const int t = threadIdx.x;
    int* xmap = tab + kHdr;
    if (blockIdx.x == 0) {
        for (int i = t; i < Dx; i += 1024) xmap[i] = coordOf(mx, i);
        return;
    }
    const bool isY = blockIdx.x == 1;
    const int n = isY ? Dy : dzCount, d0 = isY ? 0 : dz0;
    const AxisMap am = isY ? my : mz;
    int* start = tab + kHdr + Dx + (isY ? 0 : 2 * Dy + 1);
    int* src = start + n + 1;
    const int seg = (n + 1023) / 1024;
    const int i0 = min(n, t * seg), i1 = min(n, i0 + seg);
    int cnt = 0;
    int prev = i0 > 0 ? coordOf(am, d0 + i0 - 1) : 0;
    for (int i = i0; i < i1; ++i) {
        const int c = coordOf(am, d0 + i);
        cnt += (i == 0 || c != prev) ? 1 : 0;
        prev = c;
    }
    scan[t] = cnt;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = t >= o ? scan[t - o] : 0;
        __syncthreads();
        scan[t] += v;
        __syncthreads();
    }
    int k = scan[t] - cnt;   // exclusive prefix
    prev = i0 > 0 ? coordOf(am, d0 + i0 - 1) : 0;
    for (int i = i0; i < i1; ++i) {
        const int c = coordOf(am, d0 + i);
        if (i == 0 || c != prev) { start[k] = i; src[k] = c; ++k; }
        prev = c;
    }
    if (t == 1023) {
        const int total = scan[1023];
        start[total] = n;
        tab[isY ? 0 : 1] = total;
    }
}

struct RowRunParams
{
    uint8_t*       d;          // dst window base
    const uint8_t* s;          // src window base
    int Dx, Dy, dzCount;
    int Sx, Sy;
    int srcZ0;                 // global index of the src window's first plane
    int tileW, nXT;            // dst x tile (voxels), number of x tiles
    int R;                     // y-runs per item
    int stages;                // src stages in flight (<= kMaxStages)
    int srcStageRow;           // bytes reserved per src row segment in a stage (multiple of 16)
    const int* tab;
    unsigned* tapFlag;         // Float32 sources only: set to 1 when a staged src value is Inf, NaN or -0 (resample.cu)
    Mapping md, ms;
};

template <int B> struct RawT;
template <> struct RawT<1> { typedef uint8_t type; };
template <> struct RawT<2> { typedef uint16_t type; };
template <> struct RawT<4> { typedef uint32_t type; };

// Producer/consumer pipeline without CTA-wide barriers in the steady state: warp 0 keeps `stages - 1` items of
// bulk loads in flight (full[] mbarriers, transaction counted); every thread gathers its 16-byte dst chunks from
// the staged rows (tap offsets held in registers) and streams them with 128-bit stores to all dst rows of the run
// box; a stage is refilled once all warps have arrived on empty[].
template <int FS, int FD, bool MOVE>
__global__ void __launch_bounds__(kRowThreads)
resampleRowRunKernel(RowRunParams p)
{
    constexpr int BS = bytesPerVoxel(FS), BD = bytesPerVoxel(FD);
    constexpr int V = 16 / BD;   // dst voxels per 128-bit store
    typedef typename RawT<BS>::type SrcT;
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[kMaxStages], empty[kMaxStages];

    const int t = threadIdx.x;
    const int* xmap = p.tab + kHdr;
    const int* yStart = xmap + p.Dx;
    const int* ySrc = yStart + p.Dy + 1;
    const int* zStart = ySrc + p.Dy;
    const int* zSrc = zStart + p.dzCount + 1;
    const int nyr = p.tab[0], nzr = p.tab[1];
    const int nyg = (nyr + p.R - 1) / p.R;
    const long long nItems = (long long)nyg * nzr;
    const int S = p.stages;

    // this CTA's x tile and its slice of the (y-group, z-run) items
    const int xt = blockIdx.x % p.nXT;
    const long long g = blockIdx.x / p.nXT, G = gridDim.x / p.nXT;
    const int x0 = xt * p.tileW;
    const int w = min(p.tileW, p.Dx - x0);
    const int nChunks = w / V;
    const int srcByte0 = (xmap[x0] * BS) & ~15;
    const int srcBytes = (((xmap[x0 + w - 1] + 1) * BS + 15) & ~15) - srcByte0;
    // byte offsets of this thread's taps inside a staged src row segment, two 16-bit offsets per register
    uint32_t rel[kMaxChunks][V / 2];
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
        const int q = t + c * kRowThreads;
#pragma unroll
        for (int i = 0; i < V; i += 2) {
            uint32_t lo = 0, hi = 0;
            if (q < nChunks) {
                lo = (uint32_t)(xmap[x0 + q * V + i] * BS - srcByte0);
                hi = (uint32_t)(xmap[x0 + q * V + i + 1] * BS - srcByte0);
            }
            rel[c][i / 2] = lo | (hi << 16);
        }
    }

    const int stageBytes = p.R * p.srcStageRow;
    if (t == 0) {
        for (int i = 0; i < S; ++i) { mbarInit(smemAddr(&full[i]), 1); mbarInit(smemAddr(&empty[i]), kRowThreads / 32); }
        fenceBarrierInit();
    }
    // one tap: the dst voxel is a pure function of the src voxel -- for byte sources that function is a 256-entry
    // table, built here with the exact codec (the UInt8 unmap is an IEEE division per voxel otherwise)
    constexpr bool kTable = FS == F_U8 && !MOVE;
    __shared__ uint32_t convTab[kTable ? 256 : 1];
    if (kTable)
        for (int i = t; i < 256; i += kRowThreads)
            convTab[i] = mapVoxel<FD, false>(unmapVoxel<FS, false>((uint32_t)i, p.ms), p.md);
    __syncthreads();

    const long long srcRowPitch = (long long)p.Sx * BS;
    const long long srcPlanePitch = srcRowPitch * p.Sy;
    const long long dstRowPitch = (long long)p.Dx * BD;
    const long long dstPlanePitch = dstRowPitch * p.Dy;

    // warp 0: refill stage (seq % S) with item `item` (sequence number seq of this CTA)
    auto produce = [&](long long item, long long seq) {
        const int stage = (int)(seq % S);
        mbarWait(smemAddr(&empty[stage]), (uint32_t)(((seq / S) & 1) ^ 1));   // all warps are done with its last use
        const int yg = (int)(item % nyg), zr = (int)(item / nyg);
        const int r0 = yg * p.R, nr = min(p.R, nyr - r0);
        const uint32_t bar = smemAddr(&full[stage]);
        if (t == 0) mbarArriveExpectTx(bar, (uint32_t)(nr * srcBytes));
        __syncwarp();
        if (t < nr) {
            const uint8_t* sp = p.s + (long long)(zSrc[zr] - p.srcZ0) * srcPlanePitch + (long long)ySrc[r0 + t] * srcRowPitch + srcByte0;
            bulkLoad(smemAddr(smem + stage * stageBytes + t * p.srcStageRow), sp, (uint32_t)srcBytes, bar);
        }
    };

    if (t < 32) {
        for (int i = 0; i < S - 1; ++i) {
            const long long item = g + (long long)i * G;
            if (item < nItems) produce(item, i);
        }
    }

    long long k = 0;
    bool badTap = false;
    for (long long item = g; item < nItems; item += G, ++k) {
        const int stage = (int)(k % S);
        if (t < 32) {
            const long long ahead = item + (long long)(S - 1) * G;
            if (ahead < nItems) produce(ahead, k + S - 1);
        }
        mbarWait(smemAddr(&full[stage]), (uint32_t)((k / S) & 1));

        const int yg = (int)(item % nyg), zr = (int)(item / nyg);
        const int r0 = yg * p.R, nr = min(p.R, nyr - r0);
        const int z0 = zStart[zr], z1 = zStart[zr + 1];
        const uint8_t* sb = smem + stage * stageBytes;
        if (FS == F_F32 && p.tapFlag) {
            // every staged src value is looked at once (the whole source, when the axis maps are onto): Inf / NaN / -0
            for (int r = 0; r < nr; ++r) {
                const uint4* row = reinterpret_cast<const uint4*>(sb + r * p.srcStageRow);
                for (int v = t; v < srcBytes / 16; v += kRowThreads) {
                    const uint4 q = row[v];
                    badTap |= (q.x & 0x7f800000u) == 0x7f800000u || q.x == 0x80000000u
                           || (q.y & 0x7f800000u) == 0x7f800000u || q.y == 0x80000000u
                           || (q.z & 0x7f800000u) == 0x7f800000u || q.z == 0x80000000u
                           || (q.w & 0x7f800000u) == 0x7f800000u || q.w == 0x80000000u;
                }
            }
        }
        for (int r = 0; r < nr; ++r) {
            const uint8_t* srow = sb + r * p.srcStageRow;
            const int y0 = yStart[r0 + r], y1 = yStart[r0 + r + 1];
#pragma unroll
            for (int c = 0; c < kMaxChunks; ++c) {
                if (c * kRowThreads >= nChunks) break;   // uniform
                const int q = t + c * kRowThreads;
                if (q < nChunks) {
                    uint32_t out[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
                    for (int i = 0; i < V; ++i) {
                        const uint32_t off = (rel[c][i / 2] >> ((i & 1) * 16)) & 0xffffu;
                        const uint32_t raw = (uint32_t)*reinterpret_cast<const SrcT*>(srow + off);
                        vecSet<BD>(out, i, MOVE ? raw : kTable ? convTab[raw & 0xffu] : mapVoxel<FD, false>(unmapVoxel<FS, false>(raw, p.ms), p.md));
                    }
                    const uint4 v = make_uint4(out[0], out[1], out[2], out[3]);
                    uint8_t* dcol = p.d + (long long)x0 * BD + (long long)q * 16;
                    for (int z = z0; z < z1; ++z)
                        for (int y = y0; y < y1; ++y)
                            stStream(dcol + (long long)z * dstPlanePitch + (long long)y * dstRowPitch, v);
                }
            }
        }
        __syncwarp();
        if ((t & 31) == 0) mbarArrive(smemAddr(&empty[stage]));
    }
    if (FS == F_F32 && badTap) atomicOr(p.tapFlag, 1u);
}

template <int FS, int FD, bool MOVE>
cudaError_t launchRowRun(const LaunchCtx& c, const RowRunParams& p, size_t smemBytes, unsigned grid)
{
    auto kern = resampleRowRunKernel<FS, FD, MOVE>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
    if (e != cudaSuccess) return e;
    kern<<<grid, kRowThreads, smemBytes, c.stream>>>(p);
    countLaunch();
    return cudaGetLastError();
}

template <int FS>
cudaError_t launchRowRunDst(const LaunchCtx& c, const RowRunParams& p, int fd, size_t smemBytes, unsigned grid)
{
    switch (fd) {
    case F_U8:  return launchRowRun<FS, F_U8, false>(c, p, smemBytes, grid);
    case F_U16: return launchRowRun<FS, F_U16, false>(c, p, smemBytes, grid);
    case F_U32: return launchRowRun<FS, F_U32, false>(c, p, smemBytes, grid);
    case F_F32: return launchRowRun<FS, F_F32, false>(c, p, smemBytes, grid);
    }
    return cudaErrorNotSupported;
}

} // namespace

// The pipeline's launcher, shared by Resample (one-tap modes) and Scale.
// maps: the per-axis coordinate maps; spanNum / spanDen: src voxels per dst voxel along x (an upper bound of the
// slope: a dst tile of w voxels taps at most w * spanNum / spanDen + 2 src voxels)
static cudaError_t launchRowRunsMapped(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src, int dstZ0,
                                       const int32_t D[3], int srcZ0, const int32_t S[3], const AxisMap maps[3],
                                       long long spanNum, long long spanDen, bool up, unsigned* tapFlag = nullptr)
{
    const int bs = bytesPerVoxel(src.format), bd = bytesPerVoxel(dst.format);
    const int dzCount = dst.dims[2];
    auto plain = [](int f) { return f == F_U8 || f == F_U16 || f == F_U32 || f == F_F32; };
    if (!plain(src.format) || !plain(dst.format)) return cudaErrorNotSupported;
    // bulk copies need 16-byte aligned rows on both sides
    if ((reinterpret_cast<uintptr_t>(src.data) & 15u) || (reinterpret_cast<uintptr_t>(dst.data) & 15u)) return cudaErrorNotSupported;
    if (((long long)S[0] * bs) % 16 || ((long long)D[0] * bd) % 16) return cudaErrorNotSupported;
    for (int a = 0; a < 3; ++a) if (D[a] > (1 << 24) || S[a] > (1 << 24)) return cudaErrorNotSupported;

    RowRunParams p;
    p.d = dst.data; p.s = src.data;
    p.Dx = D[0]; p.Dy = D[1]; p.dzCount = dzCount;
    p.Sx = S[0]; p.Sy = S[1]; p.srcZ0 = srcZ0;
    p.md = dst.map; p.ms = src.map;
    p.tapFlag = src.format == F_F32 ? tapFlag : nullptr;

    // x tile: at most kMaxChunks 16-byte dst chunks per thread (8 KiB of dst)
    int tileW = kMaxChunks * kRowThreads * (16 / bd);
    if (tileW >= D[0]) tileW = D[0];
    p.tileW = tileW;
    p.nXT = (D[0] + tileW - 1) / tileW;
    // src bytes a tile can span: floor(tileW*Sx/Dx)+2 voxels, plus the two 16-byte roundings
    const long long spanVox = (long long)tileW * spanNum / spanDen + 2;
    long long srcRow = spanVox * bs + 32;
    if (srcRow > (long long)S[0] * bs) srcRow = (long long)S[0] * bs;
    srcRow = (srcRow + 15) / 16 * 16;
    if (srcRow > 49152) return cudaErrorNotSupported;   // extreme down-sampling ratios: direct gather is the better fit
    p.srcStageRow = (int)srcRow;
    // one stage = R src rows of about 16 KiB together; 3 stages = 2 items of loads in flight per CTA, ~4 CTAs per SM
    int R = (int)(16384 / srcRow);
    if (R < 1) R = 1;
    if (R > 8) R = 8;
    // measured (profiles/r01_notes.md: sweeps over rows x stages x grid, repeated on one box): up-sampling runs best with
    // 2 rows per item and 2 stages (small CTAs, 8 resident per SM); down-sampling with one 16 KiB item and 3 stages
    if (up && R > 2) R = 2;
    p.R = envInt("VKTB_RS_ROWS", R, 1, 8);
    p.stages = envInt("VKTB_RS_STAGES", up ? 2 : 3, 2, kMaxStages);
    const size_t smemBytes = (size_t)p.stages * p.R * srcRow;

    // plan tables: header, x map, y runs, z runs
    const size_t nTab = kHdr + (size_t)D[0] + 2 * (size_t)D[1] + 1 + 2 * (size_t)dzCount + 1;
    int* tab = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&tab, nTab * sizeof(int), c.stream);
    if (e != cudaSuccess) return e;
    p.tab = tab;
    resamplePlanKernel<<<3, 1024, 0, c.stream>>>(tab, D[0], D[1], dzCount, dstZ0, maps[0], maps[1], maps[2]);
    countLaunch();

    // persistent grid: as many CTAs per SM as shared memory allows (<= 8), a multiple of the x-tile count
    int perSm = (int)((200u << 10) / (smemBytes + 1024));
    if (perSm > 2048 / kRowThreads) perSm = 2048 / kRowThreads;
    perSm = envInt("VKTB_RS_CTAS", perSm, 1, perSm);
    if (perSm < 1) perSm = 1;
    const long long maxItems = (long long)((D[1] < S[1] ? D[1] : S[1]) + p.R - 1) / p.R * (dzCount < S[2] ? dzCount : S[2]);
    long long ctas = (long long)c.smCount * perSm;
    if (ctas > maxItems * p.nXT) ctas = maxItems * p.nXT;
    ctas = (ctas + p.nXT - 1) / p.nXT * p.nXT;
    if (ctas < p.nXT) ctas = p.nXT;
    const unsigned grid = (unsigned)ctas;

    const bool identity = src.map.identity && dst.map.identity;
    if (src.format == dst.format && identity && src.format == F_U16) {
        // unmap: raw * 2^-16 (exact); map: t * 65536 (exact), truncation of an integer-valued float: raw -> raw
        e = launchRowRun<F_U16, F_U16, true>(c, p, smemBytes, grid);
    } else if (src.format == dst.format && identity && src.format == F_U8) {
        // (uint8)(float(v) / 255.999f * 255.999f) == v for all 256 values (checked exhaustively by
        // tests/test_golden_oracle.py::test_u8_identity_round_trip against the reference build): raw -> raw
        e = launchRowRun<F_U8, F_U8, true>(c, p, smemBytes, grid);
    } else if (src.format == dst.format && identity && src.format == F_F32) {
        // unmap: raw bits; map: (v - +0) / 1 == v (NaN payloads survive, as on x86)
        e = launchRowRun<F_F32, F_F32, true>(c, p, smemBytes, grid);
    } else {
        switch (src.format) {
        case F_U8:  e = launchRowRunDst<F_U8>(c, p, dst.format, smemBytes, grid); break;
        case F_U16: e = launchRowRunDst<F_U16>(c, p, dst.format, smemBytes, grid); break;
        case F_U32: e = launchRowRunDst<F_U32>(c, p, dst.format, smemBytes, grid); break;
        case F_F32: e = launchRowRunDst<F_F32>(c, p, dst.format, smemBytes, grid); break;
        default: e = cudaErrorNotSupported;
        }
    }
    cudaFreeAsync(tab, c.stream);
    return e;
}

// One-tap resample through the TMA row pipeline.  Returns cudaErrorNotSupported when the operands do not
// qualify (unaligned rows, Int16, giant dims); the caller then uses the direct-gather kernel.
cudaError_t launchResampleRowRuns(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src,
                                  int dstZ0, const int32_t D[3], int srcZ0, const int32_t S[3], unsigned* tapFlag)
{
    // (D-1)/float(D)*S < S holds in binary32 below 2^24 (checked by the core), so every tap lies inside the src row
    AxisMap maps[3];
    for (int a = 0; a < 3; ++a) maps[a] = AxisMap{ 0, D[a], S[a], 0.f, 0.f };
    return launchRowRunsMapped(c, dst, src, dstZ0, D, srcZ0, S, maps, S[0], D[0], D[0] > S[0], tapFlag);
}

// Scale (diagonal affine gather, one-tap modes) whose EVERY dst voxel maps inside the source: a separable monotone
// gather like Resample Nearest, so it runs on the same pipeline.  m[a] > 0 and center[a] are the diagonal and the
// centre; the caller has verified the one-tap conditions.  cudaErrorNotSupported -> the affine kernel.
cudaError_t launchScaleRowRuns(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src, const float m[3],
                               const float center[3], const int32_t first[3], const int32_t last[3])
{
    AxisMap maps[3];
    for (int a = 0; a < 3; ++a) {
        if (!(m[a] > 0.f) || !std::isfinite(m[a]) || !std::isfinite(center[a]) || dst.dims[a] <= 0) return cudaErrorNotSupported;
        // q(d) = m * (d - c) + c is non-decreasing in d (every step is monotone under rounding): all dst voxels map
        // inside [first, last) iff the two ends do.  volatile: one rounding per operation, as on the device.
        auto q = [&](int d) {
            volatile float p = (float)d - center[a];
            volatile float t = m[a] * p;
            volatile float r = t + center[a];
            return (float)r;
        };
        if (!(q(0) >= (float)first[a]) || !(q(dst.dims[a] - 1) < (float)last[a])) return cudaErrorNotSupported;
        if (first[a] < 0 || last[a] > src.dims[a]) return cudaErrorNotSupported;
        maps[a] = AxisMap{ 1, dst.dims[a], src.dims[a], m[a], center[a] };
    }
    // slope bound along x as a fraction: ceil(m * 2^16) / 2^16
    const long long num = (long long)std::ceil((double)m[0] * 65536.0) + 1;
    return launchRowRunsMapped(c, dst, src, 0, dst.dims, 0, src.dims, maps, num, 65536, m[0] < 1.f);
}

} // namespace vktb