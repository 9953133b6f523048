// Resample (StructuredVolume -> StructuredVolume): nearest, reference-faithful "Linear" and the opt-in true
// trilinear filter, whole volume or z-slab window.
//
// Semantics: reference src/vkt/Resample_serial.hpp:26-71 and StructuredVolumeView::sampleLinear,
// src/vkt/StructuredVolumeView.hpp:80-119.
//   * equal dims        -> format/mapping conversion only (:32-49)
//   * src coordinate    s = d / float(D) * S in binary32 (int->float, IEEE divide, multiply), :60-62
//   * Nearest           tap (int)s, :67
//   * Linear            sampleLinear takes INTEGER parameters, so s is truncated at the call and frac == 0;
//                       taps lo=(x,y,z), hi=(x+1, clamp(y+1), clamp(z+1)) -- hi.x is NOT clamped (typo in the
//                       reference) -- and the full lerp chain runs with weight 0.
//   * Trilinear (ours)  the same body with float parameters and all six clamps: NOT what the reference computes.
//
// The coordinate maps depend on one axis only, so they are computed once per launch into small tables with
// the exact IEEE operations (coordTableKernel); the voxel kernels then contain no division.
//
// Linear on an INTEGER source format is provably identical to Nearest: every tap unmaps to a finite float, so
// each lerp(a, b, 0) = 1*a + 0*b = a + (+-0) = a (a is never -0 when neither mapping endpoint is a negative
// zero / negative denormal-scale value; checked on the host in linearIsNearest()).  Only Float32 sources (NaN,
// Inf, -0 taps) and exotic mappings need the 8 taps.
#include "common.cuh"
#include "launch.hpp"
#include "voxel_codec.cuh"

#include <cmath>

namespace vktb {

namespace {

enum { MODE_NEAREST = 0, MODE_LINEAR8 = 1, MODE_TRILINEAR = 2 };

// mode 0/1: lo = (int)s.  mode 2 (trilinear): lo = clamp((int)s), hi = clamp((int)(s + 1.f)), frac = s - float(lo)
__global__ void coordTableKernel(int* lo, int* hi, float* frac, int D, int S, int d0, int count, int trilinear)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const float s = __fmul_rn(__fdiv_rn(__int2float_rn(d0 + i), __int2float_rn(D)), __int2float_rn(S));
        int l = (int)x86Cvtt32(s);
        int h = (int)x86Cvtt32(__fadd_rn(s, 1.f));
        float f = 0.f;
        if (trilinear) {
            l = min(max(l, 0), S - 1);
            h = min(max(h, 0), S - 1);
            f = __fsub_rn(s, __int2float_rn(l));
        }
        lo[i] = l; hi[i] = h; frac[i] = f;
    }
}

struct RsParams
{
    uint8_t*       d;        // dst window base (plane dz0)
    const uint8_t* s;        // src window base (plane sz0)
    int Dx, Dy, dzCount;     // dst row length, rows per plane, planes in the window
    int Sx, Sy, Sz;          // src full dims
    int sz0, szCount;        // src window
    const int*   xmap; const int* ymap; const int* zmap;      // per-axis source coordinates (dst-window relative in z)
    const int*   xhi;  const int* yhi;  const int* zhi;       // trilinear upper taps
    const float* xfr;  const float* yfr;  const float* zfr;   // trilinear fractions
    Mapping md, ms;
};

template <int FS>
__device__ __forceinline__ float tap(const RsParams& p, long long idx /*voxel index inside the src window*/, long long count)
{
    constexpr int BS = bytesPerVoxel(FS);
    // only the reference's unclamped hi.x tap at the very last src voxel can leave the window: it is read as
    // an all-zero voxel (the reference reads one voxel past its malloc there and multiplies it by 0)
    uint32_t raw = (idx >= 0 && idx < count) ? loadRaw(p.s + idx * BS, BS) : 0u;
    return unmapVoxel<FS, false>(raw, p.ms);
}

// one thread = V consecutive dst voxels of one row (one 128-bit store)
template <int FS, int FD, int MODE>
__global__ void __launch_bounds__(kThreads)
resampleKernel(RsParams p)
{
    constexpr int BS = bytesPerVoxel(FS), BD = bytesPerVoxel(FD);
    constexpr int V = 16 / BD;
    const int chunks = (p.Dx + V - 1) / V;
    const long long rows = (long long)p.Dy * p.dzCount;
    const long long plane = (long long)p.Sx * p.Sy;
    const long long count = plane * p.szCount;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= chunks) return;
    const int x0 = c * V;
    const int n = min(V, p.Dx - x0);

    for (long long r = (long long)blockIdx.y * blockDim.y + threadIdx.y; r < rows; r += (long long)gridDim.y * blockDim.y) {
        const int zz = (int)(r / p.Dy), y = (int)(r - (long long)zz * p.Dy);
        const int sy = p.ymap[y], sz = p.zmap[zz];
        uint8_t* dp = p.d + (r * p.Dx + x0) * BD;
        uint32_t w[4] = { 0u, 0u, 0u, 0u };
        uint32_t out[V];
        if (MODE == MODE_NEAREST) {
            const uint8_t* srow = p.s + ((long long)(sz - p.sz0) * plane + (long long)sy * p.Sx) * BS;
#pragma unroll
            for (int i = 0; i < V; ++i) {
                if (i < n) {
                    float f = unmapVoxel<FS, false>(loadRaw(srow + (long long)p.xmap[x0 + i] * BS, BS), p.ms);
                    out[i] = mapVoxel<FD, false>(f, p.md);
                }
            }
        } else {
            int hy, hz; float fy, fz;
            if (MODE == MODE_LINEAR8) {
                hy = min(max(sy + 1, 0), p.Sy - 1); hz = min(max(sz + 1, 0), p.Sz - 1);
                // frac = float(s) - float(clamp(s)): +0 for every in-range coordinate
                fy = __fsub_rn(__int2float_rn(sy), __int2float_rn(min(max(sy, 0), p.Sy - 1)));
                fz = __fsub_rn(__int2float_rn(sz), __int2float_rn(sz));   // lo.z is never clamped (reference typo)
            } else {
                hy = p.yhi[y]; hz = p.zhi[zz];
                fy = p.yfr[y]; fz = p.zfr[zz];
            }
            const long long r00 = (long long)(sz - p.sz0) * plane + (long long)sy * p.Sx;
            const long long r10 = (long long)(sz - p.sz0) * plane + (long long)hy * p.Sx;
            const long long r01 = (long long)(hz - p.sz0) * plane + (long long)sy * p.Sx;
            const long long r11 = (long long)(hz - p.sz0) * plane + (long long)hy * p.Sx;
#pragma unroll
            for (int i = 0; i < V; ++i) {
                if (i < n) {
                    const int lx = p.xmap[x0 + i];
                    int hx; float fx;
                    if (MODE == MODE_LINEAR8) {
                        hx = lx + 1;   // NOT clamped (reference typo): next row / next plane / past the end
                        fx = __fsub_rn(__int2float_rn(lx), __int2float_rn(min(max(lx, 0), p.Sx - 1)));
                    } else {
                        hx = p.xhi[x0 + i];
                        fx = p.xfr[x0 + i];
                    }
                    float v0 = tap<FS>(p, r00 + lx, count), v1 = tap<FS>(p, r00 + hx, count);
                    float v2 = tap<FS>(p, r10 + lx, count), v3 = tap<FS>(p, r10 + hx, count);
                    float v4 = tap<FS>(p, r01 + lx, count), v5 = tap<FS>(p, r01 + hx, count);
                    float v6 = tap<FS>(p, r11 + lx, count), v7 = tap<FS>(p, r11 + hx, count);
                    float f = refLerp(refLerp(refLerp(v0, v1, fx), refLerp(v2, v3, fx), fy),
                                      refLerp(refLerp(v4, v5, fx), refLerp(v6, v7, fx), fy), fz);
                    out[i] = mapVoxel<FD, false>(f, p.md);
                }
            }
        }
        if (n == V && (reinterpret_cast<uintptr_t>(dp) & 15u) == 0) {
#pragma unroll
            for (int i = 0; i < V; ++i) vecSet<BD>(w, i, out[i]);
            stStream(dp, make_uint4(w[0], w[1], w[2], w[3]));
        } else {
#pragma unroll
            for (int i = 0; i < V; ++i) if (i < n) storeRaw(dp + i * BD, BD, out[i]);
        }
    }
}

// 8-tap modes when the x map is exactly lo = 2x, hi = 2x + 1 (2:1 down-sampling in x; the launcher verifies the whole
// map on the host): the 2V x taps of V consecutive dst voxels are 32 consecutive bytes of the src row, so the 8 taps
// come from 4 rows x 2 coalesced 128-bit loads, all eight in flight before the first use.  Same formats on both
// sides (same bytes per voxel).  hi.x = 2x + 1 <= Sx - 1: the reference's unclamped tap never leaves the row here.
template <int F, int MODE>
__global__ void __launch_bounds__(kThreads)
resampleHalveXKernel(RsParams p)
{
    constexpr int BPV = bytesPerVoxel(F);
    constexpr int V = 16 / BPV;
    const int chunks = p.Dx / V;                      // Dx is a multiple of V (launcher)
    const long long rows = (long long)p.Dy * p.dzCount;
    const long long rowPitch = (long long)p.Sx * BPV, planePitch = rowPitch * p.Sy;
    // byte voxels unmap through a 256-entry table (exact codec) instead of an IEEE division per tap
    __shared__ float lut[F == F_U8 ? 256 : 1];
    if (F == F_U8) {
        for (int i = threadIdx.y * blockDim.x + threadIdx.x; i < 256; i += blockDim.x * blockDim.y)
            lut[i] = unmapVoxel<F_U8, false>((uint32_t)i, p.ms);
        __syncthreads();
    }
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= chunks) return;
    const int x0 = c * V;
    float fx[V];
#pragma unroll
    for (int i = 0; i < V; ++i) fx[i] = MODE == MODE_TRILINEAR ? p.xfr[x0 + i] : 0.f;   // Linear: float(lo) - float(lo)

    for (long long r = (long long)blockIdx.y * blockDim.y + threadIdx.y; r < rows; r += (long long)gridDim.y * blockDim.y) {
        const int zz = (int)(r / p.Dy), y = (int)(r - (long long)zz * p.Dy);
        const int sy = p.ymap[y], sz = p.zmap[zz];
        int hy, hz; float fy, fz;
        if (MODE == MODE_LINEAR8) {
            hy = min(max(sy + 1, 0), p.Sy - 1); hz = min(max(sz + 1, 0), p.Sz - 1);
            fy = __fsub_rn(__int2float_rn(sy), __int2float_rn(min(max(sy, 0), p.Sy - 1)));
            fz = __fsub_rn(__int2float_rn(sz), __int2float_rn(sz));
        } else {
            hy = p.yhi[y]; hz = p.zhi[zz];
            fy = p.yfr[y]; fz = p.zfr[zz];
        }
        const uint8_t* b00 = p.s + (long long)(sz - p.sz0) * planePitch + (long long)sy * rowPitch + (long long)x0 * 2 * BPV;
        const uint8_t* b10 = p.s + (long long)(sz - p.sz0) * planePitch + (long long)hy * rowPitch + (long long)x0 * 2 * BPV;
        const uint8_t* b01 = p.s + (long long)(hz - p.sz0) * planePitch + (long long)sy * rowPitch + (long long)x0 * 2 * BPV;
        const uint8_t* b11 = p.s + (long long)(hz - p.sz0) * planePitch + (long long)hy * rowPitch + (long long)x0 * 2 * BPV;
        const uint4 q00a = ldStream(b00), q00b = ldStream(b00 + 16), q10a = ldStream(b10), q10b = ldStream(b10 + 16);
        const uint4 q01a = ldStream(b01), q01b = ldStream(b01 + 16), q11a = ldStream(b11), q11b = ldStream(b11 + 16);
        uint32_t w[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
        for (int i = 0; i < V; ++i) {
            // voxels 2i and 2i+1 of the 2V staged ones: first half of them in the a vector, second half in b
            auto pick = [&](const uint4& a, const uint4& b, int k) {
                const uint32_t raw = k < V ? vecGet<BPV>(a, k) : vecGet<BPV>(b, k - V);
                return F == F_U8 ? lut[raw] : unmapVoxel<F, false>(raw, p.ms);
            };
            const float v0 = pick(q00a, q00b, 2 * i), v1 = pick(q00a, q00b, 2 * i + 1);
            const float v2 = pick(q10a, q10b, 2 * i), v3 = pick(q10a, q10b, 2 * i + 1);
            const float v4 = pick(q01a, q01b, 2 * i), v5 = pick(q01a, q01b, 2 * i + 1);
            const float v6 = pick(q11a, q11b, 2 * i), v7 = pick(q11a, q11b, 2 * i + 1);
            const float f = refLerp(refLerp(refLerp(v0, v1, fx[i]), refLerp(v2, v3, fx[i]), fy),
                                    refLerp(refLerp(v4, v5, fx[i]), refLerp(v6, v7, fx[i]), fy), fz);
            vecSet<BPV>(w, i, mapVoxel<F, false>(f, p.md));
        }
        stStream(p.d + (r * p.Dx + x0) * BPV, make_uint4(w[0], w[1], w[2], w[3]));
    }
}

template <int F>
cudaError_t launchHalveX(const LaunchCtx& c, const RsParams& p, int mode)
{
    RowsLaunch l = rowsLaunch(p.Dx, (long long)p.Dy * p.dzCount, 16 / bytesPerVoxel(F));
    if (mode == MODE_LINEAR8) resampleHalveXKernel<F, MODE_LINEAR8><<<l.grid, l.block, 0, c.stream>>>(p);
    else resampleHalveXKernel<F, MODE_TRILINEAR><<<l.grid, l.block, 0, c.stream>>>(p);
    countLaunch();
    return cudaGetLastError();
}

// the reference's coordinate expression (Resample_serial.hpp:60-62, :67), one rounding per operation, on the host
static int srcCoordOnHost(int d, int D, int S)
{
    volatile float q = (float)d / (float)D;
    volatile float sf = q * (float)S;
    const float sv = sf;
    return (int)sv;
}

// every src index of the axis is the image of some dst index (then consecutive images differ by at most one)
static bool axisMapOnto(int D, int S)
{
    if (D < S || D > (1 << 22)) return false;
    int prev = srcCoordOnHost(0, D, S);
    if (prev != 0) return false;
    for (int d = 1; d < D; ++d) {
        const int cur = srcCoordOnHost(d, D, S);
        if (cur != prev && cur != prev + 1) return false;
        prev = cur;
    }
    return prev == S - 1;
}

// the reference's x coordinate expression (Resample_serial.hpp:60), one rounding per operation, on the host
bool xMapIsExactHalving(int Dx, int Sx)
{
    if (Sx != 2 * Dx || Dx > (1 << 20)) return false;
    for (int x = 0; x < Dx; ++x) {
        volatile float q = (float)x / (float)Dx;
        volatile float sxf = q * (float)Sx;
        const float sv = sxf;
        if ((int)sv != 2 * x || (int)(sv + 1.f) != 2 * x + 1) return false;
    }
    return true;
}

template <int FS, int FD>
cudaError_t launchMode(const LaunchCtx& c, const RsParams& p, int mode)
{
    constexpr int V = 16 / bytesPerVoxel(FD);
    RowsLaunch l = rowsLaunch(p.Dx, (long long)p.Dy * p.dzCount, V);
    switch (mode) {
    case MODE_NEAREST:   resampleKernel<FS, FD, MODE_NEAREST><<<l.grid, l.block, 0, c.stream>>>(p); break;
    case MODE_LINEAR8:   resampleKernel<FS, FD, MODE_LINEAR8><<<l.grid, l.block, 0, c.stream>>>(p); break;
    default:             resampleKernel<FS, FD, MODE_TRILINEAR><<<l.grid, l.block, 0, c.stream>>>(p); break;
    }
    countLaunch();
    return cudaGetLastError();
}

template <int FS>
cudaError_t launchDst(const LaunchCtx& c, const RsParams& p, int fd, int mode)
{
    switch (fd) {
    case F_U8:  return launchMode<FS, F_U8>(c, p, mode);
    case F_U16: return launchMode<FS, F_U16>(c, p, mode);
    case F_U32: return launchMode<FS, F_U32>(c, p, mode);
    case F_F32: return launchMode<FS, F_F32>(c, p, mode);
    }
    return cudaErrorInvalidValue;
}

// see the header comment: Linear == Nearest bit for bit under these conditions
bool linearIsNearest(const VolDesc& src)
{
    if (src.format == F_F32) return false;
    const float e[2] = { src.map.lo, src.map.hi };
    for (int i = 0; i < 2; ++i) {
        if (!std::isfinite(e[i]) || std::fabs(e[i]) >= 1e38f) return false;
        if (std::signbit(e[i]) && std::fabs(e[i]) < 1e-30f) return false;
    }
    return true;
}

} // namespace

// ---- Float32 sources under reference-faithful Linear ---------------------------------------------------------------------
// Resample_serial passes truncated coordinates to sampleLinear, so every fraction is 0 and the nested lerps are
// (1 * a) + (0 * b) at every level (StructuredVolumeView.hpp:112-118).  For a FINITE b the product is +-0, and a + (+-0)
// is a itself unless a is -0 (-0 + +0 = +0): with no Inf, no NaN and no -0 among the taps the 8-tap chain returns its
// first tap bit for bit, i.e. Linear == Nearest.  One pass over the src window decides it; the one-tap row pipeline
// (83-85 % of the HBM peak, against 56 % for the 8-tap march when up-sampling) then produces the result, and the
// exact march kernel is launched GATED on the flag: it returns at once for clean data and recomputes every voxel
// otherwise.  Costs one extra read of the source.
__global__ void __launch_bounds__(256)
floatTapScanKernel(const uint4* __restrict__ v, size_t nVec, const uint32_t* __restrict__ tail, int nTail, unsigned* flag)
{
    auto bad = [](uint32_t w) { return (w & 0x7f800000u) == 0x7f800000u || w == 0x80000000u; };   // Inf / NaN / -0
    bool any = false;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < nVec; i += 4 * stride) {               // four 128-bit loads in flight
        const uint4 a = ldStream(v + i), b = ldStream(v + i + stride), c = ldStream(v + i + 2 * stride), d = ldStream(v + i + 3 * stride);
        any |= bad(a.x) | bad(a.y) | bad(a.z) | bad(a.w) | bad(b.x) | bad(b.y) | bad(b.z) | bad(b.w)
             | bad(c.x) | bad(c.y) | bad(c.z) | bad(c.w) | bad(d.x) | bad(d.y) | bad(d.z) | bad(d.w);
    }
    for (; i < nVec; i += stride) {
        const uint4 a = ldStream(v + i);
        any |= bad(a.x) | bad(a.y) | bad(a.z) | bad(a.w);
    }
    if (blockIdx.x == 0 && (int)threadIdx.x < nTail) any |= bad(tail[threadIdx.x]);
    if (__syncthreads_or(any) && threadIdx.x == 0) atomicOr(flag, 1u);
}

static cudaError_t launchFloatTapScan(const LaunchCtx& c, const uint8_t* data, size_t nFloats, unsigned* flag)
{
    // data is 16-byte aligned (checked by the caller); whole vectors, then up to three trailing floats
    const size_t nVec = nFloats / 4;
    const int nTail = (int)(nFloats % 4);
    size_t blocks = (nVec + 256 * 4 - 1) / (256 * 4);
    const size_t cap = (size_t)c.smCount * 8;
    if (blocks > cap) blocks = cap;
    if (blocks == 0) blocks = 1;
    floatTapScanKernel<<<(unsigned)blocks, 256, 0, c.stream>>>(reinterpret_cast<const uint4*>(data), nVec,
                                                               reinterpret_cast<const uint32_t*>(data) + nVec * 4, nTail, flag);
    countLaunch();
    return cudaGetLastError();
}

cudaError_t launchResample(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src, int filter,
                           int dstZ0, const int32_t D[3], int srcZ0, const int32_t S[3])
{
    const int dzCount = dst.dims[2], szCount = src.dims[2];
    if (D[0] <= 0 || D[1] <= 0 || dzCount <= 0) return cudaSuccess;
    if (S[0] <= 0 || S[1] <= 0 || S[2] <= 0) return cudaErrorInvalidValue;
    const int bs = bytesPerVoxel(src.format);

    if (D[0] == S[0] && D[1] == S[1] && D[2] == S[2]) {
        // format / mapping conversion only; dst plane z <- src plane z
        const long long plane = (long long)S[0] * S[1];
        const uint8_t* s0 = src.data + (long long)(dstZ0 - srcZ0) * plane * bs;
        const size_t n = (size_t)plane * dzCount;
        cudaError_t e = launchConvertFlat(c, dst, dst.data, src, s0, n);
        if (e != cudaErrorNotSupported) return e;
        // unaligned window: generic converting rows kernel over the window as one volume (always by value:
        // the reference goes through float here even when format and mapping agree)
        VolDesc dv = dst, sv = src;
        dv.dims[0] = D[0]; dv.dims[1] = D[1]; dv.dims[2] = dzCount;
        sv.data = const_cast<uint8_t*>(s0); sv.dims[0] = S[0]; sv.dims[1] = S[1]; sv.dims[2] = dzCount;
        const int32_t f[3] = { 0, 0, 0 }, l[3] = { D[0], D[1], dzCount };
        return launchConvertRows(c, dv, sv, f, l, f);
    }

    int mode = filter == 0 ? MODE_NEAREST : filter == 1 ? MODE_LINEAR8 : MODE_TRILINEAR;
    if (mode == MODE_LINEAR8 && linearIsNearest(src)) mode = MODE_NEAREST;

    if (mode == MODE_NEAREST) {
        // fast path: TMA row pipeline (resample_rows.cu); falls through when rows are not 16-byte aligned
        cudaError_t fe = launchResampleRowRuns(c, dst, src, dstZ0, D, srcZ0, S);
        if (fe != cudaErrorNotSupported) return fe;
    }
    // 8-tap modes, exact 2:1 in x, same format, 16-byte aligned rows: coalesced 4-row gather (below)
    const int bd = bytesPerVoxel(dst.format);
    const bool halveX = mode != MODE_NEAREST && src.format == dst.format && src.format != F_INT16
                     && (D[0] * bd) % 16 == 0 && ((long long)S[0] * bs) % 16 == 0
                     && ((reinterpret_cast<uintptr_t>(src.data) | reinterpret_cast<uintptr_t>(dst.data)) & 15u) == 0
                     && xMapIsExactHalving(D[0], S[0]);
    if (mode == MODE_LINEAR8 && src.format == F_F32 && !halveX && !getenv("VKTB_RS_NO_GATE")
        && (reinterpret_cast<uintptr_t>(src.data) & 15u) == 0 && ((long long)S[0] * bs) % 16 == 0) {
        // Float32, reference-faithful Linear: tap check -> one-tap row pipeline -> the exact march, gated (see above).
        // When every axis map is ONTO (up-sampling; proved here with the device's float arithmetic) the row pipeline
        // stages every src row of the planes [zLo, zHi] the dst planes map to, whole, and checks the values it stages
        // itself (no extra pass over the source); the planes of the window outside [zLo, zHi] -- the +1 tap plane of
        // the last dst plane, the spill plane -- get the small separate scan.  Otherwise the whole window is scanned.
        unsigned* flag = nullptr;
        cudaError_t e = cudaMallocAsync((void**)&flag, sizeof(unsigned), c.stream);
        if (e != cudaSuccess) return e;
        e = cudaMemsetAsync(flag, 0, sizeof(unsigned), c.stream);
        const bool fused = !getenv("VKTB_RS_SEPARATE_SCAN") && axisMapOnto(D[0], S[0]) && axisMapOnto(D[1], S[1])
                        && axisMapOnto(D[2], S[2]);
        const size_t planeFloats = (size_t)S[0] * S[1];
        if (e == cudaSuccess && !fused) e = launchFloatTapScan(c, src.data, planeFloats * (size_t)szCount, flag);
        if (e == cudaSuccess && fused) {
            const int zLo = srcCoordOnHost(dstZ0, D[2], S[2]) - srcZ0, zHi = srcCoordOnHost(dstZ0 + dzCount - 1, D[2], S[2]) - srcZ0;
            const int a1 = zLo < 0 ? 0 : (zLo > szCount ? szCount : zLo);                 // window planes [0, a1)
            const int b0 = zHi + 1 < 0 ? 0 : (zHi + 1 > szCount ? szCount : zHi + 1);     // and [b0, szCount)
            if (a1 > 0) e = launchFloatTapScan(c, src.data, planeFloats * (size_t)a1, flag);
            if (e == cudaSuccess && b0 < szCount)
                e = launchFloatTapScan(c, src.data + planeFloats * (size_t)b0 * 4, planeFloats * (size_t)(szCount - b0), flag);
        }
        cudaError_t fe = e == cudaSuccess ? launchResampleRowRuns(c, dst, src, dstZ0, D, srcZ0, S, fused ? flag : nullptr) : e;
        if (fe == cudaSuccess) {
            fe = launchResampleMarch(c, dst, src, mode, dstZ0, D, srcZ0, S, flag);
            // (when the march does not qualify either, the ungated paths below overwrite the one-tap result)
        }
        cudaFreeAsync(flag, c.stream);
        if (fe != cudaErrorNotSupported) return fe;
    }
    if (mode != MODE_NEAREST && !halveX) {
        // 8-tap modes: TMA z-march (resample_march.cu); falls through when the operands do not qualify
        cudaError_t fe = launchResampleMarch(c, dst, src, mode, dstZ0, D, srcZ0, S);
        if (fe != cudaErrorNotSupported) return fe;
    }

    // coordinate tables (stream-ordered scratch)
    const size_t nTab = (size_t)D[0] + D[1] + dzCount;
    int* tab = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&tab, nTab * 3 * sizeof(int), c.stream);
    if (e != cudaSuccess) return e;
    int* hiTab = tab + nTab;
    float* frTab = reinterpret_cast<float*>(tab + 2 * nTab);
    const size_t oy = (size_t)D[0], oz = (size_t)D[0] + D[1];
    RsParams p;
    p.d = dst.data; p.s = src.data;
    p.Dx = D[0]; p.Dy = D[1]; p.dzCount = dzCount;
    p.Sx = S[0]; p.Sy = S[1]; p.Sz = S[2]; p.sz0 = srcZ0; p.szCount = szCount;
    p.xmap = tab;   p.ymap = tab + oy;   p.zmap = tab + oz;
    p.xhi = hiTab;  p.yhi = hiTab + oy;  p.zhi = hiTab + oz;
    p.xfr = frTab;  p.yfr = frTab + oy;  p.zfr = frTab + oz;
    p.md = dst.map; p.ms = src.map;
    const int tri = mode == MODE_TRILINEAR;
    coordTableKernel<<<(D[0] + 255) / 256, 256, 0, c.stream>>>(tab, hiTab, frTab, D[0], S[0], 0, D[0], tri);
    coordTableKernel<<<(D[1] + 255) / 256, 256, 0, c.stream>>>(tab + oy, hiTab + oy, frTab + oy, D[1], S[1], 0, D[1], tri);
    coordTableKernel<<<(dzCount + 255) / 256, 256, 0, c.stream>>>(tab + oz, hiTab + oz, frTab + oz, D[2], S[2], dstZ0, dzCount, tri);
    countLaunch(); countLaunch(); countLaunch();

    if (halveX) {
        switch (src.format) {
        case F_U8:  e = launchHalveX<F_U8>(c, p, mode); break;
        case F_U16: e = launchHalveX<F_U16>(c, p, mode); break;
        case F_U32: e = launchHalveX<F_U32>(c, p, mode); break;
        case F_F32: e = launchHalveX<F_F32>(c, p, mode); break;
        default: e = cudaErrorInvalidValue;
        }
        cudaFreeAsync(tab, c.stream);
        return e;
    }
    switch (src.format) {
    case F_U8:  e = launchDst<F_U8>(c, p, dst.format, mode); break;
    case F_U16: e = launchDst<F_U16>(c, p, dst.format, mode); break;
    case F_U32: e = launchDst<F_U32>(c, p, dst.format, mode); break;
    case F_F32: e = launchDst<F_F32>(c, p, dst.format, mode); break;
    default: e = cudaErrorInvalidValue;
    }
    cudaFreeAsync(tab, c.stream);
    return e;
}

} // namespace vktb
