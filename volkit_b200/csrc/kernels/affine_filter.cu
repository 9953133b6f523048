// Scale / Rotate (inverse-mapped affine gather) and ApplyFilter (small dense convolution).
//
// Scale / Rotate -- reference src/vkt/Scale_serial.hpp:15-56, Rotate_serial.hpp:41-94.  Both walk the WHOLE dst
// volume, map every dst voxel through p' = M (p - c) + c (binary32, the sum of the three products evaluated left
// to right: linalg.hpp:584-591) and, when p' lies inside [first, last) (float compares against the converted
// integers), store sampleLinear(p'.x, p'.y, p'.z) -- whose parameters are int32_t, so the coordinates are truncated
// and the "linear" filter degenerates to the reference-faithful 8-tap expression with frac = 0 (the same
// body Resample uses).  Voxels that map outside the range keep their dst value.  The matrix is built on the host
// (1 / scalingFactor on the diagonal; quaternion -> matrix with the host libm's sinf / cosf, exactly the
// reference's expressions) and arrives here as nine floats.
//
// ApplyFilter -- reference src/vkt/Filter_serial.hpp:15-64: over [0, src dims): voxels with a coordinate equal to 0
// or last - 1 receive map(0); every other voxel receives sum_{zz,yy,xx} f(xx,yy,zz) * src(x - fx/2 + xx, ...),
// accumulated in binary32 from 0 in exactly that order (multiply, round, add, round -- no FMA); the address mode is
// ignored.  The reference reads a copy of the source, so dst may alias src (the launcher copies then).  Filter dims
// are limited to 1..3 per axis: larger ones make the reference read outside the volume.
//   3x3x3, same format on both sides: z-march -- a thread owns one (x, y) column of a z chunk, keeps the unmapped
//   3x3 neighbourhoods of two planes in registers and reads nine voxels per output through the read-only L1 path
//   (neighbouring threads share them); 27 FMUL + 27 FADD per voxel in the reference's order make this kernel
//   CUDA-core bound, not HBM bound.
//   anything else: one thread per voxel, runtime codec (parity path).
// Algorithmic bytes: bpv_src + bpv_dst per voxel for both.
#include "launch.hpp"

#include <cmath>
#include <cstdlib>

namespace vktb {
namespace {

// ---------------------------------------------------------------------------------------------------------------
// affine gather
// ---------------------------------------------------------------------------------------------------------------
struct AffineParams
{
    uint8_t*       d;
    const uint8_t* s;
    int Dx, Dy, Dz;
    int Sx, Sy, Sz;
    float m[9];              // col0.xyz, col1.xyz, col2.xyz
    float c[3];
    float first[3], last[3]; // the range, converted int -> float like the reference's compares do
    Mapping md, ms;
    int diag;                // off-diagonal matrix entries are all +-0
    int fastTrunc;           // the range lies inside [0, 2^23)
    int zsep;                // m[2], m[5], m[6], m[7] are all +-0: z does not mix with x, y (a rotation about z)
};

template <int FS>
__device__ __forceinline__ float affTap(const AffineParams& p, long long idx, long long count)
{
    constexpr int BS = bytesPerVoxel(FS);
    // a tap outside the buffer (the unclamped hi.x of the volume's last voxel, an unclamped lo.z) reads as a zero voxel
    const uint32_t raw = (idx >= 0 && idx < count) ? loadRaw(p.s + idx * BS, BS) : 0u;
    return unmapVoxel<FS, false>(raw, p.ms);
}

// StructuredVolumeView::sampleLinear(int32_t, int32_t, int32_t), src/vkt/StructuredVolumeView.hpp:80-119, clamps as written
template <int FS>
__device__ __forceinline__ float sampleLinearRef(const AffineParams& p, int x, int y, int z)
{
    const float xf1 = __int2float_rn(x), yf1 = __int2float_rn(y), zf1 = __int2float_rn(z);
    int lx = (int)x86Cvtt32(xf1), ly = (int)x86Cvtt32(yf1);
    const int lz = (int)x86Cvtt32(zf1);
    const int hx = (int)x86Cvtt32(__fadd_rn(xf1, 1.f));
    int hy = (int)x86Cvtt32(__fadd_rn(yf1, 1.f)), hz = (int)x86Cvtt32(__fadd_rn(zf1, 1.f));
    lx = min(max(lx, 0), p.Sx - 1); ly = min(max(ly, 0), p.Sy - 1);
    hy = min(max(hy, 0), p.Sy - 1); hz = min(max(hz, 0), p.Sz - 1);
    const float fx = __fsub_rn(xf1, __int2float_rn(lx)), fy = __fsub_rn(yf1, __int2float_rn(ly));
    const float fz = __fsub_rn(zf1, __int2float_rn(lz));
    const long long plane = (long long)p.Sx * p.Sy, count = plane * p.Sz;
    const long long r00 = lz * plane + (long long)ly * p.Sx, r10 = lz * plane + (long long)hy * p.Sx;
    const long long r01 = hz * plane + (long long)ly * p.Sx, r11 = hz * plane + (long long)hy * p.Sx;
    const float v0 = affTap<FS>(p, r00 + lx, count), v1 = affTap<FS>(p, r00 + hx, count);
    const float v2 = affTap<FS>(p, r10 + lx, count), v3 = affTap<FS>(p, r10 + hx, count);
    const float v4 = affTap<FS>(p, r01 + lx, count), v5 = affTap<FS>(p, r01 + hx, count);
    const float v6 = affTap<FS>(p, r11 + lx, count), v7 = affTap<FS>(p, r11 + hx, count);
    return refLerp(refLerp(refLerp(v0, v1, fx), refLerp(v2, v3, fx), fy),
                   refLerp(refLerp(v4, v5, fx), refLerp(v6, v7, fx), fy), fz);
}

// float -> int of a coordinate that passed the range test.  fast: the range lies in [0, 2^23), where adding 2^23 with
// round-toward-zero leaves floor(q) in the mantissa (an FP32-pipe add instead of the conversion pipe's F2I).
__device__ __forceinline__ int truncCoord(float q, bool fast)
{
    return fast ? (int)(__float_as_uint(__fadd_rz(q, 8388608.f)) & 0x7fffffu) : (int)x86Cvtt32(q);
}

// one thread = V consecutive dst voxels of one row.
//   MODE 0  the reference's 8-tap expression (Float32 sources, ranges that leave the source, exotic mappings)
//   MODE 1  one tap, unmap -> map: the range lies inside the source and every tap of an integer source unmaps to a
//           finite float, so the 8-tap chain with frac 0 returns the lo tap bit for bit
//   MODE 2  one tap, raw voxel moved: MODE 1 with map(unmap(raw)) == raw for every raw (same format, UInt8 / UInt16
//           with the identity mapping on both sides)
// p.diag: the matrix is diagonal (Scale).  The off-diagonal products are +-0 and only flip the sign of a zero, which
// neither the range test nor the truncation can see, so q.y and q.z are computed once per row (rows that map
// outside the range are skipped) and q.x costs one multiply and one add per voxel.
template <int FS, int FD, int MODE>
__global__ void __launch_bounds__(kThreads) affineKernel(AffineParams p)
{
    constexpr int BS = bytesPerVoxel(FS), BD = bytesPerVoxel(FD);
    constexpr int V = 16 / BD;
    const int chunks = (p.Dx + V - 1) / V;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= chunks) return;
    const int x0 = c * V, n = min(V, p.Dx - x0);
    const long long rows = (long long)p.Dy * p.Dz;
    const long long plane = (long long)p.Sx * p.Sy;
    const bool fast = p.fastTrunc != 0, diag = p.diag != 0;
    const float fx0 = __int2float_rn(x0);      // x0 + i < 2^24: float(x0) + float(i) is exact
    for (long long r = (long long)blockIdx.y * blockDim.y + threadIdx.y; r < rows; r += (long long)gridDim.y * blockDim.y) {
        const int z = (int)(r / p.Dy), y = (int)(r - (long long)z * p.Dy);
        const float py = __fsub_rn(__int2float_rn(y), p.c[1]), pz = __fsub_rn(__int2float_rn(z), p.c[2]);
        float rqy = 0.f, rqz = 0.f;
        long long rowBase = 0;
        if (diag) {
            rqy = __fadd_rn(__fmul_rn(p.m[4], py), p.c[1]);
            rqz = __fadd_rn(__fmul_rn(p.m[8], pz), p.c[2]);
            if (!(rqy >= p.first[1] && rqy < p.last[1] && rqz >= p.first[2] && rqz < p.last[2])) continue;
            if (MODE != 0) rowBase = truncCoord(rqz, fast) * plane + (long long)truncCoord(rqy, fast) * p.Sx;
        }
        // z separable (rotation about z): q.z once per row, the z column's +-0 products dropped from q.x and q.y (they
        // can only flip the sign of a zero, which neither the range test nor the truncation sees)
        const bool zsep = p.zsep != 0 && !diag;
        float ax = 0.f, ay = 0.f;
        if (zsep) {
            rqz = __fadd_rn(__fmul_rn(p.m[8], pz), p.c[2]);
            if (!(rqz >= p.first[2] && rqz < p.last[2])) continue;
            ax = __fmul_rn(p.m[3], py); ay = __fmul_rn(p.m[4], py);
            if (MODE != 0) rowBase = truncCoord(rqz, fast) * plane;
        }
        uint8_t* dp = p.d + (r * p.Dx + x0) * BD;
        uint32_t out[V];
        unsigned inside = 0;
#pragma unroll
        for (int i = 0; i < V; ++i) {
            if (i >= n) continue;
            const float px = __fsub_rn(__fadd_rn(fx0, (float)i), p.c[0]);
            float qx, qy, qz;
            if (diag) {
                qx = __fadd_rn(__fmul_rn(p.m[0], px), p.c[0]); qy = rqy; qz = rqz;
                if (!(qx >= p.first[0] && qx < p.last[0])) continue;
            } else if (zsep) {
                qx = __fadd_rn(__fadd_rn(__fmul_rn(p.m[0], px), ax), p.c[0]);
                qy = __fadd_rn(__fadd_rn(__fmul_rn(p.m[1], px), ay), p.c[1]);
                qz = rqz;
                if (!(qx >= p.first[0] && qx < p.last[0] && qy >= p.first[1] && qy < p.last[1])) continue;
            } else {
                // (a*x + b*y) + c*z, then + center
                qx = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(p.m[0], px), __fmul_rn(p.m[3], py)), __fmul_rn(p.m[6], pz)), p.c[0]);
                qy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(p.m[1], px), __fmul_rn(p.m[4], py)), __fmul_rn(p.m[7], pz)), p.c[1]);
                qz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(p.m[2], px), __fmul_rn(p.m[5], py)), __fmul_rn(p.m[8], pz)), p.c[2]);
                if (!(qx >= p.first[0] && qx < p.last[0] && qy >= p.first[1] && qy < p.last[1] && qz >= p.first[2] && qz < p.last[2]))
                    continue;
            }
            const int sx = truncCoord(qx, fast);
            if (MODE == 0) {
                out[i] = mapVoxel<FD, false>(sampleLinearRef<FS>(p, sx, truncCoord(qy, fast), truncCoord(qz, fast)), p.md);
            } else {
                const long long idx = diag ? rowBase + sx
                                    : zsep ? rowBase + (long long)truncCoord(qy, fast) * p.Sx + sx
                                           : truncCoord(qz, fast) * plane + (long long)truncCoord(qy, fast) * p.Sx + sx;
                const uint32_t raw = loadRaw(p.s + idx * BS, BS);
                out[i] = MODE == 2 ? raw : mapVoxel<FD, false>(unmapVoxel<FS, false>(raw, p.ms), p.md);
            }
            inside |= 1u << i;
        }
        if (inside == (1u << V) - 1u && (reinterpret_cast<uintptr_t>(dp) & 15u) == 0) {
            uint32_t w[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
            for (int i = 0; i < V; ++i) vecSet<BD>(w, i, out[i]);
            stStream(dp, make_uint4(w[0], w[1], w[2], w[3]));
        } else if (inside) {
#pragma unroll
            for (int i = 0; i < V; ++i)
                if (inside & (1u << i)) storeRaw(dp + i * BD, BD, out[i]);
        }
    }
}

bool oneTapExact(const VolDesc& src)
{
    if (src.format == F_F32) return false;
    const float e[2] = { src.map.lo, src.map.hi };
    for (int i = 0; i < 2; ++i) {
        if (!std::isfinite(e[i]) || std::fabs(e[i]) >= 1e38f) return false;
        if (std::signbit(e[i]) && std::fabs(e[i]) < 1e-30f) return false;
    }
    return true;
}

template <int FS, int FD>
cudaError_t launchAffineFF(const LaunchCtx& c, const AffineParams& p, bool oneTap)
{
    constexpr int V = 16 / bytesPerVoxel(FD);
    RowsLaunch l = rowsLaunch(p.Dx, (long long)p.Dy * p.Dz, V);
    if (!p.diag) {
        // a general matrix maps a dst row to a slanted src line: a warp that covers 4 chunks x 8 rows gathers from a
        // compact src patch instead of a 256-voxel line (fewer sectors per request, better L1 / L2 hit rates)
        const int chunks = (p.Dx + V - 1) / V;
        const long long rows = (long long)p.Dy * p.Dz;
        long long gy = (rows + 63) / 64;
        if (gy > 65535) gy = 65535;
        l.block = dim3(4, 64, 1);
        l.grid = dim3((unsigned)((chunks + 3) / 4), (unsigned)gy, 1);
    }
    const bool rawMove = oneTap && FS == FD && (FS == F_U8 || FS == F_U16) && p.ms.identity && p.md.identity;
    if (rawMove) affineKernel<FS, FD, 2><<<l.grid, l.block, 0, c.stream>>>(p);
    else if (oneTap) affineKernel<FS, FD, 1><<<l.grid, l.block, 0, c.stream>>>(p);
    else affineKernel<FS, FD, 0><<<l.grid, l.block, 0, c.stream>>>(p);
    countLaunch();
    return cudaGetLastError();
}

template <int FS>
cudaError_t launchAffineF(const LaunchCtx& c, const AffineParams& p, int fd, bool oneTap)
{
    switch (fd) {
    case F_U8:    return launchAffineFF<FS, F_U8>(c, p, oneTap);
    case F_U16:   return launchAffineFF<FS, F_U16>(c, p, oneTap);
    case F_INT16: return launchAffineFF<FS, F_INT16>(c, p, oneTap);
    case F_U32:   return launchAffineFF<FS, F_U32>(c, p, oneTap);
    case F_F32:   return launchAffineFF<FS, F_F32>(c, p, oneTap);
    }
    return cudaErrorInvalidValue;
}

// ---------------------------------------------------------------------------------------------------------------
// ApplyFilter
// ---------------------------------------------------------------------------------------------------------------
struct FilterParams
{
    uint8_t*       d;
    const uint8_t* s;
    int nx, ny, nz;          // the range [0, src dims)
    int Dx, Dy;              // dst row length / rows per plane (setValue indexes with the dst dims)
    int fx, fy, fz;          // filter dims, 1..3
    int zChunk;
    int pairs;               // 3x3x3 march: a thread owns two output rows (filter3PairKernel)
    float w[27];             // f(xx, yy, zz) at [zz * fx * fy + yy * fx + xx]
    Mapping md, ms;
    int fs, fd;              // formats (generic kernel)
};

// nine unmapped voxels around (x, y) in plane z
template <int F, bool IDENT>
__device__ __forceinline__ void loadPlane9(float (&v)[9], const FilterParams& p, const float* lut, int x, int y, int z)
{
    constexpr int B = bytesPerVoxel(F);
    const uint8_t* row = p.s + (((long long)z * p.ny + (y - 1)) * p.nx + (x - 1)) * B;
#pragma unroll
    for (int yy = 0; yy < 3; ++yy) {
#pragma unroll
        for (int xx = 0; xx < 3; ++xx) {
            uint32_t raw;
            if (B == 1) raw = __ldg(row + xx);
            else if (B == 2) raw = __ldg(reinterpret_cast<const uint16_t*>(row) + xx);
            else raw = __ldg(reinterpret_cast<const uint32_t*>(row) + xx);
            v[yy * 3 + xx] = F == F_U8 ? lut[raw] : unmapVoxel<F, IDENT>(raw, p.ms);   // UInt8: the unmap divides
        }
        row += (long long)p.nx * B;
    }
}

__device__ __forceinline__ float accPlane9(float acc, const float (&v)[9], const float* w)
{
#pragma unroll
    for (int i = 0; i < 9; ++i) acc = __fadd_rn(acc, __fmul_rn(w[i], v[i]));
    return acc;
}

// 3x3x3, src format == dst format.  block (32, 8): a 32 x 8 tile of columns; blockIdx.z: z chunk
template <int F, bool IDS, bool IDD>
__global__ void __launch_bounds__(kThreads) filter3MarchKernel(FilterParams p)
{
    constexpr int B = bytesPerVoxel(F);
    __shared__ float lut[F == F_U8 ? 256 : 1];
    if (F == F_U8) {
        const int t = threadIdx.y * 32 + threadIdx.x;
        lut[t] = unmapVoxel<F_U8, IDS>((uint32_t)t, p.ms);
        __syncthreads();
    }
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    if (x >= p.nx || y >= p.ny) return;
    const int z0 = blockIdx.z * p.zChunk, z1 = min(z0 + p.zChunk, p.nz);
    const uint32_t zero = mapVoxel<F, IDD>(0.f, p.md);
    uint8_t* dcol = p.d + ((long long)y * p.Dx + x) * B;
    const long long dPlane = (long long)p.Dx * p.Dy * B;
    auto put = [&](int z, uint32_t raw) {
        uint8_t* q = dcol + z * dPlane;
        if (B == 1) *q = (uint8_t)raw;
        else if (B == 2) *reinterpret_cast<uint16_t*>(q) = (uint16_t)raw;
        else *reinterpret_cast<uint32_t*>(q) = raw;
    };
    if (x == 0 || y == 0 || x == p.nx - 1 || y == p.ny - 1) {
        for (int z = z0; z < z1; ++z) put(z, zero);
        return;
    }
    int zs = z0, ze = z1;
    if (zs == 0) { put(0, zero); zs = 1; }
    if (ze == p.nz) { if (p.nz - 1 >= zs) put(p.nz - 1, zero); ze = p.nz - 1; }
    if (zs >= ze) return;
    float a[9], b[9], cc[9];
    loadPlane9<F, IDS>(a, p, lut, x, y, zs - 1);
    loadPlane9<F, IDS>(b, p, lut, x, y, zs);
    int z = zs;
    // three planes per round so that the register planes rotate by renaming
    for (; z + 2 < ze; z += 3) {
        loadPlane9<F, IDS>(cc, p, lut, x, y, z + 1);
        put(z, mapVoxel<F, IDD>(accPlane9(accPlane9(accPlane9(0.f, a, p.w), b, p.w + 9), cc, p.w + 18), p.md));
        loadPlane9<F, IDS>(a, p, lut, x, y, z + 2);
        put(z + 1, mapVoxel<F, IDD>(accPlane9(accPlane9(accPlane9(0.f, b, p.w), cc, p.w + 9), a, p.w + 18), p.md));
        loadPlane9<F, IDS>(b, p, lut, x, y, z + 3);
        put(z + 2, mapVoxel<F, IDD>(accPlane9(accPlane9(accPlane9(0.f, cc, p.w), a, p.w + 9), b, p.w + 18), p.md));
    }
    for (; z < ze; ++z) {
        loadPlane9<F, IDS>(cc, p, lut, x, y, z + 1);
        put(z, mapVoxel<F, IDD>(accPlane9(accPlane9(accPlane9(0.f, a, p.w), b, p.w + 9), cc, p.w + 18), p.md));
#pragma unroll
        for (int i = 0; i < 9; ++i) { a[i] = b[i]; b[i] = cc[i]; }
    }
}

// ---- 3x3x3 as row PAIRS ------------------------------------------------------------------------------------------
// A thread owns the outputs (x, ya) and (x, ya + 1) of a z chunk: four source rows of three voxels per plane serve both
// (6 loads and unmaps per output instead of 9), and the two accumulators travel as one f32x2 register pair: per tap two
// scalar multiplies (rounded, as the reference's `f * v`) and ONE packed add (FADD2: two independent roundings).  The
// multiplies stay scalar on purpose: ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even under --fmad=false.
__device__ __forceinline__ unsigned long long pack2(float lo, float hi)
{
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b)
{
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

// twelve unmapped voxels: rows ya - 1 .. ya + 2 (the last one clamped to the volume), columns x - 1 .. x + 1
template <int F, bool IDENT>
__device__ __forceinline__ void loadPlane12(float (&v)[12], const FilterParams& p, const float* lut, int x, int ya, int z)
{
    constexpr int B = bytesPerVoxel(F);
    const long long rowBytes = (long long)p.nx * B;
    const uint8_t* row = p.s + (((long long)z * p.ny + (ya - 1)) * p.nx + (x - 1)) * B;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const uint8_t* q = (r == 3 && ya + 2 > p.ny - 1) ? row + 2 * rowBytes : row + r * rowBytes;
#pragma unroll
        for (int xx = 0; xx < 3; ++xx) {
            uint32_t raw;
            if (B == 1) raw = __ldg(q + xx);
            else if (B == 2) raw = __ldg(reinterpret_cast<const uint16_t*>(q) + xx);
            else raw = __ldg(reinterpret_cast<const uint32_t*>(q) + xx);
            v[r * 3 + xx] = F == F_U8 ? lut[raw] : unmapVoxel<F, IDENT>(raw, p.ms);
        }
    }
}

// nine taps of one plane for both outputs: acc.lo += w * v[yy][xx], acc.hi += w * v[yy + 1][xx]
__device__ __forceinline__ unsigned long long accPlanePair(unsigned long long acc, const float (&v)[12], const float* w)
{
#pragma unroll
    for (int yy = 0; yy < 3; ++yy)
#pragma unroll
        for (int xx = 0; xx < 3; ++xx)
            acc = add2(acc, pack2(__fmul_rn(w[yy * 3 + xx], v[yy * 3 + xx]), __fmul_rn(w[yy * 3 + xx], v[(yy + 1) * 3 + xx])));
    return acc;
}

template <int F, bool IDS, bool IDD>
__global__ void __launch_bounds__(128) filter3PairKernel(FilterParams p)
{
    constexpr int B = bytesPerVoxel(F);
    __shared__ float lut[F == F_U8 ? 256 : 1];
    if (F == F_U8) {
        for (int t = threadIdx.y * 32 + threadIdx.x; t < 256; t += 128) lut[t] = unmapVoxel<F_U8, IDS>((uint32_t)t, p.ms);
        __syncthreads();
    }
    const int x = blockIdx.x * 32 + threadIdx.x, j = blockIdx.y * 4 + threadIdx.y;      // block (32, 4): 72 registers x 128
    const int ya = 1 + 2 * j, yb = ya + 1;
    if (x >= p.nx) return;
    const int z0 = blockIdx.z * p.zChunk, z1 = min(z0 + p.zChunk, p.nz);
    const uint32_t zero = mapVoxel<F, IDD>(0.f, p.md);
    const long long dPlane = (long long)p.Dx * p.Dy * B;
    auto put = [&](int y, int z, uint32_t raw) {
        uint8_t* q = p.d + ((long long)y * p.Dx + x) * B + z * dPlane;
        if (B == 1) *q = (uint8_t)raw;
        else if (B == 2) *reinterpret_cast<uint16_t*>(q) = (uint16_t)raw;
        else *reinterpret_cast<uint32_t*>(q) = raw;
    };
    if (j == 0)
        for (int z = z0; z < z1; ++z) put(0, z, zero);                 // the y = 0 border row
    if (ya > p.ny - 1) return;
    const bool aBorder = ya == p.ny - 1, bInside = yb <= p.ny - 2, bBorder = yb == p.ny - 1;
    if (aBorder || x == 0 || x == p.nx - 1) {
        for (int z = z0; z < z1; ++z) {
            put(ya, z, zero);
            if (bInside || bBorder) put(yb, z, zero);
        }
        return;
    }
    int zs = z0, ze = z1;
    auto zeroPlane = [&](int z) { put(ya, z, zero); if (bInside || bBorder) put(yb, z, zero); };
    if (zs == 0) { zeroPlane(0); zs = 1; }
    if (ze == p.nz) { if (p.nz - 1 >= zs) zeroPlane(p.nz - 1); ze = p.nz - 1; }
    if (bBorder)
        for (int z = zs; z < ze; ++z) put(yb, z, zero);
    if (zs >= ze) return;
    auto emit = [&](int z, unsigned long long acc) {
        put(ya, z, mapVoxel<F, IDD>(__uint_as_float((uint32_t)acc), p.md));
        if (bInside) put(yb, z, mapVoxel<F, IDD>(__uint_as_float((uint32_t)(acc >> 32)), p.md));
    };
    float a[12], b[12], cc[12];
    loadPlane12<F, IDS>(a, p, lut, x, ya, zs - 1);
    loadPlane12<F, IDS>(b, p, lut, x, ya, zs);
    int z = zs;
    for (; z + 2 < ze; z += 3) {
        loadPlane12<F, IDS>(cc, p, lut, x, ya, z + 1);
        emit(z, accPlanePair(accPlanePair(accPlanePair(0ull, a, p.w), b, p.w + 9), cc, p.w + 18));
        loadPlane12<F, IDS>(a, p, lut, x, ya, z + 2);
        emit(z + 1, accPlanePair(accPlanePair(accPlanePair(0ull, b, p.w), cc, p.w + 9), a, p.w + 18));
        loadPlane12<F, IDS>(b, p, lut, x, ya, z + 3);
        emit(z + 2, accPlanePair(accPlanePair(accPlanePair(0ull, cc, p.w), a, p.w + 9), b, p.w + 18));
    }
    for (; z < ze; ++z) {
        loadPlane12<F, IDS>(cc, p, lut, x, ya, z + 1);
        emit(z, accPlanePair(accPlanePair(accPlanePair(0ull, a, p.w), b, p.w + 9), cc, p.w + 18));
#pragma unroll
        for (int i = 0; i < 12; ++i) { a[i] = b[i]; b[i] = cc[i]; }
    }
}

// any filter dims in 1..3, any formats: one thread per voxel
__global__ void __launch_bounds__(kThreads) filterGenericKernel(FilterParams p)
{
    const int bs = bytesPerVoxel(p.fs), bd = bytesPerVoxel(p.fd);
    const long long n = (long long)p.nx * p.ny * p.nz;
    const int hx = p.fx / 2, hy = p.fy / 2, hz = p.fz / 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(i % p.nx);
        const long long r = i / p.nx;
        const int y = (int)(r % p.ny), z = (int)(r / p.ny);
        float value = 0.f;
        if (!(x == 0 || y == 0 || z == 0 || x == p.nx - 1 || y == p.ny - 1 || z == p.nz - 1)) {
            for (int zz = 0; zz < p.fz; ++zz)
                for (int yy = 0; yy < p.fy; ++yy)
                    for (int xx = 0; xx < p.fx; ++xx) {
                        const long long si = ((long long)(z - hz + zz) * p.ny + (y - hy + yy)) * p.nx + (x - hx + xx);
                        const float v = unmapVoxelRt(p.fs, loadRaw(p.s + si * bs, bs), p.ms);
                        value = __fadd_rn(value, __fmul_rn(p.w[(zz * p.fy + yy) * p.fx + xx], v));
                    }
        }
        storeRaw(p.d + (((long long)z * p.Dy + y) * p.Dx + x) * bd, bd, mapVoxelRt(p.fd, value, p.md));
    }
}

template <int F>
cudaError_t launchFilter3(const LaunchCtx& c, const FilterParams& p, dim3 grid)
{
    const dim3 block(32, 8, 1), pairBlock(32, 4, 1);
    const bool ids = p.ms.identity && F != F_INT16, idd = p.md.identity && F != F_INT16;
    if (p.pairs) {
        if (ids && idd) filter3PairKernel<F, true, true><<<grid, pairBlock, 0, c.stream>>>(p);
        else filter3PairKernel<F, false, false><<<grid, pairBlock, 0, c.stream>>>(p);
    } else if (ids && idd) filter3MarchKernel<F, true, true><<<grid, block, 0, c.stream>>>(p);
    else filter3MarchKernel<F, false, false><<<grid, block, 0, c.stream>>>(p);
    countLaunch();
    return cudaGetLastError();
}

} // namespace

cudaError_t launchAffine(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src, const int32_t first[3],
                         const int32_t last[3], const float m[9], const float center[3])
{
    if (dst.dims[0] <= 0 || dst.dims[1] <= 0 || dst.dims[2] <= 0) return cudaSuccess;
    if (src.dims[0] <= 0 || src.dims[1] <= 0 || src.dims[2] <= 0) return cudaErrorInvalidValue;
    AffineParams p;
    p.d = dst.data; p.s = src.data;
    p.Dx = dst.dims[0]; p.Dy = dst.dims[1]; p.Dz = dst.dims[2];
    p.Sx = src.dims[0]; p.Sy = src.dims[1]; p.Sz = src.dims[2];
    for (int i = 0; i < 9; ++i) p.m[i] = m[i];
    bool insideSrc = true;
    for (int a = 0; a < 3; ++a) {
        p.c[a] = center[a];
        p.first[a] = (float)first[a]; p.last[a] = (float)last[a];
        if (first[a] < 0 || last[a] > src.dims[a]) insideSrc = false;
    }
    p.md = dst.map; p.ms = src.map;
    p.diag = 1;
    for (int i = 0; i < 9; ++i)
        if (i % 4 != 0 && m[i] != 0.f) p.diag = 0;
    for (int i = 0; i < 9; i += 4)
        if (!std::isfinite(m[i])) p.diag = 0;         // 0 * inf would be NaN, not +-0
    for (int a = 0; a < 3; ++a)
        if (!std::isfinite(center[a])) p.diag = 0;
    p.zsep = (m[2] == 0.f && m[5] == 0.f && m[6] == 0.f && m[7] == 0.f) ? 1 : 0;
    for (int i = 0; i < 9; ++i)
        if (!std::isfinite(m[i])) p.zsep = 0;
    for (int a = 0; a < 3; ++a)
        if (!std::isfinite(center[a])) p.zsep = 0;
    p.fastTrunc = 1;
    for (int a = 0; a < 3; ++a)
        if (first[a] < 0 || last[a] > (1 << 23)) p.fastTrunc = 0;
    if (dst.dims[0] >= (1 << 24)) return cudaErrorInvalidValue;
    const bool oneTap = insideSrc && oneTapExact(src);
    if (oneTap && p.diag && !(getenv("VKTB_SCALE_ROWS") && getenv("VKTB_SCALE_ROWS")[0] == '0')) {
        // a Scale whose every dst voxel maps inside the source is a separable monotone gather: Resample's bulk-copy
        // row pipeline (resample_rows.cu) with the affine coordinate maps
        const float dm[3] = { m[0], m[4], m[8] };
        const cudaError_t e = launchScaleRowRuns(c, dst, src, dm, center, first, last);
        if (e != cudaErrorNotSupported) return e;
    }
    switch (src.format) {
    case F_U8:    return launchAffineF<F_U8>(c, p, dst.format, oneTap);
    case F_U16:   return launchAffineF<F_U16>(c, p, dst.format, oneTap);
    case F_INT16: return launchAffineF<F_INT16>(c, p, dst.format, oneTap);
    case F_U32:   return launchAffineF<F_U32>(c, p, dst.format, oneTap);
    case F_F32:   return launchAffineF<F_F32>(c, p, dst.format, oneTap);
    }
    return cudaErrorInvalidValue;
}

cudaError_t launchFilter(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src, const float* weights,
                         const int32_t fdims[3])
{
    FilterParams p;
    p.nx = src.dims[0]; p.ny = src.dims[1]; p.nz = src.dims[2];
    if (p.nx <= 0 || p.ny <= 0 || p.nz <= 0) return cudaSuccess;
    for (int a = 0; a < 3; ++a)
        if (fdims[a] < 1 || fdims[a] > 3) return cudaErrorInvalidValue;
    p.d = dst.data; p.s = src.data;
    p.Dx = dst.dims[0]; p.Dy = dst.dims[1];
    p.fx = fdims[0]; p.fy = fdims[1]; p.fz = fdims[2];
    for (int i = 0; i < 27; ++i) p.w[i] = i < p.fx * p.fy * p.fz ? weights[i] : 0.f;
    p.md = dst.map; p.ms = src.map;
    p.fs = src.format; p.fd = dst.format;
    if (p.fx == 3 && p.fy == 3 && p.fz == 3 && src.format == dst.format) {
        // z chunks: enough CTAs for every SM a few times over, at least 16 planes each (two extra plane loads per chunk)
        // row pairs unless VKTB_FILTER_PAIRS=0 (kept for A/B measurements)
        const char* pe = std::getenv("VKTB_FILTER_PAIRS");
        p.pairs = !(pe && pe[0] == '0');
        const int yTiles = p.pairs ? (p.ny / 2 + 1 + 3) / 4 : (p.ny + 7) / 8;   // pairs: blocks of 4 row pairs
        const long long tiles = (long long)((p.nx + 31) / 32) * yTiles;
        int chunks = (int)(((long long)c.smCount * 16 + tiles - 1) / tiles);
        if (chunks < 1) chunks = 1;
        int zChunk = (p.nz + chunks - 1) / chunks;
        if (zChunk < 16) zChunk = 16;
        if (zChunk > p.nz) zChunk = p.nz;
        p.zChunk = zChunk;
        const dim3 grid((p.nx + 31) / 32, yTiles, (p.nz + zChunk - 1) / zChunk);
        if (grid.y <= 65535 && grid.z <= 65535) {
            switch (src.format) {
            case F_U8:    return launchFilter3<F_U8>(c, p, grid);
            case F_U16:   return launchFilter3<F_U16>(c, p, grid);
            case F_INT16: return launchFilter3<F_INT16>(c, p, grid);
            case F_U32:   return launchFilter3<F_U32>(c, p, grid);
            case F_F32:   return launchFilter3<F_F32>(c, p, grid);
            }
        }
    }
    p.zChunk = 0; p.pairs = 0;
    const size_t n = (size_t)p.nx * p.ny * p.nz;
    filterGenericKernel<<<flatGrid(n, kThreads, c.smCount, 16), kThreads, 0, c.stream>>>(p);
    countLaunch();
    return cudaGetLastError();
}

} // namespace vktb
