/* volkit-b200 -- user-supplied Transform ops on the GPU policy.
 *
 * The reference's Transform takes an arbitrary host function (TransformUnaryOp / TransformBinaryOp,
 * include/cpp/vkt/Transform.hpp:16-25; bodies src/vkt/Transform_serial.hpp:15-50, 53-103).  A host function cannot run
 * on the device, so under Device::GPU the op needs a device twin.  The user writes the op ONCE in a header:
 *
 *     #include <vktb/transform_op.cuh>
 *     VKTB_UNARY_OP(my_op, x, y, z, voxel) { voxel.bytes[0] ^= (uint8_t)(x + y + z); }
 *
 * Compiled by a host compiler this defines the plain function `my_op` (a TransformUnaryOp: pass it to
 * vkt::Transform as always).  Compiled by nvcc / NVRTC for sm_100a it defines the device function AND the kernel
 * `vktb_unary_my_op` that applies it -- the op is inlined into a coalesced, 128-bit-vectorised walk of the range, not
 * called through a pointer.  The resulting cubin / PTX / fatbin is registered once:
 *
 *     vktb_register_unary_op_image(image, size, "my_op", &handle);   vktb_bind_unary_op((void*)&my_op, handle);
 *
 * (or vktb_register_unary_op_source with the CUDA C text, compiled by NVRTC when libnvrtc is present), after which
 * the unchanged vkt::Transform(volume, my_op) runs on the GPU.  Semantics are the reference's: the op sees ABSOLUTE
 * coordinates and an 8-byte zeroed scratch holding the voxel's raw bytes; bytesPerVoxel bytes are written back.
 *
 * This header is self-contained (no other volkit-b200 header, no libdevice requirement beyond what the op uses).
 */
#ifndef VKTB_TRANSFORM_OP_CUH
#define VKTB_TRANSFORM_OP_CUH

#ifndef __CUDACC_RTC__
#include <stdint.h>
#else
typedef unsigned char uint8_t;
typedef int int32_t;
typedef unsigned int uint32_t;
typedef unsigned long long uint64_t;
#endif

#define VKTB_TRANSFORM_ABI 1

/* layout of vkt::VoxelView / vktVoxelView_t (include/cpp/vkt/Voxel.hpp:15-21): bytes, dataFormat, mappingLo/Hi */
typedef struct vktb_voxel_view
{
    uint8_t* bytes;
    int32_t  dataFormat;       /* vkt::DataFormat / enum vktb_format */
    float    mappingLo;
    float    mappingHi;
} vktb_voxel_view;

/* kernel parameters; the library fills them, the generated kernels read them (VKTB_TRANSFORM_ABI) */
typedef struct vktb_transform_vol
{
    uint8_t*  base;            /* voxel (x0, y0, z0) */
    long long pitchY, pitchZ;  /* bytes */
    int32_t   bpv, format;
    float     lo, hi;
} vktb_transform_vol;

typedef struct vktb_transform_params
{
    vktb_transform_vol v1, v2; /* v2 unused by unary ops */
    int32_t nx, ny, nz;        /* extents of [first, last) */
    int32_t x0, y0, z0;        /* first: the op receives absolute coordinates */
} vktb_transform_params;

#if defined(__CUDACC__)

/* one voxel: zeroed 8-byte scratch <- bpv raw bytes, op, bpv bytes back (Transform_serial.hpp:27-46) */
template <class Op>
__device__ __forceinline__ uint32_t vktb_apply_unary(Op op, int x, int y, int z, uint32_t raw, const vktb_transform_vol& v)
{
    uint64_t scratch = raw;
    vktb_voxel_view view = { reinterpret_cast<uint8_t*>(&scratch), v.format, v.lo, v.hi };
    op(x, y, z, view);
    return (uint32_t)scratch;
}

__device__ __forceinline__ uint32_t vktb_load_raw(const uint8_t* p, int bpv)
{
    return bpv == 1 ? *p : bpv == 2 ? *reinterpret_cast<const unsigned short*>(p) : *reinterpret_cast<const uint32_t*>(p);
}
__device__ __forceinline__ void vktb_store_raw(uint8_t* p, int bpv, uint32_t w)
{
    if (bpv == 1) *p = (uint8_t)w;
    else if (bpv == 2) *reinterpret_cast<unsigned short*>(p) = (unsigned short)w;
    else *reinterpret_cast<uint32_t*>(p) = w;
}
template <int BPV> __device__ __forceinline__ uint32_t vktb_vec_get(const uint32_t w[4], int i)
{
    return BPV == 4 ? w[i] : BPV == 2 ? (w[i >> 1] >> ((i & 1) * 16)) & 0xffffu : (w[i >> 2] >> ((i & 3) * 8)) & 0xffu;
}
template <int BPV> __device__ __forceinline__ void vktb_vec_set(uint32_t w[4], int i, uint32_t v)
{
    if (BPV == 4) w[i] = v;
    else if (BPV == 2) w[i >> 1] |= (v & 0xffffu) << ((i & 1) * 16);
    else w[i >> 2] |= (v & 0xffu) << ((i & 3) * 8);
}

/* grid: x = 16-byte chunks of a row (+1 for a misaligned head), y = rows (grid-stride).  A thread owns the aligned
 * 16-byte chunk `c` of every row it visits: full chunks are one 128-bit load, V ops and one 128-bit store.  Four
 * rows are in flight per thread (all loads issued before the first op: one row at a time leaves 16 bytes in flight
 * per thread, 0.64 of the HBM peak at 2048^3 UInt16). */
template <int BPV, class Op>
__device__ __forceinline__ void vktb_unary_chunk(const vktb_transform_params& p, Op op, long long r, int chunk, bool loadOnly,
                                                 uint4& q, uint8_t*& dp, int& n, int& ax, int& ay, int& az)
{
    constexpr int V = 16 / BPV;
    const unsigned ur = (unsigned)r;
    const int z = r < 0x7fffffffLL ? (int)(ur / (unsigned)p.ny) : (int)(r / p.ny);
    const int y = (int)(r - (long long)z * p.ny);
    uint8_t* row = p.v1.base + y * p.v1.pitchY + z * p.v1.pitchZ;
    const int shift = (int)((reinterpret_cast<unsigned long long>(row) & 15ull) / BPV);
    int xr = chunk * V - shift;
    n = V;
    if (xr < 0) { n += xr; xr = 0; }
    if (xr + n > p.nx) n = p.nx - xr;
    dp = row + (long long)xr * BPV;
    ax = p.x0 + xr; ay = p.y0 + y; az = p.z0 + z;
    if (loadOnly && n == V) q = *reinterpret_cast<const uint4*>(dp);
}

template <int BPV, class Op>
__device__ __forceinline__ void vktb_unary_finish(const vktb_transform_params& p, Op op, const uint4& q, uint8_t* dp, int n,
                                                  int ax, int ay, int az)
{
    constexpr int V = 16 / BPV;
    if (n <= 0) return;
    if (n == V) {
        const uint32_t in[4] = { q.x, q.y, q.z, q.w };
        uint32_t out[4] = { 0u, 0u, 0u, 0u };
#pragma unroll
        for (int i = 0; i < V; ++i)
            vktb_vec_set<BPV>(out, i, vktb_apply_unary(op, ax + i, ay, az, vktb_vec_get<BPV>(in, i), p.v1));
        *reinterpret_cast<uint4*>(dp) = make_uint4(out[0], out[1], out[2], out[3]);
    } else {
        for (int i = 0; i < n; ++i)
            vktb_store_raw(dp + i * BPV, BPV, vktb_apply_unary(op, ax + i, ay, az, vktb_load_raw(dp + i * BPV, BPV), p.v1));
    }
}

template <int BPV, class Op>
__device__ __forceinline__ void vktb_unary_walk(const vktb_transform_params& p, Op op)
{
    constexpr int U = 4;
    const long long rows = (long long)p.ny * p.nz;
    const int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    long long r = blockIdx.y;
    for (; r + (long long)(U - 1) * gridDim.y < rows; r += (long long)U * gridDim.y) {
        uint4 q[U]; uint8_t* dp[U]; int n[U], ax[U], ay[U], az[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            vktb_unary_chunk<BPV>(p, op, r + (long long)u * gridDim.y, chunk, true, q[u], dp[u], n[u], ax[u], ay[u], az[u]);
#pragma unroll
        for (int u = 0; u < U; ++u)
            vktb_unary_finish<BPV>(p, op, q[u], dp[u], n[u], ax[u], ay[u], az[u]);
    }
    for (; r < rows; r += gridDim.y) {
        uint4 q; uint8_t* dp; int n, ax, ay, az;
        vktb_unary_chunk<BPV>(p, op, r, chunk, true, q, dp, n, ax, ay, az);
        vktb_unary_finish<BPV>(p, op, q, dp, n, ax, ay, az);
    }
}

/* binary ops may modify both voxels and the two volumes may differ in format: one voxel per thread, x fastest */
template <class Op>
__device__ __forceinline__ void vktb_binary_walk(const vktb_transform_params& p, Op op)
{
    const long long rows = (long long)p.ny * p.nz;
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= p.nx) return;
    for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
        const int z = (int)(r / p.ny), y = (int)(r - (long long)z * p.ny);
        uint8_t* p1 = p.v1.base + y * p.v1.pitchY + z * p.v1.pitchZ + (long long)x * p.v1.bpv;
        uint8_t* p2 = p.v2.base + y * p.v2.pitchY + z * p.v2.pitchZ + (long long)x * p.v2.bpv;
        uint64_t s1 = vktb_load_raw(p1, p.v1.bpv), s2 = vktb_load_raw(p2, p.v2.bpv);
        vktb_voxel_view a = { reinterpret_cast<uint8_t*>(&s1), p.v1.format, p.v1.lo, p.v1.hi };
        vktb_voxel_view b = { reinterpret_cast<uint8_t*>(&s2), p.v2.format, p.v2.lo, p.v2.hi };
        op(p.x0 + x, p.y0 + y, p.z0 + z, a, b);
        vktb_store_raw(p1, p.v1.bpv, (uint32_t)s1);
        vktb_store_raw(p2, p.v2.bpv, (uint32_t)s2);
    }
}

#define VKTB_UNARY_OP(NAME, X, Y, Z, VOXEL)                                                                       \
    __device__ __forceinline__ void NAME(int32_t X, int32_t Y, int32_t Z, vktb_voxel_view VOXEL);                \
    struct NAME##_vktb_functor { __device__ __forceinline__ void operator()(int32_t x, int32_t y, int32_t z,    \
                                                                            vktb_voxel_view v) const { NAME(x, y, z, v); } }; \
    extern "C" __global__ void __launch_bounds__(256) vktb_unary_##NAME(vktb_transform_params p)                  \
    {                                                                                                             \
        if (p.v1.bpv == 1) vktb_unary_walk<1>(p, NAME##_vktb_functor());                                          \
        else if (p.v1.bpv == 2) vktb_unary_walk<2>(p, NAME##_vktb_functor());                                     \
        else vktb_unary_walk<4>(p, NAME##_vktb_functor());                                                        \
    }                                                                                                             \
    __device__ __forceinline__ void NAME(int32_t X, int32_t Y, int32_t Z, vktb_voxel_view VOXEL)

#define VKTB_BINARY_OP(NAME, X, Y, Z, VOXEL1, VOXEL2)                                                             \
    __device__ __forceinline__ void NAME(int32_t X, int32_t Y, int32_t Z, vktb_voxel_view VOXEL1, vktb_voxel_view VOXEL2); \
    struct NAME##_vktb_functor { __device__ __forceinline__ void operator()(int32_t x, int32_t y, int32_t z,    \
                                     vktb_voxel_view a, vktb_voxel_view b) const { NAME(x, y, z, a, b); } };      \
    extern "C" __global__ void __launch_bounds__(256) vktb_binary_##NAME(vktb_transform_params p)                 \
    {                                                                                                             \
        vktb_binary_walk(p, NAME##_vktb_functor());                                                               \
    }                                                                                                             \
    __device__ __forceinline__ void NAME(int32_t X, int32_t Y, int32_t Z, vktb_voxel_view VOXEL1, vktb_voxel_view VOXEL2)

#else  /* host compiler: the op is an ordinary function with the reference's TransformUnaryOp / BinaryOp signature */

#ifdef __cplusplus
#define VKTB_OP_LINKAGE extern "C"     /* findable by name (dlsym); still a valid TransformUnaryOp / BinaryOp */
#else
#define VKTB_OP_LINKAGE
#endif
#define VKTB_UNARY_OP(NAME, X, Y, Z, VOXEL) \
    VKTB_OP_LINKAGE void NAME(int32_t X, int32_t Y, int32_t Z, vktb_voxel_view VOXEL)
#define VKTB_BINARY_OP(NAME, X, Y, Z, VOXEL1, VOXEL2) \
    VKTB_OP_LINKAGE void NAME(int32_t X, int32_t Y, int32_t Z, vktb_voxel_view VOXEL1, vktb_voxel_view VOXEL2)

#endif
#endif /* VKTB_TRANSFORM_OP_CUH */
