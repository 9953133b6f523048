// BrickDecompose through the TMA unit (reference src/vkt/Decompose_serial.hpp:15-48: a loop of CopyRange(dest[index],
// source, first, last) calls, each clamping its reads to the source; Copy_serial.hpp:13-82).
//
// copyBoxesFlatKernel (stream_ops.cu) decodes every 16-byte dst vector into (x, y, z), clamps, funnel-shifts up to four
// global loads and merges them: correct, but 255 warp instructions per vector and 39 % of the HBM roofline for bricks
// with a halo (132-byte rows; profiles/r01_brick_ncu.md).  Here the global side is hardware:
//   * one CTA = one box x kz of its z planes.  Thread 0 issues kz tensor-map tile loads (cp.async.bulk.tensor.3d, SASS
//     UTMALDG), one per plane, box = (W, NY, 1) voxels at the brick's own (x0 rounded down to 16 bytes, y0) -- the
//     innermost tile coordinate must address a 16-byte aligned element (a misaligned one raises an illegal-instruction
//     fault), the others are free, and all may be negative -- and z clamped to the volume, so the z halo costs
//     nothing.  Out-of-volume voxels arrive as zeros;
//   * CTAs of border bricks then replicate the edge voxels into the zero-filled x / y halo IN shared memory (a few
//     hundred voxels), so that the consumer below has no clamp at all;
//   * the chunk's dst bytes are one contiguous run: every thread assembles aligned 128-bit dst vectors from shared
//     memory -- two 128-bit reads and a funnel shift per piece; a vector that straddles two brick rows merges two
//     pieces -- and streams them out.  About 70 warp instructions per vector, none of them a global load.
// Traffic: each src voxel of a brick is read once by the TMA unit (rows rounded up to 16 bytes), each dst byte written
// once: 2 bpv per brick voxel, the same algorithmic bytes as before.
#include <cstdlib>
#include <vector>

#include "common.cuh"
#include "copy_boxes.hpp"
#include "tma.cuh"

namespace vktb {

namespace {

constexpr int kBrickThreads = 256;

struct BrickParams
{
    const BoxRec* boxes;
    int sx, sy, sz;
    int kz;            // planes per CTA
    int rowPitch;      // bytes of one staged row (W * bpv, multiple of 16)
    int planePitch;    // bytes of one staged plane (multiple of 128)
    unsigned planeTx;  // bytes one tile load delivers (W * NY * bpv)
};

__device__ __forceinline__ unsigned magicDivB(unsigned e, unsigned mul, int sh)
{
    return (unsigned)(((unsigned long long)e * mul) >> sh);
}

__device__ __forceinline__ int clampB(int x, int a, int b) { int m = b < x ? b : x; return a < m ? m : a; }

__device__ __forceinline__ uint4 ldsWindow(const uint8_t* stage, unsigned a)
{
    const uint4* v = reinterpret_cast<const uint4*>(stage + (a & ~15u));
    const unsigned sh = a & 15u;
    const uint4 lo = v[0];
    if (sh == 0) return lo;
    return window16(lo, v[1], sh);
}

template <int BPV>
__global__ void __launch_bounds__(kBrickThreads)
copyBricksTmaKernel(const __grid_constant__ CUtensorMap tmap, BrickParams p)
{
    extern __shared__ __align__(128) uint8_t smemRaw[];
    uint8_t* stage = smemRaw + 128;                       // [0, 8): the mbarrier
    uint64_t* bar = reinterpret_cast<uint64_t*>(smemRaw);

    const BoxRec b = p.boxes[blockIdx.x];
    const int zBeg = blockIdx.y * p.kz;
    if (zBeg >= b.nz) return;
    const int kz = min(p.kz, b.nz - zBeg);
    const int t = threadIdx.x, nThreads = blockDim.x;
    constexpr int V = 16 / BPV;
    const int xOrigin = b.x0 & ~(V - 1);                  // floor to 16 bytes, also for a negative x0
    const unsigned padB = (unsigned)(b.x0 - xOrigin) * BPV;   // staged bytes in front of every brick row

    if (t == 0) {
        mbarInit(smemAddr(bar), 1);
        fenceBarrierInit();
        mbarArriveExpectTx(smemAddr(bar), p.planeTx * (unsigned)kz);
        for (int i = 0; i < kz; ++i)
            tensorLoad3d(smemAddr(stage + (size_t)i * p.planePitch), &tmap, xOrigin, b.y0, clampB(b.z0 + zBeg + i, 0, p.sz - 1),
                         smemAddr(bar));
    }
    __syncthreads();                                      // the barrier is initialised for everybody
    mbarWait(smemAddr(bar), 0u);

    // ---- border bricks: clamp-to-edge in x and y, done once in shared memory -------------------------------------
    const int L = clampB(-b.x0, 0, b.nx), R = clampB(b.x0 + b.nx - p.sx, 0, b.nx);      // out-of-volume columns
    const int T = clampB(-b.y0, 0, b.ny), B = clampB(b.y0 + b.ny - p.sy, 0, b.ny);      // out-of-volume rows
    if (L | R | T | B) {
        auto at = [&](int x, int y, int zr) { return stage + (size_t)zr * p.planePitch + (size_t)y * p.rowPitch + padB + (size_t)x * BPV; };
        auto cx = [&](int x) { return clampB(x, L, b.nx - R - 1); };
        auto cy = [&](int y) { return clampB(y, T, b.ny - B - 1); };
        // every out-of-volume voxel copies its doubly clamped source, which lies in the volume and is never written
        const int ncol = L + R;
        if (ncol) {
            const int items = ncol * b.ny * kz;
            for (int i = t; i < items; i += nThreads) {
                const int row = i / ncol, j = i - row * ncol;
                const int zr = row / b.ny, y = row - zr * b.ny;
                const int x = j < L ? j : b.nx - R + (j - L);
                storeRaw(at(x, y, zr), BPV, loadRaw(at(cx(x), cy(y), zr), BPV));
            }
        }
        const int nrow = T + B, wx = b.nx - L - R;
        if (nrow) {
            const int items = nrow * wx * kz;
            for (int i = t; i < items; i += nThreads) {
                const int row = i / wx, x = L + (i - row * wx);
                const int zr = row / nrow, j = row - zr * nrow;
                const int y = j < T ? j : b.ny - B + (j - T);
                storeRaw(at(x, y, zr), BPV, loadRaw(at(x, cy(y), zr), BPV));
            }
        }
        __syncthreads();
    }

    // ---- the chunk's dst bytes, cut into aligned 128-bit vectors ---------------------------------------------------
    const unsigned rowBytes = (unsigned)b.nx * BPV;
    const unsigned chunkBytes = rowBytes * (unsigned)b.ny * (unsigned)kz;
    uint8_t* dst = b.d + (size_t)zBeg * b.ny * rowBytes;
    const unsigned phase = (unsigned)(reinterpret_cast<uintptr_t>(dst) & 15u);
    const unsigned nvec = (chunkBytes + phase + 15u) / 16u;
    const unsigned gap = (unsigned)p.rowPitch - rowBytes;          // staged bytes between two brick rows

    // staged byte address of chunk byte `cb` (cb is a multiple of BPV), and how many bytes of its row follow it
    auto locate = [&](unsigned cb, unsigned& left, unsigned& y, unsigned& zr) {
        const unsigned e = cb / BPV;
        const unsigned r = magicDivB(e, b.mulX, b.shX);
        const unsigned c = (e - r * (unsigned)b.nx) * BPV;
        zr = magicDivB(r, b.mulY, b.shY);
        y = r - zr * (unsigned)b.ny;
        left = rowBytes - c;
        return zr * (unsigned)p.planePitch + y * (unsigned)p.rowPitch + padB + c;
    };

    for (unsigned k = t; k < nvec; k += nThreads) {
        const long long first = (long long)k * 16 - phase;         // chunk byte of the vector's first byte
        if (first >= 0 && first + 16 <= (long long)chunkBytes && rowBytes >= 16u) {   // (a shorter row: > 2 pieces)
            unsigned left, y, zr;
            const unsigned a = locate((unsigned)first, left, y, zr);
            uint4 q = ldsWindow(stage, a);
            if (left < 16u) {
                // the rest comes from the next brick row: its byte 0 sits `gap` staged bytes after this row's end
                // (or at the start of the next plane); read so that it lands at byte `left` of the vector
                const unsigned a2 = (y + 1u == (unsigned)b.ny) ? (zr + 1u) * (unsigned)p.planePitch + padB - left : a + gap;
                const uint4 q2 = ldsWindow(stage, a2);
                const uint32_t w1[4] = { q.x, q.y, q.z, q.w }, w2[4] = { q2.x, q2.y, q2.z, q2.w };
                uint32_t w[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int kb = min(max((int)left - 4 * i, 0), 4);
                    const uint32_t m = kb == 4 ? 0xffffffffu : ((1u << (8 * kb)) - 1u);
                    w[i] = (w1[i] & m) | (w2[i] & ~m);
                }
                q = make_uint4(w[0], w[1], w[2], w[3]);
            }
            stStream(dst + first, q);
        } else {
            // first / last partial vector of the chunk, rows shorter than a vector: voxel by voxel
            const long long lo = first < 0 ? 0 : first;
            const long long hi = first + 16 < (long long)chunkBytes ? first + 16 : (long long)chunkBytes;
            for (long long cb = lo; cb < hi; cb += BPV) {
                unsigned left, y, zr;
                const unsigned a = locate((unsigned)cb, left, y, zr);
                storeRaw(dst + cb, BPV, loadRaw(stage + a, BPV));
            }
        }
    }
}

template <int BPV>
cudaError_t launchBricks(const LaunchCtx& c, const CUtensorMap& tm, const BrickParams& p, dim3 grid, size_t smemBytes)
{
    auto kern = copyBricksTmaKernel<BPV>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
    if (e != cudaSuccess) return e;
    int threads = kBrickThreads;
    if (const char* s = getenv("VKTB_BRICK_THREADS")) { const int v = atoi(s); if (v >= 32 && v <= kBrickThreads && v % 32 == 0) threads = v; }
    kern<<<grid, threads, smemBytes, c.stream>>>(tm, p);
    countLaunch();
    return cudaGetLastError();
}

} // namespace

cudaError_t launchCopyBricksTma(const LaunchCtx& c, const BoxRec* dev, const BoxRec* host, size_t n, const uint8_t* src,
                                int bpv, const int32_t sdims[3])
{
    if (n == 0 || n > 0x7fffffffull) return cudaErrorNotSupported;
    if (getenv("VKTB_BRICK_NO_TMA")) return cudaErrorNotSupported;
    const int sx = sdims[0], sy = sdims[1], sz = sdims[2];
    // tensor map: 16-byte aligned base and strides
    if ((reinterpret_cast<uintptr_t>(src) & 15u) || ((long long)sx * bpv) % 16) return cudaErrorNotSupported;
    const int V = 16 / bpv;
    int maxNx = 0, maxNy = 0, maxNz = 0;
    for (size_t i = 0; i < n; ++i) {
        const BoxRec& b = host[i];
        // the clamp sources of the x / y halo must lie inside the staged box
        if (b.x0 >= sx || b.x0 + b.nx <= 0 || b.y0 >= sy || b.y0 + b.ny <= 0) return cudaErrorNotSupported;
        maxNx = b.nx > maxNx ? b.nx : maxNx;
        maxNy = b.ny > maxNy ? b.ny : maxNy;
        maxNz = b.nz > maxNz ? b.nz : maxNz;
    }
    const int W = (maxNx + (V - 1) + V - 1) / V * V;   // + the origin's round-down to 16 bytes
    if (W > 256 || maxNy > 256) return cudaErrorNotSupported;
    const long long planeTx = (long long)W * maxNy * bpv;
    const long long planePitch = (planeTx + 127) / 128 * 128;
    // planes per CTA: as many as fit ~56 KB (three CTAs per SM keep ~170 KB of loads in flight), preferring a count
    // that divides the brick depth (a last CTA with one plane still pays the whole prologue)
    const long long budget = 56 * 1024;
    int kzMax = (int)(budget / planePitch);
    if (kzMax < 1) {
        if (planePitch > 200 * 1024) return cudaErrorNotSupported;
        kzMax = 1;
    }
    if (kzMax > maxNz) kzMax = maxNz;
    if (kzMax > 16) kzMax = 16;
    int kz = kzMax;
    long long bestWaste = -1;
    for (int k = kzMax; k >= (kzMax + 1) / 2; --k) {
        const long long waste = (long long)((maxNz + k - 1) / k) * k - maxNz;
        if (bestWaste < 0 || waste < bestWaste) { bestWaste = waste; kz = k; }
    }
    if (const char* s = getenv("VKTB_BRICK_KZ")) {
        const int k = atoi(s);
        if (k >= 1 && k <= 16 && (long long)k * planePitch <= 200 * 1024) kz = k;
    }
    const unsigned gy = (unsigned)((maxNz + kz - 1) / kz);
    if (gy > 65535u) return cudaErrorNotSupported;

    EncodeTiledFn encode = encodeTiled();
    if (!encode) return cudaErrorNotSupported;
    CUtensorMap tm;
    const cuuint64_t gdim[3] = { (cuuint64_t)sx, (cuuint64_t)sy, (cuuint64_t)sz };
    const cuuint64_t gstr[2] = { (cuuint64_t)sx * bpv, (cuuint64_t)sx * sy * bpv };
    const cuuint32_t box[3] = { (cuuint32_t)W, (cuuint32_t)maxNy, 1u };
    const cuuint32_t estr[3] = { 1u, 1u, 1u };
    const CUtensorMapDataType dt = bpv == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : bpv == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16
                                                                                       : CU_TENSOR_MAP_DATA_TYPE_UINT32;
    if (encode(&tm, dt, 3, const_cast<uint8_t*>(src), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return cudaErrorNotSupported;

    BrickParams p;
    p.boxes = dev;
    p.sx = sx; p.sy = sy; p.sz = sz;
    p.kz = kz;
    p.rowPitch = W * bpv;
    p.planePitch = (int)planePitch;
    p.planeTx = (unsigned)planeTx;
    // + 128: the barrier; + 32: the funnel-shifted read of the last staged bytes looks one vector further
    const size_t smemBytes = 128 + (size_t)kz * planePitch + 32;
    dim3 grid((unsigned)n, gy, 1);
    switch (bpv) {
    case 1: return launchBricks<1>(c, tm, p, grid, smemBytes);
    case 2: return launchBricks<2>(c, tm, p, grid, smemBytes);
    case 4: return launchBricks<4>(c, tm, p, grid, smemBytes);
    }
    return cudaErrorNotSupported;
}

} // namespace vktb
