// Scan / ScanRange: the 3-D prefix sum ("summed-area volume") of SURVEY.md 8(f) rank 4.
//
// Reference: src/vkt/Scan_serial.hpp:18-152.  The voxel's raw bytes are read as an unsigned integer
// (IntegralVoxel<uint64_t>, src/vkt/IntegralVoxel.hpp), sums are formed in 64 bits and the low bytesPerVoxel bytes are
// stored -- arithmetic modulo 2^(8 bpv), whatever the DataFormat (Float32 voxels are summed as their bit patterns).
// The serial body copies src to dst (CopyRange, when they are different objects), moves the range to dst space and
// walks it from `first` towards `last` on each axis (right to left where first > last), combining the already
// finished neighbours by inclusion-exclusion.  Modulo 2^k that is exactly the composition of three independent 1-D
// inclusive prefix sums, along x, then y, then z -- associative integer adds, so any evaluation order gives the same
// bytes.  Three passes over dst, in place:
//     x   one warp per row: packed adds inside the 128-bit vector, a shuffle scan of the lane totals, a running
//         carry along the row; the next vector is loaded before the current one is scanned
//     y/z one thread per 16-byte column of a plane (y pass) or of an xy position (z pass): a running packed sum walks
//         the lines, coalesced across x, several loads in flight
//     x+y fused (aligned rows of up to 16 KiB, at least 2 planes per SM): one CTA per plane walks the rows with the
//         column sums in registers -- two passes over the volume instead of three
// Packed adds (__vadd4 / __vadd2) keep 16 / 8 voxels per 128-bit register for UInt8 / UInt16.
// Algorithmic traffic: 2 bpv per voxel for the ideal single pass; the passes move 4 bpv (fused x+y) or 6 bpv.
// A true single pass was built and measured (commit "Experiment: single-pass Scan as a wavefront of row bands",
// profiles/r02_scan_single_pass_experiment.log): CTAs own bands of rows for all planes, keep the z running sum in
// shared memory and hand the y prefix from band to band through release / acquire flags.  Bit-exact, but 14.1 - 17.5 ms
// against 11.1 ms for the two passes at 2048^3 UInt16 (6.1 against 2.2 ms at 1024^3): a band's plane step is 6 - 8 us
// of dependent latency (bulk-load wait, three barriers, a device-wide fence behind its stores, the flag round trip
// through L2) with only one to three CTAs per SM to hide it, where the memory time of the step is 2.6 us.
#include "common.cuh"
#include "launch.hpp"
#include "voxel_codec.cuh"

namespace vktb {

namespace {

template <int BPV> __device__ __forceinline__ uint32_t addPacked(uint32_t a, uint32_t b)
{
    return BPV == 1 ? __vadd4(a, b) : BPV == 2 ? __vadd2(a, b) : a + b;
}
// inclusive prefix sum of the elements packed in one word
template <int BPV> __device__ __forceinline__ uint32_t wordScan(uint32_t w)
{
    if (BPV == 1) { w = __vadd4(w, w << 8); w = __vadd4(w, w << 16); }
    else if (BPV == 2) w = __vadd2(w, w << 16);
    return w;
}
// the word's last element replicated into every element
template <int BPV> __device__ __forceinline__ uint32_t lastBroadcast(uint32_t w)
{
    return BPV == 1 ? __byte_perm(w, 0, 0x3333) : BPV == 2 ? __byte_perm(w, 0, 0x3232) : w;
}
template <int BPV> __device__ __forceinline__ uint4 addPacked4(const uint4& a, const uint4& b)
{
    return make_uint4(addPacked<BPV>(a.x, b.x), addPacked<BPV>(a.y, b.y), addPacked<BPV>(a.z, b.z), addPacked<BPV>(a.w, b.w));
}

struct ScanParams
{
    uint8_t* d;               // dst voxel at the START corner of the range (first + off on every axis)
    long long pitchY, pitchZ; // signed: negative where the axis runs right to left
    int nx, ny, nz;           // extents
    int stepX;                // +1 / -1
    const uint8_t* lo;        // the volume buffer [lo, hi): bounds for whole-vector accesses at the row ends
    const uint8_t* hi;
};

// ---- x pass, left to right: one warp per row ------------------------------------------------------------------
// The row's bytes [b0, b1) are covered by aligned vectors; bytes of the first / last vector outside the row are
// zeroed before the scan and not stored.
template <int BPV>
__global__ void __launch_bounds__(kThreads)
scanRowsKernel(ScanParams p)
{
    const unsigned lane = threadIdx.x & 31u;
    const long long rows = (long long)p.ny * p.nz;
    const long long warpsTotal = (long long)gridDim.x * (kThreads / 32);
    for (long long r = (long long)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5); r < rows; r += warpsTotal) {
        const int z = (int)(r / p.ny), y = (int)(r - (long long)z * p.ny);
        uint8_t* row = p.d + y * p.pitchY + z * p.pitchZ;
        const uintptr_t b0 = reinterpret_cast<uintptr_t>(row), b1 = b0 + (uintptr_t)p.nx * BPV;
        const uintptr_t a0 = b0 & ~(uintptr_t)15;
        const int nvec = (int)((b1 - a0 + 15) >> 4);

        // whole-vector load with the bytes outside [b0, b1) cleared; byte-wise where the vector leaves the buffer
        auto load = [&](int v) {
            const uintptr_t a = a0 + ((uintptr_t)v << 4);
            uint32_t w[4];
            if (a >= reinterpret_cast<uintptr_t>(p.lo) && a + 16 <= reinterpret_cast<uintptr_t>(p.hi)) {
                const uint4 q = *reinterpret_cast<const uint4*>(a);
                w[0] = q.x; w[1] = q.y; w[2] = q.z; w[3] = q.w;
                if (a < b0 || a + 16 > b1) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const long long s = (long long)b0 - (long long)(a + 4 * i);   // bytes of word i before the row
                        const long long e = (long long)(a + 4 * i + 4) - (long long)b1; // ... and after it
                        uint32_t m = 0xffffffffu;
                        if (s >= 4 || e >= 4) m = 0u;
                        else {
                            if (s > 0) m &= 0xffffffffu << (8 * s);
                            if (e > 0) m &= 0xffffffffu >> (8 * e);
                        }
                        w[i] &= m;
                    }
                }
            } else {
                w[0] = w[1] = w[2] = w[3] = 0u;
                for (int i = 0; i < 16; ++i)
                    if (a + i >= b0 && a + i < b1) w[i >> 2] |= (uint32_t)*reinterpret_cast<const uint8_t*>(a + i) << ((i & 3) * 8);
            }
            return make_uint4(w[0], w[1], w[2], w[3]);
        };

        uint32_t carry = 0u;   // sum of everything left of the current 32 vectors, replicated into every element
        uint4 cur = (int)lane < nvec ? load((int)lane) : make_uint4(0u, 0u, 0u, 0u);
        for (int v0 = 0; v0 < nvec; v0 += 32) {
            const int v = v0 + (int)lane;
            const uint4 nxt = v + 32 < nvec ? load(v + 32) : make_uint4(0u, 0u, 0u, 0u);
            // inside the vector
            uint32_t w0 = wordScan<BPV>(cur.x);
            uint32_t w1 = addPacked<BPV>(wordScan<BPV>(cur.y), lastBroadcast<BPV>(w0));
            uint32_t w2 = addPacked<BPV>(wordScan<BPV>(cur.z), lastBroadcast<BPV>(w1));
            uint32_t w3 = addPacked<BPV>(wordScan<BPV>(cur.w), lastBroadcast<BPV>(w2));
            // across the lanes: inclusive scan of the vector totals
            uint32_t t = lastBroadcast<BPV>(w3);
#pragma unroll
            for (int dlt = 1; dlt < 32; dlt <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, t, dlt);
                if ((int)lane >= dlt) t = addPacked<BPV>(t, up);
            }
            uint32_t before = __shfl_up_sync(0xffffffffu, t, 1);
            if (lane == 0) before = 0u;
            before = addPacked<BPV>(before, carry);
            carry = addPacked<BPV>(carry, __shfl_sync(0xffffffffu, t, 31));
            if (v < nvec) {
                w0 = addPacked<BPV>(w0, before); w1 = addPacked<BPV>(w1, before);
                w2 = addPacked<BPV>(w2, before); w3 = addPacked<BPV>(w3, before);
                const uintptr_t a = a0 + ((uintptr_t)v << 4);
                if (a >= b0 && a + 16 <= b1) *reinterpret_cast<uint4*>(a) = make_uint4(w0, w1, w2, w3);
                else {
                    const uint32_t w[4] = { w0, w1, w2, w3 };
                    for (int i = 0; i < 16; ++i)
                        if (a + i >= b0 && a + i < b1) *reinterpret_cast<uint8_t*>(a + i) = (uint8_t)(w[i >> 2] >> ((i & 3) * 8));
                }
            }
            cur = nxt;
        }
    }
}

// ---- fused x + y pass: one CTA per plane ------------------------------------------------------------------------
// out[y] = rowscan(in[y]) + out[y - 1]: the CTA walks the rows of its plane top to bottom; a thread owns VPT
// consecutive 16-byte vectors of every row and keeps the previous output row's vectors (= the column sums) in
// registers.  Per row: packed prefix inside the thread's vectors, shuffle scan of the thread totals, the warp totals
// through shared memory (double buffered: ONE barrier per row), next row's vectors already in flight.  One read and
// one write of the volume instead of two of each.  Launcher: x left to right, every row starts on a 16-byte boundary
// and is a whole number of vectors, nvec <= kThreads * VPT.
template <int BPV, int VPT>
__global__ void __launch_bounds__(kThreads)
scanPlaneXYKernel(ScanParams p, int nvec)
{
    __shared__ uint32_t warpTot[2][kThreads / 32];
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const int v0 = (int)threadIdx.x * VPT;
    for (int z = blockIdx.x; z < p.nz; z += gridDim.x) {
        uint8_t* base = p.d + z * p.pitchZ + (long long)v0 * 16;
        uint4 col[VPT], cur[VPT];
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            col[k] = make_uint4(0u, 0u, 0u, 0u);
            cur[k] = v0 + k < nvec ? ldPlain(base + k * 16) : make_uint4(0u, 0u, 0u, 0u);
        }
        for (int y = 0; y < p.ny; ++y) {
            uint8_t* row = base + y * p.pitchY;
            uint4 nxt[VPT];
#pragma unroll
            for (int k = 0; k < VPT; ++k)
                nxt[k] = (y + 1 < p.ny && v0 + k < nvec) ? ldPlain(row + p.pitchY + k * 16) : make_uint4(0u, 0u, 0u, 0u);
            // inside the thread: prefix over its VPT vectors
            uint4 w[VPT];
            uint32_t t = 0u;                        // running total, replicated into every element
#pragma unroll
            for (int k = 0; k < VPT; ++k) {
                w[k].x = addPacked<BPV>(wordScan<BPV>(cur[k].x), t);
                w[k].y = addPacked<BPV>(wordScan<BPV>(cur[k].y), lastBroadcast<BPV>(w[k].x));
                w[k].z = addPacked<BPV>(wordScan<BPV>(cur[k].z), lastBroadcast<BPV>(w[k].y));
                w[k].w = addPacked<BPV>(wordScan<BPV>(cur[k].w), lastBroadcast<BPV>(w[k].z));
                t = lastBroadcast<BPV>(w[k].w);
            }
            // across the lanes, then across the warps
#pragma unroll
            for (int dlt = 1; dlt < 32; dlt <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, t, dlt);
                if ((int)lane >= dlt) t = addPacked<BPV>(t, up);
            }
            uint32_t before = __shfl_up_sync(0xffffffffu, t, 1);
            if (lane == 0) before = 0u;
            if (lane == 31) warpTot[y & 1][warp] = t;
            __syncthreads();
            for (unsigned i = 0; i < warp; ++i) before = addPacked<BPV>(before, warpTot[y & 1][i]);
#pragma unroll
            for (int k = 0; k < VPT; ++k) {
                const uint4 b4 = make_uint4(before, before, before, before);
                col[k] = addPacked4<BPV>(addPacked4<BPV>(w[k], b4), col[k]);
                if (v0 + k < nvec) stStream(row + k * 16, col[k]);
                cur[k] = nxt[k];
            }
        }
        __syncthreads();       // the next plane reuses warpTot[0]
    }
}

template <int BPV, int VPT>
cudaError_t launchPlaneXY(const LaunchCtx& c, const ScanParams& p, int nvec)
{
    const long long cap = (long long)c.smCount * 8;
    scanPlaneXYKernel<BPV, VPT><<<(unsigned)(p.nz < cap ? p.nz : cap), kThreads, 0, c.stream>>>(p, nvec);
    countLaunch();
    return cudaGetLastError();
}

// ---- y / z pass: a running packed sum walks `steps` lines that lie `stepPitch` bytes apart -----------------------
// Thread = one aligned 16-byte column chunk of line family (o1, o2); every line of the range has the same 16-byte
// phase (launcher), so the chunk is the same byte window [c0, c1) of every line.
struct WalkParams
{
    uint8_t*  d;              // first line's first byte (start corner)
    int       rowBytes;       // nx * bpv
    int       steps;          // lines to accumulate over
    long long stepPitch;      // signed
    int       outer;          // independent line families
    long long outerPitch;     // signed
};

template <int BPV>
__global__ void __launch_bounds__(kThreads)
scanWalkKernel(WalkParams p, int phase /* bytes of the row start into its 16-byte line */)
{
    constexpr int U = 4;
    const int chunk = blockIdx.x * blockDim.x + threadIdx.x;
    long long c0 = (long long)chunk * 16 - phase, c1 = c0 + 16;
    const bool full = c0 >= 0 && c1 <= p.rowBytes;
    if (c0 < 0) c0 = 0;
    if (c1 > p.rowBytes) c1 = p.rowBytes;
    if (c0 >= c1) return;
    for (int o = blockIdx.y; o < p.outer; o += gridDim.y) {
        uint8_t* line = p.d + o * p.outerPitch + c0;
        if (full) {
            uint4 acc = make_uint4(0u, 0u, 0u, 0u);
            int s = 0;
            for (; s + U <= p.steps; s += U) {
                uint4 q[U];
#pragma unroll
                for (int u = 0; u < U; ++u) q[u] = ldPlain(line + (s + u) * p.stepPitch);
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    acc = addPacked4<BPV>(acc, q[u]);
                    stStream(line + (s + u) * p.stepPitch, acc);
                }
            }
            for (; s < p.steps; ++s) {
                acc = addPacked4<BPV>(acc, ldPlain(line + s * p.stepPitch));
                stStream(line + s * p.stepPitch, acc);
            }
        } else {
            // partial head / tail chunk of the row: element by element
            for (long long b = 0; b < c1 - c0; b += BPV) {
                uint32_t acc = 0u;
                for (int s = 0; s < p.steps; ++s) {
                    uint8_t* e = line + b + s * p.stepPitch;
                    acc += loadRaw(e, BPV);
                    storeRaw(e, BPV, acc);
                }
            }
        }
    }
}

// ---- generic: one thread per line, any direction, any alignment -------------------------------------------------
struct LineParams
{
    uint8_t*  d;
    int       steps;  long long stepPitch;
    int       n1;     long long pitch1;    // fast line index (adjacent threads)
    int       n2;     long long pitch2;
    int       bpv;
};

__global__ void __launch_bounds__(kThreads)
scanLinesKernel(LineParams p)
{
    const long long lines = (long long)p.n1 * p.n2;
    for (long long l = (long long)blockIdx.x * blockDim.x + threadIdx.x; l < lines; l += (long long)gridDim.x * blockDim.x) {
        const int i2 = (int)(l / p.n1), i1 = (int)(l - (long long)i2 * p.n1);
        uint8_t* e = p.d + i1 * p.pitch1 + i2 * p.pitch2;
        uint32_t acc = 0u;
        for (int s = 0; s < p.steps; ++s, e += p.stepPitch) {
            acc += loadRaw(e, p.bpv);
            storeRaw(e, p.bpv, acc);
        }
    }
}

cudaError_t lines(const LaunchCtx& c, uint8_t* d, int bpv, int steps, long long stepPitch, int n1, long long pitch1, int n2,
                  long long pitch2)
{
    if (steps <= 1) return cudaSuccess;
    LineParams p{ d, steps, stepPitch, n1, pitch1, n2, pitch2, bpv };
    const long long total = (long long)n1 * n2;
    unsigned grid = (unsigned)((total + kThreads - 1) / kThreads);
    const unsigned cap = (unsigned)c.smCount * 32u;
    if (grid > cap) grid = cap;
    scanLinesKernel<<<grid, kThreads, 0, c.stream>>>(p);
    countLaunch();
    return cudaGetLastError();
}

template <int BPV>
cudaError_t scanBpv(const LaunchCtx& c, const ScanParams& p, bool alignedBuffer)
{
    const long long rowBytes = (long long)p.nx * BPV;
    const long long ay = p.pitchY < 0 ? -p.pitchY : p.pitchY, az = p.pitchZ < 0 ? -p.pitchZ : p.pitchZ;
    cudaError_t e;
    // ---- x and y in one pass when a CTA can own a plane's rows (and there are enough planes to fill the machine)
    bool fusedXY = false;
    if (p.nx > 1 && p.ny > 1 && p.stepX > 0 && (reinterpret_cast<uintptr_t>(p.d) & 15u) == 0 && rowBytes % 16 == 0
        && ay % 16 == 0 && az % 16 == 0 && rowBytes / 16 <= kThreads * 4 && p.nz >= c.smCount * 2) {
        const int nvec = (int)(rowBytes / 16);
        e = nvec <= kThreads ? launchPlaneXY<BPV, 1>(c, p, nvec)
          : nvec <= kThreads * 2 ? launchPlaneXY<BPV, 2>(c, p, nvec) : launchPlaneXY<BPV, 4>(c, p, nvec);
        if (e != cudaSuccess) return e;
        fusedXY = true;
    }
    // ---- x
    if (p.nx > 1 && !fusedXY) {
        if (p.stepX > 0 && alignedBuffer && (long long)p.ny * p.nz < 0x7fffffffLL) {
            const long long rows = (long long)p.ny * p.nz;
            long long ctas = (rows + kThreads / 32 - 1) / (kThreads / 32);
            const long long cap = (long long)c.smCount * 8;
            scanRowsKernel<BPV><<<(unsigned)(ctas < cap ? ctas : cap), kThreads, 0, c.stream>>>(p);
            countLaunch();
            e = cudaGetLastError();
        } else {
            e = lines(c, p.d, BPV, p.nx, (long long)p.stepX * BPV, p.ny, p.pitchY, p.nz, p.pitchZ);
        }
        if (e != cudaSuccess) return e;
    }
    // ---- y, z: whole 16-byte column chunks when every line has the same 16-byte phase and x runs left to right
    // (d then addresses the row's lowest byte)
    const bool walk = p.stepX > 0 && (ay % 16) == 0 && (az % 16) == 0 && rowBytes < 0x7fffffffLL;
    const int phase = (int)(reinterpret_cast<uintptr_t>(p.d) & 15u);
    const unsigned gx = (unsigned)((rowBytes + phase + 16 * kThreads - 1) / (16 * kThreads));
    for (int axis = fusedXY ? 2 : 1; axis <= 2; ++axis) {
        const int steps = axis == 1 ? p.ny : p.nz, outer = axis == 1 ? p.nz : p.ny;
        const long long stepPitch = axis == 1 ? p.pitchY : p.pitchZ, outerPitch = axis == 1 ? p.pitchZ : p.pitchY;
        if (steps <= 1) continue;
        if (walk) {
            WalkParams w{ p.d, (int)rowBytes, steps, stepPitch, outer, outerPitch };
            dim3 grid(gx, (unsigned)(outer > 65535 ? 65535 : outer), 1);
            scanWalkKernel<BPV><<<grid, kThreads, 0, c.stream>>>(w, phase);
            countLaunch();
            e = cudaGetLastError();
        } else {
            // lines indexed by x (adjacent threads: coalesced) and the other axis; d is the START corner, x may run backwards
            e = lines(c, p.d, BPV, steps, stepPitch, p.nx, (long long)p.stepX * BPV, outer, outerPitch);
        }
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

} // namespace

// dst may be src (same data pointer: in place, no copy).  first > last on an axis = right to left (reference
// include/cpp/vkt/Scan.hpp:14-22); only meaningful in place -- the reference's CopyRange never terminates on such
// a range, so it is rejected when dst != src.
cudaError_t launchScan(const LaunchCtx& c, const VolDesc& dst, const VolDesc& src, const int32_t first[3],
                       const int32_t last[3], const int32_t off[3])
{
    const int bpv = bytesPerVoxel(dst.format);
    if (bpv == 0) return cudaErrorInvalidValue;
    const bool inPlace = dst.data == src.data;
    for (int a = 0; a < 3; ++a) {
        if (first[a] == last[a]) return cudaErrorInvalidValue;   // the reference's loops never end on an empty axis
        if (!inPlace && first[a] > last[a]) return cudaErrorInvalidValue;
    }
    if (!inPlace) {
        cudaError_t e = launchCopy(c, dst, src, first, last, off);
        if (e != cudaSuccess) return e;
    }
    ScanParams p;
    int n[3], step[3];
    long long f[3];
    for (int a = 0; a < 3; ++a) {
        f[a] = (long long)first[a] + off[a];
        step[a] = first[a] < last[a] ? 1 : -1;
        n[a] = step[a] > 0 ? last[a] - first[a] : first[a] - last[a];
    }
    p.d = dst.data + voxelIndex(dst, f[0], f[1], f[2]) * bpv;
    p.pitchY = (long long)dst.dims[0] * bpv * step[1];
    p.pitchZ = (long long)dst.dims[0] * dst.dims[1] * bpv * step[2];
    p.nx = n[0]; p.ny = n[1]; p.nz = n[2];
    p.stepX = step[0];
    p.lo = dst.data;
    p.hi = dst.data + (size_t)dst.dims[0] * dst.dims[1] * dst.dims[2] * bpv;
    const bool alignedBuffer = (reinterpret_cast<uintptr_t>(dst.data) & 15u) == 0;
    switch (bpv) {
    case 1: return scanBpv<1>(c, p, alignedBuffer);
    case 2: return scanBpv<2>(c, p, alignedBuffer);
    case 4: return scanBpv<4>(c, p, alignedBuffer);
    }
    return cudaErrorInvalidValue;
}

} // namespace vktb
