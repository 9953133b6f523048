// C-ABI algorithm entry points (include/vktb/backend.h): argument validation, then one launcher call.
#include "context.hpp"

#include <cfloat>
#include <cmath>
#include <vector>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "../../../include/vktb/backend.h"
#include "../kernels/launch.hpp"
#include "user_ops.hpp"
#include "multi.hpp"

using namespace vktb;

namespace {

bool supportedFormat(int f) { return f == F_U8 || f == F_U16 || f == F_U32 || f == F_F32 || f == F_INT16; }

size_t planeBytesOf(const VolDesc& v) { return (size_t)v.dims[0] * v.dims[1] * bytesPerVoxel(v.format); }

// The striped allocation a call fans out over: dst lives in one and every other operand is striped too (a pooled
// operand belongs to one device's memory pool, which the other devices' kernels cannot be assumed to reach).
const StripedAlloc* fanOutTarget(const VolDesc& dst, std::initializer_list<const VolDesc*> others)
{
    if (groupSize() < 2) return nullptr;
    const StripedAlloc* a = findStriped(dst.data);
    if (!a || a->stripes.size() < 2) return nullptr;
    for (const VolDesc* o : others)
        if (o && !findStriped(o->data)) return nullptr;
    return a;
}

// vktb_vol -> VolDesc; returns false (error set) on bad input
bool toDesc(const char* fn, const vktb_vol* v, VolDesc& d)
{
    if (!v) { fail((std::string(fn) + ": null volume").c_str()); return false; }
    if (!supportedFormat(v->format)) {
        // Int8 / Int32 have no case in the reference's MapVoxelImpl/UnmapVoxelImpl (VoxelMapping.hpp:26-95)
        fail((std::string(fn) + ": unsupported DataFormat (UInt8, UInt16, UInt32, Float32, Int16 only)").c_str());
        return false;
    }
    if (v->dims[0] < 0 || v->dims[1] < 0 || v->dims[2] < 0) { fail((std::string(fn) + ": negative dims").c_str()); return false; }
    const size_t n = (size_t)v->dims[0] * v->dims[1] * v->dims[2];
    if (n && !v->data) { fail((std::string(fn) + ": null data pointer").c_str()); return false; }
    d.data = (uint8_t*)v->data;
    d.dims[0] = v->dims[0]; d.dims[1] = v->dims[1]; d.dims[2] = v->dims[2];
    d.format = v->format;
    d.map = makeMapping(v->lo, v->hi);
    return true;
}

int finish(const char* fn, cudaError_t e) { return e == cudaSuccess ? 0 : fail(fn, e); }

// A range that leaves its volume's BUFFER would become an out-of-bounds device access, i.e. a sticky error that ends the
// process's CUDA context (the host reference merely scribbles over the heap: README.md:140 leaves dst bounds to the
// caller).  Every entry point therefore refuses such a call, like the hosted twins do (hosted.cu).
bool emptyRange(const int32_t first[3], const int32_t last[3])
{
    return last[0] <= first[0] || last[1] <= first[1] || last[2] <= first[2];
}

// Memory safety of [first, last) + shift in v, in the reference's own terms: it addresses voxels by the LINEAR index
// (z * dimY + y) * dimX + x without looking at the axes (StructuredVolume.cpp:311-317), so a box may legally run past a
// row end into the next row (the oracle and the kernels do the same pitch arithmetic).  What must hold is that the
// lowest and the highest linear index of the box lie inside the buffer; both corners are monotone in every coordinate.
bool inside(const VolDesc& v, const int32_t first[3], const int32_t last[3], const int32_t* shift = nullptr)
{
    const long long X = v.dims[0], Y = v.dims[1], Z = v.dims[2];
    long long lo[3], hi[3];
    for (int i = 0; i < 3; ++i) {
        const long long s = shift ? shift[i] : 0;
        lo[i] = first[i] + s;
        hi[i] = (long long)last[i] - 1 + s;
    }
    const long long lowest = (lo[2] * Y + lo[1]) * X + lo[0], highest = (hi[2] * Y + hi[1]) * X + hi[0];
    return lowest >= 0 && highest < X * Y * Z;
}

int outside(const char* fn, const char* which)
{
    return fail((std::string(fn) + ": range outside the " + which + " volume").c_str());
}

// Transform registry
const char* const kUnaryNames[kNumUnaryOps] = { "xor_coord", "invert", "checker8", "threshold" };
const char* const kBinaryNames[kNumBinaryOps] = { "or_assign", "swap" };
std::mutex g_opMutex;
std::unordered_map<const void*, int> g_unaryBound, g_binaryBound;

} // namespace

extern "C" {

int vktb_fill_range(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], float value)
{
    VKTB_API_RANGE();
    VolDesc v;
    if (!first || !last) return fail("vktb_fill_range: null range");
    if (!toDesc("vktb_fill_range", vol, v)) return -1;
    if (emptyRange(first, last)) return 0;
    if (!inside(v, first, last)) return outside("vktb_fill_range", "filled");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    if (const StripedAlloc* a = fanOutTarget(v, {})) {
        return finish("vktb_fill_range", forEachStripe(a, v.data, planeBytesOf(v), first[2], last[2],
            [&](const LaunchCtx& cg, int z0, int z1) {
                const int32_t f[3] = { first[0], first[1], z0 }, l[3] = { last[0], last[1], z1 };
                return launchFill(cg, v, f, l, value);
            }));
    }
    LaunchCtx c{ currentStream(d), d->smCount };
    return finish("vktb_fill_range", launchFill(c, v, first, last, value));
}

int vktb_copy_range(const vktb_vol* dst, const vktb_vol* src, const int32_t first[3], const int32_t last[3],
                    const int32_t off[3])
{
    VKTB_API_RANGE();
    VolDesc dv, sv;
    if (!first || !last || !off) return fail("vktb_copy_range: null range");
    if (!toDesc("vktb_copy_range", dst, dv) || !toDesc("vktb_copy_range", src, sv)) return -1;
    if (emptyRange(first, last)) return 0;
    {   // src indices are clamped (Copy_serial.hpp:38-40); the dst box [off, off + last - first) is not
        const int32_t zero[3] = { 0, 0, 0 }, ext[3] = { last[0] - first[0], last[1] - first[1], last[2] - first[2] };
        if (!inside(dv, zero, ext, off)) return outside("vktb_copy_range", "dst");
        if ((size_t)sv.dims[0] * sv.dims[1] * sv.dims[2] == 0) return fail("vktb_copy_range: empty source");
    }
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    if (const StripedAlloc* a = overlaps(dv, sv) ? nullptr : fanOutTarget(dv, { &sv })) {
        const long long shift = (long long)off[2] - first[2];            // dst plane = p.z + shift
        return finish("vktb_copy_range", forEachStripe(a, dv.data, planeBytesOf(dv), first[2] + shift, last[2] + shift,
            [&](const LaunchCtx& cg, int z0, int z1) {
                const int32_t f[3] = { first[0], first[1], (int32_t)(z0 - shift) }, l[3] = { last[0], last[1], (int32_t)(z1 - shift) };
                const int32_t o[3] = { off[0], off[1], z0 };
                return launchCopy(cg, dv, sv, f, l, o);
            }));
    }
    LaunchCtx c{ currentStream(d), d->smCount };
    return finish("vktb_copy_range", launchCopy(c, dv, sv, first, last, off));
}

int vktb_copy_range_batch(int n, const vktb_vol* dsts, const vktb_vol* src, const int32_t* firsts, const int32_t* lasts,
                          const int32_t* offs)
{
    VKTB_API_RANGE();
    if (n < 0) return fail("vktb_copy_range_batch: negative count");
    if (n == 0) return 0;
    if (!dsts || !firsts || !lasts || !offs) return fail("vktb_copy_range_batch: null argument");
    VolDesc sv;
    if (!toDesc("vktb_copy_range_batch", src, sv)) return -1;
    std::vector<VolDesc> dv((size_t)n);
    for (int i = 0; i < n; ++i)
        if (!toDesc("vktb_copy_range_batch", dsts + i, dv[(size_t)i])) return -1;
    for (int i = 0; i < n; ++i) {
        const int32_t *f = firsts + 3 * i, *l = lasts + 3 * i;
        if (emptyRange(f, l)) continue;
        const int32_t zero[3] = { 0, 0, 0 }, ext[3] = { l[0] - f[0], l[1] - f[1], l[2] - f[2] };
        if (!inside(dv[(size_t)i], zero, ext, offs + 3 * i)) return outside("vktb_copy_range_batch", "dst");
    }
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    LaunchCtx c{ currentStream(d), d->smCount };
    return finish("vktb_copy_range_batch", launchCopyBatch(c, n, dv.data(), sv, firsts, lasts, offs));
}

// No counterpart in the reference: see include/vktb/backend.h.  Nothing on this path calls cudaMalloc / cudaFree (or
// any allocating routine): the descriptors are converted into a fixed array, the box table lives in a file-scope
// device array of kernels/stream_ops.cu.
int vktb_copy_range_batch_noalloc_max(void) { return copyBatchNoAllocMaxBoxes(); }

int vktb_copy_range_batch_noalloc(int n, const vktb_vol* dsts, const vktb_vol* src, const int32_t* firsts,
                                  const int32_t* lasts, const int32_t* offs)
{
    VKTB_API_RANGE();
    const char* fn = "vktb_copy_range_batch_noalloc";
    if (n < 0) return fail("vktb_copy_range_batch_noalloc: negative count");
    if (n == 0) return 0;
    const int maxBoxes = copyBatchNoAllocMaxBoxes();
    if (n > maxBoxes) {
        char msg[200];
        snprintf(msg, sizeof msg, "vktb_copy_range_batch_noalloc: %d boxes exceed the static scratch table of %d "
                 "(use vktb_copy_range_batch, or split the batch)", n, maxBoxes);
        return fail(msg);
    }
    if (!dsts || !firsts || !lasts || !offs) return fail("vktb_copy_range_batch_noalloc: null argument");
    VolDesc sv;
    if (!toDesc(fn, src, sv)) return -1;
    static std::mutex descMutex;
    static VolDesc dv[4096];
    static_assert(sizeof(dv) / sizeof(dv[0]) >= 4096, "descriptor array");
    if (maxBoxes > 4096) return fail("vktb_copy_range_batch_noalloc: internal table size mismatch");
    std::lock_guard<std::mutex> lock(descMutex);
    for (int i = 0; i < n; ++i) {
        if (!toDesc(fn, dsts + i, dv[i])) return -1;
        if (!(dv[i].format == sv.format && dsts[i].lo == src->lo && dsts[i].hi == src->hi))
            return fail("vktb_copy_range_batch_noalloc: every box must have the source's format and mapping "
                        "(converting boxes take vktb_copy_range_batch)");
        const int32_t *f = firsts + 3 * i, *l = lasts + 3 * i;
        if (emptyRange(f, l)) continue;
        const int32_t zero[3] = { 0, 0, 0 }, ext[3] = { l[0] - f[0], l[1] - f[1], l[2] - f[2] };
        if (!inside(dv[i], zero, ext, offs + 3 * i)) return outside(fn, "dst");
    }
    DeviceCtx* d = currentDevice();     // makes the thread's device current: the scratch address is per device
    if (!d) return -1;
    LaunchCtx c{ currentStream(d), d->smCount };
    return finish(fn, launchCopyBatchNoAlloc(c, n, dv, sv, firsts, lasts, offs));
}

int vktb_scan_range(const vktb_vol* dst, const vktb_vol* src, const int32_t first[3], const int32_t last[3],
                    const int32_t off[3])
{
    VKTB_API_RANGE();
    VolDesc dv, sv;
    if (!first || !last || !off) return fail("vktb_scan_range: null range");
    if (!toDesc("vktb_scan_range", dst, dv) || !toDesc("vktb_scan_range", src, sv)) return -1;
    for (int a = 0; a < 3; ++a) {
        if (first[a] == last[a]) return fail("vktb_scan_range: empty axis (the reference's loops never end on one)");
        if (dv.data != sv.data && first[a] > last[a])
            return fail("vktb_scan_range: a right-to-left range needs dst == src (the copy of such a range is undefined)");
    }
    {   // the scanned box, whichever way each axis runs: [min(first, last + 1), max(last, first + 1)) + off in dst
        int32_t lo[3], hi[3];
        for (int a = 0; a < 3; ++a) {
            lo[a] = first[a] < last[a] ? first[a] : last[a] + 1;
            hi[a] = first[a] < last[a] ? last[a] : first[a] + 1;
        }
        if (!inside(dv, lo, hi, off)) return outside("vktb_scan_range", "dst");
        if (dv.data != sv.data && (size_t)sv.dims[0] * sv.dims[1] * sv.dims[2] == 0) return fail("vktb_scan_range: empty source");
    }
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    LaunchCtx c{ currentStream(d), d->smCount };
    return finish("vktb_scan_range", launchScan(c, dv, sv, first, last, off));
}

int vktb_arith_range(int op, const vktb_vol* dst, const vktb_vol* s1, const vktb_vol* s2, const int32_t first[3],
                     const int32_t last[3], const int32_t off[3])
{
    VKTB_API_RANGE();
    VolDesc dv, av, bv;
    if (op < 0 || op > VKTB_SAFEABSDIFF) return fail("vktb_arith_range: bad op");
    if (!first || !last || !off) return fail("vktb_arith_range: null range");
    if (!toDesc("vktb_arith_range", dst, dv) || !toDesc("vktb_arith_range", s1, av) || !toDesc("vktb_arith_range", s2, bv))
        return -1;
    if (emptyRange(first, last)) return 0;
    if (!inside(av, first, last) || !inside(bv, first, last)) return outside("vktb_arith_range", "source");
    if (!inside(dv, first, last, off)) return outside("vktb_arith_range", "dst");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    const bool shifted = off[0] || off[1] || off[2];
    const bool aliasShifted = shifted && (overlaps(dv, av) || overlaps(dv, bv));   // order-dependent: one device, as before
    if (const StripedAlloc* a = aliasShifted ? nullptr : fanOutTarget(dv, { &av, &bv })) {
        return finish("vktb_arith_range", forEachStripe(a, dv.data, planeBytesOf(dv), (long long)first[2] + off[2], (long long)last[2] + off[2],
            [&](const LaunchCtx& cg, int z0, int z1) {
                const int32_t f[3] = { first[0], first[1], z0 - off[2] }, l[3] = { last[0], last[1], z1 - off[2] };
                return launchArith(cg, op, dv, av, bv, f, l, off);
            }));
    }
    LaunchCtx c{ currentStream(d), d->smCount };
    return finish("vktb_arith_range", launchArith(c, op, dv, av, bv, first, last, off));
}

int vktb_transform_range1(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], int op)
{
    VKTB_API_RANGE();
    VolDesc v;
    const bool user = op >= kUserOpBase;
    if (user ? !userOpValid(op, false) : (op < 0 || op >= kNumUnaryOps)) return fail("vktb_transform_range1: unknown device op handle");
    if (!first || !last) return fail("vktb_transform_range1: null range");
    if (!toDesc("vktb_transform_range1", vol, v)) return -1;
    if (emptyRange(first, last)) return 0;
    if (!inside(v, first, last)) return outside("vktb_transform_range1", "transformed");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    if (const StripedAlloc* a = fanOutTarget(v, {})) {
        return finish("vktb_transform_range1", forEachStripe(a, v.data, planeBytesOf(v), first[2], last[2],
            [&](const LaunchCtx& cg, int z0, int z1) {
                const int32_t f[3] = { first[0], first[1], z0 }, l[3] = { last[0], last[1], z1 };   // absolute coordinates
                return user ? launchUserTransform(cg, op, v, nullptr, f, l) : launchTransform1(cg, v, f, l, op);
            }));
    }
    LaunchCtx c{ currentStream(d), d->smCount };
    if (user) return finish("vktb_transform_range1 (user op)", launchUserTransform(c, op, v, nullptr, first, last));
    return finish("vktb_transform_range1", launchTransform1(c, v, first, last, op));
}

int vktb_transform_range2(const vktb_vol* vol1, const vktb_vol* vol2, const int32_t first[3], const int32_t last[3], int op)
{
    VKTB_API_RANGE();
    VolDesc a, b;
    const bool user = op >= kUserOpBase;
    if (user ? !userOpValid(op, true) : (op < 0 || op >= kNumBinaryOps)) return fail("vktb_transform_range2: unknown device op handle");
    if (!first || !last) return fail("vktb_transform_range2: null range");
    if (!toDesc("vktb_transform_range2", vol1, a) || !toDesc("vktb_transform_range2", vol2, b)) return -1;
    if (emptyRange(first, last)) return 0;
    if (!inside(a, first, last) || !inside(b, first, last)) return outside("vktb_transform_range2", "transformed");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    if (const StripedAlloc* sa = fanOutTarget(a, { &b })) {
        return finish("vktb_transform_range2", forEachStripe(sa, a.data, planeBytesOf(a), first[2], last[2],
            [&](const LaunchCtx& cg, int z0, int z1) {
                const int32_t f[3] = { first[0], first[1], z0 }, l[3] = { last[0], last[1], z1 };
                return user ? launchUserTransform(cg, op, a, &b, f, l) : launchTransform2(cg, a, b, f, l, op);
            }));
    }
    LaunchCtx c{ currentStream(d), d->smCount };
    if (user) return finish("vktb_transform_range2 (user op)", launchUserTransform(c, op, a, &b, first, last));
    return finish("vktb_transform_range2", launchTransform2(c, a, b, first, last, op));
}

int vktb_resample_window(const vktb_vol* dst, const vktb_vol* src, int filter, int32_t dstZ0, const int32_t dstFull[3],
                         int32_t srcZ0, const int32_t srcFull[3])
{
    VKTB_API_RANGE();
    VolDesc dv, sv;
    if (filter < 0 || filter > VKTB_TRILINEAR) return fail("vktb_resample: bad filter mode");
    if (!dstFull || !srcFull) return fail("vktb_resample_window: null dims");
    if (!toDesc("vktb_resample", dst, dv) || !toDesc("vktb_resample", src, sv)) return -1;
    if (dv.dims[0] != dstFull[0] || dv.dims[1] != dstFull[1] || sv.dims[0] != srcFull[0] || sv.dims[1] != srcFull[1])
        return fail("vktb_resample_window: windows must span full x/y extents");
    if (dstZ0 < 0 || dstZ0 + dv.dims[2] > dstFull[2] || srcZ0 < 0 || srcZ0 + sv.dims[2] > srcFull[2])
        return fail("vktb_resample_window: window outside the full volume");
    // the src window must hold every plane the dst slab reads (the extra plane the planner requests for
    // Linear only carries the weight-0 hi.x spill tap and is optional)
    if (dv.dims[2] > 0) {
        int32_t need0, need1;
        const bool same = dstFull[0] == srcFull[0] && dstFull[1] == srcFull[1] && dstFull[2] == srcFull[2];
        if (same) { need0 = dstZ0; need1 = dstZ0 + dv.dims[2]; }
        else vktb_resample_window_src_range(filter == VKTB_LINEAR ? VKTB_TRILINEAR : filter, dstZ0, dstZ0 + dv.dims[2],
                                            dstFull[2], srcFull[2], &need0, &need1);
        if (srcZ0 > need0 || srcZ0 + sv.dims[2] < need1)
            return fail("vktb_resample_window: src window misses planes the dst slab reads");
    }
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    if (const StripedAlloc* a = overlaps(dv, sv) ? nullptr : fanOutTarget(dv, { &sv })) {
        // dst planes by stripe; the src window stays whole: a tap in the neighbour's stripe is a peer load
        const size_t pd = planeBytesOf(dv);
        return finish("vktb_resample", forEachStripe(a, dv.data, pd, 0, dv.dims[2],
            [&](const LaunchCtx& cg, int z0, int z1) {
                VolDesc w = dv;
                w.data = dv.data + (size_t)z0 * pd;
                w.dims[2] = z1 - z0;
                return launchResample(cg, w, sv, filter, dstZ0 + z0, dstFull, srcZ0, srcFull);
            }));
    }
    LaunchCtx c{ currentStream(d), d->smCount };
    return finish("vktb_resample", launchResample(c, dv, sv, filter, dstZ0, dstFull, srcZ0, srcFull));
}

int vktb_resample(const vktb_vol* dst, const vktb_vol* src, int filter)
{
    VKTB_API_RANGE();
    if (!dst || !src) return fail("vktb_resample: null volume");
    return vktb_resample_window(dst, src, filter, 0, dst->dims, 0, src->dims);
}

int vktb_flip_range(const vktb_vol* dst, const vktb_vol* src, const int32_t first[3], const int32_t last[3],
                    const int32_t off[3], int axis)
{
    VKTB_API_RANGE();
    VolDesc dv, sv;
    if (axis < 0 || axis > 2) return fail("vktb_flip_range: bad axis");
    if (!first || !last || !off) return fail("vktb_flip_range: null range");
    if (!toDesc("vktb_flip_range", dst, dv) || !toDesc("vktb_flip_range", src, sv)) return -1;
    if (bytesPerVoxel(dv.format) != bytesPerVoxel(sv.format))
        return fail("vktb_flip_range: dest and source must have the same bytes per voxel (raw bytes are exchanged)");
    if (emptyRange(first, last)) return 0;
    {   // footprint of the loops as written: indices [0, last - first) and their partners last - 1 - i = [first, last)
        const int32_t zero[3] = { 0, 0, 0 };
        const int32_t lo[3] = { first[0] < 0 ? first[0] : 0, first[1] < 0 ? first[1] : 0, first[2] < 0 ? first[2] : 0 };
        (void)zero;
        if (!inside(sv, lo, last)) return outside("vktb_flip_range", "source");
        if (!inside(dv, lo, last, off)) return outside("vktb_flip_range", "dst");
    }
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    LaunchCtx c{ currentStream(d), d->smCount };
    return finish("vktb_flip_range", launchFlip(c, dv, sv, first, last, off, axis));
}

int vktb_histogram_range(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], uint64_t* bins,
                         size_t numBins, uint64_t* droppedHost)
{
    VKTB_API_RANGE();
    VolDesc v;
    if (!first || !last || !bins) return fail("vktb_histogram_range: null argument");
    if (numBins == 0 || numBins > ((size_t)1 << 24)) return fail("vktb_histogram_range: numBins must be in [1, 2^24]");
    if (!toDesc("vktb_histogram_range", vol, v)) return -1;
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    cudaStream_t s = currentStream(d);
    LaunchCtx c{ s, d->smCount };
    unsigned long long* dropped = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&dropped, sizeof(unsigned long long), s);
    if (e != cudaSuccess) return fail("vktb_histogram_range", e);
    const StripedAlloc* a = emptyRange(first, last) ? nullptr : fanOutTarget(v, {});
    if (a) {
        // a striped volume: every device counts the planes that start in its stripe into bins of its own, the partial
        // bins travel to the calling device (peer copies) and are added there -- the one-process form of the per-rank
        // bins + all-reduce of volkit_b200/sharded.py
        const size_t rec = numBins + 1;                       // bins, then the dropped voxels
        unsigned long long* parts = nullptr;
        e = cudaMallocAsync((void**)&parts, a->stripes.size() * rec * sizeof(unsigned long long), s);
        int nParts = 0;
        if (e == cudaSuccess)
            e = forEachStripe(a, v.data, planeBytesOf(v), first[2], last[2], [&](const LaunchCtx& cg, int z0, int z1) {
                const int32_t f[3] = { first[0], first[1], z0 }, l[3] = { last[0], last[1], z1 };
                int dev = 0;
                cudaGetDevice(&dev);
                unsigned long long* mine = nullptr;
                cudaError_t pe = cudaMallocAsync((void**)&mine, rec * sizeof(unsigned long long), cg.stream);
                if (pe != cudaSuccess) return pe;
                pe = launchHistogram(cg, v, f, l, mine, numBins, mine + numBins);
                if (pe == cudaSuccess)
                    pe = cudaMemcpyPeerAsync(parts + (size_t)nParts * rec, d->ordinal, mine, dev, rec * sizeof(unsigned long long), cg.stream);
                ++nParts;
                cudaFreeAsync(mine, cg.stream);
                return pe;
            });
        if (e == cudaSuccess)
            e = launchHistCombine(c, reinterpret_cast<unsigned long long*>(bins), dropped, parts, nParts, numBins);
        if (parts) cudaFreeAsync(parts, s);
    } else {
        e = launchHistogram(c, v, first, last, reinterpret_cast<unsigned long long*>(bins), numBins, dropped);
    }
    if (e == cudaSuccess && droppedHost) {
        unsigned long long h = 0;
        e = cudaMemcpyAsync(&h, dropped, sizeof(h), cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        *droppedHost = h;
    }
    cudaFreeAsync(dropped, s);
    return finish("vktb_histogram_range", e);
}

// ---- Scale / Rotate / ApplyFilter ------------------------------------------------------------------------------
// The reference reads the source while it writes dest; with dest == source (or overlapping buffers) the serial walk
// would see its own output.  Filter works on a copy by construction (Filter_serial.hpp:29); for the affine ops the
// same copy is made here, which is what a caller who passes one volume twice can reasonably expect.
static cudaError_t privateCopyIfOverlapping(const VolDesc& dst, VolDesc& src, cudaStream_t s, void** tmp)
{
    *tmp = nullptr;
    if (!overlaps(dst, src)) return cudaSuccess;
    const size_t bytes = (size_t)src.dims[0] * src.dims[1] * src.dims[2] * bytesPerVoxel(src.format);
    cudaError_t e = cudaMallocAsync(tmp, bytes ? bytes : 1, s);
    if (e != cudaSuccess) return e;
    e = cudaMemcpyAsync(*tmp, src.data, bytes, cudaMemcpyDeviceToDevice, s);
    src.data = (uint8_t*)*tmp;
    return e;
}

int vktb_affine_range(const vktb_vol* dst, const vktb_vol* src, const int32_t first[3], const int32_t last[3],
                      const float matrix[9], const float center[3])
{
    VKTB_API_RANGE();
    VolDesc dv, sv;
    if (!first || !last || !matrix || !center) return fail("vktb_affine_range: null argument");
    if (!toDesc("vktb_affine_range", dst, dv) || !toDesc("vktb_affine_range", src, sv)) return -1;
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    cudaStream_t s = currentStream(d);
    LaunchCtx c{ s, d->smCount };
    void* tmp = nullptr;
    cudaError_t e = privateCopyIfOverlapping(dv, sv, s, &tmp);
    if (e == cudaSuccess) e = launchAffine(c, dv, sv, first, last, matrix, center);
    if (tmp) cudaFreeAsync(tmp, s);
    return finish("vktb_affine_range", e);
}

int vktb_apply_filter(const vktb_vol* dst, const vktb_vol* src, const float* weights, const int32_t fdims[3], int /*addressMode*/)
{
    VKTB_API_RANGE();
    VolDesc dv, sv;
    if (!weights || !fdims) return fail("vktb_apply_filter: null argument");
    if (!toDesc("vktb_apply_filter", dst, dv) || !toDesc("vktb_apply_filter", src, sv)) return -1;
    for (int a = 0; a < 3; ++a)
        if (fdims[a] < 1 || fdims[a] > 3)
            return fail("vktb_apply_filter: filter dims must be 1..3 per axis (the reference reads outside the volume otherwise)");
    for (int a = 0; a < 3; ++a)
        if (dv.dims[a] < sv.dims[a]) return fail("vktb_apply_filter: dest is smaller than source");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    cudaStream_t s = currentStream(d);
    LaunchCtx c{ s, d->smCount };
    void* tmp = nullptr;
    cudaError_t e = privateCopyIfOverlapping(dv, sv, s, &tmp);
    if (e == cudaSuccess) e = launchFilter(c, dv, sv, weights, fdims);
    if (tmp) cudaFreeAsync(tmp, s);
    return finish("vktb_apply_filter", e);
}

// ---- Aggregates ---------------------------------------------------------------------------------------------
static void emptyPartial(vktb_agg_partial* o)
{
    std::memset(o, 0, sizeof(*o));
    o->prod_mant = 1.0; o->min = FLT_MAX; o->max = -FLT_MAX;
}

int vktb_aggregates_partial(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], vktb_agg_partial* out)
{
    VKTB_API_RANGE();
    static_assert(sizeof(vktb_agg_partial) == 88, "vktb_agg_partial layout");
    VolDesc v;
    if (!first || !last || !out) return fail("vktb_aggregates_partial: null argument");
    if (!toDesc("vktb_aggregates_partial", vol, v)) return -1;
    emptyPartial(out);
    for (int a = 0; a < 3; ++a) {
        if (first[a] > last[a]) return fail("vktb_aggregates_partial: first > last (the reference's loops never end)");
        if (first[a] < 0 || last[a] > v.dims[a]) return fail("vktb_aggregates_partial: range outside the volume");
    }
    if (first[0] == last[0] || first[1] == last[1] || first[2] == last[2]) return 0;
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    cudaStream_t s = currentStream(d);
    LaunchCtx c{ s, d->smCount };
    AggPartial* scratch = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&scratch, aggregatesScratchBytes(c), s);
    if (e != cudaSuccess) return fail("vktb_aggregates_partial", e);
    AggPartial* totalDev = nullptr;
    AggPartial t;
    e = launchAggregates(c, v, first, last, scratch, &totalDev);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&t, totalDev, sizeof(t), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFreeAsync(scratch, s);
    if (e != cudaSuccess) return fail("vktb_aggregates_partial", e);
    const uint64_t nx = (uint64_t)(last[0] - first[0]), ny = (uint64_t)(last[1] - first[1]), nz = (uint64_t)(last[2] - first[2]);
    out->sum_d = t.sumD; out->sum_d2 = t.sumD2; out->prod_mant = t.prodMant; out->prod_exp = t.prodExp;
    out->count = nx * ny * nz;
    out->min = t.min; out->max = t.max; out->shift = t.shift;
    auto coords = [&](unsigned long long idx, int32_t* xyz) {
        const uint64_t r = idx / nx;
        xyz[0] = first[0] + (int32_t)(idx - r * nx);
        xyz[1] = first[1] + (int32_t)(r % ny);
        xyz[2] = first[2] + (int32_t)(r / ny);
    };
    if (t.minIdx != ~0ull) { out->has_min = 1; coords(t.minIdx, out->argmin); }
    if (t.maxIdx != ~0ull) { out->has_max = 1; coords(t.maxIdx, out->argmax); }
    return 0;
}

int vktb_aggregates_finish(const vktb_agg_partial* parts, int n, uint64_t numElems, vktb_aggregates* out)
{
    VKTB_API_RANGE();
    if (!out || n < 0 || (n > 0 && !parts)) return fail("vktb_aggregates_finish: bad arguments");
    std::memset(out, 0, sizeof(*out));
    out->min = FLT_MAX; out->max = -FLT_MAX;
    double s1 = 0.0, s2 = 0.0, cnt = 0.0, mant = 1.0;
    long long ex = 0;
    const double c0 = n > 0 ? (double)parts[0].shift : 0.0;
    for (int k = 0; k < n; ++k) {
        const vktb_agg_partial& p = parts[k];
        if (p.has_min && p.min < out->min) { out->min = p.min; std::memcpy(out->argmin, p.argmin, sizeof(p.argmin)); }
        if (p.has_max && p.max > out->max) { out->max = p.max; std::memcpy(out->argmax, p.argmax, sizeof(p.argmax)); }
        const double dc = (double)p.shift - c0, m = (double)p.count;     // re-base the moments on c0
        s1 += p.sum_d + m * dc;
        s2 += p.sum_d2 + 2.0 * dc * p.sum_d + m * dc * dc;
        cnt += m;
        int e2 = 0;
        mant *= p.prod_mant;
        if (std::isfinite(mant) && mant != 0.0) { mant = std::frexp(mant, &e2); ex += e2; }
        ex += p.prod_exp;
    }
    const float sumF = (float)(s1 + cnt * c0);
    out->sum = sumF;
    // the reference divides its binary32 accumulator by (double)numElems and stores a float (Aggregates_serial.hpp:62-64)
    out->mean = (float)((double)sumF / (double)numElems);
    const double dm = (double)out->mean - c0;
    double vs = s2 - 2.0 * dm * s1 + cnt * dm * dm;
    if (vs < 0.0) vs = 0.0;
    out->var = (float)((double)(float)vs / (double)numElems);           // :81
    out->stddev = sqrtf(out->var);                                       // :82
    if (ex > 4000) ex = 4000;
    if (ex < -4000) ex = -4000;
    out->prod = (float)std::ldexp(mant, (int)ex);
    return 0;
}

int vktb_aggregates_range(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], vktb_aggregates* out)
{
    VKTB_API_RANGE();
    static_assert(sizeof(vktb_aggregates) == 52, "vktb_aggregates must match vkt::Aggregates");
    if (!vol || !out) return fail("vktb_aggregates_range: null argument");
    vktb_agg_partial p;
    if (vktb_aggregates_partial(vol, first, last, &p) != 0) return -1;
    const uint64_t numElems = (uint64_t)vol->dims[0] * (uint64_t)vol->dims[1] * (uint64_t)vol->dims[2];
    return vktb_aggregates_finish(&p, 1, numElems, out);
}

// Host-side slab planner: the reference's own coordinate expression (Resample_serial.hpp:60-62) evaluated in
// binary32; it is monotone in its argument.
static int32_t srcCoordHost(int32_t d, int32_t D, int32_t S)
{
    volatile float q = (float)d / (float)D;    // volatile: one rounding per operation, no contraction
    volatile float s = q * (float)S;
    float sv = s;
    return (sv >= -2147483648.0f && sv < 2147483648.0f) ? (int32_t)sv : INT32_MIN;
}

static int planSrcRange(int filter, int32_t dz0, int32_t dz1, int32_t D, int32_t S, bool spill, int32_t* sz0, int32_t* sz1)
{
    if (!sz0 || !sz1 || D <= 0 || S <= 0 || dz1 <= dz0) return fail("vktb_resample_window_src_range: bad arguments");
    int32_t lo = srcCoordHost(dz0, D, S), hi = srcCoordHost(dz1 - 1, D, S);
    lo = lo < 0 ? 0 : (lo > S - 1 ? S - 1 : lo);
    hi = hi < 0 ? 0 : (hi > S - 1 ? S - 1 : hi);
    int32_t end = hi + 1;                                   // exclusive
    if (filter != VKTB_NEAREST) end = hi + 2;               // +1 tap plane
    if (filter == VKTB_LINEAR && spill) end = hi + 3;       // + the plane the unclamped hi.x tap can spill into
    if (end > S) end = S;
    *sz0 = lo; *sz1 = end;
    return 0;
}

// z-only form: assumes the reference's unclamped hi.x tap (Linear) may spill into the following plane
int vktb_resample_window_src_range(int filter, int32_t dz0, int32_t dz1, int32_t D, int32_t S, int32_t* sz0, int32_t* sz1)
{
    VKTB_API_RANGE();
    return planSrcRange(filter, dz0, dz1, D, S, true, sz0, sz1);
}

// full-dims form: the spill plane is requested only when it can be read at all -- the last dst column must map to
// the last src column (hx = Sx leaves the row) and some dst row must tap the last src row (so that the row after
// it is the first row of the next plane)
int vktb_resample_window_src_range3(int filter, int32_t dz0, int32_t dz1, const int32_t D[3], const int32_t S[3],
                                    int32_t* sz0, int32_t* sz1)
{
    VKTB_API_RANGE();
    if (!D || !S) return fail("vktb_resample_window_src_range3: null dims");
    bool spill = false;
    if (filter == VKTB_LINEAR && D[0] > 0 && D[1] > 0 && S[0] > 0 && S[1] > 0)
        spill = srcCoordHost(D[0] - 1, D[0], S[0]) >= S[0] - 1 && srcCoordHost(D[1] - 1, D[1], S[1]) + 1 >= S[1] - 1;
    return planSrcRange(filter, dz0, dz1, D[2], S[2], spill, sz0, sz1);
}

// ---- Headless ray-marcher ---------------------------------------------------------------------------------------
int vktb_render_default_camera(const int32_t dims[3], const float dist[3], vktb_camera* cam)
{
    VKTB_API_RANGE();
    if (!dims || !cam) return fail("vktb_render_default_camera: null argument");
    const float one[3] = { 1.f, 1.f, 1.f };
    renderDefaultCamera(dims, dist ? dist : one, cam->eye, cam->center, cam->up);
    cam->fovy_degrees = 45.f;
    return 0;
}

int vktb_render_raymarch(const vktb_vol* vol, const float dist[3], const vktb_camera* cam, const float* lutDev, int lutN,
                         float dt, int srgb, int width, int height, uint8_t* rgba8Dev, float* rgba32fDev)
{
    VKTB_API_RANGE();
    const char* fn = "vktb_render_raymarch";
    VolDesc v;
    if (!cam) return fail("vktb_render_raymarch: null camera");
    if (!toDesc(fn, vol, v)) return -1;
    if (!renderFormatSupported(v.format))
        return fail("vktb_render_raymarch: UInt8, UInt16, Int16 and UInt32 volumes only (the reference normalises a voxel by "
                    "its texel type's numeric limits, Render_kernel.hpp:31-39, which is meaningless for Float32)");
    if ((size_t)v.dims[0] * v.dims[1] * v.dims[2] == 0) return fail("vktb_render_raymarch: empty volume");
    if (width <= 0 || height <= 0 || width > 32768 || height > 32768) return fail("vktb_render_raymarch: bad viewport");
    if (!(dt > 0.f)) return fail("vktb_render_raymarch: dt must be positive (the march would never end)");
    if (lutN < 0 || (lutN > 0 && !lutDev)) return fail("vktb_render_raymarch: bad look-up table");
    if (!rgba8Dev && !rgba32fDev) return fail("vktb_render_raymarch: no output buffer");
    const float one[3] = { 1.f, 1.f, 1.f };
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    LaunchCtx c{ currentStream(d), d->smCount };
    return finish(fn, launchRayMarch(c, v, dist ? dist : one, cam->eye, cam->center, cam->up, cam->fovy_degrees, lutDev, lutN,
                                     dt, srgb, width, height, rgba8Dev, rgba32fDev));
}

// ---- Transform registry ----
int vktb_unary_op_by_name(const char* name, int* op)
{
    VKTB_API_RANGE();
    if (!name || !op) return fail("vktb_unary_op_by_name: null argument");
    for (int i = 0; i < kNumUnaryOps; ++i) if (!strcmp(name, kUnaryNames[i])) { *op = i; return 0; }
    return fail("vktb_unary_op_by_name: no such built-in device op");
}

int vktb_binary_op_by_name(const char* name, int* op)
{
    VKTB_API_RANGE();
    if (!name || !op) return fail("vktb_binary_op_by_name: null argument");
    for (int i = 0; i < kNumBinaryOps; ++i) if (!strcmp(name, kBinaryNames[i])) { *op = i; return 0; }
    return fail("vktb_binary_op_by_name: no such built-in device op");
}

int vktb_bind_unary_op(const void* hostFn, int op)
{
    VKTB_API_RANGE();
    if (!hostFn || op < 0 || (op >= kNumUnaryOps && !userOpValid(op, false))) return fail("vktb_bind_unary_op: bad arguments");
    std::lock_guard<std::mutex> lock(g_opMutex);
    g_unaryBound[hostFn] = op;
    return 0;
}

int vktb_bind_binary_op(const void* hostFn, int op)
{
    VKTB_API_RANGE();
    if (!hostFn || op < 0 || (op >= kNumBinaryOps && !userOpValid(op, true))) return fail("vktb_bind_binary_op: bad arguments");
    std::lock_guard<std::mutex> lock(g_opMutex);
    g_binaryBound[hostFn] = op;
    return 0;
}

int vktb_lookup_unary_op(const void* hostFn, int* op)
{
    VKTB_API_RANGE();
    std::lock_guard<std::mutex> lock(g_opMutex);
    auto it = g_unaryBound.find(hostFn);
    if (it == g_unaryBound.end() || !op)
        return fail("Transform: host function pointer has no device op bound (vktb_bind_unary_op); no CPU fallback on the GPU policy");
    *op = it->second;
    return 0;
}

int vktb_lookup_binary_op(const void* hostFn, int* op)
{
    VKTB_API_RANGE();
    std::lock_guard<std::mutex> lock(g_opMutex);
    auto it = g_binaryBound.find(hostFn);
    if (it == g_binaryBound.end() || !op)
        return fail("Transform: host function pointer has no device op bound (vktb_bind_binary_op); no CPU fallback on the GPU policy");
    *op = it->second;
    return 0;
}

int vktb_debug_div_proven(float lo, float hi, int* proven)
{
    VKTB_API_RANGE();
    if (!proven) return fail("vktb_debug_div_proven: null argument");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    LaunchCtx c{ currentStream(d), d->smCount };
    *proven = divisorProven(c, lo, hi);
    return 0;
}

int vktb_debug_map_unmap(int format, float lo, float hi, const float* values, uint32_t* mapped, float* unmapped, size_t n)
{
    VKTB_API_RANGE();
    if (!supportedFormat(format)) return fail("vktb_debug_map_unmap: unsupported format");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    LaunchCtx c{ currentStream(d), d->smCount };
    return finish("vktb_debug_map_unmap", launchDebugCodec(c, format, makeMapping(lo, hi), values, mapped, unmapped, n));
}

} // extern "C"
