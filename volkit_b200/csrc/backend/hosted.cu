// Host-resident volumes streamed through the device in z-slabs (vktb_hosted_* entry points).
//
// What the reference does for a CPU-resident volume under the GPU policy is ManagedBuffer::migrate()
// (reference include/cpp/vkt/ManagedBuffer.hpp:168-198): one blocking whole-buffer H2D copy per operand, the kernel,
// and later one blocking whole-buffer D2H copy -- three serial phases, each leaving either the PCIe link or the GPU
// idle, and the whole working set must fit in HBM.  Every algorithm of this path is local in z (arithmetic,
// transform: plane -> same plane; resample: dst plane <- a monotone src window, vktb_resample_window_src_range3),
// so the hosted entry points cut the range into z-slabs and run a three-stage pipeline
//
//     upload stream   : H2D of slab i+1's source planes          (copy engine 1)
//     compute stream  : the ordinary kernel on slab i            (SMs)
//     download stream : D2H of slab i-1's destination planes     (copy engine 2, PCIe is full duplex)
//
// over a ring of kSlots device slots, chained with events.  The kernels are the same launchers the device-resident
// entry points use, so results are bit-identical; volumes larger than HBM work; the volumes stay on the host.
// Host buffers should be pinned (vktb_host_alloc or cudaHostRegister) -- pageable memory is accepted, but then the
// runtime stages every copy and the overlap is lost.
#include "context.hpp"

#include <algorithm>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <string>
#include <vector>

#include "../../../include/vktb/backend.h"
#include "../kernels/launch.hpp"

using namespace vktb;

namespace {

constexpr int kSlots = 3;

struct Xfer { void* dev; void* host; size_t bytes; };

struct Slab
{
    std::vector<Xfer> in, out;
    std::function<cudaError_t(const LaunchCtx&)> launch;
};

// streams and events of the pipeline, created once per device
struct PipeState
{
    bool         ready = false;
    cudaStream_t up = nullptr, down = nullptr;
    cudaEvent_t  uploaded[kSlots] = {}, computed[kSlots] = {}, drained[kSlots] = {}, start = nullptr;
    std::mutex   busy;   // one hosted call at a time per device
};

PipeState* pipeFor(DeviceCtx* d)
{
    static std::mutex m;
    static std::map<int, PipeState*> states;
    std::lock_guard<std::mutex> lock(m);
    PipeState*& p = states[d->ordinal];
    if (!p) p = new PipeState;
    if (!p->ready) {
        if (cudaStreamCreateWithFlags(&p->up, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        if (cudaStreamCreateWithFlags(&p->down, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        for (int i = 0; i < kSlots; ++i) {
            cudaEventCreateWithFlags(&p->uploaded[i], cudaEventDisableTiming);
            cudaEventCreateWithFlags(&p->computed[i], cudaEventDisableTiming);
            cudaEventCreateWithFlags(&p->drained[i], cudaEventDisableTiming);
        }
        cudaEventCreateWithFlags(&p->start, cudaEventDisableTiming);
        p->ready = true;
    }
    return p;
}

// bytes of the largest operand per slab: 64 MiB keeps the pipeline fill (first upload, last download) below 1 % of
// a multi-GiB call while three slots of three operands stay under 1 GiB of HBM.  VKTB_HOSTED_SLAB_KB overrides
// (read per call; the tests use it to force many slabs on small volumes).
size_t slabTargetBytes()
{
    const char* e = getenv("VKTB_HOSTED_SLAB_KB");
    const long kb = e ? atol(e) : 0;
    return kb > 0 ? (size_t)kb << 10 : (size_t)64 << 20;
}

// every operand of a slot ends with at least 16 spare bytes: the funnel-shifted loads of the *Range kernels read the
// aligned vector that encloses the last source byte, which may end up to 15 bytes past the resident planes
inline size_t alignUp(size_t n) { return (n + 16 + 255) & ~(size_t)255; }

size_t planeBytes(const VolDesc& v) { return (size_t)v.dims[0] * v.dims[1] * bytesPerVoxel(v.format); }

// describe(chunk, slotBase, slab) fills in the transfers and the launch of one chunk, carving its device buffers from
// slotBase (slotBytes are reserved per slot).  Returns on completion of the last download.
int runPipeline(const char* fn, DeviceCtx* d, int nChunks, size_t slotBytes,
                const std::function<void(int, uint8_t*, Slab&)>& describe)
{
    if (nChunks <= 0) return 0;
    PipeState* p = pipeFor(d);
    if (!p) return fail(fn, cudaGetLastError());
    std::lock_guard<std::mutex> lock(p->busy);

    const int slots = std::min(kSlots, nChunks);
    void* ring = nullptr;
    if (vktb_alloc(&ring, slotBytes * slots) != 0) return -1;
    cudaStream_t compute = currentStream(d);
    LaunchCtx ctx{ compute, d->smCount };
    // the ring is allocated in compute-stream order: the copy streams may touch it only after that point
    cudaEventRecord(p->start, compute);
    cudaStreamWaitEvent(p->up, p->start, 0);
    cudaStreamWaitEvent(p->down, p->start, 0);

    cudaError_t err = cudaSuccess;
    Slab slab;
    for (int c = 0; c < nChunks && err == cudaSuccess; ++c) {
        const int s = c % slots;
        slab.in.clear(); slab.out.clear(); slab.launch = nullptr;
        describe(c, (uint8_t*)ring + (size_t)s * slotBytes, slab);

        if (c >= slots) cudaStreamWaitEvent(p->up, p->drained[s], 0);       // slot s is free again
        for (const Xfer& x : slab.in)
            if (x.bytes && err == cudaSuccess) err = cudaMemcpyAsync(x.dev, x.host, x.bytes, cudaMemcpyHostToDevice, p->up);
        cudaEventRecord(p->uploaded[s], p->up);

        cudaStreamWaitEvent(compute, p->uploaded[s], 0);
        if (err == cudaSuccess && slab.launch) err = slab.launch(ctx);
        cudaEventRecord(p->computed[s], compute);

        cudaStreamWaitEvent(p->down, p->computed[s], 0);
        for (const Xfer& x : slab.out)
            if (x.bytes && err == cudaSuccess) err = cudaMemcpyAsync(x.host, x.dev, x.bytes, cudaMemcpyDeviceToHost, p->down);
        cudaEventRecord(p->drained[s], p->down);
    }
    // results are complete in host memory on return (the contract of the D2H copy in migrate()); the ring goes back
    // to the pool in compute-stream order, after everything that used it
    cudaError_t e2 = cudaStreamSynchronize(p->down);
    cudaStreamSynchronize(p->up);
    cudaStreamSynchronize(compute);
    vktb_free(ring);
    if (err == cudaSuccess) err = e2;
    return err == cudaSuccess ? 0 : fail(fn, err);
}

bool supportedFormat(int f) { return f == F_U8 || f == F_U16 || f == F_U32 || f == F_F32 || f == F_INT16; }

bool toDesc(const char* fn, const vktb_vol* v, VolDesc& d)
{
    if (!v) { fail((std::string(fn) + ": null volume").c_str()); return false; }
    if (!supportedFormat(v->format)) { fail((std::string(fn) + ": unsupported DataFormat").c_str()); return false; }
    if (v->dims[0] < 0 || v->dims[1] < 0 || v->dims[2] < 0) { fail((std::string(fn) + ": negative dims").c_str()); return false; }
    const size_t n = (size_t)v->dims[0] * v->dims[1] * v->dims[2];
    if (n && !v->data) { fail((std::string(fn) + ": null data pointer").c_str()); return false; }
    d.data = (uint8_t*)v->data;
    d.dims[0] = v->dims[0]; d.dims[1] = v->dims[1]; d.dims[2] = v->dims[2];
    d.format = v->format;
    d.map = makeMapping(v->lo, v->hi);
    return true;
}

// [first,last) + shift inside v, per axis
bool inside(const VolDesc& v, const int32_t first[3], const int32_t last[3], const int32_t* shift)
{
    for (int i = 0; i < 3; ++i) {
        const long long s = shift ? shift[i] : 0;
        if (first[i] + s < 0 || last[i] + s > v.dims[i]) return false;
    }
    return true;
}

// the same volume with its data on the device and only the planes [z0, z0 + nz) resident: the descriptor keeps the
// full dims and a data pointer displaced by z0 planes, so absolute coordinates and ranges pass through unchanged
// (nothing outside the resident planes is ever dereferenced).
VolDesc residentWindow(const VolDesc& hostVol, uint8_t* dev, int z0)
{
    VolDesc w = hostVol;
    w.data = dev - (size_t)z0 * planeBytes(hostVol);
    return w;
}

// a slab of nz planes as a stand-alone volume
VolDesc slabVolume(const VolDesc& hostVol, uint8_t* dev, int nz)
{
    VolDesc w = hostVol;
    w.data = dev;
    w.dims[2] = nz;
    return w;
}

bool bytesOverlap(const VolDesc& a, const VolDesc& b)
{
    return a.data != b.data && overlaps(a, b);
}

} // namespace

extern "C" {

int vktb_hosted_arith_range(int op, const vktb_vol* dst, const vktb_vol* s1, const vktb_vol* s2, const int32_t first[3],
                            const int32_t last[3], const int32_t off[3])
{
    VKTB_API_RANGE();
    const char* fn = "vktb_hosted_arith_range";
    VolDesc dv, av, bv;
    if (op < 0 || op > VKTB_SAFEABSDIFF) return fail("vktb_hosted_arith_range: bad op");
    if (!first || !last || !off) return fail("vktb_hosted_arith_range: null range");
    if (!toDesc(fn, dst, dv) || !toDesc(fn, s1, av) || !toDesc(fn, s2, bv)) return -1;
    if (last[0] <= first[0] || last[1] <= first[1] || last[2] <= first[2]) return 0;
    if (!inside(av, first, last, nullptr) || !inside(bv, first, last, nullptr) || !inside(dv, first, last, off))
        return fail("vktb_hosted_arith_range: range outside a volume");
    // in-place use (dst is one of the sources, same bytes, no offset) streams like any other call; partially
    // overlapping host buffers would make slab order observable and are refused
    const bool shifted = off[0] || off[1] || off[2];
    if (bytesOverlap(dv, av) || bytesOverlap(dv, bv) || ((dv.data == av.data || dv.data == bv.data) && shifted))
        return fail("vktb_hosted_arith_range: dst overlaps a source at a different position");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;

    const size_t pa = planeBytes(av), pb = planeBytes(bv), pd = planeBytes(dv);
    const int nzTotal = last[2] - first[2];
    const int cz = (int)std::max<size_t>(1, std::min<size_t>(nzTotal, slabTargetBytes() / std::max({ pa, pb, pd, (size_t)1 })));
    const int nChunks = (nzTotal + cz - 1) / cz;
    // dst planes not fully covered by the range keep their other voxels: upload them first
    const bool dstWhole = first[0] + off[0] == 0 && first[1] + off[1] == 0
                       && last[0] - first[0] == dv.dims[0] && last[1] - first[1] == dv.dims[1];
    const size_t slotBytes = alignUp(pa * cz) + alignUp(pb * cz) + alignUp(pd * cz);

    return runPipeline(fn, d, nChunks, slotBytes, [&](int c, uint8_t* base, Slab& slab) {
        const int z = c * cz, nz = std::min(cz, nzTotal - z);
        uint8_t* da = base;
        uint8_t* db = da + alignUp(pa * cz);
        uint8_t* dd = db + alignUp(pb * cz);
        const int sz = first[2] + z, dz = first[2] + off[2] + z;
        slab.in.push_back({ da, av.data + (size_t)sz * pa, pa * nz });
        slab.in.push_back({ db, bv.data + (size_t)sz * pb, pb * nz });
        if (!dstWhole) slab.in.push_back({ dd, dv.data + (size_t)dz * pd, pd * nz });
        slab.out.push_back({ dd, dv.data + (size_t)dz * pd, pd * nz });
        const VolDesc A = slabVolume(av, da, nz), B = slabVolume(bv, db, nz), D = slabVolume(dv, dd, nz);
        const int32_t f[3] = { first[0], first[1], 0 }, l[3] = { last[0], last[1], nz }, o[3] = { off[0], off[1], 0 };
        slab.launch = [=](const LaunchCtx& ctx) { return launchArith(ctx, op, D, A, B, f, l, o); };
    });
}

int vktb_hosted_transform_range1(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], int op)
{
    VKTB_API_RANGE();
    const char* fn = "vktb_hosted_transform_range1";
    VolDesc v;
    if (op < 0 || op >= kNumUnaryOps) return fail("vktb_hosted_transform_range1: unknown device op handle");
    if (!first || !last) return fail("vktb_hosted_transform_range1: null range");
    if (!toDesc(fn, vol, v)) return -1;
    if (last[0] <= first[0] || last[1] <= first[1] || last[2] <= first[2]) return 0;
    if (!inside(v, first, last, nullptr)) return fail("vktb_hosted_transform_range1: range outside the volume");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;

    const size_t pv = planeBytes(v);
    const int nzTotal = last[2] - first[2];
    const int cz = (int)std::max<size_t>(1, std::min<size_t>(nzTotal, slabTargetBytes() / std::max(pv, (size_t)1)));
    const int nChunks = (nzTotal + cz - 1) / cz;
    return runPipeline(fn, d, nChunks, alignUp(pv * cz), [&](int c, uint8_t* base, Slab& slab) {
        const int z0 = first[2] + c * cz, nz = std::min(cz, last[2] - z0);
        slab.in.push_back({ base, v.data + (size_t)z0 * pv, pv * nz });
        slab.out.push_back({ base, v.data + (size_t)z0 * pv, pv * nz });
        const VolDesc W = residentWindow(v, base, z0);    // the op sees absolute z
        const int32_t f[3] = { first[0], first[1], z0 }, l[3] = { last[0], last[1], z0 + nz };
        slab.launch = [=](const LaunchCtx& ctx) { return launchTransform1(ctx, W, f, l, op); };
    });
}

int vktb_hosted_resample(const vktb_vol* dst, const vktb_vol* src, int filter)
{
    VKTB_API_RANGE();
    const char* fn = "vktb_hosted_resample";
    VolDesc dv, sv;
    if (filter < 0 || filter > VKTB_TRILINEAR) return fail("vktb_hosted_resample: bad filter mode");
    if (!toDesc(fn, dst, dv) || !toDesc(fn, src, sv)) return -1;
    const size_t nd = (size_t)dv.dims[0] * dv.dims[1] * dv.dims[2], ns = (size_t)sv.dims[0] * sv.dims[1] * sv.dims[2];
    if (nd == 0) return 0;
    if (ns == 0) return fail("vktb_hosted_resample: empty source");
    if (overlaps(dv, sv)) return fail("vktb_hosted_resample: dst overlaps src");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;

    const bool same = dv.dims[0] == sv.dims[0] && dv.dims[1] == sv.dims[1] && dv.dims[2] == sv.dims[2];
    const size_t pd = planeBytes(dv), ps = planeBytes(sv);
    const int Dz = dv.dims[2], Sz = sv.dims[2];
    // slab thickness: bounded by the dst slab and by the src window it pulls in (window ~ cz * Sz / Dz + 2 planes)
    size_t czD = slabTargetBytes() / std::max(pd, (size_t)1);
    size_t czS = (size_t)((double)(slabTargetBytes() / std::max(ps, (size_t)1)) * Dz / std::max(Sz, 1));
    const int cz = (int)std::max<size_t>(1, std::min<size_t>({ (size_t)Dz, czD, std::max<size_t>(czS, 1) }));
    const int nChunks = (Dz + cz - 1) / cz;

    // src window of every chunk: the planner is the reference's own coordinate expression
    std::vector<int32_t> w0(nChunks), w1(nChunks);
    int maxWin = 1;
    for (int c = 0; c < nChunks; ++c) {
        const int z0 = c * cz, z1 = std::min(Dz, z0 + cz);
        if (same) { w0[c] = z0; w1[c] = z1; }
        else if (vktb_resample_window_src_range3(filter, z0, z1, dv.dims, sv.dims, &w0[c], &w1[c]) != 0) return -1;
        maxWin = std::max(maxWin, w1[c] - w0[c]);
    }
    const size_t slotBytes = alignUp(pd * cz) + alignUp(ps * maxWin);
    return runPipeline(fn, d, nChunks, slotBytes, [&](int c, uint8_t* base, Slab& slab) {
        const int z0 = c * cz, nz = std::min(cz, Dz - z0);
        uint8_t* dd = base;
        uint8_t* ds = base + alignUp(pd * cz);
        const int wn = w1[c] - w0[c];
        slab.in.push_back({ ds, sv.data + (size_t)w0[c] * ps, ps * wn });
        slab.out.push_back({ dd, dv.data + (size_t)z0 * pd, pd * nz });
        const VolDesc D = slabVolume(dv, dd, nz), S = slabVolume(sv, ds, wn);
        const int sz0 = w0[c];
        const int32_t Df[3] = { dv.dims[0], dv.dims[1], dv.dims[2] }, Sf[3] = { sv.dims[0], sv.dims[1], sv.dims[2] };
        slab.launch = [=](const LaunchCtx& ctx) { return launchResample(ctx, D, S, filter, z0, Df, sz0, Sf); };
    });
}

} // extern "C"
