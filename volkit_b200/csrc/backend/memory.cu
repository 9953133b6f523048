// Device-buffer manager behind vktb_alloc / vktb_free / vktb_memcpy / vktb_host_alloc (see context.hpp).
#include <atomic>
#include <cstdlib>

#include <nvtx3/nvToolsExt.h>

#include "context.hpp"

#include <algorithm>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "../../../include/vktb/backend.h"
#include "../kernels/launch.hpp"
#include "multi.hpp"

namespace vktb {

ApiRange::ApiRange(const char* name) { nvtxRangePushA(name); }
ApiRange::~ApiRange() { nvtxRangePop(); }

bool injectAllocFailure()
{
    static std::atomic<long> count{0};
    const char* e = getenv("VKTB_FAIL_ALLOC_AT");
    if (!e) { count.store(0); return false; }
    return count.fetch_add(1) + 1 == atol(e);
}

// ------------------------------------------------------------------ errors ------------------------------
static thread_local std::string t_error;
static thread_local int t_device = 0;
static thread_local cudaStream_t t_streamOverride = nullptr;
static thread_local bool t_hasOverride = false;

void setError(const std::string& msg) { t_error = msg; }
int fail(const char* what, cudaError_t e)
{
    // the error is reported HERE: clear the runtime's (non-sticky) last-error slot, or a later launcher's
    // cudaGetLastError() would blame its own, successful launch for it
    cudaGetLastError();
    t_error = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    return (int)e ? (int)e : -1;
}
int fail(const char* what) { t_error = what; return -1; }

// ------------------------------------------------------------------ devices -----------------------------
static std::mutex g_devMutex;
static DeviceCtx* g_devices[64] = {};

// context of device `ord`, created on first use; makes the device current.  nullptr + error string on failure.
DeviceCtx* deviceCtx(int ord)
{
    if (ord < 0 || ord >= 64) { fail("vktb: device ordinal out of range"); return nullptr; }
    DeviceCtx* d = g_devices[ord];
    if (d && d->ready) {
        cudaSetDevice(ord);
        return d;
    }
    std::lock_guard<std::mutex> lock(g_devMutex);
    d = g_devices[ord];
    if (d && d->ready) { cudaSetDevice(ord); return d; }
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        fail("vktb: no usable CUDA device (this backend has no CPU fallback)", e != cudaSuccess ? e : cudaErrorNoDevice);
        return nullptr;
    }
    if (ord >= n) { fail("vktb: device ordinal >= device count"); return nullptr; }
    if ((e = cudaSetDevice(ord)) != cudaSuccess) { fail("cudaSetDevice", e); return nullptr; }
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, ord)) != cudaSuccess) { fail("cudaGetDeviceProperties", e); return nullptr; }
    if (prop.major != 10) {
        char msg[160];
        snprintf(msg, sizeof msg, "vktb: device %d is sm_%d%d; this library carries sm_100a code only", ord, prop.major, prop.minor);
        fail(msg);
        return nullptr;
    }
    d = new DeviceCtx();
    d->ordinal = ord;
    d->smCount = prop.multiProcessorCount;
    if ((e = cudaStreamCreateWithFlags(&d->compute, cudaStreamNonBlocking)) != cudaSuccess) { fail("cudaStreamCreate", e); return nullptr; }
    if ((e = cudaStreamCreateWithFlags(&d->copy, cudaStreamNonBlocking)) != cudaSuccess) { fail("cudaStreamCreate", e); return nullptr; }
    // keep freed device memory cached in the pool (the alloc/free pair of every migrate() becomes a pool hit)
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, ord) == cudaSuccess) {
        uint64_t threshold = UINT64_MAX;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
    }
    d->ready = true;
    g_devices[ord] = d;
    return d;
}

DeviceCtx* currentDevice() { return deviceCtx(t_device); }
int currentOrdinal() { return t_device; }
void setCurrentOrdinal(int ord) { t_device = ord; }

cudaStream_t currentStream(DeviceCtx* d) { return t_hasOverride ? t_streamOverride : d->compute; }

static int ensureStage(DeviceCtx* d)
{
    if (d->stage[0]) return 0;
    for (int i = 0; i < kStageSlots; ++i) {
        cudaError_t e = cudaHostAlloc(&d->stage[i], kStageBytes, cudaHostAllocDefault);
        if (e != cudaSuccess) return fail("cudaHostAlloc(staging)", e);
        if ((e = cudaEventCreateWithFlags(&d->stageDone[i], cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
        d->stageInFlight[i] = false;
    }
    return 0;
}

// ------------------------------------------------------------------ pinned pool -------------------------
static std::mutex g_pinMutex;
static std::map<uintptr_t, PinnedBlock> g_pinLive;                 // by base address
static std::multimap<size_t, PinnedBlock> g_pinFree;               // by capacity
static size_t g_pinInUse = 0, g_pinReserved = 0;

static size_t bucketOf(size_t bytes)
{
    // 2 MiB granularity up to 64 MiB, then 1/8-octave steps: bounded internal waste, good reuse
    const size_t g = 2u << 20;
    if (bytes <= (64u << 20)) return (bytes + g - 1) / g * g;
    size_t p = 1; while (p * 2 <= bytes) p *= 2;
    size_t step = p / 8;
    return (bytes + step - 1) / step * step;
}

int pinnedAlloc(void** out, size_t bytes)
{
    *out = nullptr;
    if (bytes == 0) bytes = 1;
    const size_t cap = bucketOf(bytes);
    PinnedBlock b;
    bool found = false;
    {
        std::lock_guard<std::mutex> lock(g_pinMutex);
        auto it = g_pinFree.lower_bound(cap);
        if (it != g_pinFree.end() && it->first <= cap + cap / 4) { b = it->second; g_pinFree.erase(it); found = true; }
    }
    if (found) {
        if (b.busy) { cudaEventSynchronize(b.busy); cudaEventDestroy(b.busy); b.busy = nullptr; }
    } else {
        cudaError_t e = cudaHostAlloc(&b.ptr, cap, cudaHostAllocPortable);
        if (e != cudaSuccess) return fail("cudaHostAlloc", e);
        b.capacity = cap;
        std::lock_guard<std::mutex> lock(g_pinMutex);
        g_pinReserved += cap;
    }
    std::lock_guard<std::mutex> lock(g_pinMutex);
    g_pinLive[reinterpret_cast<uintptr_t>(b.ptr)] = b;
    g_pinInUse += b.capacity;
    *out = b.ptr;
    return 0;
}

static PinnedBlock* findLive(const void* p)
{
    auto it = g_pinLive.upper_bound(reinterpret_cast<uintptr_t>(p));
    if (it == g_pinLive.begin()) return nullptr;
    --it;
    if (reinterpret_cast<uintptr_t>(p) < it->first + it->second.capacity) return &it->second;
    return nullptr;
}

bool pinnedOwns(const void* p)
{
    std::lock_guard<std::mutex> lock(g_pinMutex);
    return findLive(p) != nullptr;
}

void pinnedMarkBusy(const void* p, cudaStream_t s)
{
    std::lock_guard<std::mutex> lock(g_pinMutex);
    PinnedBlock* b = findLive(p);
    if (!b) return;
    if (!b->busy) cudaEventCreateWithFlags(&b->busy, cudaEventDisableTiming);
    cudaEventRecord(b->busy, s);
}

int pinnedFree(void* p)
{
    if (!p) return 0;
    std::lock_guard<std::mutex> lock(g_pinMutex);
    auto it = g_pinLive.find(reinterpret_cast<uintptr_t>(p));
    if (it == g_pinLive.end()) return fail("vktb_host_free: pointer was not returned by vktb_host_alloc");
    PinnedBlock b = it->second;
    g_pinLive.erase(it);
    g_pinInUse -= b.capacity;
    g_pinFree.insert({ b.capacity, b });
    return 0;
}

void pinnedStats(size_t* inUse, size_t* reserved)
{
    std::lock_guard<std::mutex> lock(g_pinMutex);
    if (inUse) *inUse = g_pinInUse;
    if (reserved) *reserved = g_pinReserved;
}

void pinnedTrim()
{
    std::lock_guard<std::mutex> lock(g_pinMutex);
    for (auto& kv : g_pinFree) {
        if (kv.second.busy) { cudaEventSynchronize(kv.second.busy); cudaEventDestroy(kv.second.busy); }
        cudaFreeHost(kv.second.ptr);
        g_pinReserved -= kv.second.capacity;
    }
    g_pinFree.clear();
}

} // namespace vktb

using namespace vktb;

// ================================================================ C-ABI: library / device ================
extern "C" {

int vktb_abi_version(void) { return VKTB_ABI_VERSION; }
const char* vktb_last_error(void) { return t_error.c_str(); }

int vktb_device_count(int* n)
{
    VKTB_API_RANGE();
    if (!n) return fail("vktb_device_count: null argument");
    cudaError_t e = cudaGetDeviceCount(n);
    if (e != cudaSuccess) { *n = 0; return fail("cudaGetDeviceCount", e); }
    return 0;
}

int vktb_set_device(int ordinal)
{
    VKTB_API_RANGE();
    if (ordinal < 0 || ordinal >= 64) return fail("vktb_set_device: ordinal out of range");
    t_device = ordinal;
    return currentDevice() ? 0 : -1;
}

int vktb_get_device(int* ordinal) { if (!ordinal) return fail("null argument"); *ordinal = t_device; return 0; }

uint64_t vktb_launch_count(void) { return g_launchCount.load(); }

void* vktb_get_stream(void)
{
    DeviceCtx* d = currentDevice();
    return d ? (void*)currentStream(d) : nullptr;
}

int vktb_set_stream(void* s)
{
    VKTB_API_RANGE();
    if (s) { t_streamOverride = (cudaStream_t)s; t_hasOverride = true; }
    else { t_streamOverride = nullptr; t_hasOverride = false; }
    return 0;
}

int vktb_sync(void)
{
    VKTB_API_RANGE();
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    cudaError_t e = cudaStreamSynchronize(d->copy);
    if (e == cudaSuccess) e = cudaStreamSynchronize(currentStream(d));
    return e == cudaSuccess ? 0 : fail("vktb_sync", e);
}

} // extern "C"

// ================================================================ C-ABI: memory ==========================
namespace {

// ---- staging copies on several cores ---------------------------------------------------------------------------------
// Pageable host memory reaches the device through the pinned staging ring; ONE core moves about 10 GB/s, a fifth of
// what the link takes, so a 16 MiB chunk is cut into parts copied by a small pool of persistent threads (the caller
// copies the first part itself).  VKTB_COPY_THREADS overrides the pool size (1 = the old single-threaded copy).
class CopyPool
{
public:
    // never destroyed: the detached workers wait on its condition variable for the life of the process, and
    // destroying a condition variable with waiters blocks at exit (glibc)
    static CopyPool& get() { static CopyPool* p = new CopyPool; return *p; }

    void copy(void* dst, const void* src, size_t n)
    {
        const int parts = (int)std::min<size_t>((size_t)workers_.size() + 1, n / ((size_t)1 << 20));
        if (parts <= 1) { memcpy(dst, src, n); return; }
        std::lock_guard<std::mutex> one(callMutex_);            // one split copy at a time
        const size_t per = ((n / parts) + 63) & ~(size_t)63;
        {
            std::lock_guard<std::mutex> l(m_);
            dst_ = (uint8_t*)dst; src_ = (const uint8_t*)src; n_ = n; per_ = per;
            next_ = 1; pending_ = parts - 1; parts_ = parts;
            ++generation_;
        }
        cv_.notify_all();
        memcpy(dst, src, std::min(per, n));                      // part 0 on the calling thread
        std::unique_lock<std::mutex> l(m_);
        done_.wait(l, [&] { return pending_ == 0; });
    }

private:
    CopyPool()
    {
        const char* e = getenv("VKTB_COPY_THREADS");
        int want = e ? atoi(e) : 0;
        if (want <= 0) {
            const unsigned hw = std::thread::hardware_concurrency();
            want = (int)std::min(8u, std::max(1u, hw / 4));
        }
        for (int i = 1; i < want; ++i) workers_.emplace_back([this] { run(); });
        for (auto& t : workers_) t.detach();                     // process-lifetime pool
    }

    void run()
    {
        unsigned long long seen = 0;
        for (;;) {
            int part;
            {
                std::unique_lock<std::mutex> l(m_);
                cv_.wait(l, [&] { return generation_ != seen && next_ < parts_; });
                part = next_++;
                if (next_ >= parts_) seen = generation_;
            }
            const size_t off = (size_t)part * per_;
            if (off < n_) memcpy(dst_ + off, src_ + off, std::min(per_, n_ - off));
            std::lock_guard<std::mutex> l(m_);
            if (--pending_ == 0) done_.notify_all();
        }
    }

    std::vector<std::thread> workers_;
    std::mutex m_, callMutex_;
    std::condition_variable cv_, done_;
    uint8_t* dst_ = nullptr; const uint8_t* src_ = nullptr;
    size_t n_ = 0, per_ = 0;
    int next_ = 0, parts_ = 0, pending_ = 0;
    unsigned long long generation_ = 0;
};

// pageable host -> device through the staging ring; returns once the source has been fully read
int stagedH2D(DeviceCtx* d, uint8_t* dst, const uint8_t* src, size_t bytes)
{
    std::lock_guard<std::mutex> lock(d->stageMutex);
    if (ensureStage(d)) return -1;
    cudaStream_t cs = d->copy;
    // device buffer may have been produced by stream-ordered work on the compute stream (its allocation)
    cudaEvent_t ready;
    cudaEventCreateWithFlags(&ready, cudaEventDisableTiming);
    cudaEventRecord(ready, currentStream(d));
    cudaStreamWaitEvent(cs, ready, 0);
    cudaEventDestroy(ready);
    int slot = 0;
    for (size_t off = 0; off < bytes; off += kStageBytes, slot = (slot + 1) % kStageSlots) {
        const size_t n = std::min(kStageBytes, bytes - off);
        if (d->stageInFlight[slot]) { cudaEventSynchronize(d->stageDone[slot]); d->stageInFlight[slot] = false; }
        CopyPool::get().copy(d->stage[slot], src + off, n);
        cudaError_t e = cudaMemcpyAsync(dst + off, d->stage[slot], n, cudaMemcpyHostToDevice, cs);
        if (e != cudaSuccess) return fail("cudaMemcpyAsync(H2D staged)", e);
        cudaEventRecord(d->stageDone[slot], cs);
        d->stageInFlight[slot] = true;
    }
    // later kernels on the compute stream must see the data: order them after the DMA tail
    cudaEvent_t tail;
    cudaEventCreateWithFlags(&tail, cudaEventDisableTiming);
    cudaEventRecord(tail, cs);
    cudaStreamWaitEvent(currentStream(d), tail, 0);
    cudaEventDestroy(tail);
    return 0;
}

// device -> pageable host through the staging ring; complete on return
int stagedD2H(DeviceCtx* d, uint8_t* dst, const uint8_t* src, size_t bytes)
{
    std::lock_guard<std::mutex> lock(d->stageMutex);
    if (ensureStage(d)) return -1;
    cudaStream_t s = currentStream(d);
    for (int i = 0; i < kStageSlots; ++i)
        if (d->stageInFlight[i]) { cudaEventSynchronize(d->stageDone[i]); d->stageInFlight[i] = false; }
    const size_t nChunks = (bytes + kStageBytes - 1) / kStageBytes;
    size_t issued = 0, drained = 0;
    while (drained < nChunks) {
        while (issued < nChunks && issued < drained + kStageSlots) {
            const size_t off = issued * kStageBytes, n = std::min(kStageBytes, bytes - off);
            const int slot = (int)(issued % kStageSlots);
            cudaError_t e = cudaMemcpyAsync(d->stage[slot], src + off, n, cudaMemcpyDeviceToHost, s);
            if (e != cudaSuccess) return fail("cudaMemcpyAsync(D2H staged)", e);
            cudaEventRecord(d->stageDone[slot], s);
            ++issued;
        }
        const size_t off = drained * kStageBytes, n = std::min(kStageBytes, bytes - off);
        const int slot = (int)(drained % kStageSlots);
        cudaError_t e = cudaEventSynchronize(d->stageDone[slot]);
        if (e != cudaSuccess) return fail("D2H staged wait", e);
        CopyPool::get().copy(dst + off, d->stage[slot], n);
        ++drained;
    }
    return 0;
}

// producer -> device: `read` fills one staging slot while the DMA of the previous slots runs.  *sum receives the
// sum of the producer's return values (the caller compares it with `bytes`, whatever the producer's convention).
int streamIn(DeviceCtx* d, uint8_t* dst, size_t bytes, vktb_read_fn read, void* ctx, size_t* sum)
{
    std::lock_guard<std::mutex> lock(d->stageMutex);
    if (ensureStage(d)) return -1;
    cudaStream_t cs = d->copy;
    cudaEvent_t ready;
    cudaEventCreateWithFlags(&ready, cudaEventDisableTiming);
    cudaEventRecord(ready, currentStream(d));
    cudaStreamWaitEvent(cs, ready, 0);
    cudaEventDestroy(ready);
    size_t total = 0;
    int slot = 0;
    for (size_t off = 0; off < bytes; off += kStageBytes, slot = (slot + 1) % kStageSlots) {
        const size_t n = std::min(kStageBytes, bytes - off);
        if (d->stageInFlight[slot]) { cudaEventSynchronize(d->stageDone[slot]); d->stageInFlight[slot] = false; }
        total += read(ctx, d->stage[slot], n);
        cudaError_t e = cudaMemcpyAsync(dst + off, d->stage[slot], n, cudaMemcpyHostToDevice, cs);
        if (e != cudaSuccess) return fail("cudaMemcpyAsync(H2D streamed)", e);
        cudaEventRecord(d->stageDone[slot], cs);
        d->stageInFlight[slot] = true;
    }
    cudaEvent_t tail;
    cudaEventCreateWithFlags(&tail, cudaEventDisableTiming);
    cudaEventRecord(tail, cs);
    cudaStreamWaitEvent(currentStream(d), tail, 0);
    cudaEventDestroy(tail);
    if (sum) *sum = total;
    return 0;
}

// device -> consumer: up to kStageSlots D2H copies in flight while `write` drains the oldest slot
int streamOut(DeviceCtx* d, const uint8_t* src, size_t bytes, vktb_write_fn write, void* ctx, size_t* sum)
{
    std::lock_guard<std::mutex> lock(d->stageMutex);
    if (ensureStage(d)) return -1;
    cudaStream_t s = currentStream(d);
    for (int i = 0; i < kStageSlots; ++i)
        if (d->stageInFlight[i]) { cudaEventSynchronize(d->stageDone[i]); d->stageInFlight[i] = false; }
    const size_t nChunks = (bytes + kStageBytes - 1) / kStageBytes;
    size_t issued = 0, drained = 0, total = 0;
    while (drained < nChunks) {
        while (issued < nChunks && issued < drained + kStageSlots) {
            const size_t off = issued * kStageBytes, n = std::min(kStageBytes, bytes - off);
            const int slot = (int)(issued % kStageSlots);
            cudaError_t e = cudaMemcpyAsync(d->stage[slot], src + off, n, cudaMemcpyDeviceToHost, s);
            if (e != cudaSuccess) return fail("cudaMemcpyAsync(D2H streamed)", e);
            cudaEventRecord(d->stageDone[slot], s);
            ++issued;
        }
        const size_t off = drained * kStageBytes, n = std::min(kStageBytes, bytes - off);
        const int slot = (int)(drained % kStageSlots);
        cudaError_t e = cudaEventSynchronize(d->stageDone[slot]);
        if (e != cudaSuccess) return fail("D2H streamed wait", e);
        total += write(ctx, d->stage[slot], n);
        ++drained;
    }
    if (sum) *sum = total;
    return 0;
}

bool isPinnedHost(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

} // namespace

extern "C" {

int vktb_alloc(void** ptr, size_t bytes)
{
    VKTB_API_RANGE();
    if (!ptr) return fail("vktb_alloc: null argument");
    *ptr = nullptr;
    if (injectAllocFailure()) return fail("vktb_alloc: out of memory (injected by VKTB_FAIL_ALLOC_AT)");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    if (bytes == 0) bytes = 1;   // like malloc(0) in the reference's Allocate_serial: a unique, freeable pointer
    // with a device group (vktb_set_devices) large volumes are striped over its GPUs: one pointer, all devices
    if (groupSize() > 1 && bytes >= shardMinBytes()) return stripedAlloc(ptr, bytes);
    cudaError_t e = cudaMallocAsync(ptr, bytes, currentStream(d));
    if (e != cudaSuccess) { *ptr = nullptr; return fail("vktb_alloc", e); }
    std::lock_guard<std::mutex> lock(d->allocMutex);
    d->deviceAllocs[*ptr] = bytes;
    d->deviceBytes += bytes;
    return 0;
}

int vktb_free(void* ptr)
{
    VKTB_API_RANGE();
    if (!ptr) return 0;
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    {
        int rc = 0;
        if (stripedFree(ptr, &rc)) return rc;
    }
    {
        std::lock_guard<std::mutex> lock(d->allocMutex);
        auto it = d->deviceAllocs.find(ptr);
        if (it == d->deviceAllocs.end()) return fail("vktb_free: pointer was not returned by vktb_alloc on this device");
        d->deviceBytes -= it->second;
        d->deviceAllocs.erase(it);
    }
    // staged H2D traffic into this buffer runs on the copy stream: do not recycle the memory under it
    cudaEvent_t ev;
    cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    cudaEventRecord(ev, d->copy);
    cudaStreamWaitEvent(currentStream(d), ev, 0);
    cudaEventDestroy(ev);
    cudaError_t e = cudaFreeAsync(ptr, currentStream(d));
    return e == cudaSuccess ? 0 : fail("vktb_free", e);
}

int vktb_memcpy(void* dst, const void* src, size_t bytes, int kind)
{
    VKTB_API_RANGE();
    if (bytes == 0) return 0;
    if (!dst || !src) return fail("vktb_memcpy: null pointer");
    if (kind == VKTB_H2H) { memmove(dst, src, bytes); return 0; }
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    {
        // a striped volume: every device's copy engine moves its own byte range (migrate() of a sharded volume)
        int rc = 0;
        if (stripedMemcpy(dst, src, bytes, kind, &rc)) return rc;
    }
    cudaStream_t s = currentStream(d);
    cudaError_t e;
    switch (kind) {
    case VKTB_H2D:
        if (pinnedOwns(src) || isPinnedHost(src)) {
            // pinned memory (the backend's pool or foreign): one DMA straight out of the source.  The source must be
            // consumed on return -- the caller may overwrite it, as after the reference's cudaMemcpy (Memory.cpp:70) --
            // so the call waits for this DMA.  (Round 1 returned early for pool blocks and only deferred their reuse;
            // a caller refilling the buffer raced the DMA.  Nothing is lost: the next operand's upload shares the
            // same link and the kernel needs them all.)
            e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s);
            return e == cudaSuccess ? 0 : fail("cudaMemcpyAsync(H2D pinned)", e);
        }
        return stagedH2D(d, (uint8_t*)dst, (const uint8_t*)src, bytes);
    case VKTB_D2H:
        if (isPinnedHost(dst)) {
            e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s);
            return e == cudaSuccess ? 0 : fail("cudaMemcpyAsync(D2H)", e);
        }
        return stagedD2H(d, (uint8_t*)dst, (const uint8_t*)src, bytes);
    case VKTB_D2D:
        e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, s);
        return e == cudaSuccess ? 0 : fail("cudaMemcpyAsync(D2D)", e);
    }
    return fail("vktb_memcpy: bad copy kind");
}

int vktb_stream_in(void* devDst, size_t bytes, vktb_read_fn read, void* ctx, size_t* sum)
{
    VKTB_API_RANGE();
    if (sum) *sum = 0;
    if (bytes == 0) return 0;
    if (!devDst || !read) return fail("vktb_stream_in: null argument");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    return streamIn(d, (uint8_t*)devDst, bytes, read, ctx, sum);
}

int vktb_stream_out(const void* devSrc, size_t bytes, vktb_write_fn write, void* ctx, size_t* sum)
{
    VKTB_API_RANGE();
    if (sum) *sum = 0;
    if (bytes == 0) return 0;
    if (!devSrc || !write) return fail("vktb_stream_out: null argument");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    return streamOut(d, (const uint8_t*)devSrc, bytes, write, ctx, sum);
}

int vktb_memset_range(void* dst, const void* pattern, size_t dstBytes, size_t patBytes)
{
    VKTB_API_RANGE();
    if (dstBytes == 0) return 0;
    if (!dst || !pattern || patBytes == 0) return fail("vktb_memset_range: bad arguments");
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    cudaStream_t s = currentStream(d);
    void* pat = nullptr;
    cudaError_t e = cudaMallocAsync(&pat, patBytes, s);
    if (e != cudaSuccess) return fail("vktb_memset_range", e);
    // pageable source: the runtime stages it before returning, so `pattern` may die right after this call
    e = cudaMemcpyAsync(pat, pattern, patBytes, cudaMemcpyHostToDevice, s);
    LaunchCtx c{ s, d->smCount };
    if (e == cudaSuccess) e = launchMemsetRange(c, dst, pat, dstBytes, patBytes);
    cudaFreeAsync(pat, s);
    return e == cudaSuccess ? 0 : fail("vktb_memset_range", e);
}

int vktb_host_alloc(void** ptr, size_t bytes)
{
    VKTB_API_RANGE();
    if (!ptr) return fail("vktb_host_alloc: null argument");
    if (!currentDevice()) return -1;
    return pinnedAlloc(ptr, bytes);
}

int vktb_host_free(void* ptr) { return pinnedFree(ptr); }

int vktb_host_owns(const void* ptr) { return ptr && pinnedOwns(ptr) ? 1 : 0; }

int vktb_mem_stats(size_t* devInUse, size_t* devReserved, size_t* pinInUse, size_t* pinReserved)
{
    VKTB_API_RANGE();
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    if (devInUse) { std::lock_guard<std::mutex> lock(d->allocMutex); *devInUse = d->deviceBytes; }
    if (devReserved) {
        cudaMemPool_t pool; uint64_t v = 0;
        if (cudaDeviceGetDefaultMemPool(&pool, d->ordinal) == cudaSuccess)
            cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &v);
        *devReserved = (size_t)v;
    }
    pinnedStats(pinInUse, pinReserved);
    return 0;
}

int vktb_trim_pools(void)
{
    VKTB_API_RANGE();
    DeviceCtx* d = currentDevice();
    if (!d) return -1;
    cudaStreamSynchronize(d->copy);
    cudaStreamSynchronize(currentStream(d));
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, d->ordinal) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
    pinnedTrim();
    return 0;
}

} // extern "C"
