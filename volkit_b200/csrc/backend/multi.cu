// Device group, striped (VMM) allocations, per-stripe copies and the fork / join helper: see multi.hpp.
#include "multi.hpp"

#include <cuda.h>          // driver TYPES only; the entry points come from cudaGetDriverEntryPoint (libcuda is not linked)

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>

#include "../../../include/vktb/backend.h"

namespace vktb {

namespace {

// ---- driver entry points -----------------------------------------------------------------------------------------
struct Driver
{
    CUresult (*memGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*memAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*memCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*memMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*memSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*memUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*memRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*memAddressFree)(CUdeviceptr, size_t) = nullptr;
    bool ok = false;
};

Driver& driver()
{
    static Driver d;
    static std::once_flag once;
    std::call_once(once, [] {
        auto get = [](const char* name, void** fn) {
            cudaDriverEntryPointQueryResult st;
            return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &st) == cudaSuccess && st == cudaDriverEntryPointSuccess && *fn;
        };
        d.ok = get("cuMemGetAllocationGranularity", (void**)&d.memGetAllocationGranularity)
            && get("cuMemAddressReserve", (void**)&d.memAddressReserve) && get("cuMemCreate", (void**)&d.memCreate)
            && get("cuMemMap", (void**)&d.memMap) && get("cuMemSetAccess", (void**)&d.memSetAccess)
            && get("cuMemUnmap", (void**)&d.memUnmap) && get("cuMemRelease", (void**)&d.memRelease)
            && get("cuMemAddressFree", (void**)&d.memAddressFree);
    });
    return d;
}

int failDrv(const char* what, CUresult r)
{
    setError(std::string(what) + ": CUDA driver error " + std::to_string((int)r));
    return -1;
}

// ---- the device group --------------------------------------------------------------------------------------------
struct Slot
{
    int device = 0;
    cudaStream_t compute = nullptr, copy = nullptr;   // this slot's own streams on its device
    cudaEvent_t  joined = nullptr;
    int smCount = 0;
};

std::mutex g_groupMutex;
std::vector<Slot> g_slots;                      // empty = single device
std::mutex g_allocMutex;
std::map<uintptr_t, StripedAlloc> g_striped;    // by base address

} // namespace

int groupSize() { std::lock_guard<std::mutex> l(g_groupMutex); return g_slots.empty() ? 1 : (int)g_slots.size(); }
int groupDevice(int slot) { std::lock_guard<std::mutex> l(g_groupMutex); return g_slots.empty() ? currentOrdinal() : g_slots[(size_t)slot].device; }

size_t shardMinBytes()
{
    const char* e = getenv("VKTB_SHARD_MIN_MB");
    const long mb = e ? atol(e) : 0;
    return mb > 0 ? (size_t)mb << 20 : (size_t)1 << 30;      // default: volumes of 1 GiB and more are striped (read per call)
}

int setDeviceGroup(const int* ids, int n)
{
    if (n < 1 || !ids) return fail("vktb_set_devices: need at least one device id");
    if (n > 64) return fail("vktb_set_devices: at most 64 devices");
    {
        std::lock_guard<std::mutex> l(g_allocMutex);
        if (!g_striped.empty()) return fail("vktb_set_devices: striped volumes are alive; free them before changing the group");
    }
    std::vector<Slot> slots((size_t)n);
    const int home = currentOrdinal();
    for (int i = 0; i < n; ++i) {
        DeviceCtx* d = deviceCtx(ids[i]);           // makes ids[i] current
        if (!d) return -1;
        Slot& s = slots[(size_t)i];
        s.device = ids[i]; s.smCount = d->smCount;
        cudaError_t e = cudaStreamCreateWithFlags(&s.compute, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s.copy, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s.joined, cudaEventDisableTiming);
        if (e != cudaSuccess) return fail("vktb_set_devices: stream set-up", e);
        // peer access both ways with every other device of the group: kernels read sources wherever they live
        for (int j = 0; j < n; ++j) {
            if (ids[j] == ids[i]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, ids[i], ids[j]);
            if (!can) return fail("vktb_set_devices: the devices of a group must have peer access to each other (NVLink / NVSwitch)");
            e = cudaDeviceEnablePeerAccess(ids[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail("cudaDeviceEnablePeerAccess", e);
            cudaGetLastError();
        }
    }
    (void)home;
    setCurrentOrdinal(ids[0]);                     // the group's first device is the calling thread's home
    if (!deviceCtx(ids[0])) return -1;
    std::lock_guard<std::mutex> l(g_groupMutex);
    for (Slot& s : g_slots) {                      // a previous group's streams
        cudaSetDevice(s.device);
        cudaStreamDestroy(s.compute); cudaStreamDestroy(s.copy); cudaEventDestroy(s.joined);
    }
    cudaSetDevice(ids[0]);
    if (n == 1) g_slots.clear(); else g_slots = slots;
    return 0;
}

// ---- striped allocation --------------------------------------------------------------------------------------------
int stripedAlloc(void** ptr, size_t bytes)
{
    *ptr = nullptr;
    Driver& drv = driver();
    if (!drv.ok) return fail("vktb_alloc (striped): the driver's virtual memory management entry points are unavailable");
    std::vector<Slot> slots;
    { std::lock_guard<std::mutex> l(g_groupMutex); slots = g_slots; }
    const int n = (int)slots.size();
    if (n < 2) return fail("vktb_alloc (striped): no device group");
    for (const Slot& s : slots) if (!deviceCtx(s.device)) return -1;     // primary contexts exist

    CUmemAllocationProp prop;
    std::memset(&prop, 0, sizeof prop);
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = slots[0].device;
    size_t gran = 0;
    CUresult r = drv.memGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
    if (r != CUDA_SUCCESS || gran == 0) return failDrv("cuMemGetAllocationGranularity", r);
    const size_t total = (bytes + gran - 1) / gran * gran;
    const size_t chunk = ((total / gran + (size_t)n - 1) / (size_t)n) * gran;

    CUdeviceptr va = 0;
    r = drv.memAddressReserve(&va, total, gran, 0, 0);
    if (r != CUDA_SUCCESS) return failDrv("cuMemAddressReserve", r);
    StripedAlloc a;
    a.base = reinterpret_cast<uint8_t*>(va); a.bytes = bytes; a.reserved = total;
    auto undo = [&](const char* what, CUresult rr) {
        for (size_t i = 0; i < a.handles.size(); ++i) {
            drv.memUnmap((CUdeviceptr)a.stripes[i].begin, (size_t)(a.stripes[i].end - a.stripes[i].begin));
            drv.memRelease((CUmemGenericAllocationHandle)a.handles[i]);
        }
        drv.memAddressFree(va, total);
        return failDrv(what, rr);
    };
    for (int g = 0; g < n; ++g) {
        const size_t off = std::min(total, (size_t)g * chunk), len = std::min(chunk, total - off);
        if (len == 0) break;
        prop.location.id = slots[(size_t)g].device;
        CUmemGenericAllocationHandle h = 0;
        r = drv.memCreate(&h, len, &prop, 0);
        if (r != CUDA_SUCCESS) return undo("cuMemCreate (out of device memory on one GPU of the group?)", r);
        r = drv.memMap(va + off, len, 0, h, 0);
        if (r != CUDA_SUCCESS) { drv.memRelease(h); return undo("cuMemMap", r); }
        a.handles.push_back((unsigned long long)h);
        a.stripes.push_back({ g, slots[(size_t)g].device, a.base + off, a.base + off + len });
    }
    std::vector<CUmemAccessDesc> acc;
    for (const Slot& s : slots) {
        bool seen = false;
        for (const CUmemAccessDesc& d : acc) seen = seen || d.location.id == s.device;
        if (seen) continue;
        CUmemAccessDesc d;
        std::memset(&d, 0, sizeof d);
        d.location.type = CU_MEM_LOCATION_TYPE_DEVICE; d.location.id = s.device; d.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        acc.push_back(d);
    }
    r = drv.memSetAccess(va, total, acc.data(), acc.size());
    if (r != CUDA_SUCCESS) return undo("cuMemSetAccess", r);
    cudaSetDevice(slots[0].device);
    std::lock_guard<std::mutex> l(g_allocMutex);
    g_striped[(uintptr_t)va] = a;
    *ptr = a.base;
    return 0;
}

const StripedAlloc* findStriped(const void* p)
{
    std::lock_guard<std::mutex> l(g_allocMutex);
    if (g_striped.empty()) return nullptr;
    auto it = g_striped.upper_bound((uintptr_t)p);
    if (it == g_striped.begin()) return nullptr;
    --it;
    return (uintptr_t)p < it->first + it->second.reserved ? &it->second : nullptr;
}

bool stripedFree(void* ptr, int* rc)
{
    StripedAlloc a;
    {
        std::lock_guard<std::mutex> l(g_allocMutex);
        auto it = g_striped.find((uintptr_t)ptr);
        if (it == g_striped.end()) return false;
        a = it->second;
        g_striped.erase(it);
    }
    // everything queued on any stream of the group may still use the range
    std::vector<Slot> slots;
    { std::lock_guard<std::mutex> l(g_groupMutex); slots = g_slots; }
    for (const Slot& s : slots) { cudaSetDevice(s.device); cudaStreamSynchronize(s.compute); cudaStreamSynchronize(s.copy); }
    DeviceCtx* home = currentDevice();
    if (home) { cudaStreamSynchronize(currentStream(home)); cudaStreamSynchronize(home->copy); }
    Driver& drv = driver();
    *rc = 0;
    for (size_t i = 0; i < a.handles.size(); ++i) {
        if (drv.memUnmap((CUdeviceptr)a.stripes[i].begin, (size_t)(a.stripes[i].end - a.stripes[i].begin)) != CUDA_SUCCESS) *rc = -1;
        if (drv.memRelease((CUmemGenericAllocationHandle)a.handles[i]) != CUDA_SUCCESS) *rc = -1;
    }
    if (drv.memAddressFree((CUdeviceptr)a.base, a.reserved) != CUDA_SUCCESS) *rc = -1;
    if (*rc) fail("vktb_free (striped): unmap / release failed");
    return true;
}

// ---- copies ----------------------------------------------------------------------------------------------------------
bool stripedMemcpy(void* dst, const void* src, size_t bytes, int kind, int* rc)
{
    const StripedAlloc* da = (kind == VKTB_H2D || kind == VKTB_D2D) ? findStriped(dst) : nullptr;
    const StripedAlloc* sa = (kind == VKTB_D2H || kind == VKTB_D2D) ? findStriped(src) : nullptr;
    if (!da && !sa) return false;
    *rc = 0;
    std::vector<Slot> slots;
    { std::lock_guard<std::mutex> l(g_groupMutex); slots = g_slots; }
    DeviceCtx* home = currentDevice();
    if (!home || slots.empty()) { *rc = fail("vktb_memcpy (striped): no device group"); return true; }
    // what the caller queued so far (kernels producing / consuming the range) comes first
    cudaStream_t primary = currentStream(home);
    cudaEvent_t fork;
    cudaEventCreateWithFlags(&fork, cudaEventDisableTiming);
    cudaEventRecord(fork, primary);
    // pieces follow the stripes of the striped side (dst for H2D / D2D, else src)
    const StripedAlloc* a = da ? da : sa;
    const uint8_t* ap = (const uint8_t*)(da ? dst : src);
    cudaError_t err = cudaSuccess;
    std::vector<int> used;
    for (const Stripe& st : a->stripes) {
        const uint8_t* b = std::max(ap, (const uint8_t*)st.begin);
        const uint8_t* e = std::min(ap + bytes, (const uint8_t*)st.end);
        if (b >= e) continue;
        const size_t off = (size_t)(b - ap), len = (size_t)(e - b);
        const Slot& s = slots[(size_t)st.slot];
        cudaSetDevice(s.device);
        cudaStreamWaitEvent(s.copy, fork, 0);
        const cudaMemcpyKind k = kind == VKTB_H2D ? cudaMemcpyHostToDevice : kind == VKTB_D2H ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
        cudaError_t e2 = cudaMemcpyAsync((uint8_t*)dst + off, (const uint8_t*)src + off, len, k, s.copy);
        if (e2 != cudaSuccess && err == cudaSuccess) err = e2;
        used.push_back(st.slot);
    }
    for (int slot : used) {
        const Slot& s = slots[(size_t)slot];
        cudaSetDevice(s.device);
        cudaError_t e2 = cudaStreamSynchronize(s.copy);
        if (e2 != cudaSuccess && err == cudaSuccess) err = e2;
    }
    cudaSetDevice(home->ordinal);
    cudaEventDestroy(fork);
    if (err != cudaSuccess) *rc = fail("vktb_memcpy (striped)", err);
    return true;
}

// ---- fork / join ----------------------------------------------------------------------------------------------------
cudaError_t forEachStripe(const StripedAlloc* a, const uint8_t* dstData, size_t planeBytes, long long zBegin, long long zEnd,
                          const std::function<cudaError_t(const LaunchCtx&, int, int)>& body)
{
    std::vector<Slot> slots;
    { std::lock_guard<std::mutex> l(g_groupMutex); slots = g_slots; }
    DeviceCtx* home = currentDevice();
    if (!home || slots.empty() || planeBytes == 0) return cudaErrorInvalidValue;
    cudaStream_t primary = currentStream(home);
    cudaEvent_t fork;
    cudaError_t err = cudaEventCreateWithFlags(&fork, cudaEventDisableTiming);
    if (err != cudaSuccess) return err;
    cudaEventRecord(fork, primary);
    auto ceilDiv = [](long long a_, long long b_) { return a_ <= 0 ? 0 : (a_ + b_ - 1) / b_; };
    for (const Stripe& st : a->stripes) {
        // planes whose FIRST byte lies in the stripe: they tile [0, dimZ) without gaps or overlaps
        long long z0 = ceilDiv((long long)(st.begin - dstData), (long long)planeBytes);
        long long z1 = ceilDiv((long long)(st.end - dstData), (long long)planeBytes);
        z0 = std::max(z0, zBegin); z1 = std::min(z1, zEnd);
        if (z1 <= z0) continue;
        const Slot& s = slots[(size_t)st.slot];
        cudaSetDevice(s.device);
        cudaStreamWaitEvent(s.compute, fork, 0);
        LaunchCtx ctx{ s.compute, s.smCount };
        cudaError_t e = body(ctx, (int)z0, (int)z1);
        if (e != cudaSuccess && err == cudaSuccess) err = e;
        cudaEventRecord(s.joined, s.compute);
        cudaStreamWaitEvent(primary, s.joined, 0);
    }
    cudaSetDevice(home->ordinal);
    cudaEventDestroy(fork);
    return err;
}

} // namespace vktb

using namespace vktb;

extern "C" {

int vktb_set_devices(const int* ids, int n) { return setDeviceGroup(ids, n); }

int vktb_get_devices(int* ids, int capacity, int* n)
{
    VKTB_API_RANGE();
    if (!n) return fail("vktb_get_devices: null argument");
    *n = groupSize();
    for (int i = 0; ids && i < *n && i < capacity; ++i) ids[i] = groupDevice(i);
    return 0;
}

int vktb_is_striped(const void* ptr, int* stripes)
{
    VKTB_API_RANGE();
    const StripedAlloc* a = findStriped(ptr);
    if (stripes) *stripes = a ? (int)a->stripes.size() : 0;
    return a ? 1 : 0;
}

} // extern "C"
