// Several GPUs of one box behind the SAME C-ABI, in one process (SURVEY.md 8(b) vktb_set_devices, 8(e)).
//
// vktb_set_devices(ids, n) names the device group.  From then on vktb_alloc places an allocation of at least
// shardMinBytes() as ONE virtual address range whose physical memory is striped over the group (CUDA virtual memory
// management: cuMemAddressReserve + one cuMemCreate / cuMemMap per device + cuMemSetAccess for every device), so the
// single `T* data_` of the reference's ManagedBuffer (include/cpp/vkt/ManagedBuffer.hpp:168-198) spans the GPUs and
// every device can address every byte over NVLink.  z is the slowest axis, so a contiguous byte range IS a z-slab.
//   * algorithms whose destination is a striped volume fan out: device g runs the ordinary launcher on the dst planes
//     that start in its stripe, on its own stream, forked from and joined to the caller's stream by events; sources
//     are read wherever they live -- a Resample tap or a *Range plane that belongs to the neighbour is a plain peer
//     load inside the kernel (the halo exchange is fused away, nothing is copied or planned);
//   * vktb_memcpy of a striped range drives every device's copy engine on its own byte range concurrently: this is
//     migrate() of a sharded volume (BASELINE config 5's round trip).
// The driver entry points come from cudaGetDriverEntryPoint: libcuda is still not linked.
#pragma once

#include <cstddef>
#include <cstdint>
#include <functional>
#include <vector>

#include <cuda_runtime.h>

#include "context.hpp"
#include "../kernels/launch.hpp"

namespace vktb {

struct Stripe { int slot; int device; uint8_t* begin; uint8_t* end; };

struct StripedAlloc
{
    uint8_t* base = nullptr;
    size_t   bytes = 0;        // requested
    size_t   reserved = 0;     // VA size (granularity multiple)
    std::vector<Stripe> stripes;
    std::vector<unsigned long long> handles;
};

int  groupSize();                                   // 1 = single device (the default)
int  groupDevice(int slot);
size_t shardMinBytes();

// striped allocation (nullptr: not striped / not found)
int  stripedAlloc(void** ptr, size_t bytes);
bool stripedFree(void* ptr, int* rc);               // true if ptr was a striped allocation
const StripedAlloc* findStriped(const void* p);    // the allocation that contains p

// copies between host memory and a striped range (and between device ranges), stripe by stripe, all copy engines at
// once; complete on return.  Returns false when neither pointer lies in a striped allocation (caller proceeds as usual).
bool stripedMemcpy(void* dst, const void* src, size_t bytes, int kind, int* rc);

// Fork / join: runs body(ctx, z0, z1) for every stripe of `dstAlloc` that holds dst planes of [zBegin, zEnd), where
// plane z of the dst volume starts at dstData + z * planeBytes; ctx is that device's LaunchCtx (its stream is ordered
// after everything queued on the caller's stream so far, and the caller's stream after it).
cudaError_t forEachStripe(const StripedAlloc* dstAlloc, const uint8_t* dstData, size_t planeBytes, long long zBegin,
                          long long zEnd, const std::function<cudaError_t(const LaunchCtx&, int, int)>& body);

} // namespace vktb
