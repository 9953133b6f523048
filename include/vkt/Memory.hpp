// Policy-dispatched memory primitives (reference include/cpp/vkt/Memory.hpp:15-30, src/vkt/Memory.cpp:30-82).
// Under Device::GPU they bottom out in the C-ABI's device-buffer manager (vktb_alloc / vktb_free / vktb_memcpy /
// vktb_memset_range); under Device::CPU Allocate hands out pinned pool memory when a B200 is usable (so that a later
// migrate() is a single DMA) and plain malloc otherwise.
#pragma once

#include <cstddef>

#include "common.hpp"

namespace vkt
{
    enum class CopyKind
    {
        HostToHost,
        HostToDevice,
        DeviceToHost,
        DeviceToDevice,
    };

    VKTAPI void Allocate(void** ptr, std::size_t size);
    VKTAPI void Free(void* ptr);
    // blocking contract of migrate(): on return the source may be freed (see include/vktb/backend.h)
    VKTAPI void Memcpy(void* dst, void const* src, std::size_t size, CopyKind ck);
    VKTAPI void MemsetRange(void* dst, void const* src, std::size_t dstSize, std::size_t srcSize);

    // volkit-b200 extension: the GPUs the Device::GPU policy uses (one process, several devices).  Volumes of 1 GiB and
    // more (VKTB_SHARD_MIN_MB) are then striped over them at Allocate time -- one data pointer spans the GPUs -- and
    // Fill / Copy / the arithmetic family / Transform / Resample run on all of them; see vktb_set_devices.
    VKTAPI Error SetDevices(int const* deviceIds, int numDevices);
} // vkt
