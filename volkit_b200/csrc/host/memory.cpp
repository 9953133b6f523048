// Allocate / Free / Memcpy / MemsetRange (reference src/vkt/Memory.cpp:30-82, Memory_serial.hpp:13-37,
// Memory_cuda.hpp:16-32): the four functions ManagedBuffer::migrate() is made of.
#include <cstdlib>
#include <cstring>

#include <vkt/Memory.hpp>

#include "backend_link.hpp"

namespace vkt
{
namespace host
{
    bool deviceUsable()
    {
        // decided once; a function-local static is initialised under the compiler's guard, so threads that arrive
        // together do not race (ThreadSanitizer flagged the unguarded form: tests/test_sanitizers.py)
        static const bool usable = [] {
            int n = 0;
            return vktb_device_count(&n) == 0 && n > 0 && vktb_set_device(0) == 0;
        }();
        return usable;
    }

    Error backendError(char const* what)
    {
        std::fprintf(stderr, "vkt: %s: %s\n", what, vktb_last_error());
        return InvalidValue;
    }

    static bool pinnedHostWanted()
    {
        static const bool want = [] {
            char const* e = std::getenv("VKT_B200_PINNED");
            return !(e && e[0] == '0');
        }();
        return want && deviceUsable();
    }
} // host

    void Allocate(void** ptr, std::size_t size)
    {
        ExecutionPolicy ep = GetThreadExecutionPolicy();
        *ptr = nullptr;
        if (host::onGpu(ep)) {
            if (vktb_alloc(ptr, size) != 0) host::backendError("Allocate (GPU)");
        } else if (ep.device == ExecutionPolicy::Device::CPU) {
            // pinned pool memory makes the next migrate() one DMA; plain malloc when no B200 is usable
            if (!(host::pinnedHostWanted() && vktb_host_alloc(ptr, size) == 0))
                *ptr = std::malloc(size ? size : 1);
        }
    }

    void Free(void* ptr)
    {
        if (!ptr) return;
        ExecutionPolicy ep = GetThreadExecutionPolicy();
        if (host::onGpu(ep)) {
            if (vktb_free(ptr) != 0) host::backendError("Free (GPU)");
        } else if (ep.device == ExecutionPolicy::Device::CPU) {
            if (vktb_host_owns(ptr)) vktb_host_free(ptr);
            else std::free(ptr);
        }
    }

    void Memcpy(void* dst, void const* src, std::size_t size, CopyKind ck)
    {
        if (ck == CopyKind::HostToHost) {
            std::memcpy(dst, src, size);
            return;
        }
        if (vktb_memcpy(dst, src, size, (int)ck) != 0) host::backendError("Memcpy");
    }

    Error SetDevices(int const* deviceIds, int numDevices)
    {
        if (vktb_set_devices(deviceIds, numDevices) != 0) return host::backendError("SetDevices");
        return NoError;
    }

    void MemsetRange(void* dst, void const* src, std::size_t dstSize, std::size_t srcSize)
    {
        ExecutionPolicy ep = GetThreadExecutionPolicy();
        if (host::onGpu(ep)) {
            if (vktb_memset_range(dst, src, dstSize, srcSize) != 0) host::backendError("MemsetRange (GPU)");
        } else if (ep.device == ExecutionPolicy::Device::CPU) {
            // reference Memory_serial.hpp:27-37: whole elements only
            for (std::size_t i = 0; i + srcSize <= dstSize && srcSize; i += srcSize)
                std::memcpy((char*)dst + i, src, srcSize);
        }
    }
} // vkt
