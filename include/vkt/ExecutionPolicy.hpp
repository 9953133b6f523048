// Per-thread execution policy (reference include/cpp/vkt/ExecutionPolicy.hpp:47-102).
//
// The API is deferred: changing the policy moves nothing; the next algorithm or accessor that touches a volume
// migrates it to the device the policy names (ManagedBuffer::migrate).  Policies are per thread and are not
// inherited by child threads.  Here the policy is a thread_local (the reference keeps an unsynchronised global
// map keyed by thread id, src/vkt/ExecutionPolicy.cpp:17-35, and looks it up on every voxel access).
#pragma once

#include "common.hpp"

namespace vkt
{
    struct ExecutionPolicy
    {
        enum class Device    { CPU, GPU, Unspecified };
        enum class HostAPI   { Serial, OpenMP, Auto };
        enum class DeviceAPI { CUDA, Auto };

        Device    device = Device::CPU;
        HostAPI   hostApi = HostAPI::Serial;
        DeviceAPI deviceApi = DeviceAPI::CUDA;
        Bool      printPerformance = False;   // log one timing line per algorithm call
    };

    VKTAPI void SetThreadExecutionPolicy(ExecutionPolicy policy);
    VKTAPI ExecutionPolicy GetThreadExecutionPolicy();

    // volkit-b200 extension (per thread, default off or VKT_B200_STREAMING=1): under Device::GPU, algorithms whose
    // volumes all still live on the host do not migrate() them; they stream z-slabs through the B200 (upload, kernel
    // and download overlapped, vktb_hosted_* in vktb/backend.h) and leave the volumes -- and the result -- on the
    // host.  Same voxels as the migrating call; volumes larger than device memory work.  Covers the arithmetic
    // family, unary Transform and Resample; other calls, and calls with a device-resident volume, migrate as usual.
    VKTAPI void SetHostStreaming(bool enabled);
    VKTAPI bool GetHostStreaming();
} // vkt
