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
} // vkt
