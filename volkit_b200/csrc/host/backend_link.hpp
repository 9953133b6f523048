// How the host mirror reaches the C-ABI (include/vktb/backend.h): libvkt_b200.so links libvktb200.so directly.
// Helpers shared by the host translation units.
#pragma once

#include <cstdio>

#include <vkt/ExecutionPolicy.hpp>
#include <vkt/StructuredVolume.hpp>
#include <vktb/backend.h>

namespace vkt
{
namespace host
{
    // true once a usable B200 + driver has been seen (cached); without one the CPU policy still works and every
    // GPU-policy call fails loudly
    bool deviceUsable();

    // logs "vkt: <what>: <vktb_last_error()>" to stderr and returns InvalidValue
    Error backendError(char const* what);

    inline bool onGpu(ExecutionPolicy const& ep)
    {
        return ep.device == ExecutionPolicy::Device::GPU && ep.deviceApi == ExecutionPolicy::DeviceAPI::CUDA;
    }
    inline bool onCpuSerial(ExecutionPolicy const& ep)
    {
        return ep.device == ExecutionPolicy::Device::CPU && ep.hostApi == ExecutionPolicy::HostAPI::Serial;
    }

    // POD descriptor of a volume on the CURRENT policy's device (migrates it there first)
    inline vktb_vol describe(StructuredVolume& v)
    {
        vktb_vol d;
        d.data = v.getData();
        // a migration that could not allocate leaves the bytes on the other device: hand the backend a null volume, it
        // refuses the call and the front returns InvalidValue (no kernel ever sees a pointer of the wrong device)
        if (v.residesOnHost() == onGpu(GetThreadExecutionPolicy()))
            d.data = nullptr;
        Vec3i dims = v.getDims();
        d.dims[0] = dims.x; d.dims[1] = dims.y; d.dims[2] = dims.z;
        d.format = (int32_t)v.getDataFormat();
        Vec2f m = v.getVoxelMapping();
        d.lo = m.x; d.hi = m.y;
        return d;
    }

    // descriptor of a host-resident volume as it lies (no migration): input of the vktb_hosted_* calls
    inline vktb_vol describeInPlace(StructuredVolume& v)
    {
        vktb_vol d;
        d.data = v.dataInPlace();
        Vec3i dims = v.getDims();
        d.dims[0] = dims.x; d.dims[1] = dims.y; d.dims[2] = dims.z;
        d.format = (int32_t)v.getDataFormat();
        Vec2f m = v.getVoxelMapping();
        d.lo = m.x; d.hi = m.y;
        return d;
    }
} // host
} // vkt
