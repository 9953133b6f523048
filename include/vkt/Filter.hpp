// ApplyFilter (reference include/cpp/vkt/Filter.hpp:19-48, src/vkt/Filter.cpp:25-66, Filter_serial.hpp:15-64): dest
// receives the convolution of source with a small dense filter; voxels on the volume border receive 0 and the address
// mode is ignored, as in the reference.  Filter dims must be 1..3 per axis: the reference reads outside the volume
// with larger filters, here such a call returns InvalidValue.
//
// Filter keeps its weights in host memory under every policy (they travel to the kernel as launch arguments).  The
// reference derives Filter from ManagedBuffer and its constructor memcpy()s into the buffer, which is device memory
// under the GPU policy ("this won't work if current execution policy is GPU", Filter.cpp:29).
#pragma once

#include <cstdint>
#include <vector>

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    class VKTAPI Filter
    {
    public:
        Filter(float* data, Vec3i dims);

        float* getData();
        Vec3i getDims() const;

    private:
        std::vector<float> data_;
        Vec3i dims_;
    };

    enum class AddressMode
    {
        Wrap,
        Mirror,
        Clamp,
        Border,
    };

    VKTAPI Error ApplyFilter(StructuredVolume& dest, StructuredVolume& source, Filter& filter, AddressMode am);
} // vkt
