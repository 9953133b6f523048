// OutputStream (reference include/cpp/vkt/OutputStream.hpp:14-48, src/vkt/OutputStream.cpp:20-88): writes a structured
// volume, or rows of a range of it (taken from the START of each row, OutputStream.cpp:60), to a DataSource.
// GPU policy: the volume leaves the device through the pinned staging ring while the source writes the chunks that
// have already arrived (vktb_stream_out); the volume stays on the device.
#pragma once

#include <cstddef>
#include <cstdint>

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    class VKTAPI OutputStream
    {
    public:
        OutputStream(DataSource& source);

        Error write(StructuredVolume& volume);
        Error writeRange(StructuredVolume& dst, int32_t firstX, int32_t firstY, int32_t firstZ,
                         int32_t lastX, int32_t lastY, int32_t lastZ);
        Error writeRange(StructuredVolume& dst, Vec3i first, Vec3i last);
        Error seek(std::size_t pos);
        Error flush();

    private:
        DataSource& dataSource_;
    };
} // vkt
