// InputStream (reference include/cpp/vkt/InputStream.hpp:14-41, src/vkt/InputStream.cpp:22-96): reads a structured
// volume, or rows of a range of it, from a DataSource.  read() compares the source's return value with the volume's
// size in bytes; readRange() writes each row of [first, last) at the START of row (y, z) -- the reference's offset has
// no firstX term (InputStream.cpp:62) -- and is kept that way.
// GPU policy: the reference passes volume.getData(), a device pointer, to DataSource::read.  Here the source fills
// the pinned staging ring chunk by chunk while the previous chunks' DMAs run (vktb_stream_in): no host copy of the
// volume is made and nothing is uploaded first (the stale content is discarded, not migrated).
#pragma once

#include <cstddef>
#include <cstdint>

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    class VKTAPI InputStream
    {
    public:
        InputStream(DataSource& source);

        Error read(StructuredVolume& volume);
        Error readRange(StructuredVolume& dst, int32_t firstX, int32_t firstY, int32_t firstZ,
                        int32_t lastX, int32_t lastY, int32_t lastZ);
        Error readRange(StructuredVolume& dst, Vec3i first, Vec3i last);
        Error seek(std::size_t pos);

    private:
        DataSource& dataSource_;
    };
} // vkt
