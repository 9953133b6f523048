// Histogram (reference include/cpp/vkt/Histogram.hpp:17-42, src/vkt/Histogram.cpp:23-84, Histogram_serial.hpp:18-51):
// a managed buffer of numBins 64-bit counters; ComputeHistogram zeroes it and increments bin
// (size_t)((value - lo) * (numBins / (hi - lo))) for every voxel of the range (binary32 arithmetic, value = getValue).
// Voxels whose index falls outside [0, numBins) -- undefined behaviour in the reference, which writes out of bounds --
// are skipped under both policies.  The reference has no C API for this algorithm.
#pragma once

#include <cstddef>
#include <cstdint>

#include "ManagedBuffer.hpp"
#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    class VKTAPI Histogram : public ManagedBuffer<std::size_t>
    {
    public:
        Histogram(std::size_t numBins);

        std::size_t getNumBins() const;

        // pointer to the counters ON THE CURRENT POLICY'S DEVICE (migrates first)
        std::size_t* getBinCounts();
    };

    VKTAPI Error ComputeHistogram(StructuredVolume& volume, Histogram& histogram);
    VKTAPI Error ComputeHistogramRange(StructuredVolume& volume, Histogram& histogram,
                                       int32_t firstX, int32_t firstY, int32_t firstZ,
                                       int32_t lastX, int32_t lastY, int32_t lastZ);
    VKTAPI Error ComputeHistogramRange(StructuredVolume& volume, Histogram& histogram, Vec3i first, Vec3i last);
} // vkt
