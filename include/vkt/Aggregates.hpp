// ComputeAggregates / ComputeAggregatesRange (reference include/cpp/vkt/Aggregates.hpp:14-43, src/vkt/Aggregates.cpp:23-66,
// Aggregates_serial.hpp:20-83).  CPU policy: the reference's sequential binary32 accumulation, statement for
// statement.  GPU policy: min, max, argmin, argmax exact; sum, mean, var, stddev, prod per the contract of
// vktb_aggregates_range (include/vktb/backend.h) -- a sequential binary32 accumulation has no parallel twin.
#pragma once

#include <cstdint>

#include "common.hpp"
#include "linalg.hpp"

namespace vkt
{
    struct Aggregates
    {
        float min;
        float max;
        float mean;
        float stddev;
        float var;
        float sum;
        float prod;
        Vec3i argmin;
        Vec3i argmax;
    };

    VKTAPI Error ComputeAggregates(StructuredVolume& volume, Aggregates& aggregates);
    VKTAPI Error ComputeAggregatesRange(StructuredVolume& volume, Aggregates& aggregates,
                                        int32_t firstX, int32_t firstY, int32_t firstZ,
                                        int32_t lastX, int32_t lastY, int32_t lastZ);
    VKTAPI Error ComputeAggregatesRange(StructuredVolume& volume, Aggregates& aggregates, Vec3i first, Vec3i last);
} // vkt
