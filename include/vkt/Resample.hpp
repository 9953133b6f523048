// Resample StructuredVolume -> StructuredVolume (reference include/cpp/vkt/Resample.hpp:14-31,
// src/vkt/Resample_serial.hpp:26-71): spatial + format resample.  Equal dims: conversion only.  Nearest: tap
// (int)(d / float(D) * S).  Linear: what the reference's sampleLinear computes -- the float coordinate is truncated
// at the call, so the fraction is 0; bit-compatible with the reference CPU path.
#pragma once

#include "common.hpp"

namespace vkt
{
    enum class FilterMode
    {
        Nearest,
        Linear,
    };

    VKTAPI Error Resample(StructuredVolume& dst, StructuredVolume& src, FilterMode fm);
} // vkt
