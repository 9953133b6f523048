// Handle registry for managed buffers (reference include/cpp/vkt/ManagedResource.hpp:10-20,
// src/vkt/ManagedResource.cpp:16-41; mutex-guarded here).
#pragma once

#include <cstdint>

#include "common.hpp"

namespace vkt
{
    typedef void* ManagedResource;
    typedef uint32_t ResourceHandle;

    VKTAPI ResourceHandle RegisterManagedResource(ManagedResource resource);
    VKTAPI void UnregisterManagedResource(ResourceHandle handle);
    VKTAPI ManagedResource GetManagedResource(ResourceHandle handle);
} // vkt
