// ExecutionPolicy + ManagedResource registry (reference src/vkt/ExecutionPolicy.cpp:17-36,
// src/vkt/ManagedResource.cpp:16-41).
#include <cstdlib>
#include <mutex>
#include <unordered_map>

#include <vkt/ExecutionPolicy.hpp>
#include <vkt/ManagedResource.hpp>

namespace vkt
{
    // per thread, default {CPU, Serial, CUDA, False}; not inherited by child threads (a new thread starts from the
    // default, exactly like a thread id the reference's map has never seen)
    static thread_local ExecutionPolicy t_policy;

    void SetThreadExecutionPolicy(ExecutionPolicy policy) { t_policy = policy; }
    ExecutionPolicy GetThreadExecutionPolicy() { return t_policy; }

    // host streaming (volkit-b200 extension): -1 = not set on this thread, take the environment's default
    static thread_local int t_streaming = -1;

    void SetHostStreaming(bool enabled) { t_streaming = enabled ? 1 : 0; }
    bool GetHostStreaming()
    {
        if (t_streaming < 0) {
            char const* e = std::getenv("VKT_B200_STREAMING");
            t_streaming = (e && e[0] == '1') ? 1 : 0;
        }
        return t_streaming == 1;
    }

    static std::mutex g_resourceMutex;
    static std::unordered_map<ResourceHandle, ManagedResource> g_resources;
    static ResourceHandle g_nextHandle = 0;

    ResourceHandle RegisterManagedResource(ManagedResource resource)
    {
        std::lock_guard<std::mutex> lock(g_resourceMutex);
        ResourceHandle h = g_nextHandle++;
        g_resources[h] = resource;
        return h;
    }

    void UnregisterManagedResource(ResourceHandle handle)
    {
        std::lock_guard<std::mutex> lock(g_resourceMutex);
        g_resources.erase(handle);
    }

    ManagedResource GetManagedResource(ResourceHandle handle)
    {
        std::lock_guard<std::mutex> lock(g_resourceMutex);
        auto it = g_resources.find(handle);
        return it == g_resources.end() ? nullptr : it->second;
    }
} // vkt
