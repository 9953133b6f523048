// Thread-safety of the host mirror (race detection): the reference keeps the per-thread policy map
// and the managed-resource registry in plain globals mutated without locks (src/vkt/ExecutionPolicy.cpp:17-35,
// src/vkt/ManagedResource.cpp:16-41).  Here every thread sets its own policy, creates, uses, copies and destroys volumes
// (each one registers / unregisters a resource) at the same time.  Built with -fsanitize=thread by
// tests/test_sanitizers.py: any report fails the test.  CPU policy only -- no GPU needed.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <vkt/Arithmetic.hpp>
#include <vkt/Copy.hpp>
#include <vkt/ExecutionPolicy.hpp>
#include <vkt/Fill.hpp>
#include <vkt/Resample.hpp>
#include <vkt/StructuredVolume.hpp>

using namespace vkt;

static std::atomic<int> g_failures{0};

static void worker(int id, int rounds)
{
    ExecutionPolicy ep = GetThreadExecutionPolicy();
    if (ep.device != ExecutionPolicy::Device::CPU) ++g_failures;      // a new thread starts with the default policy
    ep.printPerformance = false;                                      // every thread writes ITS OWN policy
    SetThreadExecutionPolicy(ep);
    for (int r = 0; r < rounds; ++r) {
        const int n = 8 + (id + r) % 5;
        StructuredVolume a(n, n, n, DataFormat::UInt16), b(n, n, n, DataFormat::UInt16), c(n, n, n, DataFormat::UInt16);
        if (Fill(a, 0.25f) != NoError || Fill(b, 0.5f) != NoError || SafeSum(c, a, b) != NoError) ++g_failures;
        StructuredVolume d(c);                                          // copy: a second registration
        StructuredVolume half(n / 2, n / 2, n / 2, DataFormat::UInt16);
        if (Resample(half, d, FilterMode::Nearest) != NoError) ++g_failures;
        const uint16_t* p = (const uint16_t*)half.getData();
        const uint16_t want = ((const uint16_t*)c.getData())[0];
        if (p[0] != want || want == 0) ++g_failures;
        // the policy set by this thread is still its own
        if (GetThreadExecutionPolicy().device != ExecutionPolicy::Device::CPU) ++g_failures;
    }
}

int main(int argc, char** argv)
{
    const int threads = argc > 1 ? std::atoi(argv[1]) : 8, rounds = argc > 2 ? std::atoi(argv[2]) : 50;
    std::vector<std::thread> pool;
    for (int i = 0; i < threads; ++i) pool.emplace_back(worker, i, rounds);
    for (auto& t : pool) t.join();
    if (g_failures.load()) { std::fprintf(stderr, "%d failures\n", g_failures.load()); return 1; }
    std::printf("threads ok\n");
    return 0;
}
