// Backend state: per-device streams, pooled device memory, pinned host pool, staging ring.
//
// This is the "managed-resource migration layer" half of the hot path: what vkt::Allocate / Free / Memcpy
// (reference src/vkt/Memory.cpp:30-82, Memory_cuda.hpp:16-24) bottom out in under the GPU policy.
//   * device memory   : CUDA stream-ordered pool (cudaMallocAsync) with an unlimited release threshold, so the
//                       allocate/free pair inside every migrate() is a pool hit, not a cudaMalloc
//   * pinned host pool: size-bucketed cudaHostAlloc blocks reused across migrations; a block released while
//                       a DMA still reads it carries the DMA's event and is waited on before reuse
//   * staging ring    : 4 x 16 MiB pinned buffers that pipeline pageable <-> device copies with the DMA
//   * streams         : one compute stream (kernels, D2D, pinned DMA) + one copy stream (staged H2D) per device
#pragma once

#include <cstddef>
#include <cstdint>
#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include <cuda_runtime.h>

namespace vktb {

constexpr size_t kStageBytes = 16u << 20;
constexpr int    kStageSlots = 4;

struct PinnedBlock
{
    void*       ptr = nullptr;
    size_t      capacity = 0;
    cudaEvent_t busy = nullptr;   // non-null: a DMA recorded on this event may still read/write the block
    int         device = 0;
};

struct DeviceCtx
{
    int          ordinal = -1;
    bool         ready = false;
    int          smCount = 0;
    cudaStream_t compute = nullptr;
    cudaStream_t copy = nullptr;
    void*        stage[kStageSlots] = {};
    cudaEvent_t  stageDone[kStageSlots] = {};
    bool         stageInFlight[kStageSlots] = {};
    std::mutex   stageMutex;       // one staged transfer at a time per device
    std::unordered_map<void*, size_t> deviceAllocs;
    size_t       deviceBytes = 0;
    std::mutex   allocMutex;
};

// Returns nullptr (and sets the thread's error string) when no usable GPU exists: there is no CPU fallback.
DeviceCtx* currentDevice();
DeviceCtx* deviceCtx(int ordinal);                    // any device; makes it current
int  currentOrdinal();
void setCurrentOrdinal(int ordinal);
cudaStream_t currentStream(DeviceCtx* d);

// NVTX range around every C-ABI entry point (tracing): nvtx3 is header-only and binds to a profiler's
// injection library at run time; without one a range is a load and a predictable branch
struct ApiRange
{
    explicit ApiRange(const char* name);
    ~ApiRange();
};
#define VKTB_API_RANGE() ::vktb::ApiRange vktbApiRange__(__func__)

// Fault injection: with VKTB_FAIL_ALLOC_AT=n in the environment the n-th vktb_alloc made while the
// variable is set fails like an out-of-memory condition; exercises the error path up to vkt::InvalidValue
bool injectAllocFailure();

void setError(const std::string& msg);
int  fail(const char* what, cudaError_t e);
int  fail(const char* what);

// pinned pool
int  pinnedAlloc(void** p, size_t bytes);
int  pinnedFree(void* p);
bool pinnedOwns(const void* p);                       // pointer lies inside a live pool block
void pinnedMarkBusy(const void* p, cudaStream_t s);   // a DMA on `s` uses the block until now
void pinnedStats(size_t* inUse, size_t* reserved);
void pinnedTrim();

} // namespace vktb
