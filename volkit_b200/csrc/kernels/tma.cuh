// sm_100a asynchronous-copy primitives: mbarrier, 1-D bulk copies (cp.async.bulk, SASS UBLKCP) and 3-D
// tensor-map tile loads (cp.async.bulk.tensor, SASS UTMALDG).  Inline PTX, no library.
//
// Conventions: shared-memory operands are 32-bit shared-window addresses (smemAddr()); global and shared
// addresses and byte counts of bulk copies are multiples of 16.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace vktb {

__device__ __forceinline__ uint32_t smemAddr(const void* p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier ------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t arrivals)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(arrivals) : "memory");
}
// makes mbarrier.init visible to the async proxy (TMA unit) before the first copy is issued
__device__ __forceinline__ void fenceBarrierInit()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbarArriveExpectTx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarArrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ bool mbarTryWait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity)
{
    while (!mbarTryWait(bar, parity)) { }
}

// ---- proxy fence: generic-proxy smem writes -> visible to the async proxy (before a bulk store) ----------
__device__ __forceinline__ void fenceProxyAsyncSmem()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- 1-D bulk copies -----------------------------------------------------------------------------------
__device__ __forceinline__ void bulkLoad(uint32_t dstSmem, const void* srcGlobal, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dstSmem), "l"(srcGlobal), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulkStore(void* dstGlobal, uint32_t srcSmem, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :: "l"(dstGlobal), "r"(srcSmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulkCommit()
{
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// waits until at most N of this thread's bulk groups still READ their shared-memory source
template <int N>
__device__ __forceinline__ void bulkWaitRead()
{
    asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulkWaitAll()
{
    asm volatile("cp.async.bulk.wait_group %0;" :: "n"(N) : "memory");
}

// ---- tensor-map tile loads (box of a dense x-fastest volume; out-of-bounds elements read as zero) ---------
__device__ __forceinline__ void tensorLoad3d(uint32_t dstSmem, const void* tensorMap, int x, int y, int z, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 :: "r"(dstSmem), "l"(tensorMap), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
}

__device__ __forceinline__ void tensorLoad2d(uint32_t dstSmem, const void* tensorMap, int x, int y, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 :: "r"(dstSmem), "l"(tensorMap), "r"(x), "r"(y), "r"(bar) : "memory");
}

} // namespace vktb
