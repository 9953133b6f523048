// Micro-benchmark behind ldUnaligned (kernels/common.cuh): how to feed 128-bit stores from a source that sits at
// another 16-byte phase.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o unaligned_probe unaligned_probe.cu
//   0 aligned copy (baseline)          1 two __ldg (L1-allocating) loads + funnel shift
//   2 two L1::no_allocate loads        3 one load + hi from lane+1 by shuffle (lane 31 loads its own)
//   4 two plain ld.global loads
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint4 ldNa(const void* p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void stNa(void* p, const uint4& v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 funnel(const uint4& lo, const uint4& hi, unsigned sh)
{
    uint32_t w0 = lo.x, w1 = lo.y, w2 = lo.z, w3 = lo.w, w4 = hi.x, w5 = hi.y;
    if (sh & 8u) { w0 = w2; w1 = w3; w2 = w4; w3 = w5; w4 = hi.z; w5 = hi.w; }
    if (sh & 4u) { w0 = w1; w1 = w2; w2 = w3; w3 = w4; w4 = w5; }
    const unsigned bits = (sh & 3u) * 8u;
    return make_uint4(__funnelshift_r(w0, w1, bits), __funnelshift_r(w1, w2, bits), __funnelshift_r(w2, w3, bits),
                      __funnelshift_r(w3, w4, bits));
}

template <int MODE>
__global__ void __launch_bounds__(256) probe(uint8_t* __restrict__ d, const uint8_t* __restrict__ s, size_t nvec, unsigned sh)
{
    constexpr int U = 4;
    const uint4* sv = reinterpret_cast<const uint4*>(s);   // aligned base; the data starts sh bytes in
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i + (U - 1) * stride < nvec; i += U * stride) {
        uint4 q[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t k = i + u * stride;
            if (MODE == 0) q[u] = ldNa(sv + k);
            else if (MODE == 1) q[u] = funnel(__ldg(sv + k), __ldg(sv + k + 1), sh);
            else if (MODE == 2) q[u] = funnel(ldNa(sv + k), ldNa(sv + k + 1), sh);
            else if (MODE == 4) q[u] = funnel(sv[k], sv[k + 1], sh);
            else {
                const uint4 lo = ldNa(sv + k);
                uint4 hi;
                hi.x = __shfl_down_sync(0xffffffffu, lo.x, 1); hi.y = __shfl_down_sync(0xffffffffu, lo.y, 1);
                hi.z = __shfl_down_sync(0xffffffffu, lo.z, 1); hi.w = __shfl_down_sync(0xffffffffu, lo.w, 1);
                if ((threadIdx.x & 31) == 31) hi = __ldg(sv + k + 1);
                q[u] = funnel(lo, hi, sh);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) stNa(d + (i + u * stride) * 16, q[u]);
    }
}

int main()
{
    const size_t bytes = (size_t)8 << 30, nvec = bytes / 16;
    uint8_t *d, *s;
    cudaMalloc(&d, bytes); cudaMalloc(&s, bytes + 256);
    cudaMemset(s, 1, bytes + 256);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int mode = 0; mode < 5; ++mode)
        for (unsigned sh : { 2u, 14u }) {
            float best = 1e9f;
            for (int it = 0; it < 6; ++it) {
                cudaEventRecord(a);
                const int grid = 148 * 8;
                switch (mode) {
                case 0: probe<0><<<grid, 256>>>(d, s, nvec, sh); break;
                case 1: probe<1><<<grid, 256>>>(d, s, nvec, sh); break;
                case 2: probe<2><<<grid, 256>>>(d, s, nvec, sh); break;
                case 3: probe<3><<<grid, 256>>>(d, s, nvec, sh); break;
                case 4: probe<4><<<grid, 256>>>(d, s, nvec, sh); break;
                }
                cudaEventRecord(b); cudaEventSynchronize(b);
                float ms; cudaEventElapsedTime(&ms, a, b);
                if (it >= 2 && ms < best) best = ms;
            }
            printf("mode %d shift %2u: %.3f ms  %.1f GB/s (read+write)  err=%d\n", mode, sh, best, 2.0 * bytes / best / 1e6,
                   (int)cudaGetLastError());
        }
    return 0;
}
