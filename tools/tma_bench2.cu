// tma_bench2.cu -- raw issue/transfer throughput of 1-D cp.async.bulk: one thread per CTA arms ONE mbarrier with
// the total byte count of a round, issues all copies of the round back to back, then the CTA waits once.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
constexpr int SLOTS = 16;
__device__ __forceinline__ uint32_t su32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(128) bench(const uint8_t *src, uint32_t span, int rounds, int per_round, uint32_t bytes, uint32_t align)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    const uint32_t stride = (bytes + 127) & ~127u;
    if (threadIdx.x == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(su32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t rng = blockIdx.x * 2654435761u + 12345u;
    for (int r = 0; r < rounds; ++r)
    {
        if (threadIdx.x == 0)
        {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(su32(&bar)), "r"(bytes * per_round) : "memory");
            for (int i = 0; i < per_round; ++i)
            {
                rng = rng * 1664525u + 1013904223u;
                uint32_t off = ((rng >> 8) % (span - bytes - 256)) & ~(align - 1);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(su32(smem + (i % SLOTS) * stride)), "l"(src + off), "r"(bytes), "r"(su32(&bar)) : "memory");
            }
        }
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(ok) : "r"(su32(&bar)), "r"(r & 1) : "memory");
        __syncthreads();
    }
}
int main()
{
    const uint32_t span = 2u << 20;
    uint8_t *d;
    cudaMalloc(&d, span);
    cudaMemset(d, 1, span);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int cps = 1; cps <= 4; cps *= 2)
        for (int per_round : {1, 8, 64})
            for (uint32_t bytes : {256u, 1024u, 2064u, 4096u})
                for (uint32_t align : {16u, 128u})
                {
                    const int rounds = 512 / per_round;
                    const size_t sm = (size_t)SLOTS * ((bytes + 127) & ~127u);
                    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
                    int grid = 148 * cps;
                    bench<<<grid, 128, sm>>>(d, span, rounds, per_round, bytes, align);
                    cudaEventRecord(e0);
                    for (int it = 0; it < 5; ++it)
                        bench<<<grid, 128, sm>>>(d, span, rounds, per_round, bytes, align);
                    cudaEventRecord(e1);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e) { printf("err %s\n", cudaGetErrorString(e)); return 1; }
                    float ms;
                    cudaEventElapsedTime(&ms, e0, e1);
                    double us = ms * 1e3 / 5;
                    double gb = (double)grid * 512 * bytes / (us * 1e-6) / 1e9;
                    printf("ctas/sm=%d per_round=%2d bytes=%4u align=%3u : %7.1f us  %7.0f GB/s  %.1f ns/copy/CTA\n", cps, per_round, bytes, align, us, gb,
                           us * 1e3 / 512);
                }
    return 0;
}
