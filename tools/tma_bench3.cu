// tma_bench3.cu -- cp.async.bulk issue cost: (a) issuing warps per CTA, (b) one expect_tx per copy vs per round,
// (c) per-copy mbarriers (ring) vs one barrier.  4 CTAs/SM, 2064-byte copies from a 2 MiB L2-resident buffer.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
constexpr int SLOTS = 16, MAXW = 4;
__device__ __forceinline__ uint32_t su32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(128) bench(const uint8_t *src, uint32_t span, int rounds, int per_round, uint32_t bytes, int issuers,
                                             int expect_each, const uint32_t *offs)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar[MAXW];
    const uint32_t stride = (bytes + 127) & ~127u;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0)
    {
        for (int w = 0; w < MAXW; ++w)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(su32(&bar[w])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp >= issuers)
        return;
    const uint32_t b = su32(&bar[warp]);
    uint8_t *my = smem + warp * SLOTS * stride;
    const uint32_t *o = offs + (blockIdx.x * MAXW + warp) * 64;
    for (int r = 0; r < rounds; ++r)
    {
        if (lane == 0)
        {
            if (!expect_each)
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes * per_round) : "memory");
            for (int i = 0; i < per_round; ++i)
            {
                if (expect_each)
                {
                    if (i == 0)
                        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
                    else
                        asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
                }
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(su32(my + (i % SLOTS) * stride)), "l"(src + o[(r * per_round + i) & 63]), "r"(bytes), "r"(b) : "memory");
            }
            uint32_t ok = 0;
            while (!ok)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(ok) : "r"(b), "r"(r & 1) : "memory");
        }
        __syncwarp();
    }
}
int main()
{
    const uint32_t span = 2u << 20, bytes = 2064;
    uint8_t *d;
    uint32_t *doffs;
    cudaMalloc(&d, span);
    cudaMemset(d, 1, span);
    const int grid = 148 * 4;
    uint32_t *h = new uint32_t[grid * MAXW * 64];
    uint32_t rng = 12345;
    for (int i = 0; i < grid * MAXW * 64; ++i)
    {
        rng = rng * 1664525u + 1013904223u;
        h[i] = ((rng >> 8) % (span - bytes - 256)) & ~15u;
    }
    cudaMalloc(&doffs, grid * MAXW * 64 * 4);
    cudaMemcpy(doffs, h, grid * MAXW * 64 * 4, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const size_t sm = (size_t)MAXW * SLOTS * ((bytes + 127) & ~127u);
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    for (int issuers : {1, 2, 4})
        for (int per_round : {1, 4, 16})
            for (int expect_each : {0, 1})
            {
                const int total = 256, rounds = total / per_round;
                bench<<<grid, 128, sm>>>(d, span, rounds, per_round, bytes, issuers, expect_each, doffs);
                cudaEventRecord(e0);
                for (int it = 0; it < 5; ++it)
                    bench<<<grid, 128, sm>>>(d, span, rounds, per_round, bytes, issuers, expect_each, doffs);
                cudaEventRecord(e1);
                cudaError_t e = cudaDeviceSynchronize();
                if (e) { printf("err %s\n", cudaGetErrorString(e)); return 1; }
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                double us = ms * 1e3 / 5;
                double gb = (double)grid * issuers * total * bytes / (us * 1e-6) / 1e9;
                printf("issuers/CTA=%d per_round=%2d expect_each=%d : %7.1f us  %7.0f GB/s  %.1f ns/copy/issuer\n", issuers, per_round, expect_each, us, gb,
                       us * 1e3 / total);
            }
    return 0;
}
