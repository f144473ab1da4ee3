// tma_bench.cu -- throughput of 1-D cp.async.bulk (global -> shared) from an L2-resident buffer, as a function
// of copy size and source alignment.  Each CTA keeps DEPTH copies in flight; nothing consumes the data.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
constexpr int DEPTH = 16;
__device__ __forceinline__ uint32_t su32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(128) bench(const uint8_t *src, uint32_t span, int ncopies, uint32_t bytes, uint32_t align, int consume,
                                             uint32_t *sink)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[DEPTH];
    const uint32_t stride = (bytes + 127) & ~127u;
    if (threadIdx.x == 0)
    {
        for (int s = 0; s < DEPTH; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(su32(&full[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t rng = blockIdx.x * 2654435761u + 12345u, acc = 0;
    for (int f = 0; f < ncopies + DEPTH; ++f)
    {
        const int s = f % DEPTH;
        if (f >= DEPTH)
        {
            uint32_t ok = 0, par = ((f - DEPTH) / DEPTH) & 1;
            while (!ok)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(ok) : "r"(su32(&full[s])), "r"(par) : "memory");
            if (consume)
            {
                const uint32_t *b = reinterpret_cast<const uint32_t *>(smem + s * stride) + threadIdx.x;
                for (uint32_t k = 0; k * 512 + 512 <= bytes; ++k)
                    acc += b[128 * k];
            }
            __syncthreads();
        }
        if (f < ncopies && threadIdx.x == 0)
        {
            rng = rng * 1664525u + 1013904223u;
            uint32_t off = ((rng >> 8) % (span - bytes - 256)) & ~(align - 1);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(su32(&full[s])), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(su32(smem + s * stride)), "l"(src + off), "r"(bytes), "r"(su32(&full[s])) : "memory");
        }
    }
    if (acc == 0xdeadbeef)
        *sink = acc;
}
int main()
{
    const uint32_t span = 2u << 20;
    uint8_t *d;
    uint32_t *sink;
    cudaMalloc(&d, span);
    cudaMalloc(&sink, 4);
    cudaMemset(d, 1, span);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int ncopies = 256;
    for (int consume = 0; consume < 2; ++consume)
        for (int cps = 1; cps <= 4; cps *= 2)
            for (uint32_t bytes : {512u, 1024u, 2048u, 2064u, 4096u})
                for (uint32_t align : {16u, 128u})
                {
                    const size_t sm = (size_t)DEPTH * ((bytes + 127) & ~127u);
                    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
                    int grid = 148 * cps;
                    bench<<<grid, 128, sm>>>(d, span, ncopies, bytes, align, consume, sink);
                    cudaEventRecord(e0);
                    for (int it = 0; it < 5; ++it)
                        bench<<<grid, 128, sm>>>(d, span, ncopies, bytes, align, consume, sink);
                    cudaEventRecord(e1);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e) { printf("err %s\n", cudaGetErrorString(e)); return 1; }
                    float ms;
                    cudaEventElapsedTime(&ms, e0, e1);
                    double us = ms * 1e3 / 5;
                    double gb = (double)grid * ncopies * bytes / (us * 1e-6) / 1e9;
                    printf("consume=%d ctas/sm=%d bytes=%4u align=%3u : %7.1f us  %7.0f GB/s  %.1f ns/copy/SM\n", consume, cps, bytes, align, us,
                           gb, us * 1e3 / (ncopies * cps));
                }
    return 0;
}
