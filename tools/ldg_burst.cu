// ldg_burst.cu -- how long does ONE warp wait for a burst of N independent global loads that miss L1 and hit L2?
// (k_similarity_bits at M1: a warp issues 16 x LDG.64 per 8 features and takes ~550 cycles per such group whatever else
// runs on the SM; is that latency, a per-warp limit, or a per-SM limit?)
// Each warp reads N warp-wide rows (32 lanes x VEC words, contiguous) `row_stride` bytes apart out of a buffer that the
// previous launch left in L2 (L1 is invalidated at every launch).  clock64() around issue + first use.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ldg_burst tools/ldg_burst.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

template <int VEC, int N, int NC = 1> __global__ void k(const uint32_t *__restrict__ buf, size_t warp_span_words, int row_stride_words, long long *cycles, uint32_t *sink)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const uint32_t *p = buf + (size_t)warp * warp_span_words + lane * VEC;
    uint32_t v[N][VEC];
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < N; ++i)
    {
        const uint32_t *q = p + (size_t)i * row_stride_words;
        if (NC == 0 && VEC == 2) // plain LDG.E instead of LDG.E.CONSTANT
            asm volatile("ld.global.v2.u32 {%0,%1}, [%2];" : "=r"(v[i][0]), "=r"(v[i][VEC > 1 ? 1 : 0]) : "l"(q));
        else if (NC == 2 && VEC == 2) // L1 bypass
            asm volatile("ld.global.cg.v2.u32 {%0,%1}, [%2];" : "=r"(v[i][0]), "=r"(v[i][VEC > 1 ? 1 : 0]) : "l"(q));
        else if (VEC == 1)
            asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v[i][0]) : "l"(q));
        else if (VEC == 2)
            asm volatile("ld.global.nc.v2.u32 {%0,%1}, [%2];" : "=r"(v[i][0]), "=r"(v[i][VEC > 1 ? 1 : 0]) : "l"(q));
        else
            asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v[i][0]), "=r"(v[i][VEC > 1 ? 1 : 0]), "=r"(v[i][VEC > 2 ? 2 : 0]), "=r"(v[i][VEC > 3 ? 3 : 0])
                         : "l"(q));
    }
    const long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
        for (int j = 0; j < VEC; ++j)
            acc ^= v[i][j];
    asm volatile("" ::"r"(acc) : "memory");
    const long long t2 = clock64();
    if (acc == 0x12345678u)
        sink[0] = acc;
    if (lane == 0 && warp == (gridDim.x / 2) * (blockDim.x >> 5))
        cycles[0] = t1 - t0, cycles[1] = t2 - t0;
}

template <int VEC, int N, int NC = 1> void run(const uint32_t *buf, size_t buf_words, int ctas, int warps, int row_stride_bytes, long long *d_c, uint32_t *sink)
{
    const size_t span = (size_t)N * row_stride_bytes / 4 + 64; // words per warp: all rows distinct
    if ((size_t)ctas * warps * span > buf_words)
    {
        printf("buffer too small\n");
        return;
    }
    long long h[2];
    for (int rep = 0; rep < 3; ++rep)
        k<VEC, N, NC><<<ctas, warps * 32>>>(buf, span, row_stride_bytes / 4, d_c, sink);
    cudaMemcpy(h, d_c, sizeof(h), cudaMemcpyDeviceToHost);
    printf("  %s LDG.%-3d x %2d  rows %4d B apart, %3d CTAs x %d warps: issue %5lld cycles, all data there after %5lld cycles (%.0f per load, %.1f B/cycle/warp)\n", NC == 1 ? ".nc" : NC == 0 ? "   " : ".cg", 32 * VEC, N,
           row_stride_bytes, ctas, warps, h[0], h[1], (double)h[1] / N, 128.0 * VEC * N / h[1]);
}

int main()
{
    const size_t words = 64u << 20; // 256 MiB: never fully L2-resident, but every warp's span is re-read by the same warp on the next launch
    uint32_t *buf, *sink;
    long long *d_c;
    cudaMalloc(&buf, words * 4);
    cudaMemset(buf, 1, words * 4);
    cudaMalloc(&sink, 4);
    cudaMalloc(&d_c, 16);
    for (int warps : {1, 4, 8})
    {
        printf("%d warp(s) per CTA, one CTA per SM\n", warps);
        run<2, 1>(buf, words, 148, warps, 512, d_c, sink);
        run<2, 2>(buf, words, 148, warps, 512, d_c, sink);
        run<2, 4>(buf, words, 148, warps, 512, d_c, sink);
        run<2, 8>(buf, words, 148, warps, 512, d_c, sink);
        run<2, 16>(buf, words, 148, warps, 512, d_c, sink);
        run<2, 32>(buf, words, 148, warps, 512, d_c, sink);
        run<1, 16>(buf, words, 148, warps, 512, d_c, sink);
        run<1, 32>(buf, words, 148, warps, 512, d_c, sink);
        run<4, 8>(buf, words, 148, warps, 512, d_c, sink);
        run<4, 16>(buf, words, 148, warps, 512, d_c, sink);
    }
    printf("plain and .cg loads instead of .nc, 4 warps per CTA\n");
    run<2, 16, 0>(buf, words, 148, 4, 512, d_c, sink);
    run<2, 32, 0>(buf, words, 148, 4, 512, d_c, sink);
    run<2, 16, 2>(buf, words, 148, 4, 512, d_c, sink);
    run<2, 32, 2>(buf, words, 148, 4, 512, d_c, sink);
    run<2, 32, 0>(buf, words, 37, 4, 8, d_c, sink);
    run<2, 32, 2>(buf, words, 37, 4, 8, d_c, sink);
    printf("the product's pattern: pairs of overlapping LDG.64 (second one 8 bytes further)\n");
    run<2, 32>(buf, words, 37, 4, 8, d_c, sink); // rows 8 B apart: 32 loads over the same three lines (all but the first hit or merge)
    const cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
