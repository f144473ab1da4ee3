// imma_probe.cu -- can K1's 7-tap blur and 3x3 Sobel run as banded int8 matrix products on mma.sync?
//  (1) fragment layouts of mma.sync.m16n8k16 / m16n8k32 (.s32.u8.u8 and .s32.s8.u8) checked against a CPU product;
//  (2) issue rate: cycles per MMA per warp with 1 / 2 / 4 / 8 warps per SM sub-partition issuing independent chains.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/imma_probe tools/imma_probe.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ void mma_k16_uu(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ void mma_k16_su(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.u8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ void mma_k16_us(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ void mma_k32_uu(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// A: 16 x K row-major bytes, B: K x 8 with B[k][n] at Bm[n * K + k] (column-major = each column's K bytes contiguous)
template <int K, int MODE> __global__ void k_layout(const uint8_t *A, const uint8_t *Bm, int *D)
{
    const int lane = threadIdx.x & 31, g = lane >> 2, tig = lane & 3;
    int d[4] = {0, 0, 0, 0};
    if (K == 16)
    {
        const uint32_t a0 = *reinterpret_cast<const uint32_t *>(A + g * K + 4 * tig), a1 = *reinterpret_cast<const uint32_t *>(A + (g + 8) * K + 4 * tig);
        const uint32_t b0 = *reinterpret_cast<const uint32_t *>(Bm + g * K + 4 * tig);
        if (MODE == 0)
            mma_k16_uu(d, a0, a1, b0);
        else if (MODE == 1)
            mma_k16_su(d, a0, a1, b0);
        else
            mma_k16_us(d, a0, a1, b0);
    }
    else
    {
        uint32_t a[4];
        a[0] = *reinterpret_cast<const uint32_t *>(A + g * K + 4 * tig), a[1] = *reinterpret_cast<const uint32_t *>(A + (g + 8) * K + 4 * tig);
        a[2] = *reinterpret_cast<const uint32_t *>(A + g * K + 16 + 4 * tig), a[3] = *reinterpret_cast<const uint32_t *>(A + (g + 8) * K + 16 + 4 * tig);
        const uint32_t b0 = *reinterpret_cast<const uint32_t *>(Bm + g * K + 4 * tig), b1 = *reinterpret_cast<const uint32_t *>(Bm + g * K + 16 + 4 * tig);
        mma_k32_uu(d, a, b0, b1);
    }
    D[g * 8 + 2 * tig] = d[0], D[g * 8 + 2 * tig + 1] = d[1], D[(g + 8) * 8 + 2 * tig] = d[2], D[(g + 8) * 8 + 2 * tig + 1] = d[3];
}

template <int K, int CH> __global__ void k_rate(int iters, long long *cycles, int *sink)
{
    int d[CH][4];
    uint32_t a[4] = {threadIdx.x * 0x01010101u, threadIdx.x * 0x00020003u, 0x01020304u, 0x04030201u};
    uint32_t b0 = threadIdx.x * 0x11u, b1 = threadIdx.x ^ 0x5555u;
#pragma unroll
    for (int c = 0; c < CH; ++c)
        d[c][0] = d[c][1] = d[c][2] = d[c][3] = c;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i)
    {
#pragma unroll
        for (int c = 0; c < CH; ++c)
        {
            if (K == 16)
                mma_k16_uu(d[c], a[0], a[1], b0);
            else
                mma_k32_uu(d[c], a, b0, b1);
        }
    }
    const long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c)
        s += d[c][0] + d[c][1] + d[c][2] + d[c][3];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if ((threadIdx.x & 31) == 0)
        cycles[blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)] = t1 - t0;
}

#define CK(x)                                                                                   \
    do                                                                                          \
    {                                                                                           \
        cudaError_t e = (x);                                                                    \
        if (e != cudaSuccess)                                                                   \
        {                                                                                       \
            printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e));                   \
            return 1;                                                                           \
        }                                                                                       \
    } while (0)

template <int K, int MODE> int layout_check(const char *name)
{
    uint8_t hA[16 * K], hB[8 * K];
    for (int i = 0; i < 16 * K; ++i)
        hA[i] = (uint8_t)(rand() & 0xff);
    for (int i = 0; i < 8 * K; ++i)
        hB[i] = (uint8_t)(rand() & 0xff);
    uint8_t *dA, *dB;
    int *dD, hD[128];
    CK(cudaMalloc(&dA, sizeof hA));
    CK(cudaMalloc(&dB, sizeof hB));
    CK(cudaMalloc(&dD, sizeof hD));
    CK(cudaMemcpy(dA, hA, sizeof hA, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB, sizeof hB, cudaMemcpyHostToDevice));
    k_layout<K, MODE><<<1, 32>>>(dA, dB, dD);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(hD, dD, sizeof hD, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int m = 0; m < 16; ++m)
        for (int n = 0; n < 8; ++n)
        {
            int s = 0;
            for (int k = 0; k < K; ++k)
            {
                const int av = (MODE == 1) ? (int)(int8_t)hA[m * K + k] : (int)hA[m * K + k];
                const int bv = (MODE == 2) ? (int)(int8_t)hB[n * K + k] : (int)hB[n * K + k];
                s += av * bv;
            }
            bad += s != hD[m * 8 + n];
        }
    printf("layout %-24s %s (%d of 128 differ)\n", name, bad ? "MISMATCH" : "ok", bad);
    cudaFree(dA), cudaFree(dB), cudaFree(dD);
    return bad;
}

template <int K, int CH> int rate(int warps_per_cta, int ctas_per_sm)
{
    int dev = 0, sms = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int ctas = sms * ctas_per_sm, threads = 32 * warps_per_cta, iters = 2000;
    long long *dc, *hc = (long long *)malloc(sizeof(long long) * ctas * warps_per_cta);
    int *ds;
    CK(cudaMalloc(&dc, sizeof(long long) * ctas * warps_per_cta));
    CK(cudaMalloc(&ds, sizeof(int) * ctas * threads));
    k_rate<K, CH><<<ctas, threads>>>(10, dc, ds);
    k_rate<K, CH><<<ctas, threads>>>(iters, dc, ds);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(hc, dc, sizeof(long long) * ctas * warps_per_cta, cudaMemcpyDeviceToHost));
    double mean = 0;
    for (int i = 0; i < ctas * warps_per_cta; ++i)
        mean += (double)hc[i];
    mean /= ctas * warps_per_cta;
    const double per_mma = mean / ((double)iters * CH);
    // MACs per cycle per SM: warps on the SM x (16*8*K) / cycles per MMA per warp
    printf("rate m16n8k%-2d chains %d, %2d warps/SM: %7.2f cycles per MMA per warp = %8.0f MAC/cycle/SM\n", K, CH, warps_per_cta * ctas_per_sm, per_mma,
           warps_per_cta * ctas_per_sm * 16.0 * 8 * K / per_mma);
    cudaFree(dc), cudaFree(ds), free(hc);
    return 0;
}

int main()
{
    int bad = 0;
    bad += layout_check<16, 0>("m16n8k16 u8.u8");
    bad += layout_check<16, 1>("m16n8k16 s8.u8");
    bad += layout_check<16, 2>("m16n8k16 u8.s8");
    bad += layout_check<32, 0>("m16n8k32 u8.u8");
    rate<16, 1>(4, 1), rate<16, 4>(4, 1), rate<16, 4>(8, 1), rate<16, 4>(8, 2), rate<16, 4>(8, 4);
    rate<32, 1>(4, 1), rate<32, 4>(4, 1), rate<32, 4>(8, 1), rate<32, 4>(8, 2), rate<32, 4>(8, 4);
    return bad != 0;
}
