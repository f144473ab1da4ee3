// tma_k1_stage.cu -- experiment for north_star item 1 ("... over TMA/shared-memory-staged 2D tiles"): does staging K1's
// source window with a 2-D bulk tensor copy (cp.async.bulk.tensor.2d, zero fill outside the image) beat what k_orient does,
// i.e. reading the window as per-thread 32-bit words straight into the registers of the vertical 7-tap?
//
// Both kernels walk the same 64 x 32 output tiles of a w x h x C frame with the same (64 + 10) x (32 + 10) source window
// and run the SAME consumer: the vertical {8,28,56,72,56,28,8} pass on packed u16 pairs (k_orient stage B), results
// xor-folded into one word per thread so nothing is optimised away.
//   (a) ldg : word (column, row segment) items, rows clamped to the image (BORDER_REPLICATE), exactly stage B's loads
//   (b) tma : one elected thread issues ONE tensor copy of the box {240 bytes, 42 rows} (zero fill: the replicate fix-up
//             a real kernel would still owe is NOT charged), everyone waits on the mbarrier, then the same pass from smem
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_k1_stage tools/tma_k1_stage.cu   (no -lcuda: the
// driver entry point is fetched through the runtime).  Output: us per frame for both (colour frames; the one-channel
// instantiation of the tensor-copy kernel faults with `illegal instruction` on this driver and is left out).
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <vector>

constexpr int TW = 64, TH = 32, HALO = 5, THREADS = 256;
constexpr int WIN_ROWS = TH + 2 * HALO; // 42
constexpr int BOX_BYTES = 240;          // >= 3 * (64 + 10) = 222, a multiple of 16

__device__ __forceinline__ uint32_t su32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// the vertical 7-tap of RS output rows of one word column, from RS + 6 source words
template <int RS> __device__ __forceinline__ uint32_t seven_tap(const uint32_t (&rw)[RS + 6])
{
    uint32_t acc = 0;
#pragma unroll
    for (int r = 0; r < RS; ++r)
    {
        uint32_t ev[7], od[7];
#pragma unroll
        for (int k = 0; k < 7; ++k)
            ev[k] = rw[r + k] & 0x00ff00ffu, od[k] = (rw[r + k] >> 8) & 0x00ff00ffu;
        const uint32_t se = 8u * (ev[0] + ev[6]) + 28u * (ev[1] + ev[5]) + 56u * (ev[2] + ev[4]) + 72u * ev[3];
        const uint32_t so = 8u * (od[0] + od[6]) + 28u * (od[1] + od[5]) + 56u * (od[2] + od[4]) + 72u * od[3];
        acc ^= se + (so << 3);
    }
    return acc;
}

template <int C> __global__ void __launch_bounds__(THREADS) k_ldg(const uint8_t *__restrict__ src, int w, int h, uint32_t *__restrict__ sink)
{
    constexpr int NW = (C * (TW + 2 * HALO) + 3) / 4 + 1, RS = 9, NSEG = (TH + 4) / RS; // as k_orient<C, 32>
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH, rowbytes = w * C;
    const int bstart = (C * (x0 - HALO)) & ~3;
    uint32_t acc = 0;
    for (int item = threadIdx.x; item < NW * NSEG; item += THREADS)
    {
        const int wc = item % NW, seg = item / NW;
        const int bi = clampi(bstart + 4 * wc, 0, rowbytes - 4);
        uint32_t rw[RS + 6];
#pragma unroll
        for (int j = 0; j < RS + 6; ++j)
        {
            const int sy = clampi(y0 - HALO + seg * RS + j, 0, h - 1);
            rw[j] = __ldg(reinterpret_cast<const uint32_t *>(src + (size_t)sy * rowbytes + bi));
        }
        acc ^= seven_tap<RS>(rw);
    }
    if (acc == 0x9e3779b9u)
        sink[0] = acc;
}

template <int C> __global__ void __launch_bounds__(THREADS) k_tma(const __grid_constant__ CUtensorMap map, int w, int h, uint32_t *__restrict__ sink)
{
    constexpr int NW = (C * (TW + 2 * HALO) + 3) / 4 + 1, RS = 9, NSEG = (TH + 4) / RS;
    __shared__ __align__(128) uint8_t s_win[WIN_ROWS][BOX_BYTES];
    __shared__ __align__(8) uint64_t bar;
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    if (threadIdx.x == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(su32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(su32(&bar)), "r"(WIN_ROWS * BOX_BYTES) : "memory");
        const int c0 = (C * (x0 - HALO)) & ~3, c1 = y0 - HALO; // coordinates may be negative: the hardware fills with zeros
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(su32(&s_win[0][0])),
                     "l"(&map), "r"(c0), "r"(c1), "r"(su32(&bar))
                     : "memory");
    }
    __syncthreads();
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(su32(&bar)) : "memory");
    uint32_t acc = 0;
    for (int item = threadIdx.x; item < NW * NSEG; item += THREADS)
    {
        const int wc = item % NW, seg = item / NW;
        uint32_t rw[RS + 6];
#pragma unroll
        for (int j = 0; j < RS + 6; ++j)
            rw[j] = *reinterpret_cast<const uint32_t *>(&s_win[min(seg * RS + j, WIN_ROWS - 1)][4 * wc]);
        acc ^= seven_tap<RS>(rw);
    }
    if (acc == 0x9e3779b9u)
        sink[0] = acc;
}

typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                             const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int C> int run(int w, int h, EncodeFn encode)
{
    const size_t bytes = (size_t)w * h * C;
    uint8_t *d;
    uint32_t *sink;
    cudaMalloc(&d, bytes);
    cudaMalloc(&sink, 4);
    std::vector<uint8_t> hbuf(bytes);
    uint32_t rng = 1;
    for (auto &b : hbuf)
        rng = rng * 1664525u + 1013904223u, b = (uint8_t)(rng >> 24);
    cudaMemcpy(d, hbuf.data(), bytes, cudaMemcpyHostToDevice);
    CUtensorMap map;
    const cuuint64_t gdim[2] = {(cuuint64_t)w * C, (cuuint64_t)h}, gstride[1] = {(cuuint64_t)w * C};
    const cuuint32_t box[2] = {BOX_BYTES, WIN_ROWS}, estr[2] = {1, 1};
    const CUresult r = encode(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                              CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
    {
        printf("cuTensorMapEncodeTiled failed: %d\n", (int)r);
        return 1;
    }
    const dim3 grid((w + TW - 1) / TW, (h + TH - 1) / TH);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 50;
    float ms[2];
    cudaStream_t st;
    cudaStreamCreate(&st);
    for (int variant = 0; variant < 2; ++variant)
    {
        // the launches as ONE captured graph: stream launches on this host are CPU-paced at ~4 us each, longer than the kernels
        cudaGraph_t g;
        cudaGraphExec_t ge;
        cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
        for (int it = 0; it < iters; ++it)
        {
            if (variant == 0)
                k_ldg<C><<<grid, THREADS, 0, st>>>(d, w, h, sink);
            else
                k_tma<C><<<grid, THREADS, 0, st>>>(map, w, h, sink);
        }
        cudaStreamEndCapture(st, &g);
        cudaGraphInstantiate(&ge, g, 0);
        cudaGraphLaunch(ge, st);
        cudaEventRecord(e0, st);
        cudaGraphLaunch(ge, st);
        cudaEventRecord(e1, st);
        const cudaError_t e = cudaDeviceSynchronize();
        if (e)
        {
            printf("%dx%dx%d variant %s: error: %s\n", w, h, C, variant ? "tma" : "ldg", cudaGetErrorString(e));
            return 1;
        }
        cudaEventElapsedTime(&ms[variant], e0, e1);
    }
    printf("%dx%dx%d  tiles %d  ldg-to-registers %.2f us/frame   tma-to-smem %.2f us/frame   (window + vertical 7-tap only; %d launches each, back to back)\n", w, h, C,
           grid.x * grid.y, ms[0] * 1e3 / iters, ms[1] * 1e3 / iters, iters);
    cudaFree(d);
    cudaFree(sink);
    return 0;
}

int main()
{
    EncodeFn encode = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&encode, cudaEnableDefault, &q) != cudaSuccess || !encode)
    {
        printf("cuTensorMapEncodeTiled not available\n");
        return 1;
    }
    return run<3>(1024, 1024, encode) || run<3>(512, 512, encode) || run<3>(4096, 4096, encode);
}
