// tma_probe.cu -- does the driver accept an OVERLAPPING 2-D u8 tensor map (row stride 256 B < row extent), and
// does a (256 x R) box at an arbitrary byte coordinate land as R*256 contiguous, realigned bytes in smem?
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cstdlib>
typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                             const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
constexpr int R = 4;
__global__ void probe(const __grid_constant__ CUtensorMap tm, int c0, uint8_t *out, int mode)
{
    const int tx = mode == 1 ? 256 : R * 256;
    __shared__ __align__(128) uint8_t buf[R * 256];
    __shared__ __align__(8) uint64_t bar;
    uint32_t sb = (uint32_t)__cvta_generic_to_shared(&bar), sd = (uint32_t)__cvta_generic_to_shared(buf);
    if (threadIdx.x == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sb));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sb), "r"(tx));
        if (mode == 1)
            asm volatile("cp.async.bulk.tensor.1d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2}], [%3];"
                         ::"r"(sd), "l"(&tm), "r"(c0), "r"(sb) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(sd), "l"(&tm), "r"(c0), "r"(0), "r"(sb) : "memory");
    }
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(sb), "r"(0) : "memory");
    for (int i = threadIdx.x; i < R * 256; i += blockDim.x)
        out[i] = buf[i];
}
int main(int argc, char **argv)
{
    int mode = argc > 1 ? atoi(argv[1]) : 0; // 0: overlapping 2-D, 1: 1-D (box 256), 2: 2-D with row stride 4096
    int only = argc > 2 ? atoi(argv[2]) : -1;
    const size_t N = 1 << 20;
    std::vector<uint8_t> h(N);
    for (size_t i = 0; i < N; ++i)
        h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *o;
    cudaMalloc(&d, N);
    cudaMalloc(&o, R * 256);
    cudaMemcpy(d, h.data(), N, cudaMemcpyHostToDevice);
    EncodeFn enc = nullptr;
    cudaDriverEntryPointQueryResult qr;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&enc, cudaEnableDefault, &qr);
    if (!enc) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    CUtensorMap tm;
    cuuint64_t dims[2] = {N - (R - 1) * 256, R};
    cuuint64_t strides[1] = {256};
    if (mode == 2)
        dims[0] = 4096, strides[0] = 4096, dims[1] = N / 4096;
    cuuint32_t box[2] = {256, R}, es[2] = {1, 1};
    CUresult rc = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, mode == 1 ? 1 : 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode overlapping 2D: rc=%d\n", (int)rc);
    if (rc) return 2;
    int bad_total = 0;
    for (int c0 : {0, 16, 48, 4, 8, 1, 3, 77, 1021, 65537, (int)N - R * 256 - 5})
    {
        if (only >= 0 && c0 != only) continue;
        if (mode == 2 && c0 > 3000) continue;
        probe<<<1, 128>>>(tm, c0, o, mode);
        cudaError_t e = cudaDeviceSynchronize();
        if (e) { printf("kernel: %s\n", cudaGetErrorString(e)); return 3; }
        std::vector<uint8_t> g(R * 256);
        cudaMemcpy(g.data(), o, R * 256, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int i = 0; i < R * 256; ++i)
            bad += g[i] != (mode == 0 ? h[c0 + i] : mode == 1 ? (i < 256 ? h[c0 + i] : g[i]) : h[c0 + (i / 256) * 4096 + i % 256]);
        printf("c0=%d mismatches=%d\n", c0, bad);
        bad_total += bad;
    }
    printf(bad_total ? "FAIL\n" : "PASS\n");
    return bad_total != 0;
}
