// launch_quant.cu -- is the cost of back-to-back dependent kernels on one stream quantised?  A kernel that spins for
// a given number of ns (globaltimer) on `grid` CTAs is launched 50 times between two events.
#include <cuda_runtime.h>
#include <cstdio>
__global__ void spin(unsigned long long ns)
{
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    do
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    while (t - t0 < ns);
}
// (appended) same chains as a captured CUDA graph: node-to-node cost without the CPU in the loop
struct GraphRun
{
    static void run(int grid, unsigned long long ns, cudaStream_t st, cudaEvent_t e0, cudaEvent_t e1)
    {
        cudaGraph_t g;
        cudaGraphExec_t ge;
        cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal);
        for (int i = 0; i < 50; ++i)
            spin<<<grid, 256, 0, st>>>(ns);
        cudaStreamEndCapture(st, &g);
        cudaGraphInstantiate(&ge, g, 0);
        cudaGraphLaunch(ge, st);
        cudaStreamSynchronize(st);
        cudaEventRecord(e0, st);
        cudaGraphLaunch(ge, st);
        cudaEventRecord(e1, st);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        printf("GRAPH grid=%4d spin=%5llu ns : %.2f us per node\n", grid, ns, ms * 1e3 / 50);
        cudaGraphExecDestroy(ge);
        cudaGraphDestroy(g);
    }
};

int main()
{
    cudaStream_t st;
    cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int grid : {1, 148, 1024})
        for (unsigned long long ns : {0ull, 1000ull, 2000ull, 3000ull, 4000ull, 5000ull, 6000ull, 8000ull, 10000ull, 15000ull})
        {
            for (int i = 0; i < 5; ++i)
                spin<<<grid, 256, 0, st>>>(ns);
            cudaEventRecord(e0, st);
            for (int i = 0; i < 50; ++i)
                spin<<<grid, 256, 0, st>>>(ns);
            cudaEventRecord(e1, st);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            printf("grid=%4d spin=%5llu ns : %.2f us per launch\n", grid, ns, ms * 1e3 / 50);
        }
    for (int grid : {148, 1024})
        for (unsigned long long ns : {0ull, 1000ull, 2000ull, 3000ull, 4000ull, 6000ull, 10000ull})
            GraphRun::run(grid, ns, st, e0, e1);
    return 0;
}
