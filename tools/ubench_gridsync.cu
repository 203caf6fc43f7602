// ubench_gridsync.cu - cost of a cooperative grid barrier on this GPU, and of a hand-written one (development tool).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -rdc=false tools/ubench_gridsync.cu -o tools/_bin/ubench_gridsync
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <stdio.h>
#include <stdlib.h>
namespace cg = cooperative_groups;
#define CK(x) do { cudaError_t e_ = (x); if(e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while(0)

__global__ void k_cg(int n, double* sink)
{
    cg::grid_group g = cg::this_grid();
    for(int i = 0; i < n; i++) g.sync();
    if(threadIdx.x == 0 && blockIdx.x == 0) sink[0] = n;
}

// sense-free counting barrier: one atomic per block on a monotonically increasing counter, spin on a volatile load
__device__ __forceinline__ void my_sync(unsigned* counter, unsigned& epoch)
{
    __syncthreads();
    epoch += gridDim.x;
    if(threadIdx.x == 0)
    {
        __threadfence();
        atomicAdd(counter, 1u);
        while(*(volatile unsigned*) counter < epoch) { }
        __threadfence();
    }
    __syncthreads();
}

__global__ void k_my(int n, unsigned* counter, double* sink)
{
    unsigned epoch = 0;
    for(int i = 0; i < n; i++) my_sync(counter, epoch);
    if(threadIdx.x == 0 && blockIdx.x == 0) sink[0] = n;
}

int main()
{
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    double* sink; unsigned* counter; CK(cudaMalloc(&sink, 8)); CK(cudaMalloc(&counter, 4));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for(int per_sm : {1, 2, 4}) for(int threads : {256, 1024})
    {
        if(per_sm * threads > 2048) continue;
        int blocks = sms * per_sm, n = 2000;
        void* a1[] = {&n, &sink};
        CK(cudaLaunchCooperativeKernel((const void*) k_cg, dim3(blocks), dim3(threads), a1, 0, 0));
        CK(cudaEventRecord(e0));
        CK(cudaLaunchCooperativeKernel((const void*) k_cg, dim3(blocks), dim3(threads), a1, 0, 0));
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("grid.sync   %4d blocks x %4d threads: %.2f us per barrier\n", blocks, threads, ms * 1e3 / n);
        CK(cudaMemset(counter, 0, 4));
        void* a2[] = {&n, &counter, &sink};
        CK(cudaLaunchCooperativeKernel((const void*) k_my, dim3(blocks), dim3(threads), a2, 0, 0));
        CK(cudaMemset(counter, 0, 4));
        CK(cudaEventRecord(e0));
        CK(cudaLaunchCooperativeKernel((const void*) k_my, dim3(blocks), dim3(threads), a2, 0, 0));
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("counter     %4d blocks x %4d threads: %.2f us per barrier\n", blocks, threads, ms * 1e3 / n);
    }
    return 0;
}
