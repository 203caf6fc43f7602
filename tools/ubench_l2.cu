// ubench_l2.cu - copy bandwidth as a function of the footprint: what "L2-resident" buys on this GPU (SURVEY 8d asks for the L2
// figure to be measured on the box; MEASURED_PEAKS.json only has the HBM copy number). Development / measurement tool.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 tools/ubench_l2.cu -o tools/_bin/ubench_l2
// For each footprint (read buffer + write buffer) the copy kernel runs back to back; after the first pass the whole
// footprint sits in L2 when it fits (126 MB on B200, two partitions). Bytes counted = read + written, like the HBM figure.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e_ = (x); if(e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while(0)

__global__ void __launch_bounds__(256) k_copy(const double2* __restrict__ src, double2* __restrict__ dst, size_t n)
{
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x)
        dst[i] = src[i];
}

__global__ void __launch_bounds__(256) k_read(const double2* __restrict__ src, double* __restrict__ sink, size_t n)
{
    double acc = 0;
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x)
    {
        const double2 v = src[i];
        acc += v.x + v.y;
    }
    if(acc == 1.2345e300) sink[0] = acc;
}

int main()
{
    int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    int l2 = 0; CK(cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, 0));
    printf("SMs %d, L2 %.1f MB\n", sms, l2 / 1e6);
    const size_t max_bytes = (size_t) 1 << 30;
    double2 *a, *b; double* sink;
    CK(cudaMalloc(&a, max_bytes)); CK(cudaMalloc(&b, max_bytes)); CK(cudaMalloc(&sink, 8));
    CK(cudaMemset(a, 0, max_bytes)); CK(cudaMemset(b, 0, max_bytes));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int blocks = sms * 8;
    printf("%-28s %14s %14s\n", "footprint (read + write)", "copy GB/s", "read-only GB/s");
    const double mbs[] = {8, 16, 32, 48, 64, 96, 128, 192, 256, 512, 2048};
    for(double mb : mbs)
    {
        const size_t half = (size_t)(mb * 1e6 / 2) / 16 * 16;           // bytes per buffer
        const size_t n = half / sizeof(double2);
        const int reps = mb <= 256 ? 200 : 20;
        for(int w = 0; w < 3; w++) k_copy<<<blocks, 256>>>(a, b, n);
        CK(cudaEventRecord(e0));
        for(int r = 0; r < reps; r++) k_copy<<<blocks, 256>>>(a, b, n);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double copy = 2.0 * half * reps / (ms * 1e-3) / 1e9;
        // read-only over the same total footprint
        const size_t n2 = 2 * n <= max_bytes / sizeof(double2) ? 2 * n : max_bytes / sizeof(double2);
        for(int w = 0; w < 3; w++) k_read<<<blocks, 256>>>(a, sink, n2);
        CK(cudaEventRecord(e0));
        for(int r = 0; r < reps; r++) k_read<<<blocks, 256>>>(a, sink, n2);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double rd = (double) n2 * sizeof(double2) * reps / (ms * 1e-3) / 1e9;
        printf("%10.0f MB                 %14.0f %14.0f\n", mb, copy, rd);
    }
    CK(cudaGetLastError());
    return 0;
}
