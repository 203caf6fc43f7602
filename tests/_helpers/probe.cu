// Test helper (NOT part of the product): evaluates CUDA libdevice's hypotf / atan2f / cosf on the GPU so that
// oracle/libdevice_f32.h (the CPU restatement the oracle uses) can be pinned bit-for-bit. Compiled with nvcc's
// default floating-point flags, i.e. the flags the reference is built with (reference Makefile:12).
#include <cuda_runtime.h>
#include <stddef.h>

__global__ void k_probe(int fn, size_t n, const float* a, const float* b, float* out)
{
    size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x;
    if(i >= n) return;
    if(fn == 0) out[i] = hypotf(a[i], b[i]);
    else if(fn == 1) out[i] = atan2f(a[i], b[i]);
    else out[i] = cosf(a[i]);
}

extern "C" int probe_eval(int fn, size_t n, const float* a, const float* b, float* out)
{
    float *da = nullptr, *db = nullptr, *dout = nullptr;
    if(cudaMalloc(&da, n * 4) != cudaSuccess) return -1;
    if(cudaMalloc(&db, n * 4) != cudaSuccess) return -1;
    if(cudaMalloc(&dout, n * 4) != cudaSuccess) return -1;
    cudaMemcpy(da, a, n * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(db, b ? b : a, n * 4, cudaMemcpyHostToDevice);
    k_probe<<<(unsigned)((n + 255) / 256), 256>>>(fn, n, da, db, dout);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(out, dout, n * 4, cudaMemcpyDeviceToHost);
    cudaFree(da); cudaFree(db); cudaFree(dout);
    return e == cudaSuccess ? 0 : -2;
}
