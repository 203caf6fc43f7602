// ubench_fp64.cu - what the FP64 pipe of this GPU sustains (development tool; sets the compute ceiling quoted in DESIGN.md).
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e_ = (x); if(e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while(0)

template <int MODE, int ILP>
__global__ void k(double* out, int iters, double a, double b)
{
    double v[ILP]; float f[ILP];
    #pragma unroll
    for(int i = 0; i < ILP; i++) { v[i] = threadIdx.x * 1e-3 + i; f[i] = (float) v[i]; }
    for(int it = 0; it < iters; it++)
    {
        #pragma unroll
        for(int i = 0; i < ILP; i++)
        {
            if(MODE == 0) v[i] = fma(v[i], a, b);                       // DFMA
            if(MODE == 1) v[i] = v[i] + a;                              // DADD
            if(MODE == 2) v[i] = v[i] * a;                              // DMUL
            if(MODE == 3) { f[i] = (float) v[i]; v[i] = (double) f[i] + 0.0; v[i] = __longlong_as_double(__double_as_longlong(v[i]) ^ it); }   // 2 F2F per round
            if(MODE == 4) { v[i] = fma(v[i], a, b); f[i] = fmaf(f[i], 1.0001f, 0.5f); }            // DFMA + FFMA
            if(MODE == 5) { v[i] = fma(v[i], a, b); f[i] = fmaf(f[i], 1.0001f, 0.5f); f[i] = fmaf(f[i], 1.0001f, 0.5f); f[i] = fmaf(f[i], 1.0001f, 0.5f); }   // 1 DFMA : 3 FFMA
        }
    }
    double s = 0;
    #pragma unroll
    for(int i = 0; i < ILP; i++) s += v[i] + f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE, int ILP>
static void run(const char* name, int warps_per_sm, int dp_per_round, double* d)
{
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    int clk; CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0));
    const int iters = 20000, threads = 128, blocks = sms * warps_per_sm / 4;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k<MODE, ILP><<<blocks, threads>>>(d, 100, 1.0000001, 1e-9);
    CK(cudaEventRecord(e0));
    k<MODE, ILP><<<blocks, threads>>>(d, iters, 1.0000001, 1e-9);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    double warp_instr = (double) blocks * (threads / 32) * iters * ILP * dp_per_round;
    double per_clk_sm = warp_instr / (ms * 1e-3 * clk * 1e3) / sms;
    printf("%-14s ILP %d warps/SM %2d: %8.3f ms  %.3f FP64-pipe warp-instr/clk/SM (at the nominal %d MHz)\n", name, ILP, warps_per_sm, ms, per_clk_sm, clk / 1000);
}

int main()
{
    double* d; CK(cudaMalloc(&d, 1 << 26));
    for(int w : {4, 8, 16, 32, 64})
    {
        if(w == 4)  { run<0, 8>("DFMA", 4, 1, d);  run<0, 2>("DFMA", 4, 1, d); }
        if(w == 8)  { run<0, 8>("DFMA", 8, 1, d);  run<0, 2>("DFMA", 8, 1, d); }
        if(w == 16) { run<0, 8>("DFMA", 16, 1, d); run<0, 2>("DFMA", 16, 1, d); run<0, 1>("DFMA", 16, 1, d); run<1, 8>("DADD", 16, 1, d); run<2, 8>("DMUL", 16, 1, d); run<3, 8>("F2F x2", 16, 2, d); run<4, 8>("DFMA+FFMA", 16, 1, d); run<5, 8>("DFMA+3FFMA", 16, 1, d); }
        if(w == 32) { run<0, 8>("DFMA", 32, 1, d); run<0, 1>("DFMA", 32, 1, d); run<3, 8>("F2F x2", 32, 2, d); run<5, 8>("DFMA+3FFMA", 32, 1, d); }
        if(w == 64) { run<0, 4>("DFMA", 64, 1, d); run<0, 1>("DFMA", 64, 1, d); }
    }
    return 0;
}
