// ubench_fp64_operands.cu - FP64 issue rate by operand form and in mixes with ALU / FMA-pipe work (development tool;
// results in profiles/r02_kernel_experiments.md). DFMA with three distinct register operands sustains 1.30 warp-instr/clk/SM
// on the B200, not the nominal 2.0; integer / float instructions between FP64 instructions ADD to the FP64 time.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e_ = (x); if(e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while(0)
template <int MODE, int ILP>
__global__ void __launch_bounds__(128) k(double* out, const double* in, int iters)
{
    double x[ILP], y[ILP], z[ILP]; float f[ILP]; int q[ILP];
    #pragma unroll
    for(int i = 0; i < ILP; i++) { x[i] = in[threadIdx.x + i]; y[i] = in[threadIdx.x + 64 + i]; z[i] = in[threadIdx.x + 128 + i]; f[i] = (float) x[i]; q[i] = threadIdx.x + i; }
    for(int it = 0; it < iters; it++)
    {
        #pragma unroll
        for(int i = 0; i < ILP; i++)
        {
            if(MODE == 0) x[i] = fma(y[i], z[i], x[i]);                              // DFMA, 3 distinct registers
            if(MODE == 1) x[i] = fma(x[i], y[0], z[0]);                              // DFMA, 2 shared operands
            if(MODE == 2) x[i] = x[i] + y[i];                                        // DADD 2 distinct
            if(MODE == 3) { x[i] = fma(y[i], z[i], x[i]); q[i] = (q[i] ^ it) + i; }  // DFMA + LOP/IADD (alu pipe)
            if(MODE == 4) { x[i] = fma(y[i], z[i], x[i]); f[i] = f[i] * 1.0001f; }   // DFMA + FMUL
            if(MODE == 5) { x[i] = fma(y[i], z[i], x[i]); q[i] = (q[i] ^ it) + i; f[i] = f[i] * 1.0001f; }  // DFMA + 2 ALU + FMUL
            if(MODE == 6) { x[i] = x[i] + y[i]; x[i] = x[i] * z[i]; }                // dependent DADD->DMUL
        }
    }
    double s = 0;
    #pragma unroll
    for(int i = 0; i < ILP; i++) s += x[i] + f[i] + q[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE, int ILP>
static void run(const char* name, int warps_per_sm, int dp_per_round, double* d, double* in)
{
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    int clk; CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0));
    const int iters = 20000, threads = 128, blocks = sms * warps_per_sm / 4;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k<MODE, ILP><<<blocks, threads>>>(d, in, 100);
    CK(cudaEventRecord(e0));
    k<MODE, ILP><<<blocks, threads>>>(d, in, iters);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    double warp_instr = (double) blocks * (threads / 32) * iters * ILP * dp_per_round;
    printf("%-28s ILP %d warps/SM %2d: %8.3f ms  %.3f DP warp-instr/clk/SM\n", name, ILP, warps_per_sm, ms, warp_instr / (ms * 1e-3 * clk * 1e3) / sms);
}
int main()
{
    double* d; CK(cudaMalloc(&d, 1 << 26)); double* in; CK(cudaMalloc(&in, 1 << 16)); CK(cudaMemset(in, 0, 1 << 16));
    for(int w : {8, 16, 32})
    {
        if(w == 8)  { run<0, 8>("DFMA 3 distinct", w, 1, d, in); run<0, 4>("DFMA 3 distinct", w, 1, d, in); run<0, 2>("DFMA 3 distinct", w, 1, d, in); run<2, 8>("DADD 2 distinct", w, 1, d, in); }
        if(w == 16) { run<0, 8>("DFMA 3 distinct", w, 1, d, in); run<0, 4>("DFMA 3 distinct", w, 1, d, in); run<0, 2>("DFMA 3 distinct", w, 1, d, in); run<1, 8>("DFMA shared ops", w, 1, d, in); run<2, 8>("DADD 2 distinct", w, 1, d, in); run<2, 2>("DADD 2 distinct", w, 1, d, in);
                      run<3, 8>("DFMA + 2 ALU", w, 1, d, in); run<4, 8>("DFMA + FMUL", w, 1, d, in); run<5, 8>("DFMA + 2 ALU + FMUL", w, 1, d, in); run<6, 8>("DADD->DMUL dep", w, 2, d, in); run<6, 2>("DADD->DMUL dep", w, 2, d, in); }
        if(w == 32) { run<0, 4>("DFMA 3 distinct", w, 1, d, in); run<3, 4>("DFMA + 2 ALU", w, 1, d, in); run<5, 4>("DFMA + 2 ALU + FMUL", w, 1, d, in); }
    }
    return 0;
}
