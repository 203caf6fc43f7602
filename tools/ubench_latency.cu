// ubench_latency.cu - dependent-issue latency of the instructions on the critical path of one right-hand-side
// evaluation (development tool; backs the "latency chain" paragraph of DESIGN.md). One warp on one SM runs a chain in
// which every instruction consumes the previous result; cycles per link = (clock64 difference) / links.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false tools/ubench_latency.cu -o tools/_bin/ubench_latency
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e_ = (x); if(e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while(0)

enum { DFMA, DADD, DMUL, FFMA, FMUL, F2F_DF, F2F_ROUND, RSQ, IADD, LDS_CHAIN, HYPOT };

__device__ __forceinline__ float hypot_chain(float x, float y)          // the branch-free hypot of pf_common.cuh
{
    const int ia = __float_as_int(x) & 0x7fffffff, ib = __float_as_int(y) & 0x7fffffff;
    const int lo_i = min(ia, ib), hi_i = max(ia, ib);
    const float hi = __int_as_float(hi_i), lo = __int_as_float(lo_i);
    const int e = hi_i & (int) 0xFE000000;
    const float scale = __int_as_float(e ^ 0x7E800000);
    const float ls = lo * scale, hs = hi * scale;
    const float t = fmaf(hs, hs, ls * ls);
    float rs;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(t));
    const float q = t * rs, h = rs * 0.5f;
    float r = fmaf(fmaf(-q, q, t), h, q) * __int_as_float(e | 0x00800000);
    r = lo_i == 0 ? hi : r;
    return r;
}

template <int MODE>
__global__ void k(long long* cycles, double* sink, int links, double a, double b)
{
    __shared__ int ring[64];
    for(int i = threadIdx.x; i < 64; i += 32) ring[i] = (i + 1) & 63;
    __syncwarp();
    double v = 1.0 + threadIdx.x * 1e-6;
    float f = 1.0f + threadIdx.x * 1e-3f;
    int n = threadIdx.x;
    const long long t0 = clock64();
    #pragma unroll 16
    for(int i = 0; i < links; i++)
    {
        if(MODE == DFMA) v = fma(v, a, b);
        if(MODE == DADD) v = v + b;
        if(MODE == DMUL) v = v * a;
        if(MODE == FFMA) f = fmaf(f, 1.0000001f, 1e-9f);
        if(MODE == FMUL) f = f * 1.0000001f;
        if(MODE == F2F_DF) { f = (float) v; v = (double) f; }                       // 2 links: F2F.F32.F64 then F2F.F64.F32
        if(MODE == F2F_ROUND) { f = (float) v; f = f * 1.0000001f; v = (double) f; v = v + b; }   // D->F, FMUL, F->D, DADD
        if(MODE == RSQ) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(f));
        if(MODE == IADD) n = n * 3 + 1;
        if(MODE == LDS_CHAIN) n = ring[n & 63];
        if(MODE == HYPOT) f = hypot_chain(f, 0.75f) * 1.0000001f;
    }
    const long long t1 = clock64();
    if(threadIdx.x == 0) cycles[0] = t1 - t0;
    sink[threadIdx.x] = v + f + n;
}

template <int MODE>
static void run(const char* name, int ops_per_link, long long* dc, double* ds)
{
    const int links = 4096;
    k<MODE><<<1, 32>>>(dc, ds, 64, 1.0000001, 1e-9);
    k<MODE><<<1, 32>>>(dc, ds, links, 1.0000001, 1e-9);
    long long c; CK(cudaMemcpy(&c, dc, sizeof c, cudaMemcpyDeviceToHost));
    printf("%-44s %7.1f cycles per link (%d dependent instruction%s per link)\n", name, (double) c / links, ops_per_link, ops_per_link > 1 ? "s" : "");
}

int main()
{
    long long* dc; double* ds; CK(cudaMalloc(&dc, 8)); CK(cudaMalloc(&ds, 32 * 8));
    run<DFMA>("DFMA", 1, dc, ds);
    run<DADD>("DADD", 1, dc, ds);
    run<DMUL>("DMUL", 1, dc, ds);
    run<FFMA>("FFMA", 1, dc, ds);
    run<FMUL>("FMUL", 1, dc, ds);
    run<IADD>("IMAD (n*3+1)", 1, dc, ds);
    run<F2F_DF>("F2F.F32.F64 + F2F.F64.F32", 2, dc, ds);
    run<F2F_ROUND>("F2F.F32.F64 + FMUL + F2F.F64.F32 + DADD", 4, dc, ds);
    run<RSQ>("MUFU.RSQ", 1, dc, ds);
    run<LDS_CHAIN>("LDS (pointer chase)", 1, dc, ds);
    run<HYPOT>("branch-free hypotf + FMUL", 1, dc, ds);
    return 0;
}
