// kbench.cu - development harness for the explicit step kernels (not part of the product library).
//
// Runs the tiled kernel (pf_fused.cuh, already pinned against the oracle by tests/) and variants of the streaming
// kernel (pf_stream.cuh) on the same inputs, compares the outputs bit for bit and times each with CUDA events.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -ffp-contract=off \
//        tools/kbench.cu -o gpurun_out/kbench
//   kbench [nx ny reps]
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <vector>
#include <string>
#include <algorithm>

#include "../bachelors_b200/csrc/pf_common.cuh"
#include "../bachelors_b200/csrc/pf_fused.cuh"
#include "kbench_tiled.cuh"
#include "../bachelors_b200/csrc/pf_stream.cuh"
#include "../bachelors_b200/csrc/pf_euler.cuh"

using namespace pf;

#define CK(x) do { cudaError_t e_ = (x); if(e_ != cudaSuccess) { fprintf(stderr, "%s: %s (%s:%d)\n", #x, cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while(0)

static pf_params bench_params(int nx, int ny, int bc, double S)
{
    pf_params p; memset(&p, 0, sizeof p);
    p.nx = nx; p.ny = ny; p.L0 = 4.0 * nx / 512.0;
    p.dt = 5e-6; p.L = 2; p.xi = 0.00411; p.a = 2; p.b = 1; p.alpha = 3; p.beta = 1400; p.Tm = 1;
    p.S = S; p.m0 = 4; p.theta0 = 0; p.gamma = 1;
    p.Phi_boundary = bc; p.T_boundary = bc;
    return p;
}

static uint64_t splitmix(uint64_t& s) { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }

template <typename Real>
static void make_fields(int kind, int nx, int ny, std::vector<Real>& F, std::vector<Real>& U)
{
    F.resize((size_t) nx * ny); U.resize((size_t) nx * ny);
    uint64_t seed = 0x5EED5EEDull;
    for(int y = 0; y < ny; y++) for(int x = 0; x < nx; x++)
    {
        size_t i = (size_t) y * nx + x;
        double fx = (x + 0.5) / nx, fy = (y + 0.5) / ny;
        if(kind == 0)      { double r = hypot((fx - 0.5) * nx, (fy - 0.5) * ny); F[i] = r < 0.0125 * nx ? 1 : 0; U[i] = 0; }      // dendrite seed: flat far field
        else if(kind == 1) { F[i] = (Real)(0.5 + 0.4 * sin(2 * M_PI * 8 * fx) * sin(2 * M_PI * 8 * fy)); U[i] = (Real)(0.5 * cos(2 * M_PI * 4 * fx)); }
        else               { F[i] = (Real)((splitmix(seed) >> 11) * (1.0 / 9007199254740992.0)); U[i] = (Real)((splitmix(seed) >> 11) * (1.0 / 9007199254740992.0)); }
    }
    if(kind == 2) { F[5] = -(Real) 0; U[nx + 7] = -(Real) 0; F[(size_t) nx * ny - 3] = -(Real) 0; }
}

template <typename Real>
static FusedArgs<Real> make_args(const pf_params& p, const Real* F, const Real* U, Real* oF, Real* oU, int scheme, unsigned long long* err)
{
    FusedArgs<Real> a; memset(&a, 0, sizeof a);
    a.F = F; a.U = U; a.out_F = oF; a.out_U = oU;
    a.nx = p.nx; a.ny_global = p.ny; a.row0 = 0; a.rows_local = p.ny; a.own_y0 = 0; a.own_y1 = p.ny;
    a.bc = p.Phi_boundary; a.bc_T = p.T_boundary; a.U_base = nullptr; a.wrap_y = p.Phi_boundary == PF_BOUNDARY_PERIODIC;
    a.k = make_consts<Real>(p);
    Real dt = (Real) p.dt;
    if(scheme == SCHEME_RK4) { a.w[0] = dt / 2; a.w[1] = dt / 2; a.w[2] = dt; a.w[3] = dt / 6; }
    if(scheme == SCHEME_RKM) { Real tau = dt; a.w[0] = tau / 3; a.w[1] = tau / 6; a.w[2] = tau / 8; a.w[3] = tau * 3 / 8; a.w[4] = tau / 2; a.w[5] = -tau * 3 / 2; a.w[6] = tau * 2; a.w[7] = tau / 6; }
    a.err_bits = err;
    return a;
}

template <typename Real, int SCHEME, bool ANISO>
static void launch_tiled(const FusedArgs<Real>& a)
{
    constexpr int EW = 64, TY = 8, R = SCHEME == SCHEME_RKM ? 4 : 5;
    constexpr int NST = SchemeStages<SCHEME>::value, EH = TY * R, TW = EW - 2 * NST, TH = EH - 2 * NST;
    using V2 = typename Vec2T<Real>::type;
    const size_t smem = sizeof(V2) * ((size_t) EW * EH + 2) * 2;
    auto kern = k_fused_step<Real, SCHEME, ANISO, EW, TY, R>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    dim3 grid((a.nx + TW - 1) / TW, (a.own_y1 - a.own_y0 + TH - 1) / TH), block(EW, TY);
    kern<<<grid, block, smem>>>(a);
}

template <typename Real, int SCHEME, bool ANISO, int WARPS, int MINB>
static void launch_stream(FusedArgs<Real> a, int nch)
{
    constexpr int NST = SchemeStages<SCHEME>::value, EWI = 32 * WARPS - 2 * NST;
    const size_t smem = stream_smem_bytes<Real, SCHEME, WARPS>();
    auto kern = k_stream_step<Real, SCHEME, ANISO, WARPS, MINB>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    a.strips = stream_strips(a.nx, EWI, NST, a.bc == PF_BOUNDARY_PERIODIC);
    a.nch = std::min(nch, a.own_y1 - a.own_y0);
    kern<<<a.strips * a.nch, 32 * WARPS, smem>>>(a);
}

// the Euler tile kernel (pf_euler.cuh): EW threads, R rows per tile, MINB CTAs per SM; the chunk argument is unused
template <typename Real, bool ANISO, int EW, int R, int MINB>
static void launch_euler_tile(FusedArgs<Real> a, int)
{
    const size_t smem = euler_tile_smem_bytes<Real, EW, R>();
    auto kern = k_euler_tile<Real, ANISO, false, EW, R, MINB>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    dim3 grid((a.nx + (EW - 2 * EULER_HALO) - 1) / (EW - 2 * EULER_HALO), (a.own_y1 - a.own_y0 + R - 1) / R);
    kern<<<grid, EW, smem>>>(a);
}
#define VE(Real, ANISO, EW, R, MB) Variant<Real>{ "euler tile ew" #EW " r" #R " b" #MB, launch_euler_tile<Real, ANISO, EW, R, MB>, 0 }

template <typename Real>
struct Variant { std::string name; void (*fn)(FusedArgs<Real>, int); int chunk; };

template <typename Real, int SCHEME, bool ANISO>
static void run_scheme(const char* sname, int nx, int ny, int reps, const std::vector<Variant<Real>>& variants, const std::vector<int>& bcs, const std::vector<int>& kinds)
{
    const size_t N = (size_t) nx * ny, bytes = N * sizeof(Real);
    Real *dF, *dU, *rF, *rU, *oF, *oU; unsigned long long* derr;
    CK(cudaMalloc(&dF, bytes)); CK(cudaMalloc(&dU, bytes)); CK(cudaMalloc(&rF, bytes)); CK(cudaMalloc(&rU, bytes)); CK(cudaMalloc(&oF, bytes)); CK(cudaMalloc(&oU, bytes));
    CK(cudaMalloc(&derr, 32));
    std::vector<Real> hF, hU, refF(N), refU(N), gotF(N), gotU(N);
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const char* kind_names[3] = {"seed", "smooth", "noise"};
    for(int bc : bcs) for(int kind : kinds)
    {
        pf_params p = bench_params(nx, ny, bc, ANISO ? 0.3 : 0.0);
        make_fields<Real>(kind, nx, ny, hF, hU);
        CK(cudaMemcpy(dF, hF.data(), bytes, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dU, hU.data(), bytes, cudaMemcpyHostToDevice));
        // reference: the tiled kernel
        unsigned long long ref_err[2] = {0, 0}, got_err[2];
        CK(cudaMemset(derr, 0, 16));
        FusedArgs<Real> a = make_args<Real>(p, dF, dU, rF, rU, SCHEME, derr);
        launch_tiled<Real, SCHEME, ANISO>(a);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(ref_err, derr, 16, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(refF.data(), rF, bytes, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(refU.data(), rU, bytes, cudaMemcpyDeviceToHost));
        {
            // time the tiled kernel ping-ponging (F,U) <-> (o) like consecutive steps
            FusedArgs<Real> b = make_args<Real>(p, rF, rU, oF, oU, SCHEME, derr);
            for(int i = 0; i < 3; i++) { launch_tiled<Real, SCHEME, ANISO>(a); launch_tiled<Real, SCHEME, ANISO>(b); }
            CK(cudaEventRecord(e0));
            for(int i = 0; i < reps; i++) { launch_tiled<Real, SCHEME, ANISO>(a); launch_tiled<Real, SCHEME, ANISO>(b); }
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            double us = ms * 1e3 / (2 * reps);
            printf("%-8s %-4s bc=%d %-6s %-26s %9.2f us %8.2f GLUPS\n", sname, sizeof(Real) == 8 ? "f64" : "f32", bc, kind_names[kind], "tiled(ref)", us, N / us / 1e3);
            launch_tiled<Real, SCHEME, ANISO>(a); CK(cudaDeviceSynchronize());   // restore rF/rU = step(dF,dU)
        }
        for(const auto& v : variants)
        {
            CK(cudaMemset(oF, 0xFF, bytes)); CK(cudaMemset(oU, 0xFF, bytes)); CK(cudaMemset(derr, 0, 16));
            FusedArgs<Real> g = make_args<Real>(p, dF, dU, oF, oU, SCHEME, derr);
            v.fn(g, v.chunk);
            cudaError_t le = cudaGetLastError();
            if(le != cudaSuccess) { printf("%-8s %-26s launch failed: %s\n", sname, v.name.c_str(), cudaGetErrorString(le)); continue; }
            CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(got_err, derr, 16, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(gotF.data(), oF, bytes, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(gotU.data(), oU, bytes, cudaMemcpyDeviceToHost));
            size_t bad = 0, first = (size_t) -1;
            for(size_t i = 0; i < N; i++)
                if(memcmp(&gotF[i], &refF[i], sizeof(Real)) || memcmp(&gotU[i], &refU[i], sizeof(Real))) { if(!bad) first = i; bad++; }
            bool err_ok = SCHEME != SCHEME_RKM || (got_err[0] == ref_err[0] && got_err[1] == ref_err[1]);
            FusedArgs<Real> g2 = make_args<Real>(p, oF, oU, rF, rU, SCHEME, derr);   // ping-pong partner (overwrites rF/rU: restored below)
            for(int i = 0; i < 3; i++) { v.fn(g, v.chunk); v.fn(g2, v.chunk); }
            CK(cudaEventRecord(e0));
            for(int i = 0; i < reps; i++) { v.fn(g, v.chunk); v.fn(g2, v.chunk); }
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            double us = ms * 1e3 / (2 * reps);
            printf("%-8s %-4s bc=%d %-6s %-26s %9.2f us %8.2f GLUPS  %s", sname, sizeof(Real) == 8 ? "f64" : "f32", bc, kind_names[kind], v.name.c_str(), us, N / us / 1e3,
                   bad == 0 && err_ok ? "BITEXACT" : "MISMATCH");
            if(bad) printf(" (%zu cells, first at x=%zu y=%zu: got %.17g / %.17g want %.17g / %.17g)", bad, first % nx, first / nx,
                           (double) gotF[first], (double) gotU[first], (double) refF[first], (double) refU[first]);
            if(!err_ok) printf(" (err bits %llx %llx want %llx %llx)", got_err[0], got_err[1], ref_err[0], ref_err[1]);
            printf("\n");
            fflush(stdout);
        }
    }
    cudaFree(dF); cudaFree(dU); cudaFree(rF); cudaFree(rU); cudaFree(oF); cudaFree(oU); cudaFree(derr);
}

// n_steps time steps as n_steps single-step launches and as ONE multi-step launch (StreamRun): same bits, time per step of both.
template <typename Real, int SCHEME, bool ANISO, int WARPS, int MINB>
static void run_multi(const char* sname, int nx, int ny, int n_steps, int nch, int bc, int kind, int reps)
{
    const size_t N = (size_t) nx * ny, bytes = N * sizeof(Real);
    Real *dF, *dU, *aF, *aU, *bF, *bU; unsigned* prog;
    CK(cudaMalloc(&dF, bytes)); CK(cudaMalloc(&dU, bytes)); CK(cudaMalloc(&aF, bytes)); CK(cudaMalloc(&aU, bytes)); CK(cudaMalloc(&bF, bytes)); CK(cudaMalloc(&bU, bytes));
    std::vector<Real> hF, hU, wF(N), wU(N), gF(N), gU(N);
    pf_params p = bench_params(nx, ny, bc, ANISO ? 0.3 : 0.0);
    make_fields<Real>(kind, nx, ny, hF, hU);
    constexpr int NST = SchemeStages<SCHEME>::value, EWI = 32 * WARPS - 2 * NST;
    const size_t smem = stream_smem_bytes<Real, SCHEME, WARPS>();
    auto kern = k_stream_step<Real, SCHEME, ANISO, WARPS, MINB>;
    auto kern_run = k_stream_run<Real, SCHEME, ANISO, WARPS, MINB>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    CK(cudaFuncSetAttribute(kern_run, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    CK(cudaFuncSetAttribute(kern_run, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    FusedArgs<Real> a = make_args<Real>(p, aF, aU, bF, bU, SCHEME, nullptr);
    a.strips = stream_strips(nx, EWI, NST, bc == PF_BOUNDARY_PERIODIC); a.nch = std::min(nch, ny);
    const int units = a.strips * a.nch;
    CK(cudaMalloc(&prog, (units + 1) * sizeof(unsigned))); CK(cudaMemset(prog, 0, (units + 1) * sizeof(unsigned)));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto reset = [&] { CK(cudaMemcpy(aF, hF.data(), bytes, cudaMemcpyHostToDevice)); CK(cudaMemcpy(aU, hU.data(), bytes, cudaMemcpyHostToDevice)); };
    auto single = [&](int n) { FusedArgs<Real> x = a;
        for(int i = 0; i < n; i++) { kern<<<units, 32 * WARPS, smem>>>(x); std::swap(const_cast<Real*&>(x.F), x.out_F); std::swap(const_cast<Real*&>(x.U), x.out_U); } };
    unsigned base = 1;
    auto multi = [&](int n) { StreamRun r; r.n_steps = n; r.base = base; r.progress = prog; r.units = units; base += n;
        FusedArgs<Real> x = a; void* args[] = {&x, &r};
        CK(cudaLaunchCooperativeKernel((const void*) kern_run, dim3(units), dim3(32 * WARPS), args, smem, 0)); };
    float ms1 = 0, msn = 0;
    reset(); single(n_steps); CK(cudaDeviceSynchronize());
    const Real* rF = (n_steps & 1) ? bF : aF; const Real* rU = (n_steps & 1) ? bU : aU;
    CK(cudaMemcpy(wF.data(), rF, bytes, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(wU.data(), rU, bytes, cudaMemcpyDeviceToHost));
    reset(); multi(n_steps); CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(gF.data(), rF, bytes, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(gU.data(), rU, bytes, cudaMemcpyDeviceToHost));
    unsigned stalled = 0; CK(cudaMemcpy(&stalled, prog + units, 4, cudaMemcpyDeviceToHost));
    size_t bad = 0;
    for(size_t i = 0; i < N; i++) if(memcmp(&gF[i], &wF[i], sizeof(Real)) || memcmp(&gU[i], &wU[i], sizeof(Real))) bad++;
    for(int rep = 0; rep < reps; rep++)
    {
        CK(cudaEventRecord(e0)); single(n_steps); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); float m; CK(cudaEventElapsedTime(&m, e0, e1)); ms1 = rep ? std::min(ms1, m) : m;
        CK(cudaEventRecord(e0)); multi(n_steps); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&m, e0, e1)); msn = rep ? std::min(msn, m) : m;
    }
    printf("%-6s %-4s %dx%d bc=%d kind=%d nch=%d units=%d steps=%d: per step %8.2f us (single launches) %8.2f us (one launch)  %s%s\n", sname, sizeof(Real) == 8 ? "f64" : "f32",
           nx, ny, bc, kind, a.nch, units, n_steps, ms1 * 1e3 / n_steps, msn * 1e3 / n_steps, bad ? "MISMATCH" : "BITEXACT", stalled ? " STALLED" : "");
    fflush(stdout);
    cudaFree(dF); cudaFree(dU); cudaFree(aF); cudaFree(aU); cudaFree(bF); cudaFree(bU); cudaFree(prog);
}

#define V(Real, SCHEME, ANISO, W, MB, CH) Variant<Real>{ "stream w" #W " b" #MB " nch" #CH, launch_stream<Real, SCHEME, ANISO, W, MB>, CH }

#ifdef KB_VARIANTS_FILE
#include KB_VARIANTS_FILE          // ad-hoc experiment lists: -DKB_VARIANTS_FILE='"/tmp/x.inc"'
#else
#include "kbench_variants.inc"
#endif

int main(int argc, char** argv)
{
    int nx = argc > 1 ? atoi(argv[1]) : 2048, ny = argc > 2 ? atoi(argv[2]) : 2048, reps = argc > 3 ? atoi(argv[3]) : 20;
    const char* mode = argc > 4 ? argv[4] : "all";
    run_all(nx, ny, reps, mode);
    return 0;
}
