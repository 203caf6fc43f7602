// simulation_shim.cpp - the reference's solver interface (reference simulation.h:132-145) implemented on the
// B200-native C-ABI (include/pf_solver.h). This file REPLACES the reference's simulation.cu in its build:
//
//     g++ -std=c++17 -O2 -I<reference checkout> -I<this repo>/include -c simulation_shim.cpp -o simulation.o
//     g++ main.cpp simulation.o -L<this repo>/bachelors_b200 -lpfsolver -lcudart ...          (reference Makefile:24-28)
//
// It includes the reference's own, unmodified simulation.h (found through -I), so Sim_Map / Sim_Params / Sim_Stats and the
// seven entry points keep their exact layout and linkage: main.cpp, config.h, gl.h and plot.py work unchanged.
// Error policy is the reference's: a failing device call is fatal (reference cuda_util.cuh:23-41) - the library itself
// only returns status codes; the abort lives here, on the reference's side of the boundary.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdbool.h>
#include <stdint.h>

#include "simulation.h"     // the reference's header
#include "log.h"            // the reference's logger (declarations; main.cpp holds the implementation, main.cpp:1)
#include "pf_solver.h"

namespace {

struct Shim_Solver { pf_solver* s; int nx, ny; };
enum { SHIM_MAX_SOLVERS = 8 };
Shim_Solver g_solvers[SHIM_MAX_SOLVERS];
int g_solver_count = 0;

const int SHIM_DTYPE = sizeof(Real) == 8 ? PF_F64 : PF_F32;

void shim_check(pf_solver* s, int rc, const char* what)
{
    if(rc == PF_OK) return;
    fprintf(stderr, "FATAL [pf_solver] %s failed with %d: %s\n", what, rc, pf_last_error(s));
    abort();
}

// Which handle allocated a device pointer: sim_realloc frees and sim_modify* copies through the OWNING handle (its stream, its
// staging buffer, its device), whatever other grid sizes are live.
struct Shim_Alloc { void* ptr; pf_solver* s; };
enum { SHIM_MAX_ALLOCS = 256 };
Shim_Alloc g_allocs[SHIM_MAX_ALLOCS];
int g_alloc_count = 0;

void remember_alloc(void* ptr, pf_solver* s)
{
    if(g_alloc_count == SHIM_MAX_ALLOCS) { fprintf(stderr, "FATAL [pf_solver] more than %d live maps\n", (int) SHIM_MAX_ALLOCS); abort(); }
    g_allocs[g_alloc_count++] = Shim_Alloc{ptr, s};
}

// Owner of the allocation that CONTAINS ptr (sim_modify may be handed an interior pointer); field sizes come from the handle.
pf_solver* owner_of(const void* ptr, bool forget)
{
    for(int i = 0; i < g_alloc_count; i++)
    {
        int rows = 0;
        pf_field_rows(g_allocs[i].s, &rows, NULL);
        const char* base = (const char*) g_allocs[i].ptr;
        size_t bytes = 0;
        for(int k = 0; k < g_solver_count; k++)
            if(g_solvers[k].s == g_allocs[i].s) bytes = (size_t) g_solvers[k].nx * (size_t) rows * sizeof(Real);
        if((const char*) ptr >= base && (const char*) ptr < base + (bytes ? bytes : 1))
        {
            pf_solver* s = g_allocs[i].s;
            if(forget) g_allocs[i] = g_allocs[--g_alloc_count];
            return s;
        }
    }
    fprintf(stderr, "FATAL [pf_solver] %p is not device memory of a live Sim_Map\n", ptr);
    abort();
}

bool solver_has_allocs(pf_solver* s)
{
    for(int i = 0; i < g_alloc_count; i++) if(g_allocs[i].s == s) return true;
    return false;
}

// One solver handle per grid size; the reference's maps of one run all share a size (main.cpp:164-168).
pf_solver* solver_for(int nx, int ny)
{
    for(int i = 0; i < g_solver_count; i++)
        if(g_solvers[i].nx == nx && g_solvers[i].ny == ny)
            return g_solvers[i].s;
    if(g_solver_count == SHIM_MAX_SOLVERS)
    {
        // recycle the oldest handle without live maps (sim_realloc(..., 0, 0) frees them between runs, main.cpp:568-571)
        int victim = -1;
        for(int i = 0; i < g_solver_count && victim < 0; i++) if(!solver_has_allocs(g_solvers[i].s)) victim = i;
        if(victim < 0) { fprintf(stderr, "FATAL [pf_solver] %d grid sizes with live maps\n", (int) SHIM_MAX_SOLVERS); abort(); }
        pf_destroy(g_solvers[victim].s);
        memmove(g_solvers + victim, g_solvers + victim + 1, sizeof(Shim_Solver) * (size_t)(g_solver_count - victim - 1));
        g_solver_count--;
    }
    int device = 0;
    if(const char* env = getenv("PF_DEVICE")) device = atoi(env);
    pf_solver* s = NULL;
    shim_check(NULL, pf_create(&s, device, SHIM_DTYPE, nx, ny), "pf_create");
    g_solvers[g_solver_count++] = Shim_Solver{s, nx, ny};
    return s;
}

pf_params to_pf_params(const Sim_Params& q)
{
    pf_params p;
    memset(&p, 0, sizeof p);
    p.solver = (int32_t) q.solver;
    p.nx = q.nx; p.ny = q.ny; p.time = q.time; p.iter = q.iter;
    p.T_boundary = (int32_t) q.T_boundary; p.Phi_boundary = (int32_t) q.Phi_boundary;
    p.L0 = q.L0; p.dt = q.dt; p.L = q.L; p.xi = q.xi; p.a = q.a; p.b = q.b; p.alpha = q.alpha; p.beta = q.beta;
    p.gamma = q.gamma; p.Tm = q.Tm; p.min_dt = q.min_dt; p.S = q.S; p.m0 = q.m0; p.theta0 = q.theta0;
    p.T_tolerance = q.T_tolerance; p.Phi_tolerance = q.Phi_tolerance; p.corrector_tolerance = q.corrector_tolerance;
    p.T_max_iters = q.T_max_iters; p.Phi_max_iters = q.Phi_max_iters; p.corrector_max_iters = q.corrector_max_iters;
    p.do_corrector_loop = q.do_corrector_loop; p.do_corrector_guess = q.do_corrector_guess;
    p.do_stats = q.do_stats && q.stats != NULL;
    p.do_stats_step_residual = q.do_stats_step_residual && q.stats != NULL;
    p.do_exact = q.do_exact;
    return p;
}

// _find_claim_temp_map: reference simulation.cu:103-127.
Sim_Map* claim_temp_map(const Sim_Params& params, const char* name, int nx, int ny, double time, i64 iter)
{
    Sim_Map* not_used = NULL;
    Sim_Map* not_used_with_size = NULL;
    for(int i = 0; i < params.temp_map_count; i++)
    {
        Sim_Map* curr = &params.temp_maps[i];
        if(curr->iter == iter) continue;
        if(!not_used) not_used = curr;
        if(curr->nx == nx) { not_used_with_size = curr; break; }
    }
    Sim_Map* out = not_used_with_size ? not_used_with_size : not_used;
    if(out) sim_realloc(out, name, nx, ny, time, iter);
    return out;
}

} // namespace

// reference simulation.cu:88-101 (a size of 0 frees the map; new storage is zero-filled)
void sim_realloc(Sim_Map* map, const char* name, int nx, int ny, double time, i64 iter)
{
    if(map->nx != nx || map->ny != ny)
    {
        if(map->data)
        {
            pf_solver* owner = owner_of(map->data, true);
            shim_check(owner, pf_field_free(owner, map->data), "pf_field_free");
            map->data = NULL;
        }
        if((size_t) nx * (size_t) ny > 0)
        {
            pf_solver* s = solver_for(nx, ny);
            void* dev = NULL;
            shim_check(s, pf_field_alloc(s, &dev), "pf_field_alloc");
            remember_alloc(dev, s);
            map->data = (Real*) dev;
        }
        map->nx = nx;
        map->ny = ny;
    }
    map->iter = iter;
    map->time = time;
    strncpy(map->name, name, sizeof map->name - 1);
    map->name[sizeof map->name - 1] = 0;
}

// reference simulation.cu:1091-1156
double sim_step(Sim_Map F, Sim_Map U, Sim_Map* next_F, Sim_Map* next_U, Sim_Params params)
{
    if(F.nx != params.nx || F.ny != params.ny || U.nx != params.nx || U.ny != params.ny)
    {
        fprintf(stderr, "ERROR [SOLVER] Bad sizes provided F:%ix%i U:%ix%i\n", F.nx, F.ny, U.nx, U.ny);
        return 0;
    }
    sim_realloc(next_F, "F", params.nx, params.ny, params.time, params.iter + 1);
    sim_realloc(next_U, "U", params.nx, params.ny, params.time, params.iter + 1);

    pf_solver* s = solver_for(params.nx, params.ny);
    pf_params p = to_pf_params(params);
    pf_step_info info;
    shim_check(s, pf_step(s, &p, F.data, U.data, next_F->data, next_U->data, &info), "pf_step");

    if(params.do_prints)                                                  // the console lines of the reference's solvers
    {
        if(params.solver == SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE)            // simulation.cu:484-485
            LOG("SOLVER", info.Phi_converged ? LOG_DEBUG : LOG_WARN, "rk4-adaptive %s in %i iters with error F:%e | U:%e | tau:%e",
                info.Phi_converged ? "converged" : "diverged", (int) info.Phi_iters, info.Phi_error, info.T_error, info.dt_advanced);
        if(params.solver == SOLVER_TYPE_SEMI_IMPLICIT)                    // simulation.cu:889-890, :907-908
        {
            LOG_DEBUG("SOLVER", "%lli F %s in %i iters with error %e\n", (long long) params.iter, info.Phi_converged ? "converged" : "diverged", (int) info.Phi_iters, info.Phi_error);
            LOG_DEBUG("SOLVER", "%lli U %s in %i iters with error %e\n", (long long) params.iter, info.T_converged ? "converged" : "diverged", (int) info.T_iters, info.T_error);
        }
    }

    if(params.stats)
    {
        if(params.solver == SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE)
        {
            params.stats->Phi_iters = info.Phi_iters;                    // simulation.cu:488-492
            params.stats->T_iters = (float) info.T_iters;
        }
        if(params.do_stats)
        {
            params.stats->iter = params.iter;                            // simulation.cu:1126-1142
            params.stats->time = params.time;
            params.stats->Phi_delta_L1 = (float) info.Phi_delta.L1;  params.stats->Phi_delta_L2 = (float) info.Phi_delta.L2;
            params.stats->Phi_delta_min = (float) info.Phi_delta.min; params.stats->Phi_delta_max = (float) info.Phi_delta.max;
            params.stats->T_delta_L1 = (float) info.T_delta.L1;      params.stats->T_delta_L2 = (float) info.T_delta.L2;
            params.stats->T_delta_min = (float) info.T_delta.min;     params.stats->T_delta_max = (float) info.T_delta.max;
        }
        if(params.do_stats_step_residual && info.step_res_count > 0)      // simulation.cu:979-993
        {
            int n = info.step_res_count < MAX_STEP_RESIDUALS ? info.step_res_count : MAX_STEP_RESIDUALS;
            for(int k = 0; k < n; k++)
            {
                params.stats->step_res_L1[k] = (float) info.step_res[k].L1;   params.stats->step_res_L2[k] = (float) info.step_res[k].L2;
                params.stats->step_res_min[k] = (float) info.step_res[k].min; params.stats->step_res_max[k] = (float) info.step_res[k].max;
            }
            params.stats->step_res_count = info.step_res_count;
        }
    }

    if(params.do_debug)                                                   // simulation.cu:1144-1151
    {
        Sim_Map* grad_Phi = claim_temp_map(params, "grad_Phi", params.nx, params.ny, params.time, params.iter + 1);
        Sim_Map* grad_T = claim_temp_map(params, "grad_T", params.nx, params.ny, params.time, params.iter + 1);
        Sim_Map* aniso = claim_temp_map(params, "aniso", params.nx, params.ny, params.time, params.iter + 1);
        if(grad_Phi && grad_T && aniso)
            shim_check(s, pf_debug_values(s, &p, next_F->data, next_U->data, grad_Phi->data, grad_T->data, aniso->data), "pf_debug_values");
    }
    return info.dt_advanced;
}

// reference simulation.cu:1051-1067
extern "C" void sim_modify(void* device_memory, void* host_memory, size_t size, Sim_Modify modify)
{
    pf_solver* s = owner_of(device_memory, false);
    if(modify == MODIFY_UPLOAD) shim_check(s, pf_upload_raw(s, device_memory, host_memory, size), "pf_upload_raw");
    else                        shim_check(s, pf_download_raw(s, device_memory, host_memory, size), "pf_download_raw");
}

extern "C" void sim_modify_float(Real* device_memory, float* host_memory, size_t count, Sim_Modify modify)
{
    pf_solver* s = owner_of(device_memory, false);
    if(modify == MODIFY_UPLOAD) shim_check(s, pf_upload(s, device_memory, host_memory, count, PF_F32), "pf_upload");
    else                        shim_check(s, pf_download(s, device_memory, host_memory, count, PF_F32), "pf_download");
}

extern "C" void sim_modify_double(Real* device_memory, double* host_memory, size_t count, Sim_Modify modify)
{
    pf_solver* s = owner_of(device_memory, false);
    if(modify == MODIFY_UPLOAD) shim_check(s, pf_upload(s, device_memory, host_memory, count, PF_F64), "pf_upload");
    else                        shim_check(s, pf_download(s, device_memory, host_memory, count, PF_F64), "pf_download");
}

// The reference compiles self-tests of its generic launch / reduction helpers into the binary (simulation.cu:1368-1380); those
// helpers do not exist here (this repository's tests are tests/, pytest against the CPU oracle and the reference's golden
// vectors). What can be checked from inside the host process is checked: the library answers, a step runs, and the two
// reductions reduce a known sequence to the known answers.
extern "C" bool run_tests()
{
    bool ok = pf_abi_version() == PF_ABI_VERSION && pf_device_count() > 0;
    if(ok)
    {
        pf_solver* s = solver_for(64, 64);
        double t_max = 0, t_stats = 0, mx = -1;
        ok = pf_bench_reductions(s, 64 * 64, SHIM_DTYPE, 1, &t_max, &t_stats, &mx) == PF_OK && mx > 0.99 && mx < 1.0;   // max of 4096 uniform [0,1) values
    }
    LOG("SOLVER", ok ? LOG_INFO : LOG_ERROR, "run_tests: %s (the full suite is `python -m pytest tests -m gpu` in the solver's repository)", ok ? "OK" : "FAILED");
    return ok;
}

// reference simulation.cu:1300-1358: median time and GiB/s of the max-reduction at six sizes (its CPU and thrust legs have no
// counterpart here), printed in the reference's format; plus the five-statistics reduction of pf_step's do_stats.
extern "C" bool run_benchmarks(int N_)
{
    (void) N_;
    const size_t Ns[] = {256 * 256, 512 * 512, 1024 * 1024, 2048 * 2048, 4096 * 4096, 2 * 4096 * 4096};
    const int count = (int)(sizeof Ns / sizeof *Ns);
    double t_max[8] = {0}, t_stats[8] = {0};
    const double GiB = 1024.0 * 1024.0 * 1024.0;
    pf_solver* s = solver_for(64, 64);
    for(int i = 0; i < count; i++)
    {
        double mx = 0;
        shim_check(s, pf_bench_reductions(s, Ns[i], PF_F32, 200, &t_max[i], &t_stats[i], &mx), "pf_bench_reductions");
        printf("\n\nN:%i\nCUST: time: %e s | %4.2lf GB/s   (max |x| of %zu floats)\nSTAT: time: %e s | %4.2lf GB/s   (sum, L1, L2, min, max of a - b, 2 x %zu floats)\n",
               (int) Ns[i], t_max[i], Ns[i] * 4 / GiB / t_max[i], Ns[i], t_stats[i], 2 * Ns[i] * 4 / GiB / t_stats[i], Ns[i]);
    }
    printf("cust_f32 = [");
    for(int i = 0; i < count; i++) printf("(%e, %4.2lf), ", t_max[i], Ns[i] * 4 / GiB / t_max[i]);
    printf("]\nstats_f32 = [");
    for(int i = 0; i < count; i++) printf("(%e, %4.2lf), ", t_stats[i], 2 * Ns[i] * 4 / GiB / t_stats[i]);
    printf("]\n");
    return true;
}
