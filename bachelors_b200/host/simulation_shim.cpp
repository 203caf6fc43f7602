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

// One solver handle per grid size; the reference's maps of one run all share a size (main.cpp:164-168).
pf_solver* solver_for(int nx, int ny)
{
    for(int i = 0; i < g_solver_count; i++)
        if(g_solvers[i].nx == nx && g_solvers[i].ny == ny)
            return g_solvers[i].s;
    if(g_solver_count == SHIM_MAX_SOLVERS)
    {
        // recycle the oldest handle (its fields were freed by sim_realloc(..., 0, 0) between runs, main.cpp:568-571)
        pf_destroy(g_solvers[0].s);
        memmove(g_solvers, g_solvers + 1, sizeof(Shim_Solver) * (SHIM_MAX_SOLVERS - 1));
        g_solver_count--;
    }
    int device = 0;
    if(const char* env = getenv("PF_DEVICE")) device = atoi(env);
    pf_solver* s = NULL;
    shim_check(NULL, pf_create(&s, device, SHIM_DTYPE, nx, ny), "pf_create");
    g_solvers[g_solver_count++] = Shim_Solver{s, nx, ny};
    return s;
}

pf_solver* any_solver()
{
    return g_solver_count > 0 ? g_solvers[g_solver_count - 1].s : solver_for(1, 1);
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
            shim_check(any_solver(), pf_field_free(any_solver(), map->data), "pf_field_free");
            map->data = NULL;
        }
        if((size_t) nx * (size_t) ny > 0)
        {
            pf_solver* s = solver_for(nx, ny);
            void* dev = NULL;
            shim_check(s, pf_field_alloc(s, &dev), "pf_field_alloc");
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
    pf_solver* s = any_solver();
    if(modify == MODIFY_UPLOAD) shim_check(s, pf_upload_raw(s, device_memory, host_memory, size), "pf_upload_raw");
    else                        shim_check(s, pf_download_raw(s, device_memory, host_memory, size), "pf_download_raw");
}

extern "C" void sim_modify_float(Real* device_memory, float* host_memory, size_t count, Sim_Modify modify)
{
    pf_solver* s = any_solver();
    if(modify == MODIFY_UPLOAD) shim_check(s, pf_upload(s, device_memory, host_memory, count, PF_F32), "pf_upload");
    else                        shim_check(s, pf_download(s, device_memory, host_memory, count, PF_F32), "pf_download");
}

extern "C" void sim_modify_double(Real* device_memory, double* host_memory, size_t count, Sim_Modify modify)
{
    pf_solver* s = any_solver();
    if(modify == MODIFY_UPLOAD) shim_check(s, pf_upload(s, device_memory, host_memory, count, PF_F64), "pf_upload");
    else                        shim_check(s, pf_download(s, device_memory, host_memory, count, PF_F64), "pf_download");
}

// The reference compiles its GPU self-tests into the binary (simulation.cu:1368-1380); this repository's tests live in
// tests/ (pytest, CPU oracle + golden vectors of the reference). The entry points stay so that main.cpp links.
extern "C" bool run_tests()
{
    fprintf(stderr, "run_tests: the B200 solver is tested by `python -m pytest tests -m gpu` in its own repository\n");
    return true;
}

extern "C" bool run_benchmarks(int N)
{
    (void) N;
    fprintf(stderr, "run_benchmarks: use `python bench.py` of the B200 solver repository\n");
    return true;
}
