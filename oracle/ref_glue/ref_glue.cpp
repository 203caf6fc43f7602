// Glue that turns the UNMODIFIED reference solver (compiled from /root/reference where it lies)
// into a ctypes-loadable shared library: oracle/_ref/libref_sim_{f64,f32}.so ("oracle A").
// Test infrastructure only. Nothing in the product path links or loads this.
//
// The reference's simulation.cu needs the logger implementation from its own log.h
// (JOT_ALL_IMPL, reference main.cpp:1) and exposes sim_realloc/sim_step with C++ linkage
// (reference simulation.h:132-133); ctypes needs C linkage, hence the wrappers.
#include <stdlib.h>
#include <stdbool.h>
#define JOT_ALL_IMPL
#include "log.h"
#include "simulation.h"
#include <cuda_runtime.h>

extern "C" {

int ref_sizeof_real(void) { return (int) sizeof(Real); }
int ref_sizeof_params(void) { return (int) sizeof(Sim_Params); }
int ref_sizeof_map(void) { return (int) sizeof(Sim_Map); }
int ref_sizeof_stats(void) { return (int) sizeof(Sim_Stats); }

void ref_sim_realloc(Sim_Map* map, const char* name, int nx, int ny, double time, long long iter)
{
    sim_realloc(map, name, nx, ny, time, (i64) iter);
}

double ref_sim_step(const Sim_Map* F, const Sim_Map* U, Sim_Map* next_F, Sim_Map* next_U, const Sim_Params* params)
{
    return sim_step(*F, *U, next_F, next_U, *params);
}

// n steps with the reference host loop's ping-pong (reference main.cpp:542-558). Returns summed dt.
double ref_sim_run(Sim_Map* F, Sim_Map* U, Sim_Map* next_F, Sim_Map* next_U, const Sim_Params* params, int n_steps, int* iters_done)
{
    double t = params->time;
    Sim_Params p = *params;
    int i = 0;
    for(; i < n_steps; i++)
    {
        p.iter = params->iter + i;
        p.time = t;
        t += sim_step(*F, *U, next_F, next_U, p);
        Sim_Map tmp = *F; *F = *next_F; *next_F = tmp;
        tmp = *U; *U = *next_U; *next_U = tmp;
    }
    if(iters_done) *iters_done = i;
    return t - params->time;
}

// Timed variant for the same-box GPU baseline: CUDA events around n sim_step calls.
double ref_sim_time_ms(Sim_Map* F, Sim_Map* U, Sim_Map* next_F, Sim_Map* next_U, const Sim_Params* params, int n_steps)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, 0);
    ref_sim_run(F, U, next_F, next_U, params, n_steps, NULL);
    cudaEventRecord(e1, 0);
    cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return (double) ms;
}

void ref_quiet_logs(void)
{
    // keep only warnings and worse on the console
    log_system_set_logger(NULL);
}

}
