// Glue that turns the UNMODIFIED reference solver (compiled from /root/reference where it lies)
// into a ctypes-loadable shared library: oracle/_ref/libref_sim_{f64,f32}.so ("oracle A").
// Test infrastructure only. Nothing in the product path links or loads this.
//
// The reference's simulation.cu needs the logger implementation from its own log.h
// (JOT_ALL_IMPL, reference main.cpp:1) and exposes sim_realloc/sim_step with C++ linkage
// (reference simulation.h:132-133); ctypes needs C linkage, hence the wrappers. The flat `ref_ctx_*`
// API below takes this repository's pf_params (include/pf_solver.h) and converts it to the reference's
// Sim_Params, so the Python side never mirrors a reference struct.
#include <stdlib.h>
#include <stdbool.h>
#include <stdio.h>
#include <string.h>
#define JOT_ALL_IMPL
#include "log.h"
#include "simulation.h"
#include <cuda_runtime.h>
#include "pf_solver.h"

// A logger that swallows everything but remembers the CG iteration counts the reference only logs
// (reference simulation.cu:890, :908: "%lli F converged in %i iters with error %e").
struct Capture_Logger {
    Logger base;
    int F_iters, U_iters, F_converged, U_converged;
    double F_error, U_error;
    int warnings;
};
static Capture_Logger g_capture;

static void capture_log(Logger* logger, const char* module, Log_Type type, size_t indentation, int line, const char* file,
                        const char* function, const char* format, va_list args)
{
    (void) indentation; (void) line; (void) file; (void) function;
    Capture_Logger* c = (Capture_Logger*) (void*) logger;
    if(type == LOG_WARN || type == LOG_ERROR || type == LOG_FATAL) c->warnings++;
    if(module == NULL || strcmp(module, "SOLVER") != 0) return;
    char buf[512];
    vsnprintf(buf, sizeof buf, format, args);
    long long iter = 0; char which = 0; char word[32] = {0}; int iters = 0; double err = 0;
    if(sscanf(buf, "%lld %c %31s in %d iters with error %lf", &iter, &which, word, &iters, &err) == 5)
    {
        int conv = strcmp(word, "converged") == 0;
        if(which == 'F') { c->F_iters = iters; c->F_converged = conv; c->F_error = err; }
        if(which == 'U') { c->U_iters = iters; c->U_converged = conv; c->U_error = err; }
    }
}

struct Ref_Ctx {
    Sim_Map maps[4];       // F, U, next_F, next_U
    int nx, ny;
};

static Sim_Params to_sim_params(const pf_params* p, Sim_Stats* stats)
{
    Sim_Params q;
    memset(&q, 0, sizeof q);
    q.solver = (Sim_Solver_Type) p->solver;
    q.nx = p->nx; q.ny = p->ny; q.time = p->time; q.iter = p->iter;
    q.T_boundary = (Sim_Boundary_Type) p->T_boundary;
    q.Phi_boundary = (Sim_Boundary_Type) p->Phi_boundary;
    q.L0 = p->L0; q.dt = p->dt; q.L = p->L; q.xi = p->xi; q.a = p->a; q.b = p->b; q.alpha = p->alpha; q.beta = p->beta;
    q.gamma = p->gamma; q.Tm = p->Tm; q.min_dt = p->min_dt; q.S = p->S; q.m0 = p->m0; q.theta0 = p->theta0;
    q.T_tolerance = p->T_tolerance; q.Phi_tolerance = p->Phi_tolerance; q.corrector_tolerance = p->corrector_tolerance;
    q.T_max_iters = p->T_max_iters; q.Phi_max_iters = p->Phi_max_iters; q.corrector_max_iters = p->corrector_max_iters;
    q.do_corrector_loop = p->do_corrector_loop != 0;
    q.do_corrector_guess = p->do_corrector_guess != 0;
    q.do_debug = false;
    q.do_stats = p->do_stats != 0;
    q.do_stats_step_residual = p->do_stats_step_residual != 0;
    q.do_exact = p->do_exact != 0;
    q.do_prints = true;          // so that the CG iteration counts reach the capture logger
    q.stats = stats;
    q.temp_maps = NULL; q.temp_map_count = 0;
    return q;
}

extern "C" {

int ref_sizeof_real(void) { return (int) sizeof(Real); }

void* ref_ctx_create(int nx, int ny)
{
    static bool logger_set = false;
    if(!logger_set)
    {
        memset(&g_capture, 0, sizeof g_capture);
        g_capture.base.log = capture_log;
        log_system_set_logger(&g_capture.base);
        logger_set = true;
    }
    Ref_Ctx* c = (Ref_Ctx*) calloc(1, sizeof(Ref_Ctx));
    c->nx = nx; c->ny = ny;
    const char* names[4] = {"F", "U", "next_F", "next_U"};
    for(int i = 0; i < 4; i++)
        sim_realloc(&c->maps[i], names[i], nx, ny, 0, 0);
    return c;
}

void ref_ctx_destroy(void* ctx)
{
    Ref_Ctx* c = (Ref_Ctx*) ctx;
    if(!c) return;
    for(int i = 0; i < 4; i++)
        sim_realloc(&c->maps[i], "", 0, 0, 0, 0);
    free(c);
}

// which: 0 = F, 1 = U. host_is_double selects sim_modify_double / sim_modify_float (reference simulation.cu:1042-1049).
void ref_ctx_upload(void* ctx, int which, const void* host, int host_is_double)
{
    Ref_Ctx* c = (Ref_Ctx*) ctx;
    size_t n = (size_t) c->nx * (size_t) c->ny;
    if(host_is_double) sim_modify_double(c->maps[which].data, (double*) host, n, MODIFY_UPLOAD);
    else               sim_modify_float(c->maps[which].data, (float*) host, n, MODIFY_UPLOAD);
}

void ref_ctx_download(void* ctx, int which, void* host, int host_is_double)
{
    Ref_Ctx* c = (Ref_Ctx*) ctx;
    size_t n = (size_t) c->nx * (size_t) c->ny;
    if(host_is_double) sim_modify_double(c->maps[which].data, (double*) host, n, MODIFY_DOWNLOAD);
    else               sim_modify_float(c->maps[which].data, (float*) host, n, MODIFY_DOWNLOAD);
}

// n steps with the reference host loop's ping-pong (reference main.cpp:542-558); the final state ends in maps F/U.
// Returns the summed dt. info (may be NULL) describes the last step: dt, CG iteration counts (from the log), stats.
double ref_ctx_run(void* ctx, const pf_params* params, long long n_steps, pf_step_info* info)
{
    Ref_Ctx* c = (Ref_Ctx*) ctx;
    Sim_Stats stats; memset(&stats, 0, sizeof stats);
    Sim_Params p = to_sim_params(params, &stats);
    double t = params->time, last_dt = 0;
    for(long long i = 0; i < n_steps; i++)
    {
        p.iter = params->iter + i;
        p.time = t;
        last_dt = sim_step(c->maps[0], c->maps[1], &c->maps[2], &c->maps[3], p);
        t += last_dt;
        Sim_Map tmp = c->maps[0]; c->maps[0] = c->maps[2]; c->maps[2] = tmp;
        tmp = c->maps[1]; c->maps[1] = c->maps[3]; c->maps[3] = tmp;
    }
    if(info)
    {
        memset(info, 0, sizeof *info);
        info->dt_advanced = last_dt;
        if(params->solver == PF_SOLVER_SEMI_IMPLICIT)
        {
            info->Phi_iters = g_capture.F_iters; info->T_iters = g_capture.U_iters;
            info->Phi_converged = g_capture.F_converged; info->T_converged = g_capture.U_converged;
            info->Phi_error = g_capture.F_error; info->T_error = g_capture.U_error;
        }
        else
        {
            info->Phi_iters = stats.Phi_iters; info->T_iters = (int) stats.T_iters;
        }
        info->Phi_delta.L1 = stats.Phi_delta_L1; info->Phi_delta.L2 = stats.Phi_delta_L2;
        info->Phi_delta.min = stats.Phi_delta_min; info->Phi_delta.max = stats.Phi_delta_max;
        info->T_delta.L1 = stats.T_delta_L1; info->T_delta.L2 = stats.T_delta_L2;
        info->T_delta.min = stats.T_delta_min; info->T_delta.max = stats.T_delta_max;
        info->step_res_count = stats.step_res_count;
        for(int k = 0; k < stats.step_res_count && k < PF_MAX_STEP_RESIDUALS; k++)
        {
            info->step_res[k].L1 = stats.step_res_L1[k]; info->step_res[k].L2 = stats.step_res_L2[k];
            info->step_res[k].min = stats.step_res_min[k]; info->step_res[k].max = stats.step_res_max[k];
        }
    }
    return t - params->time;
}

// Timed variant for the same-box GPU baseline ("the kernel to beat"): CUDA events around n sim_step calls
// on the legacy default stream the reference launches on. Returns milliseconds; *t_advanced the summed dt.
double ref_ctx_time_ms(void* ctx, const pf_params* params, long long n_steps, double* t_advanced)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaDeviceSynchronize();
    cudaEventRecord(e0, 0);
    double t = ref_ctx_run(ctx, params, n_steps, NULL);
    cudaEventRecord(e1, 0);
    cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if(t_advanced) *t_advanced = t;
    return (double) ms;
}

int ref_warning_count(void) { return g_capture.warnings; }

}
