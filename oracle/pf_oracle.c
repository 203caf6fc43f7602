/*
 * pf_oracle.c - CPU oracle, exported entry points. TEST INFRASTRUCTURE ONLY (see pf_oracle.h).
 * Compile with: gcc -O2 -ffp-contract=off -fopenmp -fPIC -shared (oracle/Makefile).
 */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "pf_oracle.h"
#include "libdevice_f32.h"

enum { ORC_SCRATCH_SLOTS = 12 };

struct orc_state {
    int dtype, nx, ny;
    double initial_step;       /* `static Real _initial_step`, reference simulation.cu:363 */
    int last_num_steps;        /* `static int last_num_steps`, reference simulation.cu:936 */
    void* scratch[ORC_SCRATCH_SLOTS];
};

#define REAL double
#define SFX _f64
#define ORC_IS_DOUBLE 1
#include "pf_oracle_impl.h"
#undef REAL
#undef SFX
#undef ORC_IS_DOUBLE

#define REAL float
#define SFX _f32
#define ORC_IS_DOUBLE 0
#include "pf_oracle_impl.h"
#undef REAL
#undef SFX
#undef ORC_IS_DOUBLE

orc_state* orc_create(int dtype, int nx, int ny)
{
    orc_state* s = (orc_state*) calloc(1, sizeof *s);
    s->dtype = dtype; s->nx = nx; s->ny = ny;
    return s;
}

void orc_destroy(orc_state* s)
{
    if(!s) return;
    for(int i = 0; i < ORC_SCRATCH_SLOTS; i++) free(s->scratch[i]);
    free(s);
}

void orc_reset(orc_state* s) { s->initial_step = 0; s->last_num_steps = 0; }

void orc_set_threads(int n)
{
#ifdef _OPENMP
    if(n > 0) omp_set_num_threads(n);
#else
    (void) n;
#endif
}

int orc_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

double orc_step(orc_state* s, const pf_params* p, const void* F, const void* U, void* next_F, void* next_U, pf_step_info* info)
{
    if(s->dtype == PF_F64)
        return orc_step_f64(s, p, (const double*) F, (const double*) U, (double*) next_F, (double*) next_U, info);
    return orc_step_f32(s, p, (const float*) F, (const float*) U, (float*) next_F, (float*) next_U, info);
}

double orc_run(orc_state* s, const pf_params* p, void* F, void* U, int64_t n_steps, int64_t* steps_done, pf_step_info* last_info)
{
    size_t bytes = (size_t) p->nx * (size_t) p->ny * (s->dtype == PF_F64 ? 8 : 4);
    void* nF = malloc(bytes); void* nU = malloc(bytes);
    void* cF = F; void* cU = U;
    pf_params q = *p;
    double t = p->time;
    int64_t i = 0;
    for(; i < n_steps; i++)
    {
        q.iter = p->iter + i; q.time = t;
        t += orc_step(s, &q, cF, cU, nF, nU, last_info);
        void* tmp = cF; cF = nF; nF = tmp;
        tmp = cU; cU = nU; nU = tmp;
    }
    if(cF != F) { memcpy(F, cF, bytes); memcpy(U, cU, bytes); free(cF); free(cU); }
    else { free(nF); free(nU); }
    if(steps_done) *steps_done = i;
    return t - p->time;
}

void orc_rhs(orc_state* s, const pf_params* p, int is_euler, int n_states, const double* weights,
             const void* const* Fs, const void* const* Us, void* out_F, void* out_U)
{
    if(s->dtype == PF_F64)
    {
        double w[4]; for(int i = 0; i < n_states; i++) w[i] = weights[i];
        orc_rhs_f64(p, is_euler, n_states, w, (const double* const*) Fs, (const double* const*) Us, (double*) out_F, (double*) out_U);
    }
    else
    {
        float w[4]; for(int i = 0; i < n_states; i++) w[i] = (float) weights[i];
        orc_rhs_f32(p, is_euler, n_states, w, (const float* const*) Fs, (const float* const*) Us, (float*) out_F, (float*) out_U);
    }
}

void orc_cg_solve(orc_state* s, const pf_params* p, int which, const void* aniso, const void* b, const void* x0, void* x,
                  double tolerance, int max_iters, int* iters, int* converged, double* error)
{
    int nx = p->nx, ny = p->ny, N = nx * ny;
    if(s->dtype == PF_F64)
    {
        double dt = p->dt, dx = p->L0 / nx, dy = p->L0 / ny, err = 0;
        double* r = orc_scratch_f64(s, 3); double* pv = orc_scratch_f64(s, 4); double* Ap = orc_scratch_f64(s, 5);
        if(which == 0)
        {
            orc_aniso_matrix_f64 A; A.anisotrophy = (const double*) aniso;
            A.C_minus_one = 2 * dt / (dx * dx) + 2 * dt / (dy * dy); A.X = -dt / (dx * dx); A.Y = -dt / (dy * dy);
            A.nx = nx; A.ny = ny; A.boundary = p->Phi_boundary;
            orc_cg_solve_f64(&A, (double*) x, (const double*) b, N, nx, orc_anisotrophy_matrix_multiply_f64, 1.0e-12, tolerance, max_iters,
                             (const double*) x0, r, pv, Ap, iters, converged, &err);
        }
        else
        {
            orc_cross_static_f64 A; A.C = 1 + 2 * dt / (dx * dx) + 2 * dt / (dy * dy);
            A.R = A.L = -dt / (dx * dx); A.U = A.D = -dt / (dy * dy); A.nx = nx; A.ny = ny; A.boundary = p->T_boundary;
            orc_cg_solve_f64(&A, (double*) x, (const double*) b, N, nx, orc_cross_matrix_static_multiply_f64, 1.0e-12, tolerance, max_iters,
                             (const double*) x0, r, pv, Ap, iters, converged, &err);
        }
        *error = err;
    }
    else
    {
        float dt = (float) p->dt, dx = (float) p->L0 / nx, dy = (float) p->L0 / ny, err = 0;
        float* r = orc_scratch_f32(s, 3); float* pv = orc_scratch_f32(s, 4); float* Ap = orc_scratch_f32(s, 5);
        if(which == 0)
        {
            orc_aniso_matrix_f32 A; A.anisotrophy = (const float*) aniso;
            A.C_minus_one = 2 * dt / (dx * dx) + 2 * dt / (dy * dy); A.X = -dt / (dx * dx); A.Y = -dt / (dy * dy);
            A.nx = nx; A.ny = ny; A.boundary = p->Phi_boundary;
            orc_cg_solve_f32(&A, (float*) x, (const float*) b, N, nx, orc_anisotrophy_matrix_multiply_f32, 1.0e-12f, (float) tolerance, max_iters,
                             (const float*) x0, r, pv, Ap, iters, converged, &err);
        }
        else
        {
            orc_cross_static_f32 A; A.C = 1 + 2 * dt / (dx * dx) + 2 * dt / (dy * dy);
            A.R = A.L = -dt / (dx * dx); A.U = A.D = -dt / (dy * dy); A.nx = nx; A.ny = ny; A.boundary = p->T_boundary;
            orc_cg_solve_f32(&A, (float*) x, (const float*) b, N, nx, orc_cross_matrix_static_multiply_f32, 1.0e-12f, (float) tolerance, max_iters,
                             (const float*) x0, r, pv, Ap, iters, converged, &err);
        }
        *error = err;
    }
}

void orc_cg_generic_f64(orc_matvec_f64 mul, const void* A, double* x, const double* b, int N, const double* x0,
                        double epsilon, double tolerance, int max_iters, int* iters, int* converged, double* error)
{
    double* r = (double*) malloc(sizeof(double) * (size_t) N);
    double* pv = (double*) malloc(sizeof(double) * (size_t) N);
    double* Ap = (double*) malloc(sizeof(double) * (size_t) N);
    orc_cg_solve_f64(A, x, b, N, N, mul, epsilon, tolerance, max_iters, x0, r, pv, Ap, iters, converged, error);
    free(r); free(pv); free(Ap);
}

void orc_debug_values(orc_state* s, const pf_params* p, const void* F, const void* U, void* grad_F, void* grad_U, void* aniso)
{
    if(s->dtype == PF_F64) orc_debug_values_f64(p, (const double*) F, (const double*) U, (double*) grad_F, (double*) grad_U, (double*) aniso);
    else                   orc_debug_values_f32(p, (const float*) F, (const float*) U, (float*) grad_F, (float*) grad_U, (float*) aniso);
}

void orc_stats_delta(orc_state* s, const void* a, const void* b, pf_delta_stats* out)
{
    size_t N = (size_t) s->nx * (size_t) s->ny;
    if(s->dtype == PF_F64) orc_stats_delta_f64((const double*) a, (const double*) b, N, out);
    else                   orc_stats_delta_f32((const float*) a, (const float*) b, N, out);
}

void orc_hypotf_v(const float* a, const float* b, float* out, size_t n) { for(size_t i = 0; i < n; i++) out[i] = orc_hypotf(a[i], b[i]); }
void orc_atan2f_v(const float* y, const float* x, float* out, size_t n) { for(size_t i = 0; i < n; i++) out[i] = orc_atan2f(y[i], x[i]); }
void orc_cosf_v(const float* a, float* out, size_t n) { for(size_t i = 0; i < n; i++) out[i] = orc_cosf(a[i]); }
