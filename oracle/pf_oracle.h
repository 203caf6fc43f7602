/*
 * pf_oracle.h - CPU oracle for the phase-field hot path. TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the reference's algorithm for the path (reference simulation.cu),
 * each function citing the reference file:line it follows. It exists to CHECK the CUDA product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load it. The product library (bachelors_b200/csrc) never includes, links or calls anything here.
 *
 * Pinning (see DESIGN.md "Oracle"): (1) bit-for-bit against the UNMODIFIED reference run on a B200
 * (oracle/_ref/libref_sim_*.so built by oracle/build_ref.sh; golden vectors in tests/golden/ made by
 * tools/make_golden_from_ref.py); (2) the reference's own CPU known-answer tests for CG
 * (reference cojugate_gradient.c:416-673, iteration counts 1,3,3,2,63,63); (3) libdevice float
 * functions against the GPU (tests/golden/libdevice_f32.npz).
 */
#ifndef PF_ORACLE_H
#define PF_ORACLE_H

#include "../include/pf_solver.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_state orc_state;

orc_state* orc_create(int dtype, int nx, int ny);
void       orc_destroy(orc_state* s);
void       orc_reset(orc_state* s);                 /* forget adaptive tau / corrector parity */
void       orc_set_threads(int n);                  /* OpenMP threads used by the oracle (results do not depend on it) */
int        orc_max_threads(void);

/* sim_step (reference simulation.cu:1091-1156): returns dt advanced. Host arrays of nx*ny reals. */
double orc_step(orc_state* s, const pf_params* p, const void* F, const void* U, void* next_F, void* next_U, pf_step_info* info);
/* n steps with the host ping-pong of reference main.cpp:542-558; result is left in F/U (copied back if needed). */
double orc_run(orc_state* s, const pf_params* p, void* F, void* U, int64_t n_steps, int64_t* steps_done, pf_step_info* last_info);

/* explicit_solver_solve_lin_combination (reference simulation.cu:145-243); weights are rounded to the state's real type. */
void orc_rhs(orc_state* s, const pf_params* p, int is_euler, int n_states, const double* weights,
             const void* const* Fs, const void* const* Us, void* out_F, void* out_U);

/* conjugate_gradient_solve with the two matrix-free operators of the semi-implicit solver (reference simulation.cu:528-690). */
void orc_cg_solve(orc_state* s, const pf_params* p, int which, const void* aniso, const void* b, const void* x0, void* x,
                  double tolerance, int max_iters, int* iters, int* converged, double* error);

/* conjugate_gradient_solve over a caller-supplied operator (double only): used to replay the reference's CPU
 * known-answer systems (reference cojugate_gradient.c) through the oracle's recurrence. */
typedef void (*orc_matvec_f64)(double* out, const void* A, const double* x, int N);
void orc_cg_generic_f64(orc_matvec_f64 mul, const void* A, double* x, const double* b, int N, const double* x0,
                        double epsilon, double tolerance, int max_iters, int* iters, int* converged, double* error);

void orc_debug_values(orc_state* s, const pf_params* p, const void* F, const void* U, void* grad_F, void* grad_U, void* aniso);
void orc_stats_delta(orc_state* s, const void* a, const void* b, pf_delta_stats* out);

/* libdevice restatements (oracle/libdevice_f32.h), vectorised for tests */
void orc_hypotf_v(const float* a, const float* b, float* out, size_t n);
void orc_atan2f_v(const float* y, const float* x, float* out, size_t n);
void orc_cosf_v(const float* a, float* out, size_t n);

#ifdef __cplusplus
}
#endif
#endif
