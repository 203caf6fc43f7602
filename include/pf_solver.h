/*
 * pf_solver.h - C-ABI of the B200-native phase-field (Phi, "F") + temperature (T, "U") solver.
 *
 * This is the drop-in boundary for ONE hot path of Boostibot/bachelors: the finite-volume
 * update behind the reference's `simulation.h` (reference simulation.h:132-145). Every entry
 * point below names the reference interface it replaces (file:line, relative to the
 * reference root). The reference host (main.cpp) keeps calling sim_realloc / sim_step /
 * sim_modify*; `bachelors_b200/host/simulation_shim.cpp` implements those on top of this ABI.
 *
 * Plain C: opaque handle, raw device/host pointers, sizes. No torch types, no C++ types.
 * All functions return PF_OK (0) or a negative pf_status; pf_last_error() gives the text.
 * The library never logs and never aborts (the reference aborts on any CUDA error,
 * cuda_util.cuh:23-41; here the error is returned to the caller).
 *
 * There is NO CPU fallback: without a CUDA device every call that touches the device fails
 * with PF_ERR_NO_DEVICE.
 */
#ifndef PF_SOLVER_H
#define PF_SOLVER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#  define PF_API __declspec(dllexport)
#else
#  define PF_API __attribute__((visibility("default")))
#endif

#define PF_ABI_VERSION 2

typedef enum pf_status {
    PF_OK               =  0,
    PF_ERR_INVALID      = -1,  /* bad argument / size mismatch (reference: LOG_ERROR + return 0, simulation.cu:1094-1095) */
    PF_ERR_NO_DEVICE    = -2,  /* no CUDA device / driver */
    PF_ERR_CUDA         = -3,  /* a CUDA call failed; see pf_last_error() */
    PF_ERR_OOM          = -4,
    PF_ERR_UNSUPPORTED  = -5,  /* e.g. the corrector loop or do_exact on a slab */
} pf_status;

/* Storage/arithmetic type of the fields: the reference's compile-time `Real`
 * (simulation.h:21-25: double unless USE_FLOATS) is a run-time choice here. */
typedef enum pf_dtype {
    PF_F32 = 0,
    PF_F64 = 1,
} pf_dtype;

/* Same numeric values as Sim_Boundary_Type (simulation.h:27-32). */
typedef enum pf_boundary {
    PF_BOUNDARY_PERIODIC       = 0,
    PF_BOUNDARY_DIRICHLET_ZERO = 1,
    PF_BOUNDARY_NEUMANN_ZERO   = 2,
} pf_boundary;

/* Same numeric values as Sim_Solver_Type (simulation.h:34-42). */
typedef enum pf_solver_type {
    PF_SOLVER_NONE                  = 0,
    PF_SOLVER_EXPLICIT_EULER        = 1,  /* "explicit"              */
    PF_SOLVER_EXPLICIT_RK4          = 2,  /* "explicit-rk4"          */
    PF_SOLVER_EXPLICIT_RK4_ADAPTIVE = 3,  /* "explicit-rk4-adaptive" */
    PF_SOLVER_SEMI_IMPLICIT         = 4,  /* "semi-implicit"         */
    PF_SOLVER_EXACT                 = 5,  /* "exact": writes the analytic fields at params.time (simulation.cu:1069-1089) */
} pf_solver_type;

/* The numeric content of Sim_Params (simulation.h:83-130) without the host pointers
 * (stats, temp_maps); field meanings are identical, see config.h:413-479 for the .ini keys. */
typedef struct pf_params {
    int32_t solver;              /* pf_solver_type */
    int32_t nx, ny;              /* GLOBAL grid size */
    double  time;                /* simulated time at the start of the step (bookkeeping only) */
    int64_t iter;

    int32_t T_boundary;          /* pf_boundary */
    int32_t Phi_boundary;

    double L0;                   /* physical side length; dx = L0/nx, dy = L0/ny (simulation.cu:167-168) */
    double dt;
    double L;                    /* latent heat */
    double xi, a, b, alpha, beta;
    double gamma;                /* theta-scheme weight of the semi-implicit solver */
    double Tm;
    double min_dt;               /* lower clamp of the adaptive step (simulation.cu:463) */
    double S, m0, theta0;        /* anisotropy strength / mode / phase */

    double T_tolerance, Phi_tolerance, corrector_tolerance;
    int32_t T_max_iters, Phi_max_iters, corrector_max_iters;

    int32_t do_corrector_loop;   /* bool */
    int32_t do_corrector_guess;  /* bool */
    int32_t do_stats;            /* bool: fill pf_step_info delta statistics (simulation.cu:1126-1142) */
    int32_t do_stats_step_residual; /* bool (simulation.cu:979-993) */
    int32_t do_exact;            /* bool: source term fu = exact_fu(iter*dt) in the explicit RHS (simulation.cu:179-184); single GPU */
} pf_params;

enum { PF_MAX_STEP_RESIDUALS = 20 };   /* MAX_STEP_RESIDUALS, simulation.h:56 */

/* Per-field delta statistics of one step: Reduce::Stats over (old - new), cuda_reduction.cuh:336-407.
 * L1 = sum |d| (not divided by N), L2 = sqrt(sum d^2). */
typedef struct pf_delta_stats {
    double sum, L1, L2, min, max;
} pf_delta_stats;

/* What one step did. Superset of what the reference stores in Sim_Stats (simulation.h:57-81):
 * the reference only LOGS the CG iteration counts (simulation.cu:890,908, "@TODO: stats!" :1006). */
typedef struct pf_step_info {
    double  dt_advanced;         /* value sim_step returns (simulation.cu:1155) */
    int32_t Phi_iters, T_iters;  /* explicit: 1; adaptive: attempts (simulation.cu:488-492); semi-implicit: CG iterations */
    int32_t Phi_converged, T_converged;
    double  Phi_error, T_error;  /* adaptive: epsilon_F / epsilon_U of the last attempt; semi-implicit: sqrt(r.r/N) */
    double  next_tau;            /* adaptive: step proposed for the next call (the reference's function-static, simulation.cu:363,486) */
    pf_delta_stats Phi_delta, T_delta;      /* valid iff params.do_stats */
    int32_t step_res_count;                  /* corrector loop residuals, valid iff do_stats_step_residual */
    pf_delta_stats step_res[PF_MAX_STEP_RESIDUALS];
} pf_step_info;

/* One rank's slab of a y-decomposed global grid. For a single GPU use
 * pf_create(), which is the slab {y0 = 0, ny_local = ny, no neighbours}. */
typedef struct pf_slab {
    int32_t y0;          /* first owned global row */
    int32_t ny_local;    /* number of owned rows */
    int32_t has_lo;      /* a neighbour owns rows below y0 (incl. periodic wrap between first and last rank) */
    int32_t has_hi;
    int32_t ghost;       /* ghost rows kept on each side that has a neighbour; >= pf_required_ghost(solver) */
} pf_slab;

typedef struct pf_solver pf_solver;   /* opaque */

/* ---- library ------------------------------------------------------------------------------- */
PF_API int          pf_abi_version(void);
PF_API int          pf_device_count(void);                  /* replaces cuda_one_time_setup device scan, cuda_device.cuh:14-107 */
PF_API const char*  pf_last_error(const pf_solver* s);      /* s may be NULL: last error of a failed pf_create on this thread */
PF_API size_t       pf_sizeof_real(int dtype);
PF_API int          pf_required_ghost(int solver_type);     /* ghost rows one pf_step consumes: 1 / 4 / 5 / 1 */

/* ---- lifetime ------------------------------------------------------------------------------ */
/* Creates a solver bound to CUDA device `device` for an nx x ny grid of `dtype` reals. Owns a stream,
 * the workspace arena (replaces cache_alloc/cache_free, cuda_alloc.cuh:72-169) and the state the
 * reference keeps in function statics (adaptive tau simulation.cu:363, corrector parity :936,
 * conversion staging buffer :1013). */
PF_API int  pf_create(pf_solver** out, int device, int dtype, int nx, int ny);
PF_API int  pf_create_slab(pf_solver** out, int device, int dtype, int nx, int ny_global, const pf_slab* slab);
PF_API void pf_destroy(pf_solver* s);
PF_API int  pf_reset_state(pf_solver* s);                   /* forget adaptive tau / corrector parity (the reference needs a new process) */
/* Tuning and diagnostic switches of one handle (the reference has compile-time #defines for such things, simulation.cu:103-109). */
typedef enum pf_option {
    PF_OPT_HOST_BANDS    = 0,  /* row bands of the pipelined pf_step_host (default 8, at most 32) */
    PF_OPT_SI_PER_KERNEL = 1,  /* != 0: semi-implicit step through the per-kernel CG path instead of the one cooperative launch */
    PF_OPT_SI_GRID_SYNC  = 2,  /* != 0: cooperative step with grid.sync() + partial arrays instead of the flagged all-block sums */
    PF_OPT_HALO_IN_KERNEL = 3, /* linked slabs (pf_halo_link), explicit-rk4: 1 = hand-over inside the step kernel, 0 = wait / push launches around it */
    PF_OPT_RUN_GRAPHS    = 4,  /* pf_run: batches of single-launch fixed-dt steps replayed as CUDA graphs of 64 steps (default 1) */
    PF_OPT_SI_BLOCKS_PER_SM = 5, /* semi-implicit: blocks of the CG kernels per SM, 1 .. 4 (default 4) */
} pf_option;
PF_API int  pf_set_option(pf_solver* s, int option, int value);

/* ---- fields: replaces sim_realloc (simulation.cu:88-101) ------------------------------------ */
/* A field is a dense row-major array of rows_alloc x nx reals on the device, index x + y*nx
 * (simulation.cu:51). Single GPU: rows_alloc = ny. Slab: rows_alloc = ghost_lo + ny_local + ghost_hi
 * and row 0 of the allocation is global row y0 - ghost_lo. Zero-filled like REALLOC_ZERO. */
PF_API int  pf_field_alloc(pf_solver* s, void** dev_field);
PF_API int  pf_field_free(pf_solver* s, void* dev_field);
PF_API int  pf_field_rows(const pf_solver* s, int* rows_alloc, int* first_owned_row);

/* ---- transfers: replaces sim_modify / sim_modify_float / sim_modify_double (simulation.cu:1010-1067)
 * `count` reals starting at `dev`; `host_dtype` may differ from the solver dtype, the conversion
 * then runs on the device exactly like sim_modify_T's staging path. Blocking, like the reference. */
PF_API int  pf_upload(pf_solver* s, void* dev, const void* host, size_t count, int host_dtype);
PF_API int  pf_download(pf_solver* s, const void* dev, void* host, size_t count, int host_dtype);
/* Raw byte copies on the solver's stream, blocking: sim_modify (simulation.cu:1051-1059). */
PF_API int  pf_upload_raw(pf_solver* s, void* dev, const void* host, size_t bytes);
PF_API int  pf_download_raw(pf_solver* s, const void* dev, void* host, size_t bytes);

/* ---- the step: replaces sim_step (simulation.cu:1091-1156) ---------------------------------- */
/* Advances (F, U) by one step of params->solver into (next_F, next_U); all four are device fields of
 * this solver. Returns after the step has completed on the device (the reference synchronises at
 * simulation.cu:1153). *info may be NULL. The time advanced is info->dt_advanced (== params->dt for the
 * fixed-step solvers, the accepted tau for explicit-rk4-adaptive). */
PF_API int  pf_step(pf_solver* s, const pf_params* params,
                    const void* F, const void* U, void* next_F, void* next_U, pf_step_info* info);

/* n_steps steps with the host loop's ping-pong (main.cpp:542-558) done inside the library, no host
 * synchronisation between fixed-dt steps. On return *F_out / *U_out point at whichever of the
 * two buffer pairs holds the final state; *t_advanced is the summed dt; last_info (may be NULL)
 * describes the final step. stop_time > 0 stops as soon as params->time + advanced >= stop_time. */
PF_API int  pf_run(pf_solver* s, const pf_params* params, void* F, void* U, void* F_tmp, void* U_tmp,
                   int64_t n_steps, double stop_time,
                   void** F_out, void** U_out, int64_t* steps_done, double* t_advanced, pf_step_info* last_info);

/* pf_run that also hands back after the first step that reaches pause_time (params->time + advanced >= pause_time, checked
 * after every step but not before the first; 0 = never): the host's snapshot tests (main.cpp:507-520) for solvers whose
 * step size is not known in advance - the adaptive solver's whole interval between two snapshots in one call.
 * *t_end = params->time + dt_1 + dt_2 + ... added in this order: bit for bit the sim_time of the host loop (main.cpp:553);
 * *t_last_begin = the same sum without the last step (the time of its statistics row). With params->do_stats (and no
 * do_stats_step_residual) the steps run WITHOUT statistics on the fast paths, and last_info carries the delta statistics of
 * the last step, formed after the loop from the two ping-pong buffers - the only ones a caller of this function can see. */
PF_API int  pf_run_until(pf_solver* s, const pf_params* params, void* F, void* U, void* F_tmp, void* U_tmp,
                         int64_t n_steps, double stop_time, double pause_time,
                         void** F_out, void** U_out, int64_t* steps_done, double* t_advanced, double* t_end, double* t_last_begin,
                         pf_step_info* last_info);

/* The reference host's per-step sequence with HOST fields as ONE call: sim_modify(F, UPLOAD), sim_modify(U, UPLOAD)
 * (main.cpp:164-168), sim_step (main.cpp:553), sim_modify(next_F, DOWNLOAD), sim_modify(next_U, DOWNLOAD) (main.cpp:857).
 * host_F / host_U (y(n)) and host_next_F / host_next_U (y(n+1)) are dense nx x ny_local arrays of host_dtype (pinned memory
 * lets the copies overlap); F, U, next_F, next_U are device fields of this solver that receive the same data.
 * For explicit / explicit-rk4 on the single-kernel path upload, compute and download are pipelined over row bands
 * (full-duplex PCIe); every other configuration runs upload, pf_step, download in sequence. Returns after the
 * results are in host memory. */
PF_API int  pf_step_host(pf_solver* s, const pf_params* params,
                         const void* host_F, const void* host_U, void* host_next_F, void* host_next_U, int host_dtype,
                         void* F, void* U, void* next_F, void* next_U, pf_step_info* info);

/* Asynchronous variants for host-side pipelining: enqueue on the solver's stream, return immediately.
 * Fixed-step solvers only (the adaptive solver needs its error norm on the host to answer dt). */
PF_API int  pf_step_async(pf_solver* s, const pf_params* params,
                          const void* F, const void* U, void* next_F, void* next_U);
PF_API int  pf_sync(pf_solver* s);
PF_API void* pf_stream(pf_solver* s);                        /* cudaStream_t of the solver, for callers that order their own work */
PF_API void* pf_aux_stream(pf_solver* s);                    /* second cudaStream_t: boundary rows of a slab (see *_rows_async) */

/* Row-windowed variants for overlapping the halo exchange with compute: produce only the owned
 * global rows in [y_begin, y_end), on the solver's stream or (on_aux_stream != 0) on its auxiliary stream. A slab rank
 * launches its first/last `ghost` rows on the auxiliary stream, hands them to the transport, and launches the
 * interior rows on the main stream meanwhile. Fused single-kernel path only (equal boundary types, no statistics). */
PF_API int  pf_step_rows_async(pf_solver* s, const pf_params* params, const void* F, const void* U, void* next_F, void* next_U,
                               int y_begin, int y_end, int on_aux_stream);
PF_API int  pf_rkm_attempt_rows_async(pf_solver* s, const pf_params* params, double tau,
                                      const void* F, const void* U, void* next_F, void* next_U,
                                      int y_begin, int y_end, int on_aux_stream);
PF_API int  pf_rkm_clear_errors_async(pf_solver* s, int on_aux_stream);   /* zero pf_rkm_error_bits() before the first window of an attempt */

/* ---- the adaptive step in phases, for callers that own a cross-rank reduction ----
 * explicit_solver_rk4_adaptive_step (simulation.cu:350-497) = begin; { attempt; [all-reduce MAX]; decide } until stop; finish.
 * pf_step runs exactly this sequence for one rank. A multi-rank caller runs it in lock step on every rank and
 * max-reduces the two 64-bit words at pf_rkm_error_bits() across ranks between attempt and decide: they are the bit
 * patterns of the non-negative max-norms, which order like unsigned integers, so an integer MAX all-reduce is exact
 * and order-independent and every rank takes the identical accept / step-size decision.
 * pf_rkm_decide: *stop = 0 try again with *next_tau (at most max(T_max_iters, Phi_max_iters, 1) attempts, :390),
 * 1 = accepted, 2 = min_dt floor reached (:466; the attempt is then not counted in Phi_iters/T_iters). */
PF_API int   pf_rkm_begin(pf_solver* s, const pf_params* params, const void* F, const void* U, double* tau);
PF_API int   pf_rkm_attempt_async(pf_solver* s, const pf_params* params, double tau,
                                  const void* F, const void* U, void* next_F, void* next_U);
PF_API void* pf_rkm_error_bits(pf_solver* s);                /* device uint64_t[2]: {max|e_Phi|, max|e_T|} of the last attempt */
PF_API int   pf_rkm_decide(pf_solver* s, const pf_params* params, double tau, uint64_t bits_Phi, uint64_t bits_T,
                           int* converged, int* stop, double* next_tau, double* eps_Phi, double* eps_T);
PF_API int   pf_rkm_finish_async(pf_solver* s, const pf_params* params, double used_tau, double next_tau,
                                 const void* F, const void* U, void* next_F, void* next_U);

/* ---- single building blocks, exported for parity tests and for the viewer ------------------- */
/* out = f(sum_i w_i * (F_i, U_i)): the RHS of simulation.cu:145-243 with IS_EULER=false, n_states in 1..4. */
PF_API int  pf_rhs(pf_solver* s, const pf_params* params, int n_states, const double* weights,
                   const void* const* Fs, const void* const* Us, void* out_F, void* out_U);
/* One un-preconditioned CG solve of A x = b (simulation.cu:596-690); which = 0: A_Phi with per-cell factor
 * `aniso` (simulation.cu:551-578), which = 1: A_T = I - dt*Laplace (simulation.cu:528-549). x0 may be NULL (zero). */
PF_API int  pf_cg_solve(pf_solver* s, const pf_params* params, int which, const void* aniso,
                        const void* b, const void* x0, void* x, double tolerance, int max_iters,
                        int* iters, int* converged, double* error);
/* Debug maps of calc_debug_values (simulation.cu:245-281): |grad Phi|, |grad T| (undivided differences), g(theta). */
PF_API int  pf_debug_values(pf_solver* s, const pf_params* params, const void* F, const void* U,
                            void* grad_F, void* grad_U, void* aniso);
/* Delta statistics (cuda_stats_delta, cuda_reduction.cuh:391-407) of a - b. */
PF_API int  pf_stats_delta(pf_solver* s, const void* a, const void* b, pf_delta_stats* out);

/* pf_stats_delta for callers that cannot allocate. NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc. The same kernel and the same
 * statistics as pf_stats_delta over n values of `dtype` (PF_F32 / PF_F64) at device pointers a, b on `device`, launched on
 * `stream` (a cudaStream_t, NULL = the default stream); returns after the stream has finished the call. No solver handle and no
 * cudaMalloc / cudaFree / allocating library call: the per-block partial sums live in a file-scope __device__ array of
 * PF_STATIC_STATS_MAX_BLOCKS x 40 bytes (40 KB) that every GPU loading the library holds until the process exits.
 * Limit: 1 <= n <= PF_STATIC_STATS_MAX_ELEMS (an 8192 x 8192 field), beyond that PF_ERR_INVALID with a message in
 * pf_last_error(NULL). All calls on a device share the array: calls that overlap on the GPU race; serialise them. */
enum { PF_STATIC_STATS_MAX_BLOCKS = 1024, PF_STATIC_STATS_MAX_ELEMS = 1 << 26 };
PF_API int  pf_stats_delta_static(int device, int dtype, const void* a, const void* b, size_t n, void* stream, pf_delta_stats* out);

/* ---- multi-GPU slabs ----------------------------------------------------- */
/* Pointers/sizes of the rows a rank must SEND to / RECEIVE from a neighbour before a step:
 * side 0 = lo (rows below), 1 = hi. `send` = the first/last `ghost` OWNED rows, `recv` = the ghost rows.
 * The transport (NCCL send/recv or peer copies) is the caller's; rows are contiguous nx*ghost reals. */
PF_API int  pf_halo_region(const pf_solver* s, void* field, int side, void** send_ptr, void** recv_ptr, size_t* count);

/* ---- device-initiated halo exchange over NVLink peer memory (one process per GPU of one box) -----------------------
 * A rank exports its fields and its flag block (cudaIpc), the neighbours map them; after its edge rows are computed a rank
 * copies them straight into the neighbour's ghost rows and publishes the exchange number `value` into the neighbour's flag
 * word (pf_halo_push_async: one small kernel, system-scope release); the neighbour's next edge-row launch is preceded by
 * pf_halo_wait_async (one-thread kernel, acquire). No NCCL call and no host synchronisation per step. The handles are
 * opaque 64-byte blobs to be carried between the processes by the caller (any transport). */
enum { PF_IPC_HANDLE_BYTES = 64, PF_HALO_FLAG_FROM_LO = 0, PF_HALO_FLAG_FROM_HI = 1 };
PF_API int   pf_ipc_export(pf_solver* s, void* dev_ptr, void* handle);             /* dev_ptr: a pf_field_alloc field or pf_halo_flags() */
PF_API int   pf_ipc_open(pf_solver* s, const void* handle, void** mapped);         /* in ANOTHER process of the box */
PF_API int   pf_ipc_close(pf_solver* s, void* mapped);
PF_API void* pf_halo_flags(pf_solver* s);                                           /* this rank's flag block (device memory, zeroed) */
/* side 0: my first `ghost` owned rows go to the lower neighbour, whose ghost rows start at row `peer_first_row` of its
 * allocation (its ghost_lo + ny_local); side 1: my last `ghost` owned rows go to the upper neighbour (peer_first_row = 0).
 * peer_fields[i] / peer_flags are MAPPED pointers of the neighbour's fields / flag block; peer_flag_index says which of its
 * words to publish into (PF_HALO_FLAG_FROM_HI when I am its upper neighbour, i.e. side 0). */
PF_API int   pf_halo_push_async(pf_solver* s, int side, int n_fields, const void* const* fields, void* const* peer_fields,
                                size_t peer_first_row, void* peer_flags, int peer_flag_index, unsigned value, int on_aux_stream);
PF_API int   pf_halo_wait_async(pf_solver* s, int flag_index, unsigned value, int on_aux_stream);
/* Links a slab to its neighbours for good: `fields` are this rank's n_fields fields (any of which may be the input or output of a step),
 * lo_fields / hi_fields the neighbours' SAME fields in the same order (mapped), lo_flags / hi_flags their flag blocks (NULL where the slab
 * has no neighbour), lo_first_row the row of the lower neighbour's allocation where its upper ghost rows start. From then on every
 * whole-slab pf_step_async of explicit / explicit-rk4 does the exchange itself: explicit-rk4 INSIDE its one step kernel
 * (k_stream_step_halo: the first / last chunk's units acquire the neighbours' flags before they start and forward their finished rows),
 * explicit with a wait launch before and a push launch after its kernel. Collective: every rank of the job links, then a barrier, before
 * the first step; ghost rows of the first step's inputs must be current. n_fields = 0 unlinks. */
PF_API int   pf_halo_link(pf_solver* s, int n_fields, void* const* fields, void* const* lo_fields, size_t lo_first_row, void* lo_flags,
                          void* const* hi_fields, void* hi_flags);
PF_API int   pf_halo_check(pf_solver* s);   /* after a synchronisation: PF_ERR_CUDA if a wait gave up (neighbour gone) */

/* ---- semi-implicit solver on slabs: the step in phases --------------------- */
/* semi_implicit_solver_step_based (simulation.cu:732-926) cut where ranks must talk. The library has no communication
 * code; per time step the caller runs, on every rank in lock step (all calls enqueue on pf_stream and return):
 *   pf_si_begin_async                          prepare kernel (needs 1 current ghost row of F and U)
 *   for which in {0 = Phi, 1 = T}:
 *     pf_si_cg_begin_async(which, part)        b_T (which = 1), r = b - A x0, p = r; `part` <- pf_si_partials() partial sums of r.r
 *     ALL-GATHER part -> rr_all                (n_all = sum of the ranks' pf_si_partials(), rank order)
 *     repeat max_iters times (or until pf_si_cg_poll reports done; the flag is identical on all ranks):
 *       EXCHANGE one ghost row of pf_si_direction() with the neighbours (pf_halo_region gives the rows)
 *       pf_si_cg_spmv_async(which, part)       Ap = A p, partial sums of p.Ap            -> ALL-GATHER -> pAp_all
 *       pf_si_cg_update_async(.., rr_all, pAp_all, n_all, part)   x, r update, partials of the new r.r -> ALL-GATHER -> rr_new_all
 *       pf_si_cg_pupdate_async(.., rr_all, rr_new_all, n_all, it) convergence test, p update; then rr_all <-> rr_new_all
 *   EXCHANGE the ghost rows of next_F, next_U before the next step.
 * Every rank folds the same gathered partials in the same order, so alpha, beta and the stop decision are bit-identical
 * across ranks. Against the single-GPU run only the summation tree differs (rel-L2 <= 1e-10 fp64, iterations +-1).
 * `part` and the gathered arrays are device memory of the solver's Real type. */
PF_API int   pf_si_begin_async(pf_solver* s, const pf_params* params, const void* F, const void* U, const void* U_base,
                               void* next_F, void* next_U);
PF_API int   pf_si_partials(pf_solver* s);
PF_API int   pf_si_cg_begin_async(pf_solver* s, int which, void* partials_out);
PF_API int   pf_si_cg_spmv_async(pf_solver* s, int which, void* partials_out);
PF_API int   pf_si_cg_update_async(pf_solver* s, int which, const void* rr_all, const void* pAp_all, int n_all, void* partials_out);
PF_API int   pf_si_cg_pupdate_async(pf_solver* s, int which, const void* rr_all, const void* rr_new_all, int n_all, int iter);
PF_API void* pf_si_direction(pf_solver* s);                  /* the search direction p, laid out like a field (ghost rows included) */
/* Synchronises pf_stream and reads the solve's control block: done flag, iterations (loop index at break, simulation.cu:681),
 * converged, sqrt(r.r/N). */
PF_API int   pf_si_cg_poll(pf_solver* s, int which, int* done, int* iters, int* converged, double* error);

/* ---- initial conditions on the device --------------------------------- */
/* The [initial] section of config.ini (Sim_Config, config.h:20-30). */
typedef struct pf_initial {
    double inside_phi, inside_T, outside_phi, outside_T;
    double circle_center[2];
    double circle_radius, circle_fade;
    double square_from[2], square_to[2];
} pf_initial;
/* sim_make_initial_conditions (main.cpp:93-136) evaluated on the device, straight into the solver's fields (ghost
 * rows of a slab included): replaces the CPU loop + sim_modify(UPLOAD) x2 of simulation_state_reload (main.cpp:151-173).
 * All arithmetic in double as on the host, cast to Real at the end. */
PF_API int  pf_init_fields(pf_solver* s, const pf_params* params, const pf_initial* init, void* F, void* U);

/* ---- snapshots: device -> pinned host, overlapped with stepping -------- */
enum { PF_SNAPSHOT_SLOTS = 2, PF_SNAPSHOT_MAX_FIELDS = 8 };
/* Replaces the blocking sim_modify_double(DOWNLOAD) per map of save_state (main.cpp:846-858, simulation.cu:1010-1049).
 * _begin: the OWNED rows of each field are copied (f32 solvers: converted to double, simulation.cu:1036-1047) into a
 * device staging buffer in stream order - the caller may step on at once - and from there to a pinned host buffer on
 * a copy stream. _wait blocks until slot's copy has landed and returns the host pointers (ny_local*nx doubles each,
 * valid until the slot is used again). Two slots: one can be written to disk while the other fills. */
PF_API int  pf_snapshot_begin(pf_solver* s, int slot, int n_fields, const void* const* dev_fields);
PF_API int  pf_snapshot_wait(pf_solver* s, int slot, const double** host_fields);

/* ---- output files of the reference host (pure host code: no device needed) ------------------- */
/* maps_NNNN.bin, byte for byte what save_bin_map_file writes (main.cpp:705-735; read by plot.py:26-73):
 * i32 magic 0x11223344, i32 map_count, i32 nx, i32 ny, f64 dx, f64 dy, f64 time, i64 iter, map_count x char[32]
 * names, map_count x nx*ny f64. */
PF_API int  pf_write_maps_bin(const char* path, int nx, int ny, double dx, double dy, int64_t iter, double time,
                              int n_maps, const char* const* names, const double* const* maps);
/* The same file written in pieces (slabs of several ranks): header first, then any rows of any map in any order. */
PF_API int  pf_write_maps_bin_header(const char* path, int nx, int ny, double dx, double dy, int64_t iter, double time,
                                     int n_maps, const char* const* names);
PF_API int  pf_write_maps_bin_rows(const char* path, int nx, int ny, int n_maps, int map_index, int y0, int rows, const double* data);

/* One line of stats.csv = one App_Stats sample (main.cpp:192-234); floats as the reference stores them (Sim_Stats, simulation.h:57-81). */
typedef struct pf_stats_row {
    double  time;
    int64_t iter;
    int64_t Phi_iters, T_iters;
    float   T_delta_L1, T_delta_L2, T_delta_max, T_delta_min;
    float   Phi_delta_L1, Phi_delta_L2, Phi_delta_max, Phi_delta_min;
    float   step_res_L1[PF_MAX_STEP_RESIDUALS], step_res_L2[PF_MAX_STEP_RESIDUALS];
    float   step_res_max[PF_MAX_STEP_RESIDUALS], step_res_min[PF_MAX_STEP_RESIDUALS];
} pf_stats_row;
/* save_csv_stat_file (main.cpp:741-757, 782-825): header lines unless `append`, then one line per row, every number
 * through the reference's 16-byte "%f" / "%lli" strings. */
PF_API int  pf_write_stats_csv(const char* path, int append, int nx, int ny, double dt, int step_res_count,
                               size_t n_rows, const pf_stats_row* rows);

/* ---- headless view ----------------------------------------------------- */
/* The viewer's fragment shader (gl.h:221-283: NaN -> magenta, grey arctangent ramps outside [min, max], 4-segment
 * blue-cyan-green-yellow-red inside) evaluated by a CUDA kernel over the owned rows of a field (sampled as float,
 * gl.h:396), RGBA8 into host memory (ny_local*nx*4 bytes, row y of the grid = row y of the buffer). The GL window,
 * swap chain and texture upload of the reference are not reproduced. */
PF_API int  pf_render_rgba(pf_solver* s, const void* dev_field, float display_min, float display_max, uint8_t* host_rgba);
/* Minimal PNG writer (8-bit RGBA, stored deflate blocks). flip_y != 0 puts grid row 0 at the bottom, as the viewer does. */
PF_API int  pf_write_png(const char* path, int width, int height, const uint8_t* rgba, int flip_y);

/* ---- reduction micro-benchmark: replaces run_benchmarks (simulation.cu:1300-1358, numbers in its plot.py:731-733) -------------
 * Times, on the solver's stream with CUDA events (median of `reps` after two warm-up rounds), this library's two reductions over
 * n pseudo-random values of value_dtype: (1) max |x| - 16-byte loads, warp shuffles, one integer atomicMax per warp on the IEEE bit
 * pattern, what replaces the reference's multi-pass cuda_reduce(Reduce::MAX); reads n values; (2) the five delta statistics of a - b
 * (cuda_stats_delta, cuda_reduction.cuh:391-407) as pf_step computes them; reads 2n values. *max_value returns the reduced maximum
 * (< 1, the generator is uniform in [0, 1)). */
PF_API int  pf_bench_reductions(pf_solver* s, size_t n, int value_dtype, int reps, double* seconds_max, double* seconds_stats, double* max_value);

/* ---- counters -------------------------------------------------------------------------------- */
PF_API int64_t pf_kernel_launches(const pf_solver* s);      /* kernels launched by this solver so far (bench.py: gpu_launches) */

#ifdef __cplusplus
}
#endif
#endif /* PF_SOLVER_H */
