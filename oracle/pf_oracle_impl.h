/*
 * pf_oracle_impl.h - generic body of the CPU oracle, included once per precision by pf_oracle.c
 * with REAL / SFX defined. TEST INFRASTRUCTURE ONLY (see pf_oracle.h).
 *
 * Every function names the reference lines it restates (relative to the reference root).
 *
 * Floating-point contraction. The reference is compiled by nvcc with the default --fmad=true
 * (reference Makefile:12 sets no -fmad), so its results depend on which mul+add pairs nvcc fuses.
 * The fusions written out below as explicit FMA(...) are the ones nvcc 12.9.86 makes for the
 * reference's expressions on sm_100a, read from the SASS of the reference object built by
 * oracle/build_ref.sh (cuobjdump -sass oracle/_ref/simulation_f64.o, kernel
 * cuda_tiled_for_2D_kernel<Phase_Temp, ...explicit_solver_solve_lin_combination<...>>, the two
 * semi_implicit_solver_step_based prepare kernels, the Merson error-norm reduction; fp64 and fp32 objects agree).
 * Which product of an `a*b + c*d` is fused differs from expression to expression, so each one below is
 * written the way the SASS has it. This file must be compiled with -ffp-contract=off so that nothing else is fused.
 */

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, SFX)

#if ORC_IS_DOUBLE
  #define FMA(a, b, c) fma((a), (b), (c))
  #define SQRT(a) sqrt(a)
  #define POW(a, b) pow((a), (b))
#else
  #define FMA(a, b, c) fmaf((a), (b), (c))
  #define SQRT(a) sqrtf(a)
  #define POW(a, b) powf((a), (b))
#endif

#if ORC_IS_DOUBLE
  #define EXPR(a) exp(a)
  #define ERFR(a) erf(a)
  #define HYPOTR(a, b) hypot((a), (b))
#else
  #define EXPR(a) expf(a)
  #define ERFR(a) erff(a)
  #define HYPOTR(a, b) hypotf((a), (b))
#endif

/* The analytic ("exact") mode: reference exact.h:17-96, all in Real arithmetic. Static parameters exact.h:90-96. */
#define ORC_EXACT_LAMBDA  ((REAL) 0.5)
#define ORC_EXACT_R_INI   ((REAL) 0.25)
#define ORC_EXACT_EPSILON ((REAL) 0.001)
static inline REAL FN(orc_exact_R)(REAL t) { return SQRT(ORC_EXACT_R_INI * ORC_EXACT_R_INI + 2 * ORC_EXACT_LAMBDA * t); }      /* exact.h:17-20 */
static inline REAL FN(orc_exact_U)(REAL Rt) { return -ORC_EXACT_EPSILON * (ORC_EXACT_LAMBDA + 2) / Rt; }                        /* exact.h:22-25 */
static inline REAL FN(orc_exact_T)(REAL s)                                                                                       /* exact.h:27-35 */
{
    REAL l = ORC_EXACT_LAMBDA;
    REAL pi = (REAL) 3.14159265359;
    REAL sqrtl2 = SQRT(l / 2);
    REAL integral = EXPR(-l / 2) - 1 / s * EXPR(-l / 2 * s * s) + sqrtl2 * pi * (ERFR(sqrtl2) - ERFR(s * sqrtl2));
    return -l * EXPR(l / 2) * integral;
}
static inline REAL FN(orc_exact_fu)(REAL t)                                                                                      /* exact.h:37-42 */
{
    REAL Rt = FN(orc_exact_R)(t);
    REAL l = ORC_EXACT_LAMBDA;
    return ORC_EXACT_EPSILON * l * (l + 2) / (Rt * Rt * Rt);
}
static inline REAL FN(orc_exact_u)(REAL t, REAL r)                                                                               /* exact.h:44-53 */
{
    REAL Rt = FN(orc_exact_R)(t);
    REAL ut = FN(orc_exact_U)(Rt);
    if(r > Rt) ut += FN(orc_exact_T)(r / Rt);
    return ut;
}
static inline REAL FN(orc_exact_phi)(REAL t, REAL r) { return r <= FN(orc_exact_R)(t) ? (REAL) 1 : (REAL) 0; }                  /* exact.h:55-58 */

/* exact_solver_step: reference simulation.cu:1069-1089. `((Real) xi + 0.5)*dx - L0/2` promotes to double for Real = float. */
static void FN(orc_exact_step)(const pf_params* p, REAL* next_F, REAL* next_U)
{
    const REAL dx = (REAL) p->L0 / p->nx, dy = (REAL) p->L0 / p->ny, L0 = (REAL) p->L0, t = (REAL) p->time;
    const int nx = p->nx, ny = p->ny;
    #pragma omp parallel for schedule(static)
    for(int yi = 0; yi < ny; yi++)
        for(int xi = 0; xi < nx; xi++)
        {
            REAL x = (REAL)(((REAL) xi + 0.5) * dx - L0 / 2);
            REAL y = (REAL)(((REAL) yi + 0.5) * dy - L0 / 2);
            REAL r = HYPOTR(x, y);
            next_U[(size_t) xi + (size_t) yi * (size_t) nx] = FN(orc_exact_u)(t, r);
            next_F[(size_t) xi + (size_t) yi * (size_t) nx] = FN(orc_exact_phi)(t, r);
        }
}

/* custom_hypot / custom_atan2 / custom_cos: reference simulation.cu:14-17 (float math, cast back to Real). */
static inline REAL FN(orc_custom_hypot)(REAL y, REAL x) { return (REAL) orc_hypotf((float) y, (float) x); }
static inline REAL FN(orc_custom_atan2)(REAL y, REAL x) { return (REAL) orc_atan2f((float) y, (float) x); }
static inline REAL FN(orc_custom_cos)(REAL t) { return (REAL) orc_cosf((float) t); }

/* boundary_sample: reference simulation.cu:33-85. */
static inline REAL FN(orc_boundary_sample)(const REAL* map, int x, int y, int nx, int ny, int bound)
{
    switch(bound)
    {
        case PF_BOUNDARY_PERIODIC: {
            int xm = x; if(xm < 0) xm += nx; else if(xm >= nx) xm -= nx;
            int ym = y; if(ym < 0) ym += ny; else if(ym >= ny) ym -= ny;
            return map[(size_t) xm + (size_t) ym * (size_t) nx];
        }
        case PF_BOUNDARY_DIRICHLET_ZERO: {
            if(0 <= x && x < nx && 0 <= y && y < ny)
                return map[(size_t) x + (size_t) y * (size_t) nx];
            int cx = x < 0 ? 0 : (x > nx - 1 ? nx - 1 : x);
            int cy = y < 0 ? 0 : (y > ny - 1 ? ny - 1 : y);
            REAL dirichlet = 0;
            return 2 * dirichlet - map[(size_t) cx + (size_t) cy * (size_t) nx];      /* :70 */
        }
        case PF_BOUNDARY_NEUMANN_ZERO: {
            int cx = x < 0 ? 0 : (x > nx - 1 ? nx - 1 : x);
            int cy = y < 0 ? 0 : (y > ny - 1 ? ny - 1 : y);
            return map[(size_t) cx + (size_t) cy * (size_t) nx];
        }
        default: return 0;
    }
}

/* Host-side constants of the RHS: reference simulation.cu:154-177 (all in Real arithmetic). */
typedef struct FN(orc_consts) {
    REAL a, b, alpha, beta, xi, Tm, L, dt, S0, m0, theta0, gamma;
    REAL dx, dy, one_over_2dx, one_over_2dy, one_over_dx2, one_over_dy2;
    REAL k0_factor, k2_factor, k1_factor, dt_L;
    int nx, ny, U_bound, F_bound, do_corrector_guess;
} FN(orc_consts);

static FN(orc_consts) FN(orc_make_consts)(const pf_params* p)
{
    FN(orc_consts) c;
    c.nx = p->nx; c.ny = p->ny;
    c.a = (REAL) p->a; c.b = (REAL) p->b; c.alpha = (REAL) p->alpha; c.beta = (REAL) p->beta;
    c.xi = (REAL) p->xi; c.Tm = (REAL) p->Tm; c.L = (REAL) p->L; c.dt = (REAL) p->dt;
    c.S0 = (REAL) p->S; c.m0 = (REAL) p->m0; c.theta0 = (REAL) p->theta0; c.gamma = (REAL) p->gamma;
    c.dx = (REAL) p->L0 / p->nx;
    c.dy = (REAL) p->L0 / p->ny;
    c.one_over_2dx = 1 / (2 * c.dx);
    c.one_over_2dy = 1 / (2 * c.dy);
    c.one_over_dx2 = 1 / (c.dx * c.dx);
    c.one_over_dy2 = 1 / (c.dy * c.dy);
    c.k0_factor = c.a / (c.xi * c.xi * c.alpha);
    c.k2_factor = c.b * c.beta / c.alpha;
    c.k1_factor = 1 / c.alpha;
    c.dt_L = c.dt * c.L;
    c.U_bound = p->T_boundary; c.F_bound = p->Phi_boundary;
    c.do_corrector_guess = p->do_corrector_guess != 0;
    return c;
}

/* Blended, boundary-sampled stage input at (x, y): reference simulation.cu:189-200.
 * F = 0; F += sample_i * weight_i  (each += is one fused multiply-add). */
static inline void FN(orc_gather)(const FN(orc_consts)* c, int n_states, const REAL* w,
                                  const REAL* const* Fs, const REAL* const* Us, int x, int y, REAL* outF, REAL* outU)
{
    REAL U = 0, F = 0;
    for(int i = 0; i < n_states; i++)
    {
        F = FMA(FN(orc_boundary_sample)(Fs[i], x, y, c->nx, c->ny, c->F_bound), w[i], F);
        U = FMA(FN(orc_boundary_sample)(Us[i], x, y, c->nx, c->ny, c->U_bound), w[i], U);
    }
    *outF = F; *outU = U;
}

/* The per-cell right-hand side: reference simulation.cu:201-229. Inputs are the five blended
 * (Phi, T) samples. */
static inline void FN(orc_rhs_cell)(const FN(orc_consts)* c,
    REAL C_Phi, REAL E_Phi, REAL W_Phi, REAL N_Phi, REAL S_Phi,
    REAL C_T, REAL E_T, REAL W_T, REAL N_T, REAL S_T, REAL fu, REAL* dt_Phi_out, REAL* dt_T_out)
{
    REAL grad_Phi_x = (E_Phi - W_Phi) * c->one_over_2dx;
    REAL grad_Phi_y = (N_Phi - S_Phi) * c->one_over_2dx;                 /* sic: 2dx, :209 */
    REAL grad_Phi_norm = FN(orc_custom_hypot)(grad_Phi_x, grad_Phi_y);

    REAL theta = FN(orc_custom_atan2)(grad_Phi_y, grad_Phi_x);
    REAL g_theta = FMA(-FN(orc_custom_cos)(FMA(c->m0, theta, c->theta0)), c->S0, (REAL) 1);   /* 1 - S0*cos(m0*theta + theta0) */

    REAL two_Phi = 2 * C_Phi, two_T = 2 * C_T;
    /* a*b + c*d: nvcc fuses the x product and rounds the y product (DMUL ysum, DFMA xsum in every RHS kernel's SASS) */
    REAL laplace_Phi = FMA((W_Phi - two_Phi) + E_Phi, c->one_over_dx2, ((S_Phi - two_Phi) + N_Phi) * c->one_over_dy2);
    REAL laplace_T   = FMA((W_T - two_T) + E_T,       c->one_over_dx2, ((S_T - two_T) + N_T) * c->one_over_dy2);

    REAL f0 = C_Phi * (1 - C_Phi) * (C_Phi - (REAL) 0.5);                  /* :129-132 */
    REAL k0 = g_theta * f0 * c->k0_factor;
    REAL k2 = grad_Phi_norm * c->k2_factor;
    REAL k1 = g_theta * c->k1_factor;
    REAL corr = FMA(k2, c->dt_L, (REAL) 1);

    REAL dt_Phi;
    if(c->do_corrector_guess)
        dt_Phi = FMA(-k2, FMA(c->dt, laplace_T, C_T - c->Tm), FMA(k1, laplace_Phi, k0)) / corr;   /* :225 */
    else
        dt_Phi = FMA(-k2, C_T - c->Tm, FMA(k1, laplace_Phi, k0));                                  /* :227 */

    REAL dt_T = FMA(c->L, dt_Phi, laplace_T) + fu;                         /* :229 */
    *dt_Phi_out = dt_Phi; *dt_T_out = dt_T;
}

/* explicit_solver_solve_lin_combination<IS_EULER, States...>: reference simulation.cu:145-243.
 * The reference stages the gathered values through a shared-memory tile; the arithmetic per cell
 * is independent of the tiling, so the oracle gathers the five samples directly. */
static void FN(orc_rhs)(const pf_params* p, int is_euler, int n_states, const REAL* w,
                        const REAL* const* Fs, const REAL* const* Us, REAL* out_F, REAL* out_U)
{
    FN(orc_consts) c = FN(orc_make_consts)(p);
    const int nx = c.nx, ny = c.ny;
    const REAL fu = p->do_exact ? FN(orc_exact_fu)(p->iter * c.dt) : 0;          /* :179-184 */
    #pragma omp parallel for schedule(static)
    for(int y = 0; y < ny; y++)
    {
        for(int x = 0; x < nx; x++)
        {
            REAL CF, CU, EF, EU, WF, WU, NF, NU, SF, SU;
            FN(orc_gather)(&c, n_states, w, Fs, Us, x, y, &CF, &CU);
            FN(orc_gather)(&c, n_states, w, Fs, Us, x + 1, y, &EF, &EU);
            FN(orc_gather)(&c, n_states, w, Fs, Us, x - 1, y, &WF, &WU);
            FN(orc_gather)(&c, n_states, w, Fs, Us, x, y + 1, &NF, &NU);
            FN(orc_gather)(&c, n_states, w, Fs, Us, x, y - 1, &SF, &SU);
            REAL dF, dU;
            FN(orc_rhs_cell)(&c, CF, EF, WF, NF, SF, CU, EU, WU, NU, SU, fu, &dF, &dU);
            size_t i = (size_t) x + (size_t) y * (size_t) nx;
            if(is_euler) { out_F[i] = FMA(c.dt, dF, CF); out_U[i] = FMA(c.dt, dU, CU); }     /* :233-234 */
            else         { out_F[i] = dF;               out_U[i] = dU; }
        }
    }
}

/* Scratch fields, the analogue of cache_alloc (reference cuda_alloc.cuh:72-169). */
static REAL* FN(orc_scratch)(orc_state* s, int slot)
{
    size_t N = (size_t) s->nx * (size_t) s->ny;
    if(s->scratch[slot] == NULL)
        s->scratch[slot] = calloc(N, sizeof(double));
    return (REAL*) s->scratch[slot];
}

/* explicit_solver_euler_step_based: reference simulation.cu:283-311. */
static void FN(orc_euler_step_based)(orc_state* s, const pf_params* p, const REAL* F, const REAL* U, const REAL* U_base,
                                     REAL* next_F, REAL* next_U, pf_step_info* info)
{
    REAL one = 1;
    if(U == U_base)
    {
        const REAL* Fs[1] = {F}; const REAL* Us[1] = {U};
        FN(orc_rhs)(p, 1, 1, &one, Fs, Us, next_F, next_U);
    }
    else
    {
        size_t N = (size_t) p->nx * (size_t) p->ny;
        REAL* dt_F = FN(orc_scratch)(s, 0);
        REAL* dt_U = FN(orc_scratch)(s, 1);
        REAL dt = (REAL) p->dt;
        const REAL* Fs[1] = {F}; const REAL* Us[1] = {U};
        FN(orc_rhs)(p, 0, 1, &one, Fs, Us, dt_F, dt_U);
        #pragma omp parallel for schedule(static)
        for(size_t i = 0; i < N; i++)
        {
            next_F[i] = FMA(dt, dt_F[i], F[i]);            /* :299 */
            next_U[i] = FMA(dt, dt_U[i], U_base[i]);       /* :300 */
        }
    }
    if(info) { info->Phi_iters = 1; info->T_iters = 1; }
}

/* explicit_solver_rk4_step: reference simulation.cu:313-348. */
static void FN(orc_rk4_step)(orc_state* s, const pf_params* p, const REAL* F, const REAL* U, REAL* next_F, REAL* next_U, pf_step_info* info)
{
    size_t N = (size_t) p->nx * (size_t) p->ny;
    REAL* k1_F = FN(orc_scratch)(s, 0); REAL* k2_F = FN(orc_scratch)(s, 1); REAL* k3_F = FN(orc_scratch)(s, 2); REAL* k4_F = FN(orc_scratch)(s, 3);
    REAL* k1_U = FN(orc_scratch)(s, 4); REAL* k2_U = FN(orc_scratch)(s, 5); REAL* k3_U = FN(orc_scratch)(s, 6); REAL* k4_U = FN(orc_scratch)(s, 7);
    REAL dt = (REAL) p->dt;
    {
        REAL w[1] = {1}; const REAL* Fs[1] = {F}; const REAL* Us[1] = {U};
        FN(orc_rhs)(p, 0, 1, w, Fs, Us, k1_F, k1_U);                       /* :330 */
    }
    {
        REAL w[2] = {1, dt / 2}; const REAL* Fs[2] = {F, k1_F}; const REAL* Us[2] = {U, k1_U};
        FN(orc_rhs)(p, 0, 2, w, Fs, Us, k2_F, k2_U);                       /* :331 */
    }
    {
        REAL w[2] = {1, dt / 2}; const REAL* Fs[2] = {F, k2_F}; const REAL* Us[2] = {U, k2_U};
        FN(orc_rhs)(p, 0, 2, w, Fs, Us, k3_F, k3_U);                       /* :332 */
    }
    {
        REAL w[2] = {1, dt}; const REAL* Fs[2] = {F, k3_F}; const REAL* Us[2] = {U, k3_U};
        FN(orc_rhs)(p, 0, 2, w, Fs, Us, k4_F, k4_U);                       /* :333 */
    }
    REAL dt6 = dt / 6;
    #pragma omp parallel for schedule(static)
    for(size_t i = 0; i < N; i++)
    {
        /* :336-337  F + dt/6*(k1 + 2*k2 + 2*k3 + k4); 2*k is exact so the inner fusions do not change the value */
        next_F[i] = FMA(dt6, ((k1_F[i] + 2 * k2_F[i]) + 2 * k3_F[i]) + k4_F[i], F[i]);
        next_U[i] = FMA(dt6, ((k1_U[i] + 2 * k2_U[i]) + 2 * k3_U[i]) + k4_U[i], U[i]);
    }
    if(info) { info->Phi_iters = 1; info->T_iters = 1; }
}

/* explicit_solver_rk4_adaptive_step (Runge-Kutta-Merson, RKM_ADAPTIVE_MODE 2): reference simulation.cu:350-497. */
static double FN(orc_rk4_adaptive_step)(orc_state* s, const pf_params* p, const REAL* F, const REAL* U, REAL* next_F, REAL* next_U, pf_step_info* info)
{
    size_t N = (size_t) p->nx * (size_t) p->ny;
    if(s->initial_step <= 0)                                                /* :363-365 (function-static in the reference) */
        s->initial_step = (double)(REAL) p->dt;
    REAL tau = (REAL) s->initial_step;
    REAL epsilon_F = 0, epsilon_U = 0;

    REAL* k1_F = FN(orc_scratch)(s, 0); REAL* k2_F = FN(orc_scratch)(s, 1); REAL* k3_F = FN(orc_scratch)(s, 2);
    REAL* k4_F = FN(orc_scratch)(s, 3); REAL* k5_F = FN(orc_scratch)(s, 4);
    REAL* k1_U = FN(orc_scratch)(s, 5); REAL* k2_U = FN(orc_scratch)(s, 6); REAL* k3_U = FN(orc_scratch)(s, 7);
    REAL* k4_U = FN(orc_scratch)(s, 8); REAL* k5_U = FN(orc_scratch)(s, 9);
    {
        REAL w[1] = {1}; const REAL* Fs[1] = {F}; const REAL* Us[1] = {U};
        FN(orc_rhs)(p, 0, 1, w, Fs, Us, k1_F, k1_U);                       /* :386 */
    }
    int converged = 0;
    int i = 0;
    int max_iters = p->T_max_iters > p->Phi_max_iters ? p->T_max_iters : p->Phi_max_iters;
    if(max_iters < 1) max_iters = 1;                                        /* :390 */
    REAL used_tau = tau;
    for(; i < max_iters && converged == 0; i++)
    {
        {
            REAL w[2] = {1, tau / 3}; const REAL* Fs[2] = {F, k1_F}; const REAL* Us[2] = {U, k1_U};
            FN(orc_rhs)(p, 0, 2, w, Fs, Us, k2_F, k2_U);                   /* :406 */
        }
        {
            REAL w[3] = {1, tau / 6, tau / 6}; const REAL* Fs[3] = {F, k1_F, k2_F}; const REAL* Us[3] = {U, k1_U, k2_U};
            FN(orc_rhs)(p, 0, 3, w, Fs, Us, k3_F, k3_U);                   /* :407 */
        }
        {
            REAL w[3] = {1, tau / 8, tau * 3 / 8}; const REAL* Fs[3] = {F, k1_F, k3_F}; const REAL* Us[3] = {U, k1_U, k3_U};
            FN(orc_rhs)(p, 0, 3, w, Fs, Us, k4_F, k4_U);                   /* :408 */
        }
        {
            REAL w[4] = {1, tau / 2, -tau * 3 / 2, tau * 2};
            const REAL* Fs[4] = {F, k1_F, k3_F, k4_F}; const REAL* Us[4] = {U, k1_U, k3_U, k4_U};
            FN(orc_rhs)(p, 0, 4, w, Fs, Us, k5_F, k5_U);                   /* :409 */
        }
        /* :427-435 max norm of 0.2*k1 - 0.9*k3 + 0.8*k4 - 0.1*k5; Reduce::Max starts from its identity
         * MIN(numeric_limits::min(), -inf) = -inf (cuda_reduction.cuh:421-423) and folds |.| values, i.e. plain max. */
        REAL max_F = -INFINITY, max_U = -INFINITY;
        #pragma omp parallel for schedule(static) reduction(max:max_F) reduction(max:max_U)
        for(size_t j = 0; j < N; j++)
        {
            /* nvcc contracts a*b - c*d + ... as fma(.., fma(.., fma(0.2, k1, -(0.9*k3)))): the rounded product is 0.9*k3
             * (SASS of cuda_produce_reduce_kernel<Reduce::Max, ..., explicit_solver_rk4_adaptive_step lambda>: DMUL k3, DFMA k1, DFMA k4, DFMA k5) */
            REAL eF = FMA(-(REAL) 0.1, k5_F[j], FMA((REAL) 0.8, k4_F[j], FMA((REAL) 0.2, k1_F[j], -((REAL) 0.9 * k3_F[j]))));
            REAL eU = FMA(-(REAL) 0.1, k5_U[j], FMA((REAL) 0.8, k4_U[j], FMA((REAL) 0.2, k1_U[j], -((REAL) 0.9 * k3_U[j]))));
            eF = eF >= 0 ? eF : -eF;
            eU = eU >= 0 ? eU : -eU;
            if(eF > max_F) max_F = eF;
            if(eU > max_U) max_U = eU;
        }
        epsilon_F = tau / 3 * max_F;                                        /* :437-438 */
        epsilon_U = tau / 3 * max_U;

        if((double) epsilon_F < p->Phi_tolerance && (double) epsilon_U < p->T_tolerance)   /* :456 (Real promoted to double) */
            converged = 1;

        /* :459-463. MAX/MIN mix Real and double operands, so the macros evaluate in double and the result is cast to Real. */
        double e_max = (double)(epsilon_F > epsilon_U ? epsilon_F : epsilon_U);
        REAL epsilon = (REAL)(e_max > 1e-20 ? e_max : 1e-20);
        double d_min = p->Phi_tolerance < p->T_tolerance ? p->Phi_tolerance : p->T_tolerance;
        REAL delta = (REAL)(d_min > 1e-20 ? d_min : 1e-20);
        used_tau = tau;
        tau = POW(delta / epsilon, (REAL) 0.2) * 4 / 5 * tau;
        tau = (REAL)((double) tau > p->min_dt ? (double) tau : p->min_dt);

        if((double) tau <= p->min_dt && (double) used_tau <= p->min_dt)     /* :466 */
        {
            /* the reference `break`s here, skipping the loop's i++ */
            break;
        }
    }
    REAL ut6 = used_tau / 6;
    #pragma omp parallel for schedule(static)
    for(size_t j = 0; j < N; j++)
    {
        next_F[j] = FMA(ut6, (k1_F[j] + 4 * k4_F[j]) + k5_F[j], F[j]);     /* :472 */
        next_U[j] = FMA(ut6, (k1_U[j] + 4 * k4_U[j]) + k5_U[j], U[j]);     /* :475 */
    }
    s->initial_step = (double) tau;                                         /* :486 */
    if(info)
    {
        info->Phi_iters = i; info->T_iters = i;                             /* :488-492 */
        info->Phi_converged = converged; info->T_converged = converged;
        info->Phi_error = (double) epsilon_F; info->T_error = (double) epsilon_U;
        info->next_tau = (double) tau;
    }
    return (double) used_tau;
}

/* Matrix-free operators of the semi-implicit solver. */
typedef struct FN(orc_cross_static) { REAL C, U, D, L, R; int nx, ny, boundary; } FN(orc_cross_static);       /* :499-512 */
typedef struct FN(orc_aniso_matrix) { const REAL* anisotrophy; REAL X, Y, C_minus_one; int nx, ny, boundary; } FN(orc_aniso_matrix); /* :514-525 */

/* cross_matrix_static_multiply: reference simulation.cu:528-549. */
static void FN(orc_cross_matrix_static_multiply)(REAL* out, const void* A_, const REAL* vec, int N)
{
    (void) N;
    const FN(orc_cross_static)* A = (const FN(orc_cross_static)*) A_;
    int nx = A->nx, ny = A->ny;
    #pragma omp parallel for schedule(static)
    for(int y = 0; y < ny; y++)
        for(int x = 0; x < nx; x++)
        {
            REAL val = vec[(size_t) x + (size_t) y * nx] * A->C;
            val = FMA(FN(orc_boundary_sample)(vec, x + 1, y, nx, ny, A->boundary), A->R, val);
            val = FMA(FN(orc_boundary_sample)(vec, x - 1, y, nx, ny, A->boundary), A->L, val);
            val = FMA(FN(orc_boundary_sample)(vec, x, y + 1, nx, ny, A->boundary), A->U, val);
            val = FMA(FN(orc_boundary_sample)(vec, x, y - 1, nx, ny, A->boundary), A->D, val);
            out[(size_t) x + (size_t) y * nx] = val;
        }
}

/* anisotrophy_matrix_multiply: reference simulation.cu:551-578. */
static void FN(orc_anisotrophy_matrix_multiply)(REAL* out, const void* A_, const REAL* vec, int N)
{
    (void) N;
    const FN(orc_aniso_matrix)* A = (const FN(orc_aniso_matrix)*) A_;
    int nx = A->nx, ny = A->ny;
    #pragma omp parallel for schedule(static)
    for(int y = 0; y < ny; y++)
        for(int x = 0; x < nx; x++)
        {
            size_t i = (size_t) x + (size_t) y * nx;
            REAL sc = A->anisotrophy[i];
            REAL X = A->X * sc;
            REAL Y = A->Y * sc;
            REAL C = FMA(A->C_minus_one, sc, (REAL) 1);
            REAL val = vec[i] * C;
            val = FMA(FN(orc_boundary_sample)(vec, x + 1, y, nx, ny, A->boundary), X, val);
            val = FMA(FN(orc_boundary_sample)(vec, x - 1, y, nx, ny, A->boundary), X, val);
            val = FMA(FN(orc_boundary_sample)(vec, x, y + 1, nx, ny, A->boundary), Y, val);
            val = FMA(FN(orc_boundary_sample)(vec, x, y - 1, nx, ny, A->boundary), Y, val);
            out[i] = val;
        }
}

/* cuda_dot_product (reference cuda_reduction.cuh:325-332). The reference's summation tree depends on its
 * launch geometry; the oracle sums each grid row left to right and then the row sums in order, which is
 * deterministic for any thread count. Parity on CG results is therefore by tolerance, not bit-exact. */
static REAL FN(orc_dot)(const REAL* a, const REAL* b, int N, int row)
{
    if(row <= 0) row = N;
    int rows = (N + row - 1) / row;
    REAL total = 0;
    REAL* partial = (REAL*) malloc(sizeof(REAL) * (size_t) rows);
    #pragma omp parallel for schedule(static)
    for(int r = 0; r < rows; r++)
    {
        REAL acc = 0;
        int end = (r + 1) * row < N ? (r + 1) * row : N;
        for(int i = r * row; i < end; i++)
            acc = FMA(a[i], b[i], acc);
        partial[r] = acc;
    }
    for(int r = 0; r < rows; r++)
        total += partial[r];
    free(partial);
    return total;
}

typedef void (*FN(orc_matvec))(REAL* out, const void* A, const REAL* x, int N);

/* conjugate_gradient_solve: reference simulation.cu:596-690. */
static void FN(orc_cg_solve)(const void* A, REAL* x, const REAL* b, int N, int row, FN(orc_matvec) mul,
                             REAL epsilon, REAL tolerance, int max_iters, const REAL* x0,
                             REAL* r, REAL* pv, REAL* Ap, int* iters_out, int* converged_out, REAL* error_out)
{
    REAL scaled_squared_tolerance = tolerance * tolerance * N;             /* :608 */
    REAL r_dot_r = 0;
    if(x0)
    {
        memcpy(x, x0, sizeof(REAL) * (size_t) N);                          /* :630 */
        mul(Ap, A, x0, N);                                                  /* :631 */
        #pragma omp parallel for schedule(static)
        for(int i = 0; i < N; i++) { r[i] = b[i] - Ap[i]; pv[i] = r[i]; }  /* :632-635 */
        r_dot_r = FN(orc_dot)(r, r, N, row);
    }
    else
    {
        memset(x, 0, sizeof(REAL) * (size_t) N);
        memcpy(r, b, sizeof(REAL) * (size_t) N);
        memcpy(pv, b, sizeof(REAL) * (size_t) N);
        r_dot_r = FN(orc_dot)(b, b, N, row);
    }
    int iter = 0;
    if(r_dot_r >= scaled_squared_tolerance || iter == 0)                    /* :650 (always true) */
    {
        for(; iter < max_iters; iter++)
        {
            mul(Ap, A, pv, N);
            REAL p_dot_Ap = FN(orc_dot)(pv, Ap, N, row);
            REAL alpha = r_dot_r / (p_dot_Ap > epsilon ? p_dot_Ap : epsilon);      /* :657 */
            #pragma omp parallel for schedule(static)
            for(int i = 0; i < N; i++)
            {
                x[i] = FMA(alpha, pv[i], x[i]);                             /* :660 */
                r[i] = FMA(-alpha, Ap[i], r[i]);                            /* :661 */
            }
            REAL r_dot_r_new = FN(orc_dot)(r, r, N, row);
            if(r_dot_r_new < scaled_squared_tolerance)                      /* :665 */
            {
                r_dot_r = r_dot_r_new;
                break;
            }
            REAL beta = r_dot_r_new / (r_dot_r > epsilon ? r_dot_r : epsilon);     /* :671 */
            #pragma omp parallel for schedule(static)
            for(int i = 0; i < N; i++)
                pv[i] = FMA(beta, pv[i], r[i]);                             /* :673 */
            r_dot_r = r_dot_r_new;
        }
    }
    *iters_out = iter;
    *converged_out = iter != max_iters;
    *error_out = SQRT(r_dot_r / N);
}

/* semi_implicit_solver_step_based: reference simulation.cu:732-926 (without the do_debug back-substitution :910-923). */
static void FN(orc_semi_implicit_step_based)(orc_state* s, const pf_params* p, const REAL* F, const REAL* U, const REAL* U_base,
                                             REAL* next_F, REAL* next_U, pf_step_info* info)
{
    FN(orc_consts) c = FN(orc_make_consts)(p);
    int nx = c.nx, ny = c.ny;
    int N = nx * ny;
    REAL dx = c.dx, dy = c.dy, dt = c.dt, gamma = c.gamma, L = c.L, Tm = c.Tm;
    REAL* b_F = FN(orc_scratch)(s, 0);
    REAL* b_U = FN(orc_scratch)(s, 1);
    REAL* anisotropy = FN(orc_scratch)(s, 2);
    REAL* cg_r = FN(orc_scratch)(s, 3);
    REAL* cg_p = FN(orc_scratch)(s, 4);
    REAL* cg_Ap = FN(orc_scratch)(s, 5);

    FN(orc_aniso_matrix) A_F;                                               /* :772-780 */
    A_F.anisotrophy = anisotropy;
    A_F.C_minus_one = 2 * dt / (dx * dx) + 2 * dt / (dy * dy);
    A_F.X = -dt / (dx * dx);
    A_F.Y = -dt / (dy * dy);
    A_F.nx = nx; A_F.ny = ny; A_F.boundary = p->Phi_boundary;

    FN(orc_cross_static) A_U;                                               /* :782-791 */
    A_U.C = 1 + 2 * dt / (dx * dx) + 2 * dt / (dy * dy);
    A_U.R = -dt / (dx * dx);
    A_U.L = -dt / (dx * dx);
    A_U.U = -dt / (dy * dy);
    A_U.D = -dt / (dy * dy);
    A_U.nx = nx; A_U.ny = ny; A_U.boundary = p->T_boundary;

    const REAL one_minus_gamma = 1 - gamma;
    #pragma omp parallel for schedule(static)
    for(int y = 0; y < ny; y++)
        for(int x = 0; x < nx; x++)
        {
            size_t i = (size_t) x + (size_t) y * nx;
            REAL C_Phi = F[i];
            REAL E_Phi = FN(orc_boundary_sample)(F, x + 1, y, nx, ny, c.F_bound);
            REAL W_Phi = FN(orc_boundary_sample)(F, x - 1, y, nx, ny, c.F_bound);
            REAL N_Phi = FN(orc_boundary_sample)(F, x, y + 1, nx, ny, c.F_bound);
            REAL S_Phi = FN(orc_boundary_sample)(F, x, y - 1, nx, ny, c.F_bound);
            REAL two_Phi = 2 * C_Phi;
            REAL laplace_Phi = FMA((W_Phi - two_Phi) + E_Phi, c.one_over_dx2, ((S_Phi - two_Phi) + N_Phi) * c.one_over_dy2);
            REAL f0 = C_Phi * (1 - C_Phi) * (C_Phi - (REAL) 0.5);
            if(c.do_corrector_guess)
            {
                /* :800-834 */
                REAL C_T = U[i];
                REAL E_T = FN(orc_boundary_sample)(U, x + 1, y, nx, ny, c.U_bound);
                REAL W_T = FN(orc_boundary_sample)(U, x - 1, y, nx, ny, c.U_bound);
                REAL N_T = FN(orc_boundary_sample)(U, x, y + 1, nx, ny, c.U_bound);
                REAL S_T = FN(orc_boundary_sample)(U, x, y - 1, nx, ny, c.U_bound);
                REAL grad_Phi_x = (E_Phi - W_Phi) * c.one_over_2dx;
                REAL grad_Phi_y = (N_Phi - S_Phi) * c.one_over_2dy;         /* 2dy here, :814 */
                REAL grad_Phi_norm = FN(orc_custom_hypot)(grad_Phi_x, grad_Phi_y);
                REAL theta = FN(orc_custom_atan2)(grad_Phi_y, grad_Phi_x);
                REAL g_theta = FMA(-FN(orc_custom_cos)(FMA(c.m0, theta, c.theta0)), c.S0, (REAL) 1);
                REAL two_T = 2 * C_T;
                REAL laplace_T = FMA((W_T - two_T) + E_T, c.one_over_dx2, ((S_T - two_T) + N_T) * c.one_over_dy2);
                /* (1-gamma)*k1*laplace + k0: here nvcc rounds the first product and fuses k0 = (g*f0)*k0_factor
                 * (SASS of the two prepare kernels: DMUL lap*(k1*(1-gamma)), DFMA (g*f0)*k0_factor + that) */
                REAL k2 = grad_Phi_norm * c.k2_factor;
                REAL k1 = g_theta * c.k1_factor;
                REAL corr = FMA(k2, c.dt_L, (REAL) 1);
                REAL inner = FMA(-k2, FMA(dt, laplace_T, C_T - Tm), FMA(g_theta * f0, c.k0_factor, (k1 * one_minus_gamma) * laplace_Phi));
                REAL right = FMA(dt / corr, inner, C_Phi);                  /* :828 */
                REAL factor = gamma / corr * k1;                            /* :829 */
                anisotropy[i] = factor;
                b_F[i] = right;
            }
            else
            {
                /* :838-870 */
                REAL C_T = U[i];
                REAL grad_Phi_x = (E_Phi - W_Phi) / (2 * dx);               /* division, :850-851 */
                REAL grad_Phi_y = (N_Phi - S_Phi) / (2 * dy);
                REAL grad_Phi_norm = FN(orc_custom_hypot)(grad_Phi_x, grad_Phi_y);
                REAL theta = FN(orc_custom_atan2)(grad_Phi_y, grad_Phi_x);
                REAL g_theta = FMA(-FN(orc_custom_cos)(FMA(c.m0, theta, c.theta0)), c.S0, (REAL) 1);
                REAL k2 = grad_Phi_norm * c.k2_factor;
                REAL k1 = g_theta * c.k1_factor;
                REAL inner = FMA(-k2, C_T - Tm, FMA(g_theta * f0, c.k0_factor, (k1 * one_minus_gamma) * laplace_Phi));
                REAL right = FMA(dt, inner, C_Phi);                         /* :864 */
                REAL factor = gamma * k1;                                   /* :865 */
                anisotropy[i] = factor;
                b_F[i] = right;
            }
        }

    int it_F = 0, conv_F = 0, it_U = 0, conv_U = 0; REAL err_F = 0, err_U = 0;
    /* :881-888  Solve A_F*next_F = b_F, x0 = F */
    FN(orc_cg_solve)(&A_F, next_F, b_F, N, nx, FN(orc_anisotrophy_matrix_multiply), (REAL) 1.0e-12, (REAL) p->Phi_tolerance, p->Phi_max_iters, F,
                     cg_r, cg_p, cg_Ap, &it_F, &conv_F, &err_F);

    /* :893-899 */
    const REAL dt_omg = dt * one_minus_gamma;
    #pragma omp parallel for schedule(static)
    for(int i = 0; i < N; i++)
    {
        REAL T = U_base[i];
        b_U[i] = FMA(dt_omg, T, FMA(L, next_F[i] - F[i], T));
    }
    /* :901-906  Solve A_U*next_U = b_U, x0 = U */
    FN(orc_cg_solve)(&A_U, next_U, b_U, N, nx, FN(orc_cross_matrix_static_multiply), (REAL) 1.0e-12, (REAL) p->T_tolerance, p->T_max_iters, U,
                     cg_r, cg_p, cg_Ap, &it_U, &conv_U, &err_U);
    if(info)
    {
        info->Phi_iters = it_F; info->T_iters = it_U;
        info->Phi_converged = conv_F; info->T_converged = conv_U;
        info->Phi_error = (double) err_F; info->T_error = (double) err_U;
    }
}

/* cuda_stats_delta: reference cuda_reduction.cuh:391-407 (sums row by row, see orc_dot). */
static void FN(orc_stats_delta)(const REAL* a, const REAL* b, size_t N, pf_delta_stats* out)
{
    REAL sum = 0, L1 = 0, L2 = 0;
    REAL mn = INFINITY, mx = -INFINITY;
    for(size_t i = 0; i < N; i++)
    {
        REAL v = a[i] - b[i];
        sum += v;
        L1 += v >= 0 ? v : -v;
        L2 = FMA(v, v, L2);
        if(v < mn) mn = v;
        if(v > mx) mx = v;
    }
    out->sum = (double) sum; out->L1 = (double) L1; out->L2 = (double)(REAL) sqrt((double) L2);
    out->min = (double) mn; out->max = (double) mx;
}

/* semi_implicit_and_euler_solver_step_corrector: reference simulation.cu:928-1008. */
static void FN(orc_step_corrector)(orc_state* s, const pf_params* p, const REAL* F, const REAL* U, REAL* next_F, REAL* next_U, pf_step_info* info)
{
    size_t N = (size_t) p->nx * (size_t) p->ny;
    REAL* temp_F = FN(orc_scratch)(s, 10);
    REAL* temp_U = FN(orc_scratch)(s, 11);
    REAL* steps_F[2]; REAL* steps_U[2];
    if(s->last_num_steps % 2 == 0) { steps_F[0] = next_F; steps_U[0] = next_U; steps_F[1] = temp_F; steps_U[1] = temp_U; }
    else                           { steps_F[1] = next_F; steps_U[1] = next_U; steps_F[0] = temp_F; steps_U[0] = temp_U; }

    size_t max_iters = (size_t) p->corrector_max_iters;
    if(!p->do_corrector_loop) max_iters = 0;
    if(max_iters == 0 && p->do_stats_step_residual) max_iters = 1;

    if(p->solver == PF_SOLVER_EXPLICIT_EULER)
        FN(orc_euler_step_based)(s, p, F, U, U, steps_F[0], steps_U[0], info);
    else
        FN(orc_semi_implicit_step_based)(s, p, F, U, U, steps_F[0], steps_U[0], info);
    for(size_t k = 0; k < max_iters; k++)
    {
        int curr = (int)(k % 2), next = (int)((k + 1) % 2);
        if(p->solver == PF_SOLVER_EXPLICIT_EULER)
            FN(orc_euler_step_based)(s, p, F, steps_U[curr], U, steps_F[next], steps_U[next], info);
        else
            FN(orc_semi_implicit_step_based)(s, p, F, steps_U[curr], U, steps_F[next], steps_U[next], info);
        if(p->do_stats_step_residual && info && k < PF_MAX_STEP_RESIDUALS)
        {
            FN(orc_stats_delta)(steps_F[curr], steps_F[next], N, &info->step_res[k]);
            info->step_res_count = (int) max_iters;
        }
        s->last_num_steps = (int) k;                                        /* :995 (sic: k, not k+1) */
    }
    int final_step = ((s->last_num_steps % 2) + 2) % 2;                     /* :999 */
    if(steps_F[final_step] != next_F)
    {
        memcpy(next_F, steps_F[final_step], N * sizeof(REAL));
        memcpy(next_U, steps_U[final_step], N * sizeof(REAL));
    }
}

/* sim_step: reference simulation.cu:1091-1156. */
static double FN(orc_step)(orc_state* s, const pf_params* p, const REAL* F, const REAL* U, REAL* next_F, REAL* next_U, pf_step_info* info)
{
    double dt = 0;
    pf_step_info local; memset(&local, 0, sizeof local);
    if(info == NULL) info = &local; else memset(info, 0, sizeof *info);
    switch(p->solver)
    {
        case PF_SOLVER_SEMI_IMPLICIT:
        case PF_SOLVER_EXPLICIT_EULER:
            FN(orc_step_corrector)(s, p, F, U, next_F, next_U, info); dt = p->dt; break;
        case PF_SOLVER_EXPLICIT_RK4:
            FN(orc_rk4_step)(s, p, F, U, next_F, next_U, info); dt = p->dt; break;
        case PF_SOLVER_EXPLICIT_RK4_ADAPTIVE:
            dt = FN(orc_rk4_adaptive_step)(s, p, F, U, next_F, next_U, info); break;
        case PF_SOLVER_EXACT:
            FN(orc_exact_step)(p, next_F, next_U); dt = p->dt; break;            /* :1118-1121 */
        default: break;
    }
    if(p->do_stats)
    {
        size_t N = (size_t) p->nx * (size_t) p->ny;
        FN(orc_stats_delta)(F, next_F, N, &info->Phi_delta);                /* :1128-1129 */
        FN(orc_stats_delta)(U, next_U, N, &info->T_delta);
    }
    info->dt_advanced = dt;
    return dt;
}

/* calc_debug_values: reference simulation.cu:245-281. */
static void FN(orc_debug_values)(const pf_params* p, const REAL* F, const REAL* U, REAL* grad_F, REAL* grad_U, REAL* aniso)
{
    int nx = p->nx, ny = p->ny;
    REAL S0 = (REAL) p->S, m0 = (REAL) p->m0, theta0 = (REAL) p->theta0;
    #pragma omp parallel for schedule(static)
    for(int y = 0; y < ny; y++)
        for(int x = 0; x < nx; x++)
        {
            REAL Phi_U = FN(orc_boundary_sample)(F, x, y + 1, nx, ny, p->Phi_boundary);
            REAL Phi_D = FN(orc_boundary_sample)(F, x, y - 1, nx, ny, p->Phi_boundary);
            REAL Phi_R = FN(orc_boundary_sample)(F, x + 1, y, nx, ny, p->Phi_boundary);
            REAL Phi_L = FN(orc_boundary_sample)(F, x - 1, y, nx, ny, p->Phi_boundary);
            REAL T_U = FN(orc_boundary_sample)(U, x, y + 1, nx, ny, p->T_boundary);
            REAL T_D = FN(orc_boundary_sample)(U, x, y - 1, nx, ny, p->T_boundary);
            REAL T_R = FN(orc_boundary_sample)(U, x + 1, y, nx, ny, p->T_boundary);
            REAL T_L = FN(orc_boundary_sample)(U, x - 1, y, nx, ny, p->T_boundary);
            REAL gx = Phi_R - Phi_L, gy = Phi_U - Phi_D;
            REAL tx = T_R - T_L, ty = T_U - T_D;
            REAL theta = FN(orc_custom_atan2)(gy, gx);
            size_t i = (size_t) x + (size_t) y * nx;
            grad_F[i] = FN(orc_custom_hypot)(gx, gy);
            grad_U[i] = FN(orc_custom_hypot)(tx, ty);
            aniso[i] = FMA(-FN(orc_custom_cos)(FMA(m0, theta, theta0)), S0, (REAL) 1);
        }
}

#undef FMA
#undef SQRT
#undef POW
#undef EXPR
#undef ERFR
#undef HYPOTR
#undef FN
#undef CAT
#undef CAT_
