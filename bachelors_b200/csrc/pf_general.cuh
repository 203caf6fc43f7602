// pf_general.cuh - per-stage kernels: the general (any boundary mix, any blend) right-hand side,
// elementwise stage combinations, the semi-implicit prepare kernel and the matrix-free CG kernels.
//
// The explicit solvers normally run through pf_fused.cuh; the kernels here serve pf_rhs(), mixed
// Phi/T boundary types, the Euler corrector loop, and the whole semi-implicit solver.
#pragma once
#include <type_traits>
#include "pf_common.cuh"

namespace pf {

// boundary_sample: reference simulation.cu:33-85 (single-GPU full grid).
template <typename Real>
__device__ __forceinline__ Real bsample(const Real* __restrict__ map, int x, int y, int nx, int ny, int bound)
{
    if(bound == PF_BOUNDARY_PERIODIC)
    {
        if(x < 0) x += nx; else if(x >= nx) x -= nx;
        if(y < 0) y += ny; else if(y >= ny) y -= ny;
        return map[(size_t) x + (size_t) y * (size_t) nx];
    }
    const bool inside = x >= 0 && x < nx && y >= 0 && y < ny;
    const int cx = x < 0 ? 0 : (x >= nx ? nx - 1 : x);
    const int cy = y < 0 ? 0 : (y >= ny ? ny - 1 : y);
    const Real v = map[(size_t) cx + (size_t) cy * (size_t) nx];
    if(bound == PF_BOUNDARY_DIRICHLET_ZERO && !inside) return (Real) 0 - v;
    return v;
}

// The four stencil neighbours of (x, y). Interior cells (all but the one-cell rim of the grid) read them at fixed
// offsets; the rim goes through boundary_sample. Same values either way - the split only removes the index clamping
// (some 60 integer instructions per cell) from the cells that do not need it.
template <typename Real>
__device__ __forceinline__ void sample_ewns(const Real* __restrict__ map, int x, int y, int nx, int ny, int bound,
                                            Real& e, Real& w, Real& n, Real& s)
{
    if(x > 0 && x < nx - 1 && y > 0 && y < ny - 1)
    {
        const Real* __restrict__ c = map + ((size_t) x + (size_t) y * (size_t) nx);
        e = c[1]; w = c[-1]; n = c[nx]; s = c[-nx];
    }
    else
    {
        e = bsample(map, x + 1, y, nx, ny, bound);
        w = bsample(map, x - 1, y, nx, ny, bound);
        n = bsample(map, x, y + 1, nx, ny, bound);
        s = bsample(map, x, y - 1, nx, ny, bound);
    }
}

template <typename Real>
struct BlendArgs {
    int n_states;
    Real w[4];
    const Real* F[4];
    const Real* U[4];
};

// Gathered, blended stage input at (x, y): reference simulation.cu:189-200.
template <typename Real>
__device__ __forceinline__ void gather(const BlendArgs<Real>& b, int x, int y, int nx, int ny, int F_bound, int U_bound, Real& F, Real& U)
{
    F = 0; U = 0;
    #pragma unroll
    for(int i = 0; i < 4; i++)
        if(i < b.n_states)
        {
            F = pf_fma(bsample(b.F[i], x, y, nx, ny, F_bound), b.w[i], F);
            U = pf_fma(bsample(b.U[i], x, y, nx, ny, U_bound), b.w[i], U);
        }
}

// explicit_solver_solve_lin_combination<IS_EULER, States...>: reference simulation.cu:145-243.
template <typename Real, bool ANISO>
__global__ void __launch_bounds__(256)
k_rhs_general(const BlendArgs<Real> b, const KConsts<Real> k, int nx, int ny, int F_bound, int U_bound, int is_euler,
              Real* __restrict__ out_F, Real* __restrict__ out_U)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if(x >= nx || y >= ny) return;
    Real CF, CT, EF, ET, WF, WT, NF, NT, SF, ST;
    gather(b, x, y, nx, ny, F_bound, U_bound, CF, CT);
    gather(b, x + 1, y, nx, ny, F_bound, U_bound, EF, ET);
    gather(b, x - 1, y, nx, ny, F_bound, U_bound, WF, WT);
    gather(b, x, y + 1, nx, ny, F_bound, U_bound, NF, NT);
    gather(b, x, y - 1, nx, ny, F_bound, U_bound, SF, ST);
    Real dF, dT;
    rhs_cell<Real, ANISO>(k, CF, EF, WF, NF, SF, CT, ET, WT, NT, ST, dF, dT);
    const size_t i = (size_t) x + (size_t) y * (size_t) nx;
    if(is_euler) { out_F[i] = pf_fma(k.dt, dF, CF); out_U[i] = pf_fma(k.dt, dT, CT); }
    else         { out_F[i] = dF; out_U[i] = dT; }
}

// next_F = F + dt*dt_F ; next_U = U_base + dt*dt_U : reference simulation.cu:298-301.
template <typename Real>
__global__ void k_euler_corrector_axpy(size_t N, Real dt, const Real* __restrict__ F, const Real* __restrict__ U_base,
                                       const Real* __restrict__ dF, const Real* __restrict__ dU, Real* __restrict__ nF, Real* __restrict__ nU)
{
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < N; i += (size_t) gridDim.x * blockDim.x)
    {
        nF[i] = pf_fma(dt, dF[i], F[i]);
        nU[i] = pf_fma(dt, dU[i], U_base[i]);
    }
}

// y + dt/6*(k1 + 2*k2 + 2*k3 + k4): reference simulation.cu:335-338.
template <typename Real>
__global__ void k_rk4_combine(size_t N, Real dt6, const Real* __restrict__ y, const Real* __restrict__ k1, const Real* __restrict__ k2,
                              const Real* __restrict__ k3, const Real* __restrict__ k4, Real* __restrict__ out)
{
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < N; i += (size_t) gridDim.x * blockDim.x)
        out[i] = pf_fma(dt6, pf_fma((Real) 2, k3[i], pf_fma((Real) 2, k2[i], k1[i])) + k4[i], y[i]);
}

// Merson error max-norm (reference simulation.cu:427-435) and result (simulation.cu:471-476) for one field.
template <typename Real>
__global__ void k_rkm_error(size_t N, const Real* __restrict__ k1, const Real* __restrict__ k3, const Real* __restrict__ k4,
                            const Real* __restrict__ k5, unsigned long long* err_bits)
{
    Real m = 0;
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < N; i += (size_t) gridDim.x * blockDim.x)
    {
        Real e = pf_fma(-(Real) 0.1, k5[i], pf_fma((Real) 0.8, k4[i], pf_fma((Real) 0.2, k1[i], -((Real) 0.9 * k3[i]))));
        e = e >= 0 ? e : -e;
        m = e > m ? e : m;
    }
    m = warp_max(m);
    if((threadIdx.x & 31) == 0 && m > 0) atomic_max_nonneg(err_bits, m);
}

template <typename Real>
__global__ void k_rkm_combine(size_t N, Real tau6, const Real* __restrict__ y, const Real* __restrict__ k1, const Real* __restrict__ k4,
                              const Real* __restrict__ k5, Real* __restrict__ out)
{
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < N; i += (size_t) gridDim.x * blockDim.x)
        out[i] = pf_fma(tau6, pf_fma((Real) 4, k4[i], k1[i]) + k5[i], y[i]);
}

// calc_debug_values: reference simulation.cu:245-281.
template <typename Real>
__global__ void k_debug_values(const Real* __restrict__ F, const Real* __restrict__ U, Real* __restrict__ grad_F, Real* __restrict__ grad_U,
                               Real* __restrict__ aniso, int nx, int ny, int F_bound, int U_bound, KConsts<Real> k)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if(x >= nx || y >= ny) return;
    Real gx = bsample(F, x + 1, y, nx, ny, F_bound) - bsample(F, x - 1, y, nx, ny, F_bound);
    Real gy = bsample(F, x, y + 1, nx, ny, F_bound) - bsample(F, x, y - 1, nx, ny, F_bound);
    Real tx = bsample(U, x + 1, y, nx, ny, U_bound) - bsample(U, x - 1, y, nx, ny, U_bound);
    Real ty = bsample(U, x, y + 1, nx, ny, U_bound) - bsample(U, x, y - 1, nx, ny, U_bound);
    const size_t i = (size_t) x + (size_t) y * (size_t) nx;
    grad_F[i] = (Real) pf_hypotf_bf<false>((float) gx, (float) gy);
    grad_U[i] = (Real) pf_hypotf_bf<false>((float) tx, (float) ty);
    aniso[i] = aniso_g<Real>(k, gx, gy);
}

// ---- semi-implicit -------------------------------------------------------------------------------------

// The prepare kernel: right-hand side b_F and per-cell factor s of A_Phi. Reference simulation.cu:800-834 (guess)
// and :838-870 (no guess; note the DIVISION by 2*dx / 2*dy there).
template <typename Real, bool ANISO>
__device__ __forceinline__ void si_prepare_cell(const Real* __restrict__ F, const Real* __restrict__ U, int x, int y,
                                                int nx, int ny, int F_bound, int U_bound, const KConsts<Real>& k, Real& right_out, Real& factor_out)
{
    const size_t i = (size_t) x + (size_t) y * (size_t) nx;
    const Real C_Phi = F[i];
    Real E_Phi, W_Phi, N_Phi, S_Phi;
    sample_ewns(F, x, y, nx, ny, F_bound, E_Phi, W_Phi, N_Phi, S_Phi);
    const Real two_Phi = 2 * C_Phi;
    const Real laplace_Phi = pf_fma((W_Phi - two_Phi) + E_Phi, k.one_over_dx2, ((S_Phi - two_Phi) + N_Phi) * k.one_over_dy2);
    const Real f0 = C_Phi * (1 - C_Phi) * (C_Phi - (Real) 0.5);
    const Real one_minus_gamma = 1 - k.gamma;
    const Real C_T = U[i];
    Real gx, gy;
    if(k.do_corrector_guess) { gx = (E_Phi - W_Phi) * k.one_over_2dx; gy = (N_Phi - S_Phi) * k.one_over_2dy; }
    else                     { gx = (E_Phi - W_Phi) / (2 * k.dx);     gy = (N_Phi - S_Phi) / (2 * k.dy); }
    const Real grad_Phi_norm = (Real) pf_hypotf_bf<false>((float) gx, (float) gy);
    Real gf0, k1;
    if(ANISO) { Real g = aniso_g<Real>(k, gx, gy); gf0 = g * f0; k1 = g * k.k1_factor; }
    else      { gf0 = f0; k1 = k.k1_factor; }
    const Real k2 = grad_Phi_norm * k.k2_factor;
    // here the reference's build rounds (1-gamma)*k1*laplace and fuses k0 = (g*f0)*k0_factor (oracle/pf_oracle_impl.h)
    const Real base = pf_fma(gf0, k.k0_factor, (k1 * one_minus_gamma) * laplace_Phi);
    Real right, factor;
    if(k.do_corrector_guess)
    {
        Real E_T, W_T, N_T, S_T;
        sample_ewns(U, x, y, nx, ny, U_bound, E_T, W_T, N_T, S_T);
        const Real two_T = 2 * C_T;
        const Real laplace_T = pf_fma((W_T - two_T) + E_T, k.one_over_dx2, ((S_T - two_T) + N_T) * k.one_over_dy2);
        const Real corr = pf_fma(k2, k.dt_L, (Real) 1);
        const Real inner = pf_fma(-k2, pf_fma(k.dt, laplace_T, C_T - k.Tm), base);
        right = pf_fma(k.dt / corr, inner, C_Phi);
        factor = k.gamma / corr * k1;
    }
    else
    {
        const Real inner = pf_fma(-k2, C_T - k.Tm, base);
        right = pf_fma(k.dt, inner, C_Phi);
        factor = k.gamma * k1;
    }
    right_out = right; factor_out = factor;
}

template <typename Real, bool ANISO>
__global__ void __launch_bounds__(256)
k_si_prepare(const Real* __restrict__ F, const Real* __restrict__ U, Real* __restrict__ b_F, Real* __restrict__ factor_out,
             int nx, int ny, int F_bound, int U_bound, KConsts<Real> k, int y_begin = 0, int y_end = 0x7fffffff)
{
    // rows [y_begin, y_end) of the ny rows the pointers hold: a slab passes its allocation (ghost rows included, so
    // the y-neighbours of an owned row are plain memory) and the window of its owned rows
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = y_begin + blockIdx.y * blockDim.y + threadIdx.y;
    if(x >= nx || y >= ny || y >= y_end) return;
    const size_t i = (size_t) x + (size_t) y * (size_t) nx;
    Real right, factor;
    si_prepare_cell<Real, ANISO>(F, U, x, y, nx, ny, F_bound, U_bound, k, right, factor);
    factor_out[i] = factor;
    b_F[i] = right;
}

// b_U = T + L*(Phi_next - Phi) + dt*(1-gamma)*T : reference simulation.cu:893-899.
template <typename Real>
__global__ void k_si_bU(size_t N, Real L, Real dt_omg, const Real* __restrict__ U_base, const Real* __restrict__ F,
                        const Real* __restrict__ next_F, Real* __restrict__ b_U)
{
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < N; i += (size_t) gridDim.x * blockDim.x)
    {
        const Real T = U_base[i];
        b_U[i] = pf_fma(dt_omg, T, pf_fma(L, next_F[i] - F[i], T));
    }
}

// Matrix-free operators (reference simulation.cu:499-578). which = 0: A_Phi (per-cell scaled), 1: A_T (constant).
template <typename Real>
struct CgMatrix {
    int which;
    const Real* factor;          // A_Phi: "anisotrophy" array s
    Real X, Y, C_minus_one;      // A_Phi: -dt/dx^2, -dt/dy^2, 2dt/dx^2 + 2dt/dy^2   (simulation.cu:772-776)
    Real C, R, Lc, Uc, D;        // A_T                                              (simulation.cu:782-787)
    int nx, ny, boundary;
    int y_begin, y_end;          // rows the kernels produce (all of [0, ny) on a full grid; the owned rows of a slab)
};

// (A*vec)(x, y); `sc` is the per-cell factor of A_Phi at this cell (ignored for A_T).
template <typename Real>
__device__ __forceinline__ Real cg_apply_sc(const CgMatrix<Real>& A, const Real* __restrict__ vec, int x, int y, Real sc)
{
    const int nx = A.nx, ny = A.ny;
    const size_t i = (size_t) x + (size_t) y * (size_t) nx;
    const Real c = vec[i];
    Real e, w, n, s;
    sample_ewns(vec, x, y, nx, ny, A.boundary, e, w, n, s);
    Real val;
    if(A.which == 0)
    {
        const Real X = A.X * sc, Y = A.Y * sc;
        const Real Cc = pf_fma(A.C_minus_one, sc, (Real) 1);
        val = c * Cc;
        val = pf_fma(e, X, val); val = pf_fma(w, X, val); val = pf_fma(n, Y, val); val = pf_fma(s, Y, val);
    }
    else
    {
        val = c * A.C;
        val = pf_fma(e, A.R, val); val = pf_fma(w, A.Lc, val); val = pf_fma(n, A.Uc, val); val = pf_fma(s, A.D, val);
    }
    return val;
}

// boundary_sample of the vector r + beta*p that has not been stored yet (the cooperative step fuses the p update into
// the next SpMV): same index rule and same ghost value (0 - v outside a Dirichlet edge) as bsample, v formed on the fly.
template <typename Real>
__device__ __forceinline__ Real bsample_axpy(const Real* __restrict__ p, const Real* __restrict__ r, Real beta, int x, int y, int nx, int ny, int bound)
{
    if(bound == PF_BOUNDARY_PERIODIC)
    {
        if(x < 0) x += nx; else if(x >= nx) x -= nx;
        if(y < 0) y += ny; else if(y >= ny) y -= ny;
        const size_t i = (size_t) x + (size_t) y * (size_t) nx;
        return pf_fma(beta, p[i], r[i]);
    }
    const bool inside = x >= 0 && x < nx && y >= 0 && y < ny;
    const int cx = x < 0 ? 0 : (x >= nx ? nx - 1 : x);
    const int cy = y < 0 ? 0 : (y >= ny ? ny - 1 : y);
    const size_t i = (size_t) cx + (size_t) cy * (size_t) nx;
    const Real v = pf_fma(beta, p[i], r[i]);
    if(bound == PF_BOUNDARY_DIRICHLET_ZERO && !inside) return (Real) 0 - v;
    return v;
}

// (A*(r + beta*p))(x, y) with the centre value c = r + beta*p at (x, y) supplied by the caller: cg_apply_sc on the fly.
template <typename Real>
__device__ __forceinline__ Real cg_apply_axpy(const CgMatrix<Real>& A, const Real* __restrict__ p, const Real* __restrict__ r, Real beta,
                                              Real c, int x, int y, Real sc)
{
    const int nx = A.nx, ny = A.ny;
    Real e, w, n, s;
    if(x > 0 && x < nx - 1 && y > 0 && y < ny - 1)                         // interior: fixed offsets (see sample_ewns)
    {
        const size_t i = (size_t) x + (size_t) y * (size_t) nx;
        const Real* __restrict__ pc = p + i;
        const Real* __restrict__ rc = r + i;
        e = pf_fma(beta, pc[1], rc[1]);   w = pf_fma(beta, pc[-1], rc[-1]);
        n = pf_fma(beta, pc[nx], rc[nx]); s = pf_fma(beta, pc[-nx], rc[-nx]);
    }
    else
    {
        e = bsample_axpy(p, r, beta, x + 1, y, nx, ny, A.boundary);
        w = bsample_axpy(p, r, beta, x - 1, y, nx, ny, A.boundary);
        n = bsample_axpy(p, r, beta, x, y + 1, nx, ny, A.boundary);
        s = bsample_axpy(p, r, beta, x, y - 1, nx, ny, A.boundary);
    }
    Real val;
    if(A.which == 0)
    {
        const Real X = A.X * sc, Y = A.Y * sc;
        const Real Cc = pf_fma(A.C_minus_one, sc, (Real) 1);
        val = c * Cc;
        val = pf_fma(e, X, val); val = pf_fma(w, X, val); val = pf_fma(n, Y, val); val = pf_fma(s, Y, val);
    }
    else
    {
        val = c * A.C;
        val = pf_fma(e, A.R, val); val = pf_fma(w, A.Lc, val); val = pf_fma(n, A.Uc, val); val = pf_fma(s, A.D, val);
    }
    return val;
}

template <typename Real>
__device__ __forceinline__ Real cg_apply(const CgMatrix<Real>& A, const Real* __restrict__ vec, int x, int y)
{
    return cg_apply_sc(A, vec, x, y, A.which == 0 ? A.factor[(size_t) x + (size_t) y * (size_t) A.nx] : (Real) 0);
}

// Device-resident CG control block: no host round trip per iteration (the reference blocks on a D2H copy
// for every dot product, cuda_reduction.cuh:205-208).
struct CgState {
    int done;        // set when r.r < tol^2*N (simulation.cu:665)
    int iters;       // loop index at break, or max_iters (simulation.cu:681)
    double rr;       // r.r at exit
};

enum { CG_THREADS = 256, CG_MAX_BLOCKS = 1024 };

// Sum `n` block partials in a fixed order; every block computes the identical value.
template <typename Real>
__device__ __forceinline__ Real sum_partials(const Real* __restrict__ partials, int n, Real* scratch)
{
    Real acc = 0;
    for(int i = threadIdx.x; i < n; i += CG_THREADS) acc += partials[i];
    acc = block_sum(acc, scratch);
    __shared__ Real bcast;
    if(threadIdx.x == 0) bcast = acc;
    __syncthreads();
    Real r = bcast;
    __syncthreads();
    return r;
}

// r = b - A*x0, p = r, x = x0 (or x = 0, r = p = b): reference simulation.cu:628-647.
template <typename Real>
__global__ void __launch_bounds__(CG_THREADS)
k_cg_init(CgMatrix<Real> A, const Real* __restrict__ x0, const Real* __restrict__ b, Real* __restrict__ x, Real* __restrict__ r,
          Real* __restrict__ p, Real* __restrict__ rr_partials, CgState* st, int max_iters)
{
    __shared__ Real scratch[32];
    const int nx = A.nx;
    const int segs = (nx + CG_THREADS - 1) / CG_THREADS;
    const int items = segs * (A.y_end - A.y_begin);
    Real acc = 0;
    for(int wi = blockIdx.x; wi < items; wi += gridDim.x)
    {
        const int y = A.y_begin + wi / segs, xx = (wi % segs) * CG_THREADS + threadIdx.x;
        if(xx < nx)
        {
            const size_t i = (size_t) xx + (size_t) y * (size_t) nx;
            Real rv;
            if(x0) { rv = b[i] - cg_apply(A, x0, xx, y); x[i] = x0[i]; }
            else   { rv = b[i]; x[i] = 0; }
            r[i] = rv; p[i] = rv;
            acc = pf_fma(rv, rv, acc);
        }
    }
    acc = block_sum(acc, scratch);
    if(threadIdx.x == 0) rr_partials[blockIdx.x] = acc;
    if(blockIdx.x == 0 && threadIdx.x == 0) { st->done = 0; st->iters = max_iters; st->rr = 0; }
}

// Ap = A*p and the partial sums of p.Ap: reference simulation.cu:654-656.
template <typename Real>
__global__ void __launch_bounds__(CG_THREADS)
k_cg_spmv(CgMatrix<Real> A, const Real* __restrict__ p, Real* __restrict__ Ap, Real* __restrict__ pAp_partials, const CgState* st)
{
    if(st->done) return;
    __shared__ Real scratch[32];
    const int nx = A.nx;
    const int segs = (nx + CG_THREADS - 1) / CG_THREADS;
    const int items = segs * (A.y_end - A.y_begin);
    Real acc = 0;
    for(int wi = blockIdx.x; wi < items; wi += gridDim.x)
    {
        const int y = A.y_begin + wi / segs, xx = (wi % segs) * CG_THREADS + threadIdx.x;
        if(xx < nx)
        {
            const size_t i = (size_t) xx + (size_t) y * (size_t) nx;
            const Real v = cg_apply(A, p, xx, y);
            Ap[i] = v;
            acc = pf_fma(p[i], v, acc);
        }
    }
    acc = block_sum(acc, scratch);
    if(threadIdx.x == 0) pAp_partials[blockIdx.x] = acc;
}

// alpha = r.r / max(p.Ap, eps); x += alpha p; r -= alpha Ap; partial sums of the new r.r: reference simulation.cu:657-664.
template <typename Real>
__global__ void __launch_bounds__(CG_THREADS)
k_cg_update(size_t N, Real epsilon, Real* __restrict__ x, Real* __restrict__ r, const Real* __restrict__ p, const Real* __restrict__ Ap,
            const Real* __restrict__ rr_partials, const Real* __restrict__ pAp_partials, Real* __restrict__ rr_new_partials,
            int n_partials, const CgState* st)
{
    if(st->done) return;
    __shared__ Real scratch[32];
    const Real rr = sum_partials(rr_partials, n_partials, scratch);
    const Real pAp = sum_partials(pAp_partials, n_partials, scratch);
    const Real alpha = rr / (pAp > epsilon ? pAp : epsilon);
    Real acc = 0;
    for(size_t i = blockIdx.x * (size_t) CG_THREADS + threadIdx.x; i < N; i += (size_t) gridDim.x * CG_THREADS)
    {
        x[i] = pf_fma(alpha, p[i], x[i]);
        const Real rv = pf_fma(-alpha, Ap[i], r[i]);
        r[i] = rv;
        acc = pf_fma(rv, rv, acc);
    }
    acc = block_sum(acc, scratch);
    if(threadIdx.x == 0) rr_new_partials[blockIdx.x] = acc;
}

// Convergence test, beta = r.r_new / max(r.r, eps), p = r + beta p: reference simulation.cu:665-676.
template <typename Real>
__global__ void __launch_bounds__(CG_THREADS)
k_cg_pupdate(size_t N, Real epsilon, Real scaled_squared_tolerance, const Real* __restrict__ r, Real* __restrict__ p,
             const Real* __restrict__ rr_partials, const Real* __restrict__ rr_new_partials, int n_partials, CgState* st, int iter)
{
    if(st->done) return;
    __shared__ Real scratch[32];
    const Real rr = sum_partials(rr_partials, n_partials, scratch);
    const Real rr_new = sum_partials(rr_new_partials, n_partials, scratch);
    if(rr_new < scaled_squared_tolerance)
    {
        // every block takes this branch together (identical sums); block 0 publishes the result, the flag last.
        // A block that starts after the flag is set returns at the top of the kernel, which is the same outcome.
        if(blockIdx.x == 0 && threadIdx.x == 0) { st->rr = (double) rr_new; st->iters = iter; __threadfence(); atomicExch(&st->done, 1); }
        return;
    }
    const Real beta = rr_new / (rr > epsilon ? rr : epsilon);
    for(size_t i = blockIdx.x * (size_t) CG_THREADS + threadIdx.x; i < N; i += (size_t) gridDim.x * CG_THREADS)
        p[i] = pf_fma(beta, p[i], r[i]);
    if(blockIdx.x == 0 && threadIdx.x == 0) st->rr = (double) rr_new;
}

// r.r when the loop never ran (max_iters <= 0).
template <typename Real>
__global__ void __launch_bounds__(CG_THREADS)
k_cg_finalize(const Real* __restrict__ rr_partials, int n_partials, CgState* st)
{
    __shared__ Real scratch[32];
    const Real rr = sum_partials(rr_partials, n_partials, scratch);
    if(threadIdx.x == 0) st->rr = (double) rr;
}

// ---- statistics: cuda_stats_delta, reference cuda_reduction.cuh:391-407 -------------------------------
struct StatsPartial { double sum, L1, L2, mn, mx; };

template <typename Real>
__global__ void __launch_bounds__(CG_THREADS)
k_stats_delta(size_t N, const Real* __restrict__ a, const Real* __restrict__ b, StatsPartial* partials)
{
    __shared__ Real scratch[32];
    Real sum = 0, L1 = 0, L2 = 0, mn = (Real) INFINITY, mx = -(Real) INFINITY;
    auto take = [&](const Real v)
    {
        sum += v;
        L1 += v >= 0 ? v : -v;
        L2 = pf_fma(v, v, L2);
        mn = v < mn ? v : mn;
        mx = v > mx ? v : mx;
    };
    // 16-byte loads (fields come from cudaMalloc: aligned); a thread's values are taken in index order, so the result does not
    // depend on the load width
    constexpr int V = 16 / sizeof(Real);
    using Vec = typename std::conditional<sizeof(Real) == 8, double2, float4>::type;
    const bool vec_ok = ((reinterpret_cast<size_t>(a) | reinterpret_cast<size_t>(b)) & 15) == 0;
    const size_t nv = vec_ok ? N / V : 0, stride = (size_t) gridDim.x * CG_THREADS;
    for(size_t i = blockIdx.x * (size_t) CG_THREADS + threadIdx.x; i < nv; i += stride)
    {
        const Vec va = reinterpret_cast<const Vec*>(a)[i], vb = reinterpret_cast<const Vec*>(b)[i];
        const Real* ea = reinterpret_cast<const Real*>(&va); const Real* eb = reinterpret_cast<const Real*>(&vb);
        #pragma unroll
        for(int k = 0; k < V; k++) take(ea[k] - eb[k]);
    }
    for(size_t i = nv * V + blockIdx.x * (size_t) CG_THREADS + threadIdx.x; i < N; i += stride)
        take(a[i] - b[i]);
    sum = block_sum(sum, scratch);
    L1 = block_sum(L1, scratch);
    L2 = block_sum(L2, scratch);
    mn = warp_min(mn); mx = warp_max(mx);
    __shared__ Real smn[32], smx[32];
    __syncthreads();
    if((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if(threadIdx.x == 0)
    {
        for(int w = 1; w < CG_THREADS / 32; w++) { mn = smn[w] < mn ? smn[w] : mn; mx = smx[w] > mx ? smx[w] : mx; }
        StatsPartial o; o.sum = (double) sum; o.L1 = (double) L1; o.L2 = (double) L2; o.mn = (double) mn; o.mx = (double) mx;
        partials[blockIdx.x] = o;
    }
}

// ---- the reduction micro-benchmark's kernels (pf_bench_reductions; reference run_benchmarks, simulation.cu:1300-1358) --------
// max |x| over N values: 16-byte loads, warp shuffle tree, ONE integer atomicMax per warp on the IEEE bit pattern - the reduction
// the Merson error norm uses (atomic_max_nonneg), which replaces the reference's multi-pass cuda_reduce(Reduce::MAX).
template <typename T>
__global__ void __launch_bounds__(256)
k_reduce_maxabs(size_t N, const T* __restrict__ x, unsigned long long* bits)
{
    constexpr int V = 16 / sizeof(T), UNR = 4;
    using Vec = typename std::conditional<sizeof(T) == 8, double2, float4>::type;
    T m = 0;
    const size_t nv = N / V, stride = (size_t) gridDim.x * blockDim.x;
    const Vec* xv = reinterpret_cast<const Vec*>(x);
    size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x;
    for(; i + (UNR - 1) * stride < nv; i += UNR * stride)                 // UNR independent 16-byte loads in flight per thread
    {
        Vec v[UNR];
        #pragma unroll
        for(int u = 0; u < UNR; u++) v[u] = xv[i + u * stride];
        #pragma unroll
        for(int u = 0; u < UNR; u++)
        {
            const T* e = reinterpret_cast<const T*>(&v[u]);
            #pragma unroll
            for(int k = 0; k < V; k++) { const T a = e[k] >= 0 ? e[k] : -e[k]; m = a > m ? a : m; }
        }
    }
    for(; i < nv; i += stride)
    {
        const Vec v = xv[i];
        const T* e = reinterpret_cast<const T*>(&v);
        #pragma unroll
        for(int k = 0; k < V; k++) { const T a = e[k] >= 0 ? e[k] : -e[k]; m = a > m ? a : m; }
    }
    for(size_t j = nv * V + blockIdx.x * (size_t) blockDim.x + threadIdx.x; j < N; j += stride)
    {
        const T a = x[j] >= 0 ? x[j] : -x[j]; m = a > m ? a : m;
    }
    __shared__ T sm[8];
    m = warp_max(m);
    if((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if(threadIdx.x == 0)
    {
        #pragma unroll
        for(int w = 1; w < 8; w++) m = sm[w] > m ? sm[w] : m;
        if(m > 0) atomic_max_nonneg(bits, m);                              // one atomic per block
    }
}

// uniform values in [0, 1) from an integer hash of the index (the benchmark's random_map_32, simulation.cu:1316)
template <typename T>
__global__ void k_fill_hash(size_t N, T* x, unsigned seed)
{
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < N; i += (size_t) gridDim.x * blockDim.x)
    {
        unsigned h = (unsigned) i * 2654435761u ^ seed;
        h ^= h >> 16; h *= 0x7feb352du; h ^= h >> 15; h *= 0x846ca68bu; h ^= h >> 16;
        x[i] = (T)((h >> 8) * (1.0f / 16777216.0f));
    }
}

// ---- precision conversion for uploads/downloads: sim_modify_T, reference simulation.cu:1010-1049 ---------
template <typename Dst, typename Src>
__global__ void k_convert(size_t N, const Src* __restrict__ src, Dst* __restrict__ dst)
{
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < N; i += (size_t) gridDim.x * blockDim.x)
        dst[i] = (Dst) src[i];
}

} // namespace pf
