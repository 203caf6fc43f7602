// pf_si_coop.cuh - the WHOLE semi-implicit step in one cooperative kernel.
//
// Replaces semi_implicit_solver_step_based (reference simulation.cu:732-926): the prepare stencil kernel, two
// conjugate_gradient_solve calls (each iteration = 1 SpMV launch + 3 vector launches + 2 reductions with blocking
// device-to-host copies, simulation.cu:649-678, cuda_reduction.cuh:205-208) and the b_T kernel.
//
// At the reference's sizes (<= 2048^2) a CG iteration moves a few tens of MB that sit in the 126 MB L2, so the cost of
// a step is launch and synchronisation latency, not bytes: the per-kernel path (pf_general.cuh, still used when the
// grid cannot be co-resident) needs 2 x (1 + 3 x max_iters) launches per step whatever the iteration count. Here one
// launch runs everything, phases separated by cooperative grid barriers (2 per CG iteration + 4 per step), all scalars
// (alpha, beta, the convergence test) recomputed identically by every block from the per-block partial sums in a fixed
// order. The work decomposition, the arithmetic and the summation trees are those of k_cg_init / k_cg_spmv /
// k_cg_update / k_cg_pupdate, so the results are bit-identical to the per-kernel path. The right-hand sides b_Phi and
// b_T are consumed where they are produced (r = b - A x0) and never stored.
#pragma once
#include <cooperative_groups.h>
#include "pf_general.cuh"

namespace pf {

template <typename Real>
struct SiCoopArgs {
    const Real* F; const Real* U; const Real* U_base;   // y(n); U_base differs from U only inside the corrector loop
    Real* nF; Real* nU;                                  // y(n+1)
    Real* factor; Real* r; Real* p; Real* p2; Real* Ap;   // scratch fields (p2: the second buffer of the search direction)
    Real* rr0; Real* rr1; Real* pAp;                     // per-block partial sums (CG_MAX_BLOCKS each)
    CgState* st;                                         // [0] Phi solve, [1] T solve
    CgMatrix<Real> A_F, A_U;
    KConsts<Real> k;
    int nx, ny, F_bound, U_bound;
    Real sst_F, sst_U;                                   // tolerance^2 * N (simulation.cu:608)
    int max_F, max_U;
    Real L, dt_omg;                                      // b_T = T + L (Phi' - Phi) + dt (1 - gamma) T (simulation.cu:893-899)
};

// The CG iterations after r, p, x and the partials of r.r (in rr0) have been set up: reference simulation.cu:649-690.
// Two grid barriers per iteration: the p update of iteration it (p = r + beta p, simulation.cu:676) is not a phase of its
// own but is folded into the SpMV of iteration it + 1, which forms r + beta*p at the five stencil points on the fly,
// writes the centre value into the OTHER p buffer (the old p is still being read by neighbouring blocks) and goes on with
// A p. Same values, same order of operations per cell as k_cg_pupdate followed by k_cg_spmv: bit-identical results.
template <typename Real>
__device__ __forceinline__ void coop_cg_iterations(cooperative_groups::grid_group& grid, const CgMatrix<Real>& A,
    Real* __restrict__ x, Real* __restrict__ r, Real* p, Real* p_other, Real* __restrict__ Ap,
    Real* rr0, Real* rr1, Real* pApP, CgState* st, const Real sst, const int max_iters, Real* scratch)
{
    const int nx = A.nx, ny = A.ny;
    const size_t N = (size_t) nx * (size_t) ny;
    const int segs = (nx + CG_THREADS - 1) / CG_THREADS;
    const int items = segs * ny;
    const int nb = (int) gridDim.x;
    const Real epsilon = (Real) 1.0e-12;                                    // simulation.cu:882
    Real* rrP[2] = {rr0, rr1};
    if(max_iters <= 0)
    {
        const Real rr = sum_partials(rrP[0], nb, scratch);
        if(blockIdx.x == 0 && threadIdx.x == 0) st->rr = (double) rr;
    }
    Real* pc = p;                 // the current search direction
    Real* pn = p_other;           // where the next one is written
    Real beta = 0;
    for(int it = 0; it < max_iters; it++)
    {
        // [p = r + beta p of the previous iteration,] Ap = A p, partials of p.Ap (k_cg_pupdate + k_cg_spmv)
        Real acc = 0;
        for(int wi = blockIdx.x; wi < items; wi += nb)
        {
            const int y = wi / segs, xx = (wi % segs) * CG_THREADS + threadIdx.x;
            if(xx < nx)
            {
                const size_t i = (size_t) xx + (size_t) y * (size_t) nx;
                Real pv, v;
                if(it == 0) { pv = pc[i]; v = cg_apply(A, pc, xx, y); }
                else
                {
                    pv = pf_fma(beta, pc[i], r[i]);
                    pn[i] = pv;
                    v = cg_apply_axpy(A, pc, r, beta, pv, xx, y, A.which == 0 ? A.factor[i] : (Real) 0);
                }
                Ap[i] = v;
                acc = pf_fma(pv, v, acc);
            }
        }
        if(it > 0) { Real* t = pc; pc = pn; pn = t; }
        acc = block_sum(acc, scratch);
        if(threadIdx.x == 0) pApP[blockIdx.x] = acc;
        grid.sync();
        // alpha, x += alpha p, r -= alpha Ap, partials of the new r.r (k_cg_update)
        const Real rr = sum_partials(rrP[it & 1], nb, scratch);
        const Real pAp = sum_partials(pApP, nb, scratch);
        const Real alpha = rr / (pAp > epsilon ? pAp : epsilon);
        acc = 0;
        for(size_t i = blockIdx.x * (size_t) CG_THREADS + threadIdx.x; i < N; i += (size_t) nb * CG_THREADS)
        {
            x[i] = pf_fma(alpha, pc[i], x[i]);
            const Real rv = pf_fma(-alpha, Ap[i], r[i]);
            r[i] = rv;
            acc = pf_fma(rv, rv, acc);
        }
        acc = block_sum(acc, scratch);
        if(threadIdx.x == 0) rrP[(it + 1) & 1][blockIdx.x] = acc;
        grid.sync();
        // convergence test and beta (k_cg_pupdate's scalars); the vector update waits for the next SpMV
        const Real rr_new = sum_partials(rrP[(it + 1) & 1], nb, scratch);
        if(rr_new < sst)
        {
            if(blockIdx.x == 0 && threadIdx.x == 0) { st->rr = (double) rr_new; st->iters = it; st->done = 1; }
            break;                                                          // every block takes this branch together (identical sums)
        }
        beta = rr_new / (rr > epsilon ? rr : epsilon);
        if(blockIdx.x == 0 && threadIdx.x == 0) st->rr = (double) rr_new;
    }
    grid.sync();        // nobody may still be reading the partials when the next phase overwrites them
}

template <typename Real, bool ANISO>
__global__ void __launch_bounds__(CG_THREADS, 4)
k_si_step_coop(const SiCoopArgs<Real> a)
{
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    __shared__ Real scratch[32];
    const int nx = a.nx, ny = a.ny;
    const int segs = (nx + CG_THREADS - 1) / CG_THREADS;
    const int items = segs * ny;
    const int nb = (int) gridDim.x;

    // phase 0: the prepare kernel (simulation.cu:800-870) fused with the Phi solve's set-up r = b_Phi - A_Phi Phi, p = r, x = Phi
    {
        if(blockIdx.x == 0 && threadIdx.x == 0) { a.st[0].done = 0; a.st[0].iters = a.max_F; a.st[0].rr = 0; }
        Real acc = 0;
        for(int wi = blockIdx.x; wi < items; wi += nb)
        {
            const int y = wi / segs, xx = (wi % segs) * CG_THREADS + threadIdx.x;
            if(xx < nx)
            {
                const size_t i = (size_t) xx + (size_t) y * (size_t) nx;
                Real right, factor;
                si_prepare_cell<Real, ANISO>(a.F, a.U, xx, y, nx, ny, a.F_bound, a.U_bound, a.k, right, factor);
                a.factor[i] = factor;
                const Real rv = right - cg_apply_sc(a.A_F, a.F, xx, y, factor);
                a.nF[i] = a.F[i];
                a.r[i] = rv; a.p[i] = rv;
                acc = pf_fma(rv, rv, acc);
            }
        }
        acc = block_sum(acc, scratch);
        if(threadIdx.x == 0) a.rr0[blockIdx.x] = acc;
    }
    grid.sync();
    coop_cg_iterations<Real>(grid, a.A_F, a.nF, a.r, a.p, a.p2, a.Ap, a.rr0, a.rr1, a.pAp, a.st + 0, a.sst_F, a.max_F, scratch);

    // b_T (simulation.cu:893-899) fused with the T solve's set-up
    {
        if(blockIdx.x == 0 && threadIdx.x == 0) { a.st[1].done = 0; a.st[1].iters = a.max_U; a.st[1].rr = 0; }
        Real acc = 0;
        for(int wi = blockIdx.x; wi < items; wi += nb)
        {
            const int y = wi / segs, xx = (wi % segs) * CG_THREADS + threadIdx.x;
            if(xx < nx)
            {
                const size_t i = (size_t) xx + (size_t) y * (size_t) nx;
                const Real T = a.U_base[i];
                const Real bU = pf_fma(a.dt_omg, T, pf_fma(a.L, a.nF[i] - a.F[i], T));
                const Real rv = bU - cg_apply_sc(a.A_U, a.U, xx, y, (Real) 0);
                a.nU[i] = a.U[i];
                a.r[i] = rv; a.p[i] = rv;
                acc = pf_fma(rv, rv, acc);
            }
        }
        acc = block_sum(acc, scratch);
        if(threadIdx.x == 0) a.rr0[blockIdx.x] = acc;
    }
    grid.sync();
    coop_cg_iterations<Real>(grid, a.A_U, a.nU, a.r, a.p, a.p2, a.Ap, a.rr0, a.rr1, a.pAp, a.st + 1, a.sst_U, a.max_U, scratch);
}

} // namespace pf
