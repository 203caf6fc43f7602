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
    unsigned long long* ll;                              // flagged partial sums (GridSum below); NULL: grid.sync + the arrays above
    unsigned ll_epoch;                                   // first epoch of this launch (never reused while the slots live)
    CgState* st;                                         // [0] Phi solve, [1] T solve
    CgMatrix<Real> A_F, A_U;
    KConsts<Real> k;
    int nx, ny, F_bound, U_bound;
    Real sst_F, sst_U;                                   // tolerance^2 * N (simulation.cu:608)
    int max_F, max_U;
    Real L, dt_omg;                                      // b_T = T + L (Phi' - Phi) + dt (1 - gamma) T (simulation.cu:893-899)
};

// One sum over all blocks that is also the grid barrier between two phases.
//
// Plain form: store the block's partial, cooperative grid.sync(), every block re-reads all partials (sum_partials).
// That is ~8 dependent trips to L2 per phase (drain the stores, the arrival atomic, the polling loop, the re-read),
// and at <= 1024^2 those trips - not bytes, not flops - are most of a semi-implicit step.
//
// Flagged form (ll != NULL): the partial travels WITH its flag. A block publishes {value, epoch} into its own slot
// (8-byte words: 32 bits of the value + the 32-bit epoch, two words for a double; an aligned 8-byte access is single-copy
// atomic, so a reader never sees a value without its flag) after a __threadfence() that orders the block's field stores
// before it. Block 0 polls all slots until they carry the current epoch, sums them in the fixed order of sum_partials and
// publishes the result the same way into a result slot, on which one thread of every other block spins. Seeing the
// result is the barrier, and the value arrives with it: no arrival atomics (592 on one address), no re-read of the
// partials by every block. (Every block polling every slot - the barrier with one trip fewer - was measured too: the
// 592^2 x 16 B of polling traffic per barrier made a 1024^2 step 305 us instead of 222.) Two slot sets alternate by epoch
// parity: a block can publish epoch e + 2 only after it has seen the result of e + 1, which block 0 forms from slots
// that every block writes only after it has seen the result of e, i.e. after block 0 has finished reading the slots of e.
// The grid must be co-resident (cooperative launch), as for grid.sync(). Same values added in the same order:
// bit-identical to the plain form (pf_set_option PF_OPT_SI_GRID_SYNC; tests/test_gpu_parity.py runs both).
// Measured, one B200, fp64 us/step plain -> flagged: 1024^2 222 -> 214, 512^2 148 -> 142, 256^2 112 -> 92.
// __threadfence() is a sequentially consistent fence (fence.sc.gpu); release / acquire is all the flag protocol needs
__device__ __forceinline__ void ll_fence()
{
    asm volatile("fence.acq_rel.gpu;" ::: "memory");
}
__device__ __forceinline__ void ll_store(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.volatile.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ll_load(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// both words of a 16-byte slot in one access (a slot is 16-byte aligned; each word carries the tag, so tearing would be seen)
__device__ __forceinline__ void ll_store2(unsigned long long* p, unsigned long long a, unsigned long long b)
{
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" :: "l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void ll_load2(const unsigned long long* p, unsigned long long& a, unsigned long long& b)
{
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}

template <typename Real>
struct GridSum {
    cooperative_groups::grid_group& grid;
    unsigned long long* ll;
    unsigned epoch;
    int nb;
    Real* scratch;

    // block_val: this block's partial (valid in thread 0, as block_sum returns it); arr: the plain form's partial array
    __device__ __forceinline__ Real operator()(Real block_val, Real* arr)
    {
        if(ll == nullptr)
        {
            if(threadIdx.x == 0) arr[blockIdx.x] = block_val;
            grid.sync();
            return sum_partials(arr, nb, scratch);
        }
        const unsigned e = epoch++;
        const unsigned long long tag = (unsigned long long) e << 32;
        unsigned long long* set = ll + (size_t)(e & 1u) * (2 * CG_MAX_BLOCKS);
        if(threadIdx.x == 0)
        {
            ll_fence();                                                     // release: the block's field stores (ordered before by __syncthreads) first
            if constexpr (sizeof(Real) == 8)
            {
                const unsigned long long bits = (unsigned long long) __double_as_longlong((double) block_val);
                ll_store2(set + 2 * blockIdx.x, tag | (bits & 0xffffffffull), tag | (bits >> 32));
            }
            else ll_store(set + 2 * blockIdx.x, tag | (unsigned long long) __float_as_uint((float) block_val));
        }
        unsigned long long* out = ll + 4 * CG_MAX_BLOCKS + 2 * (e & 1u);    // behind the two slot sets
        __shared__ Real bcast;
        if(blockIdx.x == 0)
        {
            Real acc = 0;
            if(nb <= CG_THREADS)
            {
                // one slot per thread: spin on it
                if((int) threadIdx.x < nb)
                {
                    const unsigned long long* slot = set + 2 * threadIdx.x;
                    if constexpr (sizeof(Real) == 8)
                    {
                        unsigned long long a, b;
                        do { ll_load2(slot, a, b); } while((unsigned)(a >> 32) != e || (unsigned)(b >> 32) != e);
                        acc += (Real) __longlong_as_double((long long)((b << 32) | (a & 0xffffffffull)));
                    }
                    else
                    {
                        unsigned long long a;
                        do { a = ll_load(slot); } while((unsigned)(a >> 32) != e);
                        acc += (Real) __uint_as_float((unsigned) a);
                    }
                }
            }
            else
            {
                // Up to 4 slots per thread (nb <= 1024): the pending ones are re-read in every round and a warp leaves the loop as a
                // whole (vote). Lanes that leave a spin loop one by one run what follows - the next slot's trip to L2 - one after
                // the other. Measured (tools/si_phase_cost.py, fp64, us per CG iteration of both solves, slot after slot -> vote):
                // 512^2 24.4 -> 21.8, 1024^2 42.3 -> 39.6 (592 blocks).
                constexpr int SPT = CG_MAX_BLOCKS / CG_THREADS;
                Real val[SPT];
                unsigned pending = 0;
                #pragma unroll
                for(int q = 0; q < SPT; q++)
                {
                    val[q] = 0;
                    if((int) threadIdx.x + q * CG_THREADS < nb) pending |= 1u << q;
                }
                do
                {
                    #pragma unroll
                    for(int q = 0; q < SPT; q++)
                        if(pending & (1u << q))
                        {
                            const unsigned long long* slot = set + 2 * ((int) threadIdx.x + q * CG_THREADS);
                            if constexpr (sizeof(Real) == 8)
                            {
                                unsigned long long a, b;
                                ll_load2(slot, a, b);
                                if((unsigned)(a >> 32) == e && (unsigned)(b >> 32) == e)
                                {
                                    val[q] = (Real) __longlong_as_double((long long)((b << 32) | (a & 0xffffffffull)));
                                    pending &= ~(1u << q);
                                }
                            }
                            else
                            {
                                const unsigned long long a = ll_load(slot);
                                if((unsigned)(a >> 32) == e) { val[q] = (Real) __uint_as_float((unsigned) a); pending &= ~(1u << q); }
                            }
                        }
                } while(!__all_sync(0xffffffffu, pending == 0));
                #pragma unroll
                for(int q = 0; q < SPT; q++)
                    if((int) threadIdx.x + q * CG_THREADS < nb) acc += val[q];  // the order of sum_partials
            }
            acc = block_sum(acc, scratch);                                  // its __syncthreads orders all threads' polls before thread 0's fence
            if(threadIdx.x == 0)
            {
                // acquire the partial flags, release the result: block B's field stores -> B's flag -> (this fence) -> result flag
                // -> block C's acquire is then a chain of release/acquire pairs in the PTX memory model, not a hardware property
                ll_fence();
                bcast = acc;
                if constexpr (sizeof(Real) == 8)
                {
                    const unsigned long long bits = (unsigned long long) __double_as_longlong((double) acc);
                    ll_store2(out, tag | (bits & 0xffffffffull), tag | (bits >> 32));
                }
                else ll_store(out, tag | (unsigned long long) __float_as_uint((float) acc));
            }
        }
        else if(threadIdx.x == 0)
        {
            if constexpr (sizeof(Real) == 8)
            {
                unsigned long long a, b;
                do { ll_load2(out, a, b); } while((unsigned)(a >> 32) != e || (unsigned)(b >> 32) != e);
                bcast = (Real) __longlong_as_double((long long)((b << 32) | (a & 0xffffffffull)));
            }
            else
            {
                unsigned long long a;
                do { a = ll_load(out); } while((unsigned)(a >> 32) != e);
                bcast = (Real) __uint_as_float((unsigned) a);
            }
        }
        if(threadIdx.x == 0) ll_fence();                                    // acquire: loads after the barrier see what was stored before the flags
        __syncthreads();
        const Real r = bcast;
        __syncthreads();
        return r;
    }
};

// The CG iterations after r, p, x and the partials of r.r (in rr0) have been set up: reference simulation.cu:649-690.
// Two grid barriers per iteration: the p update of iteration it (p = r + beta p, simulation.cu:676) is not a phase of its
// own but is folded into the SpMV of iteration it + 1, which forms r + beta*p at the five stencil points on the fly,
// writes the centre value into the OTHER p buffer (the old p is still being read by neighbouring blocks) and goes on with
// A p. Same values, same order of operations per cell as k_cg_pupdate followed by k_cg_spmv: bit-identical results.
template <typename Real>
__device__ __forceinline__ void coop_cg_iterations(GridSum<Real>& gsum, Real rr, const CgMatrix<Real>& A,
    Real* __restrict__ x, Real* __restrict__ r, Real* p, Real* p_other, Real* __restrict__ Ap,
    Real* rr0, Real* rr1, Real* pApP, CgState* st, const Real sst, const int max_iters, Real* scratch)
{
    // rr: r.r of the initial residual (the caller's GridSum after the set-up phase - also the barrier in front of the first SpMV)
    const int nx = A.nx, ny = A.ny;
    const size_t N = (size_t) nx * (size_t) ny;
    const int segs = (nx + CG_THREADS - 1) / CG_THREADS;
    const int items = segs * ny;
    const int nb = (int) gridDim.x;
    const Real epsilon = (Real) 1.0e-12;                                    // simulation.cu:882
    Real* rrP[2] = {rr0, rr1};
    if(max_iters <= 0)
    {
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
        const Real pAp = gsum(acc, pApP);
        // alpha, x += alpha p, r -= alpha Ap, partials of the new r.r (k_cg_update)
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
        // convergence test and beta (k_cg_pupdate's scalars); the vector update waits for the next SpMV
        const Real rr_new = gsum(acc, rrP[(it + 1) & 1]);
        if(rr_new < sst)
        {
            if(blockIdx.x == 0 && threadIdx.x == 0) { st->rr = (double) rr_new; st->iters = it; st->done = 1; }
            break;                                                          // every block takes this branch together (identical sums)
        }
        beta = rr_new / (rr > epsilon ? rr : epsilon);
        rr = rr_new;                                                        // the same partials summed in the same order: what the next iteration would recompute
        if(blockIdx.x == 0 && threadIdx.x == 0) st->rr = (double) rr_new;
    }
    // plain form: nobody may still be reading the partial arrays when the next phase overwrites them. (Flagged form: the slot
    // sets alternate; and every block has finished its last SpMV - the only reader of other blocks' p and r - before its
    // p.Ap flag, which everyone has seen: the next phase may overwrite r and p.)
    if(gsum.ll == nullptr) gsum.grid.sync();
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
    GridSum<Real> gsum{grid, a.ll, a.ll_epoch, nb, scratch};
    Real rr_init;

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
        rr_init = gsum(acc, a.rr0);
    }
    coop_cg_iterations<Real>(gsum, rr_init, a.A_F, a.nF, a.r, a.p, a.p2, a.Ap, a.rr0, a.rr1, a.pAp, a.st + 0, a.sst_F, a.max_F, scratch);

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
        rr_init = gsum(acc, a.rr0);
    }
    coop_cg_iterations<Real>(gsum, rr_init, a.A_U, a.nU, a.r, a.p, a.p2, a.Ap, a.rr0, a.rr1, a.pAp, a.st + 1, a.sst_U, a.max_U, scratch);
}

} // namespace pf
