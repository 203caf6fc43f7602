// pf_fused.cuh - the argument block of the single-launch explicit steps (one launch = ALL stages of one time step).
//
// Shared by the streaming kernels (pf_stream.cuh: explicit-rk4, Merson) and the Euler tile kernel (pf_euler.cuh), which replace,
// for one time step, the reference's 1 (Euler) / 5 (RK4) / >= 7 (Merson) launches of explicit_solver_solve_lin_combination +
// combine + reductions (reference simulation.cu:283-497). The 2-D tiled kernel that this file held in round 1 (extended tiles
// with a halo of #stages cells, all stages in registers / shared memory) is no longer part of the product: it lives on as the
// cross-check of tools/kbench.cu (tools/kbench_tiled.cuh).
//
// Multi-GPU slabs: `row0` is the global row of local row 0; rows outside [0, rows_local) do not exist locally
// and rows outside [0, ny_global) are outside the domain unless the boundary is periodic.
#pragma once
#include "pf_common.cuh"

namespace pf {

enum { SCHEME_EULER = 0, SCHEME_RK4 = 1, SCHEME_RKM = 2 };

template <typename Real>
struct FusedArgs {
    const Real* F; const Real* U;      // y(n)
    Real* out_F; Real* out_U;          // y(n+1)
    int nx, ny_global;
    int row0, rows_local;              // local allocation = global rows [row0, row0 + rows_local)
    int own_y0, own_y1;                // global rows this launch must produce
    int bc;                            // boundary type of Phi (tiled kernel of this file: of both fields)
    int bc_T;                          // streaming and Euler tile kernels: boundary type of T (both periodic or both not)
    const Real* U_base;                // Euler tile kernel, corrector iteration: y(n+1) = {F, U_base} + dt f(F, U); NULL: the plain step
    int wrap_y;                        // single GPU + periodic: wrap rows modulo ny_global
    int strips, nch;                   // streaming kernel (pf_stream.cuh): column strips x row chunks = work units; chunk c produces rows
                                       // own_y0 + [rows*c/nch, rows*(c+1)/nch) (balanced: lengths differ by at most one row)
    KConsts<Real> k;
    // scheme weights, computed on the host in Real arithmetic exactly as the reference does:
    //  RK4 (simulation.cu:330-337):  w[0]=dt/2 w[1]=dt/2 w[2]=dt w[3]=dt/6
    //  RKM (simulation.cu:406-409, :472): w[0]=tau/3 w[1]=tau/6 w[2]=tau/8 w[3]=tau*3/8 w[4]=tau/2 w[5]=-tau*3/2 w[6]=tau*2 w[7]=tau/6
    Real w[8];
    unsigned long long* err_bits;      // RKM: [0] = max|e_Phi| bits, [1] = max|e_T| bits (atomicMax)
    // RKM, one launch per attempt (pf_stream.cuh): the last unit to finish hands the two norms and err_seq to the host through
    // mapped pinned memory (err_host[0..2]) and re-arms err_bits / err_ticket; NULL: the words stay on the device for the caller
    unsigned long long* err_host; unsigned* err_ticket; unsigned long long err_seq;
    // slabs, k_stream_step_halo (pf_stream.cuh): the halo hand-over inside the step kernel. Side 0 = lower neighbour, 1 = upper.
    // The units of the first / last chunk wait until this rank's flag word `side` has reached halo_e (the neighbour's rows of the
    // previous exchange have landed in the ghost rows they read), and after their rows are stored copy the slab's first / last
    // halo_g rows into peer_F/U[side] (mapped peer memory, already offset to the first ghost row to fill); the last of them
    // publishes halo_e + 1 into peer_flag[side]. A NULL peer_flag[side]: no neighbour on that side.
    unsigned halo_e; int halo_g;
    unsigned* halo_flags;
    Real* peer_F[2]; Real* peer_U[2]; unsigned* peer_flag[2];
};

template <int SCHEME> struct SchemeStages { static constexpr int value = SCHEME == SCHEME_EULER ? 1 : (SCHEME == SCHEME_RK4 ? 4 : 5); };

} // namespace pf
