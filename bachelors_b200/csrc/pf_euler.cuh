// pf_euler.cuh - the explicit Euler step (reference simulation.cu:283-311, RHS :145-243) as a bandwidth kernel.
//
// The streaming kernel of pf_stream.cuh is built for the multi-stage schemes: a skewed row pipeline with one barrier per row
// and ONE row of loads in flight per thread. For the single-stage Euler step that is the wrong shape - there is nothing to
// pipeline, and a step that moves 32 B per cell for ~30 FP64 instructions lives or dies by the bytes it keeps in flight
// (measured, profiles/r02_kernel_experiments.md section 5: 33.9 us at 2048^2 fp64 where a copy of the same bytes takes 20.8).
//
// Here a CTA of EW threads owns a tile of EW columns x R rows, thread = column. Every thread first issues ALL of its loads -
// 2 fields x (R + 2) rows, independent, coalesced across the warp - then publishes its R rows as {Phi, T} pairs in shared
// memory for the E/W neighbours (EULER_HALO threads on each side of the tile only load and publish), ONE barrier, and
// the R cells of the column are computed from registers (C, N, S) and shared memory (E, W) and stored. Bytes in flight per SM:
// CTAs x EW x 2(R + 2) x sizeof(Real) (6 x 128 x 20 x 8 B = 123 KB at R = 8, fp64) instead of 16 KB.
// Traffic: rows are re-read (R + 2)/R times and columns EW/(EW - 8) times; the neighbouring tiles load the same lines at about
// the same time, so L2 serves the overlap. Arithmetic, FMA placement and the -0 handling are those of the streaming kernel's
// Euler instantiation (rhs_cell_lean on canon_zero(y), y(n+1) = fma(dt, k, y)): bit-identical results.
//
// Boundaries (reference boundary_sample, simulation.cu:33-85): periodic edges wrap the index at load time; outside a
// non-periodic edge the ghost value is formed from the edge cell (Neumann: the same value, Dirichlet: 0 - value).
// Slabs: row0 / rows_local / own_y0..1
// as in FusedArgs; one ghost row per side.
#pragma once
#include "pf_common.cuh"
#include "pf_fused.cuh"

namespace pf {

enum { EULER_HALO = 4 };      // columns of halo on each side of a tile: 1 is needed, 4 keeps every row segment 32-byte aligned (fp64)

template <typename Real, int EW, int R>
constexpr size_t euler_tile_smem_bytes() { return sizeof(typename Vec2T<Real>::type) * (size_t) R * EW; }

// GENERAL = false: the plain step on the lean right-hand side (no corrector guess, no source term).
// GENERAL = true: the whole parameter space of the reference's Euler path - the corrector-guess form of the RHS (simulation.cu:221-225,
// run-time flag), the source term fu of do_exact, and the corrector iteration's form y(n+1) = {F, U_base} + dt * f(F, U)
// (simulation.cu:288-302; a.U_base != NULL) - in the same single launch.
// Both: Phi and T may have different NON-periodic boundary types (a.bc, a.bc_T); periodic must be both or none.
template <typename Real, bool ANISO, bool GENERAL, int EW, int R, int MINB>
__global__ void __launch_bounds__(EW, MINB)
k_euler_tile(const FusedArgs<Real> a)
{
    using V2 = typename Vec2T<Real>::type;
    extern __shared__ __align__(16) unsigned char pf_smem_raw[];
    V2* const sm = reinterpret_cast<V2*>(pf_smem_raw);
    constexpr int EWI = EW - 2 * EULER_HALO;

    const int nx = a.nx, NY = a.ny_global, bc = a.bc, bcT = a.bc_T;
    const bool periodic = bc == PF_BOUNDARY_PERIODIC;
    const int tx = (int) threadIdx.x;
    const int gx = (int) blockIdx.x * EWI - EULER_HALO + tx;               // this thread's column (halo threads: beside the tile, maybe outside the grid)
    const int gy0 = a.own_y0 + (int) blockIdx.y * R;                        // first row of the tile
    const bool interior = tx >= EULER_HALO && tx < EW - EULER_HALO && gx < nx;
    // column to load: wrapped (periodic) or clamped; outside a non-periodic edge the ghost rule applies to the clamped value
    const bool out_x = !periodic && (gx < 0 || gx >= nx);
    int lxi = gx;
    if(periodic) { lxi %= nx; if(lxi < 0) lxi += nx; }
    else         lxi = lxi < 0 ? 0 : (lxi >= nx ? nx - 1 : lxi);
    const unsigned lx = (unsigned) lxi;
    // rows that exist both locally and in the domain: loads are clamped into them
    const int rlo = periodic ? a.row0 : max(a.row0, 0);
    const int rhi = (periodic ? a.row0 + a.rows_local : min(a.row0 + a.rows_local, NY)) - 1;

    // 1. all loads of this thread's column: rows gy0 - 1 .. gy0 + R (wrapped / clamped)
    Real yF[R + 2], yT[R + 2];
    unsigned off[R + 2];
    #pragma unroll
    for(int j = 0; j < R + 2; j++)
    {
        int g = gy0 - 1 + j;
        if(a.wrap_y) { while(g < 0) g += NY; while(g >= NY) g -= NY; }
        g = g < rlo ? rlo : (g > rhi ? rhi : g);
        off[j] = (unsigned)(g - a.row0) * (unsigned) nx + lx;
    }
    #pragma unroll
    for(int j = 0; j < R + 2; j++) { yF[j] = a.F[off[j]]; yT[j] = a.U[off[j]]; }

    // 2. canonical values (the reference's blend "0 + y*1" turns -0 into +0), ghost values outside non-periodic edges, publish
    Real rawF[GENERAL ? R : 1];                                            // the corrector form adds to the RAW Phi (simulation.cu:299)
    if(GENERAL)
    {
        #pragma unroll
        for(int j = 0; j < R; j++) rawF[j] = yF[j + 1];
    }
    #pragma unroll
    for(int j = 0; j < R + 2; j++) { yF[j] = canon_zero(yF[j]); yT[j] = canon_zero(yT[j]); }
    if(!periodic)
    {
        // a row outside the domain was loaded clamped, i.e. as the edge row itself: mirror it
        #pragma unroll
        for(int j = 0; j < R + 2; j++)
        {
            const int g = gy0 - 1 + j;
            if(g < 0 || g >= NY) { yF[j] = ghost_of(yF[j], bc); yT[j] = ghost_of(yT[j], bcT); }
        }
    }
    #pragma unroll
    for(int j = 0; j < R; j++)
    {
        V2 v; v.x = yF[j + 1]; v.y = yT[j + 1];
        if(out_x) { v.x = ghost_of(v.x, bc); v.y = ghost_of(v.y, bcT); }    // the column beside a non-periodic edge: the edge column mirrored
        sm[j * EW + tx] = v;
    }
    __syncthreads();

    // 3. the R cells of this column
    if(!interior) return;
    #pragma unroll
    for(int j = 0; j < R; j++)
    {
        const int g = gy0 + j;
        if(g < a.own_y1)
        {
            const V2 e = sm[j * EW + tx + 1], w = sm[j * EW + tx - 1];
            Real kF, kT;
            const unsigned idx = (unsigned)(g - a.row0) * (unsigned) nx + (unsigned) gx;
            if(GENERAL)
            {
                rhs_cell<Real, ANISO>(a.k, yF[j + 1], e.x, w.x, yF[j + 2], yF[j], yT[j + 1], e.y, w.y, yT[j + 2], yT[j], kF, kT);
                if(a.U_base)
                {
                    a.out_F[idx] = pf_fma(a.k.dt, kF, rawF[j]);             // simulation.cu:299-300
                    a.out_U[idx] = pf_fma(a.k.dt, kT, a.U_base[idx]);
                    continue;
                }
            }
            else
                rhs_cell_lean<Real, ANISO>(a.k, yF[j + 1], e.x, w.x, yF[j + 2], yF[j], yT[j + 1], e.y, w.y, yT[j + 2], yT[j], kF, kT);
            a.out_F[idx] = pf_fma(a.k.dt, kF, yF[j + 1]);                   // simulation.cu:233-234
            a.out_U[idx] = pf_fma(a.k.dt, kT, yT[j + 1]);
        }
    }
}

} // namespace pf
