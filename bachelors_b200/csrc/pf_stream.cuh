// pf_stream.cuh - streaming temporally blocked explicit integrators: ALL stages of one time step in ONE kernel,
// marching down the rows with a skewed software pipeline.
//
// Replaces, for one time step, the reference's 1 (Euler) / 5 (RK4) / >=7 (Merson) launches of
// explicit_solver_solve_lin_combination + combine + reductions (reference simulation.cu:283-497).
//
// One CTA of EW = 32*WARPS threads owns a column strip: thread cx owns one global column (H = #stages columns on each
// side of the strip are halo, except towards a non-periodic domain edge, where none is needed) and walks down a chunk of rows. In iteration `it` the thread
// evaluates stage s on row g1 - (s-1), g1 = G0 + it being the row of stage 1: stage s+1 consumes, as its N neighbour,
// what stage s has produced a moment ago, in the same iteration.
// The N/S neighbours and the running combinations of the k's live in REGISTERS (rings indexed by row modulo 3 = the
// unroll factor, all indices compile-time). Shared memory carries (i) the E/W neighbours: one row
// of {Phi,T} pairs per stage in a 3-deep ring, and (ii) a ring of the last YR rows of canon_zero(y(n)), from which
// stage 1 takes its E/W neighbours and stages >= 3 re-read their own column's y instead of holding it in registers.
// ONE __syncthreads per row. y(n) is read from global memory once per strip (two rows ahead of its use) and y(n+1)
// is written once: HBM traffic per cell per step = read {Phi,T} + write {Phi,T} = 4 reals (x EW/(EW-2H) strip
// overlap, served by L2), the algorithmic minimum. Redundant arithmetic: EW/(EW-2H) in x; in y
// only the pipeline fill (2(NST-1) rows per chunk, idle stages skipped) - no 2-D halo recomputation.
//
// Garbage discipline: halo columns and out-of-domain rows are computed from clamped loads; wrong values move inward
// one cell per stage, so after NST stages exactly the interior columns are clean, and at non-periodic edges the
// ghost rule below overrides every arm that points outside before it can reach a valid cell.
//
// Boundaries: when a stencil arm of an edge cell points outside a non-periodic edge, the ghost value of
// boundary_sample (reference simulation.cu:33-85) is formed from the edge cell's own stage input (Neumann: same
// value, Dirichlet: 0 - value) - the reference applies boundary_sample to every blended state, and the blend is
// linear. In x the edge thread itself stores that ghost pair into the exchange slot of the column outside (whose
// own thread does not store), so the stencil code has no x-boundary logic at all. In y the ghost replaces the N/S
// register operand, which only happens in the `CHECKED` form of the loop body: iterations are grouped by U, and a
// group runs the straight-line form (no activity tests, no boundary selects, one basic block per row) whenever all
// stages are active and no stage row is a domain edge row. Periodic edges wrap the global index at load time.
// Phi and T may have different NON-periodic boundary types (a.bc, a.bc_T); periodic must be both or none, and the corrector
// guess is not served (otherwise the host falls back to the tiled kernel / the general per-stage kernels).
//
// Multi-GPU slabs: `row0` is the global row of local row 0; the launch produces rows [own_y0, own_y1) and reads
// H rows beyond them, which the slab layout provides as ghost rows.
#pragma once
#include "pf_common.cuh"
#include "pf_fused.cuh"

namespace pf {

constexpr __host__ __device__ int pmod(int a, int m) { return ((a % m) + m) % m; }
constexpr __host__ __device__ int pow2_at_least(int n) { int p = 1; while(p < n) p *= 2; return p; }

// Geometry of the pipeline for a scheme.
template <int SCHEME>
struct StreamGeom {
    static constexpr int NST = SchemeStages<SCHEME>::value;
    static constexpr int U = 3;                                       // unroll factor = size of the register rings and of the exchange rings (indexed by row mod 3)
    static constexpr int YR = pow2_at_least(NST + 1);                 // rows of canon y kept in shared memory: g1+1 .. g1-(NST-1)
    static constexpr int FILL = 2 * (NST - 1);                        // first iteration in which every stage is active
    static constexpr int L2_AHEAD = 8;                                // rows ahead of the register prefetch that are pulled into L2
};

// Forces two values to be computed where this call stands (the compiler may not sink them into a later branch).
__device__ __forceinline__ void keep_here(double& a, double& b) { asm volatile("" : "+d"(a), "+d"(b)); }
__device__ __forceinline__ void keep_here(float& a, float& b)   { asm volatile("" : "+f"(a), "+f"(b)); }

// v with the sign bit set when `bit` is 1 (v is +0 in that case: restores a raw -0).
__device__ __forceinline__ double with_sign(double v, unsigned bit)
{
    return __hiloint2double(__double2hiint(v) | (int)(bit << 31), __double2loint(v));
}
__device__ __forceinline__ float with_sign(float v, unsigned bit) { return __int_as_float(__float_as_int(v) | (int)(bit << 31)); }

// Per-thread pipeline state. Every array is indexed with compile-time constants only (after unrolling), so it lives
// in registers and unused slots cost nothing. All rings are indexed by (row mod 3).
template <typename Real, int SCHEME>
struct StreamState {
    using G = StreamGeom<SCHEME>;
    Real yF[G::U], yT[G::U];               // canon_zero(y(n)) of rows g1-1, g1, g1+1 (and the retiring g1+2)
    Real cF[G::NST][G::U], cT[G::NST][G::U];   // c[s-1] = input of stage s (s >= 2)
    Real aF[G::NST][G::U], aT[G::NST][G::U];   // a[s][row]: what stage s leaves for stage s+1 (RK4: k1+2k2+..; RKM: k1, later k1+4k4)
    Real bF[G::NST][G::U], bT[G::NST][G::U];   // RKM: k3, later the error combination
    unsigned negmask;                      // raw y was -0: bits (2d, 2d+1) = (Phi, T) of row g1 + 1 - d during an iteration (shifted by 2 every row)
    Real errF, errT;                       // RKM: running max of the error combination over the cells this thread stored
};

template <typename Real>
struct StreamCtx {
    using V2 = typename Vec2T<Real>::type;
    const Real* F; const Real* U;      // y(n) of the step being computed (a multi-step launch swaps the roles of the buffers every step)
    Real* out_F; Real* out_U;          // y(n+1)
    V2* ybuf;                          // ring of canon y rows, offset to this thread's slot
    V2* cbuf;                          // U-deep ring of the stage-input rows of stages 2..NST, offset to this thread's slot
    unsigned lx;                       // clamped / wrapped column for loads
    unsigned gxu;                      // this thread's column (stores; only used where st_ok)
    bool wr_self;                      // this thread publishes its own values (false for the column just outside a non-periodic edge)
    bool x_lo_edge, x_hi_edge;         // non-periodic domain edge columns: they publish the ghost pair for the column outside
    bool st_ok;                        // this thread's column is an interior column inside the domain
    bool edge_warp;                    // warp-uniform: some thread of this warp is a non-periodic domain edge column
    int G0;                            // global row of stage 1 in iteration 0
    int Y0, Y1;                        // rows this CTA produces
    int ylo, yhi;                      // rows a stage may be evaluated on (domain rows; everything when periodic)
};

template <typename Real>
__device__ __forceinline__ void stream_store_pair(typename Vec2T<Real>::type* p, const bool on, const Real vF, const Real vT)
{
    typename Vec2T<Real>::type v; v.x = vF; v.y = vT;
    if(on) *p = v;
}

// Domain edge columns also publish the ghost pair of boundary_sample into the slot of the column outside.
template <typename Real>
__device__ __forceinline__ void stream_store_ghost(const FusedArgs<Real>& a, const StreamCtx<Real>& c, typename Vec2T<Real>::type* p, const Real vF, const Real vT)
{
    typename Vec2T<Real>::type g; g.x = vF; g.y = vT;
    if(a.bc == PF_BOUNDARY_DIRICHLET_ZERO) g.x = (Real) 0 - vF;                // uniform branches; Neumann repeats the edge cell
    if(a.bc_T == PF_BOUNDARY_DIRICHLET_ZERO) g.y = (Real) 0 - vT;
    if(c.x_lo_edge) p[-1] = g;
    if(c.x_hi_edge) p[+1] = g;
}

// One stage of one iteration. PH = it mod U, S = stage (1-based). CHECKED: test whether the stage is active in this
// iteration and apply the y-boundary ghosts; otherwise the caller guarantees both are unnecessary.
// `old` is the state as it was when the iteration began: everything a stage consumes comes from there, except its
// N neighbour, which is what stage S-1 has just written to `st`.
template <typename Real, int SCHEME, int ANISO, int EW, bool CHECKED, int PH, int S>
__device__ __forceinline__ void stream_stage(const FusedArgs<Real>& a, const StreamCtx<Real>& c,
                                             const StreamState<Real, SCHEME>& old, StreamState<Real, SCHEME>& st, const int it)
{
    using V2 = typename Vec2T<Real>::type;
    using G = StreamGeom<SCHEME>;
    constexpr int NST = G::NST, U = G::U, ROWSTRIDE = EW + 2;
    constexpr int rel = PH - (S - 1);
    constexpr int jc = pmod(rel, U), jn = pmod(rel + 1, U), js = pmod(rel - 1, U);       // ring slots of rows r, r+1, r-1
    const int row = c.G0 + it - (S - 1);
    if(CHECKED)
    {
        const bool act = row >= c.Y0 - (NST - S) && row < c.Y1 + (NST - S) && row >= c.ylo && row < c.yhi;     // block-uniform
        if(!act)
        {
            // an idle stage still "writes" its ring slots, so that they are dead between their last use and this point
            if(S < NST) { st.cF[S][jc] = 0; st.cT[S][jc] = 0; st.aF[S][jc] = 0; st.aT[S][jc] = 0; st.bF[S][jc] = 0; st.bT[S][jc] = 0; }
            return;
        }
    }

    Real CF, CT, NF, NT, SF, ST, y0F, y0T;
    V2 e, w;
    const V2* const yrow = c.ybuf + (row & (G::YR - 1)) * ROWSTRIDE;
    if(S == 1)
    {
        CF = old.yF[jc]; CT = old.yT[jc]; NF = old.yF[jn]; NT = old.yT[jn]; SF = old.yF[js]; ST = old.yT[js];
        e = yrow[1]; w = yrow[-1];
        y0F = CF; y0T = CT;
    }
    else
    {
        CF = old.cF[S - 1][jc]; CT = old.cT[S - 1][jc]; SF = old.cF[S - 1][js]; ST = old.cT[S - 1][js];
        NF = st.cF[S - 1][jn]; NT = st.cT[S - 1][jn];                                    // fresh: stage S-1, this iteration
        const V2* const crow = c.cbuf + (jc * (NST - 1) + (S - 2)) * ROWSTRIDE;          // published when stage S-1 ran on this row
        e = crow[1]; w = crow[-1];
        if(S == 2) { y0F = old.yF[pmod(PH - 1, U)]; y0T = old.yT[pmod(PH - 1, U)]; }   // stage 1's S row: still in registers
        else { const V2 y0 = yrow[0]; y0F = y0.x; y0T = y0.y; }                          // canon_zero(y) of this row, own column
    }
    if(CHECKED && a.bc != PF_BOUNDARY_PERIODIC)
    {
        if(row == a.ny_global - 1) { NF = ghost_of(CF, a.bc); NT = ghost_of(CT, a.bc_T); }
        if(row == 0)               { SF = ghost_of(CF, a.bc); ST = ghost_of(CT, a.bc_T); }
    }
    Real kF, kT;
    // ANISO: bit 0 = anisotropy, bit 1 = the corrector-guess form of the right-hand side (simulation.cu:221-225)
    rhs_cell_lean<Real, (ANISO & 1) != 0, (ANISO & 2) != 0>(a.k, CF, e.x, w.x, NF, SF, CT, e.y, w.y, NT, ST, kF, kT);   // kT without the reference's "+ fu"

    Real nF = 0, nT = 0;
    if(SCHEME == SCHEME_EULER)
    {
        nF = pf_fma(a.k.dt, kF, CF);                                        // simulation.cu:233-234
        nT = pf_fma(a.k.dt, kT, CT);                                        // C_T is never -0, so the missing "+ fu" cannot show
    }
    else if(SCHEME == SCHEME_RK4)
    {
        if(S == 1)      { st.aF[1][jc] = kF; st.aT[1][jc] = kT;
                          nF = pf_fma(kF, a.w[0], y0F); nT = pf_fma(kT, a.w[0], y0T); }
        else if(S == 2) { st.aF[2][jc] = pf_fma((Real) 2, kF, old.aF[1][jc]); st.aT[2][jc] = pf_fma((Real) 2, kT, old.aT[1][jc]);     // k1 + 2*k2 (2*k exact)
                          nF = pf_fma(kF, a.w[1], y0F); nT = pf_fma(kT, a.w[1], y0T); }
        else if(S == 3) { st.aF[3][jc] = pf_fma((Real) 2, kF, old.aF[2][jc]); st.aT[3][jc] = pf_fma((Real) 2, kT, old.aT[2][jc]);
                          nF = pf_fma(kF, a.w[2], y0F); nT = pf_fma(kT, a.w[2], y0T); }
        else            { kF = old.aF[3][jc] + kF; kT = old.aT[3][jc] + kT; }  // simulation.cu:336-337, finished below with the raw y
    }
    else
    {
        if(S == 1)      { st.aF[1][jc] = kF; st.aT[1][jc] = kT;                                                         // k1
                          nF = pf_fma(kF, a.w[0], y0F); nT = pf_fma(kT, a.w[0], y0T); }                                 // y + tau/3 k1
        else if(S == 2) { st.aF[2][jc] = old.aF[1][jc]; st.aT[2][jc] = old.aT[1][jc];
                          nF = pf_fma(kF, a.w[1], pf_fma(old.aF[1][jc], a.w[1], y0F));                                  // y + tau/6 k1 + tau/6 k2
                          nT = pf_fma(kT, a.w[1], pf_fma(old.aT[1][jc], a.w[1], y0T)); }
        else if(S == 3) { st.aF[3][jc] = old.aF[2][jc]; st.aT[3][jc] = old.aT[2][jc];
                          st.bF[3][jc] = kF; st.bT[3][jc] = kT;                                                         // k3
                          nF = pf_fma(kF, a.w[3], pf_fma(old.aF[2][jc], a.w[2], y0F));                                  // y + tau/8 k1 + 3tau/8 k3
                          nT = pf_fma(kT, a.w[3], pf_fma(old.aT[2][jc], a.w[2], y0T)); }
        else if(S == 4) { nF = pf_fma(kF, a.w[6], pf_fma(old.bF[3][jc], a.w[5], pf_fma(old.aF[3][jc], a.w[4], y0F)));   // y + tau/2 k1 - 3tau/2 k3 + 2tau k4
                          nT = pf_fma(kT, a.w[6], pf_fma(old.bT[3][jc], a.w[5], pf_fma(old.aT[3][jc], a.w[4], y0T)));
                          // error combination so far: 0.2*k1 - 0.9*k3 + 0.8*k4 (simulation.cu:428); nvcc rounds 0.9*k3 and fuses the rest
                          st.bF[4][jc] = pf_fma((Real) 0.8, kF, pf_fma((Real) 0.2, old.aF[3][jc], -((Real) 0.9 * old.bF[3][jc])));
                          st.bT[4][jc] = pf_fma((Real) 0.8, kT, pf_fma((Real) 0.2, old.aT[3][jc], -((Real) 0.9 * old.bT[3][jc])));
                          st.aF[4][jc] = pf_fma((Real) 4, kF, old.aF[3][jc]); st.aT[4][jc] = pf_fma((Real) 4, kT, old.aT[3][jc]); }   // k1 + 4*k4 (exact product)
        else            { Real eF = pf_fma(-(Real) 0.1, kF, old.bF[4][jc]);
                          Real eT = pf_fma(-(Real) 0.1, kT, old.bT[4][jc]);
                          eF = eF >= 0 ? eF : -eF;
                          eT = eT >= 0 ? eT : -eT;
                          if(c.st_ok) { st.errF = eF > st.errF ? eF : st.errF; st.errT = eT > st.errT ? eT : st.errT; }
                          kF = old.aF[4][jc] + kF; kT = old.aT[4][jc] + kT; }                                           // simulation.cu:472, finished below
    }

    if(S < NST)
    {
        st.cF[S][jc] = nF; st.cT[S][jc] = nT;
        stream_store_pair<Real>(c.cbuf + (jc * (NST - 1) + (S - 1)) * ROWSTRIDE, c.wr_self, nF, nT);
    }
    else
    {
        if(SCHEME != SCHEME_EULER)
        {
            // y(n+1) = y + w*(combination) with the RAW y (a -0 stays -0 here, as in the reference's final combine)
            constexpr int d = 1 + (NST - 1);
            const Real rawF = with_sign(y0F, (old.negmask >> (2 * d)) & 1u);
            const Real rawT = with_sign(y0T, (old.negmask >> (2 * d + 1)) & 1u);
            const Real wl = SCHEME == SCHEME_RK4 ? a.w[3] : a.w[7];
            // The reference adds fu = 0 to every k_T, which turns a -0 into +0 and changes nothing else. Stage inputs are
            // insensitive to that sign (they add k*w to a canonical y, which is never -0), error norms take |.|, and a
            // combination of k's is -0 only if all of them are: adding 0 once, here, gives the same bits.
            kT = kT + a.k.fu;
            nF = pf_fma(wl, kF, rawF);
            nT = pf_fma(wl, kT, rawT);
        }
        // keep the last stage in the straight-line code of the row (the halo threads compute it too, only the store is predicated)
        keep_here(nF, nT);
        // element indices fit 32 bits (pf_create checks nx*ny); predicated stores instead of a branch keep the row one basic block
        // A branch, not predicated stores (measured: predicated st.global costs 2-4 %). Element indices fit 32 bits (pf_create).
        if(c.st_ok)
        {
            const unsigned idx = (unsigned)(row - a.row0) * (unsigned) a.nx + c.gxu;
            c.out_F[idx] = nF;
            c.out_U[idx] = nT;
        }
    }
}

template <typename Real, int SCHEME, int ANISO, int EW, bool CHECKED, int PH, int S>
struct StreamStages {
    static __device__ __forceinline__ void run(const FusedArgs<Real>& a, const StreamCtx<Real>& c,
                                               const StreamState<Real, SCHEME>& old, StreamState<Real, SCHEME>& st, const int it)
    {
        stream_stage<Real, SCHEME, ANISO, EW, CHECKED, PH, S>(a, c, old, st, it);
        if constexpr (S < SchemeStages<SCHEME>::value)
            StreamStages<Real, SCHEME, ANISO, EW, CHECKED, PH, S + 1>::run(a, c, old, st, it);
    }
};

// Local memory row of global row g: wrapped (single GPU, periodic) or clamped into the rows that exist locally and
// in the domain, so that loads of rows that only feed garbage stay inside the allocation.
template <typename Real>
__device__ __forceinline__ unsigned stream_mem_row(const FusedArgs<Real>& a, int g, const int rlo, const int rhi)
{
    if(a.wrap_y) { while(g < 0) g += a.ny_global; while(g >= a.ny_global) g -= a.ny_global; }
    g = g < rlo ? rlo : (g > rhi ? rhi : g);
    return (unsigned)(g - a.row0);
}

// One row of the march. PH = it mod U.
template <typename Real, int SCHEME, int ANISO, int EW, bool EDGE, bool COHERENT, bool CHECKED, int PH>
__device__ __forceinline__ void stream_iteration(const FusedArgs<Real>& a, const StreamCtx<Real>& c, StreamState<Real, SCHEME>& st,
                                                 const int it, const int rlo, const int rhi)
{
    using V2 = typename Vec2T<Real>::type;
    using G = StreamGeom<SCHEME>;
    constexpr int NST = G::NST, U = G::U, ROWSTRIDE = EW + 2;
    constexpr int j1 = pmod(PH + 1, U), j2 = pmod(PH + 2, U);
    const int g1 = c.G0 + it;
    // 1. prefetch the raw y of row g1 + 2, retired at the end of this iteration
    const unsigned ro = (CHECKED ? stream_mem_row(a, g1 + 1 + 1, rlo, rhi) : (unsigned)(g1 + 1 + 1 - a.row0)) * (unsigned) a.nx + c.lx;
    // COHERENT (multi-step launches): L2 only - another SM rewrites these rows every other step, and L1 is not coherent across SMs
    const Real tF = COHERENT ? __ldcg(c.F + ro) : c.F[ro], tT = COHERENT ? __ldcg(c.U + ro) : c.U[ro];
    if(!CHECKED && g1 + 2 + G::L2_AHEAD <= rhi)
    {
        // pull the rows this CTA reads a few iterations from now into L2 (measured: -1.5 % on the RK4 step at 2048^2)
        const unsigned po = ro + (unsigned) G::L2_AHEAD * (unsigned) a.nx;
        asm volatile("prefetch.global.L2 [%0];" :: "l"(c.F + po));
        asm volatile("prefetch.global.L2 [%0];" :: "l"(c.U + po));
    }
    // 2. publish y of row g1 + 1: the E/W neighbours of stage 1 in the NEXT iteration, and the y the later stages re-read
    V2* const ypub = c.ybuf + ((g1 + 1) & (G::YR - 1)) * ROWSTRIDE;
    stream_store_pair<Real>(ypub, c.wr_self, st.yF[j1], st.yT[j1]);
    // 3. all stages
    {
        const StreamState<Real, SCHEME> old = st;
        StreamStages<Real, SCHEME, ANISO, EW, CHECKED, PH, 1>::run(a, c, old, st, it);
    }
    // 3b. strips with a non-periodic domain edge column (EDGE), in the warp that holds it: the ghost pairs of everything
    // published in this row. The other strips run an instantiation without this block: its mere presence (a branch that cuts
    // the row's basic block in two) costs every CTA 6 % (measured, profiles/r02_cta_balance.md).
    if(EDGE && c.edge_warp)
    {
        stream_store_ghost<Real>(a, c, ypub, st.yF[j1], st.yT[j1]);
        #pragma unroll
        for(int S = 1; S < NST; S++)
            stream_store_ghost<Real>(a, c, c.cbuf + (pmod(PH - (S - 1), U) * (NST - 1) + (S - 1)) * ROWSTRIDE,
                                     st.cF[S][pmod(PH - (S - 1), U)], st.cT[S][pmod(PH - (S - 1), U)]);
    }
    // 4. retire the prefetch of row g1 + 2
    st.yF[j2] = canon_zero(tF); st.yT[j2] = canon_zero(tT);
    if(SCHEME != SCHEME_EULER)
        st.negmask = (st.negmask << 2) | (unsigned) is_negzero(tF) | ((unsigned) is_negzero(tT) << 1);
    // 5. one barrier per row: this iteration's writes become visible, the slots they will overwrite next have been read
    __syncthreads();
}

template <typename Real, int SCHEME, int ANISO, int EW, bool EDGE, bool COHERENT, bool CHECKED, int PH>
struct StreamIters {
    static __device__ __forceinline__ void run(const FusedArgs<Real>& a, const StreamCtx<Real>& c, StreamState<Real, SCHEME>& st,
                                               const int it0, const int rlo, const int rhi)
    {
        stream_iteration<Real, SCHEME, ANISO, EW, EDGE, COHERENT, CHECKED, PH>(a, c, st, it0 + PH, rlo, rhi);
        if constexpr (PH + 1 < StreamGeom<SCHEME>::U)
            StreamIters<Real, SCHEME, ANISO, EW, EDGE, COHERENT, CHECKED, PH + 1>::run(a, c, st, it0, rlo, rhi);
    }
};

// Number of column strips of width EW = EWI + 2H threads that cover nx columns (see the kernel for the ownership rule).
static inline __host__ __device__ int stream_strips(int nx, int EWI, int H, bool periodic)
{
    if(periodic) return (nx + EWI - 1) / EWI;
    const int n = (nx - 2 * H + EWI - 1) / EWI;                         // (n-1)*EWI + EW >= nx
    return n < 1 ? 1 : n;
}

template <typename Real, int SCHEME, int WARPS>
constexpr size_t stream_smem_bytes()
{
    using G = StreamGeom<SCHEME>;
    return sizeof(typename Vec2T<Real>::type) * (size_t)(G::YR + G::U * (G::NST - 1)) * (32 * WARPS + 2);
}

// Work decomposition: a CTA = one unit (1-D grid) = a strip of EW columns x a chunk of rows; the chunks of a strip are
// balanced (lengths differ by at most one row). Units in dispatch order: the two edge strips of a non-periodic grid first
// (their CTAs do a little more per row, and the CTAs dispatched first get the SM's oldest, i.e. fastest, resident slot), then
// the interior strips.
template <typename Real>
__device__ __forceinline__ void stream_unit_to_tile(const FusedArgs<Real>& a, const bool periodic, const int unit, int& strip, int& chunk)
{
    if(periodic || a.strips < 3) { strip = unit % a.strips; chunk = unit / a.strips; }
    else if(unit < 2 * a.nch)    { strip = (unit & 1) ? a.strips - 1 : 0; chunk = unit >> 1; }
    else                         { const int v = unit - 2 * a.nch; strip = 1 + v % (a.strips - 2); chunk = v / (a.strips - 2); }
}
template <typename Real>
__device__ __forceinline__ int stream_tile_to_unit(const FusedArgs<Real>& a, const bool periodic, const int strip, const int chunk)
{
    if(periodic || a.strips < 3) return strip + a.strips * chunk;
    if(strip == 0) return 2 * chunk;
    if(strip == a.strips - 1) return 2 * chunk + 1;
    return 2 * a.nch + (strip - 1) + (a.strips - 2) * chunk;
}

// One time step of one unit: y(n) = (F, U) -> y(n+1) = (out_F, out_U).
template <typename Real, int SCHEME, int ANISO, int WARPS, bool COHERENT>
__device__ __forceinline__ void stream_unit(const FusedArgs<Real>& a, const Real* in_F, const Real* in_U, Real* out_F, Real* out_U,
                                            typename Vec2T<Real>::type* const smem, const int strip, const int chunk)
{
    using V2 = typename Vec2T<Real>::type;
    using G = StreamGeom<SCHEME>;
    constexpr int NST = G::NST, U = G::U;
    constexpr int H = NST, EW = 32 * WARPS, EWI = EW - 2 * H, ROWSTRIDE = EW + 2;
    constexpr int RING_ELEMS = (G::YR + U * (NST - 1)) * ROWSTRIDE;
    const int cx = (int) threadIdx.x;

    const int nx = a.nx, NY = a.ny_global;
    const bool periodic = a.bc == PF_BOUNDARY_PERIODIC;
    const int own_rows = a.own_y1 - a.own_y0;

    StreamCtx<Real> c;
    c.F = in_F; c.U = in_U; c.out_F = out_F; c.out_U = out_U;
    // Periodic: strip k owns columns [k*EWI, (k+1)*EWI), its threads start H columns to the left (wrapped).
    // Non-periodic: nothing outside the domain is needed (the edge thread publishes the ghost pair itself), so thread 0 of
    // strip 0 IS column 0, and every column of the first / last strip up to the domain edge is clean:
    // strip k owns [k*EWI + (k ? H : 0), (k+1)*EWI + H), the last one everything up to nx - 1 (see stream_strips).
    const int gx0 = strip * EWI - (periodic ? H : 0);                  // column of thread 0
    const int gx = gx0 + cx;
    if(periodic) { int lx = gx % nx; if(lx < 0) lx += nx; c.lx = (unsigned) lx; }
    else         c.lx = (unsigned)(gx < 0 ? 0 : (gx >= nx ? nx - 1 : gx));
    c.x_lo_edge = !periodic && gx == 0;
    c.x_hi_edge = !periodic && gx == nx - 1;
    c.wr_self = periodic || (gx != -1 && gx != nx);
    const int cx_lo = (!periodic && gx0 == 0) ? 0 : H;                  // left domain edge inside the strip: no garbage enters from the left
    const int cx_hi = (!periodic && gx0 + EW >= nx) ? EW : EW - H;      // right domain edge inside the strip: none from the right
    c.st_ok = cx >= cx_lo && cx < cx_hi && gx < nx;
    c.edge_warp = __any_sync(0xffffffffu, c.x_lo_edge || c.x_hi_edge);
    const bool edge_strip = !periodic && (gx0 <= 0 || gx0 + EW >= nx);      // block-uniform
    c.gxu = (unsigned) gx;
    c.Y0 = a.own_y0 + (int)((long long) own_rows * chunk / a.nch);
    c.Y1 = a.own_y0 + (int)((long long) own_rows * (chunk + 1) / a.nch);
    const int n_it = (c.Y1 - c.Y0) + G::FILL;
    c.G0 = c.Y0 - (NST - 1);
    c.ylo = periodic ? INT_MIN : 0;
    c.yhi = periodic ? INT_MAX : NY;
    c.ybuf = smem + cx + 1;
    c.cbuf = smem + G::YR * ROWSTRIDE + cx + 1;
    // rows that exist both locally and in the domain (loads are clamped into them)
    const int rlo = periodic ? a.row0 : max(a.row0, 0);
    const int rhi = (periodic ? a.row0 + a.rows_local : min(a.row0 + a.rows_local, NY)) - 1;

    StreamState<Real, SCHEME> st;
    #pragma unroll
    for(int j = 0; j < U; j++)
    {
        st.yF[j] = st.yT[j] = 0;
        #pragma unroll
        for(int s = 0; s < NST; s++) { st.cF[s][j] = 0; st.cT[s][j] = 0; }
    }
    #pragma unroll
    for(int s = 0; s < NST; s++)
    {
        #pragma unroll
        for(int j = 0; j < U; j++) { st.aF[s][j] = st.aT[s][j] = 0; st.bF[s][j] = st.bT[s][j] = 0; }
    }
    st.negmask = 0; st.errF = 0; st.errT = 0;

    // zero the rings (pads included), then preload y rows G0-1, G0, G0+1
    for(int i = cx; i < RING_ELEMS; i += EW) { V2 z; z.x = 0; z.y = 0; smem[i] = z; }
    #pragma unroll
    for(int d = -1; d <= 1; d++)
    {
        const unsigned ro = stream_mem_row(a, c.G0 + d, rlo, rhi) * (unsigned) nx + c.lx;
        const Real f = COHERENT ? __ldcg(c.F + ro) : c.F[ro], u = COHERENT ? __ldcg(c.U + ro) : c.U[ro];
        const int j = d < 0 ? U - 1 : d;
        st.yF[j] = canon_zero(f); st.yT[j] = canon_zero(u);
        st.negmask = (st.negmask << 2) | (unsigned) is_negzero(f) | ((unsigned) is_negzero(u) << 1);
    }
    __syncthreads();
    {
        V2* const ypub = c.ybuf + (c.G0 & (G::YR - 1)) * ROWSTRIDE;
        stream_store_pair<Real>(ypub, c.wr_self, st.yF[0], st.yT[0]);
        if(c.edge_warp) stream_store_ghost<Real>(a, c, ypub, st.yF[0], st.yT[0]);
    }
    __syncthreads();

    // Iterations run in groups of U. A group takes the straight-line form when, in all of its iterations, every stage
    // is active, no stage row is a non-periodic domain edge row (or outside the domain), and the prefetched row
    // needs neither clamping nor wrapping.
    int lo_it = max(G::FILL, rlo - 1 - 1 - c.G0);
    int hi_it = min((c.Y1 - c.Y0) + 2 * NST - 2, rhi - 1 - c.G0);
    if(!periodic) { lo_it = max(lo_it, 1 - c.G0 + (NST - 1)); hi_it = min(hi_it, NY - 1 - c.G0); }
    if(edge_strip)
        for(int it0 = 0; it0 < n_it; it0 += U)
        {
            if(it0 >= lo_it && it0 + U <= hi_it) StreamIters<Real, SCHEME, ANISO, EW, true, COHERENT, false, 0>::run(a, c, st, it0, rlo, rhi);
            else                                 StreamIters<Real, SCHEME, ANISO, EW, true, COHERENT, true, 0>::run(a, c, st, it0, rlo, rhi);
        }
    else
        for(int it0 = 0; it0 < n_it; it0 += U)
        {
            if(it0 >= lo_it && it0 + U <= hi_it) StreamIters<Real, SCHEME, ANISO, EW, false, COHERENT, false, 0>::run(a, c, st, it0, rlo, rhi);
            else                                 StreamIters<Real, SCHEME, ANISO, EW, false, COHERENT, true, 0>::run(a, c, st, it0, rlo, rhi);
        }

    if(SCHEME == SCHEME_RKM)
    {
        const Real eF = warp_max(st.errF), eT = warp_max(st.errT);
        if((cx & 31) == 0)
        {
            if(eF > 0) atomic_max_nonneg(a.err_bits + 0, eF);
            if(eT > 0) atomic_max_nonneg(a.err_bits + 1, eT);
        }
        if(a.err_host)
        {
            // the host's step-size controller spins on err_host[2] instead of a device-to-host copy + stream synchronisation
            __syncthreads();
            if(cx == 0)
            {
                __threadfence();                                            // this unit's atomics before its ticket
                if(atomicAdd(a.err_ticket, 1u) == (unsigned)(a.strips * a.nch) - 1u)
                {
                    const unsigned long long e0 = atomicExch(a.err_bits + 0, 0ull), e1 = atomicExch(a.err_bits + 1, 0ull);
                    *a.err_ticket = 0;
                    volatile unsigned long long* h = a.err_host;
                    h[0] = e0; h[1] = e1;
                    __threadfence_system();
                    h[2] = a.err_seq;
                }
            }
        }
    }
}

// ---- multi-step launches -----------------------------------------------------------------------------------------------
// n_steps > 1: ONE launch advances n_steps fixed-size time steps. All units are co-resident (cooperative launch: one wave) and
// keep their tile; the buffers swap roles every step. There is no grid-wide barrier: a unit may start step m as soon as its
// 8 neighbouring tiles (and itself) have finished step m - 1 - their rows / columns are the only ones it reads (halo H <= the
// tile extents) and, the ping-pong being what it is, the only readers of what it is about to overwrite. progress[u] counts the
// steps unit u has finished (base + k; base grows from launch to launch, so the words are never reset).
// Writer: all stores, __syncthreads, thread 0: fence + st.release. Reader: ld.acquire by the polling threads, __syncthreads,
// then ld.global.cg (the fields are never read through L1, which is not coherent across SMs).
struct StreamRun {
    int n_steps;
    unsigned base;
    unsigned* progress;                // [units] + 1 word: set to 1 by a unit that gave up waiting (co-residency violated)
    int units;
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p)
{
    unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void st_release_u32(unsigned* p, unsigned v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// The two cold parts of a multi-step launch, as functions of their own: inlined around the march they would take part in its
// register allocation and scheduling (measured on the slab kernel: 8 %).
__device__ __noinline__ void stream_run_wait(const unsigned* progress, const int watch, const unsigned want, const int units)
{
    if(watch >= 0)
    {
        unsigned spins = 0;
        while((int)(ld_acquire_u32(progress + watch) - want) < 0)
        {
            __nanosleep(40);
            if(++spins > (1u << 24)) { const_cast<unsigned*>(progress)[units] = 1; break; }   // ~ a second: the grid is not co-resident; the host reports it
        }
    }
    __syncthreads();
}
__device__ __noinline__ void stream_run_done(unsigned* mine, const unsigned value)
{
    __syncthreads();
    if(threadIdx.x == 0) { __threadfence(); st_release_u32(mine, value); }
}

// One time step per launch.
template <typename Real, int SCHEME, int ANISO, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB)
k_stream_step(const FusedArgs<Real> a)
{
    using V2 = typename Vec2T<Real>::type;
    extern __shared__ __align__(16) unsigned char pf_smem_raw[];
    int strip, chunk;
    stream_unit_to_tile(a, a.bc == PF_BOUNDARY_PERIODIC, (int) blockIdx.x, strip, chunk);
    stream_unit<Real, SCHEME, ANISO, WARPS, false>(a, a.F, a.U, a.out_F, a.out_U, reinterpret_cast<V2*>(pf_smem_raw), strip, chunk);
}

// One time step of a slab with the halo hand-over to the neighbouring GPUs inside the kernel (see FusedArgs::halo_* and
// pf_halo.cuh): no separate wait / push launches, so a step is ONE launch and the hand-over of the lower rows - produced in
// the first iterations of the first chunk - overlaps the rest of the step. Requires chunks of at least halo_g rows.
// The two cold parts are separate functions (not inlined) so that they do not take part in the register allocation and
// scheduling of the march.
__device__ __noinline__ void stream_halo_wait(unsigned* flags, const unsigned e, const unsigned sides)
{
    if(threadIdx.x < 2 && (sides >> threadIdx.x & 1u))
    {
        unsigned v, spins = 0;
        for(;;)
        {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + threadIdx.x) : "memory");
            if((int)(v - e) >= 0) break;
            __nanosleep(100);
            if(++spins > (1u << 24)) { flags[2] = 1u + threadIdx.x; break; }   // HALO_FLAG_STALLED: the neighbour is gone
        }
    }
    __syncthreads();
}

template <typename Real>
__device__ __noinline__ void stream_halo_forward(const Real* out_F, const Real* out_U, Real* peer_F, Real* peer_U, const unsigned src0, const unsigned nx,
                                                 const int rows, const bool own, const unsigned gx, unsigned* ticket, const unsigned strips,
                                                 unsigned* peer_flag, const unsigned value)
{
    if(own)
        for(int r = 0; r < rows; r++)
        {
            peer_F[(unsigned) r * nx + gx] = __ldcg(out_F + src0 + (unsigned) r * nx + gx);
            peer_U[(unsigned) r * nx + gx] = __ldcg(out_U + src0 + (unsigned) r * nx + gx);
        }
    __syncthreads();
    if(threadIdx.x == 0)
    {
        __threadfence_system();                                             // this unit's peer stores before its ticket
        if(atomicAdd(ticket, 1u) == strips - 1u)
        {
            *ticket = 0;
            __threadfence_system();
            asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(peer_flag), "r"(value) : "memory");
        }
    }
}

template <typename Real, int SCHEME, int ANISO, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB)
k_stream_step_halo(const FusedArgs<Real> a)
{
    using V2 = typename Vec2T<Real>::type;
    extern __shared__ __align__(16) unsigned char pf_smem_raw[];
    constexpr int H = SchemeStages<SCHEME>::value, EW = 32 * WARPS, EWI = EW - 2 * H;
    int strip, chunk;
    stream_unit_to_tile(a, a.bc == PF_BOUNDARY_PERIODIC, (int) blockIdx.x, strip, chunk);
    {
        const unsigned sides = (chunk == 0 && a.peer_flag[0] ? 1u : 0u) | (chunk == a.nch - 1 && a.peer_flag[1] ? 2u : 0u);
        if(sides) stream_halo_wait(a.halo_flags, a.halo_e, sides);
    }
    stream_unit<Real, SCHEME, ANISO, WARPS, false>(a, a.F, a.U, a.out_F, a.out_U, reinterpret_cast<V2*>(pf_smem_raw), strip, chunk);
    {
        // the columns this strip stored (thread = column, as in stream_unit): every thread forwards what it wrote itself
        const bool periodic = a.bc == PF_BOUNDARY_PERIODIC;
        int strip2, chunk2;
        stream_unit_to_tile(a, periodic, (int) blockIdx.x, strip2, chunk2);
        const int gx0 = strip2 * EWI - (periodic ? H : 0), cx = (int) threadIdx.x, gx = gx0 + cx;
        const int cx_lo = (!periodic && gx0 == 0) ? 0 : H, cx_hi = (!periodic && gx0 + EW >= a.nx) ? EW : EW - H;
        const bool own = cx >= cx_lo && cx < cx_hi && gx < a.nx;
        if(chunk2 == 0 && a.peer_flag[0])
            stream_halo_forward<Real>(a.out_F, a.out_U, a.peer_F[0], a.peer_U[0], (unsigned)(a.own_y0 - a.row0) * (unsigned) a.nx, (unsigned) a.nx,
                                      a.halo_g, own, (unsigned) gx, a.halo_flags + 3, (unsigned) a.strips, a.peer_flag[0], a.halo_e + 1u);
        if(chunk2 == a.nch - 1 && a.peer_flag[1])
            stream_halo_forward<Real>(a.out_F, a.out_U, a.peer_F[1], a.peer_U[1], (unsigned)(a.own_y1 - a.halo_g - a.row0) * (unsigned) a.nx, (unsigned) a.nx,
                                      a.halo_g, own, (unsigned) gx, a.halo_flags + 4, (unsigned) a.strips, a.peer_flag[1], a.halo_e + 1u);
    }
}

// r.n_steps time steps per launch (cooperative launch: all units co-resident).
template <typename Real, int SCHEME, int ANISO, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB)
k_stream_run(const FusedArgs<Real> a, const StreamRun r)
{
    using V2 = typename Vec2T<Real>::type;
    extern __shared__ __align__(16) unsigned char pf_smem_raw[];
    V2* const smem = reinterpret_cast<V2*>(pf_smem_raw);
    const bool periodic = a.bc == PF_BOUNDARY_PERIODIC;
    const int unit = (int) blockIdx.x;
    int strip, chunk;
    stream_unit_to_tile(a, periodic, unit, strip, chunk);
    // the tile this thread polls (threads 0..7: the 8 neighbours; -1: none, e.g. beyond a non-periodic edge)
    int watch = -1;
    if(threadIdx.x < 8)
    {
        const int k = (int) threadIdx.x + (threadIdx.x >= 4 ? 1 : 0);      // 0..8 without the centre
        int s2 = strip + k % 3 - 1, c2 = chunk + k / 3 - 1;
        if(periodic) s2 = (s2 + a.strips) % a.strips;
        if(a.wrap_y) c2 = (c2 + a.nch) % a.nch;
        if(s2 >= 0 && s2 < a.strips && c2 >= 0 && c2 < a.nch) watch = stream_tile_to_unit(a, periodic, s2, c2);
        if(watch == unit) watch = -1;
    }
    for(int step = 0; step < r.n_steps; step++)
    {
        if(step > 0) stream_run_wait(r.progress, watch, r.base + (unsigned) step, r.units);
        const bool odd = step & 1;
        stream_unit<Real, SCHEME, ANISO, WARPS, true>(a, odd ? a.out_F : a.F, odd ? a.out_U : a.U,
                                                 odd ? const_cast<Real*>(a.F) : a.out_F, odd ? const_cast<Real*>(a.U) : a.out_U, smem, strip, chunk);
        stream_run_done(r.progress + unit, r.base + (unsigned) step + 1u);
    }
}

} // namespace pf
