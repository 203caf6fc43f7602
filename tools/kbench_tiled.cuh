// kbench_tiled.cuh - the 2-D tiled single-launch kernel of round 1 (formerly bachelors_b200/csrc/pf_fused.cuh), kept as the
// independent cross-check of tools/kbench.cu: not part of the product library.
//
// One CTA owns an extended tile of EW x EH cells = interior (EW-2H) x (EH-2H) plus a halo of H = #stages cells on every side.
// Thread (cx, ty) owns column cx, rows [ty*R, ty*R+R) of the extended tile and keeps, in REGISTERS, for each of its R cells: the
// raw state y, the current stage input, and the scheme's running combination of the k's. Only the stage inputs {Phi,T} travel
// through shared memory (two ping-pong tiles of interleaved {Phi,T} pairs). Stage s is evaluated on the tile shrunk by s cells,
// so after H stages exactly the interior is final. Requires Phi_boundary == T_boundary.
#pragma once
#include "../bachelors_b200/csrc/pf_common.cuh"
#include "../bachelors_b200/csrc/pf_fused.cuh"

namespace pf {

// The stage loop. FAST = the extended tile lies strictly inside the domain and inside the local rows, and no y(n)
// value of the tile is -0: then every cell is computed unconditionally (cells in the shrinking rim hold garbage that
// never reaches a valid cell), there is no boundary logic, and canon_zero(y) == y. Otherwise the general form runs.
// Both forms evaluate the identical arithmetic on every valid cell.
template <typename Real, int SCHEME, bool ANISO, int EW, int TY, int R, bool FAST>
__device__ __forceinline__ void fused_stages(const FusedArgs<Real>& a, typename Vec2T<Real>::type* const buf0, typename Vec2T<Real>::type* const buf1,
    const int cx, const int ry0, const int gx, const int gy0, const unsigned in_mask,
    const Real (&yF)[R], const Real (&yT)[R], Real (&cF)[R], Real (&cT)[R], Real (&bF)[R], Real (&bT)[R])
{
    constexpr int NST = SchemeStages<SCHEME>::value;
    constexpr int EH = TY * R;
    using V2 = typename Vec2T<Real>::type;
    const int nx = a.nx, NY = a.ny_global;
    const int bc = a.bc;
    const bool periodic = bc == PF_BOUNDARY_PERIODIC;
    const bool x_lo_edge = !periodic && gx == 0;
    const bool x_hi_edge = !periodic && gx == nx - 1;
    Real aF[R], aT[R];      // running combination (RK4: k1+2k2+2k3; RKM: k1, later k1+4k4)
    #pragma unroll
    for(int j = 0; j < R; j++) { aF[j] = 0; aT[j] = 0; }

    #pragma unroll
    for(int s = 1; s <= NST; s++)
    {
        const V2* rb = (s & 1) ? buf0 : buf1;
        V2* wb = (s & 1) ? buf1 : buf0;
        const bool col_ok = FAST ? true : (cx >= s && cx < EW - s);
        Real nF[R], nT[R];
        #pragma unroll
        for(int j = 0; j < R; j++)
        {
            nF[j] = cF[j]; nT[j] = cT[j];
            const int ry = ry0 + j;
            const bool row_ok = ry >= s && ry < EH - s;          // warp-uniform
            if(row_ok && (FAST || (col_ok && ((in_mask >> j) & 1u))))
            {
                const Real CF = cF[j], CT = cT[j];
                V2 e = rb[ry * EW + cx + 1];
                V2 w = rb[ry * EW + cx - 1];
                Real NF, NT, SF, ST;
                if(j < R - 1) { NF = cF[j + 1]; NT = cT[j + 1]; } else { V2 n = rb[(ry + 1) * EW + cx]; NF = n.x; NT = n.y; }
                if(j > 0)     { SF = cF[j - 1]; ST = cT[j - 1]; } else { V2 q = rb[(ry - 1) * EW + cx]; SF = q.x; ST = q.y; }
                if(!FAST && !periodic)
                {
                    const int gy = gy0 + j;
                    if(x_hi_edge)    { e.x = ghost_of(CF, bc); e.y = ghost_of(CT, bc); }
                    if(x_lo_edge)    { w.x = ghost_of(CF, bc); w.y = ghost_of(CT, bc); }
                    if(gy == NY - 1) { NF = ghost_of(CF, bc); NT = ghost_of(CT, bc); }
                    if(gy == 0)      { SF = ghost_of(CF, bc); ST = ghost_of(CT, bc); }
                }
                Real kF, kT;
                rhs_cell<Real, ANISO>(a.k, CF, e.x, w.x, NF, SF, CT, e.y, w.y, NT, ST, kF, kT);

                const Real y0F = FAST ? yF[j] : canon_zero(yF[j]), y0T = FAST ? yT[j] : canon_zero(yT[j]);
                if(SCHEME == SCHEME_EULER)
                {
                    // simulation.cu:233-234: C + dt*k, C being the gathered (blended) value
                    nF[j] = pf_fma(a.k.dt, kF, CF);
                    nT[j] = pf_fma(a.k.dt, kT, CT);
                }
                else if(SCHEME == SCHEME_RK4)
                {
                    if(s == 1)      { aF[j] = kF; aT[j] = kT;
                                      nF[j] = pf_fma(kF, a.w[0], y0F); nT[j] = pf_fma(kT, a.w[0], y0T); }
                    else if(s == 2) { aF[j] = pf_fma((Real) 2, kF, aF[j]); aT[j] = pf_fma((Real) 2, kT, aT[j]);   // k1 + 2*k2 (2*k exact)
                                      nF[j] = pf_fma(kF, a.w[1], y0F); nT[j] = pf_fma(kT, a.w[1], y0T); }
                    else if(s == 3) { aF[j] = pf_fma((Real) 2, kF, aF[j]); aT[j] = pf_fma((Real) 2, kT, aT[j]);
                                      nF[j] = pf_fma(kF, a.w[2], y0F); nT[j] = pf_fma(kT, a.w[2], y0T); }
                    else            { // simulation.cu:336-337: F + dt/6*(k1 + 2*k2 + 2*k3 + k4)
                                      nF[j] = pf_fma(a.w[3], aF[j] + kF, yF[j]);
                                      nT[j] = pf_fma(a.w[3], aT[j] + kT, yT[j]); }
                }
                else
                {
                    if(s == 1)      { aF[j] = kF; aT[j] = kT;                                                     // k1
                                      nF[j] = pf_fma(kF, a.w[0], y0F); nT[j] = pf_fma(kT, a.w[0], y0T); }         // y + tau/3 k1
                    else if(s == 2) { nF[j] = pf_fma(kF, a.w[1], pf_fma(aF[j], a.w[1], y0F));                     // y + tau/6 k1 + tau/6 k2
                                      nT[j] = pf_fma(kT, a.w[1], pf_fma(aT[j], a.w[1], y0T)); }
                    else if(s == 3) { bF[j] = kF; bT[j] = kT;                                                     // k3
                                      nF[j] = pf_fma(kF, a.w[3], pf_fma(aF[j], a.w[2], y0F));                     // y + tau/8 k1 + 3tau/8 k3
                                      nT[j] = pf_fma(kT, a.w[3], pf_fma(aT[j], a.w[2], y0T)); }
                    else if(s == 4) { nF[j] = pf_fma(kF, a.w[6], pf_fma(bF[j], a.w[5], pf_fma(aF[j], a.w[4], y0F)));   // y + tau/2 k1 - 3tau/2 k3 + 2tau k4
                                      nT[j] = pf_fma(kT, a.w[6], pf_fma(bT[j], a.w[5], pf_fma(aT[j], a.w[4], y0T)));
                                      // error combination so far: 0.2*k1 - 0.9*k3 + 0.8*k4 (simulation.cu:428)
                                      // (nvcc rounds the product 0.9*k3 and fuses the rest; see oracle/pf_oracle_impl.h)
                                      bF[j] = pf_fma((Real) 0.8, kF, pf_fma((Real) 0.2, aF[j], -((Real) 0.9 * bF[j])));
                                      bT[j] = pf_fma((Real) 0.8, kT, pf_fma((Real) 0.2, aT[j], -((Real) 0.9 * bT[j])));
                                      aF[j] = pf_fma((Real) 4, kF, aF[j]); aT[j] = pf_fma((Real) 4, kT, aT[j]); } // k1 + 4*k4 (exact product)
                    else            { Real eF = pf_fma(-(Real) 0.1, kF, bF[j]);
                                      Real eT = pf_fma(-(Real) 0.1, kT, bT[j]);
                                      bF[j] = eF >= 0 ? eF : -eF;
                                      bT[j] = eT >= 0 ? eT : -eT;
                                      nF[j] = pf_fma(a.w[7], aF[j] + kF, yF[j]);                                  // simulation.cu:472
                                      nT[j] = pf_fma(a.w[7], aT[j] + kT, yT[j]); }
                }
                if(s < NST)
                {
                    V2 v; v.x = nF[j]; v.y = nT[j];
                    wb[ry * EW + cx] = v;
                }
            }
        }
        #pragma unroll
        for(int j = 0; j < R; j++) { cF[j] = nF[j]; cT[j] = nT[j]; }
        if(s < NST) __syncthreads();
    }
}

template <typename Real, int SCHEME, bool ANISO, int EW, int TY, int R>
__global__ void __launch_bounds__(EW * TY, 1)
k_fused_step(const FusedArgs<Real> a)
{
    constexpr int NST = SchemeStages<SCHEME>::value;
    constexpr int H = NST;
    constexpr int EH = TY * R;
    using V2 = typename Vec2T<Real>::type;
    extern __shared__ __align__(16) unsigned char pf_smem_raw[];
    // one pad element in front and behind each tile: the FAST form reads cx-1 / cx+1 at the tile's corner cells
    V2* const buf0 = reinterpret_cast<V2*>(pf_smem_raw) + 1;
    V2* const buf1 = buf0 + EW * EH + 2;

    const int cx = threadIdx.x;
    const int ry0 = threadIdx.y * R;
    const int nx = a.nx, NY = a.ny_global;
    const bool periodic = a.bc == PF_BOUNDARY_PERIODIC;

    const int tile_x0 = (int) blockIdx.x * (EW - 2 * H) - H;
    const int tile_y0 = a.own_y0 + (int) blockIdx.y * (EH - 2 * H) - H;
    const int gx = tile_x0 + cx;
    const int gy0 = tile_y0 + ry0;

    // block-uniform: the whole extended tile is inside the domain and inside this rank's rows, no wrap anywhere
    const bool interior = tile_x0 >= 0 && tile_x0 + EW <= nx && tile_y0 >= 0 && tile_y0 + EH <= NY &&
                          tile_y0 - a.row0 >= 0 && tile_y0 - a.row0 + EH <= a.rows_local;

    Real yF[R], yT[R];      // raw y(n)
    Real cF[R], cT[R];      // this thread's stage inputs
    Real bF[R], bT[R];      // RKM only: k3, later the error combination
    unsigned in_mask = 0;
    int lx = gx;
    bool fast = false;

    if(interior)
    {
        int negzero = 0;
        const size_t base = (size_t)(gy0 - a.row0) * (size_t) nx + (size_t) gx;
        #pragma unroll
        for(int j = 0; j < R; j++)
        {
            const Real f = a.F[base + (size_t) j * (size_t) nx], u = a.U[base + (size_t) j * (size_t) nx];
            yF[j] = f; yT[j] = u;
            bF[j] = bT[j] = 0;
        }
        // -0 compares equal to +0, so test the bit patterns
        #pragma unroll
        for(int j = 0; j < R; j++) negzero |= (int) is_negzero(yF[j]) | (int) is_negzero(yT[j]);
        #pragma unroll
        for(int j = 0; j < R; j++)
        {
            cF[j] = canon_zero(yF[j]); cT[j] = canon_zero(yT[j]);
            V2 v; v.x = cF[j]; v.y = cT[j];
            buf0[(ry0 + j) * EW + cx] = v;
        }
        in_mask = (1u << R) - 1u;
        fast = __syncthreads_or(negzero) == 0;
    }
    else
    {
        bool x_in;
        if(periodic) { lx = gx % nx; if(lx < 0) lx += nx; x_in = true; }
        else         { x_in = gx >= 0 && gx < nx; lx = gx < 0 ? 0 : (gx >= nx ? nx - 1 : gx); }
        #pragma unroll
        for(int j = 0; j < R; j++)
        {
            const int gy = gy0 + j;
            int ly; bool y_in;
            if(a.wrap_y) { ly = gy % NY; if(ly < 0) ly += NY; y_in = true; }
            else         { ly = gy - a.row0; y_in = ly >= 0 && ly < a.rows_local && (periodic || (gy >= 0 && gy < NY)); }
            const bool in = x_in && y_in;
            Real f = 0, u = 0;
            if(in)
            {
                const size_t idx = (size_t) ly * (size_t) nx + (size_t) lx;
                f = a.F[idx]; u = a.U[idx];
                in_mask |= 1u << j;
            }
            yF[j] = f; yT[j] = u;
            cF[j] = canon_zero(f); cT[j] = canon_zero(u);
            bF[j] = bT[j] = 0;
            V2 v; v.x = cF[j]; v.y = cT[j];
            buf0[(ry0 + j) * EW + cx] = v;
        }
        __syncthreads();
    }

    if(fast) fused_stages<Real, SCHEME, ANISO, EW, TY, R, true>(a, buf0, buf1, cx, ry0, gx, gy0, in_mask, yF, yT, cF, cT, bF, bT);
    else     fused_stages<Real, SCHEME, ANISO, EW, TY, R, false>(a, buf0, buf1, cx, ry0, gx, gy0, in_mask, yF, yT, cF, cT, bF, bT);

    // write the interior: cF/cT now hold y(n+1) for every cell that survived all H stages
    Real err_F = 0, err_T = 0;
    const bool col_int = cx >= H && cx < EW - H;
    #pragma unroll
    for(int j = 0; j < R; j++)
    {
        const int ry = ry0 + j;
        const int gy = gy0 + j;
        if(col_int && gx >= 0 && gx < nx && ry >= H && ry < EH - H && ((in_mask >> j) & 1u) && gy >= a.own_y0 && gy < a.own_y1)
        {
            int ly;
            if(a.wrap_y) { ly = gy % NY; if(ly < 0) ly += NY; } else ly = gy - a.row0;
            const size_t idx = (size_t) ly * (size_t) nx + (size_t) lx;
            a.out_F[idx] = cF[j];
            a.out_U[idx] = cT[j];
            if(SCHEME == SCHEME_RKM)
            {
                err_F = bF[j] > err_F ? bF[j] : err_F;
                err_T = bT[j] > err_T ? bT[j] : err_T;
            }
        }
    }
    if(SCHEME == SCHEME_RKM)
    {
        err_F = warp_max(err_F);
        err_T = warp_max(err_T);
        if(((threadIdx.x + threadIdx.y * EW) & 31) == 0)
        {
            if(err_F > 0) atomic_max_nonneg(a.err_bits + 0, err_F);
            if(err_T > 0) atomic_max_nonneg(a.err_bits + 1, err_T);
        }
    }
}

} // namespace pf
