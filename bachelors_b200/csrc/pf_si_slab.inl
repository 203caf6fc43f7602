// pf_si_slab.inl - the semi-implicit step (reference simulation.cu:732-926) in phases, for y-slabs on several GPUs
// (SURVEY section 8e). The library holds no communication code: between the phases the caller
//   * exchanges ONE ghost row of the search direction p with its neighbours (before every SpMV), and
//   * all-gathers the per-block partial sums of the dot products, so that every rank folds the SAME list of
//     partials in the SAME order (sum_partials): alpha, beta and the convergence test come out bit-identical on all
//     ranks and the ranks stay in lock step without exchanging a decision.
// The kernels are the per-kernel CG path of the single-GPU solver (pf_general.cuh) restricted to the owned rows; the
// ghost rows of the slab allocation stand in for boundary_sample's y-neighbours, the global boundary rule applies
// only where the slab touches the domain edge. Included at the end of pf_solver.cu.

namespace {

template <typename Real>
struct SiSlab {
    Real *b_F, *b_U, *factor, *r, *p, *Ap;
    size_t off, n_own;          // first owned element, owned elements
    int y0, y1, rows, blocks;
};

template <typename Real>
int si_slab_views(pf_solver* s, SiSlab<Real>& v)
{
    void* q[6];
    for(int i = 0; i < 6; i++) if(int rc = get_scratch(s, 2 + i, &q[i])) return rc;
    v.b_F = (Real*) q[0]; v.b_U = (Real*) q[1]; v.factor = (Real*) q[2]; v.r = (Real*) q[3]; v.p = (Real*) q[4]; v.Ap = (Real*) q[5];
    v.rows = s->rows_alloc; v.y0 = s->ghost_lo; v.y1 = s->ghost_lo + s->slab.ny_local;
    v.off = (size_t) s->ghost_lo * (size_t) s->nx;
    v.n_own = (size_t) s->slab.ny_local * (size_t) s->nx;
    v.blocks = cg_blocks(s, s->nx, s->slab.ny_local);
    return PF_OK;
}

template <typename Real>
CgMatrix<Real> si_slab_matrix(const pf_solver* s, const SiSlab<Real>& v, int which)
{
    CgMatrix<Real> A = make_matrix<Real>(s->si.p, which, which == 0 ? v.factor : nullptr);   // coefficients from the GLOBAL grid
    A.ny = v.rows; A.y_begin = v.y0; A.y_end = v.y1;
    return A;
}

template <typename Real>
int si_slab_begin(pf_solver* s)
{
    SiSlab<Real> v; if(int rc = si_slab_views<Real>(s, v)) return rc;
    const pf_params& p = s->si.p;
    const KConsts<Real> k = make_consts<Real>(p);
    dim3 block(32, 8), grid((unsigned)((s->nx + 31) / 32), (unsigned)((s->slab.ny_local + 7) / 8));
    const Real *F = (const Real*) s->si.F, *U = (const Real*) s->si.U;
    if(p.S != 0) k_si_prepare<Real, true><<<grid, block, 0, s->stream>>>(F, U, v.b_F, v.factor, s->nx, v.rows, p.Phi_boundary, p.T_boundary, k, v.y0, v.y1);
    else         k_si_prepare<Real, false><<<grid, block, 0, s->stream>>>(F, U, v.b_F, v.factor, s->nx, v.rows, p.Phi_boundary, p.T_boundary, k, v.y0, v.y1);
    PF_LAUNCH_CHECK(s);
    return PF_OK;
}

template <typename Real>
int si_slab_cg_begin(pf_solver* s, int which, void* partials_out)
{
    SiSlab<Real> v; if(int rc = si_slab_views<Real>(s, v)) return rc;
    const pf_params& p = s->si.p;
    const CgMatrix<Real> A = si_slab_matrix<Real>(s, v, which);
    const Real *x0, *b; Real* x; int max_iters;
    if(which == 0) { x0 = (const Real*) s->si.F; b = v.b_F; x = (Real*) s->si.nF; max_iters = p.Phi_max_iters; }
    else
    {
        const Real dt = (Real) p.dt, gamma = (Real) p.gamma;                 // b_U, simulation.cu:893-899
        k_si_bU<Real><<<elementwise_blocks(s, v.n_own, 256), 256, 0, s->stream>>>(v.n_own, (Real) p.L, dt * (1 - gamma),
            (const Real*) s->si.U_base + v.off, (const Real*) s->si.F + v.off, (const Real*) s->si.nF + v.off, v.b_U + v.off);
        PF_LAUNCH_CHECK(s);
        x0 = (const Real*) s->si.U; b = v.b_U; x = (Real*) s->si.nU; max_iters = p.T_max_iters;
    }
    k_cg_init<Real><<<v.blocks, CG_THREADS, 0, s->stream>>>(A, x0, b, x, v.r, v.p, (Real*) partials_out, s->d_cg + which, max_iters);
    PF_LAUNCH_CHECK(s);
    return PF_OK;
}

template <typename Real>
int si_slab_cg_spmv(pf_solver* s, int which, void* partials_out)
{
    SiSlab<Real> v; if(int rc = si_slab_views<Real>(s, v)) return rc;
    const CgMatrix<Real> A = si_slab_matrix<Real>(s, v, which);
    k_cg_spmv<Real><<<v.blocks, CG_THREADS, 0, s->stream>>>(A, v.p, v.Ap, (Real*) partials_out, s->d_cg + which);
    PF_LAUNCH_CHECK(s);
    return PF_OK;
}

template <typename Real>
int si_slab_cg_update(pf_solver* s, int which, const void* rr_all, const void* pAp_all, int n_all, void* partials_out)
{
    SiSlab<Real> v; if(int rc = si_slab_views<Real>(s, v)) return rc;
    Real* x = (Real*)(which == 0 ? s->si.nF : s->si.nU);
    k_cg_update<Real><<<v.blocks, CG_THREADS, 0, s->stream>>>(v.n_own, (Real) 1.0e-12, x + v.off, v.r + v.off, v.p + v.off, v.Ap + v.off,
        (const Real*) rr_all, (const Real*) pAp_all, (Real*) partials_out, n_all, s->d_cg + which);
    PF_LAUNCH_CHECK(s);
    return PF_OK;
}

template <typename Real>
int si_slab_cg_pupdate(pf_solver* s, int which, const void* rr_all, const void* rr_new_all, int n_all, int iter)
{
    SiSlab<Real> v; if(int rc = si_slab_views<Real>(s, v)) return rc;
    const pf_params& p = s->si.p;
    const Real tol = (Real)(which == 0 ? p.Phi_tolerance : p.T_tolerance);
    const Real sst = tol * tol * (Real)(int)((size_t) p.nx * (size_t) p.ny);          // simulation.cu:608 with the GLOBAL N
    k_cg_pupdate<Real><<<v.blocks, CG_THREADS, 0, s->stream>>>(v.n_own, (Real) 1.0e-12, sst, v.r + v.off, v.p + v.off,
        (const Real*) rr_all, (const Real*) rr_new_all, n_all, s->d_cg + which, iter);
    PF_LAUNCH_CHECK(s);
    return PF_OK;
}

} // namespace

#define PF_SI_DISPATCH(fn, ...) (s->dtype == PF_F64 ? fn<double>(__VA_ARGS__) : fn<float>(__VA_ARGS__))

extern "C" {

int pf_si_begin_async(pf_solver* s, const pf_params* params, const void* F, const void* U, const void* U_base, void* next_F, void* next_U)
{
    if(int rc = bind(s)) return rc;
    if(int rc = check_step_args(s, params, F, U, next_F, next_U)) return rc;
    if(params->solver != PF_SOLVER_SEMI_IMPLICIT) return fail(s, PF_ERR_INVALID, "pf_si_begin_async: solver must be semi-implicit");
    if(params->do_corrector_loop) return fail(s, PF_ERR_UNSUPPORTED, "pf_si_begin_async: the corrector loop is single-GPU only");
    if((s->slab.has_lo || s->slab.has_hi) && s->slab.ghost < 1) return fail(s, PF_ERR_INVALID, "pf_si_begin_async: the slab needs one ghost row");
    s->si.p = *params; s->si.F = F; s->si.U = U; s->si.U_base = U_base ? U_base : U; s->si.nF = next_F; s->si.nU = next_U; s->si.active = true;
    return PF_SI_DISPATCH(si_slab_begin, s);
}

int pf_si_partials(pf_solver* s)
{
    if(!s) return 0;
    return cg_blocks(s, s->nx, s->slab.ny_local);
}

#define PF_SI_CHECK(which) do { if(int rc = bind(s)) return rc; \
    if(!s->si.active) return fail(s, PF_ERR_INVALID, "%s: call pf_si_begin_async first", __func__); \
    if((which) != 0 && (which) != 1) return fail(s, PF_ERR_INVALID, "%s: which must be 0 (Phi) or 1 (T)", __func__); } while(0)

int pf_si_cg_begin_async(pf_solver* s, int which, void* partials_out)
{
    PF_SI_CHECK(which);
    if(!partials_out) return fail(s, PF_ERR_INVALID, "pf_si_cg_begin_async: NULL partials");
    return PF_SI_DISPATCH(si_slab_cg_begin, s, which, partials_out);
}

int pf_si_cg_spmv_async(pf_solver* s, int which, void* partials_out)
{
    PF_SI_CHECK(which);
    if(!partials_out) return fail(s, PF_ERR_INVALID, "pf_si_cg_spmv_async: NULL partials");
    return PF_SI_DISPATCH(si_slab_cg_spmv, s, which, partials_out);
}

int pf_si_cg_update_async(pf_solver* s, int which, const void* rr_all, const void* pAp_all, int n_all, void* partials_out)
{
    PF_SI_CHECK(which);
    if(!rr_all || !pAp_all || !partials_out || n_all <= 0) return fail(s, PF_ERR_INVALID, "pf_si_cg_update_async: bad argument");
    return PF_SI_DISPATCH(si_slab_cg_update, s, which, rr_all, pAp_all, n_all, partials_out);
}

int pf_si_cg_pupdate_async(pf_solver* s, int which, const void* rr_all, const void* rr_new_all, int n_all, int iter)
{
    PF_SI_CHECK(which);
    if(!rr_all || !rr_new_all || n_all <= 0) return fail(s, PF_ERR_INVALID, "pf_si_cg_pupdate_async: bad argument");
    return PF_SI_DISPATCH(si_slab_cg_pupdate, s, which, rr_all, rr_new_all, n_all, iter);
}

void* pf_si_direction(pf_solver* s)
{
    if(!s || bind(s)) return nullptr;
    void* q = nullptr;
    return get_scratch(s, 2 + 4, &q) ? nullptr : q;
}

int pf_si_cg_poll(pf_solver* s, int which, int* done, int* iters, int* converged, double* error)
{
    PF_SI_CHECK(which);
    PF_CUDA(s, cudaMemcpyAsync(s->h_cg + which, s->d_cg + which, sizeof(CgState), cudaMemcpyDeviceToHost, s->stream));
    PF_CUDA(s, cudaStreamSynchronize(s->stream));
    const pf_params& p = s->si.p;
    const int max_iters = which == 0 ? p.Phi_max_iters : p.T_max_iters;
    const double N = (double) p.nx * (double) p.ny;
    if(done) *done = s->h_cg[which].done;
    if(iters) *iters = s->h_cg[which].iters;
    if(converged) *converged = s->h_cg[which].iters != max_iters;
    if(error) *error = s->dtype == PF_F64 ? sqrt(s->h_cg[which].rr / N) : (double) sqrtf((float) s->h_cg[which].rr / (float) N);
    return PF_OK;
}

} // extern "C"
