"""Slab decomposition of the grid along y (rows are contiguous: index x + y*nx, reference simulation.cu:51).

The reference is single-GPU; this is the multi-GPU layer the B200 build adds. Rank g of G owns
global rows [g*ny/G, (g+1)*ny/G) plus `ghost` rows on every side that has a neighbour; `ghost` = the number of
right-hand-side evaluations one step of the solver makes (pf_required_ghost), because the explicit integrators are
temporally blocked: all stages of a step run in one kernel and recompute the overlap instead of exchanging per stage.
So there is ONE halo exchange per time step, of 2 fields x ghost rows x nx reals per neighbour.
"""
import ctypes as C

import numpy as np

from ._abi import BOUNDARY_PERIODIC, SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE, pf_slab
from . import solver as S


def slab_rows(ny, G, g):
    """Owned global rows [y0, y1) of rank g (balanced: the first ny % G ranks get one row more)."""
    base, rem = divmod(ny, G)
    y0 = g * base + min(g, rem)
    return y0, y0 + base + (1 if g < rem else 0)


def slab_layout(ny, G, g, ghost, periodic=False):
    y0, y1 = slab_rows(ny, G, g)
    multi = G > 1
    return pf_slab(y0=y0, ny_local=y1 - y0, has_lo=int(multi and (g > 0 or periodic)), has_hi=int(multi and (g < G - 1 or periodic)),
                   ghost=ghost if multi else 0)


def neighbours(G, g, periodic):
    """(lo, hi) neighbour ranks or None."""
    if G == 1:
        return None, None
    lo = g - 1 if g > 0 else (G - 1 if periodic else None)
    hi = g + 1 if g < G - 1 else (0 if periodic else None)
    return lo, hi


def step_slabs_lockstep(solvers, params, state):
    """One time step of every slab held by THIS process (single-device emulation of the multi-rank step, used by
    the tests): state[g] = [F, U, next_F, next_U]. Returns the dt advanced. Ghost rows of the inputs must be current."""
    L = S.lib()
    if params.solver != SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE:
        for s, f in zip(solvers, state):
            s.step_async(params, f[0], f[1], f[2], f[3])
        for s in solvers:
            s.sync()
        return params.dt
    taus = []
    for s, f in zip(solvers, state):
        tau = C.c_double(0)
        s._ck(L.pf_rkm_begin(s._h, C.byref(params), f[0], f[1], C.byref(tau)))
        taus.append(tau.value)
    assert len(set(taus)) == 1
    tau = taus[0]
    max_iters = max(params.T_max_iters, params.Phi_max_iters, 1)
    used = tau
    for _ in range(max_iters):
        bits = np.zeros(2, dtype=np.uint64)
        for s, f in zip(solvers, state):
            s._ck(L.pf_rkm_attempt_async(s._h, C.byref(params), tau, f[0], f[1], f[2], f[3]))
        for s in solvers:
            s.sync()
            local = np.zeros(2, dtype=np.uint64)
            s._ck(L.pf_download_raw(s._h, L.pf_rkm_error_bits(s._h), local.ctypes.data, 16))
            bits = np.maximum(bits, local)                      # the cross-rank MAX all-reduce
        conv, stop, nxt = C.c_int(0), C.c_int(0), C.c_double(0)
        solvers[0]._ck(L.pf_rkm_decide(solvers[0]._h, C.byref(params), tau, int(bits[0]), int(bits[1]),
                                       C.byref(conv), C.byref(stop), C.byref(nxt), None, None))
        used, tau = tau, nxt.value
        if stop.value:
            break
    for s, f in zip(solvers, state):
        s._ck(L.pf_rkm_finish_async(s._h, C.byref(params), used, tau, f[0], f[1], f[2], f[3]))
        s.sync()
    return used


# ---- semi-implicit solver on slabs -------------------------------------------------------------------
def semi_implicit_step_phases(solvers, params, state, bufs, gather, exchange_direction, poll_every=2):
    """One semi-implicit time step of the slabs THIS process holds (one per rank under NCCL; all of them in the
    single-device emulation of the tests), following the phase protocol of include/pf_solver.h ("semi-implicit solver on
    slabs"). The library does the arithmetic; this function only sequences the phases and the two kinds of communication:

      gather(solvers, src_ptrs, dst_ptrs) : all-gather of the per-block partial sums, `width` reals per rank (zero padded),
                                            rank order, into every rank's dst; called after the local launches are enqueued
      exchange_direction(solvers)         : fill the single ghost row of each rank's search direction p from its neighbours

    state[g] = [F, U, next_F, next_U]; bufs[g] = {"part", "rr0", "rr1", "pAp"} device pointers of width / n_all reals
    ("part" zero-filled beyond pf_si_partials() once, by the caller). Returns ((Phi_iters, T_iters), (Phi_conv, T_conv),
    (Phi_error, T_error)), identical on every rank. Ghost rows of F and U must be current on entry; the caller exchanges
    next_F / next_U afterwards.
    """
    L = S.lib()
    n_all = bufs[0]["n_all"]
    for s, f in zip(solvers, state):
        s._ck(L.pf_si_begin_async(s._h, C.byref(params), f[0], f[1], f[1], f[2], f[3]))
    iters, conv, errs = [], [], []
    for which, max_iters in ((0, params.Phi_max_iters), (1, params.T_max_iters)):
        for s, b in zip(solvers, bufs):
            s._ck(L.pf_si_cg_begin_async(s._h, which, b["part"]))
        cur, new = "rr0", "rr1"
        gather(solvers, [b["part"] for b in bufs], [b[cur] for b in bufs])
        done = False
        for it in range(max_iters):
            exchange_direction(solvers)
            for s, b in zip(solvers, bufs):
                s._ck(L.pf_si_cg_spmv_async(s._h, which, b["part"]))
            gather(solvers, [b["part"] for b in bufs], [b["pAp"] for b in bufs])
            for s, b in zip(solvers, bufs):
                s._ck(L.pf_si_cg_update_async(s._h, which, b[cur], b["pAp"], n_all, b["part"]))
            gather(solvers, [b["part"] for b in bufs], [b[new] for b in bufs])
            for s, b in zip(solvers, bufs):
                s._ck(L.pf_si_cg_pupdate_async(s._h, which, b[cur], b[new], n_all, it))
            cur, new = new, cur
            if (it + 1) % poll_every == 0 and it + 1 < max_iters:
                d = C.c_int(0)
                flags = []
                for s in solvers:
                    s._ck(L.pf_si_cg_poll(s._h, which, C.byref(d), None, None, None))
                    flags.append(d.value)
                assert len(set(flags)) == 1, "ranks disagree on CG convergence"      # same gathered sums -> same decision
                if flags[0]:
                    done = True
                    break
        res = []
        for s in solvers:
            d, i, c, e = C.c_int(0), C.c_int(0), C.c_int(0), C.c_double(0)
            s._ck(L.pf_si_cg_poll(s._h, which, C.byref(d), C.byref(i), C.byref(c), C.byref(e)))
            res.append((i.value, c.value, e.value))
        assert len(set(res)) == 1, res
        iters.append(res[0][0]); conv.append(res[0][1]); errs.append(res[0][2])
    return tuple(iters), tuple(conv), tuple(errs)


def semi_implicit_buffers(solvers, world_widths=None):
    """Allocates the partial-sum buffers of semi_implicit_step_phases as solver fields (zero-filled by pf_field_alloc).
    world_widths: pf_si_partials() of EVERY rank of the job, rank order (default: the local solvers are the whole job)."""
    L = S.lib()
    widths = world_widths if world_widths is not None else [L.pf_si_partials(s._h) for s in solvers]
    width = max(widths)
    n_all = width * len(widths)
    bufs = []
    for s in solvers:
        assert n_all <= s.nx * s.rows_alloc, "grid too small for the partial-sum buffers"
        bufs.append({"part": s.field_alloc(), "rr0": s.field_alloc(), "rr1": s.field_alloc(), "pAp": s.field_alloc(),
                     "width": width, "n_all": n_all})
    return bufs


def make_host_gather(width):
    """The all-gather of the single-device emulation: through the host (tests only)."""
    L = S.lib()

    def gather(solvers, src, dst):
        parts = []
        for s, p in zip(solvers, src):
            s.sync()
            a = np.empty(width, dtype=S.np_dtype(s.dtype))
            s._ck(L.pf_download_raw(s._h, p, a.ctypes.data, a.nbytes))
            parts.append(a)
        allp = np.concatenate(parts)
        for s, d in zip(solvers, dst):
            s._ck(L.pf_upload_raw(s._h, d, allp.ctypes.data, allp.nbytes))
    return gather


def host_exchange_rows(solvers, slabs, fields, ny, periodic):
    """Ghost rows of one field per slab, filled from the owners' rows through the host (tests only)."""
    nx = solvers[0].nx
    full = np.empty((ny, nx), dtype=S.np_dtype(solvers[0].dtype))
    local = []
    for sl, s, f in zip(slabs, solvers, fields):
        a = s.download(f)
        local.append(a)
        full[sl.y0:sl.y0 + sl.ny_local] = a[s.first_owned_row:s.first_owned_row + sl.ny_local]
    for sl, s, f, a in zip(slabs, solvers, fields, local):
        rows = np.arange(s.rows_alloc) - s.first_owned_row + sl.y0
        rows = rows % ny if periodic else np.clip(rows, 0, ny - 1)
        s.upload(f, full[rows])
    return full
