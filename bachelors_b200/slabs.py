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
