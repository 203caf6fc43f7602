"""Multi-GPU stepping: one process per GPU (torch.distributed), the grid cut into y-slabs (slabs.py).

The reference is single-GPU; this layer is added by the B200 build and must reproduce the
single-grid results bit for bit (the explicit integrators' only cross-cell reductions are max-norms).

Per time step and rank:
  1. the first / last `ghost` owned rows are produced first, on the solver's auxiliary (high-priority) stream;
  2. they are handed to the neighbours, while
  3. the interior rows are produced on the solver's main stream.
`ghost` = number of right-hand-side evaluations per step (the integrators are temporally blocked), so there is ONE
exchange per step. The adaptive solver adds one 16-byte integer MAX all-reduce per attempt (slabs.py / pf_solver.h).

Transport of step 2, fixed-step explicit solvers: DEVICE-INITIATED over NVLink peer memory (pf_halo_push_async /
pf_halo_wait_async, csrc/pf_halo.cuh): a small kernel stores the rows straight into the neighbour's ghost rows through a
cudaIpc mapping and publishes the exchange number into a flag word there; the neighbour's next edge launch acquires it.
No NCCL call and no host synchronisation per step. (NCCL's send/recv kernels cannot get SM slots beside the interior kernel's
one wave of long-lived CTAs, so that exchange really ran after the interior: 43 us of a 518 us step at 2 GPUs.) NCCL
send/recv remains for the initial ghost fill, the adaptive and the semi-implicit solver.

torch is plumbing only: device tensors for the fields (their data_ptr() goes through the C-ABI), streams, NCCL.
"""
import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from . import solver as S
from ._abi import (BOUNDARY_PERIODIC, PF_F64, SOLVER_TYPE_EXPLICIT_EULER, SOLVER_TYPE_EXPLICIT_RK4,
                   SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE, pf_step_info)
from .slabs import neighbours, slab_layout, slab_rows


def exchange_halos(fields, ghost, first_owned, ny_local, lo, hi, group=None):
    """Fills the ghost rows of every tensor in `fields` (shape (rows_alloc, nx)) from the neighbours' edge rows and sends
    ours. lo / hi are neighbour ranks or None. Works for CPU tensors over gloo and CUDA tensors over NCCL.
    Returns the list of outstanding work handles (already waited on for CUDA stream semantics)."""
    if ghost == 0 or (lo is None and hi is None):
        return []
    ops = []
    own0, own1 = first_owned, first_owned + ny_local
    # sends first, lo then hi; receives hi then lo: with two ranks and periodic boundaries both neighbours are the same
    # peer, and messages between one pair match in posting order (my to-lo rows are the peer's from-hi rows).
    for f in fields:
        if lo is not None:
            ops.append(dist.P2POp(dist.isend, f[own0:own0 + ghost], lo, group))
    for f in fields:
        if hi is not None:
            ops.append(dist.P2POp(dist.isend, f[own1 - ghost:own1], hi, group))
    for f in fields:
        if hi is not None:
            ops.append(dist.P2POp(dist.irecv, f[own1:own1 + ghost], hi, group))
    for f in fields:
        if lo is not None:
            ops.append(dist.P2POp(dist.irecv, f[own0 - ghost:own0], lo, group))
    works = dist.batch_isend_irecv(ops)
    for w in works:
        w.wait()
    return works


def allreduce_max_bits(bits, group=None):
    """MAX all-reduce of the error-norm bit patterns (int64 view: non-negative IEEE values order like integers)."""
    dist.all_reduce(bits, op=dist.ReduceOp.MAX, group=group)
    return bits


class _DevArray:
    """Exposes raw device memory to torch through __cuda_array_interface__."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class SlabRunner:
    """One rank's slab of a global nx x ny grid; steps in lock step with the other ranks of `group`."""

    def __init__(self, params, dtype=PF_F64, rank=None, world=None, device=None, group=None, overlap=True, p2p=True, halo_in_kernel=None):
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world is None else world
        self.group = group
        self.params = params.copy()
        self.dtype = dtype
        self.device = torch.cuda.current_device() if device is None else device
        self.overlap = overlap
        L = S.lib()
        self.ghost = L.pf_required_ghost(params.solver) if self.world > 1 else 0
        self.periodic = params.Phi_boundary == BOUNDARY_PERIODIC
        self.layout = slab_layout(params.ny, self.world, self.rank, self.ghost, self.periodic)
        self.lo, self.hi = neighbours(self.world, self.rank, self.periodic)
        self.solver = S.Solver(params.nx, params.ny, dtype, device=self.device, slab=self.layout)
        if halo_in_kernel is not None:
            self.solver.set_option(3, int(halo_in_kernel))          # PF_OPT_HALO_IN_KERNEL
        s = self.solver
        tdt = torch.float64 if dtype == PF_F64 else torch.float32
        dev = torch.device("cuda", self.device)
        # the four ping-pong fields are the library's allocations (cudaMalloc: exportable to the neighbours' processes), seen
        # through torch tensors so that NCCL can address their rows too
        typestr = "<f8" if dtype == PF_F64 else "<f4"
        self._field_ptrs = [s.field_alloc() for _ in range(4)]
        self.F, self.U, self.nF, self.nU = (torch.as_tensor(_DevArray(q, (s.rows_alloc, params.nx), typestr), device=dev) for q in self._field_ptrs)
        self._ptr_index = {t.data_ptr(): i for i, t in enumerate((self.F, self.U, self.nF, self.nU))}
        self.main = torch.cuda.ExternalStream(s.stream, device=dev)
        self.aux = torch.cuda.ExternalStream(L.pf_aux_stream(s._h), device=dev)
        self.comm = torch.cuda.Stream(device=dev)
        self.ev_halo = torch.cuda.Event()
        self.ev_aux = torch.cuda.Event()
        self.ev_main = torch.cuda.Event()
        self.ev_halo.record(self.comm)
        self.ev_aux.record(self.aux)
        self.ev_main.record(self.main)
        self.bits = torch.as_tensor(_DevArray(L.pf_rkm_error_bits(s._h), (2,), "<i8"), device=dev)
        self.h_bits = torch.zeros(2, dtype=torch.int64).pin_memory()
        self.time = float(params.time)
        self.iter = int(params.iter)
        self.last_attempts = 0
        y0, y1 = self.layout.y0, self.layout.y0 + self.layout.ny_local
        g = self.ghost
        self.windows_edge = []
        if self.layout.has_lo:
            self.windows_edge.append((y0, min(y0 + g, y1)))
        if self.layout.has_hi:
            self.windows_edge.append((max(y1 - g, y0 + (g if self.layout.has_lo else 0)), y1))
        self.window_mid = (y0 + (g if self.layout.has_lo else 0), y1 - (g if self.layout.has_hi else 0))
        self.ev_edge = torch.cuda.Event()
        self.ev_edge.record(self.aux)
        self.p2p = bool(p2p and overlap and self.world > 1 and params.solver in (SOLVER_TYPE_EXPLICIT_EULER, SOLVER_TYPE_EXPLICIT_RK4, SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE))
        if self.p2p:
            self._p2p_setup()

    # -- device-initiated halo exchange: peer mappings --------------------------------------------------------------
    def _p2p_setup(self):
        L, s = S.lib(), self.solver
        flags = L.pf_halo_flags(s._h)
        if not flags:
            raise S.PfError(-4, "pf_halo_flags: out of memory")

        def export(ptr):
            h = C.create_string_buffer(64)
            s._ck(L.pf_ipc_export(s._h, ptr, h))
            return h.raw

        mine = {"fields": [export(q) for q in self._field_ptrs], "flags": export(flags),
                "ghost_lo": s.first_owned_row, "ny_local": self.layout.ny_local}
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=self.group)
        opened = {}

        def open_peer(r):
            if r not in opened:
                def op(blob):
                    out = C.c_void_p()
                    s._ck(L.pf_ipc_open(s._h, C.create_string_buffer(blob, 64), C.byref(out)))
                    return out.value
                opened[r] = {"fields": [op(b) for b in everyone[r]["fields"]], "flags": op(everyone[r]["flags"]),
                             "ghost_lo": everyone[r]["ghost_lo"], "ny_local": everyone[r]["ny_local"]}
            return opened[r]

        lo = open_peer(self.lo) if self.lo is not None else None     # my first rows -> its UPPER ghost rows (row ghost_lo + ny_local of its allocation)
        hi = open_peer(self.hi) if self.hi is not None else None     # my last rows -> its LOWER ghost rows (row 0)
        n = len(self._field_ptrs)
        arr = lambda v: (C.c_void_p * n)(*v)
        s._ck(L.pf_halo_link(s._h, n, arr(self._field_ptrs), arr(lo["fields"]) if lo else None, (lo["ghost_lo"] + lo["ny_local"]) if lo else 0,
                             lo["flags"] if lo else None, arr(hi["fields"]) if hi else None, hi["flags"] if hi else None))
        self._opened = opened
        dist.barrier(group=self.group)

    def _step_p2p(self, p, F, U, nF, nU):
        """ONE call, and for explicit-rk4 ONE kernel: the linked solver (pf_halo_link) waits for the neighbours' rows and hands its own
        edge rows over from inside the step kernel. Nothing here waits on the host or on NCCL.
        The slab is NOT cut into edge and interior launches: the step kernel is one wave of long-lived CTAs, and an edge launch that
        takes some of its slots for its first 20 us delays the CTAs that could not start by as much - the whole step ends that much
        later (measured at 2 GPUs: 518 us against 474 us for the slab alone, with NCCL and with peer stores alike)."""
        L, s = S.lib(), self.solver
        self.main.wait_event(self.ev_halo)
        s._ck(L.pf_step_async(s._h, C.byref(p), F, U, nF, nU))
        self.ev_main.record(self.main)

    # -- data movement -----------------------------------------------------------------------------------------
    def owned(self, t):
        f = self.solver.first_owned_row
        return t[f:f + self.layout.ny_local]

    def upload_global(self, F_host, U_host):
        """F_host / U_host: this rank's OWNED rows (ny_local, nx) as numpy arrays; ghosts come from the first exchange."""
        for t, h in ((self.F, F_host), (self.U, U_host)):
            self.owned(t).copy_(torch.from_numpy(np.ascontiguousarray(h)).to(t.dtype))
        torch.cuda.synchronize(self.device)
        self.exchange(self.F, self.U)
        torch.cuda.synchronize(self.device)

    def download_owned(self):
        self.sync()
        return self.owned(self.F).cpu().numpy(), self.owned(self.U).cpu().numpy()

    def exchange(self, F, U):
        with torch.cuda.stream(self.comm):
            exchange_halos([F, U], self.ghost, self.solver.first_owned_row, self.layout.ny_local, self.lo, self.hi, self.group)
            self.ev_halo.record(self.comm)

    def sync(self):
        for st in (self.main, self.aux, self.comm):
            st.synchronize()
        if self.p2p:
            self.solver._ck(S.lib().pf_halo_check(self.solver._h))

    # -- stepping ----------------------------------------------------------------------------------------------------
    def _ptrs(self):
        return self.F.data_ptr(), self.U.data_ptr(), self.nF.data_ptr(), self.nU.data_ptr()

    def _launch(self, fn_edge, fn_mid):
        """Edge rows on aux, exchange on comm, interior rows on main. The inputs' ghost rows are ready at ev_halo."""
        self.aux.wait_event(self.ev_halo); self.aux.wait_event(self.ev_main)
        self.main.wait_event(self.ev_halo); self.main.wait_event(self.ev_aux)
        for (a, b) in self.windows_edge:
            if b > a:
                fn_edge(a, b)
        self.ev_aux.record(self.aux)
        a, b = self.window_mid
        if b > a:
            fn_mid(a, b)
        self.ev_main.record(self.main)
        self.comm.wait_event(self.ev_aux)
        if not self.overlap:
            self.comm.wait_event(self.ev_main)
        self.exchange(self.nF, self.nU)

    def step(self):
        """One time step of the configured solver; returns dt advanced (host float)."""
        p = self.params
        p.time, p.iter = self.time, self.iter
        L, s = S.lib(), self.solver
        F, U, nF, nU = self._ptrs()
        if self.p2p and p.solver != SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE:
            self._step_p2p(p, F, U, nF, nU)
            dt = p.dt
            self.last_attempts = 1
        elif p.solver in (SOLVER_TYPE_EXPLICIT_EULER, SOLVER_TYPE_EXPLICIT_RK4):
            self._launch(lambda a, b: s._ck(L.pf_step_rows_async(s._h, C.byref(p), F, U, nF, nU, a, b, 1)),
                         lambda a, b: s._ck(L.pf_step_rows_async(s._h, C.byref(p), F, U, nF, nU, a, b, 0)))
            dt = p.dt
            self.last_attempts = 1
        elif p.solver == SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE and self.world > 1:
            dt = self._step_adaptive(p, F, U, nF, nU)
        elif p.solver == SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE:
            # one rank: the library's own attempt / decide loop (one 16-byte read-back per attempt, no cross-stream hops)
            info = pf_step_info()
            s._ck(L.pf_step(s._h, C.byref(p), F, U, nF, nU, C.byref(info)))
            dt = info.dt_advanced
            self.last_attempts = info.Phi_iters
        elif self.world > 1:
            dt = self._step_semi_implicit(p, F, U, nF, nU)
        else:
            # one sim_step-shaped call: returns after the device finished (the CG iteration counts come back with it)
            info = pf_step_info()
            s._ck(L.pf_step(s._h, C.byref(p), F, U, nF, nU, C.byref(info)))
            dt = info.dt_advanced
            self.last_attempts = 1
            self.last_cg_iters = (info.Phi_iters, info.T_iters)
        self.F, self.nF = self.nF, self.F
        self.U, self.nU = self.nU, self.U
        self.time += dt
        self.iter += 1
        return dt

    def _step_adaptive(self, p, F, U, nF, nU):
        L, s = S.lib(), self.solver
        tau = C.c_double(0)
        s._ck(L.pf_rkm_begin(s._h, C.byref(p), F, U, C.byref(tau)))
        tau = tau.value
        used = tau
        max_iters = max(p.T_max_iters, p.Phi_max_iters, 1)
        attempts = 0
        for _ in range(max_iters):
            t = tau
            if self.p2p:
                # linked slab: ONE attempt kernel for the whole slab, which also hands the candidate rows to the neighbours
                self.main.wait_event(self.ev_halo)
                s._ck(L.pf_rkm_attempt_async(s._h, C.byref(p), t, F, U, nF, nU))
                self.ev_main.record(self.main)
                self.ev_aux.record(self.aux)
            else:
                # zero the error words before either stream's kernels can touch them
                self.aux.wait_event(self.ev_main)
                s._ck(L.pf_rkm_clear_errors_async(s._h, 1))
                self.ev_aux.record(self.aux)
                self._launch(lambda a, b: s._ck(L.pf_rkm_attempt_rows_async(s._h, C.byref(p), t, F, U, nF, nU, a, b, 1)),
                             lambda a, b: s._ck(L.pf_rkm_attempt_rows_async(s._h, C.byref(p), t, F, U, nF, nU, a, b, 0)))
            with torch.cuda.stream(self.comm):
                self.comm.wait_event(self.ev_main)
                self.comm.wait_event(self.ev_aux)
                if self.world > 1:
                    allreduce_max_bits(self.bits, self.group)
                self.h_bits.copy_(self.bits, non_blocking=True)
            self.comm.synchronize()
            conv, stop, nxt = C.c_int(0), C.c_int(0), C.c_double(0)
            s._ck(L.pf_rkm_decide(s._h, C.byref(p), tau, int(self.h_bits[0]), int(self.h_bits[1]),
                                  C.byref(conv), C.byref(stop), C.byref(nxt), None, None))
            used, tau = tau, nxt.value
            if stop.value != 2:
                attempts += 1
            if stop.value:
                break
        s._ck(L.pf_rkm_finish_async(s._h, C.byref(p), used, tau, F, U, nF, nU))
        self.last_attempts = attempts
        return used

    # -- semi-implicit solver on slabs: phases of the library + all-gathered partial sums + one-row exchange of p ------------
    def _si_setup(self):
        L, s = S.lib(), self.solver
        dev = torch.device("cuda", self.device)
        tdt = torch.float64 if self.dtype == PF_F64 else torch.float32
        mine = torch.tensor([L.pf_si_partials(s._h)], dtype=torch.int64, device=dev)
        widths = [torch.zeros_like(mine) for _ in range(self.world)]
        dist.all_gather(widths, mine, group=self.group)
        width = max(int(w.item()) for w in widths)
        n_all = width * self.world
        t = {k: torch.zeros(n, dtype=tdt, device=dev) for k, n in (("part", width), ("rr0", n_all), ("rr1", n_all), ("pAp", n_all))}
        self._si_t = t
        self._si_by_ptr = {v.data_ptr(): v for v in t.values()}
        self._si_bufs = [{k: v.data_ptr() for k, v in t.items()} | {"width": width, "n_all": n_all}]
        typestr = "<f8" if self.dtype == PF_F64 else "<f4"
        self._si_dir = torch.as_tensor(_DevArray(L.pf_si_direction(s._h), (s.rows_alloc, self.params.nx), typestr), device=dev)
        torch.cuda.synchronize(self.device)

    def _step_semi_implicit(self, p, F, U, nF, nU):
        from .slabs import semi_implicit_step_phases
        if not hasattr(self, "_si_bufs"):
            self._si_setup()
        s = self.solver

        def gather(_solvers, src, dst):
            dist.all_gather_into_tensor(self._si_by_ptr[dst[0]], self._si_by_ptr[src[0]], group=self.group)

        def exchange_direction(_solvers):
            exchange_halos([self._si_dir], 1, s.first_owned_row, self.layout.ny_local, self.lo, self.hi, self.group)

        # everything in the order of the solver's own stream: its kernels, the collectives, the row exchanges
        self.main.wait_event(self.ev_halo)
        with torch.cuda.stream(self.main):
            iters, conv, errs = semi_implicit_step_phases([s], p, [[F, U, nF, nU]], self._si_bufs, gather, exchange_direction)
            exchange_halos([self.nF, self.nU], self.ghost, s.first_owned_row, self.layout.ny_local, self.lo, self.hi, self.group)
            self.ev_halo.record(self.main)
            self.ev_main.record(self.main)
        self.last_attempts = 1
        self.last_cg_iters = iters
        self.last_cg_errors = errs
        return p.dt

    def close(self):
        self.sync()
        if self.p2p:
            # nobody may free memory a neighbour can still be pushing into
            dist.barrier(group=self.group)
            L, s = S.lib(), self.solver
            for q in self._opened.values():
                for m in q["fields"] + [q["flags"]]:
                    L.pf_ipc_close(s._h, m)
            L.pf_halo_link(s._h, 0, None, None, 0, None, None, None)
            self._opened, self.p2p = {}, False
        self.solver.close()
