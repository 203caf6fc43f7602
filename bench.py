#!/usr/bin/env python
"""Benchmark of the hot path: GLUPS (1e9 cell updates per second) of one time step of the phase-field solver.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--solver explicit-rk4] [--dtype f64]

N = 1: workload = benchmark_configs/config_<solver>_2048.ini (default explicit-rk4: BASELINE.json's metric config), fp64.
N > 1 (launched by torchrun, one rank per GPU): the 16384-wide synthetic family of BASELINE.json configs[4], cut into
y-slabs of 2048 rows per GPU (16384 x 2048*N cells; N = 8 is the 16384^2 grid), weak scaling, halo exchange over NCCL.
A "step" is one time step of the solver over the whole grid. Synthetic data: the config's own dendrite seed.

One JSON line on stdout (rank 0). See DESIGN.md "Measurement" for how every key is produced.
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from bachelors_b200._abi import PF_F32, PF_F64, SOLVER_NAMES                # noqa: E402
from bachelors_b200.config import read_config                               # noqa: E402
from bachelors_b200.initial import make_initial_conditions                  # noqa: E402

PER_GPU_ROWS = 2048
WIDE_NX = 16384


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic(solver, dtype_name, size=2048):
    """Per-launch DRAM bytes of the dominant kernel from the committed ncu capture (profiles/traffic.json), or None."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get(f"{solver}_{dtype_name}_{size}")
    return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._halt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(0.004)

    def finish(self):
        self._halt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def physical_gpu_index(local):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local])
        except Exception:
            return local
    return local


def workload(args, world):
    """-> (cfg, description). The config file supplies every parameter; for N > 1 only the grid is widened."""
    size = args.size or (1024 if args.solver == "semi-implicit" else 2048)      # BASELINE.json configs: semi-implicit is quoted at 1024^2
    cfg = read_config(os.path.join(ROOT, "benchmark_configs", f"config_{args.solver}_{size}.ini"))
    if world == 1 and not args.wide:
        return cfg, f"config_{args.solver}_{size}.ini ({args.dtype})"
    p = cfg.params
    dx = p.L0 / p.nx
    p.nx, p.ny = WIDE_NX, PER_GPU_ROWS * world
    # the reference has one L0 for both axes (dx = L0/nx, dy = L0/ny); keep dx = 4/512 and let dy follow
    p.L0 = dx * p.nx
    cfg.init_circle_center = (p.L0 / 2, p.L0 / 2)
    return cfg, f"synthetic {args.solver} {p.nx}x{p.ny} ({args.dtype}), {PER_GPU_ROWS}-row slab per GPU"


def config_dict(args, world, p, desc, rs):
    """`config` of the JSON line: identical in both arms (the driver compares them)."""
    return {"workload": desc, "solver": args.solver, "nx": p.nx, "ny": p.ny, "parallelism": f"y-slabs x{world}",
            "l2_policy": ("working set (2 fields in + 2 out) %.0f MB per GPU vs 126 MB L2; consecutive time steps, no flush"
                          % (4 * rs * p.nx * p.ny / world / 1e6)),
            "halo_overlap": (not args.no_overlap) if world > 1 else None}


def load_inst_counts(solver, dtype_name, size):
    """Executed warp instructions per launch of the dominant kernel from the committed ncu capture (profiles/inst_counts.json)."""
    path = os.path.join(ROOT, "profiles", "inst_counts.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get(f"{solver}_{dtype_name}_{size}")
    return None


# ------------------------------------------------------------------------------------------------------------------------------
# reference arm: the CPU oracle port (the reference has no CPU implementation of this path, see DESIGN.md "Oracle")
# ------------------------------------------------------------------------------------------------------------------------------
def cpu_window(cfg, dtype, w):
    """A centred w x w window of the workload as its own grid with L0 scaled by w/nx: dx, dy and therefore every cell's
    arithmetic are those of the full grid (a bounded sample of the same workload)."""
    p = cfg.params.copy()
    npdt = np.float64 if dtype == PF_F64 else np.float32
    F0, U0 = make_initial_conditions(cfg, npdt)
    if w and w < min(p.nx, p.ny):
        x0, y0 = (p.nx - w) // 2, (p.ny - w) // 2
        F0 = np.ascontiguousarray(F0[y0:y0 + w, x0:x0 + w]); U0 = np.ascontiguousarray(U0[y0:y0 + w, x0:x0 + w])
        p.L0 = p.L0 * w / p.nx
        p.nx = p.ny = w
    return p, F0, U0


def cpu_port_rate(cfg, dtype, budget_s, max_steps, w=None):
    """Times oracle steps on the host cores. -> (GLUPS, cores, sample text)."""
    from oracle.oracle import Oracle, lib as olib
    olib().orc_set_threads(len(os.sched_getaffinity(0)))      # every core this process may use, whatever OMP_NUM_THREADS says
    cores = olib().orc_max_threads()
    p, F, U = cpu_window(cfg, dtype, w)
    o = Oracle(dtype, p.nx, p.ny)
    F, U, _, _ = o.step(p, F, U)                     # warm-up (allocates the oracle's scratch fields)
    n, t0 = 0, time.perf_counter()
    while n < max_steps and (time.perf_counter() - t0) < budget_s:
        F, U, _, _ = o.step(p, F, U)
        n += 1
    secs = time.perf_counter() - t0
    sample = f"{n} step(s) of {p.nx}x{p.ny} {('fp64' if dtype == PF_F64 else 'fp32')} in {secs:.1f} s on {cores} OpenMP thread(s)"
    return p.nx * p.ny * n / secs / 1e9, cores, sample


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return 0
    dtype = PF_F64 if args.dtype == "f64" else PF_F32
    rs = 8 if dtype == PF_F64 else 4
    cfg_job, desc = workload(args, world)              # the CUDA arm's workload at this N: the line's config
    from oracle.oracle import Oracle, lib as olib
    # all the host threads this process may use: torchrun exports OMP_NUM_THREADS=1 to its workers, and at N > 1 rank 0 is the
    # only rank that works in this arm
    olib().orc_set_threads(len(os.sched_getaffinity(0)))
    cores = olib().orc_max_threads()
    npdt = np.float64 if dtype == PF_F64 else np.float32
    if world == 1:
        # the WHOLE grid of the workload every step (2048^2: ~60 ms per explicit-rk4 step on 16 cores)
        p = cfg_job.params.copy()
        F, U = make_initial_conditions(cfg_job, npdt)
        what = f"each step = one {args.solver} step of the whole {p.nx}x{p.ny} grid of the workload"
    else:
        # bounded sample per step: ONE rank's slab of the job (16384 x 2048 cells = 1/N of the cells, the same arithmetic per cell)
        wide = argparse.Namespace(**vars(args)); wide.wide = True
        cfg1, _ = workload(wide, 1)
        p = cfg1.params.copy()
        F, U = make_initial_conditions(cfg1, npdt)
        what = f"each step = one {args.solver} step of one rank's {p.nx}x{p.ny} slab (1/{world} of the job's cells)"
    o = Oracle(dtype, p.nx, p.ny)
    for _ in range(max(args.warmup, 1)):
        F, U, _, _ = o.step(p, F, U)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        F, U, _, _ = o.step(p, F, U)
    secs = time.perf_counter() - t0
    glups = p.nx * p.ny * args.steps / secs / 1e9
    sample = f"{what}, {cores} OpenMP threads"
    line = {
        "impl": "reference", "metric": "GLUPS", "value": glups, "unit": "1e9 cell updates/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic (config's dendrite seed; no dataset)",
        "config": config_dict(args, world, cfg_job.params, desc, rs),
        "note": "the reference has no CPU implementation of this path (CUDA only); this arm times the CPU oracle port of it "
                "(oracle/, C + OpenMP) on the host cores",
        "cpu_baseline": {"value": glups, "unit": "1e9 cell updates/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": glups, "unit": "1e9 cell updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------------------------------------
# the CUDA arm
# ------------------------------------------------------------------------------------------------------------------------------
def time_device_steps(runner, torch, steps, world, dist):
    """K steps bracketed by barrier + synchronize, timed with CUDA events on the solver's stream; max over ranks."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    runner.sync()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record(runner.main)
    t_adv, attempts = 0.0, 0
    runner.cg_iters_total = 0
    for _ in range(steps):
        t_adv += runner.step()
        attempts += runner.last_attempts
        runner.cg_iters_total += sum(getattr(runner, "last_cg_iters", (0, 0)))
    # the step's tail may live on the aux / comm streams: make main wait for them before the closing event
    runner.main.wait_event(runner.ev_aux)
    runner.main.wait_event(runner.ev_halo)
    e1.record(runner.main)
    runner.sync()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
    return ms, t_adv, attempts


def time_e2e(runner, torch, steps, world, dist):
    """The reference-facing call with HOST buffers: every step hands y(n) in pinned host memory to pf_step_host (= sim_modify
    UPLOAD x2 + sim_step + sim_modify DOWNLOAD x2, main.cpp:164-168/:553/:857) and gets y(n+1) back in host memory; the output
    buffers of one step are the input buffers of the next. Wall clock bracketed by synchronize; max over ranks.
    N > 1: each rank moves its own slab; the ghost rows of the uploaded state are exchanged before the step."""
    own_F, own_U = runner.owned(runner.F), runner.owned(runner.U)
    hA = [torch.empty(own_F.shape, dtype=own_F.dtype).pin_memory() for _ in range(2)]
    hB = [torch.empty(own_F.shape, dtype=own_F.dtype).pin_memory() for _ in range(2)]
    hA[0].copy_(own_F); hA[1].copy_(own_U)
    nbytes = hA[0].numel() * hA[0].element_size() * 2
    from bachelors_b200 import solver as S
    from bachelors_b200._abi import pf_step_info
    s, L = runner.solver, S.lib()
    dt_code = runner.dtype
    p = runner.params
    runner.sync()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        if world == 1:
            info = pf_step_info()
            s._ck(L.pf_step_host(s._h, C.byref(p), hA[0].data_ptr(), hA[1].data_ptr(), hB[0].data_ptr(), hB[1].data_ptr(), dt_code,
                                 runner.F.data_ptr(), runner.U.data_ptr(), runner.nF.data_ptr(), runner.nU.data_ptr(), C.byref(info)))
            runner.F, runner.nF = runner.nF, runner.F
            runner.U, runner.nU = runner.nU, runner.U
        else:
            s._ck(L.pf_upload(s._h, runner.owned(runner.F).data_ptr(), hA[0].data_ptr(), hA[0].numel(), dt_code))
            s._ck(L.pf_upload(s._h, runner.owned(runner.U).data_ptr(), hA[1].data_ptr(), hA[1].numel(), dt_code))
            runner.exchange(runner.F, runner.U)          # the uploaded rows' halos (host data changed the owned rows)
            runner.step()
            runner.sync()
            s._ck(L.pf_download(s._h, runner.owned(runner.F).data_ptr(), hB[0].data_ptr(), hB[0].numel(), dt_code))
            s._ck(L.pf_download(s._h, runner.owned(runner.U).data_ptr(), hB[1].data_ptr(), hB[1].numel(), dt_code))
        hA, hB = hB, hA
    torch.cuda.synchronize()
    secs = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([secs], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs = float(t.item())
    return secs, nbytes, nbytes


def run_cuda_arm(args):
    import torch
    import torch.distributed as dist
    from bachelors_b200 import solver as S
    from bachelors_b200.distributed import SlabRunner
    from bachelors_b200.slabs import slab_rows

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.gpus > 1 and world == 1:
        raise SystemExit("launch N > 1 with torch.distributed.run (one rank per GPU)")
    S.lib()                                             # fails loudly if the CUDA library is missing
    if not torch.cuda.is_available():
        raise SystemExit("no CUDA device: this benchmark has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dtype = PF_F64 if args.dtype == "f64" else PF_F32
    npdt = np.float64 if dtype == PF_F64 else np.float32
    cfg, desc = workload(args, world)
    p = cfg.params
    y0, y1 = slab_rows(p.ny, world, rank)
    F0, U0 = make_initial_conditions(cfg, npdt, rows=(y0, y1))
    if args.field != "seed":
        # the dendrite seed has an exactly flat far field; throughput is also reported on a smooth and a seeded-noise state
        from bachelors_b200.testing import seeded_fields
        gF, gU = seeded_fields(p.nx, p.ny, npdt, kind=args.field)
        F0, U0 = np.ascontiguousarray(gF[y0:y1]), np.ascontiguousarray(gU[y0:y1])
        desc += f", field = {args.field}"

    runner = SlabRunner(p, dtype, rank=rank, world=world, device=local, overlap=not args.no_overlap, p2p=not args.no_p2p,
                        halo_in_kernel=None if args.halo_in_kernel < 0 else bool(args.halo_in_kernel))
    runner.upload_global(F0, U0)
    launches0 = runner.solver.kernel_launches
    for _ in range(args.warmup):
        runner.step()
    runner.sync()

    sampler = ClockSampler(physical_gpu_index(local))
    sampler.start()
    launches1 = runner.solver.kernel_launches
    ms, t_adv, attempts = time_device_steps(runner, torch, args.steps, world, dist)
    launches = runner.solver.kernel_launches - launches1
    clocks = sampler.finish()

    cells = p.nx * p.ny
    glups = cells * args.steps / (ms * 1e-3) / 1e9
    rs = 8 if dtype == PF_F64 else 4
    peak, peak_src = load_peaks()
    # dominant kernel: the streaming step kernel (pf_stream.cuh), one launch per step (adaptive: one per attempt) producing this rank's rows
    launches_per_step = launches / args.steps
    alg_bytes_per_launch = 4 * rs * (cells / world)                    # read Phi, T + write Phi', T' 
    avg_launch_s = (ms * 1e-3) / max(attempts if args.solver == "explicit-rk4-adaptive" else args.steps, 1)
    achieved = alg_bytes_per_launch / avg_launch_s / 1e9
    kernel_name, alg_per_cell = "k_stream_step", 4 * rs
    if args.solver == "semi-implicit":
        # 4 s for prepare / b_T plus >= 10 s per CG iteration (SpMV 2-3, x/r update 6, p update 3 field passes);
        # the "launch" of the roofline is the whole step (prepare + both solves)
        cg_per_step = runner.cg_iters_total / max(args.steps, 1)
        alg_per_cell = (4 + 10 * cg_per_step) * rs
        achieved = alg_per_cell * (cells / world) / (ms * 1e-3 / args.steps) / 1e9       # per GPU, against one GPU's peak
        kernel_name = ("k_si_step_coop (whole step in one cooperative launch)" if world == 1
                       else "k_si_prepare + k_cg_spmv/update/pupdate (whole step, phase API)")
    traffic = load_traffic(args.solver, args.dtype, p.nx) if world == 1 and not args.wide and args.field == "seed" else None

    e2e_steps = max(3, min(args.steps, args.e2e_steps))
    e2e_secs, h2d, d2h = time_e2e(runner, torch, e2e_steps, world, dist)
    e2e_glups = cells * e2e_steps / e2e_secs / 1e9

    tot_launches = launches
    if world > 1:
        t = torch.tensor([launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(t)
        tot_launches = int(t.item())

    line = {
        "metric": "GLUPS", "value": glups, "unit": "1e9 cell updates/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": args.dtype, "data": ("synthetic (config's dendrite seed; no dataset)" if args.field == "seed" else f"synthetic ({args.field} field of SURVEY 8d; no dataset)"),
        "config": config_dict(args, world, p, desc, rs),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "kernel": kernel_name, "peak_source": peak_src,
                     "algorithmic_bytes_per_cell_step": alg_per_cell, "launches_per_step": launches_per_step},
        "e2e": {"value": e2e_glups, "unit": "1e9 cell updates/s", "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world,
                "steps": e2e_steps, "what": "pf_step_host every step: y(n) from pinned host memory in, y(n+1) to pinned host memory out (upload + step + download; banded and overlapped for explicit / explicit-rk4 at N = 1)"},
        "gpu_launches": tot_launches,
        "clocks": clocks,
        "sim_time_advanced": t_adv,
    }
    sec = secondary_roofline(args, p, world, avg_launch_s, clocks, runner.solver)
    if sec:
        line["roofline_secondary"] = sec
    if args.solver == "explicit-rk4-adaptive":
        line["attempts_per_step"] = attempts / args.steps
    if args.solver == "semi-implicit":
        line["cg_iterations_per_step"] = runner.cg_iters_total / args.steps

    if world > 1 and not args.no_parity:
        line["parity_check"] = parity_check(args, cfg, dtype, runner, F0, U0, torch, dist, rank, world, local)
    if world > 1 and not args.no_weak_base:
        base = weak_base(args, dtype, torch, dist, rank, local)
        if rank == 0:
            line["weak_base"] = base
            line["weak_efficiency"] = glups / (world * base["value"])
    if rank == 0 and world == 1 and not args.no_cpu:
        v, cores, sample = cpu_port_rate(cfg, dtype, budget_s=args.cpu_budget, max_steps=400, w=None if not args.wide else 2048)
        line["cpu_baseline"] = {"value": v, "unit": "1e9 cell updates/s", "cores": cores, "kind": "port", "sample": sample}
    if rank == 0 and world == 1 and not args.no_ref_gpu:
        ref = same_box_reference(cfg, dtype, args)
        if ref:
            line["reference_gpu_same_box"] = ref
    runner.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)
    return 0


def parity_check(args, cfg, dtype, runner, F0, U0, torch, dist, rank, world, local, steps=3):
    """N > 1, after the timed region: the slabs and ONE GPU holding the whole grid each take `steps` steps from the same initial
    state; every rank's owned rows must equal the single-grid rows bit for bit (compared on rank 0 as 64-bit integers, and by
    SHA-256 of the 2 x ghost rows next to every slab boundary - the rows the halo exchange feeds)."""
    import hashlib
    from bachelors_b200 import solver as S
    from bachelors_b200.distributed import SlabRunner
    from bachelors_b200.slabs import slab_rows
    p = cfg.params
    fresh = SlabRunner(p, dtype, rank=rank, world=world, device=local, overlap=not args.no_overlap, p2p=not args.no_p2p)
    fresh.upload_global(F0, U0)
    dev = torch.device("cuda", local)
    tdt = torch.float64 if dtype == PF_F64 else torch.float32
    full = None
    if rank == 0:
        full = [torch.empty((p.ny, p.nx), dtype=tdt, device=dev) for _ in range(4)]
    # the initial state of the whole grid, gathered on rank 0 from the slabs' owned rows
    for f_i, t in enumerate((fresh.F, fresh.U)):
        own = fresh.owned(t).contiguous()
        if rank == 0:
            y0, y1 = slab_rows(p.ny, world, 0)
            full[f_i][y0:y1].copy_(own)
            for r in range(1, world):
                y0, y1 = slab_rows(p.ny, world, r)
                dist.recv(full[f_i][y0:y1], src=r)
        else:
            dist.send(own, dst=0)
    # the sends read the input fields on NCCL's stream; the solver's streams do not wait for it, and step 2 overwrites those fields
    torch.cuda.synchronize()
    dist.barrier()
    for _ in range(steps):
        fresh.step()
    fresh.sync()
    ok, digests, worst_rel, mismatches = True, [], 0.0, []
    if rank == 0:
        with S.Solver(p.nx, p.ny, dtype, device=local) as one:
            q = p.copy()
            F, U, nF, nU = (t.data_ptr() for t in full)
            cur = [0, 1, 2, 3]
            for i in range(steps):
                q.iter, q.time = i, i * p.dt
                one.step(q, full[cur[0]].data_ptr(), full[cur[1]].data_ptr(), full[cur[2]].data_ptr(), full[cur[3]].data_ptr())
                cur = [cur[2], cur[3], cur[0], cur[1]]
            one.sync()
        ref_F, ref_U = full[cur[0]], full[cur[1]]
    itype = torch.int64 if dtype == PF_F64 else torch.int32
    g = max(runner.ghost, 1) * 2
    for f_i, t in enumerate((fresh.F, fresh.U)):
        own = fresh.owned(t).contiguous()
        if rank == 0:
            ref = ref_F if f_i == 0 else ref_U
            buf = torch.empty_like(own)
            for r in range(world):
                y0, y1 = slab_rows(p.ny, world, r)
                got = own if r == 0 else buf[:y1 - y0]
                if r:
                    dist.recv(got, src=r)
                want = ref[y0:y1]
                if args.solver == "semi-implicit":
                    # the CG dot products are summed in another order on slabs: parity by north_star's tolerance, not by bits
                    rel = float(torch.linalg.vector_norm((got - want).double()) / torch.linalg.vector_norm(want.double()).clamp_min(1e-300))
                    worst_rel = max(worst_rel, rel)
                    continue
                same = bool(torch.equal(got.view(itype), want.view(itype)))
                if not same:
                    bad_rows = torch.nonzero((got.view(itype) != want.view(itype)).any(dim=1)).flatten()
                    mismatches.append({"rank": r, "field": "FU"[f_i], "rows": int(bad_rows.numel()), "first": int(bad_rows[0]) + y0, "last": int(bad_rows[-1]) + y0})
                ok = ok and same
                for a, b in ((y0, min(y0 + g, y1)), (max(y1 - g, y0), y1)):
                    hg = hashlib.sha256(got[a - y0:b - y0].cpu().numpy().tobytes()).hexdigest()
                    hw = hashlib.sha256(want[a - y0:b - y0].cpu().numpy().tobytes()).hexdigest()
                    ok = ok and hg == hw
                    digests.append(hg[:16])
        else:
            dist.send(own, dst=0)
    torch.cuda.synchronize()
    fresh.close()
    del full
    torch.cuda.empty_cache()
    dist.barrier()
    if rank != 0:
        return None
    if args.solver == "semi-implicit":
        tol = 1e-10 if dtype == PF_F64 else 1e-5
        return {"steps": steps, "bit_identical": False, "rel_l2": worst_rel, "within_tolerance": worst_rel <= tol, "tolerance": tol, "ranks": world,
                "what": f"{steps} steps on {world} slabs vs the whole grid on rank 0's GPU: worst relative L2 difference over fields and slabs "
                        "(the CG dot products are summed in another order on slabs)"}
    return {"steps": steps, "bit_identical": bool(ok), "ranks": world,
            "what": f"{steps} steps of the {p.nx}x{p.ny} grid on {world} slabs vs the whole grid on rank 0's GPU: all owned rows compared as "
                    f"integers on the device; SHA-256 of the {g} rows at both ends of every slab (first 16 hex digits below)",
            "sha256_boundary_rows": digests, "mismatches": mismatches}


def weak_base(args, dtype, torch, dist, rank, local):
    """N > 1: ONE rank's workload (the 16384 x 2048 slab) timed alone on rank 0's GPU in the same run: the base the weak-scaling
    efficiency is measured against (the driver's N = 1 line is the 2048^2 config, a different per-GPU workload)."""
    from bachelors_b200.distributed import SlabRunner
    out = None
    if rank == 0:
        wide = argparse.Namespace(**vars(args)); wide.wide = True
        cfg1, desc1 = workload(wide, 1)
        p1 = cfg1.params
        npdt = np.float64 if dtype == PF_F64 else np.float32
        F1, U1 = make_initial_conditions(cfg1, npdt)
        r1 = SlabRunner(p1, dtype, rank=0, world=1, device=local)
        r1.upload_global(F1, U1)
        for _ in range(args.warmup):
            r1.step()
        r1.sync()
        ms, _, _ = time_device_steps(r1, torch, args.steps, 1, dist)
        out = {"value": p1.nx * p1.ny * args.steps / (ms * 1e-3) / 1e9, "ms_per_step": ms / args.steps, "workload": desc1 + ", rank 0 alone, same run"}
        r1.close()
    dist.barrier()
    return out


def secondary_roofline(args, p, world, avg_launch_s, clocks, solver):
    """The ceilings that bind the multi-stage kernels: FP64 pipe and issue slots. Executed warp instructions
    per launch come from the committed ncu capture of the same kernel and grid (profiles/inst_counts.json); the FP64 peak is the
    rate measured on this pool's B200 with tools/ubench_fp64.cu (profiles/r02_kernel_experiments.md section 4)."""
    if world != 1 or args.wide or args.field != "seed":
        return None
    ic = load_inst_counts(args.solver, args.dtype, p.nx)
    if not ic:
        return None
    mhz = (clocks or {}).get("sm_mhz") or 1965.0
    sms = 148
    clk = avg_launch_s * mhz * 1e6                                         # SM cycles per launch
    fp64 = ic["fp64_warp_inst"] / (clk * sms)                              # warp instructions / clk / SM
    issue = ic["warp_inst"] / (clk * sms * 4)                              # warp instructions / clk / scheduler
    return [
        {"bound": "fp64_pipe", "achieved": fp64, "peak": ic.get("fp64_peak_per_clk_sm", 1.9), "unit": "FP64 warp-instructions/clk/SM",
         "frac": fp64 / ic.get("fp64_peak_per_clk_sm", 1.9), "source": ic.get("source")},
        {"bound": "issue", "achieved": issue, "peak": 1.0, "unit": "warp-instructions/clk/scheduler", "frac": issue, "source": ic.get("source")},
    ]


def same_box_reference(cfg, dtype, args):
    """The UNMODIFIED reference kernels (oracle/_ref, built for sm_100a) timed on the same GPU: 'the kernel to beat'.
    Extra information only; never part of `value`."""
    try:
        from oracle import ref_sim
        if not ref_sim.available(dtype):
            return None
        p = cfg.params
        npdt = np.float64 if dtype == PF_F64 else np.float32
        F0, U0 = make_initial_conditions(cfg, npdt)
        with ref_sim.RefSim(dtype, p.nx, p.ny, fresh=True) as ref:
            ref.upload(F0, U0)
            ref.time_ms(p, 3)
            n = 20
            ms, _ = ref.time_ms(p, n)
        return {"ms_per_step": ms / n, "value": p.nx * p.ny * n / (ms * 1e-3) / 1e9, "unit": "1e9 cell updates/s",
                "what": "reference simulation.cu compiled unmodified for sm_100a, same config, CUDA events around sim_step calls"}
    except Exception as e:                                   # never let the extra leg break the bench line
        return {"error": str(e)[:200]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--solver", default="explicit-rk4", choices=[k for k in SOLVER_NAMES if k not in ("none", "exact")])
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    ap.add_argument("--size", type=int, default=0, help="N = 1: grid size of the benchmark config (128 ... 2048; default 2048, semi-implicit 1024)")
    ap.add_argument("--field", default="seed", choices=["seed", "smooth", "noise"],
                    help="initial state: the config's dendrite seed (default, flat far field), or the smooth / seeded-noise fields of SURVEY 8d")
    ap.add_argument("--wide", action="store_true", help="N = 1 on the 16384 x 2048 slab of the multi-GPU family (weak-scaling base)")
    ap.add_argument("--no-overlap", action="store_true", help="exchange halos only after the interior finished")
    ap.add_argument("--no-p2p", action="store_true", help="N > 1: halo rows through NCCL send/recv instead of the device-initiated peer stores")
    ap.add_argument("--halo-in-kernel", type=int, default=-1, help="N > 1, explicit-rk4: 1 / 0 = halo hand-over inside / around the step kernel (default: the library's)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the 3-step bit comparison against the whole grid on one GPU")
    ap.add_argument("--no-weak-base", action="store_true", help="N > 1: skip timing one rank's slab alone on rank 0")
    ap.add_argument("--no-ref-gpu", action="store_true")
    ap.add_argument("--cpu-budget", type=float, default=12.0)
    ap.add_argument("--e2e-steps", type=int, default=20)
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    # The contract is ONE JSON line on stdout. Libraries write banners to fd 1 (NCCL prints its version line there), so
    # everything but the final line goes to stderr: fd 1 is pointed at fd 2 for the run and restored for the print.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    captured = []
    import builtins
    _print = builtins.print

    def print_line(*a, **k):
        if k.get("file") in (None, sys.stdout) and len(a) == 1 and isinstance(a[0], str) and a[0].startswith("{"):
            captured.append(a[0])
        else:
            _print(*a, **k)
    builtins.print = print_line
    try:
        rc = run_reference_arm(args) if args.impl == "reference" else run_cuda_arm(args)
    finally:
        builtins.print = _print
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    for line in captured:
        print(line, flush=True)
    return rc


if __name__ == "__main__":
    sys.exit(main())
