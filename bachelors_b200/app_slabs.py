"""The headless host on several GPUs (SURVEY section 8e + 8f): the reference's non-interactive loop (main.cpp:495-564) with the
grid cut into y-slabs, one process per GPU (launched by torchrun; bachelors_b200/distributed.py does the stepping).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 -m bachelors_b200.app \
        benchmark_configs/config_explicit-rk4-adaptive_16384.ini

  * initial conditions: every rank evaluates its own rows (ghost rows included) on its device - no scatter from a host;
  * stepping: SlabRunner.step() in lock step (one halo exchange per step; the adaptive solver's error norm and the
    semi-implicit solver's CG sums are the only reductions);
  * snapshots: every rank stages its owned rows on its device and copies them to pinned memory on a copy stream, rank 0
    writes the header of `maps_NNNN.bin`, then every rank writes its rows into the same file at their offset
    (`pf_write_maps_bin_rows`) from a worker thread while the solvers step on. The file is byte for byte the one the
    single-GPU host (and the reference, main.cpp:705-735) writes: tests/test_gpu_multi.py compares them.

All ranks see the same simulation time (fixed dt, or the all-reduced accepted step), so they take the same branches of the
loop without talking about it.
"""
import os
import threading
import time

import torch
import torch.distributed as dist

from . import solver as S
from ._abi import PF_F64, pf_stats_row
from .app import initial_of, save_folder_name
from .config import read_config
from .distributed import SlabRunner


class SlabSnapshotWriter:
    """This rank's rows of every snapshot; two slots in flight, as the single-GPU writer."""

    def __init__(self, runner, cfg, folder):
        self.runner, self.cfg, self.folder = runner, cfg, folder
        self.threads = [None, None]
        self.next_slot = 0
        self.errors = []
        self.written = []

    def submit(self, index, it, sim_time):
        r, p = self.runner, self.cfg.params
        slot = self.next_slot
        self.next_slot ^= 1
        if self.threads[slot] is not None:
            self.threads[slot].join()
        r.sync()                                            # the step that produced F, U ran on three streams
        r.solver.snapshot_begin(slot, [r.F.data_ptr(), r.U.data_ptr()])
        path = os.path.join(self.folder, "maps_%04i.bin" % index)
        if r.rank == 0:
            S.write_maps_bin_header(path, p.nx, p.ny, p.L0 / p.nx, p.L0 / p.ny, it, sim_time, ["F", "U"])
        dist.barrier(group=r.group)                         # the header (and with it the file's size) exists before any rows land
        th = threading.Thread(target=self._write, args=(slot, path), daemon=True)
        self.threads[slot] = th
        th.start()

    def _write(self, slot, path):
        try:
            r, p = self.runner, self.cfg.params
            hF, hU = r.solver.snapshot_wait(slot)
            S.write_maps_bin_rows(path, p.nx, p.ny, 2, 0, r.layout.y0, hF)
            S.write_maps_bin_rows(path, p.nx, p.ny, 2, 1, r.layout.y0, hU)
            self.written.append(path)
        except Exception as e:
            self.errors.append(e)

    def finish(self):
        for th in self.threads:
            if th is not None:
                th.join()
        if self.errors:
            raise self.errors[0]


def run_config_slabs(config_path, dtype=PF_F64, overrides=(), log=None, init_time=None, group=None):
    """Runs one config to `stop_after` on all ranks of `group`; every rank returns the same dict as app.run_config."""
    cfg = read_config(config_path, overrides)
    p = cfg.params
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    log = log if (log and rank == 0) else (lambda msg: None)
    box = [time.time() if init_time is None else init_time]
    dist.broadcast_object_list(box, src=0, group=group)     # one folder name for all ranks
    folder = save_folder_name(cfg, box[0])
    if rank == 0:
        os.makedirs(folder, exist_ok=True)
    dist.barrier(group=group)

    runner = SlabRunner(p, dtype, group=group)
    try:
        sol = runner.solver
        sol.init_fields(p, initial_of(cfg), runner.F.data_ptr(), runner.U.data_ptr())     # owned rows and ghost rows
        writer = SlabSnapshotWriter(runner, cfg, folder)
        state = {"stats_files": 0, "config_written": False, "snapshots": 0}
        stats_rows = []

        def save_state(index, it, sim_time, with_stats):
            writer.submit(index, it, sim_time)
            if rank != 0:
                return
            if with_stats:                                                      # SAVE_STATS, main.cpp:865-893
                S.write_stats_csv(os.path.join(folder, "stats.csv"), state["stats_files"] != 0, p.nx, p.ny, p.dt, 0, stats_rows)
                stats_rows.clear()
                state["stats_files"] += 1
            if not state["config_written"]:                                     # SAVE_CONFIG, main.cpp:895-900
                with open(os.path.join(folder, "config.ini"), "w") as f:
                    f.write(cfg.entire_config_file)
                state["config_written"] = True

        if cfg.snapshot_initial_conditions:
            save_state(0, 0, 0.0, with_stats=False)

        stop = cfg.simul_stop_time
        it, sim_time = 0, 0.0
        snapshot_every_i = snapshot_times_i = 0
        end_reached = False
        collect = bool(cfg.app_collect_stats)
        last_stats_save = 0.0
        torch.cuda.synchronize()
        dist.barrier(group=group)
        t0 = time.perf_counter()
        while True:                                                             # main.cpp:507-559
            save_this_iter = False
            if sim_time >= (snapshot_every_i + 1) * cfg.snapshot_every:
                snapshot_every_i += 1
                save_this_iter = True
            if cfg.snapshot_times and sim_time >= (snapshot_times_i + 1) * stop / cfg.snapshot_times and not end_reached:
                snapshot_times_i += 1
                save_this_iter = True
            if stop - sim_time < 1e-16 and not end_reached:
                end_reached = True
                if cfg.snapshot_times > 0:
                    save_this_iter = True
            if save_this_iter:
                state["snapshots"] += 1
                log("saving snapshot %i (iter %i, t = %g)" % (state["snapshots"], it, sim_time))
                save_state(state["snapshots"], it, sim_time, with_stats=True)
            if end_reached:
                break
            runner.time, runner.iter = sim_time, it
            step_time = sim_time                                                # params.time of the step: what the reference's stats record
            sim_time += runner.step()
            if collect and sim_time >= last_stats_save + cfg.app_collect_stats_every:     # main.cpp:554-555 (every rank keeps the same clock)
                last_stats_save = sim_time                                                 # main.cpp:233
                if rank == 0:
                    # iteration counts only: the delta norms of the reference's stats (simulation.cu:1126-1142) are per-grid sums
                    # this host does not reduce across ranks
                    row = pf_stats_row()
                    row.time, row.iter = step_time, it
                    cg = getattr(runner, "last_cg_iters", None)
                    row.Phi_iters, row.T_iters = cg if cg else (runner.last_attempts, runner.last_attempts)
                    stats_rows.append(row)
            it += 1
        runner.sync()
        writer.finish()
        dist.barrier(group=group)
        runtime = time.perf_counter() - t0
        log("Finished! runtime: %.2fs | iters: %i | average step time: %.4f ms" % (runtime, it, runtime / max(it, 1) * 1000))
        files = [os.path.join(folder, "maps_%04i.bin" % i) for i in range(0 if cfg.snapshot_initial_conditions else 1, state["snapshots"] + 1)]
        return {"folder": folder, "iters": it, "sim_time": sim_time, "runtime_s": runtime, "snapshots": state["snapshots"],
                "files": files, "world": world}
    finally:
        runner.close()
