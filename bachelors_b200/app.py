"""Headless host: the non-interactive branch of the reference's main loop (reference main.cpp:238-284, 495-564)
over the C-ABI, with the rows SURVEY section 8f puts around the stepping path:

  * initial conditions evaluated on the device (`pf_init_fields`; main.cpp:93-136, 151-173);
  * stepping in batches up to the next snapshot (`pf_run`: no host round trip between the steps of a fixed-dt solver);
  * snapshots that overlap the stepping: `pf_snapshot_begin` stages the fields on the device in stream order and copies
    them to pinned host memory on a copy stream while the solver steps on; a writer thread turns each landed slot into
    `maps_NNNN.bin` (two slots: one fills while the other is written);
  * `stats.csv` and `config.ini` in the snapshot folder, as `save_state` writes them (main.cpp:827-903);
  * optionally a PNG of Phi per snapshot through the colour-map kernel (the viewer's shader, gl.h:221-283).

    python -m bachelors_b200.app config.ini [more.ini ...] [--dtype f64|f32] [--png]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 -m bachelors_b200.app config.ini
                                      (N GPUs of one box: y-slabs, one snapshot file written by all ranks; app_slabs.py)

The file formats are the reference's, so the reference's `plot.py` reads the output unchanged.
"""
import argparse
import os
import sys
import threading
import time

from . import solver as S
from ._abi import (PF_F32, PF_F64, SOLVER_NAME_OF, SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE, pf_initial, pf_stats_row)
from .config import read_config


def initial_of(cfg):
    """pf_initial from the [initial] section of a parsed config."""
    ic = pf_initial()
    ic.inside_phi, ic.inside_T = cfg.init_inside_phi, cfg.init_inside_T
    ic.outside_phi, ic.outside_T = cfg.init_outside_phi, cfg.init_outside_T
    ic.circle_center[0], ic.circle_center[1] = cfg.init_circle_center
    ic.circle_radius, ic.circle_fade = cfg.init_circle_radius, cfg.init_circle_fade
    ic.square_from[0], ic.square_from[1] = cfg.init_square_from
    ic.square_to[0], ic.square_to[1] = cfg.init_square_to
    return ic


def save_folder_name(cfg, init_time):
    """make_save_folder (main.cpp:760-780); tm_mon is used as is (0-based), as the reference does."""
    t = time.localtime(init_time)
    return "%s%s%s%04i-%02i-%02i__%02i-%02i-%02i__%s%s" % (
        cfg.snapshot_folder, "/" if cfg.snapshot_folder else "", cfg.snapshot_prefix,
        t.tm_year, t.tm_mon - 1, t.tm_mday, t.tm_hour, t.tm_min, t.tm_sec,
        SOLVER_NAME_OF[cfg.params.solver], cfg.snapshot_postfix)


class SnapshotWriter:
    """Turns landed snapshot slots into files on a worker thread; at most PF_SNAPSHOT_SLOTS in flight."""

    def __init__(self, sol, cfg, folder, png=False):
        self.sol, self.cfg, self.folder, self.png = sol, cfg, folder, png
        self.threads = [None, None]
        self.next_slot = 0
        self.errors = []
        self.written = []

    def submit(self, index, it, sim_time, F, U):
        slot = self.next_slot
        self.next_slot ^= 1
        if self.threads[slot] is not None:
            self.threads[slot].join()                       # the slot's previous file must be on disk before it refills
        rgba = None
        if self.png:                                        # rendered from the live field before the solver steps on
            rgba = self.sol.render_rgba(F, self.cfg.app_display_min, self.cfg.app_display_max)
        self.sol.snapshot_begin(slot, [F, U])
        th = threading.Thread(target=self._write, args=(slot, index, it, sim_time, rgba), daemon=True)
        self.threads[slot] = th
        th.start()

    def _write(self, slot, index, it, sim_time, rgba):
        try:
            p = self.cfg.params
            hF, hU = self.sol.snapshot_wait(slot)
            path = os.path.join(self.folder, "maps_%04i.bin" % index)
            S.write_maps_bin(path, p.L0 / p.nx, p.L0 / p.ny, it, sim_time, {"F": hF, "U": hU})
            self.written.append(path)
            if rgba is not None:
                S.write_png(os.path.join(self.folder, "phi_%04i.png" % index), rgba, flip_y=True)
        except Exception as e:                              # surfaced by finish()
            self.errors.append(e)

    def finish(self):
        for th in self.threads:
            if th is not None:
                th.join()
        if self.errors:
            raise self.errors[0]


def _stats_row(params, info):
    """app_collect_stats (main.cpp:192-234) from the step's pf_step_info."""
    r = pf_stats_row()
    r.time, r.iter = params.time, params.iter
    r.Phi_iters, r.T_iters = info.Phi_iters, info.T_iters
    for name, d in (("T", info.T_delta), ("Phi", info.Phi_delta)):
        setattr(r, name + "_delta_L1", d.L1)
        setattr(r, name + "_delta_L2", d.L2)
        setattr(r, name + "_delta_max", d.max)
        setattr(r, name + "_delta_min", d.min)
    for k in range(info.step_res_count):
        r.step_res_L1[k], r.step_res_L2[k] = info.step_res[k].L1, info.step_res[k].L2
        r.step_res_max[k], r.step_res_min[k] = info.step_res[k].max, info.step_res[k].min
    return r


def run_config(config_path, dtype=PF_F64, device=0, overrides=(), png=False, log=None, init_time=None):
    """Runs one config to `stop_after`; returns a dict (folder, iters, sim_time, runtime, snapshots written)."""
    cfg = read_config(config_path, overrides)
    p = cfg.params
    log = log or (lambda msg: None)
    init_time = time.time() if init_time is None else init_time
    folder = save_folder_name(cfg, init_time)
    os.makedirs(folder, exist_ok=True)

    with S.Solver(p.nx, p.ny, dtype, device=device) as sol:
        F, U, F2, U2 = (sol.field_alloc() for _ in range(4))
        sol.init_fields(p, initial_of(cfg), F, U)
        writer = SnapshotWriter(sol, cfg, folder, png=png)
        state = {"stats_files": 0, "config_written": False, "snapshots": 0}
        stats_rows, step_res_count = [], [0]

        def save_state(index, it, sim_time, with_stats):
            writer.submit(index, it, sim_time, F, U)
            if with_stats:                                                      # SAVE_STATS, main.cpp:865-893
                S.write_stats_csv(os.path.join(folder, "stats.csv"), state["stats_files"] != 0, p.nx, p.ny, p.dt,
                                  step_res_count[0], stats_rows)
                stats_rows.clear()
                state["stats_files"] += 1
            if not state["config_written"]:                                     # SAVE_CONFIG, main.cpp:895-900
                with open(os.path.join(folder, "config.ini"), "w") as f:
                    f.write(cfg.entire_config_file)
                state["config_written"] = True

        if cfg.snapshot_initial_conditions:                                     # main.cpp:283-284
            save_state(0, 0, 0.0, with_stats=False)

        stop = cfg.simul_stop_time
        it, sim_time = 0, 0.0
        snapshot_every_i = snapshot_times_i = 0
        end_reached = False
        collect = bool(cfg.app_collect_stats)
        # per-step residuals of the corrector loop need every step's info; plain delta statistics are wanted only for the step that
        # crosses the next collection time (main.cpp:554-555): pf_run_until hands back there, with that step's statistics
        per_step = collect and bool(cfg.app_collect_step_residuals)
        batched = p.solver != SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE and not per_step
        last_stats_save = 0.0                                                   # app->last_stats_save: set by every collection (main.cpp:233)
        t0 = time.perf_counter()

        def save_due(t):
            """The three tests at the top of the reference's loop (main.cpp:507-530), without side effects."""
            next_every = (snapshot_every_i + 1) * cfg.snapshot_every
            next_times = (snapshot_times_i + 1) * stop / cfg.snapshot_times if cfg.snapshot_times else float("inf")
            return t >= next_every or t >= next_times or stop - t < 1e-16

        while True:
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

            params = p.copy()
            params.iter, params.time = it, sim_time
            params.do_stats = int(collect)
            params.do_stats_step_residual = int(cfg.app_collect_step_residuals)
            if batched:
                # steps until the loop head would save again: the same additions the stepping performs
                n, t = 0, sim_time
                while True:
                    t += p.dt
                    n += 1
                    if save_due(t) or (collect and t >= last_stats_save + cfg.app_collect_stats_every):
                        break
                four = (F, U, F2, U2)
                F, U, done, _adv, info = sol.run(params, F, U, F2, U2, n)
                F2, U2 = [f for f in four if f not in (F, U)]           # the other two allocations are the scratch pair now
                assert done == n
                it += n
                sim_time = t
                if collect and sim_time >= last_stats_save + cfg.app_collect_stats_every:      # main.cpp:554-555
                    q = params.copy(); q.iter, q.time = it - 1, sol.last_run_last_step_time
                    stats_rows.append(_stats_row(q, info))
                    last_stats_save = sim_time
            elif p.solver == SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE and not per_step:
                # the step sizes are the solver's: pf_run_until steps until one of the three tests at the loop head would fire
                # (its t_end is this loop's sim_time: the same additions in the same order)
                next_every = (snapshot_every_i + 1) * cfg.snapshot_every
                next_times = (snapshot_times_i + 1) * stop / cfg.snapshot_times if cfg.snapshot_times else float("inf")
                pause = min(next_every, next_times)
                if collect:
                    pause = min(pause, last_stats_save + cfg.app_collect_stats_every)
                four = (F, U, F2, U2)
                F, U, done, _adv, info = sol.run(params, F, U, F2, U2, (1 << 40) if pause > sim_time else 1, stop_time=stop,
                                              pause_time=pause if pause != float("inf") and pause > 0 else 0.0)
                F2, U2 = [f for f in four if f not in (F, U)]
                it += done
                sim_time = sol.last_run_end_time
                if collect and sim_time >= last_stats_save + cfg.app_collect_stats_every:      # main.cpp:554-555
                    q = params.copy(); q.iter, q.time = it - 1, sol.last_run_last_step_time
                    stats_rows.append(_stats_row(q, info))
                    last_stats_save = sim_time
            else:
                info = sol.step(params, F, U, F2, U2)
                sim_time += info.dt_advanced
                if collect and sim_time >= last_stats_save + cfg.app_collect_stats_every:      # main.cpp:554-555
                    step_res_count[0] = max(step_res_count[0], info.step_res_count)
                    stats_rows.append(_stats_row(params, info))
                    last_stats_save = sim_time                                             # main.cpp:233
                F, U, F2, U2 = F2, U2, F, U
                it += 1
        sol.sync()
        writer.finish()
        runtime = time.perf_counter() - t0
        log("Finished! runtime: %.2fs | iters: %i | average step time: %.4f ms" % (runtime, it, runtime / max(it, 1) * 1000))
        return {"folder": folder, "iters": it, "sim_time": sim_time, "runtime_s": runtime, "snapshots": state["snapshots"],
                "files": sorted(writer.written)}


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("configs", nargs="*", default=["config.ini"])
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--png", action="store_true", help="also write phi_NNNN.png per snapshot (colour-map kernel)")
    ap.add_argument("--set", action="append", default=[], metavar="'[section]\\nkey = value'", help="config override")
    a = ap.parse_args(argv)
    rc = 0
    world = int(os.environ.get("WORLD_SIZE", "1")) if "RANK" in os.environ else 0
    if world > 0:                                               # launched by torchrun: one process per GPU, y-slabs
        import torch
        import torch.distributed as dist
        from .app_slabs import run_config_slabs
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank = int(os.environ.get("RANK", "0"))
    for path in a.configs:
        try:
            dtype = PF_F64 if a.dtype == "f64" else PF_F32
            log = lambda m: print("[app] " + m, flush=True)
            if world > 0:
                out = run_config_slabs(path, dtype, overrides=a.set, log=log)
            else:
                out = run_config(path, dtype, a.device, overrides=a.set, png=a.png, log=log)
            if rank == 0:
                print("[app] %s -> %s (%i iters, %i snapshots)" % (path, out["folder"], out["iters"], out["snapshots"]))
        except Exception as e:                                  # main.cpp:253-258: log, skip to the next config
            print("[app] failed to run config '%s': %s" % (path, e), file=sys.stderr)
            rc = 1
    if world > 0:
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
