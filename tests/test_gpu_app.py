"""GPU tests of the rows around the stepping path: device-side initial conditions, overlapped
snapshots, the headless host loop with the reference's output files, the colour-map kernel."""
import os

import numpy as np
import pytest

from bachelors_b200 import solver as S
from bachelors_b200._abi import PF_F32, PF_F64, pf_slab
from bachelors_b200.app import initial_of, run_config
from bachelors_b200.config import read_config
from bachelors_b200.initial import make_initial_conditions
from bachelors_b200.testing import seeded_fields
from conftest import ROOT
from oracle.colormap import shade
from oracle.oracle import Oracle
from test_output_cpu import load_bin_map_file

pytestmark = pytest.mark.gpu

NP = {PF_F64: np.float64, PF_F32: np.float32}


def cfg_of(name, *overrides):
    return read_config(os.path.join(ROOT, "benchmark_configs", name), overrides)


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
@pytest.mark.parametrize("name", ["config_explicit_128.ini", "config_explicit-rk4_512.ini", "config_semi-implicit_1024.ini"])
def test_device_initial_conditions_equal_the_host_loop(name, dtype):
    """Benchmark family (sharp circle): bit-identical to the host restatement of main.cpp:93-136."""
    cfg = cfg_of(name)
    p = cfg.params
    hF, hU = make_initial_conditions(cfg, NP[dtype])
    with S.Solver(p.nx, p.ny, dtype) as s:
        F, U = s.field_alloc(), s.field_alloc()
        s.init_fields(p, initial_of(cfg), F, U)
        assert np.array_equal(s.download(F), hF) and np.array_equal(s.download(U), hU)
    assert hF.sum() > 0                                              # the seed is there


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
def test_device_initial_conditions_faded_circle_and_square(dtype):
    """A faded rim and a square: the rim goes through hypot, where libm and the device may differ in the last bit."""
    cfg = cfg_of("config_explicit_128.ini", "[initial]\ncircle_fade = 12\ncircle_radius = 0.2\nsquare_from = 0.1 0.15\nsquare_to = 0.3 0.4\n"
                 "inside_T = 0.5\noutside_T = -0.25\ninside_phi = 0.9\noutside_phi = 0.1")
    p = cfg.params
    hF, hU = make_initial_conditions(cfg, NP[dtype])
    with S.Solver(p.nx, p.ny, dtype) as s:
        F, U = s.field_alloc(), s.field_alloc()
        s.init_fields(p, initial_of(cfg), F, U)
        dF, dU = s.download(F), s.download(U)
    tol = 1e-14 if dtype == PF_F64 else 1e-6
    assert np.abs(dF - hF).max() <= tol and np.abs(dU - hU).max() <= tol
    assert (dF != hF).mean() < 0.01                                  # all but a few rim cells are bit-identical
    rim = (hF > 0.1001) & (hF < 0.8999)
    assert rim.sum() > 50                                            # the fade really produced intermediate values


def test_device_initial_conditions_on_a_slab_fill_ghost_rows():
    cfg = cfg_of("config_explicit-rk4_128.ini", "[simulation]\nPhi_boundary = periodic\nT_boundary = periodic\n[initial]\ncircle_center = 0.5 0.02\ncircle_radius = 0.2")
    p = cfg.params
    full = make_initial_conditions(cfg, np.float64)[0]
    slab = pf_slab(y0=0, ny_local=40, has_lo=1, has_hi=1, ghost=4)     # first slab of a periodic ring: its lo ghosts wrap
    with S.Solver(p.nx, p.ny, PF_F64, slab=slab) as s:
        F, U = s.field_alloc(), s.field_alloc()
        s.init_fields(p, initial_of(cfg), F, U)
        got = s.download(F)
    rows = [(y % p.ny) for y in range(-4, 44)]
    assert got.shape == (48, p.nx) and np.array_equal(got, full[rows])


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
def test_snapshot_is_taken_in_stream_order_and_overlaps_stepping(dtype):
    nx, ny = 192, 160
    cfg = cfg_of("config_explicit-rk4_128.ini")
    p = cfg.params.copy(); p.nx, p.ny, p.L0 = nx, ny, nx / 128.0
    F0, U0 = seeded_fields(nx, ny, NP[dtype], kind="interface")
    with S.Solver(nx, ny, dtype) as s:
        F, U, F2, U2 = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        s.step_async(p, F, U, F2, U2)                                # y1 in (F2, U2), not yet synchronised
        s.snapshot_begin(0, [F2, U2])
        for _ in range(6):                                           # keeps overwriting both pairs while the copy is in flight
            s.step_async(p, F2, U2, F, U)
            s.step_async(p, F, U, F2, U2)
        s.snapshot_begin(1, [F2, U2])
        a = [x.copy() for x in s.snapshot_wait(0)]
        b = [x.copy() for x in s.snapshot_wait(1)]
        s.sync()
        end = s.download(F2), s.download(U2)
    with S.Solver(nx, ny, dtype) as s:
        F, U, F2, U2 = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        s.step(p, F, U, F2, U2)
        y1 = s.download(F2), s.download(U2)
    assert a[0].dtype == np.float64 and a[0].shape == (ny, nx)
    assert np.array_equal(a[0], y1[0].astype(np.float64)) and np.array_equal(a[1], y1[1].astype(np.float64))
    assert np.array_equal(b[0], end[0].astype(np.float64)) and np.array_equal(b[1], end[1].astype(np.float64))
    assert not np.array_equal(a[0], b[0])


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
def test_colour_map_kernel_matches_the_shader_restatement(dtype):
    nx, ny = 256, 64
    rng = np.random.default_rng(11)
    v = rng.uniform(-0.5, 1.5, (ny, nx))
    v[0, :8] = [np.nan, np.inf, -np.inf, 0.0, 1.0, 0.25, 0.5, 0.75]
    v[1] = np.linspace(0, 1, nx)
    v[2] = np.linspace(-30, 30, nx)
    v = v.astype(NP[dtype])
    with S.Solver(nx, ny, dtype) as s:
        f = s.field_from(v)
        got = s.render_rgba(f, 0.0, 1.0)
        got_other_range = s.render_rgba(f, -0.25, 0.5)
        flat = s.render_rgba(f, 0.5, 0.5)
    for g, (lo, hi) in ((got, (0.0, 1.0)), (got_other_range, (-0.25, 0.5)), (flat, (0.5, 0.5))):
        want = shade(v, lo, hi)
        inside = ~np.isnan(v) & (v >= NP[dtype](lo)) & (v <= NP[dtype](hi))
        # inside the range: plain float arithmetic, bit-exact; the grey ramps go through atan (device vs libm: <= 1 level)
        assert np.array_equal(g[inside], want[inside])
        assert np.abs(g.astype(int) - want.astype(int)).max() <= 1
        assert (g != want).mean() < 0.01
    assert tuple(got[0, 0]) == (255, 0, 255, 255)


def oracle_steps(cfg, dtype, steps):
    p = cfg.params
    F, U = make_initial_conditions(cfg, NP[dtype])
    o = Oracle(dtype, p.nx, p.ny)
    q = p.copy()
    t = 0.0
    for i in range(steps):
        q.iter, q.time = i, t
        F, U, dt, _ = o.step(q, F, U)
        t += dt
    return F, U


@pytest.mark.parametrize("solver_name,dtype", [("explicit-rk4", PF_F64), ("explicit", PF_F32), ("explicit-rk4-adaptive", PF_F64),
                                               ("semi-implicit", PF_F64)])
def test_headless_run_writes_the_reference_files(tmp_path, solver_name, dtype):
    """The reference's non-interactive loop (main.cpp:495-564) to t = 2e-4 at 128^2: snapshot count and iteration
    stamps as the reference's loop gives them, files readable by plot.py's loader, fields equal to the oracle's."""
    over = "[simulation]\nstop_after = 0.0002\n[snapshot]\ntimes = 4\nfolder = %s\n[program]\ncollect_stats = %s" % (
        tmp_path, "true" if solver_name == "semi-implicit" else "false")
    path = os.path.join(ROOT, "benchmark_configs", "config_%s_128.ini" % solver_name)
    out = run_config(path, dtype, overrides=[over], png=True)
    cfg = read_config(path, [over])
    names = sorted(os.listdir(out["folder"]))
    n_snap = out["snapshots"]
    assert [n for n in names if n.startswith("maps_")] == ["maps_%04d.bin" % i for i in range(n_snap + 1)]      # 0 = initial conditions
    assert "config.ini" in names and "stats.csv" in names and "phi_0000.png" in names
    assert open(os.path.join(out["folder"], "config.ini")).read() == open(path).read()
    first = load_bin_map_file(os.path.join(out["folder"], "maps_0000.bin"))
    iF, iU = make_initial_conditions(cfg, NP[dtype])
    assert first["iter"] == 0 and first["time"] == 0.0 and first["dx"] == cfg.params.L0 / cfg.params.nx
    assert np.array_equal(first["maps"]["F"], iF.astype(np.float64)) and np.array_equal(first["maps"]["U"], iU.astype(np.float64))
    last = load_bin_map_file(os.path.join(out["folder"], "maps_%04d.bin" % n_snap))
    assert last["iter"] == out["iters"] and abs(last["time"] - 0.0002) < 2e-5
    if solver_name != "explicit-rk4-adaptive":
        assert out["iters"] == 40 and n_snap == 4
        mid = load_bin_map_file(os.path.join(out["folder"], "maps_0002.bin"))
        assert mid["iter"] == 20
    oF, oU = oracle_steps(cfg, dtype, out["iters"])
    if solver_name == "semi-implicit":
        rel = np.linalg.norm(last["maps"]["F"] - oF) / np.linalg.norm(oF)
        assert rel <= 1e-10
        lines = open(os.path.join(out["folder"], "stats.csv")).read().strip().split("\n")
        assert lines[0].startswith("128,128,") and len(lines) == 2 + 40          # one row per step (collect_stats_every = 0)
        assert [int(l.split(",")[1]) for l in lines[2:]] == list(range(40))
        assert all(int(l.split(",")[2]) >= 1 for l in lines[2:])                 # CG iterations of the Phi solve
    else:
        assert np.array_equal(last["maps"]["F"], oF.astype(np.float64)) and np.array_equal(last["maps"]["U"], oU.astype(np.float64))


def test_collect_stats_every_throttles_rows_like_the_reference(tmp_path):
    """app_collect_stats sets last_stats_save = sim_time (main.cpp:233), so `collect_stats_every` > 0 leaves one row per
    interval, not one per step; every row carries the time and iteration of the step's params (main.cpp:543-544)."""
    every = 2.5e-5                                                               # 5 steps of dt = 5e-6
    over = "[simulation]\nstop_after = 0.0002\n[snapshot]\ntimes = 4\nfolder = %s\n[program]\ncollect_stats = true\ncollect_stats_every = %g" % (tmp_path, every)
    path = os.path.join(ROOT, "benchmark_configs", "config_explicit_128.ini")
    out = run_config(path, PF_F64, overrides=[over])
    cfg = read_config(path, [over])
    # replay of the reference loop's test on the same additions of dt (main.cpp:553-555)
    want, t, last = [], 0.0, 0.0
    for it in range(out["iters"]):
        t_step = t
        t += cfg.params.dt
        if t >= last + every:
            want.append(it)
            last = t
        del t_step
    lines = open(os.path.join(out["folder"], "stats.csv")).read().strip().split("\n")
    assert [int(l.split(",")[1]) for l in lines[2:]] == want and 0 < len(want) < out["iters"]


def test_full_size_adaptive_run_takes_the_published_number_of_steps(tmp_path):
    """the reference's adaptive solver needs 9522 steps (A100; 9523 on the other GPU) to t = 0.04 at
    2048^2 with tolerances 5e-9 (benchmark_ellapsed_times.txt:112,199). The whole run, 11 snapshots included."""
    over = "[snapshot]\nfolder = %s" % tmp_path
    out = run_config(os.path.join(ROOT, "benchmark_configs", "config_explicit-rk4-adaptive_2048.ini"), PF_F64, overrides=[over])
    assert abs(out["iters"] - 9522) <= 2
    assert out["snapshots"] == 10 and len(out["files"]) == 11
    last = load_bin_map_file(out["files"][-1])
    assert last["nx"] == 2048 and last["iter"] == out["iters"] and last["time"] >= 0.04 - 1e-16
    F = last["maps"]["F"]
    assert np.isfinite(F).all() and F.min() > -0.1 and F.max() < 1.1
    # the seed is centred: the dendrite stays mirror-symmetric to rounding ((W - 2C) + E is not bitwise symmetric in E, W)
    assert np.allclose(F, F[::-1], rtol=1e-6, atol=1e-9) and np.allclose(F, F[:, ::-1], rtol=1e-6, atol=1e-9)
    assert F.sum() > load_bin_map_file(out["files"][0])["maps"]["F"].sum() * 4   # and it grew
