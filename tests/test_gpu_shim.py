"""The drop-in boundary at run time: bachelors_b200/host/simulation_shim.cpp (the file that replaces the reference's
simulation.cu in its build) compiled against the reference's OWN simulation.h / log.h and driven through the reference's
entry points - sim_realloc, sim_modify_float/_double, sim_step with Sim_Map / Sim_Params / Sim_Stats by value
(reference simulation.h:46-145) - by the same glue that drives the unmodified reference solver (oracle/ref_glue).

oracle/build_ref.sh builds it into oracle/_ref/libshim_sim_{f64,f32}.so where /root/reference exists; the prebuilt
libraries travel to the GPU box. What is checked: the reference-shaped call gives bit for bit what the C-ABI gives, the
struct layouts of the reference's header really are what the shim reads, stats / iteration counts / console lines
arrive where the reference's host expects them, and - side by side - the unmodified reference's own results."""
import numpy as np
import pytest

from bachelors_b200 import solver as S
from bachelors_b200._abi import PF_F32, PF_F64, default_params
from bachelors_b200.testing import seeded_fields
from oracle import ref_sim

pytestmark = pytest.mark.gpu
NP = {PF_F64: np.float64, PF_F32: np.float32}


def need_shim(dtype):
    if not ref_sim.available(dtype, "shim"):
        pytest.skip("oracle/_ref/libshim_sim_*.so not built (needs /root/reference at build time)")


def direct(p, F0, U0, dtype, steps):
    with S.Solver(p.nx, p.ny, dtype) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        q, t, info = p.copy(), p.time, None
        for i in range(steps):
            q.iter, q.time = p.iter + i, t
            info = s.step(q, F, U, nF, nU)
            t += info.dt_advanced
            F, nF, U, nU = nF, F, nU, U
        return s.download(F), s.download(U), t - p.time, info


# every adaptive case gets a grid size of its own: the shim keeps one solver handle per size for the life of the process,
# and with it the adaptive step size - as the reference keeps it in a function static (simulation.cu:363)
SIZES = {("explicit", PF_F64): (136, 120), ("explicit", PF_F32): (136, 120),
         ("explicit-rk4", PF_F64): (200, 136), ("explicit-rk4", PF_F32): (200, 136),
         ("explicit-rk4-adaptive", PF_F64): (152, 104), ("explicit-rk4-adaptive", PF_F32): (168, 104),
         ("semi-implicit", PF_F64): (160, 128), ("semi-implicit", PF_F32): (160, 128)}


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
@pytest.mark.parametrize("solver", ["explicit", "explicit-rk4", "explicit-rk4-adaptive", "semi-implicit"])
def test_reference_entry_points_give_the_c_abi_result(solver, dtype):
    need_shim(dtype)
    nx, ny = SIZES[(solver, dtype)]
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.3)
    p.iter, p.time = 3, 0.25
    if solver == "explicit-rk4-adaptive":
        p.T_tolerance = p.Phi_tolerance = 1e-7 if dtype == PF_F64 else 1e-4
    F0, U0 = seeded_fields(nx, ny, NP[dtype], kind="interface")
    steps = 5
    wF, wU, wt, winfo = direct(p, F0, U0, dtype, steps)
    with ref_sim.RefSim(dtype, nx, ny, impl="shim") as sim:
        sim.upload(F0, U0)
        t, info = sim.run(p, steps)
        gF, gU = sim.download()
        assert sim.warnings == 0
    assert t == wt and info.dt_advanced == winfo.dt_advanced
    assert np.array_equal(gF, wF) and np.array_equal(gU, wU)
    assert not np.array_equal(gF, F0)
    if solver == "semi-implicit":
        # the counts travel as the reference's console lines (simulation.cu:889-890, :907-908), parsed back by the glue's logger
        assert (info.Phi_iters, info.T_iters) == (winfo.Phi_iters, winfo.T_iters) and info.Phi_iters >= 1
        assert info.Phi_converged == 1 and info.T_converged == 1
        assert abs(info.Phi_error - winfo.Phi_error) <= 1e-6 * winfo.Phi_error + 1e-300      # printed with %e: 7 digits
    if solver == "explicit-rk4-adaptive":
        assert (info.Phi_iters, info.T_iters) == (winfo.Phi_iters, winfo.T_iters)            # Sim_Stats.Phi_iters / T_iters (:488-492)


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
@pytest.mark.parametrize("solver", ["explicit", "explicit-rk4"])
def test_shim_and_unmodified_reference_side_by_side(solver, dtype):
    """Same glue, same inputs, two libraries: the reference's simulation.cu and this repository's replacement of it."""
    need_shim(dtype)
    if not ref_sim.available(dtype):
        pytest.skip("reference library not built")
    nx, ny = 256, 192
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.3)
    F0, U0 = seeded_fields(nx, ny, NP[dtype], kind="interface")
    out = {}
    for impl in ("ref", "shim"):
        with ref_sim.RefSim(dtype, nx, ny, fresh=True, impl=impl) as sim:
            sim.upload(F0, U0)
            t, _ = sim.run(p, 8)
            out[impl] = sim.download() + [t]
    assert out["ref"][2] == out["shim"][2]
    assert np.array_equal(out["ref"][0], out["shim"][0]) and np.array_equal(out["ref"][1], out["shim"][1])


def test_semi_implicit_through_the_shim_against_the_reference():
    need_shim(PF_F64)
    if not ref_sim.available(PF_F64):
        pytest.skip("reference library not built")
    nx, ny = 192, 192
    p = default_params(solver="semi-implicit", nx=nx, ny=ny, L0=nx / 128.0, S=0.3)
    F0, U0 = seeded_fields(nx, ny, np.float64, kind="interface")
    out = {}
    for impl in ("ref", "shim"):
        with ref_sim.RefSim(PF_F64, nx, ny, fresh=True, impl=impl) as sim:
            sim.upload(F0, U0)
            _, info = sim.run(p, 4)
            out[impl] = sim.download() + [info]
    for k in (0, 1):
        rel = np.linalg.norm(out["shim"][k] - out["ref"][k]) / np.linalg.norm(out["ref"][k])
        assert rel <= 1e-10
    assert abs(out["shim"][2].Phi_iters - out["ref"][2].Phi_iters) <= 1 and abs(out["shim"][2].T_iters - out["ref"][2].T_iters) <= 1


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
def test_stats_arrive_in_the_reference_struct(dtype):
    """Sim_Stats (simulation.h:57-83) filled through params.stats as simulation.cu:1126-1142 does."""
    need_shim(dtype)
    nx, ny = 144, 112
    p = default_params(solver="explicit-rk4", nx=nx, ny=ny, L0=nx / 128.0, S=0.3)
    p.do_stats = 1
    F0, U0 = seeded_fields(nx, ny, NP[dtype], kind="interface")
    _, _, _, winfo = direct(p, F0, U0, dtype, 3)
    with ref_sim.RefSim(dtype, nx, ny, impl="shim") as sim:
        sim.upload(F0, U0)
        _, info = sim.run(p, 3)
    f32 = np.float32
    for got, want in ((info.Phi_delta, winfo.Phi_delta), (info.T_delta, winfo.T_delta)):
        assert (f32(got.L1), f32(got.L2), f32(got.min), f32(got.max)) == (f32(want.L1), f32(want.L2), f32(want.min), f32(want.max))
        assert want.L2 > 0
    if ref_sim.available(dtype):
        with ref_sim.RefSim(dtype, nx, ny, fresh=True) as ref:
            ref.upload(F0, U0)
            _, rinfo = ref.run(p, 3)
        tol = 1e-5
        assert abs(rinfo.Phi_delta.L2 - info.Phi_delta.L2) <= tol * abs(rinfo.Phi_delta.L2)
        assert abs(rinfo.T_delta.max - info.T_delta.max) <= tol * abs(rinfo.T_delta.max)


def test_maps_are_zero_filled_resized_and_freed_like_the_reference():
    """sim_realloc (simulation.cu:88-101): new storage is zeroed, a second size gets a second handle, size 0 frees."""
    need_shim(PF_F64)
    for nx, ny in ((96, 80), (208, 64)):
        with ref_sim.RefSim(PF_F64, nx, ny, impl="shim") as sim:
            F, U = sim.download()
            assert F.shape == (ny, nx) and not F.any() and not U.any()
            a = np.arange(nx * ny, dtype=np.float32).reshape(ny, nx)               # float host array into a double map
            sim.upload(a, a * 2)
            F, U = sim.download(np.float32)
            assert np.array_equal(F, a) and np.array_equal(U, a * 2)


def test_two_grid_sizes_live_at_once():
    """sim_realloc frees and sim_modify* copies through the handle that OWNS a map (the shim's allocation table), so maps of
    different sizes can be live, stepped and freed in any order."""
    need_shim(PF_F64)
    pa = default_params(solver="explicit-rk4", nx=144, ny=96, L0=144 / 128.0)
    pb = default_params(solver="explicit-rk4", nx=208, ny=176, L0=208 / 128.0)
    Fa, Ua = seeded_fields(pa.nx, pa.ny, np.float64, kind="interface")
    Fb, Ub = seeded_fields(pb.nx, pb.ny, np.float64, kind="smooth")
    wa = direct(pa, Fa, Ua, PF_F64, 3)
    wb = direct(pb, Fb, Ub, PF_F64, 3)
    a = ref_sim.RefSim(PF_F64, pa.nx, pa.ny, impl="shim")
    b = ref_sim.RefSim(PF_F64, pb.nx, pb.ny, impl="shim")          # now the NEWEST handle is b's
    a.upload(Fa, Ua)                                                  # must go through a's handle
    b.upload(Fb, Ub)
    a.run(pa, 3); b.run(pb, 3)
    gFa, gUa = a.download()
    a.close()                                                         # frees a's maps while b's are live
    gFb, gUb = b.download()
    b.close()
    assert np.array_equal(gFa, wa[0]) and np.array_equal(gUa, wa[1])
    assert np.array_equal(gFb, wb[0]) and np.array_equal(gUb, wb[1])


def test_run_tests_and_run_benchmarks_have_bodies(capfd):
    """run_tests / run_benchmarks (reference simulation.h:143-144): the reduction micro-benchmark at the reference's six sizes."""
    need_shim(PF_F32)
    import ctypes
    L = ref_sim.load(PF_F32, impl="shim")
    L.run_tests.restype = ctypes.c_bool
    L.run_benchmarks.restype = ctypes.c_bool
    L.run_benchmarks.argtypes = [ctypes.c_int]
    assert L.run_tests()
    assert L.run_benchmarks(0)
    out = capfd.readouterr().out
    assert "cust_f32 = [" in out and out.count("N:") == 6


def test_reduction_benchmark_values():
    """pf_bench_reductions: the maximum of the hashed uniform sequence is just below 1; times are positive and ordered by size."""
    with S.Solver(64, 64, PF_F64) as s:
        t_small = s.bench_reductions(256 * 256, PF_F32, reps=5)
        t_big = s.bench_reductions(4096 * 4096, PF_F32, reps=5)
        assert 0.99 < t_small[2] < 1.0 and 0.999 < t_big[2] < 1.0
        assert 0 < t_small[0] < t_big[0] and 0 < t_small[1] < t_big[1]
        t64 = s.bench_reductions(1 << 20, PF_F64, reps=3)
        assert 0.99 < t64[2] < 1.0
