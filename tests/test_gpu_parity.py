"""Parity of the CUDA path (through the C-ABI, bachelors_b200/libpfsolver.so) with the CPU oracle, with the golden
outputs of the reference, and - when oracle/_ref was built - with the UNMODIFIED reference running on the same GPU.

Bars (BASELINE.json north_star): rel-L2 <= 1e-10 (fp64) / 1e-5 (fp32), CG iteration counts +-1, bit-stable runs.
The explicit solvers are held to the stronger bar of bit-identity with the oracle.
"""
import hashlib
import os

import numpy as np
import pytest

from bachelors_b200 import solver as S
from bachelors_b200._abi import (PF_ERR_INVALID, PF_ERR_UNSUPPORTED, PF_F32, PF_F64, PF_OPT_HOST_BANDS, PF_OPT_SI_GRID_SYNC,
                                 PF_OPT_SI_PER_KERNEL, PF_OPT_RUN_GRAPHS, default_params, pf_slab)
from bachelors_b200.config import read_config
from bachelors_b200.initial import make_initial_conditions
from bachelors_b200.testing import seeded_fields
from conftest import GOLDEN, ROOT
from golden_cases import all_cases, case_inputs
from oracle import ref_sim
from oracle.oracle import Oracle

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def tol_of(dtype):
    return 1e-10 if dtype == PF_F64 else 1e-5


def cuda_replay(case_or_params, F0=None, U0=None, dtype=None, steps=None):
    if isinstance(case_or_params, dict):
        p, F0, U0, dtype = case_inputs(case_or_params)
        steps = case_or_params["steps"]
    else:
        p = case_or_params
    with S.Solver(p.nx, p.ny, dtype) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        t = 0.0
        infos = []
        q = p.copy()
        for i in range(steps):
            q.iter = i; q.time = t
            info = s.step(q, F, U, nF, nU)
            t += info.dt_advanced
            infos.append((info.dt_advanced, info.Phi_iters, info.T_iters, info.Phi_converged, info.T_converged, info.Phi_error, info.T_error))
            F, nF = nF, F; U, nU = nU, U
        return s.download(F), s.download(U), t, np.array(infos, dtype=np.float64)


def oracle_replay(p, F0, U0, dtype, steps):
    o = Oracle(dtype, p.nx, p.ny)
    F, U, t, infos = F0, U0, 0.0, []
    q = p.copy()
    for i in range(steps):
        q.iter = i; q.time = t
        F, U, dt, info = o.step(q, F, U)
        t += dt
        infos.append((dt, info.Phi_iters, info.T_iters, info.Phi_converged, info.T_converged, info.Phi_error, info.T_error))
    return F, U, t, np.array(infos, dtype=np.float64)


def check_against(case, got, want, want_name):
    F, U, t, infos = got
    wF, wU, wt, winfos = want
    dtype = PF_F64 if case["dtype"] == "f64" else PF_F32
    eF, eU = rel_l2(F, wF), rel_l2(U, wU)
    assert eF <= tol_of(dtype) and eU <= tol_of(dtype), (want_name, eF, eU)
    assert abs(t - wt) <= (1e-12 if dtype == PF_F64 else 1e-6) * max(abs(wt), 1e-30), (want_name, t, wt)
    if case["solver"] == "semi-implicit":
        assert np.all(np.abs(infos[:, 1:3] - winfos[:, 1:3]) <= 1), (want_name, infos[:, 1:3], winfos[:, 1:3])
    if case["solver"] == "explicit-rk4-adaptive":
        assert np.array_equal(infos[:, 1], winfos[:, 1]), (want_name, infos[:, 1], winfos[:, 1])
    return eF, eU


# ---- the golden cases: CUDA vs oracle, vs the reference's stored outputs ---------------------------------------------
@pytest.mark.parametrize("case", all_cases(), ids=lambda c: c["name"])
def test_cuda_matches_oracle_and_golden(case):
    got = cuda_replay(case)
    p, F0, U0, dtype = case_inputs(case)
    want = oracle_replay(p, F0, U0, dtype, case["steps"])
    check_against(case, got, want, "oracle")
    if case["solver"] != "semi-implicit":
        # same arithmetic, same order: bit-identical (reductions in the explicit solvers are max-norms, order-free)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), "not bit-identical to the oracle"
        assert np.array_equal(got[3][:, 0], want[3][:, 0])
    path = os.path.join(GOLDEN, case["name"] + ".npz")
    assert os.path.exists(path), f"{path} missing"
    g = np.load(path)
    check_against(case, got, (g["F"], g["U"], float(g["t"]), g["step_info"]), "golden(reference on B200)")


LIVE = [c for c in all_cases() if c["name"].split("_f")[0] in
        ("euler_neumann", "euler_dirichlet_aniso", "rk4_neumann_aniso_noise", "rk4_periodic", "rkm_neumann", "si_neumann", "si_theta_half_guess")]


@pytest.mark.parametrize("case", LIVE, ids=lambda c: c["name"])
def test_cuda_matches_live_reference(case):
    """The unmodified reference (oracle A) on the same GPU, longer runs than the golden fixtures."""
    p, F0, U0, dtype = case_inputs(case)
    if not ref_sim.available(dtype):
        pytest.skip("oracle/_ref not built")
    steps = case["steps"] * 5
    with ref_sim.RefSim(dtype, p.nx, p.ny, fresh=True) as ref:
        ref.upload(F0, U0)
        infos, t = [], 0.0
        q = p.copy()
        for i in range(steps):
            q.iter = i; q.time = t
            dt, info = ref.run(q, 1)
            t += dt
            infos.append((dt, info.Phi_iters, info.T_iters, info.Phi_converged, info.T_converged, info.Phi_error, info.T_error))
        rF, rU = ref.download()
    got = cuda_replay(p, F0, U0, dtype, steps)
    eF, eU = check_against(case, got, (rF, rU, t, np.array(infos)), "live reference")
    if case["solver"] in ("explicit", "explicit-rk4"):
        assert eF == 0 and eU == 0, (eF, eU)


# ---- ragged / degenerate sizes against the oracle ---------------------------------------------------------------------
SIZES = [(1, 1), (1, 7), (5, 1), (2, 3), (63, 5), (65, 33), (129, 31), (57, 97), (257, 66)]


@pytest.mark.parametrize("bc", ["neumann", "dirichlet", "periodic"])
@pytest.mark.parametrize("solver", ["explicit", "explicit-rk4", "explicit-rk4-adaptive"])
def test_ragged_sizes_bit_identical(solver, bc):
    for dtype, npdt in ((PF_F64, np.float64), (PF_F32, np.float32)):
        for nx, ny in SIZES:
            tolv = 5e-9 if dtype == PF_F64 else 5e-5
            p = default_params(solver=solver, nx=nx, ny=ny, L0=max(nx, ny) / 128.0, S=0.25, Phi_boundary=bc, T_boundary=bc,
                               Phi_tolerance=tolv, T_tolerance=tolv)
            F0, U0 = seeded_fields(nx, ny, npdt, kind="noise", seed=nx * 1000 + ny)
            got = cuda_replay(p, F0, U0, dtype, 2)
            want = oracle_replay(p, F0, U0, dtype, 2)
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (nx, ny, dtype, rel_l2(got[0], want[0]))
            assert np.array_equal(got[3][:, :3], want[3][:, :3])


@pytest.mark.parametrize("bc", ["neumann", "dirichlet", "periodic"])
def test_semi_implicit_ragged_sizes(bc):
    for dtype, npdt in ((PF_F64, np.float64), (PF_F32, np.float32)):
        for nx, ny in [(1, 1), (3, 5), (65, 33), (300, 7), (257, 66)]:
            tolv = 5e-9 if dtype == PF_F64 else 1e-5
            p = default_params(solver="semi-implicit", nx=nx, ny=ny, L0=max(nx, ny) / 128.0, S=0.25, Phi_boundary=bc, T_boundary=bc,
                               Phi_tolerance=tolv, T_tolerance=tolv)
            F0, U0 = seeded_fields(nx, ny, npdt, kind="interface")
            got = cuda_replay(p, F0, U0, dtype, 2)
            want = oracle_replay(p, F0, U0, dtype, 2)
            assert rel_l2(got[0], want[0]) <= tol_of(dtype) and rel_l2(got[1], want[1]) <= tol_of(dtype), (nx, ny, dtype)
            assert np.all(np.abs(got[3][:, 1:3] - want[3][:, 1:3]) <= 1), (nx, ny, got[3][:, 1:3], want[3][:, 1:3])


# ---- building blocks ------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
def test_rhs_blends_match_oracle(dtype):
    npdt = np.float64 if dtype == PF_F64 else np.float32
    nx, ny = 70, 45
    rng = np.random.default_rng(5)
    for Fb, Ub in (("neumann", "neumann"), ("periodic", "dirichlet"), ("dirichlet", "periodic")):
        p = default_params(solver="explicit-rk4", nx=nx, ny=ny, L0=nx / 128.0, S=0.3, Phi_boundary=Fb, T_boundary=Ub)
        fields = [seeded_fields(nx, ny, npdt, kind="noise", seed=i + 1) for i in range(4)]
        for n in (1, 2, 3, 4):
            w = [1.0] + list(rng.uniform(-1e-5, 1e-5, n - 1))
            o = Oracle(dtype, nx, ny)
            wF, wU = o.rhs(p, w, [f[0] for f in fields[:n]], [f[1] for f in fields[:n]])
            with S.Solver(nx, ny, dtype) as s:
                Fs = [s.field_from(f[0]) for f in fields[:n]]
                Us = [s.field_from(f[1]) for f in fields[:n]]
                oF, oU = s.field_alloc(), s.field_alloc()
                s.rhs(p, w, Fs, Us, oF, oU)
                gF, gU = s.download(oF), s.download(oU)
            assert np.array_equal(gF, wF) and np.array_equal(gU, wU), (Fb, Ub, n)


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
def test_cg_solve_matches_oracle(dtype):
    npdt = np.float64 if dtype == PF_F64 else np.float32
    nx, ny = 96, 80
    for bc in ("neumann", "dirichlet", "periodic"):
        p = default_params(solver="semi-implicit", nx=nx, ny=ny, L0=nx / 128.0, Phi_boundary=bc, T_boundary=bc, dt=2e-5)
        rng = np.random.default_rng(11)
        factor = rng.uniform(0.2, 0.5, (ny, nx)).astype(npdt)
        b, x0 = seeded_fields(nx, ny, npdt, kind="smooth")
        for which in (0, 1):
            for use_x0 in (True, False):
                tolv = 1e-9 if dtype == PF_F64 else 1e-5
                o = Oracle(dtype, nx, ny)
                wx, wit, wconv, werr = o.cg_solve(p, which, factor if which == 0 else None, b, x0 if use_x0 else None, tolv, 50)
                with S.Solver(nx, ny, dtype) as s:
                    db, dx0, dfac, dx = s.field_from(b), s.field_from(x0), s.field_from(factor), s.field_alloc()
                    it, conv, err = s.cg_solve(p, which, dfac if which == 0 else None, db, dx0 if use_x0 else None, dx, tolv, 50)
                    gx = s.download(dx)
                assert abs(it - wit) <= 1 and conv == wconv, (bc, which, use_x0, it, wit)
                assert rel_l2(gx, wx) <= tol_of(dtype)
                assert wit >= 1 and np.isfinite(err)


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
def test_debug_values_and_stats(dtype):
    npdt = np.float64 if dtype == PF_F64 else np.float32
    nx, ny = 77, 51
    p = default_params(nx=nx, ny=ny, L0=nx / 128.0, S=0.3, Phi_boundary="periodic", T_boundary="neumann")
    F0, U0 = seeded_fields(nx, ny, npdt, kind="interface")
    o = Oracle(dtype, nx, ny)
    wg, wh, wa = o.debug_values(p, F0, U0)
    ws = o.stats_delta(F0, U0)
    with S.Solver(nx, ny, dtype) as s:
        F, U = s.field_from(F0), s.field_from(U0)
        g, h, a = s.field_alloc(), s.field_alloc(), s.field_alloc()
        s.debug_values(p, F, U, g, h, a)
        assert np.array_equal(s.download(g), wg) and np.array_equal(s.download(h), wh) and np.array_equal(s.download(a), wa)
        st = s.stats_delta(F, U)
    rt = 1e-12 if dtype == PF_F64 else 1e-4
    assert st.min == ws.min and st.max == ws.max
    assert abs(st.L1 - ws.L1) <= rt * abs(ws.L1) and abs(st.L2 - ws.L2) <= rt * abs(ws.L2) and abs(st.sum - ws.sum) <= rt * abs(ws.L1)


def test_step_stats_and_residuals_match_oracle():
    nx, ny = 64, 48
    p = default_params(solver="explicit", nx=nx, ny=ny, L0=nx / 128.0, do_stats=1, do_stats_step_residual=1, do_corrector_loop=1, corrector_max_iters=2)
    F0, U0 = seeded_fields(nx, ny, kind="interface")
    o = Oracle(PF_F64, nx, ny)
    wF, wU, _, winfo = o.step(p, F0, U0)
    with S.Solver(nx, ny, PF_F64) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        info = s.step(p, F, U, nF, nU)
        assert np.array_equal(s.download(nF), wF) and np.array_equal(s.download(nU), wU)
    assert info.step_res_count == winfo.step_res_count == 2
    for a, b in ((info.Phi_delta, winfo.Phi_delta), (info.T_delta, winfo.T_delta), (info.step_res[0], winfo.step_res[0]), (info.step_res[1], winfo.step_res[1])):
        assert a.min == b.min and a.max == b.max and abs(a.L2 - b.L2) <= 1e-12 * max(abs(b.L2), 1e-300) and abs(a.L1 - b.L1) <= 1e-12 * max(abs(b.L1), 1e-300)


def test_upload_download_convert():
    nx, ny = 50, 30
    a64, _ = seeded_fields(nx, ny, np.float64, kind="noise")
    with S.Solver(nx, ny, PF_F32) as s:
        f = s.field_from(a64)                               # double host -> float device (sim_modify_double on a USE_FLOATS build)
        assert np.array_equal(s.download(f), a64.astype(np.float32))
        assert np.array_equal(s.download(f, dtype=np.float64), a64.astype(np.float32).astype(np.float64))
    with S.Solver(nx, ny, PF_F64) as s:
        f = s.field_from(a64.astype(np.float32))            # float host -> double device (sim_modify_float)
        assert np.array_equal(s.download(f), a64.astype(np.float32).astype(np.float64))
        assert np.array_equal(s.download(f, dtype=np.float32), a64.astype(np.float32))
        z = s.field_alloc()
        assert not s.download(z).any()                      # REALLOC_ZERO


# ---- stepping entry points ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("solver", ["explicit", "explicit-rk4", "explicit-rk4-adaptive", "semi-implicit"])
def test_run_equals_repeated_steps_and_is_bit_stable(solver):
    nx, ny = 200, 136
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.2)
    F0, U0 = seeded_fields(nx, ny, kind="interface")
    ref = cuda_replay(p, F0, U0, PF_F64, 7)
    digests = set()
    for _ in range(3):
        with S.Solver(nx, ny, PF_F64) as s:
            F, U, tF, tU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
            Fo, Uo, done, t, info = s.run(p, F, U, tF, tU, 7)
            gF, gU = s.download(Fo), s.download(Uo)
        assert done == 7 and abs(t - ref[2]) <= 1e-15
        assert np.array_equal(gF, ref[0]) and np.array_equal(gU, ref[1])
        digests.add(hashlib.sha256(gF.tobytes() + gU.tobytes()).hexdigest())
    assert len(digests) == 1


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32], ids=["f64", "f32"])
@pytest.mark.parametrize("bc", ["neumann", "periodic", "dirichlet"])
@pytest.mark.parametrize("solver,size,steps", [("explicit-rk4", (400, 336), 7), ("explicit-rk4", (512, 512), 12), ("explicit-rk4", (333, 1025), 9),
                                               ("explicit-rk4", (1024, 1024), 6), ("explicit-rk4", (2048, 2048), 5), ("explicit-rk4", (37, 640), 4)])
def test_run_is_one_multi_step_launch_with_the_bits_of_single_steps(solver, size, steps, bc, dtype):
    """pf_run of explicit-rk4 = ONE cooperative launch that advances all the steps (k_stream_run: every unit waits only for its
    8 neighbouring tiles, no grid barrier, buffers swap roles every step): the same bits as `steps` single-step launches, for odd
    and even step counts, every boundary type, grids of one and of many strips."""
    nx, ny = size
    if size == (2048, 2048) and (bc != "neumann" or dtype != PF_F64):
        pytest.skip("full size once")
    # L0 such that the finer of dx, dy is 1/128 (the reference has one L0 for both axes): the run stays stable, no NaNs to compare
    p = default_params(solver=solver, nx=nx, ny=ny, L0=max(nx, ny) / 128.0, S=0.2 if nx < 1000 else 0.0, Phi_boundary=bc, T_boundary=bc)
    npdt = np.float64 if dtype == PF_F64 else np.float32
    F0, U0 = seeded_fields(nx, ny, npdt, kind="interface")
    ref = cuda_replay(p, F0, U0, dtype, steps)
    assert np.isfinite(ref[0]).all() and np.isfinite(ref[1]).all()
    with S.Solver(nx, ny, dtype) as s:
        F, U, tF, tU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        l0 = s.kernel_launches
        Fo, Uo, done, t, info = s.run(p, F, U, tF, tU, steps)
        launches = s.kernel_launches - l0                    # 1: the multi-step path ran; small grids (chunks shorter than 2 x 4 rows) keep single-step launches
        assert launches == 1 if nx * ny >= 1024 * 1024 else launches in (1, steps)
        assert done == steps and Fo == (tF if steps % 2 else F) and abs(t - ref[2]) <= 1e-15
        assert np.array_equal(s.download(Fo), ref[0]) and np.array_equal(s.download(Uo), ref[1])
        # a second run continues from the first one's state (the progress words are not reset between launches)
        Fo2, Uo2, done2, _, _ = s.run(p, Fo, Uo, tF if Fo == F else F, tU if Uo == U else U, 3)
        ref2 = cuda_replay(p, ref[0], ref[1], dtype, 3)
        assert done2 == 3 and np.array_equal(s.download(Fo2), ref2[0]) and np.array_equal(s.download(Uo2), ref2[1])


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32], ids=["f64", "f32"])
@pytest.mark.parametrize("solver,size,steps", [("explicit", (128, 128), 134), ("explicit", (300, 171), 200), ("explicit", (512, 512), 65),
                                               ("explicit-rk4", (96, 40), 150)])
def test_run_replays_single_launch_steps_as_cuda_graphs(solver, size, steps, dtype):
    """pf_run of the fixed-step solvers whose step is one launch (explicit Euler; explicit-rk4 on grids too small for the
    multi-step launch): batches of 64 steps are captured once and replayed as a CUDA graph (PF_OPT_RUN_GRAPHS, default on).
    Same bits, same step count, same time and the same number of kernels as with the graphs switched off, also when a second
    run continues with the buffers in the other orientation and when a stop time cuts the batch."""
    nx, ny = size
    p = default_params(solver=solver, nx=nx, ny=ny, L0=max(nx, ny) / 128.0, S=0.2)
    npdt = np.float64 if dtype == PF_F64 else np.float32
    F0, U0 = seeded_fields(nx, ny, npdt, kind="interface")

    def run(graphs):
        with S.Solver(nx, ny, dtype) as s:
            s.set_option(PF_OPT_RUN_GRAPHS, graphs)
            F, U, tF, tU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
            l0 = s.kernel_launches
            Fo, Uo, done, t, _ = s.run(p, F, U, tF, tU, steps)
            q = p.copy(); q.iter, q.time = p.iter + done, p.time + t
            oF, oU = (tF, tU) if Fo == F else (F, U)
            Fo2, Uo2, done2, t2, _ = s.run(q, Fo, Uo, oF, oU, 10 ** 6, stop_time=q.time + 70.5 * p.dt)     # 71 steps, then the stop time
            return s.download(Fo2), s.download(Uo2), done, done2, t, t2, s.kernel_launches - l0, Fo == (tF if steps % 2 else F)

    gF, gU, gdone, gdone2, gt, gt2, glaunch, gorient = run(1)
    pF, pU, pdone, pdone2, pt, pt2, plaunch, porient = run(0)
    assert (gdone, gdone2) == (steps, 71) == (pdone, pdone2) and gorient and porient
    assert gt == pt and gt2 == pt2
    assert glaunch == plaunch == steps + 71
    assert np.isfinite(gF).all() and np.array_equal(gF, pF) and np.array_equal(gU, pU)


@pytest.mark.parametrize("solver,size,steps", [("explicit-rk4", (1200, 1030), 7), ("explicit", (300, 171), 70), ("explicit-rk4-adaptive", (200, 136), 5),
                                               ("semi-implicit", (160, 128), 4)])
def test_run_with_statistics_keeps_the_fast_paths(solver, size, steps):
    """pf_run / pf_run_until with do_stats: the steps run without statistics (explicit-rk4: still ONE multi-step launch, explicit:
    still the graph replay) and last_info carries the delta statistics of the LAST step, formed after the loop from the two
    ping-pong buffers - bit for bit what pf_step with do_stats returns for that step, with the same fields."""
    nx, ny = size
    p = default_params(solver=solver, nx=nx, ny=ny, L0=max(nx, ny) / 128.0, S=0.2)
    if solver == "explicit-rk4-adaptive":
        p.T_tolerance = p.Phi_tolerance = 1e-7
    p.do_stats = 1
    F0, U0 = seeded_fields(nx, ny, np.float64, kind="interface")
    with S.Solver(nx, ny, PF_F64) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        q, t = p.copy(), p.time
        for i in range(steps):
            q.iter, q.time = p.iter + i, t
            info = s.step(q, F, U, nF, nU)
            t += info.dt_advanced
            F, nF, U, nU = nF, F, nU, U
        want = (s.download(F), s.download(U), t)
        want_stats = [(d.sum, d.L1, d.L2, d.min, d.max) for d in (info.Phi_delta, info.T_delta)]
    with S.Solver(nx, ny, PF_F64) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        l0 = s.kernel_launches
        Fo, Uo, done, _, rinfo = s.run(p, F, U, nF, nU, steps)
        launches = s.kernel_launches - l0
        assert done == steps and s.last_run_end_time == want[2]
        assert np.array_equal(s.download(Fo), want[0]) and np.array_equal(s.download(Uo), want[1])
        got_stats = [(d.sum, d.L1, d.L2, d.min, d.max) for d in (rinfo.Phi_delta, rinfo.T_delta)]
        assert got_stats == want_stats and want_stats[0][1] > 0
        if solver == "explicit-rk4":
            assert launches == 1 + 2                                       # the multi-step launch + the two statistics kernels
        if solver == "explicit":
            assert launches == steps + 2


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32], ids=["f64", "f32"])
@pytest.mark.parametrize("variant", ["guess_periodic", "neumann_dirichlet", "dirichlet_neumann_aniso_guess", "corrector_loop", "corrector_loop_mixed", "exact_source"])
def test_euler_parameter_space_is_one_launch_per_rhs(variant, dtype):
    """The Euler tile kernel (pf_euler.cuh, GENERAL instantiation) serves the reference's whole Euler parameter space - corrector
    guess (simulation.cu:221-225), different non-periodic boundary types of Phi and T (:186-187), the corrector loop (:928-1008), the
    source term of do_exact - with ONE launch per right-hand-side evaluation, bit-identical to the CPU oracle."""
    kw = dict(guess_periodic=dict(do_corrector_guess=1, Phi_boundary="periodic", T_boundary="periodic"),
              neumann_dirichlet=dict(Phi_boundary="neumann", T_boundary="dirichlet"),
              dirichlet_neumann_aniso_guess=dict(Phi_boundary="dirichlet", T_boundary="neumann", S=0.3, do_corrector_guess=1),
              corrector_loop=dict(do_corrector_loop=1, corrector_max_iters=3, do_corrector_guess=1),
              corrector_loop_mixed=dict(do_corrector_loop=1, corrector_max_iters=2, Phi_boundary="dirichlet", T_boundary="neumann", S=0.2),
              exact_source=dict(do_exact=1))[variant]
    nx, ny, steps = 157, 93, 4
    p = default_params(solver="explicit", nx=nx, ny=ny, L0=nx / 128.0, **kw)
    npdt = np.float64 if dtype == PF_F64 else np.float32
    F0, U0 = seeded_fields(nx, ny, npdt, kind="interface")
    want = oracle_replay(p, F0, U0, dtype, steps)
    with S.Solver(nx, ny, dtype) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        l0 = s.kernel_launches
        q = p.copy()
        for i in range(steps):
            q.iter, q.time = i, i * p.dt
            s.step(q, F, U, nF, nU)
            F, nF = nF, F; U, nU = nU, U
        launches = s.kernel_launches - l0
        gF, gU = s.download(F), s.download(U)
    iters = p.corrector_max_iters if p.do_corrector_loop else 0
    assert launches == steps * (1 + iters), launches
    assert np.array_equal(gF, want[0]) and np.array_equal(gU, want[1])


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32], ids=["f64", "f32"])
@pytest.mark.parametrize("solver", ["explicit-rk4", "explicit-rk4-adaptive"])
@pytest.mark.parametrize("bcs", [("neumann", "dirichlet"), ("dirichlet", "neumann")])
def test_streaming_kernel_takes_different_boundary_types_for_phi_and_t(solver, bcs, dtype):
    """Independent non-periodic boundary types of Phi and T (simulation.cu:186-187) stay on the one-launch streaming kernel:
    one launch per step / attempt, bit-identical to the CPU oracle."""
    nx, ny, steps = 261, 150, 3
    tol = 1e-7 if dtype == PF_F64 else 1e-4
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.2, Phi_boundary=bcs[0], T_boundary=bcs[1], Phi_tolerance=tol, T_tolerance=tol)
    npdt = np.float64 if dtype == PF_F64 else np.float32
    F0, U0 = seeded_fields(nx, ny, npdt, kind="interface")
    want = oracle_replay(p, F0, U0, dtype, steps)
    with S.Solver(nx, ny, dtype) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        l0, attempts, t = s.kernel_launches, 0, 0.0
        q = p.copy()
        for i in range(steps):
            q.iter, q.time = i, t
            info = s.step(q, F, U, nF, nU)
            t += info.dt_advanced
            attempts += info.Phi_iters
            F, nF = nF, F; U, nU = nU, U
        assert s.kernel_launches - l0 == attempts
        assert np.array_equal(s.download(F), want[0]) and np.array_equal(s.download(U), want[1]) and abs(t - want[2]) <= 1e-15


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32], ids=["f64", "f32"])
@pytest.mark.parametrize("solver", ["explicit-rk4", "explicit-rk4-adaptive"])
@pytest.mark.parametrize("bc,S_aniso,kind", [("periodic", 0.0, "noise"), ("neumann", 0.25, "interface"), ("dirichlet", 0.0, "interface"),
                                             (("dirichlet", "neumann"), 0.2, "noise")])
def test_streaming_kernel_takes_the_corrector_guess_rhs(solver, bc, S_aniso, kind, dtype):
    """The corrector-guess form of the right-hand side (simulation.cu:221-225) runs on instantiations of the one-launch streaming
    kernel (round 1: a separate 2-D tiled kernel): one launch per step / attempt, bit-identical to the CPU oracle, with noisy
    fields (signed zeros included), anisotropy on and off, every boundary type, grids of several strips."""
    nx, ny, steps = 261, 150, 3
    tol = 1e-7 if dtype == PF_F64 else 1e-4
    bcF, bcT = bc if isinstance(bc, tuple) else (bc, bc)
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=S_aniso, Phi_boundary=bcF, T_boundary=bcT, do_corrector_guess=1,
                       Phi_tolerance=tol, T_tolerance=tol)
    npdt = np.float64 if dtype == PF_F64 else np.float32
    F0, U0 = seeded_fields(nx, ny, npdt, kind=kind)
    want = oracle_replay(p, F0, U0, dtype, steps)
    with S.Solver(nx, ny, dtype) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        l0, attempts, t = s.kernel_launches, 0, 0.0
        q = p.copy()
        for i in range(steps):
            q.iter, q.time = i, t
            info = s.step(q, F, U, nF, nU)
            t += info.dt_advanced
            attempts += info.Phi_iters
            F, nF = nF, F; U, nU = nU, U
        assert s.kernel_launches - l0 == attempts
        assert np.array_equal(s.download(F), want[0]) and np.array_equal(s.download(U), want[1]) and abs(t - want[2]) <= 1e-15


def test_run_stops_at_stop_time():
    nx = ny = 64
    p = default_params(solver="explicit-rk4", nx=nx, ny=ny, L0=0.5)
    F0, U0 = seeded_fields(nx, ny, kind="interface")
    with S.Solver(nx, ny, PF_F64) as s:
        F, U, tF, tU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        Fo, Uo, done, t, _ = s.run(p, F, U, tF, tU, 1000, stop_time=10.5 * p.dt)
        assert done == 11 and Fo == tF


def test_async_steps_then_sync():
    nx, ny = 96, 96
    p = default_params(solver="explicit-rk4", nx=nx, ny=ny, L0=nx / 128.0)
    F0, U0 = seeded_fields(nx, ny, kind="interface")
    want = cuda_replay(p, F0, U0, PF_F64, 4)
    with S.Solver(nx, ny, PF_F64) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        for _ in range(4):
            s.step_async(p, F, U, nF, nU)
            F, nF = nF, F; U, nU = nU, U
        s.sync()
        assert np.array_equal(s.download(F), want[0])
        assert s.kernel_launches == 4                        # one fused kernel per RK4 step


def test_errors_are_returned_not_fatal():
    with S.Solver(32, 32, PF_F64) as s:
        F, U, nF, nU = (s.field_alloc() for _ in range(4))
        with pytest.raises(S.PfError) as e:
            s.step(default_params(nx=64, ny=32), F, U, nF, nU)           # reference: "Bad sizes provided", simulation.cu:1094
        assert e.value.code == PF_ERR_INVALID
        with pytest.raises(S.PfError) as e:
            s.step_async(default_params(nx=32, ny=32, solver="explicit-rk4-adaptive"), F, U, nF, nU)   # dt is a host decision
        assert e.value.code == PF_ERR_UNSUPPORTED
        with pytest.raises(S.PfError) as e:
            s.step(default_params(nx=32, ny=32), F, U, F, nU)
        assert e.value.code == PF_ERR_INVALID
        s.step(default_params(nx=32, ny=32), F, U, nF, nU)               # the handle stays usable


# ---- slabs (the multi-GPU decomposition), emulated on one device --------------------------------------------------------
@pytest.mark.parametrize("bc", ["neumann", "periodic", "dirichlet"])
@pytest.mark.parametrize("solver", ["explicit", "explicit-rk4", "explicit-rk4-adaptive"])
def test_slabs_reproduce_the_single_grid_bit_for_bit(solver, bc):
    from bachelors_b200.slabs import slab_layout
    nx, ny, G, steps = 150, 131, 3, 3
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.2, Phi_boundary=bc, T_boundary=bc, Phi_tolerance=1e-7, T_tolerance=1e-7)
    F0, U0 = seeded_fields(nx, ny, kind="interface")
    want = cuda_replay(p, F0, U0, PF_F64, steps)
    L = S.lib()
    ghost = L.pf_required_ghost(p.solver)
    slabs = [slab_layout(ny, G, g, ghost, periodic=(bc == "periodic")) for g in range(G)]
    solvers = [S.Solver(nx, ny, PF_F64, slab=sl) for sl in slabs]
    try:
        def padded(a, sl, s):
            rows = (np.arange(s.rows_alloc) - s.first_owned_row + sl.y0) % ny
            return a[rows]
        st = []
        for sl, s in zip(slabs, solvers):
            st.append([s.field_from(padded(F0, sl, s)), s.field_from(padded(U0, sl, s)), s.field_alloc(), s.field_alloc()])
        t = 0.0
        q = p.copy()
        for i in range(steps):
            q.iter = i; q.time = t
            # every rank must take the same adaptive decision: step all slabs with a shared error norm
            from bachelors_b200.slabs import step_slabs_lockstep
            dt = step_slabs_lockstep(solvers, q, st)
            t += dt
            for g, (sl, s) in enumerate(zip(slabs, solvers)):
                st[g][0], st[g][2] = st[g][2], st[g][0]
                st[g][1], st[g][3] = st[g][3], st[g][1]
            # halo exchange through the host (the transport is the caller's: NCCL / peer copies on a real multi-GPU run)
            full_F = np.empty((ny, nx)); full_U = np.empty((ny, nx))
            for sl, s, f in zip(slabs, solvers, st):
                a = s.download(f[0]); b = s.download(f[1])
                full_F[sl.y0:sl.y0 + sl.ny_local] = a[s.first_owned_row:s.first_owned_row + sl.ny_local]
                full_U[sl.y0:sl.y0 + sl.ny_local] = b[s.first_owned_row:s.first_owned_row + sl.ny_local]
            for sl, s, f in zip(slabs, solvers, st):
                s.upload(f[0], padded(full_F, sl, s)); s.upload(f[1], padded(full_U, sl, s))
        assert abs(t - want[2]) <= 1e-18
        assert np.array_equal(full_F, want[0]) and np.array_equal(full_U, want[1])
    finally:
        for s in solvers:
            s.close()


# ---- full-size, size-independent properties (BASELINE.json sizes) ---------------------------------------------------------
@pytest.mark.parametrize("solver,dtype", [("explicit", PF_F64), ("explicit-rk4", PF_F64), ("explicit-rk4", PF_F32), ("explicit-rk4-adaptive", PF_F64)])
def test_full_size_2048_properties(solver, dtype):
    cfg = read_config(os.path.join(ROOT, "benchmark_configs", f"config_{solver}_2048.ini"))
    p = cfg.params
    npdt = np.float64 if dtype == PF_F64 else np.float32
    F0, U0 = make_initial_conditions(cfg, npdt)
    assert F0.shape == (2048, 2048)
    with S.Solver(p.nx, p.ny, dtype) as s:
        F, U, tF, tU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        Fo, Uo, done, t, info = s.run(p, F, U, tF, tU, 40)
        gF, gU = s.download(Fo).astype(np.float64), s.download(Uo).astype(np.float64)
    assert done == 40 and np.isfinite(gF).all() and np.isfinite(gU).all()
    # enthalpy sum(T - L*Phi) is conserved under Neumann boundaries
    h0 = (U0.astype(np.float64) - p.L * F0.astype(np.float64)).sum(); h1 = (gU - p.L * gF).sum()
    assert abs(h1 - h0) <= (1e-9 if dtype == PF_F64 else 2e-3) * abs(h0), (h0, h1)
    # mirror symmetry of the centred seed, to rounding
    assert np.allclose(gF, gF[::-1, :], rtol=1e-6, atol=1e-9) and np.allclose(gF, gF[:, ::-1], rtol=1e-6, atol=1e-9)
    # far from the seed nothing happened; at the seed something did
    assert not gF[:512].any() and np.abs(gF - F0).max() > 1e-3
    # the centre 256^2 window equals the oracle run on the whole grid's window? (local dependence: 40 steps x 4-5 stages < 256 cells)
    if dtype == PF_F64 and solver != "explicit-rk4-adaptive":
        w = 512
        c0 = 1024 - w // 2
        sub = default_params(**{k: getattr(p, k) for k, _ in p._fields_})
        sub.nx = sub.ny = w; sub.L0 = p.L0 * w / p.nx
        o = Oracle(PF_F64, w, w)
        oF, oU, _, _ = o.run(sub, F0[c0:c0 + w, c0:c0 + w], U0[c0:c0 + w, c0:c0 + w], 40)
        inner = slice(200, w - 200)
        assert np.array_equal(oF[inner, inner], gF[c0:c0 + w, c0:c0 + w][inner, inner])
        assert np.array_equal(oU[inner, inner], gU[c0:c0 + w, c0:c0 + w][inner, inner])


def test_semi_implicit_1024_iteration_counts():
    cfg = read_config(os.path.join(ROOT, "benchmark_configs", "config_semi-implicit_1024.ini"))
    p = cfg.params
    F0, U0 = make_initial_conditions(cfg)
    with S.Solver(p.nx, p.ny, PF_F64) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        its = []
        for i in range(5):
            info = s.step(p, F, U, nF, nU)
            its.append((info.Phi_iters, info.T_iters))
            F, nF = nF, F; U, nU = nU, U
        gF = s.download(F)
    o = Oracle(PF_F64, p.nx, p.ny)
    oF, oU, wits = F0, U0, []
    for i in range(5):
        oF, oU, _, info = o.step(p, oF, oU)
        wits.append((info.Phi_iters, info.T_iters))
    assert np.all(np.abs(np.array(its) - np.array(wits)) <= 1), (its, wits)
    assert rel_l2(gF, oF) <= 1e-10


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32], ids=["f64", "f32"])
@pytest.mark.parametrize("size", [(300, 171), (512, 203), (1024, 1024), (256, 37)], ids=lambda s: "%dx%d" % s)
@pytest.mark.parametrize("variant", ["neumann", "periodic_aniso", "dirichlet_theta_half", "corrector_loop"])
def test_semi_implicit_cooperative_step_equals_per_kernel_path(variant, size, dtype):
    """The one-launch cooperative step (pf_si_coop.cuh) and the per-kernel CG path (pf_set_option PF_OPT_SI_PER_KERNEL) share
    decomposition, arithmetic and summation trees: fields, iteration counts and residuals must agree bit for bit.
    PF_OPT_SI_GRID_SYNC replaces the flagged all-block sums of the cooperative step (GridSum, pf_si_coop.cuh) by grid.sync() +
    partial arrays - a third identical run."""
    kw = dict(neumann=dict(),
              periodic_aniso=dict(Phi_boundary="periodic", T_boundary="periodic", S=0.3),
              dirichlet_theta_half=dict(Phi_boundary="dirichlet", T_boundary="dirichlet", gamma=0.5, do_corrector_guess=1),
              corrector_loop=dict(do_corrector_loop=1, corrector_max_iters=2))[variant]
    nx, ny = size
    if size == (1024, 1024) and variant not in ("neumann", "periodic_aniso"):
        pytest.skip("full size: two variants are enough")
    p = default_params(solver="semi-implicit", nx=nx, ny=ny, L0=nx / 128.0, Phi_tolerance=1e-9, T_tolerance=1e-9, Phi_max_iters=30, T_max_iters=30, **kw)
    npdt = np.float64 if dtype == PF_F64 else np.float32
    F0, U0 = seeded_fields(nx, ny, npdt, kind="interface")

    def run(options=()):
        with S.Solver(nx, ny, dtype) as s:
            for opt in options:
                s.set_option(opt, 1)
            F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
            infos, launches0 = [], s.kernel_launches
            for _ in range(4):
                info = s.step(p, F, U, nF, nU)
                infos.append((info.Phi_iters, info.T_iters, info.Phi_converged, info.T_converged, info.Phi_error, info.T_error))
                F, nF = nF, F; U, nU = nU, U
            return s.download(F), s.download(U), infos, s.kernel_launches - launches0

    cF, cU, cinfo, claunches = run()
    gF_, gU_, ginfo, glaunches = run([PF_OPT_SI_GRID_SYNC])             # cooperative grid.sync() + partial arrays instead of the flagged sums
    assert np.array_equal(cF, gF_) and np.array_equal(cU, gU_) and cinfo == ginfo and claunches == glaunches
    kF, kU, kinfo, klaunches = run([PF_OPT_SI_GRID_SYNC, PF_OPT_SI_PER_KERNEL])
    assert np.array_equal(cF, kF) and np.array_equal(cU, kU)
    assert cinfo == kinfo, (cinfo, kinfo)
    assert claunches < klaunches / 10, (claunches, klaunches)          # the cooperative path really ran
    assert max(i[0] for i in cinfo) >= 2                               # and the solves really iterated


@pytest.mark.parametrize("solver,dtype,host_dtype", [("explicit-rk4", PF_F64, PF_F64), ("explicit", PF_F64, PF_F64), ("explicit-rk4", PF_F32, PF_F32),
                                                     ("explicit-rk4", PF_F64, PF_F32), ("explicit-rk4-adaptive", PF_F64, PF_F64), ("semi-implicit", PF_F64, PF_F64)])
@pytest.mark.parametrize("bc", ["neumann", "periodic"])
def test_step_host_equals_upload_step_download(solver, dtype, host_dtype, bc):
    """pf_step_host (banded upload / compute / download pipeline for the fixed-step explicit solvers) returns exactly what the
    three separate calls return, over several steps with the outputs fed back as inputs. The device fields are poisoned before
    every call, so a band that computes before all of its inputs are uploaded (periodic: band 0 reads the wrapped last rows)
    cannot pass on stale device contents."""
    nx, ny = 333, 700
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.2, Phi_tolerance=1e-7, T_tolerance=1e-7, Phi_boundary=bc, T_boundary=bc)
    npdt = np.float64 if host_dtype == PF_F64 else np.float32
    F0, U0 = seeded_fields(nx, ny, npdt, kind="interface")
    poison = np.full((ny, nx), np.nan, dtype=npdt)
    with S.Solver(nx, ny, dtype) as a, S.Solver(nx, ny, dtype) as b:
        a.set_option(PF_OPT_HOST_BANDS, 10)
        aF, aU, anF, anU = a.field_alloc(), a.field_alloc(), a.field_alloc(), a.field_alloc()
        bF, bU, bnF, bnU = b.field_alloc(), b.field_alloc(), b.field_alloc(), b.field_alloc()
        hF, hU = F0.copy(), U0.copy()
        gF, gU = F0.copy(), U0.copy()
        for it in range(3):
            p.do_stats = int(it == 2)                                   # the pipelined path keeps running; statistics after the last band
            oF, oU = np.empty_like(hF), np.empty_like(hU)
            L = S.lib()
            info = S.pf_step_info()
            import ctypes
            a.upload(aF, poison); a.upload(aU, poison)
            a._ck(L.pf_step_host(a._h, ctypes.byref(p), hF.ctypes.data, hU.ctypes.data, oF.ctypes.data, oU.ctypes.data, host_dtype, aF, aU, anF, anU, ctypes.byref(info)))
            b.upload(bF, gF); b.upload(bU, gU)
            winfo = b.step(p, bF, bU, bnF, bnU)
            wF, wU = b.download(bnF, dtype=npdt), b.download(bnU, dtype=npdt)
            assert np.array_equal(oF, wF) and np.array_equal(oU, wU)
            assert info.dt_advanced == winfo.dt_advanced and info.Phi_iters == winfo.Phi_iters
            if p.do_stats and host_dtype == dtype:
                for d, w in ((info.Phi_delta, winfo.Phi_delta), (info.T_delta, winfo.T_delta)):
                    assert (d.sum, d.L1, d.L2, d.min, d.max) == (w.sum, w.L1, w.L2, w.min, w.max) and w.L1 > 0
            hF, hU, gF, gU = oF, oU, wF, wU
            aF, anF, aU, anU = anF, aF, anU, aU
            bF, bnF, bU, bnU = bnF, bF, bnU, bU


# ---- libdevice ----------------------------------------------------------------------------------------------------------------
def test_libdevice_restatement_bit_identical_on_gpu():
    from _helpers.probe import Probe
    from oracle import oracle
    pr = Probe()
    rng = np.random.default_rng(99)
    n = 1 << 21
    a = (rng.standard_normal(n) * 10.0 ** rng.integers(-10, 8, n)).astype(np.float32)
    b = (rng.standard_normal(n) * 10.0 ** rng.integers(-10, 8, n)).astype(np.float32)
    b[::17] = 0; a[::19] = 0; a[5::1001] = np.inf
    c = rng.uniform(-200, 200, n).astype(np.float32)
    c[::13] *= 1e4
    for name, got, want in (("hypotf", oracle.hypotf(a, b), pr.hypotf(a, b)), ("atan2f", oracle.atan2f(a, b), pr.atan2f(a, b)),
                            ("cosf", oracle.cosf(c), pr.cosf(c))):
        same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
        assert same.all(), (name, int((~same).sum()), a[~same][:3], b[~same][:3], c[~same][:3])


# ---- semi-implicit solver on slabs: phases + gathered partial sums -----------------------------------
@pytest.mark.parametrize("bc", ["neumann", "periodic", "dirichlet"])
@pytest.mark.parametrize("dtype,S_aniso", [(PF_F64, 0.0), (PF_F64, 0.25), (PF_F32, 0.0)])
def test_semi_implicit_slabs_match_the_single_grid(bc, dtype, S_aniso):
    """Three slabs on one device, stepped in lock step through the pf_si_* phases with the all-gather and the one-row
    exchange of the search direction done through the host. Against the single-grid solver only the summation tree of the
    dot products differs: rel-L2 <= 1e-10 (fp64) / 1e-5 (fp32), CG iteration counts within +-1 (north_star)."""
    from bachelors_b200.slabs import (host_exchange_rows, make_host_gather, semi_implicit_buffers, semi_implicit_step_phases,
                                      slab_layout)
    nx, ny, G, steps = 150, 131, 3, 4
    npdt = np.float64 if dtype == PF_F64 else np.float32
    p = default_params(solver="semi-implicit", nx=nx, ny=ny, L0=nx / 128.0, S=S_aniso, Phi_boundary=bc, T_boundary=bc, gamma=0.75,
                       Phi_tolerance=1e-9 if dtype == PF_F64 else 1e-5, T_tolerance=1e-9 if dtype == PF_F64 else 1e-5)
    F0, U0 = seeded_fields(nx, ny, npdt, kind="interface")
    want_F, want_U, _, want_info = cuda_replay(p, F0, U0, dtype, steps)
    periodic = bc == "periodic"
    slabs = [slab_layout(ny, G, g, 1, periodic=periodic) for g in range(G)]
    solvers = [S.Solver(nx, ny, dtype, slab=sl) for sl in slabs]
    L = S.lib()
    try:
        def padded(a, sl, s):
            rows = np.arange(s.rows_alloc) - s.first_owned_row + sl.y0
            return a[rows % ny if periodic else np.clip(rows, 0, ny - 1)]
        st = [[s.field_from(padded(F0, sl, s)), s.field_from(padded(U0, sl, s)), s.field_alloc(), s.field_alloc()] for sl, s in zip(slabs, solvers)]
        bufs = semi_implicit_buffers(solvers)
        gather = make_host_gather(bufs[0]["width"])

        def exchange_direction(solvers_):
            for s in solvers_:
                s.sync()
            host_exchange_rows(solvers_, slabs, [L.pf_si_direction(s._h) for s in solvers_], ny, periodic)

        q = p.copy()
        got_info = []
        for i in range(steps):
            q.iter = i; q.time = i * p.dt
            iters, conv, errs = semi_implicit_step_phases(solvers, q, st, bufs, gather, exchange_direction, poll_every=1)
            got_info.append(iters)
            for f in st:
                f[0], f[2] = f[2], f[0]; f[1], f[3] = f[3], f[1]
            full_F = host_exchange_rows(solvers, slabs, [f[0] for f in st], ny, periodic)
            full_U = host_exchange_rows(solvers, slabs, [f[1] for f in st], ny, periodic)
        tol = tol_of(dtype)
        assert rel_l2(full_F, want_F) <= tol and rel_l2(full_U, want_U) <= tol
        for (gi_F, gi_T), w in zip(got_info, want_info):
            assert abs(gi_F - int(w[1])) <= 1 and abs(gi_T - int(w[2])) <= 1
        if dtype == PF_F64:
            assert max(max(i) for i in got_info) >= 2                # the solves really iterate
    finally:
        for s in solvers:
            s.close()
