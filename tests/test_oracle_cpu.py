"""Pins the CPU oracle (oracle/) without a GPU:
  * the reference's own CPU known-answer tests for conjugate gradients (reference cojugate_gradient.c:416-673),
  * golden outputs of the UNMODIFIED reference run on a B200 (tests/golden/*.npz, made by tools/make_golden_from_ref.py),
  * GPU libdevice outputs for the float functions the reference uses (tests/golden/libdevice_f32.npz),
  * the reference's published behavioural answers (adaptive step counts, benchmark_ellapsed_times.txt).
"""
import ctypes as C
import hashlib
import os
import subprocess

import numpy as np
import pytest

from bachelors_b200._abi import PF_F32, PF_F64, default_params, pf_step_info
from bachelors_b200.config import read_config
from bachelors_b200.initial import make_initial_conditions
from bachelors_b200.testing import seeded_fields
from conftest import GOLDEN, ROOT
from golden_cases import all_cases, case_inputs
from oracle import oracle
from oracle.oracle import Oracle, cg_generic

REF_CG_ITERS = [1, 3, 3, 2, 63, 63]      # printed by the reference's CPU program, in program order


# ---- conjugate gradients: the reference's CPU known answers ------------------------------------------------
def dense_op(A):
    A = np.asarray(A, dtype=np.float64)
    return lambda x: A @ x


def cross_op(C_, U, D, L, R, n, m):
    """cross_matrix_multiply with is_padded = false / zero padding (reference cojugate_gradient.c:291-327):
    the vector is treated as 1-D, neighbours i+-1 and i+-m, dropped when out of [0, N)."""
    N = n * m

    def mul(x):
        out = x * C_
        out[:-1] += x[1:] * R
        out[1:] += x[:-1] * L
        out[:-m] += x[m:] * U
        out[m:] += x[:-m] * D
        return out
    return mul, N


def test_cg_known_answers_match_reference_iteration_counts():
    got = []
    # dense 2x2 (:480-495)
    A = [[4, 1], [1, 3]]; b = np.array([1.0, 2.0])
    x, it, conv, _ = cg_generic(dense_op(A), 2, b, epsilon=1e-10, tolerance=1e-5, max_iters=10)
    assert conv and np.allclose(np.array(A) @ x, b, atol=1e-5)
    got.append(it)
    # dense 4x4 tridiagonal (:496-513)
    A = np.diag([4.0] * 4) + np.diag([1.0] * 3, 1) + np.diag([1.0] * 3, -1); b = np.array([1.0, 2, 3, 4])
    x, it, conv, _ = cg_generic(dense_op(A), 4, b, epsilon=1e-10, tolerance=1e-5, max_iters=10)
    assert conv and np.allclose(A @ x, b, atol=1e-4)
    got.append(it)
    # 2x2 cross, C=4 L=R=1 (:514-530)
    mul, N = cross_op(4, 0, 0, 1, 1, 2, 2)
    x, it, conv, _ = cg_generic(mul, N, np.array([1.0, 2, 3, 4]), epsilon=1e-10, tolerance=1e-5, max_iters=10)
    assert conv
    got.append(it)
    # padded 2x2 cross, all off-diagonals 1, b = x + y (:532-563)
    mul, N = cross_op(4, 1, 1, 1, 1, 2, 2)
    b = np.array([0.0, 1, 1, 2])
    x, it, conv, _ = cg_generic(mul, N, b, epsilon=1e-10, tolerance=1e-5, max_iters=10)
    assert conv and np.allclose(mul(x), b, atol=1e-4)
    got.append(it)
    # 32x32 five-point system diag 4 / off 1, b = x + y, tol 1e-5, max 100 (:565-614) and its dense twin (:616-655)
    n = m = 32
    mul, N = cross_op(4, 1, 1, 1, 1, n, m)
    yy, xx = np.mgrid[0:n, 0:m]
    b = (xx + yy).astype(np.float64).ravel()
    x, it, conv, _ = cg_generic(mul, N, b, epsilon=1e-10, tolerance=1e-5, max_iters=100)
    assert conv and np.allclose(mul(x), b, atol=1e-3)
    got.append(it)
    A = np.zeros((N, N))
    i = np.arange(N)
    A[i, i] = 4
    A[i[:-1], i[:-1] + 1] = 1; A[i[1:], i[1:] - 1] = 1
    A[i[:-m], i[:-m] + m] = 1; A[i[m:], i[m:] - m] = 1
    x2, it2, conv2, _ = cg_generic(dense_op(A), N, b, epsilon=1e-10, tolerance=1e-5, max_iters=100)
    assert conv2 and np.allclose(x, x2, atol=1e-6)
    got.append(it2)
    assert got == REF_CG_ITERS


def test_reference_cg_program_prints_the_pinned_counts():
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_cg_main")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/ref_cg_main not built (needs /root/reference)")
    out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout
    its = [int(l.split()[2]) for l in out.splitlines() if l.startswith("Solved with")]
    assert its == REF_CG_ITERS


def test_pde_operators_match_dense_matrices():
    """A_T = I - dt*Laplace and A_Phi (per-cell factor) of the oracle against explicitly assembled dense matrices."""
    nx, ny = 7, 5
    for bc in ("neumann", "dirichlet", "periodic"):
        p = default_params(solver="semi-implicit", nx=nx, ny=ny, L0=nx / 128.0, dt=1e-5, Phi_boundary=bc, T_boundary=bc)
        o = Oracle(PF_F64, nx, ny)
        rng = np.random.default_rng(3)
        s = rng.uniform(0.2, 0.5, (ny, nx))
        dx = p.L0 / nx; dy = p.L0 / ny
        for which in (0, 1):
            A = np.zeros((nx * ny, nx * ny))
            for y in range(ny):
                for x in range(nx):
                    i = x + y * nx
                    sc = s[y, x] if which == 0 else 1.0
                    A[i, i] += 1 + sc * (2 * p.dt / dx**2 + 2 * p.dt / dy**2)
                    for (xn, yn, w) in ((x + 1, y, p.dt / dx**2), (x - 1, y, p.dt / dx**2), (x, y + 1, p.dt / dy**2), (x, y - 1, p.dt / dy**2)):
                        sign = 1.0
                        if bc == "periodic":
                            xn %= nx; yn %= ny
                        elif not (0 <= xn < nx and 0 <= yn < ny):
                            xn = min(max(xn, 0), nx - 1); yn = min(max(yn, 0), ny - 1)
                            sign = -1.0 if bc == "dirichlet" else 1.0
                        A[i, xn + yn * nx] += -sc * w * sign
            # values of order 1e4 keep p.Ap far above the reference's epsilon = 1e-12 guard (simulation.cu:657)
            xt = 1e4 * rng.standard_normal((ny, nx))
            b = (A @ xt.ravel()).reshape(ny, nx)
            x, it, conv, err = o.cg_solve(p, which, s if which == 0 else None, b, None, 1e-5, 200)
            assert conv and it <= nx * ny + 5, (bc, which, it, err)
            assert np.allclose(x, xt, rtol=0, atol=1e-3), (bc, which, np.abs(x - xt).max())


# ---- libdevice float functions ------------------------------------------------------------------------------
def test_libdevice_restatement_matches_gpu_outputs():
    path = os.path.join(GOLDEN, "libdevice_f32.npz")
    assert os.path.exists(path), "tests/golden/libdevice_f32.npz missing: run tools/make_golden_from_ref.py on a GPU box"
    g = np.load(path)
    for name, got in (("hypot", oracle.hypotf(g["a"], g["b"])), ("atan2", oracle.atan2f(g["a"], g["b"])), ("cos", oracle.cosf(g["c"]))):
        want = g[name]
        same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
        assert same.all(), (name, int((~same).sum()), g["a"][~same][:4], g["b"][~same][:4], g["c"][~same][:4], got[~same][:4], want[~same][:4])


def test_libdevice_restatement_close_to_libm():
    rng = np.random.default_rng(1)
    a = rng.standard_normal(20000).astype(np.float32) * 50
    b = rng.standard_normal(20000).astype(np.float32) * 50
    assert np.allclose(oracle.hypotf(a, b), np.hypot(a.astype(np.float64), b.astype(np.float64)), rtol=3e-7)
    assert np.allclose(oracle.atan2f(a, b), np.arctan2(a.astype(np.float64), b.astype(np.float64)), rtol=0, atol=1e-6)
    assert np.allclose(oracle.cosf(a), np.cos(a.astype(np.float64)), rtol=0, atol=3e-7)
    big = np.array([1e6, -3e7, 123456.7, 1e30], dtype=np.float32)
    assert np.allclose(oracle.cosf(big), np.cos(big.astype(np.float64)), rtol=0, atol=3e-7)


# ---- golden outputs of the reference itself ---------------------------------------------------------------------
def oracle_replay(case):
    p, F0, U0, dtype = case_inputs(case)
    o = Oracle(dtype, case["nx"], case["ny"])
    F, U = F0, U0
    t = 0.0
    infos = []
    q = p.copy()
    for i in range(case["steps"]):
        q.iter = i; q.time = t
        F, U, dt, info = o.step(q, F, U)
        t += dt
        infos.append((dt, info.Phi_iters, info.T_iters, info.Phi_converged, info.T_converged, info.Phi_error, info.T_error))
    return F, U, t, np.array(infos, dtype=np.float64)


def rel_l2(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


@pytest.mark.parametrize("case", all_cases(), ids=lambda c: c["name"])
def test_oracle_matches_reference_golden(case):
    path = os.path.join(GOLDEN, case["name"] + ".npz")
    assert os.path.exists(path), f"{path} missing: run tools/make_golden_from_ref.py on a GPU box and commit its output"
    g = np.load(path)
    F, U, t, infos = oracle_replay(case)
    tol = 1e-10 if case["dtype"] == "f64" else 1e-5           # BASELINE.json north_star tolerances
    eF, eU = rel_l2(F, g["F"]), rel_l2(U, g["U"])
    assert eF <= tol and eU <= tol, (eF, eU)
    gi = g["step_info"]
    assert abs(t - float(g["t"])) <= 1e-12 * max(1.0, abs(t)) if case["dtype"] == "f64" else abs(t - float(g["t"])) <= 1e-6 * abs(t)
    if case["solver"] == "semi-implicit":
        # CG iteration counts within +-1 (north_star)
        assert np.all(np.abs(infos[:, 1] - gi[:, 1]) <= 1) and np.all(np.abs(infos[:, 2] - gi[:, 2]) <= 1), (infos[:, 1:3], gi[:, 1:3])
    if case["solver"] == "explicit-rk4-adaptive":
        assert np.array_equal(infos[:, 1], gi[:, 1]), (infos[:, 1], gi[:, 1])      # attempts per step
    if case["solver"] in ("explicit", "explicit-rk4") and not case["params"].get("do_corrector_loop"):
        # the oracle places its fused multiply-adds where nvcc does: bit-identical to the reference
        assert np.array_equal(F, g["F"]) and np.array_equal(U, g["U"]), (eF, eU)


# ---- behaviour pinned by the reference's published runs -----------------------------------------------------
def run_to_stop(cfg, dtype=PF_F64):
    p = cfg.params
    F, U = make_initial_conditions(cfg)
    o = Oracle(dtype, p.nx, p.ny)
    L = oracle.lib()
    cF, cU = F.copy(), U.copy()
    nF, nU = np.empty_like(cF), np.empty_like(cU)
    info = pf_step_info()
    q = p.copy()
    t, n = 0.0, 0
    while cfg.simul_stop_time - t >= 1e-16:                    # reference main.cpp:518
        q.iter = n; q.time = t
        t += L.orc_step(o._h, C.byref(q), cF.ctypes.data, cU.ctypes.data, nF.ctypes.data, nU.ctypes.data, C.byref(info))
        n += 1
        cF, nF = nF, cF; cU, nU = nU, cU
    return cF, cU, t, n


# config_explicit-rk4-adaptive_128.ini to t = 0.04, fp64: final time and SHA-256 of the final fields (8920 steps)
WHOLE_RUN_ADAPTIVE_128 = {"time": "0x1.47b1d1b9adf0ep-5",
                          "F": "d341587de9cfb2471f3c975183dd145d9106d16b0c258fd3100b5152929cbf1f",
                          "U": "eb09bd7b4b57df3c24466adfd9713710a34142384df23e36cee88ccd0cd829fd"}


# the fixed-step solvers' whole 128^2 runs (8000 steps): end state of the CPU oracle = CUDA path = unmodified reference
WHOLE_RUN_128 = {
    "explicit": {"time": "0x1.47ae147ae14f4p-5",
                 "F": "e5ddc9ca3c6649b27f6a52ced58ad3afbbc3d65a1388d84a11c1e77fb6547276",
                 "U": "ca99d7f13cfbfa47bb22153387b1f2b555ba257de937774a02d3c2f6fc665688"},
    "explicit-rk4": {"time": "0x1.47ae147ae14f4p-5",
                     "F": "068559de0edf098baf05311d0d74dc58a88cce16b26233170b7e472b8ad17be9",
                     "U": "1f43afdf361212caeee09b3636b6c7eeccdbb82ea6f7afab84e21ef9edb3a447"},
}


@pytest.mark.parametrize("solver", ["explicit", "explicit-rk4"])
def test_whole_fixed_step_run_128_end_state(solver):
    cfg = read_config(os.path.join(ROOT, "benchmark_configs", "config_%s_128.ini" % solver))
    F, U, t, n = run_to_stop(cfg)
    W = WHOLE_RUN_128[solver]
    assert n == 8000 and t.hex() == W["time"]
    assert hashlib.sha256(F.tobytes()).hexdigest() == W["F"] and hashlib.sha256(U.tobytes()).hexdigest() == W["U"]


def test_adaptive_step_count_128_matches_published():
    """benchmark_ellapsed_times.txt:6 (A100) 8920 iters, :151 (other GPU) 8922 iters at 128^2."""
    cfg = read_config(os.path.join(ROOT, "benchmark_configs", "config_explicit-rk4-adaptive_128.ini"))
    F, U, t, n = run_to_stop(cfg)
    assert n == 8920, n                                            # the A100 count; also the unmodified reference's on the B200 (tests/test_gpu_whole_runs.py)
    assert t >= 0.04 and t < 0.04 + 1e-5
    # the end state of the whole run, as the CUDA path and the unmodified reference produce it on the GPU (same file)
    assert t.hex() == WHOLE_RUN_ADAPTIVE_128["time"]
    assert hashlib.sha256(F.tobytes()).hexdigest() == WHOLE_RUN_ADAPTIVE_128["F"] and hashlib.sha256(U.tobytes()).hexdigest() == WHOLE_RUN_ADAPTIVE_128["U"]
    assert 0 < F.sum() < F.size and np.isfinite(F).all() and np.isfinite(U).all()
    # the dendrite grew: the seed disc of radius 0.05 covers ~129 cells of 128^2 (L0 = 1)
    assert F.sum() > 500


def test_fixed_dt_solvers_take_8000_steps():
    cfg = read_config(os.path.join(ROOT, "benchmark_configs", "config_explicit_128.ini"))
    n, t = 0, 0.0
    while cfg.simul_stop_time - t >= 1e-16:
        t += cfg.params.dt; n += 1
    assert n == 8000


# ---- properties ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("solver", ["explicit", "explicit-rk4", "explicit-rk4-adaptive", "semi-implicit"])
def test_results_do_not_depend_on_thread_count(solver):
    nx, ny = 40, 36
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.2)
    F0, U0 = seeded_fields(nx, ny, kind="interface")
    outs = []
    for th in (1, 3):
        oracle.lib().orc_set_threads(th)
        o = Oracle(PF_F64, nx, ny)
        outs.append(o.run(p, F0, U0, 3)[:2])
    oracle.lib().orc_set_threads(oracle.lib().orc_max_threads())
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("bc", ["neumann", "periodic"])
@pytest.mark.parametrize("solver", ["explicit", "explicit-rk4", "explicit-rk4-adaptive"])
def test_enthalpy_is_conserved(solver, bc):
    """dT/dt = Laplace T + L dPhi/dt  =>  sum(T - L*Phi) is invariant under zero-flux or periodic boundaries."""
    nx, ny = 48, 40
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.3, Phi_boundary=bc, T_boundary=bc)
    F0, U0 = seeded_fields(nx, ny, kind="interface")
    o = Oracle(PF_F64, nx, ny)
    F, U, _, _ = o.run(p, F0, U0, 20)
    h0 = (U0 - p.L * F0).sum(); h1 = (U - p.L * F).sum()
    assert abs(h1 - h0) <= 1e-9 * max(1.0, abs(h0)), (h0, h1)
    assert not np.array_equal(F, F0)


def test_symmetric_seed_stays_symmetric():
    cfg = read_config(os.path.join(ROOT, "benchmark_configs", "config_explicit-rk4_128.ini"))
    F0, U0 = make_initial_conditions(cfg)
    o = Oracle(PF_F64, cfg.params.nx, cfg.params.ny)
    F, U, _, _ = o.run(cfg.params, F0, U0, 50)
    # the stencil's summation order is not mirror-symmetric, so symmetry holds to rounding only
    for a in (F, U):
        for b in (a[::-1, :], a[:, ::-1], a.T):
            assert np.allclose(a, b, rtol=1e-9, atol=1e-13)
    assert F.sum() > 130
