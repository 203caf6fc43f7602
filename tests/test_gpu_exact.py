"""The analytic ("exact") mode, SURVEY section 8f-5: reference exact.h:17-96, simulation.cu:179-184 (source term fu),
:1069-1089 (solver = exact), main.cpp:104-110 (initial profile), config.h:492-509 (parameter overrides).

Parity: the explicit solvers with do_exact go through the per-stage kernels and are bit-identical to the CPU oracle and
to the unmodified reference on the same GPU (fu is a host-computed constant of the step); the analytic fields go through
exp / erf / hypot of two different math libraries: rel-L2 <= 1e-10 (fp64) / 1e-5 (fp32)."""
import os

import numpy as np
import pytest

from bachelors_b200 import solver as S
from bachelors_b200._abi import PF_F32, PF_F64, SOLVER_NAMES
from bachelors_b200.app import initial_of
from bachelors_b200.config import read_config
from bachelors_b200.initial import make_initial_conditions
from conftest import ROOT
from oracle import ref_sim
from oracle.oracle import Oracle

pytestmark = pytest.mark.gpu
NP = {PF_F64: np.float64, PF_F32: np.float32}


def exact_cfg(solver, size=128):
    return read_config(os.path.join(ROOT, "benchmark_configs", "config_explicit_%d.ini" % size),
                       ["[simulation]\ndo_exact = true\nsolver = %s\nT_tolerance = 1e-7\nPhi_tolerance = 1e-7" % solver])


def rel_l2(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def test_config_overrides_of_the_analytic_case():
    cfg = exact_cfg("explicit")
    p = cfg.params
    h = p.L0 / p.nx
    assert (p.Tm, p.L, p.a, p.b, p.alpha, p.S) == (0, 1, 1, 1, 1, 0) and p.beta == 1 / 0.001
    assert p.dt == (1.0 / 16) / 4 * h * h and p.xi == p.L0 / p.nx * 11 / 10 and cfg.init_circle_radius == 0.25
    assert exact_cfg("exact").params.dt == 0.000005                  # solver = exact keeps the configured dt (config.h:500-501)


def cuda_steps(p, F0, U0, dtype, steps):
    with S.Solver(p.nx, p.ny, dtype) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        q, t, dts = p.copy(), 0.0, []
        for i in range(steps):
            q.iter, q.time = i + 7, t                                # fu = exact_fu(iter*dt): a different constant every step
            info = s.step(q, F, U, nF, nU)
            t += info.dt_advanced; dts.append(info.dt_advanced)
            F, nF, U, nU = nF, F, nU, U
        return s.download(F), s.download(U), dts


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
@pytest.mark.parametrize("solver", ["explicit", "explicit-rk4", "explicit-rk4-adaptive"])
def test_source_term_steps_are_bit_identical_to_oracle_and_reference(solver, dtype):
    cfg = exact_cfg(solver)
    p = cfg.params
    F0, U0 = make_initial_conditions(cfg, NP[dtype])
    steps = 6
    gF, gU, gdts = cuda_steps(p, F0, U0, dtype, steps)
    o = Oracle(dtype, p.nx, p.ny)
    F, U, q, t, odts = F0, U0, p.copy(), 0.0, []
    for i in range(steps):
        q.iter, q.time = i + 7, t
        F, U, dt, _ = o.step(q, F, U)
        t += dt; odts.append(dt)
    assert gdts == odts
    assert np.array_equal(gF, F) and np.array_equal(gU, U)
    q = p.copy(); q.do_exact = 0
    nF, nU, _ = cuda_steps(q, F0, U0, dtype, steps)
    assert not np.array_equal(nU, gU)                                # the source term is really there
    if ref_sim.available(dtype) and solver != "explicit-rk4-adaptive":
        with ref_sim.RefSim(dtype, p.nx, p.ny, fresh=True) as ref:
            ref.upload(F0, U0)
            q, t = p.copy(), 0.0
            for i in range(steps):
                q.iter, q.time = i + 7, t
                dt, _ = ref.run(q, 1)
                t += dt
            rF, rU = ref.download()
        assert np.array_equal(gF, rF) and np.array_equal(gU, rU)


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
def test_exact_solver_writes_the_analytic_fields(dtype):
    cfg = exact_cfg("exact")
    p = cfg.params
    assert p.solver == SOLVER_NAMES["exact"]
    F0, U0 = make_initial_conditions(cfg, NP[dtype])
    tol = 1e-10 if dtype == PF_F64 else 1e-5
    o = Oracle(dtype, p.nx, p.ny)
    with S.Solver(p.nx, p.ny, dtype) as s:
        F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        for t in (0.0, 0.003, 0.05):
            q = p.copy(); q.time = t
            info = s.step(q, F, U, nF, nU)
            assert info.dt_advanced == p.dt
            gF, gU = s.download(nF), s.download(nU)
            wF, wU, _, _ = o.step(q, F0, U0)
            assert np.array_equal(gF, wF) or (gF != wF).sum() <= 8       # the 0/1 disc; a rim cell may sit within an ulp of R(t)
            assert rel_l2(gU, wU) <= tol
            assert gF.sum() > 0 and np.isfinite(gU).all()
        if ref_sim.available(dtype):
            with ref_sim.RefSim(dtype, p.nx, p.ny, fresh=True) as ref:
                ref.upload(F0, U0)
                ref.run(q, 1)
                rF, rU = ref.download()
            assert (gF != rF).sum() <= 8 and rel_l2(gU, rU) <= tol


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
def test_exact_initial_conditions_on_the_device(dtype):
    cfg = exact_cfg("explicit", 256)
    p = cfg.params
    hF, hU = make_initial_conditions(cfg, NP[dtype])
    with S.Solver(p.nx, p.ny, dtype) as s:
        F, U = s.field_alloc(), s.field_alloc()
        s.init_fields(p, initial_of(cfg), F, U)
        dF, dU = s.download(F), s.download(U)
    tol = 1e-12 if dtype == PF_F64 else 1e-5
    assert np.abs(dF - hF).max() <= tol * 10 and rel_l2(dU, hU) <= tol
    assert 0 < ((hF > 0) & (hF < 1)).sum()                           # the xi-wide fade exists
