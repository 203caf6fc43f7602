"""The parity cases shared by tools/make_golden_from_ref.py (which runs the UNMODIFIED reference on a B200 and
stores its outputs under tests/golden/) and by the tests that replay them through the CPU oracle and the CUDA path.

A case is fully described by plain data: inputs are regenerated from the seed, only outputs are stored.
"""
import numpy as np

from bachelors_b200._abi import PF_F32, PF_F64, default_params
from bachelors_b200.testing import seeded_fields


def _case(name, solver, nx, ny, steps, dtype="f64", kind="interface", **params):
    return dict(name=name, solver=solver, nx=nx, ny=ny, steps=steps, dtype=dtype, kind=kind, params=params)


def all_cases():
    cases = []
    for dt_name in ("f64", "f32"):
        sfx = "_" + dt_name
        # explicit Euler: all boundary types, anisotropy on/off, corrector guess on/off, ragged sizes
        cases += [
            _case("euler_neumann" + sfx, "explicit", 72, 56, 5, dt_name),
            _case("euler_dirichlet_aniso" + sfx, "explicit", 67, 45, 5, dt_name, S=0.3, Phi_boundary="dirichlet", T_boundary="dirichlet"),
            _case("euler_periodic_guess" + sfx, "explicit", 64, 64, 5, dt_name, kind="noise", do_corrector_guess=1,
                  Phi_boundary="periodic", T_boundary="periodic"),
            _case("euler_mixed_bc" + sfx, "explicit", 50, 70, 4, dt_name, S=0.2, Phi_boundary="periodic", T_boundary="neumann"),
            _case("euler_corrector_loop" + sfx, "explicit", 40, 40, 3, dt_name, do_corrector_loop=1, corrector_max_iters=3, do_corrector_guess=1),
            # classic RK4
            _case("rk4_neumann" + sfx, "explicit-rk4", 72, 56, 3, dt_name),
            _case("rk4_neumann_aniso_noise" + sfx, "explicit-rk4", 61, 83, 3, dt_name, kind="noise", S=0.3),
            _case("rk4_periodic" + sfx, "explicit-rk4", 96, 40, 3, dt_name, S=0.1, Phi_boundary="periodic", T_boundary="periodic"),
            _case("rk4_dirichlet_guess" + sfx, "explicit-rk4", 33, 47, 3, dt_name, do_corrector_guess=1,
                  Phi_boundary="dirichlet", T_boundary="dirichlet"),
            _case("rk4_mixed_bc" + sfx, "explicit-rk4", 48, 48, 2, dt_name, Phi_boundary="neumann", T_boundary="dirichlet"),
            # Runge-Kutta-Merson adaptive (tolerances chosen so that the first attempts are rejected)
            _case("rkm_neumann" + sfx, "explicit-rk4-adaptive", 72, 56, 4, dt_name,
                  Phi_tolerance=5e-9 if dt_name == "f64" else 5e-6, T_tolerance=5e-9 if dt_name == "f64" else 5e-6),
            _case("rkm_periodic_aniso" + sfx, "explicit-rk4-adaptive", 64, 48, 3, dt_name, S=0.25, Phi_boundary="periodic", T_boundary="periodic",
                  Phi_tolerance=1e-7 if dt_name == "f64" else 1e-5, T_tolerance=1e-7 if dt_name == "f64" else 1e-5, dt=2e-5),
            # semi-implicit (CG iteration counts are part of the golden record)
            _case("si_neumann" + sfx, "semi-implicit", 64, 64, 3, dt_name,
                  Phi_tolerance=5e-9 if dt_name == "f64" else 1e-5, T_tolerance=5e-9 if dt_name == "f64" else 1e-5),
            _case("si_theta_half_guess" + sfx, "semi-implicit", 56, 72, 2, dt_name, gamma=0.5, do_corrector_guess=1, S=0.2,
                  Phi_tolerance=5e-9 if dt_name == "f64" else 1e-5, T_tolerance=5e-9 if dt_name == "f64" else 1e-5,
                  Phi_max_iters=40, T_max_iters=40),
            _case("si_periodic_maxiters" + sfx, "semi-implicit", 48, 48, 2, dt_name, Phi_boundary="periodic", T_boundary="periodic",
                  Phi_tolerance=1e-14 if dt_name == "f64" else 1e-9, T_tolerance=1e-14 if dt_name == "f64" else 1e-9,
                  Phi_max_iters=6, T_max_iters=6),
        ]
    return cases


def case_by_name(name):
    for c in all_cases():
        if c["name"] == name:
            return c
    raise KeyError(name)


def case_inputs(case):
    """-> (pf_params, F0, U0, pf dtype)."""
    dtype = PF_F64 if case["dtype"] == "f64" else PF_F32
    npdt = np.float64 if dtype == PF_F64 else np.float32
    nx, ny = case["nx"], case["ny"]
    # the benchmark family keeps dx = 4/512 (reference config.h:481-488): L0 scales with nx
    kw = dict(solver=case["solver"], nx=nx, ny=ny, L0=nx / 128.0)
    kw.update(case["params"])
    p = default_params(**kw)
    F0, U0 = seeded_fields(nx, ny, npdt, kind=case["kind"])
    return p, F0, U0, dtype
