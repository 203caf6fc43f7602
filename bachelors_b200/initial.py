"""Initial conditions on the host: mirror of sim_make_initial_conditions (reference main.cpp:93-136)."""
import numpy as np


def make_initial_conditions(cfg, dtype=np.float64, rows=None):
    """Returns (F, U) arrays of shape (ny, nx). `rows=(y_begin, y_end)` builds only a slab of global rows."""
    p = cfg.params
    nx, ny, L0 = p.nx, p.ny, p.L0
    y0, y1 = (0, ny) if rows is None else rows
    if p.do_exact:
        return _exact_initial_conditions(cfg, dtype, rows)
    xs = (np.arange(nx, dtype=np.float64) + 0.5) / nx * L0
    ys = (np.arange(y0, y1, dtype=np.float64) + 0.5) / ny * L0
    X, Y = np.meshgrid(xs, ys)
    # `Real lo, hi` in the reference (main.cpp:113-114): rounded to the field type before use
    lo = float(np.dtype(dtype).type(cfg.init_circle_radius - p.xi * cfg.init_circle_fade / 2))
    hi = float(np.dtype(dtype).type(cfg.init_circle_radius + p.xi * cfg.init_circle_fade / 2))
    r = np.hypot(cfg.init_circle_center[0] - X, cfg.init_circle_center[1] - Y)
    within = ((cfg.init_square_from[0] <= X) & (X < cfg.init_square_to[0]) &
              (cfg.init_square_from[1] <= Y) & (Y < cfg.init_square_to[1]))
    with np.errstate(divide="ignore", invalid="ignore"):
        circle = 1 - (r - lo) / (hi - lo)             # fade 0 -> +-inf (or nan exactly on the rim), as in the reference
    circle = np.where(circle > 1, 1.0, circle)
    circle = np.where(circle < 0, 0.0, circle)
    cube = np.where(within, 1.0, 0.0)
    factor = np.where(cube > circle, cube, circle)
    F = factor * cfg.init_inside_phi + (1 - factor) * cfg.init_outside_phi
    U = factor * cfg.init_inside_T + (1 - factor) * cfg.init_outside_T
    return F.astype(dtype), U.astype(dtype)


def _exact_initial_conditions(cfg, dtype, rows):
    """do_exact branch of sim_make_initial_conditions (main.cpp:104-110) with exact.h:17-96 restated in numpy, in the
    field type's arithmetic (Real): Phi = linear fade of width xi around R_ini, T = exact_u(0, r)."""
    from scipy.special import erf
    p = cfg.params
    R = np.dtype(dtype).type
    nx, ny, L0 = p.nx, p.ny, p.L0
    y0, y1 = (0, ny) if rows is None else rows
    xs = (np.arange(nx, dtype=np.float64) + 0.5) / nx * L0
    ys = (np.arange(y0, y1, dtype=np.float64) + 0.5) / ny * L0
    X, Y = np.meshgrid(xs, ys)
    r = np.hypot(X - L0 / 2, Y - L0 / 2).astype(dtype)                      # `Real r = hypot(double, double)`
    lam, R_ini, eps, one, two = R(0.5), R(0.25), R(0.001), R(1), R(2)
    xi = R(p.xi)
    lo, hi = R_ini - one * xi / two, R_ini + one * xi / two                  # exact.h:73-75 (fade = 1)
    with np.errstate(divide="ignore", invalid="ignore"):
        F = np.where(r < lo, one, np.where(r > hi, R(0), one - (r - lo) / (hi - lo))).astype(dtype)
    Rt = np.sqrt(R_ini * R_ini + two * lam * R(0))                           # exact_R(0), exact.h:17-20
    Ut = -eps * (lam + two) / Rt                                             # exact.h:22-25
    s = (r / Rt).astype(dtype)
    pi = R(3.14159265359)
    sqrtl2 = np.sqrt(lam / two)
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):      # exact_T, exact.h:27-35
        integral = np.exp(-lam / two) - one / s * np.exp(-lam / two * s * s) + sqrtl2 * pi * (erf(sqrtl2) - erf(s * sqrtl2))
        T = (-lam * np.exp(lam / two) * integral).astype(dtype)
    U = np.where(r > Rt, Ut + T, Ut).astype(dtype)
    return np.ascontiguousarray(F), np.ascontiguousarray(U)
