"""ctypes mirrors of the plain-C structs in include/pf_solver.h (keep in sync with PF_ABI_VERSION)."""
import ctypes as C

PF_ABI_VERSION = 2

PF_OK, PF_ERR_INVALID, PF_ERR_NO_DEVICE, PF_ERR_CUDA, PF_ERR_OOM, PF_ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5

PF_F32, PF_F64 = 0, 1

# pf_option
PF_OPT_HOST_BANDS, PF_OPT_SI_PER_KERNEL, PF_OPT_SI_GRID_SYNC, PF_OPT_HALO_IN_KERNEL, PF_OPT_RUN_GRAPHS, PF_OPT_SI_BLOCKS_PER_SM = 0, 1, 2, 3, 4, 5

# Sim_Boundary_Type, reference simulation.h:27-32
BOUNDARY_PERIODIC, BOUNDARY_DIRICHLET_ZERO, BOUNDARY_NEUMANN_ZERO = 0, 1, 2
BOUNDARY_NAMES = {"periodic": 0, "dirichlet": 1, "neumann": 2}      # simulation.h:147-156

# Sim_Solver_Type, reference simulation.h:34-42
SOLVER_TYPE_NONE, SOLVER_TYPE_EXPLICIT_EULER, SOLVER_TYPE_EXPLICIT_RK4, SOLVER_TYPE_EXPLICIT_RK4_ADAPTIVE, \
    SOLVER_TYPE_SEMI_IMPLICIT, SOLVER_TYPE_EXACT = range(6)
SOLVER_NAMES = {                                                        # simulation.h:158-170
    "none": 0, "explicit": 1, "explicit-rk4": 2, "explicit-rk4-adaptive": 3, "semi-implicit": 4, "exact": 5,
}
SOLVER_NAME_OF = {v: k for k, v in SOLVER_NAMES.items()}

PF_MAX_STEP_RESIDUALS = 20


class pf_params(C.Structure):
    _fields_ = [
        ("solver", C.c_int32), ("nx", C.c_int32), ("ny", C.c_int32),
        ("time", C.c_double), ("iter", C.c_int64),
        ("T_boundary", C.c_int32), ("Phi_boundary", C.c_int32),
        ("L0", C.c_double), ("dt", C.c_double), ("L", C.c_double), ("xi", C.c_double),
        ("a", C.c_double), ("b", C.c_double), ("alpha", C.c_double), ("beta", C.c_double),
        ("gamma", C.c_double), ("Tm", C.c_double), ("min_dt", C.c_double),
        ("S", C.c_double), ("m0", C.c_double), ("theta0", C.c_double),
        ("T_tolerance", C.c_double), ("Phi_tolerance", C.c_double), ("corrector_tolerance", C.c_double),
        ("T_max_iters", C.c_int32), ("Phi_max_iters", C.c_int32), ("corrector_max_iters", C.c_int32),
        ("do_corrector_loop", C.c_int32), ("do_corrector_guess", C.c_int32),
        ("do_stats", C.c_int32), ("do_stats_step_residual", C.c_int32), ("do_exact", C.c_int32),
    ]

    def copy(self):
        out = pf_params()
        C.memmove(C.byref(out), C.byref(self), C.sizeof(pf_params))
        return out


class pf_delta_stats(C.Structure):
    _fields_ = [("sum", C.c_double), ("L1", C.c_double), ("L2", C.c_double), ("min", C.c_double), ("max", C.c_double)]


class pf_step_info(C.Structure):
    _fields_ = [
        ("dt_advanced", C.c_double),
        ("Phi_iters", C.c_int32), ("T_iters", C.c_int32),
        ("Phi_converged", C.c_int32), ("T_converged", C.c_int32),
        ("Phi_error", C.c_double), ("T_error", C.c_double),
        ("next_tau", C.c_double),
        ("Phi_delta", pf_delta_stats), ("T_delta", pf_delta_stats),
        ("step_res_count", C.c_int32),
        ("step_res", pf_delta_stats * PF_MAX_STEP_RESIDUALS),
    ]


class pf_slab(C.Structure):
    _fields_ = [("y0", C.c_int32), ("ny_local", C.c_int32), ("has_lo", C.c_int32), ("has_hi", C.c_int32), ("ghost", C.c_int32)]


class pf_initial(C.Structure):
    """[initial] section of config.ini (reference config.h:20-30)."""
    _fields_ = [("inside_phi", C.c_double), ("inside_T", C.c_double), ("outside_phi", C.c_double), ("outside_T", C.c_double),
                ("circle_center", C.c_double * 2), ("circle_radius", C.c_double), ("circle_fade", C.c_double),
                ("square_from", C.c_double * 2), ("square_to", C.c_double * 2)]


class pf_stats_row(C.Structure):
    """One line of stats.csv (reference main.cpp:192-234)."""
    _fields_ = [("time", C.c_double), ("iter", C.c_int64), ("Phi_iters", C.c_int64), ("T_iters", C.c_int64),
                ("T_delta_L1", C.c_float), ("T_delta_L2", C.c_float), ("T_delta_max", C.c_float), ("T_delta_min", C.c_float),
                ("Phi_delta_L1", C.c_float), ("Phi_delta_L2", C.c_float), ("Phi_delta_max", C.c_float), ("Phi_delta_min", C.c_float),
                ("step_res_L1", C.c_float * PF_MAX_STEP_RESIDUALS), ("step_res_L2", C.c_float * PF_MAX_STEP_RESIDUALS),
                ("step_res_max", C.c_float * PF_MAX_STEP_RESIDUALS), ("step_res_min", C.c_float * PF_MAX_STEP_RESIDUALS)]


PF_SNAPSHOT_SLOTS, PF_SNAPSHOT_MAX_FIELDS = 2, 8


def default_params(**kw):
    """Parameters of the reference's benchmark family (benchmark_configs/config_*_512.ini at scale 1)."""
    p = pf_params()
    p.solver = SOLVER_TYPE_EXPLICIT_EULER
    p.nx = p.ny = 512
    p.T_boundary = p.Phi_boundary = BOUNDARY_NEUMANN_ZERO
    p.L0 = 4.0; p.dt = 0.000005; p.L = 2.0; p.xi = 0.0043; p.a = 2.0; p.b = 1.0; p.alpha = 3.0; p.beta = 1400.0
    p.gamma = 1.0; p.Tm = 1.0; p.min_dt = 0.0; p.S = 0.0; p.m0 = 6.0; p.theta0 = 0.0
    p.T_tolerance = p.Phi_tolerance = 5e-9; p.corrector_tolerance = 0.0
    p.T_max_iters = p.Phi_max_iters = 20; p.corrector_max_iters = 3
    for k, v in kw.items():
        if isinstance(v, str) and k == "solver":
            v = SOLVER_NAMES[v]
        if isinstance(v, str) and k.endswith("boundary"):
            v = BOUNDARY_NAMES[v]
        setattr(p, k, v)
    return p
