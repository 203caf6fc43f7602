"""Reader for the reference's config.ini / benchmark_configs/*.ini.

Host-side mirror of `allen_cahn_read_config` (reference config.h:396-519): same sections, same keys,
same optional/mandatory split, same `scale` and parsing rules (reference config.h:143-197: `key = value`
or `key : value`, `;`/`#` comments also after a value, `[section]` headers, later duplicates win).
"""
import re
from dataclasses import dataclass, field

from ._abi import pf_params, SOLVER_NAMES, BOUNDARY_NAMES

_WS = " \t\n\r\v\f"


class ConfigError(ValueError):
    pass


def parse_ini(text):
    """-> {(section, key): value}; mirrors match_key_value_ini_pair / key_value_ini_append (config.h:143-220)."""
    out = {}
    section = ""
    for line in text.split("\n"):
        s = line.strip(_WS)
        if not s or s[0] in "#;":
            continue
        # key = up to first whitespace or marker
        m = re.match(r"([^ \t\r\v\f=:]+)[ \t\r\v\f]*([=:]*)[ \t\r\v\f]*(.*)$", s)
        if not m:
            continue
        key, marker, value = m.group(1), m.group(2), m.group(3)
        if key[0] == "[" and key[-1] == "]":
            section = key[1:-1]
            continue
        if marker not in ("=", ":"):
            continue                                   # KEY_VALUE_ERROR: logged and ignored by the reference
        cut = len(value)
        for ch in ";#":
            i = value.find(ch)
            if i != -1:
                cut = min(cut, i)
        out[(section, key)] = value[:cut].rstrip(_WS)
    return out


def _scan_double(s):
    m = re.match(r"\s*([+-]?(?:\d+\.?\d*(?:[eE][+-]?\d+)?|\.\d+(?:[eE][+-]?\d+)?|inf|nan))", s, re.I)   # sscanf("%lf")
    if not m:
        raise ValueError(s)
    return float(m.group(1))


def _scan_int(s):
    m = re.match(r"\s*([+-]?(?:0[xX][0-9a-fA-F]+|\d+))", s)                                               # sscanf("%i")
    if not m:
        raise ValueError(s)
    return int(m.group(1), 0) if not re.match(r"\s*[+-]?0\d", s) else int(m.group(1), 8)


def _scan_bool(s):
    if s in ("true", "1"):
        return True
    if s in ("false", "0"):
        return False
    raise ValueError(s)


def _scan_vec2(s):
    parts = s.split()
    if len(parts) < 2:
        raise ValueError(s)
    return (_scan_double(parts[0]), _scan_double(parts[1]))


@dataclass
class SimConfig:
    """Sim_Config of reference config.h:10-57 (params is the C-ABI pf_params instead of Sim_Params)."""
    entire_config_file: str = ""
    params: pf_params = field(default_factory=pf_params)
    scale: float = 1.0
    simul_stop_time: float = 0.0
    init_inside_phi: float = 0.0
    init_inside_T: float = 0.0
    init_outside_phi: float = 0.0
    init_outside_T: float = 0.0
    init_circle_center: tuple = (0.0, 0.0)
    init_circle_radius: float = 0.0
    init_circle_fade: float = 0.0
    init_square_from: tuple = (0.0, 0.0)
    init_square_to: tuple = (0.0, 0.0)
    snapshot_every: float = 0.0
    snapshot_times: int = 0
    snapshot_initial_conditions: bool = False
    snapshot_folder: str = ""
    snapshot_prefix: str = ""
    snapshot_postfix: str = ""
    app_run_simulation: bool = True
    app_run_tests: bool = False
    app_run_benchmarks: bool = False
    app_collect_stats: bool = False
    app_collect_step_residuals: bool = False
    app_print_in_noninteractive: bool = False
    app_collect_stats_every: float = 0.0
    app_interactive_mode: bool = False
    app_linear_filtering: bool = False
    app_display_min: float = 0.0
    app_display_max: float = 1.0


def read_config_text(text, overrides=()):
    kv = parse_ini(text)
    for o in overrides:
        kv.update(parse_ini(o))
    cfg = SimConfig(entire_config_file=text)
    p = cfg.params
    missing = []

    def get(section, key, conv, optional=False):
        v = kv.get((section, key))
        if v is None:
            if not optional:
                missing.append(f"[{section}] {key}")
            return None
        try:
            return conv(v)
        except ValueError:
            missing.append(f"[{section}] {key} = {v!r}")
            return None

    def setp(attr, section, key, conv, optional=False):
        v = get(section, key, conv, optional)
        if v is not None:
            setattr(p, attr, v)

    def setc(attr, section, key, conv, optional=False):
        v = get(section, key, conv, optional)
        if v is not None:
            setattr(cfg, attr, v)

    def solver(s):
        if s not in SOLVER_NAMES or s == "none":                 # "exact" is accepted (simulation.h:158-170)
            raise ValueError(s)
        return SOLVER_NAMES[s]

    def boundary(s):
        if s not in BOUNDARY_NAMES:
            raise ValueError(s)
        return BOUNDARY_NAMES[s]

    S = "simulation"                                            # config.h:413-444
    for attr, key in (("dt", "dt"), ("L0", "L0"), ("L", "L"), ("xi", "xi"), ("a", "a"), ("b", "b"), ("alpha", "alpha"),
                      ("beta", "beta"), ("Tm", "Tm"), ("S", "S"), ("m0", "m"), ("theta0", "theta0"), ("gamma", "gamma"),
                      ("T_tolerance", "T_tolerance"), ("Phi_tolerance", "Phi_tolerance"),
                      ("corrector_tolerance", "corrector_tolerance")):
        setp(attr, S, key, _scan_double)
    setp("do_exact", S, "do_exact", _scan_bool)
    setp("solver", S, "solver", solver)
    setp("Phi_boundary", S, "Phi_boundary", boundary)
    setp("T_boundary", S, "T_boundary", boundary)
    setc("simul_stop_time", S, "stop_after", _scan_double)
    setp("nx", S, "mesh_size_x", _scan_int)
    setp("ny", S, "mesh_size_y", _scan_int)
    for attr in ("T_max_iters", "Phi_max_iters", "corrector_max_iters"):
        setp(attr, S, attr, _scan_int)
    setp("do_corrector_loop", S, "do_corrector_loop", _scan_bool)
    setp("do_corrector_guess", S, "do_corrector_guess", _scan_bool)
    setp("min_dt", S, "min_dt", _scan_double, optional=True)

    I = "initial"                                               # config.h:447-456
    setc("init_inside_phi", I, "inside_phi", _scan_double)
    setc("init_inside_T", I, "inside_T", _scan_double)
    setc("init_outside_phi", I, "outside_phi", _scan_double)
    setc("init_outside_T", I, "outside_T", _scan_double)
    setc("init_circle_center", I, "circle_center", _scan_vec2)
    setc("init_circle_radius", I, "circle_radius", _scan_double)
    setc("init_circle_fade", I, "circle_fade", _scan_double)
    setc("init_square_from", I, "square_from", _scan_vec2)
    setc("init_square_to", I, "square_to", _scan_vec2)

    N = "snapshot"                                              # config.h:458-465
    setc("snapshot_every", N, "every", _scan_double)
    setc("snapshot_times", N, "times", _scan_int)
    setc("snapshot_initial_conditions", N, "snapshot_initial_conditions", _scan_bool)
    setc("snapshot_folder", N, "folder", str)
    setc("snapshot_prefix", N, "prefix", str)
    setc("snapshot_postfix", N, "postfix", str)

    P = "program"                                               # config.h:467-479
    setc("app_run_simulation", P, "run_simulation", _scan_bool)
    setc("app_run_tests", P, "run_tests", _scan_bool)
    setc("app_run_benchmarks", P, "run_benchmarks", _scan_bool)
    setc("app_interactive_mode", P, "interactive", _scan_bool)
    setc("app_print_in_noninteractive", P, "print_in_noninteractive", _scan_bool)
    setc("app_linear_filtering", P, "linear_filtering", _scan_bool)
    setc("app_collect_stats", P, "collect_stats", _scan_bool)
    setc("app_collect_step_residuals", P, "collect_step_residual", _scan_bool)
    setc("app_collect_stats_every", P, "collect_stats_every", _scan_double)
    setc("app_display_min", P, "display_min", _scan_double)
    setc("app_display_max", P, "display_max", _scan_double)

    scale = get(S, "scale", _scan_double, optional=True)        # config.h:481-490
    if scale is not None:
        cfg.scale = scale
        p.nx = int(p.nx * scale)
        p.ny = int(p.ny * scale)
        p.L0 *= scale
        cfg.init_circle_center = (cfg.init_circle_center[0] * scale, cfg.init_circle_center[1] * scale)
    else:
        cfg.scale = 1.0

    if p.do_exact:                                              # config.h:492-509: the analytic test case overrides the physics
        A = 1.0 / 16
        h = max(p.L0 / p.nx, p.L0 / p.ny)
        p.Tm = 0
        p.L = 1
        if p.solver != SOLVER_NAMES["exact"]:
            p.dt = A / 4 * h * h
        p.a = 1
        p.b = 1
        p.alpha = 1
        p.beta = 1 / 0.001
        p.S = 0
        p.xi = p.L0 / p.nx * 11 / 10
        cfg.init_circle_radius = 0.25
    if missing:
        raise ConfigError("couldnt find or parse some config entries: " + ", ".join(missing))
    return cfg


def read_config(path, overrides=()):
    with open(path, "r") as f:
        return read_config_text(f.read(), overrides)
