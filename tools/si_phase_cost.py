"""Per-iteration and set-up cost of the cooperative semi-implicit step: steps with a tolerance of 0 and a fixed number of CG
iterations k per solve (so every solve runs exactly k iterations), timed with CUDA events on the solver's stream, with the
flagged all-block sums (default) and with grid.sync() (PF_OPT_SI_GRID_SYNC). Slope of time over k = cost of one iteration of
both solves (4 phases, 4 sums); intercept = launch + prepare + the set-up phases.   python tools/si_phase_cost.py [size ...]"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from bachelors_b200 import solver as S
from bachelors_b200._abi import PF_F64, PF_OPT_SI_GRID_SYNC, default_params
from bachelors_b200.testing import seeded_fields


def time_steps(s, p, F, U, nF, nU, n):
    st = torch.cuda.ExternalStream(s.stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(10):
        s.step_async(p, F, U, nF, nU)
    s.sync()
    e0.record(st)
    for _ in range(n):
        s.step_async(p, F, U, nF, nU)          # same input every time: the cost does not depend on the values
    e1.record(st)
    s.sync()
    return e0.elapsed_time(e1) / n * 1e3


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [256, 512, 1024, 2048]
    out = {}
    for n in sizes:
        F0, U0 = seeded_fields(n, n, np.float64, kind="interface")
        for name, opts in (("flagged", []), ("grid_sync", [PF_OPT_SI_GRID_SYNC])):
            with S.Solver(n, n, PF_F64) as s:
                for o in opts:
                    s.set_option(o, 1)
                F, U, nF, nU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
                row = []
                for k in (0, 1, 2, 3, 4, 6, 8):
                    p = default_params(solver="semi-implicit", nx=n, ny=n, L0=n / 128.0, Phi_tolerance=0.0, T_tolerance=0.0,
                                       Phi_max_iters=k, T_max_iters=k)
                    row.append((k, round(time_steps(s, p, F, U, nF, nU, 200), 2)))
                slope = (row[-1][1] - row[2][1]) / (row[-1][0] - row[2][0])
                out[f"{n}_{name}"] = {"us_by_iterations": row, "us_per_iteration_of_both_solves": round(slope, 2)}
                print(n, name, row, "slope", round(slope, 2), flush=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
