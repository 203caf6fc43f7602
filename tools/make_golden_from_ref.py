#!/usr/bin/env python
"""Runs the UNMODIFIED reference solver (oracle/_ref/libref_sim_*.so, "oracle A") on a B200 for every case of
tests/golden_cases.py and stores its outputs as small fixtures: <out>/<case>.npz (Phi, T after `steps` steps, the
time advanced, per-step info). Run on the GPU box:   python tools/make_golden_from_ref.py gpurun_out/golden
then copy gpurun_out/golden/*.npz into tests/golden/ and commit them. Also stores libdevice_f32.npz: GPU outputs of
hypotf / atan2f / cosf (through tests/_helpers/libprobe.so) on seeded inputs, which pin oracle/libdevice_f32.h.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from golden_cases import all_cases, case_inputs          # noqa: E402
from oracle.ref_sim import RefSim                         # noqa: E402


def libdevice_inputs():
    rng = np.random.default_rng(20260411)
    n = 1 << 16
    a = np.concatenate([rng.standard_normal(n) * 10.0 ** rng.integers(-8, 6, n), [0.0, -0.0, 1.0, -1.0, 1e-30, 3e38, np.inf]]).astype(np.float32)
    b = np.concatenate([rng.standard_normal(n) * 10.0 ** rng.integers(-8, 6, n), [0.0, 0.0, -0.0, 2.0, 1e-38, 3e38, 1.0]]).astype(np.float32)
    c = np.concatenate([rng.uniform(-40, 40, n), [0.0, -0.0, 1e5, 2e5, -3e7, 1e30, 105615.0]]).astype(np.float32)
    return a, b, c


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    for case in all_cases():
        p, F0, U0, dtype = case_inputs(case)
        # a private copy of the library per case: the reference keeps tau / corrector parity in function statics
        with RefSim(dtype, case["nx"], case["ny"], fresh=True) as ref:
            ref.upload(F0, U0)
            infos = []
            t = 0.0
            q = p.copy()
            for i in range(case["steps"]):
                q.iter = i; q.time = t
                dt, info = ref.run(q, 1)
                t += dt
                infos.append((dt, info.Phi_iters, info.T_iters, info.Phi_converged, info.T_converged, info.Phi_error, info.T_error))
            F, U = ref.download()
        np.savez_compressed(os.path.join(out_dir, case["name"] + ".npz"), F=F, U=U, t=np.float64(t),
                            step_info=np.array(infos, dtype=np.float64))
        print(f"{case['name']:36s} t={t:.9e} last info={infos[-1]}", flush=True)

    from _helpers.probe import Probe
    a, b, c = libdevice_inputs()
    pr = Probe()
    np.savez_compressed(os.path.join(out_dir, "libdevice_f32.npz"), a=a, b=b, c=c,
                        hypot=pr.hypotf(a, b), atan2=pr.atan2f(a, b), cos=pr.cosf(c))
    print("libdevice_f32.npz written")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
