#!/usr/bin/env python
"""The reference's reduction micro-benchmark (run_benchmarks, simulation.cu:1300-1358; its numbers: BASELINE.md 1d) on this
library's reductions, through the C-ABI (pf_bench_reductions). Prints a markdown table: median seconds and GB/s against the
measured HBM copy rate.   usage: python tools/bench_reductions.py [> profiles/rNN_reductions.md]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bachelors_b200 import solver as S                 # noqa: E402
from bachelors_b200._abi import PF_F32, PF_F64         # noqa: E402

peak = 6650.0
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
sizes = [256 * 256, 512 * 512, 1024 * 1024, 2048 * 2048, 4096 * 4096, 2 * 4096 * 4096]
print("# Reduction micro-benchmark (`pf_bench_reductions`, the reference's `run_benchmarks` sizes)\n")
print(f"Median of 200 runs, CUDA events on the solver's stream; GB = 1e9 bytes; HBM peak {peak:.0f} GB/s (MEASURED_PEAKS.json).")
print("Sizes up to 2048^2 floats (16.8 MB) are L2-resident after the first run, as in the reference's loop.\n")
with S.Solver(64, 64, PF_F64) as s:
    for dt, name, es in ((PF_F32, "f32", 4), (PF_F64, "f64", 8)):
        print(f"## {name}\n\n| N | max abs: s | GB/s | of HBM peak | 5 delta statistics (2 arrays): s | GB/s | of HBM peak |\n|---|---|---|---|---|---|---|")
        for n in sizes:
            tm, ts, mx = s.bench_reductions(n, dt, reps=200)
            gm, gs = n * es / tm / 1e9, 2 * n * es / ts / 1e9
            print(f"| {n} | {tm:.3e} | {gm:.0f} | {gm / peak:.2f} | {ts:.3e} | {gs:.0f} | {gs / peak:.2f} |")
        print()
