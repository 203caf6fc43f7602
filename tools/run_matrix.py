"""Whole benchmark runs of every config of the reference's matrix (benchmark_configs/) through the
headless host (bachelors_b200/app.py = the reference's non-interactive loop, main.cpp:495-564): t = 0 .. 0.04, snapshots
and stats.csv written as the reference writes them. Prints one markdown table next to the published A100 runtimes
(reference benchmark_ellapsed_times.txt). Development / measurement tool.

    python tools/run_matrix.py [--sizes 128 256 512 1024 2048] [--dtype f64] [--out gpurun_out/matrix.md]
"""
import argparse
import os
import shutil
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from bachelors_b200._abi import PF_F32, PF_F64  # noqa: E402
from bachelors_b200.app import run_config  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# runtime s of the reference on its A100 node (benchmark_ellapsed_times.txt:4-7, :21-24, :52-55, :81-84, :110-113)
PUBLISHED = {
    "explicit": {128: 0.20, 256: 0.35, 512: 0.66, 1024: 1.64, 2048: 3.50},
    "explicit-rk4": {128: 0.76, 256: 1.42, 512: 2.88, 1024: 8.73, 2048: 18.73},
    "explicit-rk4-adaptive": {128: 1.59, 256: 2.85, 512: 5.39, 1024: 16.92, 2048: 35.87},
    "semi-implicit": {128: 3.86, 256: 4.45, 512: 5.67, 1024: 12.43, 2048: 34.05},
}
PUBLISHED_ITERS = {"explicit-rk4-adaptive": {128: 8920, 256: 9522, 512: 9522, 1024: 9522, 2048: 9522}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", type=int, nargs="+", default=[128, 256, 512, 1024, 2048])
    ap.add_argument("--solvers", nargs="+", default=list(PUBLISHED))
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    dtype = PF_F64 if a.dtype == "f64" else PF_F32
    rows = []
    for solver in a.solvers:
        for size in a.sizes:
            path = os.path.join(ROOT, "benchmark_configs", "config_%s_%d.ini" % (solver, size))
            if not os.path.exists(path):
                continue
            tmp = tempfile.mkdtemp(prefix="pf_matrix_")
            try:
                out = run_config(path, dtype, overrides=["[snapshot]\nfolder = %s" % tmp])
            finally:
                shutil.rmtree(tmp, ignore_errors=True)
            pub = PUBLISHED.get(solver, {}).get(size)
            pub_it = PUBLISHED_ITERS.get(solver, {}).get(size, 8000)
            rows.append((solver, size, out["iters"], out["runtime_s"], out["runtime_s"] / max(out["iters"], 1) * 1e3, out["snapshots"], pub, pub_it))
            print("[matrix] %-22s %5d^2: %6d iters %8.3f s (%.4f ms/step), published %s s" % (solver, size, out["iters"], out["runtime_s"],
                  rows[-1][4], pub), flush=True)
    lines = ["| solver | grid | iterations (published) | runtime s, this repo (1 x B200, %s) | ms/step | snapshots written | published runtime s (A100) | ratio |" % a.dtype,
             "|---|---|---|---|---|---|---|---|"]
    for solver, size, iters, rt, ms, snaps, pub, pub_it in rows:
        lines.append("| %s | %d² | %d (%d) | %.3f | %.4f | %d + initial | %s | %s |" % (
            solver, size, iters, pub_it, rt, ms, snaps, "%.2f" % pub if pub else "—", "%.0f x" % (pub / rt) if pub else "—"))
    text = "\n".join(lines) + "\n"
    print(text)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            f.write(text)


if __name__ == "__main__":
    main()
