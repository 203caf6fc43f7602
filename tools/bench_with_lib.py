#!/usr/bin/env python
"""Same-box A/B of library variants: runs bench.py (or any other script of this repo) against ANOTHER build of libpfsolver.so.

    nvcc <the flags of __graft_entry__.NVCC_FLAGS> [-DVARIANT ...] bachelors_b200/csrc/pf_solver.cu -o tools/_bin/libpf_variant.so -lcudart
    python tools/bench_with_lib.py tools/_bin/libpf_variant.so bench.py --solver explicit-rk4 --no-cpu --no-ref-gpu
    python tools/bench_with_lib.py tools/_bin/libpf_variant.so tools/si_phase_cost.py 256 1024

All variants of one comparison go into ONE gpurun call, alternating, each at least twice (box-to-box differences are larger than most
of the effects measured this way: profiles/r02_kernel_experiments.md section 6, profiles/r02_si_phase_cost.md)."""
import os
import runpy
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from bachelors_b200 import solver as S  # noqa: E402

if __name__ == "__main__":
    if len(sys.argv) < 3:
        sys.exit(__doc__)
    S.LIB_PATH = os.path.abspath(sys.argv[1])
    script = sys.argv[2] if os.path.isabs(sys.argv[2]) else os.path.join(ROOT, sys.argv[2])
    sys.argv = [script] + sys.argv[3:]
    sys.path.insert(0, os.path.dirname(script))
    runpy.run_path(script, run_name="__main__")
