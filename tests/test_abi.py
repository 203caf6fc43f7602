"""The C-ABI library loads and exports every symbol include/pf_solver.h declares (no compute without a GPU)."""
import ctypes as C
import os
import re

import pytest

from bachelors_b200 import solver
from bachelors_b200._abi import (PF_ERR_NO_DEVICE, PF_F64, pf_params, pf_step_info, pf_slab, pf_delta_stats, pf_initial,
                                  pf_stats_row)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "pf_solver.h")).read()
    return sorted(set(re.findall(r"PF_API\s+[\w\s\*]+?\b(pf_\w+)\s*\(", text)))


def test_header_symbols_exported():
    names = declared_symbols()
    assert len(names) >= 25
    L = solver.lib()
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(solver.EXPORTS) == names


def test_no_other_default_visibility_symbols():
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", solver.LIB_PATH], capture_output=True, text=True, check=True).stdout
    syms = [l.split()[-1] for l in out.splitlines() if " T " in l]
    extra = [s for s in syms if not s.startswith("pf_") and s not in ("_init", "_fini")]
    assert not extra, extra[:10]


def test_abi_version_and_sizes():
    L = solver.lib()
    assert L.pf_abi_version() == 2
    assert L.pf_sizeof_real(PF_F64) == 8 and L.pf_sizeof_real(0) == 4
    assert [L.pf_required_ghost(s) for s in (1, 2, 3, 4)] == [1, 4, 5, 1]


def test_struct_layouts_match_header():
    """sizeof of the ctypes mirrors == sizeof in C (compiled from the header with gcc)."""
    import subprocess
    import tempfile
    src = '#include "pf_solver.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu %zu %zu\\n", sizeof(pf_params), sizeof(pf_step_info), sizeof(pf_slab), sizeof(pf_delta_stats), sizeof(pf_initial), sizeof(pf_stats_row));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "s")
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe], check=True)
        sizes = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [C.sizeof(pf_params), C.sizeof(pf_step_info), C.sizeof(pf_slab), C.sizeof(pf_delta_stats), C.sizeof(pf_initial),
                     C.sizeof(pf_stats_row)]


def test_create_fails_loudly_without_device():
    L = solver.lib()
    if L.pf_device_count() > 0:
        pytest.skip("a device is present")
    with pytest.raises(solver.PfError) as e:
        solver.Solver(64, 64, PF_F64)
    assert e.value.code == PF_ERR_NO_DEVICE
    assert "no CPU fallback" in str(e.value)


def test_product_never_touches_oracle():
    """The product package must not import, link or call anything under oracle/."""
    pkg = os.path.join(ROOT, "bachelors_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".c")) or f == "Makefile":
                text = open(os.path.join(base, f), errors="ignore").read()
                for line in text.splitlines():
                    code = line.split("//")[0].split("#")[0] if not f.endswith(".py") else line.split("#")[0]
                    assert not re.search(r"(import|from|include|-l|/)\s*\"?\.{0,2}/?oracle\b", code), (f, line)
