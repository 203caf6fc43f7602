"""pf_stats_delta_static (include/pf_solver.h): the delta statistics through the entry point that allocates nothing - its
scratch is a file-scope __device__ array whose address the library takes with cudaGetSymbolAddress. No counterpart in the
reference. Checked against pf_stats_delta (the entry point with the solver
handle's cudaMalloc'ed scratch) on the same device fields: same kernel, same grid, same order of the final sum, so the five
numbers must be the same bits. Two calls in a row on DIFFERENT inputs, the second smaller than the first: the shared array
still holds the first call's partials beyond the second call's blocks, and they must not leak into the result.

The file is named so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

from bachelors_b200 import solver as S
from bachelors_b200._abi import PF_F32, PF_F64, PF_OK, PF_ERR_INVALID, pf_delta_stats

pytestmark = pytest.mark.gpu
NP = {PF_F64: np.float64, PF_F32: np.float32}
FIELDS = ("sum", "L1", "L2", "min", "max")


def as_tuple(d):
    return tuple(getattr(d, k) for k in FIELDS)


def numpy_stats(a, b):
    d = a.astype(np.float64) - b.astype(np.float64)
    return (d.sum(), np.abs(d).sum(), np.sqrt((d * d).sum()), d.min(), d.max())


@pytest.mark.parametrize("dtype", [PF_F64, PF_F32])
def test_static_scratch_entry_point_equals_pf_stats_delta_on_two_calls_in_a_row(dtype):
    rng = np.random.default_rng(20260511 + dtype)
    # first a field with more blocks of partials than the second: stale slots stay behind in the shared array
    for nx, ny in ((1000, 640), (96, 40)):
        a = rng.standard_normal((ny, nx)).astype(NP[dtype])
        b = (a + 1e-3 * rng.standard_normal((ny, nx))).astype(NP[dtype])
        with S.Solver(nx, ny, dtype) as s:
            da, db = s.field_from(a), s.field_from(b)
            want = as_tuple(s.stats_delta(da, db))
            for _ in range(2):                                             # and twice on the same input
                got = as_tuple(S.stats_delta_static(s.device, dtype, da, db, nx * ny))
                assert got == want                                          # bit for bit
            # on the solver's own stream as well
            got = as_tuple(S.stats_delta_static(s.device, dtype, da, db, nx * ny, stream=s.stream))
            assert got == want
        ref = numpy_stats(a, b)
        rel = 1e-12 if dtype == PF_F64 else 2e-4                            # fp32 sums over 640000 values
        for g, r in zip(got, ref):
            assert abs(g - r) <= rel * max(1.0, abs(r)) * (1 if dtype == PF_F64 else 10)


def test_static_scratch_entry_point_rejects_what_it_cannot_hold():
    with S.Solver(64, 64, PF_F64) as s:
        da = s.field_from(np.ones((64, 64)))
        for n in (0, S.PF_STATIC_STATS_MAX_ELEMS + 1):
            with pytest.raises(S.PfError) as e:
                S.stats_delta_static(s.device, PF_F64, da, da, n)
            assert e.value.code == PF_ERR_INVALID and "PF_STATIC_STATS_MAX_ELEMS" in str(e.value)
        with pytest.raises(S.PfError):
            S.stats_delta_static(s.device, 7, da, da, 64)
        with pytest.raises(S.PfError):
            S.stats_delta_static(s.device, PF_F64, None, da, 64)
        out = pf_delta_stats()
        assert S.lib().pf_stats_delta_static(s.device, PF_F64, da, da, C.c_size_t(64 * 64), None, C.byref(out)) == PF_OK
        assert (out.sum, out.L1, out.L2, out.min, out.max) == (0.0, 0.0, 0.0, 0.0, 0.0)
