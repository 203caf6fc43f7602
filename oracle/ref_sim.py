"""ctypes binding of "oracle A": the UNMODIFIED reference solver built by oracle/build_ref.sh into
oracle/_ref/libref_sim_{f64,f32}.so. TEST INFRASTRUCTURE ONLY - needs a GPU (the reference has no CPU path).

The reference keeps the adaptive step size and the corrector parity in function statics
(reference simulation.cu:363, :936), so a fresh library instance is needed for a clean state:
`RefSim(..., fresh=True)` loads a private copy of the .so.
"""
import ctypes as C
import os
import shutil
import tempfile

import numpy as np

from bachelors_b200._abi import PF_F32, PF_F64, pf_params, pf_step_info

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")
_libs = {}
_tmpdirs = []


def lib_path(dtype, impl="ref"):
    """impl = "ref": the unmodified reference solver; "shim": the same glue and reference headers over this repository's
    bachelors_b200/host/simulation_shim.cpp + libpfsolver.so (the drop-in as the reference's host would link it)."""
    assert impl in ("ref", "shim")
    return os.path.join(_REF_DIR, "lib%s_sim_%s.so" % (impl, "f64" if dtype == PF_F64 else "f32"))


def available(dtype=PF_F64, impl="ref"):
    return os.path.exists(lib_path(dtype, impl))


def _declare(L):
    L.ref_sizeof_real.restype = C.c_int
    L.ref_ctx_create.restype = C.c_void_p
    L.ref_ctx_create.argtypes = [C.c_int, C.c_int]
    L.ref_ctx_destroy.argtypes = [C.c_void_p]
    L.ref_ctx_upload.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    L.ref_ctx_download.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    L.ref_ctx_run.restype = C.c_double
    L.ref_ctx_run.argtypes = [C.c_void_p, C.POINTER(pf_params), C.c_longlong, C.POINTER(pf_step_info)]
    L.ref_ctx_time_ms.restype = C.c_double
    L.ref_ctx_time_ms.argtypes = [C.c_void_p, C.POINTER(pf_params), C.c_longlong, C.POINTER(C.c_double)]
    L.ref_warning_count.restype = C.c_int
    return L


def load(dtype, fresh=False, impl="ref"):
    path = lib_path(dtype, impl)
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path}: run oracle/build_ref.sh where /root/reference exists")
    if impl == "shim":
        # the shim resolves libpfsolver.so relative to its own location (rpath $ORIGIN/../../bachelors_b200): load it in place.
        # Its adaptive step size lives in the pf_solver handle, not in a function static: no private copy needed.
        fresh = False
    if not fresh:
        if (dtype, impl) not in _libs:
            _libs[(dtype, impl)] = _declare(C.CDLL(path))
        return _libs[(dtype, impl)]
    d = tempfile.mkdtemp(prefix="refsim_")
    _tmpdirs.append(d)
    private = os.path.join(d, os.path.basename(path))
    shutil.copy(path, private)
    return _declare(C.CDLL(private))


class RefSim:
    """The reference's sim_realloc / sim_modify / sim_step behind numpy arrays of shape (ny, nx)."""

    def __init__(self, dtype, nx, ny, fresh=False, impl="ref"):
        self.L = load(dtype, fresh, impl)
        self.dtype, self.nx, self.ny = dtype, nx, ny
        self.np_dtype = np.float64 if dtype == PF_F64 else np.float32
        assert self.L.ref_sizeof_real() == (8 if dtype == PF_F64 else 4)
        self._h = self.L.ref_ctx_create(nx, ny)

    def close(self):
        if getattr(self, "_h", None):
            self.L.ref_ctx_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def upload(self, F, U):
        for which, a in ((0, F), (1, U)):
            a = np.ascontiguousarray(a)
            assert a.shape == (self.ny, self.nx) and a.dtype in (np.float32, np.float64)
            self.L.ref_ctx_upload(self._h, which, a.ctypes.data, int(a.dtype == np.float64))

    def download(self, dtype=None):
        dtype = dtype or self.np_dtype
        out = []
        for which in (0, 1):
            a = np.empty((self.ny, self.nx), dtype=dtype)
            self.L.ref_ctx_download(self._h, which, a.ctypes.data, int(a.dtype == np.float64))
            out.append(a)
        return out

    def run(self, params, n_steps):
        info = pf_step_info()
        t = self.L.ref_ctx_run(self._h, C.byref(params), n_steps, C.byref(info))
        return t, info

    def time_ms(self, params, n_steps):
        t = C.c_double(0)
        ms = self.L.ref_ctx_time_ms(self._h, C.byref(params), n_steps, C.byref(t))
        return ms, t.value

    @property
    def warnings(self):
        return self.L.ref_warning_count()
