"""ctypes binding of oracle/_build/liboracle.so - the CPU oracle. TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from bachelors_b200._abi import PF_F32, PF_F64, pf_params, pf_step_info, pf_delta_stats

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build(force=False):
    """Compile the C restatement (gcc). Building the checker is not using it."""
    if force or not os.path.exists(_LIB_PATH) or any(
        os.path.getmtime(os.path.join(_HERE, f)) > os.path.getmtime(_LIB_PATH)
        for f in ("pf_oracle.c", "pf_oracle_impl.h", "pf_oracle.h", "libdevice_f32.h", "Makefile")
    ):
        subprocess.run(["make", "-C", _HERE, "-s"], check=True)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int, C.c_int, C.c_int]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_reset.argtypes = [C.c_void_p]
        L.orc_set_threads.argtypes = [C.c_int]
        L.orc_max_threads.restype = C.c_int
        L.orc_step.restype = C.c_double
        L.orc_step.argtypes = [C.c_void_p, C.POINTER(pf_params), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(pf_step_info)]
        L.orc_run.restype = C.c_double
        L.orc_run.argtypes = [C.c_void_p, C.POINTER(pf_params), C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.POINTER(pf_step_info)]
        L.orc_rhs.argtypes = [C.c_void_p, C.POINTER(pf_params), C.c_int, C.c_int, C.POINTER(C.c_double),
                              C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p]
        L.orc_cg_solve.argtypes = [C.c_void_p, C.POINTER(pf_params), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_double, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.orc_debug_values.argtypes = [C.c_void_p, C.POINTER(pf_params)] + [C.c_void_p] * 5
        L.orc_stats_delta.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(pf_delta_stats)]
        for name in ("orc_hypotf_v", "orc_atan2f_v"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_cosf_v.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        _lib = L
    return _lib


MATVEC_F64 = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.c_void_p, C.POINTER(C.c_double), C.c_int)


def _np_dtype(dtype):
    return np.float64 if dtype == PF_F64 else np.float32


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """Host-side stepping through the oracle; arrays are numpy, shape (ny, nx)."""

    def __init__(self, dtype, nx, ny):
        self.dtype, self.nx, self.ny = dtype, nx, ny
        self.np_dtype = _np_dtype(dtype)
        self._h = lib().orc_create(dtype, nx, ny)

    def close(self):
        if self._h:
            lib().orc_destroy(self._h)
            self._h = None

    __del__ = close

    def reset(self):
        lib().orc_reset(self._h)

    def _chk(self, a):
        a = np.ascontiguousarray(a, dtype=self.np_dtype)
        assert a.shape == (self.ny, self.nx), (a.shape, (self.ny, self.nx))
        return a

    def step(self, params, F, U):
        F, U = self._chk(F), self._chk(U)
        nF, nU = np.empty_like(F), np.empty_like(U)
        info = pf_step_info()
        dt = lib().orc_step(self._h, C.byref(params), _ptr(F), _ptr(U), _ptr(nF), _ptr(nU), C.byref(info))
        return nF, nU, dt, info

    def run(self, params, F, U, n_steps):
        F, U = self._chk(F).copy(), self._chk(U).copy()
        info = pf_step_info()
        done = C.c_int64(0)
        t = lib().orc_run(self._h, C.byref(params), _ptr(F), _ptr(U), n_steps, C.byref(done), C.byref(info))
        return F, U, t, info

    def rhs(self, params, weights, Fs, Us, is_euler=False):
        n = len(weights)
        Fs = [self._chk(a) for a in Fs]
        Us = [self._chk(a) for a in Us]
        w = (C.c_double * n)(*[float(x) for x in weights])
        pf = (C.c_void_p * n)(*[a.ctypes.data for a in Fs])
        pu = (C.c_void_p * n)(*[a.ctypes.data for a in Us])
        oF, oU = np.empty_like(Fs[0]), np.empty_like(Us[0])
        lib().orc_rhs(self._h, C.byref(params), int(is_euler), n, w, pf, pu, _ptr(oF), _ptr(oU))
        return oF, oU

    def cg_solve(self, params, which, aniso, b, x0, tolerance, max_iters):
        b = self._chk(b)
        an = self._chk(aniso) if aniso is not None else None
        x0a = self._chk(x0) if x0 is not None else None
        x = np.empty_like(b)
        it, conv, err = C.c_int(0), C.c_int(0), C.c_double(0)
        lib().orc_cg_solve(self._h, C.byref(params), which, _ptr(an) if an is not None else None, _ptr(b),
                           _ptr(x0a) if x0a is not None else None, _ptr(x), tolerance, max_iters,
                           C.byref(it), C.byref(conv), C.byref(err))
        return x, it.value, bool(conv.value), err.value

    def debug_values(self, params, F, U):
        F, U = self._chk(F), self._chk(U)
        g, h, a = np.empty_like(F), np.empty_like(F), np.empty_like(F)
        lib().orc_debug_values(self._h, C.byref(params), _ptr(F), _ptr(U), _ptr(g), _ptr(h), _ptr(a))
        return g, h, a

    def stats_delta(self, a, b):
        a, b = self._chk(a), self._chk(b)
        out = pf_delta_stats()
        lib().orc_stats_delta(self._h, _ptr(a), _ptr(b), C.byref(out))
        return out


def cg_generic(matvec, x_shape_n, b, x0=None, epsilon=1e-10, tolerance=1e-5, max_iters=10):
    """conjugate_gradient_solve over a Python operator (double): matvec(x: ndarray) -> ndarray."""
    L = lib()
    L.orc_cg_generic_f64.argtypes = [MATVEC_F64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                     C.c_double, C.c_double, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double)]
    N = x_shape_n

    def cb(out, A, x, n):
        xv = np.ctypeslib.as_array(x, shape=(n,))
        ov = np.ctypeslib.as_array(out, shape=(n,))
        ov[:] = matvec(xv)

    cbp = MATVEC_F64(cb)
    b = np.ascontiguousarray(b, dtype=np.float64)
    x = np.zeros(N)
    x0a = np.ascontiguousarray(x0, dtype=np.float64) if x0 is not None else None
    it, conv, err = C.c_int(0), C.c_int(0), C.c_double(0)
    L.orc_cg_generic_f64(cbp, None, _ptr(x), _ptr(b), N, _ptr(x0a) if x0a is not None else None,
                         epsilon, tolerance, max_iters, C.byref(it), C.byref(conv), C.byref(err))
    return x, it.value, bool(conv.value), err.value


def hypotf(a, b):
    a = np.ascontiguousarray(a, np.float32); b = np.ascontiguousarray(b, np.float32)
    o = np.empty_like(a); lib().orc_hypotf_v(_ptr(a), _ptr(b), _ptr(o), a.size); return o


def atan2f(y, x):
    y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
    o = np.empty_like(y); lib().orc_atan2f_v(_ptr(y), _ptr(x), _ptr(o), y.size); return o


def cosf(a):
    a = np.ascontiguousarray(a, np.float32)
    o = np.empty_like(a); lib().orc_cosf_v(_ptr(a), _ptr(o), a.size); return o
