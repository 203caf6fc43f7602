"""ctypes binding of tests/_helpers/libprobe.so (GPU libdevice probe; test helper only)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "libprobe.so")


def build(force=False):
    src = os.path.join(_HERE, "probe.cu")
    if force or not os.path.exists(LIB) or os.path.getmtime(src) > os.path.getmtime(LIB):
        subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-Xcompiler", "-fPIC", "-shared",
                        src, "-o", LIB, "-lcudart"], check=True)
    return LIB


class Probe:
    def __init__(self):
        self.L = C.CDLL(build())
        self.L.probe_eval.argtypes = [C.c_int, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]

    def _run(self, fn, a, b=None):
        a = np.ascontiguousarray(a, np.float32)
        out = np.empty_like(a)
        bp = None
        if b is not None:
            b = np.ascontiguousarray(b, np.float32)
            bp = b.ctypes.data
        rc = self.L.probe_eval(fn, a.size, a.ctypes.data, bp, out.ctypes.data)
        if rc:
            raise RuntimeError(f"probe_eval failed: {rc}")
        return out

    def hypotf(self, a, b):
        return self._run(0, a, b)

    def atan2f(self, y, x):
        return self._run(1, y, x)

    def cosf(self, a):
        return self._run(2, a)
