"""ctypes binding of the C-ABI in include/pf_solver.h (bachelors_b200/libpfsolver.so).

There is no CPU fallback: if the CUDA library is missing or no device is present, this raises.
"""
import ctypes as C
import os

import numpy as np

from ._abi import (PF_ABI_VERSION, PF_F32, PF_F64, PF_OK, pf_params, pf_step_info, pf_delta_stats, pf_slab, pf_initial,
                   pf_stats_row)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpfsolver.so")

# every symbol include/pf_solver.h declares
EXPORTS = [
    "pf_abi_version", "pf_device_count", "pf_last_error", "pf_sizeof_real", "pf_required_ghost",
    "pf_create", "pf_create_slab", "pf_destroy", "pf_reset_state", "pf_set_option",
    "pf_field_alloc", "pf_field_free", "pf_field_rows", "pf_upload", "pf_download", "pf_upload_raw", "pf_download_raw",
    "pf_aux_stream", "pf_step_rows_async", "pf_rkm_attempt_rows_async", "pf_rkm_clear_errors_async",
    "pf_rkm_begin", "pf_rkm_attempt_async", "pf_rkm_error_bits", "pf_rkm_decide", "pf_rkm_finish_async",
    "pf_step", "pf_step_host", "pf_run", "pf_run_until", "pf_step_async", "pf_sync", "pf_stream",
    "pf_rhs", "pf_cg_solve", "pf_debug_values", "pf_stats_delta", "pf_stats_delta_static", "pf_halo_region", "pf_kernel_launches", "pf_bench_reductions",
    "pf_init_fields", "pf_snapshot_begin", "pf_snapshot_wait", "pf_write_maps_bin", "pf_write_maps_bin_header",
    "pf_write_maps_bin_rows", "pf_write_stats_csv", "pf_render_rgba", "pf_write_png",
    "pf_si_begin_async", "pf_si_partials", "pf_si_cg_begin_async", "pf_si_cg_spmv_async", "pf_si_cg_update_async",
    "pf_si_cg_pupdate_async", "pf_si_direction", "pf_si_cg_poll",
    "pf_ipc_export", "pf_ipc_open", "pf_ipc_close", "pf_halo_flags", "pf_halo_push_async", "pf_halo_wait_async", "pf_halo_check", "pf_halo_link",
]


class PfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"pf_solver error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    """Loads libpfsolver.so. Fails loudly when the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(nvcc, sm_100a). This package has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, dbl, sz = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_size_t
    P = C.POINTER
    sig = {
        "pf_abi_version": (i32, []), "pf_device_count": (i32, []), "pf_last_error": (C.c_char_p, [vp]),
        "pf_sizeof_real": (sz, [i32]), "pf_required_ghost": (i32, [i32]),
        "pf_create": (i32, [P(vp), i32, i32, i32, i32]), "pf_create_slab": (i32, [P(vp), i32, i32, i32, i32, P(pf_slab)]),
        "pf_destroy": (None, [vp]), "pf_reset_state": (i32, [vp]), "pf_set_option": (i32, [vp, i32, i32]),
        "pf_field_alloc": (i32, [vp, P(vp)]), "pf_field_free": (i32, [vp, vp]), "pf_field_rows": (i32, [vp, P(i32), P(i32)]),
        "pf_upload": (i32, [vp, vp, vp, sz, i32]), "pf_download": (i32, [vp, vp, vp, sz, i32]),
        "pf_upload_raw": (i32, [vp, vp, vp, sz]), "pf_download_raw": (i32, [vp, vp, vp, sz]),
        "pf_aux_stream": (vp, [vp]),
        "pf_step_rows_async": (i32, [vp, P(pf_params), vp, vp, vp, vp, i32, i32, i32]),
        "pf_rkm_attempt_rows_async": (i32, [vp, P(pf_params), dbl, vp, vp, vp, vp, i32, i32, i32]),
        "pf_rkm_clear_errors_async": (i32, [vp, i32]),
        "pf_rkm_begin": (i32, [vp, P(pf_params), vp, vp, P(dbl)]),
        "pf_rkm_attempt_async": (i32, [vp, P(pf_params), dbl, vp, vp, vp, vp]),
        "pf_rkm_error_bits": (vp, [vp]),
        "pf_rkm_decide": (i32, [vp, P(pf_params), dbl, C.c_uint64, C.c_uint64, P(i32), P(i32), P(dbl), P(dbl), P(dbl)]),
        "pf_rkm_finish_async": (i32, [vp, P(pf_params), dbl, dbl, vp, vp, vp, vp]),
        "pf_step": (i32, [vp, P(pf_params), vp, vp, vp, vp, P(pf_step_info)]),
        "pf_step_host": (i32, [vp, P(pf_params), vp, vp, vp, vp, i32, vp, vp, vp, vp, P(pf_step_info)]),
        "pf_run": (i32, [vp, P(pf_params), vp, vp, vp, vp, i64, dbl, P(vp), P(vp), P(i64), P(dbl), P(pf_step_info)]),
        "pf_run_until": (i32, [vp, P(pf_params), vp, vp, vp, vp, i64, dbl, dbl, P(vp), P(vp), P(i64), P(dbl), P(dbl), P(dbl), P(pf_step_info)]),
        "pf_step_async": (i32, [vp, P(pf_params), vp, vp, vp, vp]), "pf_sync": (i32, [vp]), "pf_stream": (vp, [vp]),
        "pf_rhs": (i32, [vp, P(pf_params), i32, P(dbl), P(vp), P(vp), vp, vp]),
        "pf_cg_solve": (i32, [vp, P(pf_params), i32, vp, vp, vp, vp, dbl, i32, P(i32), P(i32), P(dbl)]),
        "pf_debug_values": (i32, [vp, P(pf_params), vp, vp, vp, vp, vp]),
        "pf_stats_delta": (i32, [vp, vp, vp, P(pf_delta_stats)]),
        "pf_stats_delta_static": (i32, [i32, i32, vp, vp, sz, vp, P(pf_delta_stats)]),
        "pf_halo_region": (i32, [vp, vp, i32, P(vp), P(vp), P(sz)]),
        "pf_kernel_launches": (i64, [vp]),
        "pf_bench_reductions": (i32, [vp, sz, i32, i32, P(dbl), P(dbl), P(dbl)]),
        "pf_init_fields": (i32, [vp, P(pf_params), P(pf_initial), vp, vp]),
        "pf_snapshot_begin": (i32, [vp, i32, i32, P(vp)]), "pf_snapshot_wait": (i32, [vp, i32, P(vp)]),
        "pf_write_maps_bin": (i32, [C.c_char_p, i32, i32, dbl, dbl, i64, dbl, i32, P(C.c_char_p), P(vp)]),
        "pf_write_maps_bin_header": (i32, [C.c_char_p, i32, i32, dbl, dbl, i64, dbl, i32, P(C.c_char_p)]),
        "pf_write_maps_bin_rows": (i32, [C.c_char_p, i32, i32, i32, i32, i32, i32, vp]),
        "pf_write_stats_csv": (i32, [C.c_char_p, i32, i32, i32, dbl, i32, sz, P(pf_stats_row)]),
        "pf_render_rgba": (i32, [vp, vp, C.c_float, C.c_float, vp]),
        "pf_write_png": (i32, [C.c_char_p, i32, i32, vp, i32]),
        "pf_si_begin_async": (i32, [vp, P(pf_params), vp, vp, vp, vp, vp]), "pf_si_partials": (i32, [vp]),
        "pf_si_cg_begin_async": (i32, [vp, i32, vp]), "pf_si_cg_spmv_async": (i32, [vp, i32, vp]),
        "pf_si_cg_update_async": (i32, [vp, i32, vp, vp, i32, vp]), "pf_si_cg_pupdate_async": (i32, [vp, i32, vp, vp, i32, i32]),
        "pf_si_direction": (vp, [vp]), "pf_si_cg_poll": (i32, [vp, i32, P(i32), P(i32), P(i32), P(dbl)]),
        "pf_ipc_export": (i32, [vp, vp, vp]), "pf_ipc_open": (i32, [vp, vp, P(vp)]), "pf_ipc_close": (i32, [vp, vp]),
        "pf_halo_flags": (vp, [vp]),
        "pf_halo_push_async": (i32, [vp, i32, i32, P(vp), P(vp), sz, vp, i32, C.c_uint, i32]),
        "pf_halo_wait_async": (i32, [vp, i32, C.c_uint, i32]), "pf_halo_check": (i32, [vp]),
        "pf_halo_link": (i32, [vp, i32, P(vp), P(vp), sz, vp, P(vp), vp]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype = res
        f.argtypes = args
    if L.pf_abi_version() != PF_ABI_VERSION:
        raise ImportError(f"libpfsolver.so ABI {L.pf_abi_version()} != binding {PF_ABI_VERSION}: rebuild")
    _lib = L
    return L


def np_dtype(dtype):
    return np.float64 if dtype == PF_F64 else np.float32


def pf_dtype_of(a):
    if a.dtype == np.float64:
        return PF_F64
    if a.dtype == np.float32:
        return PF_F32
    raise TypeError(a.dtype)


class Solver:
    """One pf_solver handle (one GPU, one grid or slab). Thin: every method is one C-ABI call."""

    def __init__(self, nx, ny, dtype=PF_F64, device=0, slab=None):
        self._h = C.c_void_p()
        L = lib()
        self.nx, self.ny, self.dtype, self.device = nx, ny, dtype, device
        self._slab = slab
        if slab is None:
            rc = L.pf_create(C.byref(self._h), device, dtype, nx, ny)
        else:
            rc = L.pf_create_slab(C.byref(self._h), device, dtype, nx, ny, C.byref(slab))
        if rc != PF_OK:
            self._h = C.c_void_p()
            raise PfError(rc, L.pf_last_error(None).decode())
        rows, first = C.c_int(0), C.c_int(0)
        L.pf_field_rows(self._h, C.byref(rows), C.byref(first))
        self.rows_alloc, self.first_owned_row = rows.value, first.value
        self._fields = []

    # -- lifetime
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            for f in list(self._fields):
                lib().pf_field_free(self._h, f)
            self._fields.clear()
            lib().pf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc != PF_OK:
            raise PfError(rc, lib().pf_last_error(self._h).decode())

    # -- fields
    def field_alloc(self):
        p = C.c_void_p()
        self._ck(lib().pf_field_alloc(self._h, C.byref(p)))
        self._fields.append(p.value)
        return p.value

    def field_free(self, f):
        self._ck(lib().pf_field_free(self._h, f))
        self._fields.remove(f)

    def upload(self, dev, host, offset_elems=0):
        host = np.ascontiguousarray(host)
        rs = 8 if self.dtype == PF_F64 else 4
        self._ck(lib().pf_upload(self._h, dev + offset_elems * rs, host.ctypes.data, host.size, pf_dtype_of(host)))

    def download(self, dev, shape=None, dtype=None, offset_elems=0, out=None):
        if out is None:
            shape = shape if shape is not None else (self.rows_alloc, self.nx)
            out = np.empty(shape, dtype=dtype if dtype is not None else np_dtype(self.dtype))
        rs = 8 if self.dtype == PF_F64 else 4
        self._ck(lib().pf_download(self._h, dev + offset_elems * rs, out.ctypes.data, out.size, pf_dtype_of(out)))
        return out

    def field_from(self, host):
        f = self.field_alloc()
        self.upload(f, host)
        return f

    # -- stepping
    def step(self, params, F, U, next_F, next_U):
        info = pf_step_info()
        self._ck(lib().pf_step(self._h, C.byref(params), F, U, next_F, next_U, C.byref(info)))
        return info

    def step_host(self, params, host_F, host_U, host_next_F, host_next_U, F, U, next_F, next_U):
        """upload + step + download in one call (numpy arrays or raw host pointers of the solver's dtype)."""
        def ptr(a):
            return a.ctypes.data if hasattr(a, "ctypes") else a
        info = pf_step_info()
        self._ck(lib().pf_step_host(self._h, C.byref(params), ptr(host_F), ptr(host_U), ptr(host_next_F), ptr(host_next_U),
                                    self.dtype, F, U, next_F, next_U, C.byref(info)))
        return info

    def step_async(self, params, F, U, next_F, next_U):
        self._ck(lib().pf_step_async(self._h, C.byref(params), F, U, next_F, next_U))

    def run(self, params, F, U, F_tmp, U_tmp, n_steps, stop_time=0.0, pause_time=0.0):
        Fo, Uo, done, t = C.c_void_p(), C.c_void_p(), C.c_int64(0), C.c_double(0)
        info = pf_step_info()
        t_end, t_last = C.c_double(0), C.c_double(0)
        self._ck(lib().pf_run_until(self._h, C.byref(params), F, U, F_tmp, U_tmp, n_steps, stop_time, pause_time,
                                    C.byref(Fo), C.byref(Uo), C.byref(done), C.byref(t), C.byref(t_end), C.byref(t_last), C.byref(info)))
        self.last_run_last_step_time = t_last.value                         # the time the last step started from
        self.last_run_end_time = t_end.value                                # params.time + dt + dt + ...: the host loop's sim_time
        return Fo.value, Uo.value, done.value, t.value, info

    def sync(self):
        self._ck(lib().pf_sync(self._h))

    def reset_state(self):
        self._ck(lib().pf_reset_state(self._h))

    def set_option(self, option, value):
        self._ck(lib().pf_set_option(self._h, option, int(value)))

    @property
    def stream(self):
        return lib().pf_stream(self._h)

    @property
    def kernel_launches(self):
        return lib().pf_kernel_launches(self._h)

    # -- building blocks
    def rhs(self, params, weights, Fs, Us, out_F, out_U):
        n = len(weights)
        w = (C.c_double * n)(*[float(x) for x in weights])
        pf = (C.c_void_p * n)(*Fs)
        pu = (C.c_void_p * n)(*Us)
        self._ck(lib().pf_rhs(self._h, C.byref(params), n, w, pf, pu, out_F, out_U))

    def cg_solve(self, params, which, aniso, b, x0, x, tolerance, max_iters):
        it, conv, err = C.c_int(0), C.c_int(0), C.c_double(0)
        self._ck(lib().pf_cg_solve(self._h, C.byref(params), which, aniso, b, x0, x, tolerance, max_iters,
                                   C.byref(it), C.byref(conv), C.byref(err)))
        return it.value, bool(conv.value), err.value

    def debug_values(self, params, F, U, grad_F, grad_U, aniso):
        self._ck(lib().pf_debug_values(self._h, C.byref(params), F, U, grad_F, grad_U, aniso))

    def stats_delta(self, a, b):
        out = pf_delta_stats()
        self._ck(lib().pf_stats_delta(self._h, a, b, C.byref(out)))
        return out

    def bench_reductions(self, n, value_dtype=PF_F32, reps=50):
        """-> (seconds of max|x| over n values, seconds of the five delta statistics over 2n values, the maximum)."""
        a, b, m = C.c_double(0), C.c_double(0), C.c_double(0)
        self._ck(lib().pf_bench_reductions(self._h, n, value_dtype, reps, C.byref(a), C.byref(b), C.byref(m)))
        return a.value, b.value, m.value

    # -- rows around the stepping path
    def init_fields(self, params, init, F, U):
        """Initial conditions evaluated on the device (main.cpp:93-136)."""
        self._ck(lib().pf_init_fields(self._h, C.byref(params), C.byref(init), F, U))

    def snapshot_begin(self, slot, fields):
        arr = (C.c_void_p * len(fields))(*fields)
        self._ck(lib().pf_snapshot_begin(self._h, slot, len(fields), arr))
        self._snap_n = getattr(self, "_snap_n", {})
        self._snap_n[slot] = len(fields)

    def snapshot_wait(self, slot, owned_rows=None):
        """-> list of (rows, nx) float64 views of the slot's pinned host buffers (valid until the slot is reused)."""
        n = self._snap_n[slot]
        ptrs = (C.c_void_p * n)()
        self._ck(lib().pf_snapshot_wait(self._h, slot, ptrs))
        rows = owned_rows if owned_rows is not None else self.owned_rows
        out = []
        for p in ptrs:
            buf = (C.c_double * (rows * self.nx)).from_address(p)
            out.append(np.frombuffer(buf, dtype=np.float64).reshape(rows, self.nx))
        return out

    @property
    def owned_rows(self):
        return self.ny if self._slab is None else self._slab.ny_local

    def render_rgba(self, field, display_min, display_max):
        out = np.empty((self.owned_rows, self.nx, 4), dtype=np.uint8)
        self._ck(lib().pf_render_rgba(self._h, field, display_min, display_max, out.ctypes.data))
        return out

    def halo_region(self, field, side):
        send, recv, n = C.c_void_p(), C.c_void_p(), C.c_size_t(0)
        self._ck(lib().pf_halo_region(self._h, field, side, C.byref(send), C.byref(recv), C.byref(n)))
        return send.value, recv.value, n.value


def stats_delta_static(device, dtype, a, b, n, stream=None):
    """pf_stats_delta_static: the delta statistics of a - b (device pointers, n values) through the entry point that allocates
    nothing (scratch in a file-scope __device__ array; n <= PF_STATIC_STATS_MAX_ELEMS). No counterpart in the reference."""
    out = pf_delta_stats()
    rc = lib().pf_stats_delta_static(device, dtype, a, b, n, stream, C.byref(out))
    if rc != PF_OK:
        raise PfError(rc, lib().pf_last_error(None).decode())
    return out


PF_STATIC_STATS_MAX_ELEMS = 1 << 26


# ---- output files of the reference host (pure host code in the library: usable without a GPU) -------------------
def _ck_io(rc):
    if rc != PF_OK:
        raise PfError(rc, lib().pf_last_error(None).decode())


def write_maps_bin(path, dx, dy, it, time, maps):
    """maps: {name: (ny, nx) float64 array}; byte-compatible with save_bin_map_file (main.cpp:705-735)."""
    names = list(maps)
    arrs = [np.ascontiguousarray(maps[n], dtype=np.float64) for n in names]
    ny, nx = arrs[0].shape
    cn = (C.c_char_p * len(names))(*[n.encode() for n in names])
    cp = (C.c_void_p * len(names))(*[a.ctypes.data for a in arrs])
    _ck_io(lib().pf_write_maps_bin(path.encode(), nx, ny, dx, dy, it, time, len(names), cn, cp))


def write_maps_bin_header(path, nx, ny, dx, dy, it, time, names):
    cn = (C.c_char_p * len(names))(*[n.encode() for n in names])
    _ck_io(lib().pf_write_maps_bin_header(path.encode(), nx, ny, dx, dy, it, time, len(names), cn))


def write_maps_bin_rows(path, nx, ny, n_maps, map_index, y0, data):
    data = np.ascontiguousarray(data, dtype=np.float64)
    _ck_io(lib().pf_write_maps_bin_rows(path.encode(), nx, ny, n_maps, map_index, y0, data.shape[0], data.ctypes.data))


def write_stats_csv(path, append, nx, ny, dt, step_res_count, rows):
    arr = (pf_stats_row * len(rows))(*rows)
    _ck_io(lib().pf_write_stats_csv(path.encode(), int(append), nx, ny, dt, step_res_count, len(rows), arr))


def write_png(path, rgba, flip_y=True):
    rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
    h, w, _ = rgba.shape
    _ck_io(lib().pf_write_png(path.encode(), w, h, rgba.ctypes.data, int(flip_y)))
