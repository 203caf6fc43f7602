"""Seeded synthetic inputs shared by tests, smoke() and bench.py (SURVEY section 8d "input value distributions")."""
import numpy as np

_MASK = (1 << 64) - 1


def splitmix64_stream(seed, n):
    """n draws of the reference's random_splitmix_64 (reference cuda_random.cuh:89-95), vectorised."""
    idx = np.arange(1, n + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + idx * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def bits_to_f64(bits):
    """random_bits_to_f64 (reference cuda_random.cuh:126-131): top 52 bits into the mantissa of [1,2), minus 1."""
    mant = (bits >> np.uint64(12)) | np.uint64(0x3FF0000000000000)
    return mant.view(np.float64) - 1.0


def seeded_fields(nx, ny, dtype=np.float64, kind="noise", seed=0x5EED5EED, L0=None):
    """(Phi, T) test fields of shape (ny, nx).

    noise : Phi, T ~ U[0,1) from splitmix64, one draw per cell in index order, Phi first then T.
    smooth: Phi = 1/2 + 0.4 sin(2pi 8x/L0) sin(2pi 8y/L0), T = 1/2 cos(2pi 4x/L0)  (x, y cell centres in units of L0).
    interface: a tanh-profiled disc plus weak noise - every term of the right-hand side is active.
    """
    n = nx * ny
    if kind == "noise":
        r = bits_to_f64(splitmix64_stream(seed, 2 * n))
        F = r[:n].reshape(ny, nx)
        U = r[n:].reshape(ny, nx)
    else:
        x = (np.arange(nx) + 0.5) / nx
        y = (np.arange(ny) + 0.5) / ny
        X, Y = np.meshgrid(x, y)
        if kind == "smooth":
            F = 0.5 + 0.4 * np.sin(2 * np.pi * 8 * X) * np.sin(2 * np.pi * 8 * Y)
            U = 0.5 * np.cos(2 * np.pi * 4 * X) + 0.0 * Y
        elif kind == "interface":
            rr = np.hypot(X - 0.5, (Y - 0.5) * ny / nx)
            F = 0.5 * (1 - np.tanh((rr - 0.2) / 0.02))
            r = bits_to_f64(splitmix64_stream(seed, n)).reshape(ny, nx)
            F = np.clip(F + 0.01 * (r - 0.5), 0, 1)
            U = 0.3 * F + 0.1 * np.cos(2 * np.pi * 3 * X)
        else:
            raise ValueError(kind)
    return np.ascontiguousarray(F, dtype=dtype), np.ascontiguousarray(U, dtype=dtype)
