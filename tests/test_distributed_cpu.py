"""World-size-2/3 gloo tests (CPU) of the host-side multi-GPU logic: slab layout, neighbour ranks, the halo-exchange
plan and the bit-pattern MAX all-reduce of bachelors_b200/distributed.py. The per-rank compute is a numpy stand-in
(a k-stage explicit diffusion step that consumes k ghost rows, like the temporally blocked integrators); the CUDA
kernels themselves are covered on one device by tests/test_gpu_parity.py::test_slabs_reproduce_the_single_grid_bit_for_bit
and across devices by tests/test_gpu_multi.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bachelors_b200.slabs import neighbours, slab_layout, slab_rows


def test_slab_rows_partition_the_grid():
    for ny in (1, 7, 128, 2048, 16384, 1000):
        for G in (1, 2, 3, 4, 8):
            if G > ny:
                continue
            rows = [slab_rows(ny, G, g) for g in range(G)]
            assert rows[0][0] == 0 and rows[-1][1] == ny
            assert all(rows[i][1] == rows[i + 1][0] for i in range(G - 1))
            sizes = [b - a for a, b in rows]
            assert max(sizes) - min(sizes) <= 1


def test_layout_and_neighbours():
    assert neighbours(1, 0, True) == (None, None)
    assert neighbours(4, 0, False) == (None, 1) and neighbours(4, 3, False) == (2, None)
    assert neighbours(4, 0, True) == (3, 1) and neighbours(4, 3, True) == (2, 0)
    assert neighbours(2, 0, True) == (1, 1)
    sl = slab_layout(100, 4, 0, 4, periodic=False)
    assert (sl.y0, sl.ny_local, sl.has_lo, sl.has_hi, sl.ghost) == (0, 25, 0, 1, 4)
    sl = slab_layout(100, 4, 0, 4, periodic=True)
    assert (sl.has_lo, sl.has_hi) == (1, 1)
    sl = slab_layout(100, 1, 0, 4, periodic=True)
    assert (sl.has_lo, sl.has_hi, sl.ghost) == (0, 0, 0)


def _standin_step(u, stages, bc_lo_open, bc_hi_open, periodic_x=False):
    """`stages` sweeps of u += 0.1*laplace(u) on a (rows, nx) array. Rows beyond an 'open' (cut) edge are unknown: they
    are treated as Neumann there, which contaminates one more row per sweep - exactly what the ghost rows absorb."""
    for _ in range(stages):
        p = np.pad(u, 1, mode="edge")
        u = u + 0.1 * ((p[1:-1, :-2] - 2 * u + p[1:-1, 2:]) + (p[:-2, 1:-1] - 2 * u + p[2:, 1:-1]))
    return u


def _worker(rank, world, port, ny, nx, ghost, steps, periodic, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from bachelors_b200.distributed import allreduce_max_bits, exchange_halos
    try:
        rng = np.random.default_rng(7)
        full = rng.random((ny, nx))
        sl = slab_layout(ny, world, rank, ghost, periodic)
        lo, hi = neighbours(world, rank, periodic)
        glo = ghost if sl.has_lo else 0
        ghi = ghost if sl.has_hi else 0
        rows = (np.arange(sl.y0 - glo, sl.y0 + sl.ny_local + ghi)) % ny
        field = torch.from_numpy(full[rows].copy())
        other = torch.from_numpy((2 * full[rows]).copy())       # a second field travels in the same batch
        for _ in range(steps):
            new = torch.from_numpy(_standin_step(field.numpy(), ghost, sl.has_lo, sl.has_hi))
            new2 = torch.from_numpy(_standin_step(other.numpy(), ghost, sl.has_lo, sl.has_hi))
            # poison the ghost rows: they must be rewritten by the exchange
            if glo:
                new[:glo] = np.nan; new2[:glo] = np.nan
            if ghi:
                new[-ghi:] = np.nan; new2[-ghi:] = np.nan
            exchange_halos([new, new2], ghost, glo, sl.ny_local, lo, hi)
            field, other = new, new2
        own = field[glo:glo + sl.ny_local].numpy()
        own2 = other[glo:glo + sl.ny_local].numpy()
        # the adaptive solver's reduction: MAX over ranks of the bit patterns of non-negative doubles
        local = np.array([np.abs(own).max(), float(rank + 1) * 1e-300], dtype=np.float64)
        bits = torch.from_numpy(local.view(np.int64).copy())
        allreduce_max_bits(bits)
        np.savez(os.path.join(out, f"r{rank}.npz"), own=own, own2=own2, y0=sl.y0, bits=bits.numpy())
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


@pytest.mark.parametrize("world,periodic", [(2, False), (2, True), (3, False), (3, True)])
def test_halo_exchange_reproduces_single_grid(tmp_path, world, periodic):
    ny, nx, ghost, steps = 37, 11, 4, 3
    mp.spawn(_worker, args=(world, _free_port(), ny, nx, ghost, steps, periodic, str(tmp_path)), nprocs=world, join=True)
    rng = np.random.default_rng(7)
    full = rng.random((ny, nx))
    want = full.copy()
    for _ in range(steps):
        if periodic:
            p = np.concatenate([want[-ghost:], want, want[:ghost]])
            want = _standin_step(p, ghost, True, True)[ghost:-ghost]
        else:
            want = _standin_step(want, ghost, False, False)
    got = np.empty_like(full)
    got2 = np.empty_like(full)
    gmax = 0.0
    for r in range(world):
        z = np.load(tmp_path / f"r{r}.npz")
        y0 = int(z["y0"])
        got[y0:y0 + z["own"].shape[0]] = z["own"]
        got2[y0:y0 + z["own"].shape[0]] = z["own2"]
        gmax = max(gmax, np.abs(z["own"]).max())
        bits = z["bits"]
    assert np.array_equal(got, want)
    assert np.allclose(got2, 2 * want, rtol=1e-13)
    assert bits.view(np.float64)[0] == gmax and bits.view(np.float64)[1] == world * 1e-300
