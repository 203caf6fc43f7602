"""SlabRunner (bachelors_b200/distributed.py) on real devices: world size 1 in-process, world size 2 under torchrun when
the box has two GPUs. The slabs must reproduce the single-grid CUDA result bit for bit."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from bachelors_b200 import solver as S
from bachelors_b200._abi import PF_F64, default_params
from bachelors_b200.testing import seeded_fields
from conftest import ROOT

pytestmark = pytest.mark.gpu

WORKER = r'''
import os, sys, json
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from bachelors_b200._abi import PF_F64, default_params
from bachelors_b200.testing import seeded_fields
from bachelors_b200.distributed import SlabRunner
from bachelors_b200.slabs import slab_rows
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
solver, bc, nx, ny, steps, out = sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6]), sys.argv[7]
p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.2, Phi_boundary=bc, T_boundary=bc, Phi_tolerance=1e-7, T_tolerance=1e-7)
F0, U0 = seeded_fields(nx, ny, kind="interface")
y0, y1 = slab_rows(ny, world, rank)
r = SlabRunner(p, PF_F64, rank=rank, world=world, device=local)
r.upload_global(F0[y0:y1], U0[y0:y1])
t = 0.0
for _ in range(steps):
    t += r.step()
F, U = r.download_owned()
np.savez(os.path.join(out, f"r{rank}.npz"), F=F, U=U, y0=y0, t=t)
r.close()
dist.barrier(); dist.destroy_process_group()
'''


def single_grid(solver, bc, nx, ny, steps):
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.2, Phi_boundary=bc, T_boundary=bc, Phi_tolerance=1e-7, T_tolerance=1e-7)
    F0, U0 = seeded_fields(nx, ny, kind="interface")
    with S.Solver(nx, ny, PF_F64) as s:
        F, U, tF, tU = s.field_from(F0), s.field_from(U0), s.field_alloc(), s.field_alloc()
        Fo, Uo, done, t, _ = s.run(p, F, U, tF, tU, steps)
        return s.download(Fo), s.download(Uo), t


@pytest.mark.parametrize("solver", ["explicit", "explicit-rk4", "explicit-rk4-adaptive"])
def test_slab_runner_world1_equals_solver(solver):
    import torch
    from bachelors_b200.distributed import SlabRunner
    nx, ny, steps = 300, 200, 5
    p = default_params(solver=solver, nx=nx, ny=ny, L0=nx / 128.0, S=0.2, Phi_tolerance=1e-7, T_tolerance=1e-7)
    F0, U0 = seeded_fields(nx, ny, kind="interface")
    r = SlabRunner(p, PF_F64, rank=0, world=1, device=0)
    r.upload_global(F0, U0)
    t = sum(r.step() for _ in range(steps))
    F, U = r.download_owned()
    r.close()
    wF, wU, wt = single_grid(solver, "neumann", nx, ny, steps)
    assert np.array_equal(F, wF) and np.array_equal(U, wU) and abs(t - wt) < 1e-18


@pytest.mark.parametrize("solver,bc", [("explicit", "neumann"), ("explicit-rk4", "neumann"), ("explicit-rk4", "periodic"),
                                       ("explicit-rk4-adaptive", "neumann"), ("explicit-rk4-adaptive", "periodic")])
def test_slab_runner_two_gpus(tmp_path, solver, bc):
    if S.lib().pf_device_count() < 2:
        pytest.skip("needs two GPUs")
    nx, ny, steps = 520, 301, 6
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", str(script), ROOT, solver, bc, str(nx), str(ny), str(steps), str(tmp_path)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-3000:]
    wF, wU, wt = single_grid(solver, bc, nx, ny, steps)
    for r in range(2):
        z = np.load(tmp_path / f"r{r}.npz")
        y0 = int(z["y0"])
        assert np.array_equal(z["F"], wF[y0:y0 + z["F"].shape[0]]), (solver, bc, r)
        assert np.array_equal(z["U"], wU[y0:y0 + z["U"].shape[0]])
        assert abs(float(z["t"]) - wt) < 1e-18


@pytest.mark.parametrize("bc", ["neumann", "periodic"])
def test_slab_runner_two_gpus_semi_implicit(tmp_path, bc):
    """The semi-implicit solver over NCCL: pf_si_* phases, all-gathered partial sums, one-row exchange of the search
    direction per CG iteration. Against the single grid only the summation tree differs: rel-L2 <= 1e-10 (fp64)."""
    if S.lib().pf_device_count() < 2:
        pytest.skip("needs two GPUs")
    nx, ny, steps = 520, 301, 5
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29543", str(script), ROOT, "semi-implicit", bc, str(nx), str(ny), str(steps), str(tmp_path)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-3000:]
    wF, wU, wt = single_grid("semi-implicit", bc, nx, ny, steps)
    got_F = np.concatenate([np.load(tmp_path / f"r{r}.npz")["F"] for r in range(2)])
    got_U = np.concatenate([np.load(tmp_path / f"r{r}.npz")["U"] for r in range(2)])
    assert got_F.shape == wF.shape
    assert np.linalg.norm(got_F - wF) / np.linalg.norm(wF) <= 1e-10
    assert np.linalg.norm(got_U - wU) / np.linalg.norm(wU) <= 1e-10
    assert not np.array_equal(got_F, seeded_fields(nx, ny, kind="interface")[0])


FULL_WORKER = r'''
import os, sys, json, hashlib
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from bachelors_b200._abi import PF_F64
from bachelors_b200.app import initial_of
from bachelors_b200.config import read_config
from bachelors_b200.distributed import SlabRunner
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
root, solver, n, steps, out = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
cfg = read_config(os.path.join(root, "benchmark_configs", "config_%s_2048.ini" % solver), ["[simulation]\nscale = %g" % (n / 512.0)])
p = cfg.params
assert p.nx == n and p.ny == n
r = SlabRunner(p, PF_F64, rank=rank, world=world, device=local)
r.solver.init_fields(p, initial_of(cfg), r.F.data_ptr(), r.U.data_ptr())      # ghost rows included: no first exchange needed
ts = [r.step() for _ in range(steps)]
r.sync()
f0 = r.solver.first_owned_row
shas = [hashlib.sha256(t[f0:f0 + r.layout.ny_local].cpu().numpy().tobytes()).hexdigest() for t in (r.F, r.U)]
json.dump({"y0": r.layout.y0, "rows": r.layout.ny_local, "sha": shas, "ts": ts}, open(os.path.join(out, "r%d.json" % rank), "w"))
r.close()
dist.barrier(); dist.destroy_process_group()
'''


@pytest.mark.parametrize("solver", ["explicit-rk4-adaptive"])
def test_full_size_16384_slabs_equal_the_single_gpu_run(tmp_path, solver):
    """BASELINE.json configs[4] at its full size: the synthetic 16384^2 adaptive run cut into y-slabs over every GPU of the box
    gives, slab by slab, the SHA-256 of the single-GPU run of the same grid (fields and accepted step sizes)."""
    import hashlib
    from bachelors_b200.app import initial_of
    from bachelors_b200.config import read_config
    G = S.lib().pf_device_count()
    G = 8 if G >= 8 else (4 if G >= 4 else (2 if G >= 2 else 1))
    if G < 2:
        pytest.skip("needs at least two GPUs")
    n, steps = 16384, 3
    script = tmp_path / "full_worker.py"
    script.write_text(FULL_WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % G, "--master-addr", "127.0.0.1",
           "--master-port", "29547", str(script), ROOT, solver, str(n), str(steps), str(tmp_path)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stderr[-3000:]
    parts = [json.load(open(tmp_path / ("r%d.json" % r))) for r in range(G)]
    cfg = read_config(os.path.join(ROOT, "benchmark_configs", "config_%s_2048.ini" % solver), ["[simulation]\nscale = %g" % (n / 512.0)])
    p = cfg.params
    with S.Solver(n, n, PF_F64) as s:
        F, U, tF, tU = (s.field_alloc() for _ in range(4))
        s.init_fields(p, initial_of(cfg), F, U)
        q, t, ts = p.copy(), 0.0, []
        for i in range(steps):
            q.iter, q.time = i, t
            info = s.step(q, F, U, tF, tU)
            t += info.dt_advanced; ts.append(info.dt_advanced)
            F, tF, U, tU = tF, F, tU, U
        wF, wU = s.download(F), s.download(U)
    assert wF.shape == (n, n) and wF.sum() > 0
    for part in parts:
        assert part["ts"] == ts
        y0, rows = part["y0"], part["rows"]
        assert hashlib.sha256(wF[y0:y0 + rows].tobytes()).hexdigest() == part["sha"][0]
        assert hashlib.sha256(wU[y0:y0 + rows].tobytes()).hexdigest() == part["sha"][1]
    assert sum(part["rows"] for part in parts) == n


# ---- the headless host on slabs (bachelors_b200/app_slabs.py) ---------------------------------------------------------------
def run_app_under_torchrun(nproc, config, override, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % nproc, "--master-addr", "127.0.0.1",
           "--master-port", str(port), "-m", "bachelors_b200.app", os.path.join(ROOT, "benchmark_configs", config), "--set", override]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    return res.stdout


@pytest.mark.parametrize("solver", ["explicit-rk4", "explicit-rk4-adaptive", "semi-implicit"])
def test_headless_host_on_slabs_writes_the_single_gpu_files(tmp_path, solver):
    """`torchrun -m bachelors_b200.app` on every GPU of the box (one process on a one-GPU box: the same code path with
    world size 1) against the single-GPU host: same snapshot schedule, every maps_NNNN.bin byte for byte the same file
    for the explicit solvers, within 1e-10 for the semi-implicit one; config.ini and stats.csv next to them."""
    from bachelors_b200.app import run_config
    n = min(S.lib().pf_device_count(), 8)
    config = "config_%s_256.ini" % solver
    over = "[simulation]\nstop_after = 0.0005\n[snapshot]\ntimes = 5\nfolder = %s"
    (tmp_path / "one").mkdir(); (tmp_path / "many").mkdir()
    one = run_config(os.path.join(ROOT, "benchmark_configs", config), PF_F64, overrides=[over % (tmp_path / "one")])
    out = run_app_under_torchrun(n, config, over % (tmp_path / "many"), 29571)
    assert "%i iters, %i snapshots" % (one["iters"], one["snapshots"]) in out
    folders = os.listdir(tmp_path / "many")
    assert len(folders) == 1                                          # all ranks agreed on one folder name
    many = tmp_path / "many" / folders[0]
    names = sorted(os.listdir(many))
    assert names == sorted(os.listdir(one["folder"])) and "config.ini" in names and "stats.csv" in names
    maps = [x for x in names if x.startswith("maps_")]
    assert len(maps) == one["snapshots"] + 1
    for name in maps:
        a = open(os.path.join(one["folder"], name), "rb").read()
        b = open(many / name, "rb").read()
        assert len(a) == len(b)
        if solver != "semi-implicit" or n == 1:
            assert a == b, name
        else:
            assert a[:112] == b[:112]                                 # header: sizes, dx, dy, time, iter, names
            fa, fb = np.frombuffer(a[112:]), np.frombuffer(b[112:])
            assert np.linalg.norm(fa - fb) <= 1e-10 * np.linalg.norm(fa)
