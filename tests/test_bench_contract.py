"""The measurement contract of bench.py: one JSON line per run with the keys the driver reads. The reference arm (CPU oracle
port on the host cores) runs without a GPU; the CUDA arm is a GPU test with a short timed region."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
             "dtype", "data", "config", "e2e", "cpu_baseline"}


def run_bench(*args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=900, cwd=ROOT, env=e)
    assert res.returncode == 0, res.stderr[-3000:]
    lines = [l for l in res.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, res.stdout[-2000:]                         # ONE JSON line
    return json.loads(lines[0])


def test_reference_arm_line_without_a_gpu():
    d = run_bench("--impl", "reference", "--steps", "2", "--warmup", "1", env={"OMP_NUM_THREADS": "1"})   # torchrun's export must not bind it
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["metric"] == "GLUPS" and d["higher_is_better"] is True and d["vs_baseline"] is None and d["n_gpus"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["steps"] == 2
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] == d["value"] and cb["cores"] == len(os.sched_getaffinity(0))
    assert "whole 2048x2048 grid" in cb["sample"]                       # the same configuration as the CUDA arm at N = 1, not a window
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]
    # both arms print the same `config` (the driver compares them): what the CUDA arm would print for this command line
    import argparse
    import bench
    args = argparse.Namespace(solver="explicit-rk4", dtype="f64", size=0, wide=False, no_overlap=False)
    cfg, desc = bench.workload(args, 1)
    assert d["config"] == bench.config_dict(args, 1, cfg.params, desc, 8)


def test_reference_arm_other_ranks_exit_without_work():
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, timeout=300, cwd=ROOT, env=dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1"))
    assert res.returncode == 0 and res.stdout.strip() == ""


@pytest.mark.gpu
def test_cuda_arm_line():
    d = run_bench("--steps", "30", "--warmup", "3", "--cpu-budget", "1", "--e2e-steps", "3")
    assert BASE_KEYS | {"roofline", "clocks", "gpu_launches"} <= set(d) and d.get("impl", "cuda") == "cuda"
    assert d["metric"] == "GLUPS" and d["dtype"] == "f64" and d["n_gpus"] == 1 and d["scaling"] == "weak"
    assert "config_explicit-rk4_2048.ini" in d["config"]["workload"]
    cells = 2048 * 2048
    assert abs(d["value"] - cells / (d["ms_per_step"] * 1e-3) / 1e9) <= 1e-6 * d["value"]
    assert d["gpu_launches"] == 30                                       # one kernel per explicit-rk4 step
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-12 and 0 < r["frac"] < 1
    assert abs(r["achieved"] - 32 * cells / (d["ms_per_step"] * 1e-3) / 1e9) <= 1e-6 * r["achieved"]     # 4 s bytes per cell, fp64
    assert r["traffic"] is None or r["traffic"] > 0
    sec = {x["bound"]: x for x in d["roofline_secondary"]}               # the ceilings that bind RK4: FP64 pipe and issue slots
    assert set(sec) == {"fp64_pipe", "issue"} and all(0 < x["frac"] < 1 and abs(x["frac"] - x["achieved"] / x["peak"]) < 1e-12 for x in sec.values())
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] > 0 and cb["cores"] >= 1 and cb["sample"]
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] == 2 * cells * 8 and e["d2h_bytes_per_step"] == 2 * cells * 8
    assert e["value"] < d["value"]                                       # host buffers in and out every step cannot beat the resident number
    c = d["clocks"]
    assert c["sm_mhz"] > 0 and c["sm_max_mhz"] >= c["sm_mhz"] and isinstance(c["reasons"], list)
