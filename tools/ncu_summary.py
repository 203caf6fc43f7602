#!/usr/bin/env python
"""Summarises an `ncu --set full` report (read here, no GPU needed) into profiles/<name>.md and records the per-launch
DRAM traffic of the kernel in profiles/traffic.json (the source of bench.py's roofline.traffic).

    python tools/ncu_summary.py gpurun_out/prof_rk4_f64.ncu-rep r01_rk4_f64_2048 explicit-rk4_f64_2048
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.sum", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_lsu.sum", "smsp__inst_executed.sum",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__maximum_warps_per_active_cycle_pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
    "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct", "smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct",
    "smsp__warp_issue_stalled_wait_per_warp_active.pct", "smsp__warp_issue_stalled_not_selected_per_warp_active.pct",
    "smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.avg.per_cycle_elapsed",
]


def main(rep, name, traffic_key=None):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    header, units = rows[0], rows[1]
    launches = rows[2:]
    col = {h: i for i, h in enumerate(header)}
    lines = [f"# ncu summary: {name}", "", f"source: `{rep}` (ncu --set full --clock-control none; per-launch values are cold-cache and serialised)", ""]
    traffic = []
    for r in launches:
        kname = r[col["Kernel Name"]]
        lines.append(f"## launch {r[col['ID']]}: `{kname[:160]}`  grid {r[col.get('Grid Size', 0)] if 'Grid Size' in col else ''} block {r[col.get('Block Size', 0)] if 'Block Size' in col else ''}")
        lines.append("")
        lines.append("| metric | value | unit |")
        lines.append("|---|---|---|")
        for k in KEYS:
            if k in col:
                lines.append(f"| {k} | {r[col[k]]} | {units[col[k]]} |")
        extra = [h for h in header if ("pipe_fp64" in h or "stalled" in h and "per_warp_active" in h) and h not in KEYS]
        for k in extra[:40]:
            lines.append(f"| {k} | {r[col[k]]} | {units[col[k]]} |")
        lines.append("")

        def num(k):
            try:
                v = float(r[col[k]].replace(",", ""))
                u = units[col[k]].lower()
                mult = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
                return v * mult
            except Exception:
                return None
        rd, wr = num("dram__bytes_read.sum"), num("dram__bytes_write.sum")
        if rd is not None and wr is not None:
            traffic.append(rd + wr)
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", name + ".md"), "w") as f:
        f.write("\n".join(lines) + "\n")
    if traffic_key and traffic:
        path = os.path.join(ROOT, "profiles", "traffic.json")
        d = json.load(open(path)) if os.path.exists(path) else {}
        d[traffic_key] = sum(traffic) / len(traffic)
        d[traffic_key + "__source"] = f"profiles/{name}.md (dram__bytes_read.sum + dram__bytes_write.sum, mean of {len(traffic)} launches)"
        json.dump(d, open(path, "w"), indent=1)
    print("\n".join(lines[:70]))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
