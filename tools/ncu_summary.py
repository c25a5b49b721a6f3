#!/usr/bin/env python3
"""Summarise ncu captures brought back in gpurun_out/ into profiles/ (tracked).

    python tools/ncu_summary.py r1 median sd quantiles mean

Writes profiles/<tag>_<name>_raw.csv (selected raw metrics of the captured launch),
profiles/<tag>_launches.csv (copy of the bench.py launch list) and updates profiles/traffic.json with
the dram bytes of each captured launch plus the algorithmic bytes of that launch (bench.py scales the
ratio to its own launch size and says so).
"""
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = (
    "Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_dynamic",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg.per_second",
    "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
)
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}


def main():
    tag, names = sys.argv[1], sys.argv[2:]
    out_dir = os.path.join(ROOT, "profiles")
    os.makedirs(out_dir, exist_ok=True)
    tj = os.path.join(out_dir, "traffic.json")
    traffic = json.load(open(tj)) if os.path.exists(tj) else {}
    for name in names:
        rep = os.path.join(ROOT, "gpurun_out", f"prof_{name}_{tag}.ncu-rep")
        raw = os.path.join(ROOT, "gpurun_out", f"raw_{name}_{tag}.csv")  # converted on the GPU box (64 MiB limit)
        if os.path.exists(raw):
            txt = open(raw).read()
        elif os.path.exists(rep):
            txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        else:
            print("missing", rep)
            continue
        rows = list(csv.reader(io.StringIO(txt)))
        hdr, units = rows[0], rows[1]
        with open(os.path.join(out_dir, f"{tag}_{name}_raw.csv"), "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["launch", "metric", "unit", "value"])
            for li, vals in enumerate(rows[2:]):
                d = {}
                for h, u, v in zip(hdr, units, vals):
                    if h in KEEP:
                        w.writerow([li, h, u, v])
                        d[h] = (u, v)
                rd = float(d["dram__bytes_read.sum"][1]) * UNIT[d["dram__bytes_read.sum"][0]]
                wr = float(d["dram__bytes_write.sum"][1]) * UNIT[d["dram__bytes_write.sum"][0]]
                meta = {}
                mp = os.path.join(ROOT, "gpurun_out", f"plain_{name}_{tag}.log")
                if os.path.exists(mp):
                    meta["plain_run"] = open(mp).read().strip()[:300]
                traffic[name] = {"tag": tag, "kernel": d["Kernel Name"][1], "dram_read_bytes": rd,
                                 "dram_write_bytes": wr, "grid": d["Grid Size"][1], "block": d["Block Size"][1],
                                 "duration_ms_under_ncu": float(d["gpu__time_duration.sum"][1]), **meta}
        print("wrote", name)
    ll = os.path.join(ROOT, "gpurun_out", f"launches_{tag}.csv")
    if os.path.exists(ll):
        shutil.copy(ll, os.path.join(out_dir, f"{tag}_launches.csv"))
    with open(tj, "w") as f:
        json.dump(traffic, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
