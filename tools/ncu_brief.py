#!/usr/bin/env python3
"""Print the handful of raw ncu metrics that decide where a kernel is bound (one column per .ncu-rep).

    python tools/ncu_brief.py gpurun_out/prof_a.ncu-rep gpurun_out/prof_b.ncu-rep > profiles/r2_x_brief.txt
"""
import csv
import io
import subprocess
import sys

KEYS = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
STALL = "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio"
for s in ["long_scoreboard", "short_scoreboard", "math_pipe_throttle", "mio_throttle", "lg_throttle", "barrier", "wait",
          "not_selected", "no_instruction", "dispatch_stall", "branch_resolving"]:
    KEYS.append(STALL % s)
cols = []
for rep in sys.argv[1:]:
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    cols.append({h: (v, u) for h, u, v in zip(hdr, units, vals)})
print("%-40s" % "metric" + "".join("%22s" % r.split("prof_")[-1].replace(".ncu-rep", "")[:21] for r in sys.argv[1:]))
for k in KEYS:
    short = k.replace("smsp__average_warps_issue_stalled_", "stall_").replace("_per_issue_active.ratio", "")
    short = short.replace(".avg.pct_of_peak_sustained_active", "%").replace(".avg.pct_of_peak_sustained_elapsed", "%")
    print("%-40s" % short[:39] + "".join("%22s" % ((c.get(k, ("-", ""))[0][:16] + " " + c.get(k, ("", ""))[1][:5]).strip()[:21]) for c in cols))
