#!/usr/bin/env python
"""Markdown table of the metrics that matter from an `ncu --set full` report (first launch of each kernel / grid shape):
usage: python tools/summarize_full.py gpurun_out/<tag>_full.ncu-rep > profiles/<tag>_ncu_full.md   (runs `ncu -i ... --page raw --csv`)"""
import csv
import io
import re
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, body = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
WANT = [("time [us]", "gpu__time_duration.sum", 1e-3), ("DRAM read [MB]", "dram__bytes_read.sum", None), ("DRAM write [MB]", "dram__bytes_write.sum", None),
        ("DRAM throughput %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1), ("L2 throughput %", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1),
        ("L1/TEX throughput %", "l1tex__throughput.avg.pct_of_peak_sustained_active", 1), ("issue active %", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1),
        ("achieved occupancy %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1), ("L1 hit %", "l1tex__t_sector_hit_rate.pct", 1),
        ("L2 hit %", "lts__t_sector_hit_rate.pct", 1), ("registers/thread", "launch__registers_per_thread", 1), ("warp instructions", "smsp__inst_executed.sum", 1),
        ("stall long scoreboard / issue", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", 1),
        ("stall short scoreboard / issue", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", 1),
        ("stall barrier / issue", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", 1),
        ("stall math pipe / issue", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", 1),
        ("stall wait / issue", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", 1),
        ("stall not selected / issue", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", 1),
        ("stall lg throttle / issue", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", 1)]


def val(r, name, scale):
    if name not in col:
        return "-"
    v = r[col[name]].replace(",", "")
    try:
        x = float(v)
    except ValueError:
        return v or "-"
    u = units[col[name]]
    if scale is None:      # bytes -> MB whatever the unit ncu chose
        x *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1.0)
    elif name == "gpu__time_duration.sum":
        x *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(u, 1.0)
    return ("%.0f" % x) if abs(x) >= 1000 else ("%.2f" % x)


seen, cols = set(), []
for r in body:
    name = re.sub(r"^void ", "", r[col["Kernel Name"]]).replace("mf::", "").split("(")[0]
    key = name + " grid" + r[col["Grid Size"]].replace(" ", "")
    if key not in seen:
        seen.add(key); cols.append((key, r))
print("ncu --set full --clock-control none --cache-control none, one 256^3 plume step (tools/profile_step.py window); first launch of each\n"
      "kernel / grid shape.  Source report: %s (not committed).\n" % sys.argv[1])
print("| metric | " + " | ".join(k for k, _ in cols) + " |")
print("|---|" + "---:|" * len(cols))
for label, name, scale in WANT:
    print("| %s | " % label + " | ".join(val(r, name, scale) for _, r in cols) + " |")
