#!/usr/bin/env python
"""Per-kernel table from an ncu --csv capture with several metrics (tools/profile_step.py window):
usage: python tools/summarize_step.py gpurun_out/step.csv > profiles/<round>_ncu_step.md"""
import csv
import re
import sys
from collections import OrderedDict, defaultdict

rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
per = {}
for r in csv.DictReader(rows):
    d = per.setdefault(int(r["ID"]), {"name": r["Kernel Name"], "grid": r["Grid Size"]})
    d[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
agg = OrderedDict()
for i, d in sorted(per.items()):
    n = re.sub(r"^void ", "", d["name"])
    n = re.sub(r"^mf::", "", n).split("(")[0]
    n = re.sub(r"<.*", "", n)
    a = agg.setdefault(n + " grid" + d["grid"].replace(" ", ""), defaultdict(float))
    a["n"] += 1
    for k, v in d.items():
        if k not in ("name", "grid"):
            a[k] += v
T = "gpu__time_duration.sum"
tot = sum(a[T] for a in agg.values())
print("one step, %d launches, sum of kernel durations %.3f ms (ncu serialises launches: no PDL overlap; --cache-control none)\n" % (
    sum(a["n"] for a in agg.values()), tot / 1e6))
print("| kernel / grid | launches | total us | avg us | DRAM MB/launch | issue active % | warps active % | L1 hit % | L2 hit % | warp inst/launch |")
print("|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][T]):
    n = a["n"]
    print("| %s | %d | %.1f | %.1f | %.2f | %.1f | %.1f | %.1f | %.1f | %.0f |" % (
        k, n, a[T] / 1e3, a[T] / 1e3 / n, (a["dram__bytes_read.sum"] + a["dram__bytes_write.sum"]) / n / 1e6,
        a["smsp__issue_active.avg.pct_of_peak_sustained_active"] / n, a["sm__warps_active.avg.pct_of_peak_sustained_active"] / n,
        a["l1tex__t_sector_hit_rate.pct"] / n, a["lts__t_sector_hit_rate.pct"] / n, a["smsp__inst_executed.sum"] / n))
