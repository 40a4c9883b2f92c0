#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel launches / total time / share.
usage: python tools/summarize_launches.py gpurun_out/launches.csv > profiles/<round>_launches.md"""
import csv
import re
import sys
from collections import OrderedDict

rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
agg = OrderedDict()
for r in csv.DictReader(rows):
    if r["Metric Name"] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"^void ", "", r["Kernel Name"])
    name = re.sub(r"^mf::", "", name).split("(")[0].split("<")[0]
    ns = float(r["Metric Value"].replace(",", ""))
    a = agg.setdefault(name, [0, 0.0, r["Block Size"], r["Grid Size"]])
    a[0] += 1
    a[1] += ns
tot = sum(a[1] for a in agg.values())
print("| kernel | launches | total ms | share | avg us | block | grid (first launch) |")
print("|---|---:|---:|---:|---:|---|---|")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("| %s | %d | %.3f | %.1f%% | %.1f | %s | %s |" % (k, a[0], a[1] / 1e6, 100 * a[1] / tot, a[1] / a[0] / 1e3, a[2], a[3]))
print("\ntotal %.3f ms over %d launches" % (tot / 1e6, sum(a[0] for a in agg.values())))
