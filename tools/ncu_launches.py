#!/usr/bin/env python
"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel:
    python tools/ncu_launches.py gpurun_out/launches.csv > profiles/rNN_bench_launches.md"""
import csv
import sys
from collections import defaultdict

path = sys.argv[1]
rows = []
with open(path, newline="") as fh:
    lines = [l for l in fh if l.startswith('"')]
rd = csv.reader(lines)
hdr = next(rd)
ix = {k: i for i, k in enumerate(hdr)}
tot = defaultdict(lambda: [0, 0.0, "", ""])
for r in rd:
    if len(r) < len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    name = r[ix["Kernel Name"]]
    val = float(r[ix["Metric Value"]].replace(",", ""))
    unit = r[ix["Metric Unit"]]
    us = val / 1e3 if unit in ("ns", "nsecond") else (val * 1e3 if unit in ("ms", "msecond") else val)
    t = tot[name]
    t[0] += 1
    t[1] += us
    t[2], t[3] = r[ix["Block Size"]], r[ix["Grid Size"]]
total = sum(t[1] for t in tot.values())
print(f"ncu launch list ({path}); gpu__time_duration.sum, --clock-control none; cold-cache, serialised\n")
print("| kernel | launches | total us | share | block | grid (last) |\n|---|---|---|---|---|---|")
for name, t in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{name[:110]}` | {t[0]} | {t[1]:.1f} | {100 * t[1] / total:.1f}% | {t[2]} | {t[3]} |")
