#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total, share, average.
Usage: ncu_launch_summary.py launches.csv"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[start]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot = collections.OrderedDict()
for r in rows[start + 1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    v = v / 1e3 if r[ui] == "ns" else v * 1e3 if r[ui] == "ms" else v
    name = r[ki].split("(")[0][:64]
    tot.setdefault(name, [0, 0.0])
    tot[name][0] += 1
    tot[name][1] += v
T = sum(v[1] for v in tot.values())
print(f"{'kernel':64s} {'n':>5s} {'total ms':>10s} {'share':>7s} {'avg us':>10s}")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1])[:16]:
    print(f"{k:64s} {v[0]:5d} {v[1] / 1e3:10.3f} {100 * v[1] / T:6.1f}% {v[1] / v[0]:10.1f}")
