#!/usr/bin/env python
"""Aggregate an ncu cuda,sass source export by code region given as name:file:lo-hi arguments.
Usage: ncu_region_summary.py export.csv name:file:lo-hi [...]"""
import csv
import sys

path = sys.argv[1]
regions = []
for a in sys.argv[2:]:
    name, f, rng = a.split(":")
    lo, hi = rng.split("-")
    regions.append((name, f, int(lo), int(hi)))
cur = None
hdr = None
agg = {}
tot = [0, 0, 0]
for r in csv.reader(open(path)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        i_s, i_i, i_t = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
        continue
    if hdr is None or not r[0].isdigit():
        continue
    try:
        s, i, t = int(r[i_s]), int(r[i_i]), int(r[i_t])
    except ValueError:
        continue
    ln = int(r[0])
    key = cur
    for name, f, lo, hi in regions:
        if f == cur and lo <= ln <= hi:
            key = name
            break
    a = agg.setdefault(key, [0, 0, 0])
    a[0] += s
    a[1] += i
    a[2] += t
    tot[0] += s
    tot[1] += i
    tot[2] += t
print(f"warp instructions {tot[1]}  avg active threads {tot[2] / max(tot[1], 1):.1f}")
for k, (s, i, t) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{k:28s} samples {100 * s / tot[0]:5.1f}%  instr {100 * i / tot[1]:5.1f}%  threads {t / max(i, 1):5.1f}")
