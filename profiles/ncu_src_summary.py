#!/usr/bin/env python
"""Summarise an `ncu --page source --csv --print-source cuda,sass` export: samples / instructions per source line
and the stall-reason mix.  Usage: ncu_src_summary.py file.csv [top_n]"""
import csv
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
cur = None
hdr = None
lines = []
stall_tot = {}
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
        i_s, i_i = hdr.index("# Samples"), hdr.index("Instructions Executed")
        stalls = [(h, k) for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if hdr is None or r[0] in ("-", "") or not r[0].isdigit():
        continue  # sass rows have '-' as line number
    try:
        s, n = int(r[i_s]), int(r[i_i])
    except ValueError:
        continue
    lines.append((s, n, cur, r[0], r[1].strip()[:100]))
    for h, k in stalls:
        try:
            stall_tot[h] = stall_tot.get(h, 0) + int(r[k])
        except ValueError:
            pass
ts, ti = sum(l[0] for l in lines), sum(l[1] for l in lines)
print(f"total samples {ts}  warp instructions {ti}")
T = sum(stall_tot.values()) or 1
print("stall mix:", {k: round(100 * v / T, 1) for k, v in sorted(stall_tot.items(), key=lambda kv: -kv[1])[:8]})
for s, n, f, ln, src in sorted(lines, key=lambda l: -l[0])[:top]:
    print(f"{100 * s / ts:5.1f}% smp {100 * n / ti:5.1f}% ins  {f}:{ln:>4s}  {src}")
