#!/usr/bin/env python
"""Key metrics of every captured launch in an `ncu -i X.ncu-rep --page raw --csv` export."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed.avg.per_cycle_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "smsp__inst_executed.sum", "lts__t_bytes.sum"]
idx = [hdr.index(w) if w in hdr else None for w in want]
for r in rows[2:]:
    print("---")
    for w, i in zip(want, idx):
        if i is not None:
            print(f"  {w}: {r[i]} [{units[i]}]")
