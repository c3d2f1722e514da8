#!/bin/bash
# Profiling recipe of this repo (run under gpurun, one GPU).  Usage: bash profiles/run_ncu.sh <round-tag>
# 1) plain run must exit 0; 2) launch list of this library's kernels (per-launch device time, cold cache, serialised: compare SHARES);
# 3) one --set full capture of one whole iteration (launch order: step0, backward(step0) on the side stream, step1, step2, backward(step2), backward(step1)).
set -e
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-train"
$CMD > $OUT/${TAG}_plain.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_trace_step|k_backward' -s 18 -c 6 -o $OUT/${TAG}_prof $CMD > $OUT/${TAG}_ncu_full.log 2>&1
tail -2 $OUT/${TAG}_plain.log
