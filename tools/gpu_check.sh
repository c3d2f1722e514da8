#!/bin/bash
# Standard GPU check of a kernel change (run under gpurun): parity tests, then the bench line.
# Usage: bash tools/gpu_check.sh <tag> [extra bench args]
TAG=${1:-check}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; tail -4 gpurun_out/${TAG}_tests.log
python bench.py --steps 20 --warmup 5 --no-cpu "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || tail -5 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json"))
r=d["roofline"]
print("value %.1f e2e %.1f ms/step %.3f | phase %s | serial %s | dom %s frac %.3f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], r["phase_ms"], r["serial_kernel_ms"], r["kernel"], r["frac"]))
PY
