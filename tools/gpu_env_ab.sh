#!/bin/bash
# A/B of environment settings on one library: bash tools/gpu_env_ab.sh <lib.so|default> "ENV1=.. ENV2=.." "..." ...
lib=$1; shift
mkdir -p gpurun_out
if [ "$lib" != default ]; then export EGR_B200_LIB=$lib; fi
for envs in "$@"; do
  env $envs python bench.py --steps 20 --warmup 5 --no-cpu --no-train > gpurun_out/ab.json 2> gpurun_out/ab.err || tail -5 gpurun_out/ab.err
  python - "$envs" <<PY
import json, sys
d=json.load(open("gpurun_out/ab.json"))
r=d["roofline"]
print("%-40s value %.1f e2e %.1f ms/step %.3f | %s | serial %s" % (sys.argv[1], d["value"], d["e2e"]["value"], d["ms_per_step"], r["phase_ms"], r["serial_kernel_ms"]))
PY
done
