#!/bin/bash
# A/B of library variants on the named configs: bash tools/gpu_ab_configs.sh "<configs>" <lib.so|default> ...
cfgs=$1; shift
mkdir -p gpurun_out
for lib in "$@"; do
  if [ "$lib" = default ]; then unset EGR_B200_LIB; else export EGR_B200_LIB=$lib; fi
  for c in $cfgs; do
    python bench.py --config $c --steps 6 --warmup 3 --no-cpu --no-train > gpurun_out/abc.json 2> gpurun_out/abc.err || tail -5 gpurun_out/abc.err
    python - "$lib" $c <<PY
import json, sys
d=json.load(open("gpurun_out/abc.json"))
r=d["roofline"]
print("%-26s %s value %.1f e2e %.1f ms/step %.3f | %s | serial %s" % (sys.argv[1], sys.argv[2], d["value"], d["e2e"]["value"], d["ms_per_step"], r["phase_ms"], r["serial_kernel_ms"]))
PY
  done
done
