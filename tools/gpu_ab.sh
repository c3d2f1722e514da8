#!/bin/bash
# A/B of library variants on the bench line (run under gpurun): bash tools/gpu_ab.sh <lib.so|default> ...
mkdir -p gpurun_out
for lib in "$@"; do
  if [ "$lib" = default ]; then unset EGR_B200_LIB; else export EGR_B200_LIB=$lib; fi
  python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/ab.json 2> gpurun_out/ab.err || tail -5 gpurun_out/ab.err
  python - "$lib" <<PY
import json, sys
d=json.load(open("gpurun_out/ab.json"))
r=d["roofline"]
print("%-28s value %.1f e2e %.1f ms/step %.3f | %s | serial %s" % (sys.argv[1], d["value"], d["e2e"]["value"], d["ms_per_step"], r["phase_ms"], r["serial_kernel_ms"]))
PY
done
