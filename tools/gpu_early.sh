#!/bin/bash
mkdir -p gpurun_out
for lib in tools/tmp/libegr_tol64.so tools/tmp/libegr_tol96.so tools/tmp/libegr_tol128.so; do
 if [ "$lib" = default ]; then unset EGR_B200_LIB; else export EGR_B200_LIB=$lib; fi
 for c in c3 c4; do
  for e in 1e-5; do
    python bench.py --config $c --steps 6 --warmup 3 --no-cpu --no-train --early-stop $e > gpurun_out/es.json 2> gpurun_out/es.err || tail -5 gpurun_out/es.err
    python - $c $e $lib <<PY
import json, sys
d=json.load(open("gpurun_out/es.json"))
r=d["roofline"]
print("%s %s early=%s value %.1f ms/step %.3f | %s" % (sys.argv[3][-14:], sys.argv[1], sys.argv[2], d["value"], d["ms_per_step"], r["serial_kernel_ms"]))
PY
  done
 done
done
