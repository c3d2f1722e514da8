#!/bin/bash
# Multi-GPU bench lines (run under gpurun --gpus N): bash tools/gpu_multi.sh <N> <tag> [bench args]
N=$1; TAG=$2; shift; shift
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 20 --warmup 5 "$@" 2> gpurun_out/${TAG}.err | grep '^{' > gpurun_out/${TAG}.json || tail -5 gpurun_out/${TAG}.err
python - gpurun_out/${TAG}.json <<PY
import json, sys
d=json.load(open(sys.argv[1]))
t=d.get("train") or {}
print("N=%d value %.1f e2e %.1f ms/step %.3f | %s | train: iter/s %.1f views/s %.1f ms/iter %.3f %s" % (d["n_gpus"], d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["phase_ms"], t.get("iter_per_s",0), t.get("views_per_s",0), t.get("ms_per_iter",0), t.get("phase_ms")))
PY
