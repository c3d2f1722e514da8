"""Build a variant of the CUDA library with extra nvcc flags (kernel experiments, instrumentation):
    python tools/build_variant.py tools/tmp/libegr_clocks.so -DEGR_TASK_CLOCKS
then run anything with EGR_B200_LIB=<that path>.  tools/tmp/ is git-ignored (it still travels to the GPU box)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import egr_b200  # noqa: E402

out = os.path.abspath(sys.argv[1])
os.makedirs(os.path.dirname(out), exist_ok=True)
print(egr_b200.build.build(out=out, extra_flags=sys.argv[2:], verbose=False))
