"""In-tree build of the CUDA extension: ``csrc/*.cu`` -> ``csrc/libegr_b200.so`` for sm_100a.

No torch, no pybind: the product of the build is a plain C-ABI shared library (``include/egr_b200.h``)
loaded with ctypes.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# EGR_B200_LIB: load an already built variant of the library instead (kernel experiments: a second build of the same
# sources with other -D switches, see tools/build_variant.py); it is never rebuilt automatically.
_LIB_OVERRIDE = os.environ.get("EGR_B200_LIB")
LIB_PATH = _LIB_OVERRIDE or os.path.join(CSRC, "libegr_b200.so")
SOURCES = ["egr_capi.cu", "egr_bvh.cu", "egr_trace.cu", "egr_backward.cu", "egr_prune.cu", "egr_edit.cu", "egr_denoise.cu"]
HEADERS = ["egr_common.cuh", "egr_trace.cuh", os.path.join("..", "..", "include", "egr_b200.h"),
           os.path.join("..", "..", "include", "egr_detmath.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr", "-Xcudafe", "--diag_suppress=177",
] + os.environ.get("EGR_NVCC_EXTRA", "").split()


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the CUDA extension cannot be built")


def needs_build() -> bool:
    if _LIB_OVERRIDE:
        return False
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False, out: str | None = None, extra_flags=()) -> str:
    """``out`` / ``extra_flags``: build a variant (other -D switches) to another path, objects kept apart."""
    if out is not None:
        return _build_to(out, list(extra_flags), verbose)
    if _LIB_OVERRIDE:
        return LIB_PATH
    if not force and not needs_build():
        return LIB_PATH
    return _build_to(LIB_PATH, [], verbose)


def _build_to(lib_path: str, extra_flags, verbose: bool) -> str:
    nvcc = _nvcc()
    tag = "" if lib_path == os.path.join(CSRC, "libegr_b200.so") else "_" + os.path.basename(lib_path).replace(".so", "")
    objdir = os.path.join(CSRC, "build" + tag)
    os.makedirs(objdir, exist_ok=True)
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, *extra_flags, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        objs.append(obj)
    for src, pr in procs:
        out, _ = pr.communicate()
        if pr.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{out.decode(errors='replace')}")
    cmd = [nvcc, "-shared", "-o", lib_path, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-ldl"]
    subprocess.check_call(cmd)
    return lib_path


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
