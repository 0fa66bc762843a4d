"""In-tree build of libcatenary_b200.so for sm_100a (nvcc cross-compiles without a GPU).

    python build.py [--force] [--verbose]

1. codegen.py: artifacts/model.json -> csrc/gen/dyn_eqs.cuh (PySR equations + scaler constants)
2. nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -> libcatenary_b200.so
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB = os.path.join(_HERE, "libcatenary_b200.so")
SOURCES = ["catenary_b200.cu"]
HEADERS = ["cat_math.cuh", "cat_kernels_geom.cuh", "cat_kernels_dyn.cuh", "cat_kernels_mpc.cuh", "cat_kernels_ctrl.cuh", "cat_kernels_expr.cuh",
           os.path.join("gen", "dyn_eqs.cuh"), os.path.join("..", "..", "include", "catenary_b200.h")]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=...)")


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    sys.path.insert(0, _HERE)
    try:
        import codegen
    finally:
        sys.path.pop(0)
    codegen.generate()
    if not force and not _stale():
        return LIB
    cmd = [_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "-shared", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-cudart", "static",
           "-DCAT_BUILD_SHARED", "-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stderr[-4000:])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
