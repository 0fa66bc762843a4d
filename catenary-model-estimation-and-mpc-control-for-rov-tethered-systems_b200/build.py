"""In-tree build of libcatenary_b200.so for sm_100a (nvcc cross-compiles without a GPU).

    python build.py [--force] [--verbose]

1. codegen.py: artifacts/model.json -> csrc/gen/dyn_eqs.cuh (PySR equations + scaler constants),
   artifacts/icdf_table.json -> csrc/gen/icdf_table.cuh (sampler table)
2. nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -c, one process per translation unit, in parallel
   -> csrc/build/*.o -> nvcc -shared -> libcatenary_b200.so
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
LIB = os.path.join(_HERE, "libcatenary_b200.so")
# translation units: (object name, source, extra defines).  k_mpc_rollout is instantiated in its own units, one per
# (compute type, equation source), so the heavy part of the build runs in parallel (2 min 11 s serial on 8 cores).
UNITS = [("catenary_b200.o", "catenary_b200.cu", [])] + [
    (f"mpc_inst_{t}_{k}.o", "mpc_inst.cu", [f"-DCAT_INST_T={t}", f"-DCAT_INST_KIND={k}"])
    for t in ("double", "float") for k in (0, 1, 2)]
HEADERS = ["cat_host.cuh", "cat_math.cuh", "cat_kernels_geom.cuh", "cat_kernels_dyn.cuh", "cat_kernels_mpc.cuh",
           "cat_mpc_launch.cuh", "cat_kernels_ctrl.cuh", "cat_kernels_expr.cuh", "cat_kernels_feat.cuh",
           os.path.join("gen", "dyn_eqs.cuh"), os.path.join("gen", "icdf_table.cuh"),
           os.path.join("..", "..", "include", "catenary_b200.h")]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "-Xcompiler", "-fvisibility=hidden", "-DCAT_BUILD_SHARED"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=...)")


def _newest_dep(src):
    deps = [os.path.join(CSRC, f) for f in [src] + HEADERS] + [os.path.abspath(__file__)]
    return max(os.path.getmtime(d) for d in deps if os.path.exists(d))


def _compile(unit, force, verbose, objdir=OBJ, extra=()):
    obj, src, defs = unit
    out = os.path.join(objdir, obj)
    if not force and os.path.exists(out) and os.path.getmtime(out) >= _newest_dep(src):
        return obj, 0, ""
    cmd = ([_nvcc()] + FLAGS + defs + list(extra) + (["-Xptxas", "-v"] if verbose else []) +
           ["-c", os.path.join(CSRC, src), "-o", out])
    res = subprocess.run(cmd, capture_output=True, text=True)
    return obj, res.returncode, res.stdout + res.stderr


def build(force=False, verbose=False, extra=(), lib=LIB, objdir=OBJ):
    """``extra`` / ``lib`` / ``objdir``: a variant build for A/B timing (tools/ab.sh), e.g.
    build(extra=["-DCAT_MPC_ASYNC_F64=1"], lib="ab/1_async64.so", objdir="/tmp/obj_async64")."""
    sys.path.insert(0, _HERE)
    try:
        import codegen
    finally:
        sys.path.pop(0)
    codegen.generate()
    os.makedirs(objdir, exist_ok=True)
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(len(UNITS), os.cpu_count() or 2)) as pool:
        results = list(pool.map(lambda u: _compile(u, force, verbose, objdir, extra), UNITS))
    for obj, rc, log in results:
        if verbose or rc != 0:
            sys.stderr.write(f"--- {obj}\n{log}")
        if rc != 0:
            raise RuntimeError(f"nvcc failed on {obj}:\n" + log[-4000:])
    objs = [os.path.join(objdir, u[0]) for u in UNITS]
    if force or not os.path.exists(lib) or any(os.path.getmtime(o) > os.path.getmtime(lib) for o in objs):
        cmd = [_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-cudart", "static", "-o", lib] + objs
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("link failed:\n" + (res.stdout + res.stderr)[-4000:])
    return lib


if __name__ == "__main__":
    _extra = [a for a in sys.argv[1:] if a.startswith("-D")]
    _lib = next((a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--lib=")), LIB)
    _obj = next((a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--objdir=")), OBJ)
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, extra=_extra, lib=_lib, objdir=_obj))
