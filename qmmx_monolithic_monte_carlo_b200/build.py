"""In-tree build of libqmmx_mc.so (the C-ABI CUDA library) for sm_100a.

nvcc cross-compiles without a GPU; the .so is git-ignored but travels with the tree to the GPU box.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
INCLUDE = os.path.join(ROOT, "include")
LIB = os.path.join(PKG, "libqmmx_mc.so")
_OBJ_DIR = os.path.join(PKG, "build")

SOURCES = ["qmmx_mc.cu", "kernels.cuh", "engine.cuh", "scan_lean.cuh", "scan_tile.cuh", "tables.h", "tables_host.cpp", "tables_host.h"]

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "--fmad=false",                       # float64/fp32 decisions must not be contracted (parity contract)
    "-Xcompiler", "-fPIC,-ffp-contract=off",
]
CXX_FLAGS = ["-O2", "-fPIC", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-Wall"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libqmmx_mc.so cannot be built (there is no CPU fallback)")


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + [os.path.join(INCLUDE, "qmmx_mc.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False, out: str = None, defines=()) -> str:
    """Compile csrc/ into libqmmx_mc.so when sources are newer than the library.
    `out` / `defines` build a tuning variant (e.g. -DQMMX_BLOCKS_PER_SM=3) beside the product library."""
    if out is None and not force and not is_stale():
        return LIB
    os.makedirs(_OBJ_DIR, exist_ok=True)
    obj = os.path.join(_OBJ_DIR, "tables_host.o")
    cxx = os.environ.get("CXX", "g++")
    cmd1 = [cxx] + CXX_FLAGS + ["-I", INCLUDE, "-c", os.path.join(CSRC, "tables_host.cpp"), "-o", obj]
    target = out or LIB
    cmd2 = [_nvcc()] + NVCC_FLAGS + ["-D" + d for d in defines] + (["-Xptxas", "-v"] if verbose else []) + [
        "-I", INCLUDE, "-shared", "-o", target + ".tmp", os.path.join(CSRC, "qmmx_mc.cu"), obj]
    for cmd in (cmd1, cmd2):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose:
            sys.stderr.write(r.stderr)
        if r.returncode != 0:
            raise RuntimeError("build failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
    os.replace(target + ".tmp", target)
    return target


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
