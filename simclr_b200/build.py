"""Builds libntxent_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m simclr_b200.build            # build if sources are newer than the library
    python -m simclr_b200.build --force
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libntxent_b200.so")
SOURCES = ["ntx_api.cu", "ntx_l2norm.cu", "ntx_simt.cu", "ntx_finalize.cu", "ntx_tc.cu", "ntx_project.cu", "ntx_project_bwd.cu", "ntx_larc.cu", "ntx_augment.cu"]
HEADERS = ["ntx_common.cuh", "ntx_umma.cuh", os.path.join("..", "..", "include", "ntxent_b200.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",   # explicit: -arch=sm_100a would add a plain compute_100 pass
    "-O3", "-lineinfo", "-std=c++17", "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC", "-shared",               # static cudart: no dependence on which libcudart is loaded
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libntxent_b200.so cannot be built")


def stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, defines=(), out: str = LIB) -> str:
    """`defines`/`out` build instrumented variants (e.g. -DNTX_TIMING) next to the product library."""
    if not force and not defines and not stale():
        return LIB
    cmd = [find_nvcc(), *NVCC_FLAGS, *[f"-D{d}" for d in defines]]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += [os.path.join(CSRC, f) for f in SOURCES] + ["-o", out]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    if verbose:
        print(proc.stderr)
    return out


if __name__ == "__main__":
    if "--timing" in sys.argv:
        print(build(force=True, defines=("NTX_TIMING",), out=os.path.join(PKG, "libntxent_b200_timing.so")))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
