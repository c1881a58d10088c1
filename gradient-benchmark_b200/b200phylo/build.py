"""Build ``libb200phylo.so`` in-tree with nvcc for sm_100a (``python -m b200phylo.build``).

nvcc cross-compiles without a GPU.  The library lands next to this file so it travels
with the source tree; it is git-ignored.
"""
from __future__ import annotations

import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.normpath(os.path.join(_HERE, "..", "csrc"))
_OUT = os.path.join(_HERE, "libb200phylo.so")
_SOURCES = ["b2p_api.cu", "b2p_timetree.cu", "b2p_noalloc.cu"]
_DEPS = sorted(f for f in os.listdir(_CSRC) if f.endswith((".cu", ".cuh", ".h"))) + [
    os.path.join("..", "..", "include", "b200phylo.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "--shared", "-Xcompiler", "-fPIC",
]


def needs_build() -> bool:
    if not os.path.exists(_OUT):
        return True
    t = os.path.getmtime(_OUT)
    return any(os.path.getmtime(os.path.join(_CSRC, d)) > t for d in _DEPS)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return _OUT
    nvcc = os.environ.get("NVCC", "nvcc")
    extra = os.environ.get("B2P_NVCC_EXTRA", "").split()  # e.g. -DB2P_TRACE_CLOCKS (debug build)
    cmd = [nvcc, *NVCC_FLAGS, *extra, "-o", _OUT, *[os.path.join(_CSRC, s) for s in _SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    subprocess.check_call(cmd, cwd=_CSRC)
    return _OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
