"""Builds libgran_b200/liblgb200.so with nvcc for sm_100a (see libgran_b200/csrc/Makefile)."""
from __future__ import annotations

import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "liblgb200.so")


def build(verbose: bool = False, jobs: int = 8) -> str:
    """Compile every CUDA translation unit (-gencode arch=compute_100a,code=sm_100a -lineinfo) and link the
    C-ABI shared library in-tree.  nvcc cross-compiles without a GPU."""
    cmd = ["make", "-C", CSRC, f"-j{jobs}"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise RuntimeError("building liblgb200.so failed")
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("make succeeded but liblgb200.so is missing")
    return LIB_PATH
