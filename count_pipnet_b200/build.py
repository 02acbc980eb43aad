"""Builds libcpnb200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

`python -m count_pipnet_b200.build` or `build_library()`; `__graft_entry__.build()` calls this.
"""
from __future__ import annotations

import os
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libcpnb200.so")
SOURCES = ["api.cu"]
HEADERS = ["ptx.cuh", "gemm_tcgen05.cuh", "gemm_tma_epi.cuh", "gelu_fast.cuh", "convnext_kernels.cuh", "dwconv_stats.cuh", "mlp_fused.cuh", "dwconv_tma.cuh", "head_kernels.cuh", "eval_kernels.cuh", "layout_kernels.cuh",
           os.path.join("..", "..", "include", "cpn_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
]
LINK_FLAGS = ["-lcudart_static", "-ldl", "-lrt", "-lpthread"]


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or "/usr/local/cuda/bin/nvcc"
    return cand if os.path.exists(cand) else "nvcc"


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB_PATH
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH + ".tmp"] + \
        [os.path.join(CSRC, s) for s in SOURCES] + LINK_FLAGS
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed building libcpnb200.so: " + " ".join(cmd))
    os.replace(LIB_PATH + ".tmp", LIB_PATH)  # a snapshot of the tree never sees a half-written library
    if verbose:
        sys.stderr.write(res.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
