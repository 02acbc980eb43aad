"""Builds libcpnb200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

`python -m count_pipnet_b200.build [--force] [-v] [--dev]` or `build_library()`; `__graft_entry__.build()` calls this.
Every translation unit of csrc/ is compiled to an object (in parallel, only when it or a header it includes changed) and
the objects are linked into the shared library. `--dev` adds -DCPN_DEV_KNOBS (work-skipping timing experiments and
clock64 traces; never part of the product build) and writes libcpnb200_dev.so instead.
"""
from __future__ import annotations

import concurrent.futures
import os
import re
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
OBJ_DIR = os.path.join(PKG_DIR, "build_tmp")
LIB_PATH = os.path.join(PKG_DIR, "libcpnb200.so")
DEV_LIB_PATH = os.path.join(PKG_DIR, "libcpnb200_dev.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]
LINK_FLAGS = ["-shared", "-Xcompiler", "-fPIC", "-lcudart_static", "-ldl", "-lrt", "-lpthread"]
_INCLUDE = re.compile(r'^\s*#\s*include\s*"([^"]+)"', re.M)


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or "/usr/local/cuda/bin/nvcc"
    return cand if os.path.exists(cand) else "nvcc"


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _deps(path: str, seen=None) -> set:
    """The file and every header it includes with quotes, transitively."""
    seen = set() if seen is None else seen
    path = os.path.normpath(path)
    if path in seen or not os.path.exists(path):
        return seen
    seen.add(path)
    with open(path) as f:
        for inc in _INCLUDE.findall(f.read()):
            _deps(os.path.join(os.path.dirname(path), inc), seen)
    return seen


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _obj(src: str, dev: bool) -> str:
    return os.path.join(OBJ_DIR, os.path.splitext(src)[0] + ("_dev.o" if dev else ".o"))


def needs_build(dev: bool = False) -> bool:
    lib = DEV_LIB_PATH if dev else LIB_PATH
    return any(_stale(lib, _deps(os.path.join(CSRC, s))) for s in sources())


def _compile(src: str, dev: bool, verbose: bool):
    cmd = [_nvcc()] + NVCC_FLAGS + (["-DCPN_DEV_KNOBS"] if dev else []) + (["-Xptxas", "-v"] if verbose else []) + \
        ["-c", os.path.join(CSRC, src), "-o", _obj(src, dev)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    return src, cmd, res


def build_library(force: bool = False, verbose: bool = False, dev: bool = False) -> str:
    lib = DEV_LIB_PATH if dev else LIB_PATH
    if not force and not needs_build(dev):
        return lib
    os.makedirs(OBJ_DIR, exist_ok=True)
    todo = [s for s in sources() if force or _stale(_obj(s, dev), _deps(os.path.join(CSRC, s)))]
    with concurrent.futures.ThreadPoolExecutor(max_workers=max(1, min(len(todo), os.cpu_count() or 1))) as pool:
        for src, cmd, res in pool.map(lambda s: _compile(s, dev, verbose), todo):
            if res.returncode != 0:
                sys.stderr.write(res.stdout + res.stderr)
                raise RuntimeError("nvcc failed compiling " + src + ": " + " ".join(cmd))
            if verbose:
                sys.stderr.write(res.stderr)
    cmd = [_nvcc()] + ["-gencode", "arch=compute_100a,code=sm_100a"] + LINK_FLAGS + ["-o", lib + ".tmp"] + \
        [_obj(s, dev) for s in sources()]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed linking libcpnb200.so: " + " ".join(cmd))
    os.replace(lib + ".tmp", lib)  # a snapshot of the tree never sees a half-written library
    return lib


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv, dev="--dev" in sys.argv))
