"""In-tree build of ``libhive_b200.so`` (hand-written sm_100a kernels + the C ABI of include/hive_b200.h).

``nvcc`` cross-compiles without a GPU; the resulting ``.so`` is git-ignored but travels to the GPU box with the
snapshot.  Usage: ``python -m hive_b200.build [--force] [--verbose]``.
"""
from __future__ import annotations

import glob
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "libhive_b200.so")

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    # no implicit FMA contraction: every fused multiply-add in the kernels is written explicitly, because parity with
    # the oracle depends on where roundings happen (nvcc contracts even __fadd2_rn(__fmul2_rn(..)) otherwise)
    "-fmad=false",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O2", "-shared", "-Xptxas", "-warn-spills",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libhive_b200.so")


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _inputs():
    return sources() + sorted(glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) +
                              glob.glob(os.path.join(HERE, "..", "include", "*.h")))


def source_hash() -> str:
    """sha256 (first 16 hex digits) over the names and contents of every file the library is compiled from, plus the
    compiler flags.  The same string is compiled into the library (``hv_build_id``), so a binary can be attributed to
    the sources that are lying next to it -- bench.py prints both."""
    h = hashlib.sha256(" ".join(NVCC_FLAGS).encode())
    for path in _inputs():
        h.update(os.path.basename(path).encode())
        with open(path, "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def built_id() -> str | None:
    """Build id of the library on disk (sidecar written next to it by ``build``), None if there is none."""
    try:
        with open(LIB_PATH + ".id") as f:
            return f.read().strip()
    except OSError:
        return None


def is_stale() -> bool:
    return not os.path.exists(LIB_PATH) or built_id() != source_hash()


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB_PATH
    bid = source_hash()
    cmd = [_nvcc()] + NVCC_FLAGS + [f'-DHV_BUILD_ID="{bid}"'] + (["-Xptxas", "-v"] if verbose else []) + \
        ["-o", LIB_PATH] + sources()
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed building libhive_b200.so")
    with open(LIB_PATH + ".id", "w") as f:
        f.write(bid + "\n")
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
