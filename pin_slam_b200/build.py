"""Builds libpinq.so (hand-written sm_100a CUDA kernels behind the C-ABI of include/pinq.h) in-tree.

nvcc cross-compiles without a GPU; the .so is git-ignored but travels with the tree to the GPU box.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libpinq.so")
STAMP = os.path.join(HERE, ".libpinq.stamp")
SOURCES = ["table.cu", "index.cu", "sort.cu", "brick.cu", "query.cu", "query_bwd.cu", "gn.cu", "mapping.cu", "peer.cu", "pool.cu", "frame.cu", "mc.cu"]
HEADERS = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "query_common.cuh"), os.path.join(CSRC, "mma.cuh"), os.path.join(CSRC, "brick.cuh"), os.path.join(CSRC, "decode_tc.cuh"), os.path.join(CSRC, "tc.cuh"), os.path.join(CSRC, "bwd_tc.cuh"),
           os.path.join(ROOT, "include", "pinq.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "--shared", "-cudart", "shared",
    # exact-arithmetic contract of the index path: IEEE div/sqrt, no fast-math anywhere
    "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
]


def _digest() -> str:
    h = hashlib.sha256()
    for p in [os.path.join(CSRC, s) for s in SOURCES] + HEADERS + [__file__]:
        with open(p, "rb") as f:
            h.update(f.read())
    h.update((" ".join(NVCC_FLAGS) + os.environ.get("PINQ_NVCC_EXTRA", "")).encode())
    return h.hexdigest()


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def is_current() -> bool:
    """libpinq.so exists and was built from exactly the sources / flags in the tree."""
    if not (os.path.exists(LIB) and os.path.exists(STAMP)):
        return False
    with open(STAMP) as f:
        return f.read().strip() == _digest()


def build(force: bool = False, verbose: bool = False) -> str:
    digest = _digest()
    if not force and is_current():
        return LIB
    cmd = [nvcc_path()] + NVCC_FLAGS + ["-I" + os.path.join(ROOT, "include"), "-I" + CSRC]
    cmd += os.environ.get("PINQ_NVCC_EXTRA", "").split()  # kernel experiments only (scripts/variants.py)
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += [os.path.join(CSRC, s) for s in SOURCES] + ["-o", LIB]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout + proc.stderr)
        raise RuntimeError("nvcc failed building libpinq.so")
    if verbose:
        sys.stderr.write(proc.stderr)
    with open(STAMP, "w") as f:
        f.write(digest)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
