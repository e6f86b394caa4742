"""Builds librm_b200.so (sm_100a) in-tree with nvcc.  ``python -m ray_marching_b200.csrc.build``"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SOURCES = ["rm_host.cu", "rm_forward.cu", "rm_backward.cu", "rm_probe.cu"]
LIB = os.path.join(HERE, "librm_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def source_digest() -> str:
    """sha1 over every kernel source, the public header and the flags: freshness by CONTENT (file
    times do not survive the copy to the GPU box)."""
    import hashlib
    h = hashlib.sha1(" ".join(FLAGS).encode())
    deps = sorted(os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith((".cu", ".cuh", ".h")))
    deps.append(os.path.join(HERE, "..", "..", "include", "rm_b200.h"))
    for path in deps:
        with open(path, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def is_fresh() -> bool:
    stamp = LIB + ".digest"
    return (os.path.exists(LIB) and os.path.exists(stamp)
            and open(stamp).read().strip() == source_digest())


def build(force: bool = False, verbose: bool = False) -> str:
    srcs = [os.path.join(HERE, s) for s in SOURCES]
    if not force and is_fresh():
        return LIB
    if not os.path.exists(NVCC):
        raise RuntimeError(f"nvcc not found at {NVCC}; cannot build {LIB} (missing, or older than its sources)")
    tmp = LIB + f".tmp{os.getpid()}"
    cmd = [NVCC, "-shared", "-o", tmp] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + srcs
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    os.replace(tmp, LIB)
    with open(LIB + ".digest", "w") as f:
        f.write(source_digest() + "\n")
    if verbose:
        print(res.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
