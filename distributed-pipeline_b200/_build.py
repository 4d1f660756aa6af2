"""Builds ``libdpk.so`` in-tree with nvcc for sm_100a.

Run as ``python distributed-pipeline_b200/_build.py`` or through
``__graft_entry__.build()``.  The shared object sits next to this file so it
travels with the repository snapshot; nothing is cached outside the tree.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdpk.so")
SOURCES = ["dpk.cu", "dpk_sampler.cu", "dpk_noalloc.cu", "dpk_selftest.cu", "dpk_nvls.cu"]
HEADERS = ["dpk_kernels.cuh", "dpk_internal.h", os.path.join("..", "..", "include", "dpk.h")]


def _nccl_include():
    """Header of the NCCL copy torch loads (2.28.x); system header as a fallback."""
    for base in sys.path:
        cand = os.path.join(base, "nvidia", "nccl", "include")
        if os.path.exists(os.path.join(cand, "nccl.h")):
            return cand
    if os.path.exists("/usr/include/nccl.h"):
        return "/usr/include"
    raise RuntimeError("nccl.h not found (needed for types only)")


def nvcc_path():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; libdpk.so cannot be built")
    return nvcc


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not is_stale():
        return LIB
    cmd = [
        nvcc_path(),
        "-gencode", "arch=compute_100a,code=sm_100a",
        "-lineinfo", "-O3", "-std=c++17",
        # torch's kernels round after every op; FMAs are written out by hand
        "-fmad=false",
        "-Xcompiler", "-fPIC", "-shared",
        "-I", _nccl_include(),
    ]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += [os.path.join(CSRC, s) for s in SOURCES]
    cmd += ["-o", LIB + ".tmp", "-ldl"]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + proc.stdout + proc.stderr)
    os.replace(LIB + ".tmp", LIB)
    if verbose:
        print(proc.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
