"""Build the in-tree native libraries with nvcc for sm_100a.

    python -m serverless_learn_b200.build [--force]

Outputs (git-ignored, but shipped to the GPU box with the tree):
    serverless_learn_b200/libsl_gpu.so     CUDA kernels + the C ABI of include/sl_gpu.h
    serverless_learn_b200/libsl_roles.so   host-side role cores (include/sl_roles.h)
    serverless_learn_b200/sl_job           example program: the three roles in one C++ process
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]

TARGETS = {
    "libsl_gpu.so": dict(
        srcs=["sl_gpu.cu"], deps=["kernels.cuh", "kernels_wide.cuh", "kernels_multi.cuh", "kernels_cluster.cuh", "../../include/sl_gpu.h"],
        cmd=lambda out, srcs: [NVCC, *ARCH, "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
                               "-shared", "-cudart", "static", *os.environ.get("SL_NVCC_DEFS", "").split(),
                               "-o", out, *srcs, "-ldl"]),
    "libsl_roles.so": dict(
        srcs=["roles.cc"], deps=["roles.h", "wire.h", "../../include/sl_roles.h",
                                 "../../include/sl_gpu.h"],
        cmd=lambda out, srcs: ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-o", out,
                               *srcs, "-L" + HERE, "-lsl_gpu", "-Wl,-rpath,$ORIGIN", "-lpthread"]),
    # example program: the three roles in one C++ process (examples/sl_job.cc)
    "sl_job": dict(
        srcs=["../../examples/sl_job.cc"], deps=["roles.h", "../../include/sl_roles.h", "../../include/sl_gpu.h"],
        cmd=lambda out, srcs: ["g++", "-O2", "-std=c++17", "-Wall", "-o", out, *srcs, "-L" + HERE, "-lsl_roles",
                               "-lsl_gpu", "-Wl,-rpath,$ORIGIN", "-lpthread"]),
}


def _stale(out: str, files: list[str]) -> bool:
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.exists(f) and os.path.getmtime(f) > t for f in files)


def build(force: bool = False, verbose: bool = False) -> None:
    for name, spec in TARGETS.items():
        srcs = [os.path.join(CSRC, s) for s in spec["srcs"]]
        if not all(os.path.exists(s) for s in srcs):
            continue
        deps = srcs + [os.path.normpath(os.path.join(CSRC, d)) for d in spec["deps"]]
        out = os.path.join(HERE, name)
        if not (force or _stale(out, deps)):
            continue
        cmd = spec["cmd"](out, srcs)
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        subprocess.run(cmd, check=True, cwd=CSRC)


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
