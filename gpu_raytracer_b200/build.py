"""Builds libgpu_raytracer_b200.so in-tree with nvcc for sm_100a.

``-fmad=false`` is part of the numerical contract: nvcc never
contracts a*b+c, the only fused operations are explicit fmaf() calls, which is
what makes the kernels bit-comparable with the -ffp-contract=off CPU oracle.
"""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libgpu_raytracer_b200.so")
SOURCES = ["api.cu", "scene_gen.cu", "ray_gen.cu", "morton.cu", "radix_sort.cu",
           "refit.cu", "traverse.cu", "wide.cu", "scene_io.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false",
              "-std=c++17", "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared",
              "-diag-suppress", "177"]


def _stale() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, "..", "include", "gpu_raytracer.h"))
    deps.append(os.path.abspath(__file__))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source into one shared library; returns its path."""
    if not force and not _stale():
        return SO
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, *NVCC_FLAGS, "-ccbin", "/usr/bin/g++"] + (["-Xptxas", "-v"] if verbose else [])
    cmd += ["-o", SO] + [os.path.join(CSRC, s) for s in SOURCES]
    env = dict(os.environ)
    env.pop("CXX", None)
    env.pop("CC", None)
    res = subprocess.run(cmd, cwd=CSRC, env=env, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return SO


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
