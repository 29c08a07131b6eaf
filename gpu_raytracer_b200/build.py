"""Builds libgpu_raytracer_b200.so in-tree with nvcc for sm_100a.

``-fmad=false`` is part of the numerical contract: nvcc never
contracts a*b+c, the only fused operations are explicit fmaf() calls, which is
what makes the kernels bit-comparable with the -ffp-contract=off CPU oracle.

Every ``csrc/*.cu`` is its own translation unit: the files are compiled to
objects side by side (one nvcc process each, only the stale ones) and linked into
one shared library.
"""
from __future__ import annotations

import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
SO = os.path.join(HERE, "libgpu_raytracer_b200.so")
SOURCES = ["api.cu", "scene_gen.cu", "ray_gen.cu", "morton.cu", "radix_sort.cu",
           "refit.cu", "traverse.cu", "wide.cu", "wide8.cu", "compact.cu", "scene_io.cu", "multi.cu", "noalloc.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false",
              "-std=c++17", "-Xcompiler", "-fPIC,-fvisibility=hidden",
              "-diag-suppress", "177"]


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hs.append(os.path.join(HERE, "..", "include", "gpu_raytracer.h"))
    hs.append(os.path.abspath(__file__))
    return hs


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source (stale ones only) and link one shared library; returns its path."""
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    env = dict(os.environ)
    env.pop("CXX", None)
    env.pop("CC", None)
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _headers()
    jobs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, s[:-3] + ".o")
        if force or _newer(obj, [src] + hdrs):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        cmd = [nvcc, *NVCC_FLAGS, "-ccbin", "/usr/bin/g++"] + (["-Xptxas", "-v"] if verbose else [])
        cmd += ["-c", "-o", obj, src]
        return src, subprocess.run(cmd, cwd=CSRC, env=env, capture_output=True, text=True)

    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            for src, res in ex.map(compile_one, jobs):
                if res.returncode != 0:
                    raise RuntimeError(f"nvcc failed on {src}:\n" + res.stdout + res.stderr)
                if verbose:
                    print(f"==== {os.path.basename(src)}\n{res.stderr}")
    objs = [os.path.join(OBJ, s[:-3] + ".o") for s in SOURCES]
    if jobs or force or _newer(SO, objs):
        cmd = [nvcc, "-shared", "-ccbin", "/usr/bin/g++", "-gencode", "arch=compute_100a,code=sm_100a",
               "-o", SO] + objs
        res = subprocess.run(cmd, cwd=CSRC, env=env, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("link failed:\n" + res.stdout + res.stderr)
    return SO


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
