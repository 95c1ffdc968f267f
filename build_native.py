#!/usr/bin/env python
"""Build libdtf_b200.so (all CUDA kernels + the native host runtime) for sm_100a, in-tree.

    python build_native.py [--force] [--verbose]

nvcc cross-compiles without a GPU.  The library lands in distributedtf_b200/lib/ so that it
travels with the repository snapshot to the GPU box (it is git-ignored, not gpurun-ignored).
"""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(ROOT, "csrc")
OUT_DIR = os.path.join(ROOT, "distributedtf_b200", "lib")
# DTF_BUILD_TAG=name builds a side-by-side variant (e.g. DTF_TIMING=1 DTF_BUILD_TAG=timing -> libdtf_b200_timing.so,
# loaded with DTF_LIB=...); the production library has no tag
_TAG = os.environ.get("DTF_BUILD_TAG", "")
OBJ_DIR = os.path.join(ROOT, "build", "obj" + ("_" + _TAG if _TAG else ""))
LIB = os.path.join(OUT_DIR, "libdtf_b200" + ("_" + _TAG if _TAG else "") + ".so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ARCH + ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-I" + os.path.join(ROOT, "include"),
                     "-I" + SRC, "--expt-relaxed-constexpr"] + (["-DDTF_TIMING"] if os.environ.get("DTF_TIMING") else []) + \
    os.environ.get("DTF_NVCC_EXTRA", "").split()


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found")
    return exe


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    headers = [os.path.join(SRC, f) for f in os.listdir(SRC) if f.endswith((".cuh", ".h"))]
    headers += [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
    sources = sorted(f for f in os.listdir(SRC) if f.endswith((".cu", ".cpp")))
    jobs = []
    objs = []
    for s in sources:
        src = os.path.join(SRC, s)
        obj = os.path.join(OBJ_DIR, os.path.splitext(s)[0] + ".o")
        objs.append(obj)
        if not force and _newer(obj, [src] + headers):
            continue
        cmd = [nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-x", "cu", "-c", src, "-o", obj]
        jobs.append((s, cmd))

    def run(job):
        name, cmd = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        return name, r.returncode, r.stdout + r.stderr

    if jobs:
        with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for name, rc, log in ex.map(run, jobs):
                if verbose or rc != 0:
                    print(f"--- {name}\n{log}")
                if rc != 0:
                    raise RuntimeError(f"nvcc failed on {name}")
    if jobs or force or not _newer(LIB, objs):
        cmd = [nvcc()] + ARCH + ["-shared", "-o", LIB] + objs + ["-lcudart", "-lcuda"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            print(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return LIB


BENCH_SRC = os.path.join(ROOT, "bench", "csrc")
BENCH_LIB = os.path.join(OUT_DIR, "libdtf_bench.so")


def build_bench() -> str:
    """Micro-benchmark kernels (bench/csrc): their own library, never linked into libdtf_b200.so."""
    os.makedirs(OUT_DIR, exist_ok=True)
    srcs = sorted(os.path.join(BENCH_SRC, f) for f in os.listdir(BENCH_SRC) if f.endswith(".cu"))
    if _newer(BENCH_LIB, srcs):
        return BENCH_LIB
    cmd = [nvcc()] + NVCC_FLAGS + ["-shared", "-o", BENCH_LIB] + srcs + ["-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        print(r.stdout + r.stderr)
        raise RuntimeError("bench library build failed")
    return BENCH_LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
    if "--bench" in sys.argv:
        print(build_bench())
