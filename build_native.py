"""Builds the native sm_100a library in-tree.

    python build_native.py [--force] [--verbose]

Compiles every ``csrc/*.cu`` with ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` into
``build/*.o`` (in parallel, skipped when up to date) and links
``distributed-machine-learning-pipeline_b200/_native/libdmlp_b200.so``. nvcc cross-compiles without
a GPU, so this also runs on the CPU-only dev box. ptxas resource usage (``-Xptxas -v``) is saved
to ``build/ptxas_<name>.log``.
"""
from __future__ import annotations

import argparse
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(ROOT, "csrc")
BUILD = os.path.join(ROOT, "build")
PKG_DIR = os.path.join(ROOT, "distributed-machine-learning-pipeline_b200")
OUT_SO = os.path.join(PKG_DIR, "_native", "libdmlp_b200.so")

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall",
    "-Xptxas", "-v", "--expt-relaxed-constexpr",
]


def find_nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; the native library needs the CUDA 12.9 toolkit")
    return cand


def _newest_header_mtime() -> float:
    t = 0.0
    for d in (CSRC, os.path.join(ROOT, "include")):
        for f in os.listdir(d):
            if f.endswith((".h", ".cuh", ".hpp")):
                t = max(t, os.path.getmtime(os.path.join(d, f)))
    return t


def _compile_one(nvcc: str, src: str, obj: str, verbose: bool) -> None:
    cmd = [nvcc, *ARCH_FLAGS, *NVCC_FLAGS, "-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    name = os.path.splitext(os.path.basename(src))[0]
    with open(os.path.join(BUILD, f"ptxas_{name}.log"), "w") as f:
        f.write(res.stdout)
        f.write(res.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{res.stdout}\n{res.stderr}")


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile + link if anything changed. Returns the path of the shared library."""
    nvcc = find_nvcc()
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(os.path.dirname(OUT_SO), exist_ok=True)
    srcs = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))
    hdr_t = _newest_header_mtime()
    jobs, objs = [], []
    for src in srcs:
        obj = os.path.join(BUILD, os.path.splitext(os.path.basename(src))[0] + ".o")
        objs.append(obj)
        stale = force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_t)
        if stale:
            jobs.append((src, obj))
    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            futs = [ex.submit(_compile_one, nvcc, s, o, verbose) for s, o in jobs]
            for f in futs:
                f.result()
    need_link = bool(jobs) or not os.path.exists(OUT_SO) or any(
        os.path.getmtime(o) > os.path.getmtime(OUT_SO) for o in objs)
    if need_link:
        cmd = [nvcc, *ARCH_FLAGS, "-shared", "-o", OUT_SO, *objs, "-lcudart_static", "-lpthread", "-ldl", "-lrt"]
        if verbose:
            print(" ".join(cmd), flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return OUT_SO


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
    sys.exit(0)
