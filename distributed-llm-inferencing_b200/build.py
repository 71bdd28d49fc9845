"""Build the C-ABI shared library (libdli_b200.so) in-tree with nvcc for sm_100a.

Usage:  python distributed-llm-inferencing_b200/build.py [--force] [--verbose]

The library is the product's only compute path: the Python host fails loudly if
it is missing (see _lib.py).  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libdli_b200.so")
OBJ_DIR = os.path.join(HERE, "build")
SOURCES = [
    "abi.cu",
    "norm.cu",
    "rope_kv.cu",
    "attn_decode.cu",
    "attn_prefill.cu",
    "attn_prefill_tc.cu",
    "gemv.cu",
    "gemv_mma.cu",
    "gemm_tcgen05.cu",
    "elementwise.cu",
    "sample.cu",
]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC"]


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(HERE, "..", "include", "dli.h"))
    return hs


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(src, verbose):
    obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
    path = os.path.join(CSRC, src)
    if not _stale(obj, [path] + _headers()):
        return obj, ""
    cmd = [NVCC, *ARCH, *FLAGS, "-c", path, "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{p.stdout}\n{p.stderr}")
    return obj, p.stderr


def build(force=False, verbose=False):
    os.makedirs(OBJ_DIR, exist_ok=True)
    if force:
        for f in os.listdir(OBJ_DIR):
            os.remove(os.path.join(OBJ_DIR, f))
        if os.path.exists(OUT):
            os.remove(OUT)
    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        results = list(ex.map(lambda s: _compile(s, verbose), SOURCES))
    objs = [o for o, _ in results]
    if verbose:
        for _, log in results:
            if log:
                print(log)
    if _stale(OUT, objs):
        cmd = [NVCC, *ARCH, "-shared", "-o", OUT, *objs, ]
        p = subprocess.run(cmd, capture_output=True, text=True)
        if p.returncode != 0:
            raise RuntimeError(f"link failed:\n{p.stdout}\n{p.stderr}")
    return OUT


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose))
