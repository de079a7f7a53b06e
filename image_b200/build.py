"""Builds image_b200/lib/libb200img.so in-tree with nvcc for sm_100a.

Usage: python -m image_b200.build [--force]
The .so is git-ignored but travels to the GPU box with the repository snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "lib", "obj")
SO = os.path.join(HERE, "lib", "libb200img.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CUFLAGS = ["-O3", "-std=c++17", "-lineinfo", "--use_fast_math=false", "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math,-Wall",
           "--fmad=false", "-Xptxas", "-v"]
# --fmad=false: contraction is never left to the compiler; kernels fuse explicitly (fmaf) where
# the FAST flag allows it, so STRICT paths cannot be fused behind our back.
CUFLAGS = [f for f in CUFLAGS if f != "--use_fast_math=false"]
CXXFLAGS = ["-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-fno-fast-math", "-Wall"]


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(HERE, "..", "include", "b200img.h"))
    return hs


def _compile(src: str, force: bool) -> str:
    obj = os.path.join(OBJ, src + ".o")
    path = os.path.join(CSRC, src)
    deps = [path] + _headers()
    if not force and os.path.exists(obj) and all(os.path.getmtime(d) <= os.path.getmtime(obj) for d in deps):
        return obj
    if src.endswith(".cu"):
        cmd = [NVCC] + ARCH + CUFLAGS + ["-c", path, "-o", obj]
    else:
        cmd = ["g++"] + CXXFLAGS + ["-c", path, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log = os.path.join(OBJ, src + ".log")
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"compile failed: {src}\n{r.stdout}\n{r.stderr}")
    return obj


def build(force: bool = False) -> str:
    srcs = _sources()
    deps = [os.path.join(CSRC, f) for f in srcs] + _headers()
    if not force and os.path.exists(SO) and all(os.path.getmtime(d) <= os.path.getmtime(SO) for d in deps):
        return SO  # up to date (the object directory need not exist: it does not travel to the GPU box)
    os.makedirs(OBJ, exist_ok=True)
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, force), srcs))
    if force or not os.path.exists(SO) or any(os.path.getmtime(o) > os.path.getmtime(SO) for o in objs):
        cmd = [NVCC] + ARCH + ["-shared", "-o", SO] + objs + ["-Xlinker", "-z,defs", "-lcudart_static", "-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed\n{r.stdout}\n{r.stderr}")
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
