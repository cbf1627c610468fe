"""Builds libaten_b200.so (the CUDA kernels + C ABI) in-tree for sm_100a with nvcc.

    python -m aten_b200.build [--force]

nvcc cross-compiles without a GPU.  -fmad=false is required: the traversal arithmetic must round
exactly like the reference's x86 build ("Bit-exactness")."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "atn_kernels.cu")
import glob
# every file the translation unit can include: a stale library would be compared with a fresh restatement
DEPS = sorted(glob.glob(os.path.join(HERE, "csrc", "*")) + glob.glob(os.path.join(HERE, "..", "include", "*.h")))
OUT = os.path.join(HERE, "libaten_b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared", "-cudart", "shared"]


def nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def build(force: bool = False, verbose: bool = False, defines=(), out: str = OUT) -> str:
    """``defines``/``out``: build a kernel variant next to the product library (tuning experiments)."""
    if not force and os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in DEPS):
        return out
    cmd = [nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [f"-D{d}" for d in defines] + ["-o", out, SRC]
    env = dict(os.environ)
    env.pop("CXX", None)     # the image exports CXX=/opt/gcc/bin/g++, which nvcc must not pick up
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
    if verbose or r.returncode != 0:
        print(r.stdout)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building libaten_b200.so")
    return out


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv or "--verbose" in sys.argv)
    print(OUT)
