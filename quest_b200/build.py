"""Build recipe for the product library (hand-written CUDA for sm_100a) -> quest_b200/lib/libQuEST.so.

Explicit nvcc commands, in-tree outputs (the .so travels to the GPU box with the repo snapshot).  nvcc
cross-compiles without a GPU.  ``-fmad=false`` keeps the amplitude arithmetic free of FMA contraction so results
equal the reference CPU build bit for bit; these kernels are HBM-bound, so it costs nothing.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CU_SOURCES = ["backend_cuda.cu", "qb_comm.cu", "qb_noalloc.cu", "qb_io.cu"]
CXX_SOURCES = ["quest_api.cpp", "quest_common.cpp", "quest_qasm.cpp", "quest_validation.cpp", "mt19937.cpp"]


def _newer(target, deps):
    if not os.path.isfile(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_product(prec: int = 2, force: bool = False, verbose: bool = False, defines=(), variant: str = "") -> str:
    """`defines` / `variant`: an experimental build with extra -D flags goes to lib/variants/lib<variant>.so (used by
    tools/ab_variants.sh through QUEST_B200_LIB); the product itself is always built without any."""
    os.makedirs(LIBDIR, exist_ok=True)
    out = os.path.join(LIBDIR, "libQuEST.so" if prec == 2 else "libQuEST_f32.so")
    if variant:
        os.makedirs(os.path.join(LIBDIR, "variants"), exist_ok=True)
        out = os.path.join(LIBDIR, "variants", f"lib{variant}.so")
        force = True
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
    if not force and not _newer(out, deps):
        return out
    cmd = [NVCC, *ARCH, "-O3", "-std=c++17", "-lineinfo", "-fmad=false", "-shared", "-cudart", "static",
           "-Xcompiler", "-fPIC,-fvisibility=default,-Wall,-Wno-unused-function",
           "-Xlinker", "-Bsymbolic", f"-DQuEST_PREC={prec}",
           "-I" + os.path.join(ROOT, "include"), "-I" + CSRC]
    cmd += [f"-D{d}" for d in defines]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += [os.path.join(CSRC, f) for f in CU_SOURCES + CXX_SOURCES]  # .cu through nvcc, .cpp straight to g++
    cmd += ["-o", out, "-ldl", "-lpthread"]
    print(" ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)
    return out


if __name__ == "__main__":
    if "--variant" in sys.argv:  # python build.py --variant NAME -DFOO=1 -DBAR=2
        name = sys.argv[sys.argv.index("--variant") + 1]
        build_product(2, verbose="-v" in sys.argv, defines=[a[2:] for a in sys.argv if a.startswith("-D")], variant=name)
        sys.exit(0)
    build_product(2, force="--force" in sys.argv, verbose="-v" in sys.argv)
    if "--f32" in sys.argv:
        build_product(1, force="--force" in sys.argv)
