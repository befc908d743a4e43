"""Builds librt_b200.so (the CUDA kernels + the C ABI of include/rt.h) in-tree for sm_100a.

    python -m real_time_ray_tracer_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU.  Flags that define the arithmetic: --fmad=false (fused
multiply-add only where the source says so), -prec-div/-prec-sqrt=true, -ftz=false."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "librt_b200.so")
SOURCES = ["rt_kernels.cu", "rt_ao_phased.cu", "rt_ao_pulled.cu", "rt_capi.cu"]
DEPS = SOURCES + ["rt_kernels.cuh", "rt_math.cuh", "rt_trace_common.cuh", os.path.join("..", "..", "include", "rt.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC", "-shared",
]


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build(force: bool = False, verbose: bool = False, out: str = None) -> str:
    """``out``: write a tuning variant (RT_NVCC_EXTRA=-D...) next to the product library instead of replacing it"""
    if out is not None:
        force = True
    elif not force and not needs_build():
        return LIB
    nvcc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "bin", "nvcc")
    extra = os.environ.get("RT_NVCC_EXTRA", "").split()   # tuning experiments (-DPP_BATCH=2 ...), never needed for a build
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", out or LIB] + SOURCES
    r = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if verbose:
        print(" ".join(cmd))
        print(r.stderr)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building librt_b200.so")
    return out or LIB


if __name__ == "__main__":
    _out = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else None
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, out=_out))
