"""Builds libpsana_ray_cuda.so in-tree for sm_100a (nvcc cross-compiles without a GPU).

    python psana-ray_b200/build.py [--force] [--verbose]

The library lands in psana-ray_b200/lib/ (git-ignored, but shipped to the GPU box by gpurun).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libpsana_ray_cuda.so")
SOURCES = ["capi.cu", "calib_map.cu", "common_mode.cu"]
HEADERS = [os.path.join(CSRC, "psr_kernels.cuh"), os.path.join(CSRC, "sortnet.cuh"), os.path.join(HERE, "..", "include", "psana_ray_cuda.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--fmad=false",                 # belt and braces: every float op is an explicit _rn intrinsic anyway
    "-Xcompiler", "-fPIC", "-shared",
    "-Xlinker", "-Bsymbolic",       # internal calls bind inside the library, whatever else the process has loaded
    "-Xptxas", "-v",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def up_to_date() -> bool:
    if not os.path.exists(LIB):
        return False
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + HEADERS + [os.path.abspath(__file__)]
    return all(os.path.getmtime(d) <= t for d in deps)


def build_variant(name: str, defines: dict) -> str:
    """A tuning variant of the library (-D knobs) under lib/variants/, for sweeps on the GPU box."""
    vdir = os.path.join(LIBDIR, "variants")
    os.makedirs(vdir, exist_ok=True)
    out = os.path.join(vdir, f"libpsana_ray_cuda.{name}.so")
    flags = [f for f in NVCC_FLAGS if f not in ("-Xptxas", "-v")]
    cmd = [find_nvcc()] + flags + [f"-D{k}={v}" for k, v in defines.items()] + [os.path.join(CSRC, s) for s in SOURCES] + ["-o", out]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + (res.stdout + res.stderr)[-4000:])
    return out


def build_library(force: bool = False, verbose: bool = False) -> str:
    if not force and up_to_date():
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    cmd = [find_nvcc()] + NVCC_FLAGS + [os.path.join(CSRC, s) for s in SOURCES] + ["-o", LIB]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = res.stdout + res.stderr
    with open(os.path.join(LIBDIR, "build.log"), "w") as fh:
        fh.write(" ".join(cmd) + "\n" + log)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + log[-8000:])
    if verbose:
        print(log)
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
