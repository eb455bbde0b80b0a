"""Builds libbball.so (CUDA kernels + C ABI) in-tree for sm_100a."""
from __future__ import annotations

import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbball.so")
SOURCES = [os.path.join(_HERE, "csrc", "bb_engine.cu")]
DEPS = SOURCES + [
    os.path.join(_HERE, "csrc", "bb_world.cuh"),
    os.path.join(_HERE, "csrc", "bb_pair.cuh"),
    os.path.join(_HERE, "csrc", "bb_rollout.cuh"),
    os.path.join(_HERE, "csrc", "bb_ppo.cuh"),
    os.path.join(_HERE, "..", "include", "bb_ppo.h"),
    os.path.join(_HERE, "..", "include", "bb_rollout.h"),
    os.path.join(_HERE, "..", "include", "bb_math.h"),
    os.path.join(_HERE, "..", "include", "bb_engine.h"),
]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",  # bit-exact float parity with the CPU oracle: no implicit FMA contraction
    "-Xcompiler", "-fPIC,-ffp-contract=off",
    "-shared",
]


def find_nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; cannot build libbball.so")
    return nvcc


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(d) > t for d in DEPS)


def build_engine(force: bool = False, verbose: bool = False, extra=(), out: str = LIB_PATH) -> str:
    if not force and out == LIB_PATH and not needs_build():
        return LIB_PATH
    cmd = [find_nvcc()] + NVCC_FLAGS + list(extra) + (["-Xptxas", "-v"] if verbose else []) + SOURCES + ["-o", out]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return out


if __name__ == "__main__":
    print(build_engine(force=True, verbose=True))
