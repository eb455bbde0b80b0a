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
    os.path.join(_HERE, "csrc", "bb_umma.cuh"),
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


HASH_PATH = LIB_PATH + ".hash"


def source_hash(extra=()) -> str:
    """sha256 over every source / header the library is built from + the compiler flags."""
    import hashlib

    h = hashlib.sha256()
    for d in DEPS:
        with open(d, "rb") as f:
            h.update(os.path.basename(d).encode() + b"\0" + f.read())
    h.update(" ".join(NVCC_FLAGS + list(extra)).encode())
    return h.hexdigest()


def needs_build() -> bool:
    """True unless libbball.so was built from exactly the sources in the tree (content hash, not mtime: the
    snapshot that carries a prebuilt library to the GPU box does not preserve a meaningful mtime order)."""
    if not os.path.exists(LIB_PATH) or not os.path.exists(HASH_PATH):
        return True
    with open(HASH_PATH) as f:
        return f.read().strip() != source_hash()


def build_engine(force: bool = False, verbose: bool = False, extra=(), out: str = LIB_PATH) -> str:
    if not force and out == LIB_PATH and not needs_build():
        return LIB_PATH
    cmd = [find_nvcc()] + NVCC_FLAGS + list(extra) + (["-Xptxas", "-v"] if verbose else []) + SOURCES + ["-o", out]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    if out == LIB_PATH:
        with open(HASH_PATH, "w") as f:
            f.write(source_hash(extra) + "\n")
    return out


# ---- binding glue: a CPython callable bound to one C-ABI entry point (csrc/bb_fastcall.c; gcc, no CUDA) ----
FASTCALL_SRC = os.path.join(_HERE, "csrc", "bb_fastcall.c")


def fastcall_path() -> str:
    import sysconfig

    return os.path.join(_HERE, "_fastcall" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_fastcall(force: bool = False) -> str:
    """gcc -O2 -shared -fPIC bb_fastcall.c -> madrona_basketball_b200/_fastcall.<abi>.so (rebuilt when older than its source)."""
    import sysconfig

    out = fastcall_path()
    if not force and os.path.exists(out) and os.path.getsize(out) > 0:
        with open(FASTCALL_SRC, "rb") as f, open(out + ".hash", "a+") as h:
            import hashlib

            want = hashlib.sha256(f.read()).hexdigest()
            h.seek(0)
            if h.read().strip() == want:
                return out
    gcc = shutil.which("gcc") or shutil.which("cc")
    if not gcc:
        raise RuntimeError("gcc not found; cannot build _fastcall")
    cmd = [gcc, "-O2", "-shared", "-fPIC", "-I", sysconfig.get_paths()["include"], FASTCALL_SRC, "-o", out]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("gcc failed:\n" + res.stdout + res.stderr)
    import hashlib

    with open(FASTCALL_SRC, "rb") as f, open(out + ".hash", "w") as h:
        h.write(hashlib.sha256(f.read()).hexdigest() + "\n")
    return out


if __name__ == "__main__":
    print(build_engine(force=True, verbose=True))
    print(build_fastcall(force=True))
