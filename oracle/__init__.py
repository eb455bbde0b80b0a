"""ctypes front-end of the CPU oracle (oracle/bb_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(madrona_basketball_b200) never imports this module.  PARITY UNPINNED: see
bb_oracle.h.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Dict

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

SLOTS: Dict[str, int] = {
    # reference ExportID order, src/types.hpp:10-42
    "reset": 0, "game_state": 1, "action": 2, "action_mask": 3, "agent_pos": 4,
    "observations": 5, "reward": 6, "done": 7, "agent_entity_id": 8,
    "agent_possession": 9, "orientation": 10, "agent_team": 11, "agent_stats": 12,
    "basketball_pos": 13, "ball_physics": 14, "ball_entity_id": 15, "ball_grabbed": 16,
    "ball_velocity": 17, "hoop_pos": 18,
    # internal components
    "agent_velocity": 100, "inbounding": 101, "grab_cooldown": 102, "attributes": 103,
    "cur_step": 104, "world_clock": 105, "rng_count": 106,
}

FLAG_TAG_MASK = 1


class _Config(ctypes.Structure):
    _fields_ = [
        ("num_worlds", ctypes.c_int64),
        ("discrete_x", ctypes.c_int64),
        ("discrete_y", ctypes.c_int64),
        ("start_x", ctypes.c_float),
        ("start_y", ctypes.c_float),
        ("rand_seed", ctypes.c_uint32),
        ("flags", ctypes.c_uint32),
    ]


def build(force: bool = False) -> None:
    """Compile liboracle.so / liboracle_libm.so with oracle/Makefile."""
    if force:
        subprocess.run(["make", "-C", _HERE, "clean"], check=True, capture_output=True)
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)


_LIBS: Dict[str, ctypes.CDLL] = {}


def build_native() -> str:
    """The glibc-libm build once more, tuned for THIS host (-O3 -march=native, still -ffp-contract=off): the binary the
    CPU arm of bench.py times, so that the baseline is not handicapped by a generic -O2 build.  Compiled on the machine
    that runs it (the file name carries a hash of the CPU's feature flags: a library built in the container is never
    loaded on a GPU box with a different CPU)."""
    import hashlib

    flags = ""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith(("flags", "model name")):
                    flags += line
                    if line.startswith("flags"):
                        break
    except OSError:
        pass
    tag = hashlib.sha256(flags.encode()).hexdigest()[:12]
    out_dir = os.path.join(_HERE, "_native")
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, f"liboracle_libm_native_{tag}.so")
    src = os.path.join(_HERE, "bb_oracle.cpp")
    if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
        cmd = ["g++", "-std=c++17", "-O3", "-march=native", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-DBB_USE_LIBM",
               src, "-o", path, "-lm"]
        subprocess.run(cmd, check=True, capture_output=True)
    return path


def _load(libm) -> ctypes.CDLL:
    name = "native" if libm == "native" else ("liboracle_libm.so" if libm else "liboracle.so")
    if name in _LIBS:
        return _LIBS[name]
    path = build_native() if libm == "native" else os.path.join(_HERE, name)
    if not os.path.exists(path):
        build()
    lib = ctypes.CDLL(path)
    lib.orc_create.restype = ctypes.c_void_p
    lib.orc_create.argtypes = [ctypes.POINTER(_Config)]
    lib.orc_destroy.argtypes = [ctypes.c_void_p]
    lib.orc_step.argtypes = [ctypes.c_void_p, ctypes.c_int]
    lib.orc_tensor.restype = ctypes.c_void_p
    lib.orc_tensor.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int),
                               ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int)]
    lib.orc_fill_actions.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int64, ctypes.c_int64]
    lib.orc_run_synthetic.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int64, ctypes.c_int64,
                                      ctypes.c_int64, ctypes.c_int]
    lib.orc_checksum.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.orc_threefry2x32.argtypes = [ctypes.c_uint32] * 4 + [ctypes.POINTER(ctypes.c_uint32)]
    lib.orc_rng_uniform.restype = ctypes.c_float
    lib.orc_rng_uniform.argtypes = [ctypes.c_uint32, ctypes.c_uint32]
    lib.orc_math_eval.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64]
    _LIBS[name] = lib
    return lib


def rotate_vec(q, v) -> np.ndarray:
    """Quat{w,x,y,z}.rotateVec(v) of the oracle's Madrona primitive restatement (float32)."""
    qa = np.ascontiguousarray(q, dtype=np.float32)
    va = np.ascontiguousarray(v, dtype=np.float32)
    out = np.empty(3, np.float32)
    lib = _load(False)
    lib.orc_rotate_vec.argtypes = [ctypes.c_void_p] * 3
    lib.orc_rotate_vec(qa.ctypes.data, va.ctypes.data, out.ctypes.data)
    return out


def court_constants() -> Dict[str, float]:
    out = np.empty(12, np.float32)
    lib = _load(False)
    lib.orc_court_constants.argtypes = [ctypes.c_void_p]
    lib.orc_court_constants(out.ctypes.data)
    names = ["COURT_LENGTH_M", "COURT_WIDTH_M", "WORLD_WIDTH_M", "WORLD_HEIGHT_M", "COURT_MIN_X", "COURT_MAX_X", "COURT_MIN_Y",
             "COURT_MAX_Y", "HOOP_FROM_BASELINE_M", "ARC_RADIUS_M", "CORNER_3_FROM_SIDELINE_M", "CORNER_3_LENGTH_FROM_BASELINE_M"]
    return {k: float(v) for k, v in zip(names, out)}


MATH_FNS = {"sinf": 0, "cosf": 1, "atanf": 2, "atan2f": 3, "acosf": 4, "expf": 5, "erff": 6}


def math_eval(fn: str, x: np.ndarray, y: np.ndarray | None = None, libm: bool = False) -> np.ndarray:
    """Elementwise evaluation of one libm call site in the spec or the glibc build."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.ascontiguousarray(y if y is not None else x, dtype=np.float32)
    out = np.empty_like(x)
    _load(libm).orc_math_eval(MATH_FNS[fn], x.ctypes.data, y.ctypes.data, out.ctypes.data, x.size)
    return out


def threefry2x32(k0: int, k1: int, c0: int, c1: int):
    out = (ctypes.c_uint32 * 2)()
    _load(False).orc_threefry2x32(k0, k1, c0, c1, out)
    return int(out[0]), int(out[1])


def rng_uniform(seed: int, draw_index: int) -> float:
    return float(_load(False).orc_rng_uniform(seed, draw_index))


class Oracle:
    """CPU restatement of one Manager (src/mgr.cpp) instance.

    Constructor arguments mirror bindings.cpp:18-26 as scripts/env.py:22-35 passes them.
    """

    def __init__(self, num_worlds: int, discrete_x: int = 32, discrete_y: int = 17,
                 start_x: float = 31.515 / 2.0, start_y: float = 16.764 / 2.0,
                 rand_seed: int = 0, flags: int = FLAG_TAG_MASK, libm=False):
        self._lib = _load(libm)
        cfg = _Config(num_worlds, discrete_x, discrete_y, start_x, start_y, rand_seed, flags)
        self._h = self._lib.orc_create(ctypes.byref(cfg))
        self.num_worlds = num_worlds
        self._views: Dict[str, np.ndarray] = {}

    def close(self) -> None:
        if self._h:
            self._lib.orc_destroy(self._h)
            self._h = None
            self._views.clear()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def tensor(self, name: str) -> np.ndarray:
        """Writable numpy view of a component column (same shape/dtype as the reference export)."""
        if name in self._views:
            return self._views[name]
        dtype = ctypes.c_int()
        ndim = ctypes.c_int()
        dims = (ctypes.c_int64 * 4)()
        ptr = self._lib.orc_tensor(self._h, SLOTS[name], ctypes.byref(dtype), dims, ctypes.byref(ndim))
        if not ptr:
            raise KeyError(name)
        shape = tuple(int(dims[i]) for i in range(ndim.value))
        np_dtype = np.int32 if dtype.value == 0 else np.float32
        count = int(np.prod(shape))
        buf = (ctypes.c_byte * (count * 4)).from_address(ptr)
        arr = np.frombuffer(buf, dtype=np_dtype, count=count).reshape(shape)
        self._views[name] = arr
        return arr

    def step(self, num_threads: int = 1) -> None:
        self._lib.orc_step(self._h, num_threads)

    def fill_actions(self, seed: int, step: int, world_offset: int = 0) -> None:
        self._lib.orc_fill_actions(self._h, seed, step, world_offset)

    def run_synthetic(self, seed: int, first_step: int, nsteps: int, world_offset: int = 0,
                      num_threads: int = 1) -> None:
        self._lib.orc_run_synthetic(self._h, seed, first_step, nsteps, world_offset, num_threads)

    def checksum(self) -> np.ndarray:
        out = np.empty(self.num_worlds, dtype=np.uint64)
        self._lib.orc_checksum(self._h, out.ctypes.data_as(ctypes.c_void_p))
        return out

    def state(self) -> Dict[str, np.ndarray]:
        """Copy of every column (exported + internal)."""
        return {k: self.tensor(k).copy() for k in SLOTS}
