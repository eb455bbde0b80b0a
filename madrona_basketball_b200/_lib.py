"""ctypes binding of libbball.so (include/bb_engine.h).  No fallback: if the CUDA
library is missing or cannot be loaded, importing the engine fails loudly."""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BB_LIB_PATH") or os.path.join(_HERE, "libbball.so")  # BB_LIB_PATH: tuning builds

# every symbol include/bb_engine.h declares
SYMBOLS = [
    "bb_create", "bb_destroy", "bb_step", "bb_step_n", "bb_step_with_actions", "bb_step_host", "bb_export",
    "bb_set_action", "bb_trigger_reset", "bb_fill_actions", "bb_checksum", "bb_launch_count",
    "bb_num_worlds", "bb_last_error", "bb_math_eval", "bb_serve_begin", "bb_serve_step", "bb_serve_end",
    "bb_serve_last_gpu_ns", "bb_step_host_u8", "bb_serve_step_u8", "bb_serve_last_marks", "bb_step_n_with_actions",
]


class BBConfig(ctypes.Structure):
    _fields_ = [
        ("num_worlds", ctypes.c_int64),
        ("discrete_x", ctypes.c_int64),
        ("discrete_y", ctypes.c_int64),
        ("start_x", ctypes.c_float),
        ("start_y", ctypes.c_float),
        ("max_episode_length", ctypes.c_int64),
        ("rand_seed", ctypes.c_uint32),
        ("gpu_id", ctypes.c_int32),
        ("flags", ctypes.c_uint32),
    ]


_lib = None


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the B200 step engine has no CPU or PyTorch fallback. "
            "Build it with `python -m madrona_basketball_b200.build` (needs nvcc).")
    lib = ctypes.CDLL(LIB_PATH)
    vp, i32, i64, u32 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_uint32
    lib.bb_create.restype = vp
    lib.bb_create.argtypes = [ctypes.POINTER(BBConfig)]
    lib.bb_destroy.restype = None
    lib.bb_destroy.argtypes = [vp]
    lib.bb_step.restype = ctypes.c_int
    lib.bb_step.argtypes = [vp, vp]
    lib.bb_step_n.restype = ctypes.c_int
    lib.bb_step_n.argtypes = [vp, i32, vp]
    lib.bb_step_n_with_actions.restype = ctypes.c_int
    lib.bb_step_n_with_actions.argtypes = [vp, i32, vp, vp]
    lib.bb_step_with_actions.restype = ctypes.c_int
    lib.bb_step_with_actions.argtypes = [vp, vp, vp]
    lib.bb_step_host.restype = ctypes.c_int
    lib.bb_step_host.argtypes = [vp, vp, vp, vp, vp]
    lib.bb_serve_begin.restype = ctypes.c_int
    lib.bb_serve_begin.argtypes = [vp]
    lib.bb_serve_step.restype = ctypes.c_int
    lib.bb_serve_step.argtypes = [vp, vp, vp, vp, vp]
    lib.bb_serve_end.restype = ctypes.c_int
    lib.bb_serve_end.argtypes = [vp]
    lib.bb_serve_last_marks.restype = ctypes.c_int
    lib.bb_serve_last_marks.argtypes = [vp, ctypes.POINTER(ctypes.c_longlong)]
    for name in ("bb_step_host_u8", "bb_serve_step_u8"):
        getattr(lib, name).restype = ctypes.c_int
        getattr(lib, name).argtypes = [vp, vp, vp, vp, vp]
    lib.bb_serve_last_gpu_ns.restype = ctypes.c_longlong
    lib.bb_serve_last_gpu_ns.argtypes = [vp]
    lib.bb_export.restype = ctypes.c_int
    lib.bb_export.argtypes = [vp, ctypes.c_int, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_int),
                              ctypes.POINTER(i64), ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)]
    lib.bb_set_action.restype = ctypes.c_int
    lib.bb_set_action.argtypes = [vp] + [i32] * 8
    lib.bb_trigger_reset.restype = ctypes.c_int
    lib.bb_trigger_reset.argtypes = [vp, i32]
    lib.bb_fill_actions.restype = ctypes.c_int
    lib.bb_fill_actions.argtypes = [vp, vp, u32, i64, i64, vp]
    lib.bb_checksum.restype = ctypes.c_int
    lib.bb_checksum.argtypes = [vp, vp, vp]
    lib.bb_math_eval.restype = ctypes.c_int
    lib.bb_math_eval.argtypes = [ctypes.c_int, vp, vp, vp, i64, vp]
    lib.bb_launch_count.restype = i64
    lib.bb_launch_count.argtypes = [vp]
    lib.bb_num_worlds.restype = i64
    lib.bb_num_worlds.argtypes = [vp]
    lib.bb_last_error.restype = ctypes.c_char_p
    lib.bb_last_error.argtypes = []
    _lib = lib
    return lib


def last_error() -> str:
    return load().bb_last_error().decode()
