"""Python mirror of the reference's nanobind class (src/bindings.cpp:14-101).

Same class name, constructor keywords, 22 method names, tensor shapes and dtypes,
so scripts/env.py, ppo.py and self_play.py of the reference run unmodified.  All
compute happens in libbball.so (CUDA, sm_100a); torch is used only to wrap the
engine's device memory as zero-copy tensors and to name the current stream.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, Optional, Tuple

import numpy as np

from . import _lib
from .madrona import ExecMode

FLAG_TAG_MASK = 1  # include/bb_engine.h: BB_FLAG_TAG_MASK

_FASTCALL = False


def _fastcall_module():
    """The bound-call glue (csrc/bb_fastcall.c), or None when it has not been built (the ctypes closures are used then).
    BB_NO_FASTCALL=1 disables it (A/B runs)."""
    global _FASTCALL
    if _FASTCALL is False:
        _FASTCALL = None
        if not os.environ.get("BB_NO_FASTCALL"):
            try:
                from . import _fastcall
                _FASTCALL = _fastcall
            except ImportError:
                pass
    return _FASTCALL

# include/bb_engine.h enum bb_slot
SLOTS: Dict[str, int] = {
    "reset": 0, "game_state": 1, "action": 2, "action_mask": 3, "agent_pos": 4,
    "observations": 5, "reward": 6, "done": 7, "agent_entity_id": 8,
    "agent_possession": 9, "orientation": 10, "agent_team": 11, "agent_stats": 12,
    "basketball_pos": 13, "ball_physics": 14, "ball_entity_id": 15, "ball_grabbed": 16,
    "ball_velocity": 17, "hoop_pos": 18,
    "agent_velocity": 100, "inbounding": 101, "grab_cooldown": 102, "attributes": 103,
    "cur_step": 104, "world_clock": 105, "rng_count": 106,
}


class Tensor:
    """madrona::py::Tensor stand-in: a non-owning view of engine memory.

    `.to_torch()` returns a zero-copy torch.Tensor on the simulator's device
    (reference: Tensor(dev_ptr, type, dims, gpu_id), mgr.cpp:121-126).  The
    simulator object must outlive the tensors, as in the reference.
    """

    def __init__(self, owner: "SimpleGridworldSimulator", ptr: int, dtype: int, shape: Tuple[int, ...], device: int):
        self._owner = owner  # keeps the engine alive while views exist
        self._ptr = ptr
        self._dtype = dtype
        self._shape = shape
        self._device = device
        self._torch = None

    @property
    def shape(self) -> Tuple[int, ...]:
        return self._shape

    @property
    def __cuda_array_interface__(self):
        return {
            "shape": self._shape,
            "typestr": "<i4" if self._dtype == 0 else "<f4",
            "data": (self._ptr, False),
            "version": 3,
            "strides": None,
            "stream": None,
        }

    def to_torch(self):
        import torch

        if self._torch is None:
            self._torch = torch.as_tensor(self, device=torch.device("cuda", self._device))
        return self._torch

    def device_ptr(self) -> int:
        return self._ptr


class SimpleGridworldSimulator:
    """Drop-in for `madrona_basketball.SimpleGridworldSimulator` (bindings.cpp:17-53)."""

    def __init__(self, discrete_x: int, discrete_y: int, start_x: float, start_y: float,
                 max_episode_length: int, exec_mode, num_worlds: int, gpu_id: int = -1,
                 *, _flags: int = FLAG_TAG_MASK):
        mode = exec_mode.value if isinstance(exec_mode, ExecMode) else int(exec_mode)
        if mode != ExecMode.CUDA.value:
            raise RuntimeError(
                "madrona_basketball_b200 is a CUDA-only engine (sm_100a): ExecMode.CPU is not provided. "
                "Construct with exec_mode=madrona.ExecMode.CUDA (scripts: --use-gpu).")
        self._lib = _lib.load()
        cfg = _lib.BBConfig(int(num_worlds), int(discrete_x), int(discrete_y), float(start_x), float(start_y),
                            int(max_episode_length), 0,  # bindings.cpp:37: randSeed = 0
                            int(gpu_id), int(_flags))
        self._h = self._lib.bb_create(ctypes.byref(cfg))
        if not self._h:
            raise RuntimeError("bb_create failed: " + _lib.last_error())
        self.num_worlds = int(num_worlds)
        self.gpu_id = max(int(gpu_id), 0)
        self._tensors: Dict[int, Tensor] = {}
        self._host_ptrs: Dict[int, tuple] = {}

    # ------------------------------------------------------------------ lifetime
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.bb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int, what: str) -> None:
        if rc != 0:
            raise RuntimeError(f"{what} failed: {_lib.last_error()}")

    def _stream(self) -> int:
        import torch

        return torch.cuda.current_stream(self.gpu_id).cuda_stream

    # ------------------------------------------------------------------ stepping
    def step(self) -> None:
        """Manager::step (mgr.cpp:243-246), enqueued on torch's current stream."""
        self._check(self._lib.bb_step(self._h, self._stream()), "bb_step")

    def step_n(self, n: int) -> None:
        """Extension: n consecutive step() calls as ONE kernel launch (bb_step_n: every warp takes its worlds through the n
        steps in a row; BB_STEP_N_GRAPH=1: a CUDA graph of n step kernels)."""
        self._check(self._lib.bb_step_n(self._h, int(n), self._stream()), "bb_step_n")

    def step_n_with_actions(self, actions) -> None:
        """Extension: `for a in actions: action_tensor.copy_(a); step()` as one kernel launch (bb_step_n_with_actions);
        actions: int32 [n, N, 2, 6] on the device."""
        import torch

        assert actions.is_cuda and actions.is_contiguous() and actions.dtype == torch.int32
        assert actions.dim() == 4 and actions[0].numel() == self.num_worlds * 12
        self._check(self._lib.bb_step_n_with_actions(self._h, int(actions.shape[0]), actions.data_ptr(), self._stream()),
                    "bb_step_n_with_actions")

    def step_with_actions(self, actions) -> None:
        """Extension: `action_tensor.copy_(actions); step()` fused into the kernel's loads."""
        assert actions.is_cuda and actions.is_contiguous() and actions.numel() == self.num_worlds * 12
        self._check(self._lib.bb_step_with_actions(self._h, actions.data_ptr(), self._stream()), "bb_step_with_actions")

    def bind_step_with_actions(self, actions):
        """step_with_actions(actions) with the argument checks, pointer and stream look-ups done ONCE: returns a
        zero-argument callable that enqueues the step on the stream that is current now.  For loops that launch faster
        than a Python-level call with tensor indexing can (a step kernel lasts a few microseconds)."""
        assert actions.is_cuda and actions.is_contiguous() and actions.numel() == self.num_worlds * 12
        fn, h, ptr, st = self._lib.bb_step_with_actions, self._h, actions.data_ptr(), self._stream()
        check = self._check

        def call():
            if fn(h, ptr, st) != 0:
                check(-1, "bb_step_with_actions")
        return call

    def step_host(self, actions: Optional[np.ndarray], reward: Optional[np.ndarray] = None,
                  done: Optional[np.ndarray] = None, obs: Optional[np.ndarray] = None) -> None:
        """Extension: host-buffer step (H2D actions, step, D2H results), synchronous."""
        n, ptr = self.num_worlds, self._host_ptr
        if actions is not None and actions.dtype == np.uint8:   # packed rows (bb_step_host_u8)
            self._check(self._lib.bb_step_host_u8(self._h, ptr(actions, np.uint8, n * 12), ptr(reward, np.float32, n * 2),
                                                  ptr(done, np.float32, n * 2), ptr(obs, np.float32, n * 256)), "bb_step_host_u8")
            return
        self._check(self._lib.bb_step_host(self._h, ptr(actions, np.int32, n * 12), ptr(reward, np.float32, n * 2),
                                           ptr(done, np.float32, n * 2), ptr(obs, np.float32, n * 256)), "bb_step_host")

    def serve_begin(self) -> None:
        """Extension: start the persistent server kernel (bb_serve_begin)."""
        self._check(self._lib.bb_serve_begin(self._h), "bb_serve_begin")

    def _host_ptr(self, a: Optional[np.ndarray], dtype, n: int):
        """Address of a host buffer, validated once per array object (the per-step path is latency-bound)."""
        if a is None:
            return None
        hit = self._host_ptrs.get(id(a))
        if hit is not None and hit[0] is a:
            return hit[1]
        assert a.dtype == dtype and a.flags["C_CONTIGUOUS"] and a.size == n, (a.dtype, a.shape)
        if len(self._host_ptrs) >= 256:
            self._host_ptrs.clear()
        self._host_ptrs[id(a)] = (a, a.ctypes.data)   # holds a reference: the id cannot be recycled
        return self._host_ptrs[id(a)][1]

    def serve_step(self, actions: np.ndarray, reward: Optional[np.ndarray] = None,
                   done: Optional[np.ndarray] = None, obs: Optional[np.ndarray] = None) -> None:
        """Extension: one host-buffer step through the persistent server (pinned buffers only)."""
        n = self.num_worlds
        if actions.dtype == np.uint8:   # packed rows (bb_serve_step_u8)
            rc = self._lib.bb_serve_step_u8(self._h, self._host_ptr(actions, np.uint8, n * 12), self._host_ptr(reward, np.float32, n * 2),
                                            self._host_ptr(done, np.float32, n * 2), self._host_ptr(obs, np.float32, n * 256))
        else:
            rc = self._lib.bb_serve_step(self._h, self._host_ptr(actions, np.int32, n * 12), self._host_ptr(reward, np.float32, n * 2),
                                         self._host_ptr(done, np.float32, n * 2), self._host_ptr(obs, np.float32, n * 256))
        if rc != 0:
            self._check(rc, "bb_serve_step")

    def bind_serve_step(self, actions: np.ndarray, reward: Optional[np.ndarray] = None, done: Optional[np.ndarray] = None,
                        obs: Optional[np.ndarray] = None):
        """serve_step with its buffer checks and pointer look-ups done once: a zero-argument callable for loops where the
        ~2 us of per-call Python argument handling matter (the served step itself takes ~18 us at 8192 worlds)."""
        n = self.num_worlds
        u8 = actions.dtype == np.uint8
        fn = self._lib.bb_serve_step_u8 if u8 else self._lib.bb_serve_step
        args = (self._h, self._host_ptr(actions, np.uint8 if u8 else np.int32, n * 12), self._host_ptr(reward, np.float32, n * 2),
                self._host_ptr(done, np.float32, n * 2), self._host_ptr(obs, np.float32, n * 256))
        check = self._check
        fast = _fastcall_module()
        if fast is not None:   # csrc/bb_fastcall.c: the same C-ABI call with the same pointers, without ctypes' ~1.5 us per call
            addr = lambda f: ctypes.cast(f, ctypes.c_void_p).value
            return fast.bind5(addr(fn), addr(self._lib.bb_last_error), *(int(a or 0) for a in args))

        def call():
            if fn(*args) != 0:
                check(-1, "bb_serve_step")
        return call

    def serve_end(self) -> None:
        self._check(self._lib.bb_serve_end(self._h), "bb_serve_end")

    def set_action(self, world_idx: int, agent_idx: int, move_speed: int, move_angle: int, rotate: int,
                   grab: int, pass_: int = 0, shoot: int = 0, **kw) -> None:
        """Manager::setAction (mgr.cpp:270-293); keyword `pass` accepted via **kw like the nanobind signature."""
        if "pass" in kw:
            pass_ = kw.pop("pass")
        self._check(self._lib.bb_set_action(self._h, world_idx, agent_idx, move_speed, move_angle, rotate, grab,
                                            pass_, shoot), "bb_set_action")

    def trigger_reset(self, world_idx: int) -> None:
        """Manager::triggerReset (mgr.cpp:297-311)."""
        self._check(self._lib.bb_trigger_reset(self._h, world_idx), "bb_trigger_reset")

    # ------------------------------------------------------------------ extensions used by bench / tests
    def fill_actions(self, seed: int, step: int, world_offset: int = 0, out=None) -> None:
        ptr = out.data_ptr() if out is not None else None
        self._check(self._lib.bb_fill_actions(self._h, ptr, seed, step, world_offset, self._stream()), "bb_fill_actions")

    def checksum(self):
        import torch

        out = torch.empty(self.num_worlds, dtype=torch.int64, device=torch.device("cuda", self.gpu_id))
        self._check(self._lib.bb_checksum(self._h, out.data_ptr(), self._stream()), "bb_checksum")
        return out

    def launch_count(self) -> int:
        return int(self._lib.bb_launch_count(self._h))

    # ------------------------------------------------------------------ tensors
    def _tensor(self, slot: int) -> Tensor:
        if slot in self._tensors:
            return self._tensors[slot]
        ptr = ctypes.c_void_p()
        dtype = ctypes.c_int()
        ndim = ctypes.c_int()
        dev = ctypes.c_int()
        dims = (ctypes.c_int64 * 4)()
        self._check(self._lib.bb_export(self._h, slot, ctypes.byref(ptr), ctypes.byref(dtype), dims,
                                        ctypes.byref(ndim), ctypes.byref(dev)), "bb_export")
        t = Tensor(self, ptr.value, dtype.value, tuple(int(dims[i]) for i in range(ndim.value)), dev.value)
        self._tensors[slot] = t
        return t

    def tensor(self, name: str) -> Tensor:
        """Any column by name, including the internal ones (slots >= 100)."""
        return self._tensor(SLOTS[name])

    # the 19 accessors of bindings.cpp:71-99
    def reset_tensor(self): return self._tensor(0)
    def game_state_tensor(self): return self._tensor(1)
    def action_tensor(self): return self._tensor(2)
    def action_mask_tensor(self): return self._tensor(3)
    def agent_pos_tensor(self): return self._tensor(4)
    def observations_tensor(self): return self._tensor(5)
    def reward_tensor(self): return self._tensor(6)
    def done_tensor(self): return self._tensor(7)
    def agent_entity_id_tensor(self): return self._tensor(8)
    def agent_possession_tensor(self): return self._tensor(9)
    def orientation_tensor(self): return self._tensor(10)
    def agent_team_tensor(self): return self._tensor(11)
    def agent_stats_tensor(self): return self._tensor(12)
    def basketball_pos_tensor(self): return self._tensor(13)
    def ball_physics_tensor(self): return self._tensor(14)
    def ball_entity_id_tensor(self): return self._tensor(15)
    def ball_grabbed_tensor(self): return self._tensor(16)
    def ball_velocity_tensor(self): return self._tensor(17)
    def hoop_pos_tensor(self): return self._tensor(18)
