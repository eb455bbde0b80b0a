"""Fused rollout path (include/bb_rollout.h): SURVEY.md section 8(f) rows n1 + n2.

Host-side mirror of what scripts/ppo.py does around the step path — `rollout()` (ppo.py:60-141),
`compute_advantages` (ppo.py:144-166), `RolloutBuffer` (scripts/buffers.py), `PPOStats.update`
(scripts/ppo_stats.py:21-36) — with the per-step work moved into libbball.so: one kernel for the
whole `Agent.forward` + sampling + buffer writes, the fused step kernel, and one GAE kernel.
torch is used for storage (parameters, buffers) and the stream only.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional

from . import _lib

OBS_DIM, NUM_LOGITS, NUM_ACTION_DIMS = 128, 19, 6
ACTION_BUCKETS = [2, 8, 3, 2, 2, 2]  # scripts/env.py:102

_fp = ctypes.c_void_p


class BBPolicy(ctypes.Structure):
    _fields_ = [("hidden", ctypes.c_int32), ("obs_mean", _fp), ("obs_inv_std", _fp), ("obs_clamp", ctypes.c_float),
                ("w1", _fp), ("b1", _fp), ("ln1_g", _fp), ("ln1_b", _fp), ("w2", _fp), ("b2", _fp), ("ln2_g", _fp),
                ("ln2_b", _fp), ("wa", _fp), ("ba", _fp), ("wc", _fp), ("bc", _fp)]


class BBRolloutBuffers(ctypes.Structure):
    _fields_ = [("horizon", ctypes.c_int32), ("obs", _fp), ("actions", _fp), ("values", _fp), ("log_probs", _fp),
                ("rewards", _fp), ("not_dones", _fp), ("next_value", _fp), ("cur_reward", _fp), ("ep_length", _fp),
                ("ep_sums", _fp)]


def _bind(lib):
    if getattr(lib, "_rollout_bound", False):
        return lib
    vp, i32, i64, u32 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_uint32
    lib.bb_policy_forward.restype = ctypes.c_int
    lib.bb_policy_forward.argtypes = [ctypes.POINTER(BBPolicy), vp, i64, i64, ctypes.c_int, u32, u32, i64, vp, i64, vp, vp, vp, vp]
    lib.bb_rollout.restype = ctypes.c_int
    lib.bb_rollout.argtypes = [vp, ctypes.POINTER(BBPolicy), ctypes.POINTER(BBPolicy), ctypes.POINTER(BBRolloutBuffers),
                               i32, u32, u32, i64, vp]
    lib.bb_gae.restype = ctypes.c_int
    lib.bb_gae.argtypes = [ctypes.POINTER(BBRolloutBuffers), i64, ctypes.c_double, ctypes.c_double, vp, vp, ctypes.c_float,
                           vp, vp, vp, vp]
    lib._rollout_bound = True
    return lib


# state_dict keys of scripts/agent.py Agent(num_layers=2): checkpoints of the reference load as they are
_KEYS = {"w1": "backbone.0.weight", "b1": "backbone.0.bias", "ln1_g": "backbone.1.weight", "ln1_b": "backbone.1.bias",
         "w2": "backbone.3.weight", "b2": "backbone.3.bias", "ln2_g": "backbone.4.weight", "ln2_b": "backbone.4.bias",
         "wa": "actor.weight", "ba": "actor.bias", "wc": "critic.weight", "bc": "critic.bias"}


class PolicyWeights:
    """Device pointers to an Agent's parameters (scripts/agent.py:108-131).

    Built from a state_dict-like mapping whose tensors live on the simulator's device.  If the mapping holds
    the live parameters of a torch module (``dict(module.state_dict(keep_vars=True))`` or ``from_module``),
    optimiser steps are seen without copies; call ``refresh()`` after the observation normaliser changed.
    """

    def __init__(self, tensors: Dict[str, "object"], obs_clamp: float = 5.0):
        import torch

        self._t = {}
        for field, key in _KEYS.items():
            t = tensors[key].detach()
            if t.dtype != torch.float32 or not t.is_contiguous():
                t = t.to(torch.float32).contiguous()
            assert t.is_cuda, f"{key} must be on the simulator's CUDA device"
            self._t[field] = t
        self.hidden = int(self._t["w1"].shape[0])
        assert self._t["w1"].shape == (self.hidden, OBS_DIM) and self._t["w2"].shape == (self.hidden, self.hidden)
        assert self._t["wa"].shape == (NUM_LOGITS, self.hidden) and self._t["wc"].shape == (1, self.hidden)
        assert 1 <= self.hidden <= 32
        self._src = tensors
        self.obs_clamp = float(obs_clamp)
        self.c = BBPolicy()
        self.refresh()

    @classmethod
    def from_module(cls, agent) -> "PolicyWeights":
        return cls(dict(agent.state_dict(keep_vars=True)), obs_clamp=float(getattr(agent.obs_norm, "clamp", 5.0)))

    def refresh(self) -> None:
        """Re-derive the float32 normaliser vectors (agent.py:28-30,36) from obs_norm.mean / obs_norm.var."""
        import torch

        mean = self._src["obs_norm.mean"].detach().to(torch.float32)
        var = self._src["obs_norm.var"].detach().to(torch.float32) + 1e-5
        self._t["obs_mean"] = mean.contiguous()
        self._t["obs_inv_std"] = torch.rsqrt(var).contiguous()
        c = self.c
        c.hidden = self.hidden
        c.obs_clamp = self.obs_clamp
        for field in list(_KEYS) + ["obs_mean", "obs_inv_std"]:
            setattr(c, field, self._t[field].data_ptr())


class RolloutBuffers:
    """scripts/buffers.py RolloutBuffer as plain torch tensors + the C view of them."""

    def __init__(self, n_steps: int, n_envs: int, device):
        import torch

        z = lambda *s, dt=torch.float32: torch.zeros(s, dtype=dt, device=device)
        self.horizon, self.n_envs = n_steps, n_envs
        self.obs = z(n_steps, n_envs, OBS_DIM)
        self.actions = z(n_steps, n_envs, NUM_ACTION_DIMS)
        self.values, self.log_probs = z(n_steps, n_envs), z(n_steps, n_envs)
        self.rewards, self.not_dones = z(n_steps, n_envs), z(n_steps, n_envs)
        self.next_value = z(n_envs)
        self.advantages, self.returns, self.values_unnorm = z(n_steps, n_envs), z(n_steps, n_envs), z(n_steps, n_envs)
        self.cur_reward, self.ep_length = z(n_envs), z(n_envs)
        self.ep_sums = z(3, dt=torch.float64)   # finished episodes: sum of returns, sum of lengths, count
        c = self.c = BBRolloutBuffers()
        c.horizon = n_steps
        for f in ("obs", "actions", "values", "log_probs", "rewards", "not_dones", "next_value", "cur_reward", "ep_length", "ep_sums"):
            setattr(c, f, getattr(self, f).data_ptr())

    def episode_stats(self) -> Dict[str, float]:
        """Mean return / length over the episodes finished so far (one D2H read; call it when logging)."""
        s = self.ep_sums.tolist()
        cnt = max(s[2], 1.0)
        return {"mean_reward": s[0] / cnt, "mean_episode_length": s[1] / cnt, "episodes": s[2]}


def policy_forward(weights: PolicyWeights, obs, stochastic: bool = True, seed: int = 0, draw: int = 0, row_offset: int = 0,
                   actions_out=None, want_logits: bool = False):
    """Agent.forward (agent.py:134-154) over the rows of `obs` ([n,128] f32 view, any row stride).
    Returns (actions i32 [n,6], log_prob f32 [n], value f32 [n][, logits f32 [n,19]])."""
    import torch

    lib = _bind(_lib.load())
    n = obs.shape[0]
    assert obs.is_cuda and obs.dtype == torch.float32 and obs.shape[1] == OBS_DIM and obs.stride(1) == 1
    dev = obs.device
    actions = actions_out if actions_out is not None else torch.empty((n, NUM_ACTION_DIMS), dtype=torch.int32, device=dev)
    assert actions.dtype == torch.int32 and actions.stride(1) == 1
    logp = torch.empty(n, dtype=torch.float32, device=dev)
    value = torch.empty(n, dtype=torch.float32, device=dev)
    logits = torch.empty((n, NUM_LOGITS), dtype=torch.float32, device=dev) if want_logits else None
    rc = lib.bb_policy_forward(ctypes.byref(weights.c), obs.data_ptr(), obs.stride(0), n, int(stochastic), seed, draw, row_offset,
                               actions.data_ptr(), actions.stride(0), logp.data_ptr(), value.data_ptr(),
                               logits.data_ptr() if want_logits else None, torch.cuda.current_stream(dev).cuda_stream)
    if rc != 0:
        raise RuntimeError("bb_policy_forward failed: " + _lib.last_error())
    return (actions, logp, value, logits) if want_logits else (actions, logp, value)


class FusedRollout:
    """ppo.py `rollout()` + `compute_advantages`' GAE on a SimpleGridworldSimulator, without per-step host work."""

    def __init__(self, sim, trainee: PolicyWeights, agent_idx: int, frozen: Optional[PolicyWeights] = None,
                 seed: int = 321, world_offset: int = 0):
        self.sim, self.trainee, self.frozen = sim, trainee, frozen
        self.agent_idx, self.seed, self.world_offset = int(agent_idx), int(seed), int(world_offset)
        self.draws = 0
        self._lib = _bind(_lib.load())

    def collect(self, buf: RolloutBuffers) -> None:
        """One rollout of buf.horizon steps (incl. the env.reset() step); asynchronous on torch's current stream."""
        assert buf.n_envs == self.sim.num_worlds
        rc = self._lib.bb_rollout(self.sim._h, ctypes.byref(self.trainee.c),
                                  ctypes.byref(self.frozen.c) if self.frozen is not None else None, ctypes.byref(buf.c),
                                  self.agent_idx, self.seed, self.draws & 0x7FFFFFFF, self.world_offset, self.sim._stream())
        if rc != 0:
            raise RuntimeError("bb_rollout failed: " + _lib.last_error())
        self.draws += buf.horizon

    def gae(self, buf: RolloutBuffers, gamma: float, gae_lambda: float, value_mean, value_var, value_clamp: float = 5.0) -> None:
        """ppo.py:144-166: fills buf.values_unnorm, buf.advantages, buf.returns (un-normalised)."""
        import torch

        assert value_mean.dtype == torch.float64 and value_var.dtype == torch.float64 and value_mean.is_cuda
        rc = self._lib.bb_gae(ctypes.byref(buf.c), buf.n_envs, float(gamma), float(gae_lambda), value_mean.data_ptr(),
                              value_var.data_ptr(), float(value_clamp), buf.values_unnorm.data_ptr(), buf.advantages.data_ptr(),
                              buf.returns.data_ptr(), self.sim._stream())
        if rc != 0:
            raise RuntimeError("bb_gae failed: " + _lib.last_error())
