"""Fused PPO update (include/bb_ppo.h): host-side mirror of scripts/ppo.py `update_policy` (ppo.py:180-216).

A minibatch step is `bb_ppo_grad` (forward + clipped-PPO loss + backward of the 5 972-parameter policy in one
kernel, deterministic block-order reduction) -> optional flat all-reduce across ranks -> `bb_ppo_adam`
(clip_grad_norm_ + Adam).  The parameters of the torch module are re-bound as views of ONE flat float32
vector, so the kernels, `state_dict()` (reference checkpoint keys) and `PolicyWeights` all see the same memory.
torch is used for storage, the permutation (`torch.randperm`, ppo.py:186) and the collective only.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional

from . import _lib
from .rollout import BBPolicy, PolicyWeights, _bind as _bind_rollout

NUM_PARAMS, NUM_STATS = 5972, 4
# flat layout = order of agent.parameters() (agent.py:111-127); offsets are include/bb_ppo.h BB_PPO_OFF_*
PARAM_ORDER = ["backbone.0.weight", "backbone.0.bias", "backbone.1.weight", "backbone.1.bias", "backbone.3.weight",
               "backbone.3.bias", "backbone.4.weight", "backbone.4.bias", "actor.weight", "actor.bias", "critic.weight",
               "critic.bias"]

_fp = ctypes.c_void_p


class BBRunningNorm(ctypes.Structure):
    _fields_ = [("mean", _fp), ("var", _fp), ("count", _fp)]


MOMENT_COLS = 131   # include/bb_ppo.h BB_PPO_MOMENT_COLS


class BBPpoBatch(ctypes.Structure):
    _fields_ = [("obs", _fp), ("actions", _fp), ("log_probs", _fp), ("values", _fp), ("advantages", _fp), ("returns", _fp),
                ("indices", _fp), ("num_rows", ctypes.c_int64), ("clip_coef", ctypes.c_float), ("ent_coef", ctypes.c_float),
                ("vf_coef", ctypes.c_float), ("indices32", _fp)]


def _bind(lib):
    if getattr(lib, "_ppo_bound", False):
        return lib
    _bind_rollout(lib)
    vp, f32 = ctypes.c_void_p, ctypes.c_float
    lib.bb_ppo_workspace_bytes.restype = ctypes.c_int64
    lib.bb_ppo_workspace_bytes.argtypes = []
    lib.bb_ppo_grad.restype = ctypes.c_int
    lib.bb_ppo_grad.argtypes = [ctypes.POINTER(BBPolicy), ctypes.POINTER(BBPpoBatch), vp, vp, vp, vp]
    lib.bb_ppo_adam.restype = ctypes.c_int
    lib.bb_ppo_adam.argtypes = [vp, vp, vp, vp, vp, ctypes.c_int64, f32, f32, f32, f32, f32, f32, vp, vp]
    lib.bb_ppo_moments_workspace_bytes.restype = ctypes.c_int64
    lib.bb_ppo_moments_workspace_bytes.argtypes = []
    lib.bb_ppo_moments.restype = ctypes.c_int
    lib.bb_ppo_moments.argtypes = [vp, vp, vp, vp, ctypes.c_int64, vp, vp, vp]
    lib.bb_ppo_finish_advantages.restype = ctypes.c_int
    lib.bb_ppo_finish_advantages.argtypes = [vp, ctypes.c_double, ctypes.POINTER(BBRunningNorm), ctypes.POINTER(BBRunningNorm), vp, vp,
                                             f32, vp, vp, vp, ctypes.c_int64, vp, vp]
    lib.bb_randperm_workspace_bytes.restype = ctypes.c_int64
    lib.bb_randperm_workspace_bytes.argtypes = [ctypes.c_int64]
    lib.bb_randperm.restype = ctypes.c_int
    lib.bb_randperm.argtypes = [ctypes.c_int64, ctypes.c_uint32, ctypes.c_uint32, vp, ctypes.c_int64, vp, vp]
    lib.bb_p2p_create.restype = vp
    lib.bb_p2p_create.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_int64, vp]
    lib.bb_p2p_connect.restype = ctypes.c_int
    lib.bb_p2p_connect.argtypes = [vp, vp]
    lib.bb_p2p_destroy.restype = None
    lib.bb_p2p_destroy.argtypes = [vp]
    lib.bb_ppo_adam_p2p.restype = ctypes.c_int
    lib.bb_ppo_adam_p2p.argtypes = [vp, vp, vp, vp, vp, vp, ctypes.c_int64, f32, f32, f32, f32, f32, vp, vp, ctypes.c_int32, vp]
    lib._ppo_bound = True
    return lib


def flatten_parameters_(agent):
    """Re-binds the module's parameters as views of one flat float32 vector (in PARAM_ORDER) and returns it."""
    import torch

    named = dict(agent.named_parameters())
    assert list(named) == PARAM_ORDER, f"unexpected parameter list {list(named)}"
    dev = named[PARAM_ORDER[0]].device
    flat = torch.empty(NUM_PARAMS, dtype=torch.float32, device=dev)
    off = 0
    with torch.no_grad():
        for k in PARAM_ORDER:
            p = named[k]
            n = p.numel()
            flat[off:off + n].copy_(p.detach().reshape(-1))
            p.data = flat[off:off + n].view(p.shape)
            off += n
    assert off == NUM_PARAMS, "Agent must be Agent(input_dim=128, num_channels=32, num_layers=2)"
    return flat


class FusedPPO:
    """clip-PPO minibatch steps on the flat parameter vector of an `Agent` (policy_torch.Agent / reference Agent)."""

    def __init__(self, agent, lr: float = 3e-4, betas=(0.9, 0.999), eps: float = 1e-8, clip_coef: float = 0.2,
                 ent_coef: float = 0.01, vf_coef: float = 1.0, max_grad_norm: float = 1.0, seed: int = 0):
        import torch

        self._lib = _bind(_lib.load())
        self.agent = agent
        self.flat = flatten_parameters_(agent)
        dev = self.device = self.flat.device
        assert dev.type == "cuda", "the fused PPO update has no CPU path"
        self.weights = PolicyWeights.from_module(agent)      # pointers into self.flat + the normaliser vectors
        self.exp_avg = torch.zeros_like(self.flat)
        self.exp_avg_sq = torch.zeros_like(self.flat)
        self.step = torch.zeros(1, dtype=torch.int32, device=dev)
        self.grad = torch.zeros_like(self.flat)
        self.stats = torch.zeros(NUM_STATS, dtype=torch.float32, device=dev)
        self.grad_norm = torch.zeros(1, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            nbytes = int(self._lib.bb_ppo_workspace_bytes())
        if nbytes <= 0:
            raise RuntimeError("bb_ppo_workspace_bytes failed: " + _lib.last_error())
        self.workspace = torch.empty(nbytes // 4, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            mbytes = int(self._lib.bb_ppo_moments_workspace_bytes())
        self.mom_workspace = torch.empty(mbytes // 8, dtype=torch.float64, device=dev)
        self.mom_sums = torch.zeros(2 * MOMENT_COLS, dtype=torch.float64, device=dev)
        self.mom_scratch = torch.zeros(4, dtype=torch.float32, device=dev)
        self.lr, self.betas, self.eps = float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.clip_coef, self.ent_coef, self.vf_coef = float(clip_coef), float(ent_coef), float(vf_coef)
        self.max_grad_norm = float(max_grad_norm)
        self.seed, self._perms = int(seed) & 0xFFFFFFFF, 0   # bb_randperm stream: one draw per epoch
        self._perm_ws, self._perm_out = None, None
        self._p2p, self._p2p_timeout_ms = None, 2000

    def enable_p2p(self, timeout_ms: int = 2000) -> None:
        """Multi-GPU: sum the gradient over NVLink peer memory inside the optimiser kernel (bb_ppo_adam_p2p) instead of one
        NCCL all-reduce per minibatch step.  Collective call: every rank of the (single-node) default process group makes it
        once, after torch.distributed is initialised; the 64-byte CUDA IPC handles travel through all_gather_object."""
        import torch
        import torch.distributed as dist

        assert dist.is_available() and dist.is_initialized(), "enable_p2p needs an initialised process group"
        ws, rank = dist.get_world_size(), dist.get_rank()
        handle = (ctypes.c_ubyte * 64)()
        with torch.cuda.device(self.device):
            ctx = self._lib.bb_p2p_create(rank, ws, NUM_PARAMS, handle)
        ok = [None] * ws
        dist.all_gather_object(ok, bytes(handle) if ctx else None)
        if any(h is None for h in ok):
            if ctx:
                self._lib.bb_p2p_destroy(ctx)
            raise RuntimeError("bb_p2p_create failed on some rank: " + _lib.last_error())
        blob = (ctypes.c_ubyte * (64 * ws)).from_buffer_copy(b"".join(ok))
        with torch.cuda.device(self.device):
            rc = self._lib.bb_p2p_connect(ctx, blob)
        err = _lib.last_error() if rc != 0 else ""
        rcs = [None] * ws
        dist.all_gather_object(rcs, rc)          # also the barrier: every buffer is mapped before the first publish
        if any(r != 0 for r in rcs):
            self._lib.bb_p2p_destroy(ctx)
            raise RuntimeError("bb_p2p_connect failed on some rank (CUDA IPC / peer access unavailable?): " + err)
        self._p2p, self._p2p_timeout_ms = ctx, int(timeout_ms)
        self.p2p_error = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._p2p_error_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._p2p_error_event = None

    def disable_p2p(self) -> None:
        """Collective: unmaps the peers' buffers (all ranks must have finished their updates)."""
        import torch
        import torch.distributed as dist

        if self._p2p is None:
            return
        torch.cuda.synchronize(self.device)
        self.check_p2p()
        dist.barrier()
        self._lib.bb_p2p_destroy(self._p2p)
        self._p2p = None

    def check_p2p(self) -> None:
        """Raises if an earlier bb_ppo_adam_p2p gave up waiting for a peer (its Adam step was NOT applied)."""
        if self._p2p is None or self._p2p_error_event is None:
            return
        self._p2p_error_event.synchronize()
        if int(self._p2p_error_host[0]) != 0:
            raise RuntimeError(f"bb_ppo_adam_p2p: a peer rank did not publish its gradient within {self._p2p_timeout_ms} ms; "
                               "the parameters of the ranks have diverged")

    def apply_p2p(self) -> None:
        """all-reduce (mean over ranks, rank-order sum over NVLink peer memory) + clip_grad_norm_ + Adam in one kernel."""
        rc = self._lib.bb_ppo_adam_p2p(self._p2p, self.flat.data_ptr(), self.grad.data_ptr(), self.exp_avg.data_ptr(),
                                       self.exp_avg_sq.data_ptr(), self.step.data_ptr(), NUM_PARAMS, self.max_grad_norm, self.lr,
                                       self.betas[0], self.betas[1], self.eps, self.grad_norm.data_ptr(), self.p2p_error.data_ptr(),
                                       self._p2p_timeout_ms, self._stream())
        if rc != 0:
            raise RuntimeError("bb_ppo_adam_p2p failed: " + _lib.last_error())

    def randperm(self, n: int):
        """torch.randperm(n) of ppo.py:186 as bb_randperm (Threefry keys + radix sort): int32 [n], a new draw per call."""
        import torch

        if self._perm_out is None or self._perm_out.numel() != n:
            nbytes = int(self._lib.bb_randperm_workspace_bytes(n))
            if nbytes <= 0:
                raise RuntimeError("bb_randperm_workspace_bytes failed")
            self._perm_ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            self._perm_out = torch.empty(n, dtype=torch.int32, device=self.device)
        rc = self._lib.bb_randperm(n, self.seed, self._perms & 0xFFFFFFFF, self._perm_ws.data_ptr(), self._perm_ws.numel(),
                                   self._perm_out.data_ptr(), self._stream())
        if rc != 0:
            raise RuntimeError("bb_randperm failed: " + _lib.last_error())
        self._perms += 1
        return self._perm_out

    def _stream(self):
        import torch

        return torch.cuda.current_stream(self.device).cuda_stream

    def gradient(self, obs, actions, log_probs, values, advantages, returns, indices=None):
        """d loss / d parameters of one minibatch (ppo.py:189-211) into self.grad (flat); loss sums into self.stats.
        All tensors are the flattened rollout buffers ([R,128], [R,6], [R]...); `indices` (int64 [m]) selects the rows."""
        import torch

        for t in (obs, actions, log_probs, values, advantages, returns):
            assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
        b = BBPpoBatch()
        b.obs, b.actions, b.log_probs = obs.data_ptr(), actions.data_ptr(), log_probs.data_ptr()
        b.values, b.advantages, b.returns = values.data_ptr(), advantages.data_ptr(), returns.data_ptr()
        b.indices, b.indices32 = None, None
        if indices is not None:
            assert indices.dtype in (torch.int64, torch.int32) and indices.is_contiguous() and indices.is_cuda
            if indices.dtype == torch.int64:
                b.indices = indices.data_ptr()
            else:
                b.indices32 = indices.data_ptr()
            b.num_rows = indices.numel()
        else:
            b.num_rows = obs.shape[0]
        b.clip_coef, b.ent_coef, b.vf_coef = self.clip_coef, self.ent_coef, self.vf_coef
        rc = self._lib.bb_ppo_grad(ctypes.byref(self.weights.c), ctypes.byref(b), self.workspace.data_ptr(), self.grad.data_ptr(),
                                   self.stats.data_ptr(), self._stream())
        if rc != 0:
            raise RuntimeError("bb_ppo_grad failed: " + _lib.last_error())
        self._last_rows = int(b.num_rows)
        return self.grad

    def apply(self, grad_scale: float = 1.0) -> None:
        """clip_grad_norm_ + Adam step (ppo.py:212-214) with self.grad * grad_scale."""
        rc = self._lib.bb_ppo_adam(self.flat.data_ptr(), self.grad.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                                   self.step.data_ptr(), NUM_PARAMS, float(grad_scale), self.max_grad_norm, self.lr,
                                   self.betas[0], self.betas[1], self.eps, self.grad_norm.data_ptr(), self._stream())
        if rc != 0:
            raise RuntimeError("bb_ppo_adam failed: " + _lib.last_error())

    def finish_advantages(self, buf) -> None:
        """The part of compute_advantages after the GAE scan (ppo.py:167-176; same contract as
        ppo_dist.finish_advantages): expects buf.values_unnorm / advantages / returns from FusedRollout.gae; updates
        agent.obs_norm and agent.value_norm (statistics over the rows of ALL ranks, one all-reduce of 262 doubles),
        leaves advantages normalised and values_unnorm / returns in the value normaliser's scale, and refreshes the
        float32 normaliser vectors self.weights points the kernels at.  Four kernels, no host synchronisation."""
        import torch
        import torch.distributed as dist

        a = self.agent
        for nrm in (a.obs_norm, a.value_norm):
            assert nrm.mean.dtype == torch.float64 and nrm.mean.is_cuda and nrm.count.dtype == torch.float64
        rows = buf.horizon * buf.n_envs
        st = self._stream()
        rc = self._lib.bb_ppo_moments(buf.obs.data_ptr(), buf.values_unnorm.data_ptr(), buf.returns.data_ptr(),
                                      buf.advantages.data_ptr(), rows, self.mom_workspace.data_ptr(), self.mom_sums.data_ptr(), st)
        if rc != 0:
            raise RuntimeError("bb_ppo_moments failed: " + _lib.last_error())
        ws = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
        if ws > 1:
            dist.all_reduce(self.mom_sums, op=dist.ReduceOp.SUM)   # every rank holds the same number of rows
        on, vn = BBRunningNorm(), BBRunningNorm()
        on.mean, on.var, on.count = a.obs_norm.mean.data_ptr(), a.obs_norm.var.data_ptr(), a.obs_norm.count.data_ptr()
        vn.mean, vn.var, vn.count = a.value_norm.mean.data_ptr(), a.value_norm.var.data_ptr(), a.value_norm.count.data_ptr()
        w = self.weights
        rc = self._lib.bb_ppo_finish_advantages(self.mom_sums.data_ptr(), float(rows * ws), ctypes.byref(on), ctypes.byref(vn),
                                                w._t["obs_mean"].data_ptr(), w._t["obs_inv_std"].data_ptr(),
                                                float(getattr(a.value_norm, "clamp", 5.0)), buf.values_unnorm.data_ptr(),
                                                buf.returns.data_ptr(), buf.advantages.data_ptr(), rows, self.mom_scratch.data_ptr(), st)
        if rc != 0:
            raise RuntimeError("bb_ppo_finish_advantages failed: " + _lib.last_error())

    def update(self, buf, num_minibatches: int = 4, update_epochs: int = 4, generator=None,
               refresh_normalizer: bool = True, permutation: str = "fused") -> Dict[str, "object"]:
        """ppo.py:180-216 on this rank's rollout buffers (rollout.RolloutBuffers after finish_advantages).  With
        torch.distributed initialised the flat gradient is summed across ranks (one all-reduce per step) and averaged."""
        import torch
        import torch.distributed as dist

        ws = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
        total = buf.horizon * buf.n_envs
        mb = total // num_minibatches
        obs = buf.obs.view(-1, buf.obs.shape[-1])
        act = buf.actions.view(-1, buf.actions.shape[-1])
        lp, v = buf.log_probs.view(-1), buf.values.view(-1)
        adv, ret = buf.advantages.view(-1), buf.returns.view(-1)
        p2p = ws > 1 and self._p2p is not None
        if p2p:
            self.check_p2p()     # the flag copied at the end of the previous update
        if refresh_normalizer:   # the observation normaliser moved since the rollout (ppo.py:168); finish_advantages() already did it
            self.weights.refresh()
        for _ in range(update_epochs):
            # one permutation of the rollout rows per epoch (ppo.py:186): bb_randperm, or torch's own stream
            perm = self.randperm(total) if permutation == "fused" else torch.randperm(total, device=self.device, generator=generator)
            for start in range(0, total, mb):   # ppo.py:188: a remainder minibatch when total % num_minibatches != 0
                self.gradient(obs, act, lp, v, adv, ret, perm[start:start + mb])
                if p2p:
                    self.apply_p2p()
                    continue
                if ws > 1:
                    dist.all_reduce(self.grad, op=dist.ReduceOp.SUM)
                self.apply(1.0 / ws)
        if p2p:
            self._p2p_error_host.copy_(self.p2p_error, non_blocking=True)
            self._p2p_error_event = torch.cuda.Event()
            self._p2p_error_event.record(torch.cuda.current_stream(self.device))
            self.check_p2p()     # before anything (the next rollout) consumes the weights: a timed-out step raises HERE
        m = float(self._last_rows)   # rows of the last minibatch (the remainder one when total % num_minibatches != 0)
        return {"pg_loss": self.stats[0] / m, "v_loss": self.stats[1] / m * self.vf_coef,
                "ent_loss": -self.stats[2] / m * self.ent_coef, "clip_frac": self.stats[3] / m, "grad_norm": self.grad_norm[0]}
