"""Multi-GPU PPO around the sharded step path (new capability: the reference is
single-process, scripts/ppo.py:226-231).

One process per GPU, worlds sharded by rank (shard.py), policy replicated.  The step and the rollout have
no collective.  What must agree across ranks is reduced with as few calls as possible, because every
payload is tiny (the policy has ~6 k parameters = 24 KB; NVLink bandwidth is irrelevant, call count is not):

  * gradients          one flat all-reduce per optimiser step (ppo.py:211-214)
  * normaliser moments one flat all-reduce per RunningMeanStd.update (agent.py:41-51)
  * advantage moments  one flat all-reduce per iteration (ppo.py:173-174)
  * episode statistics shard.reduce_episode_stats

All functions degrade to the single-process arithmetic when torch.distributed is not initialised; with the
gloo backend they run on CPU (tests/test_ppo_dist_cpu.py), with nccl on the GPUs.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterable, List, Tuple

import torch
import torch.distributed as dist


def _on() -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def flat_all_reduce_(tensors: List[torch.Tensor], average: bool = False) -> None:
    """Sum (or average) a list of tensors across ranks with ONE collective: pack, all-reduce, unpack in place."""
    if not _on() or not tensors:
        return
    flat = torch.cat([t.reshape(-1).to(torch.float64 if t.dtype == torch.float64 else torch.float32) for t in tensors])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    if average:
        flat /= dist.get_world_size()
    off = 0
    for t in tensors:
        n = t.numel()
        t.copy_(flat[off:off + n].view_as(t).to(t.dtype))
        off += n


def global_moments(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, float]:
    """(mean, unbiased variance, count) over dim 0 of the rows of ALL ranks — what torch.var_mean(unbiased=True)
    gives on the concatenated batch.  Per-rank (count, mean, M2) are merged with Chan's formula in float64."""
    x = x.reshape(-1, x.size(-1)).to(torch.float64)
    n = float(x.size(0))
    mean = x.mean(dim=0)
    m2 = ((x - mean) ** 2).sum(dim=0)
    if not _on():
        return mean, m2 / max(n - 1.0, 1.0), n
    d = x.size(1)
    pack = torch.cat([torch.tensor([n], dtype=torch.float64, device=x.device), n * mean, m2, n * mean * mean])
    dist.all_reduce(pack, op=dist.ReduceOp.SUM)
    n_all = float(pack[0].item())
    g_mean = pack[1:1 + d] / n_all
    # sum_i [M2_i + n_i (mean_i - g)^2] = sum M2_i + sum n_i mean_i^2 - n_all g^2
    g_m2 = pack[1 + d:1 + 2 * d] + pack[1 + 2 * d:1 + 3 * d] - n_all * g_mean * g_mean
    return g_mean, g_m2.clamp_min(0.0) / max(n_all - 1.0, 1.0), n_all


@torch.no_grad()
def update_normalizer(norm, x: torch.Tensor) -> None:
    """RunningMeanStd.update (agent.py:41-51) over the rows of every rank."""
    mean, var, count = global_moments(x)
    norm.update_from_moments(mean, var, count)


@torch.no_grad()
def normalize_advantages_(adv: torch.Tensor) -> None:
    """ppo.py:173-174 with global mean / unbiased std."""
    mean, var, _ = global_moments(adv.reshape(-1, 1))
    adv.sub_(mean.to(adv.dtype)).div_(var.sqrt().to(adv.dtype) + 1e-8)


def all_reduce_gradients(params: Iterable[torch.nn.Parameter]) -> None:
    """Average .grad across ranks with one flat all-reduce (the DDP step, bucket = the whole model)."""
    grads = [p.grad for p in params if p.grad is not None]
    flat_all_reduce_(grads, average=True)


def broadcast_module_(module: torch.nn.Module, src: int = 0) -> None:
    """Make parameters and buffers identical on every rank (rank `src` wins), one flat broadcast per dtype."""
    if not _on():
        return
    by_dtype = {}
    for t in list(module.parameters()) + list(module.buffers()):
        by_dtype.setdefault(t.dtype, []).append(t.data)
    for ts in by_dtype.values():
        flat = torch.cat([t.reshape(-1) for t in ts])
        dist.broadcast(flat, src=src)
        off = 0
        for t in ts:
            t.copy_(flat[off:off + t.numel()].view_as(t))
            off += t.numel()


@dataclass
class PPOConfig:   # scripts/ppo.py:41-55 defaults
    learning_rate: float = 3e-4
    gamma: float = 0.998
    gae_lambda: float = 0.95
    num_minibatches: int = 4
    update_epochs: int = 4
    clip_coef: float = 0.2
    ent_coef: float = 0.01
    vf_coef: float = 1.0
    max_grad_norm: float = 1.0


@torch.no_grad()
def finish_advantages(agent, buf) -> None:
    """The part of compute_advantages after the GAE scan (ppo.py:167-176), with global statistics.
    Expects buf.values_unnorm / advantages / returns from FusedRollout.gae; leaves advantages normalised,
    buf.values_unnorm and buf.returns in the value normaliser's scale (the reference's `values`, `returns`)."""
    update_normalizer(agent.obs_norm, buf.obs.view(-1, buf.obs.shape[-1]))
    update_normalizer(agent.value_norm, buf.values_unnorm.view(-1, 1))
    update_normalizer(agent.value_norm, buf.returns.view(-1, 1))
    normalize_advantages_(buf.advantages)
    buf.values_unnorm.copy_(agent.value_norm(buf.values_unnorm))
    buf.returns.copy_(agent.value_norm(buf.returns))


def update_policy(agent, optimizer, buf, cfg: PPOConfig, generator=None) -> dict:
    """ppo.py:180-216 on this rank's shard of the rollout; gradients are averaged across ranks before the clip
    and the optimiser step, so every replica takes the same step.  Returns the last minibatch's losses."""
    total = buf.horizon * buf.n_envs
    mb = total // cfg.num_minibatches
    obs = buf.obs.view(-1, buf.obs.shape[-1])
    act = buf.actions.view(-1, buf.actions.shape[-1])
    lp_old, v_old = buf.log_probs.view(-1), buf.values.view(-1)
    adv, ret = buf.advantages.view(-1), buf.returns.view(-1)
    stats = {}
    params = list(agent.parameters())
    for _ in range(cfg.update_epochs):
        perm = torch.randperm(total, device=obs.device, generator=generator)
        for start in range(0, total, mb):
            idx = perm[start:start + mb]
            lp, ent, v = agent.get_stats(obs[idx], act[idx])
            ratio = torch.exp(lp - lp_old[idx])
            a = adv[idx]
            pg_loss = torch.max(-a * ratio, -a * torch.clamp(ratio, 1 - cfg.clip_coef, 1 + cfg.clip_coef)).mean()
            vo, r = v_old[idx], ret[idx]
            v_clip = vo + (v - vo).clamp(-cfg.clip_coef, cfg.clip_coef)
            v_loss = 0.5 * torch.max((v - r) ** 2, (v_clip - r) ** 2).mean() * cfg.vf_coef
            ent_loss = -ent.mean() * cfg.ent_coef
            loss = pg_loss + v_loss + ent_loss
            optimizer.zero_grad(set_to_none=False)
            loss.backward()
            all_reduce_gradients(params)
            torch.nn.utils.clip_grad_norm_(params, cfg.max_grad_norm)
            optimizer.step()
            stats = {"pg_loss": pg_loss.detach(), "v_loss": v_loss.detach(), "ent_loss": ent_loss.detach()}
    return stats
