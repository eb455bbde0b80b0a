"""Differentiable torch policy with the reference's layout (scripts/agent.py:108-182, scripts/action.py).

The fused kernels (rollout.py) cover inference; training needs autograd, which stays in torch exactly as
in the reference (ppo.py:180-216).  Parameter / buffer names equal the reference's state_dict keys, so
checkpoints move both ways, and `PolicyWeights.from_module(agent)` points the kernels at these tensors.
"""
from __future__ import annotations

import torch
import torch.nn as nn

ACTION_BUCKETS = [2, 8, 3, 2, 2, 2]  # scripts/env.py:102


class RunningMeanStd(nn.Module):
    """agent.py:18-51.  `update` takes precomputed batch moments so that callers can make them global
    across ranks first (ppo_dist.global_moments)."""

    def __init__(self, dim: int, clamp: float = 0.0):
        super().__init__()
        self.epsilon, self.clamp = 1e-5, clamp
        self.register_buffer("mean", torch.zeros(dim, dtype=torch.float64))
        self.register_buffer("var", torch.ones(dim, dtype=torch.float64))
        self.register_buffer("count", torch.ones((), dtype=torch.float64))

    def forward(self, x, unnorm: bool = False):
        mean = self.mean.to(torch.float32)
        var = self.var.to(torch.float32) + self.epsilon
        if unnorm:
            if self.clamp:
                x = torch.clamp(x, min=-self.clamp, max=self.clamp)
            return mean + torch.sqrt(var) * x
        x = (x - mean) * torch.rsqrt(var)
        return torch.clamp(x, min=-self.clamp, max=self.clamp) if self.clamp else x

    @torch.no_grad()
    def update_from_moments(self, mean, var, count) -> None:
        """agent.py:41-51 with (var, mean, count) of the batch given (unbiased variance, like torch.var_mean)."""
        mean, var = mean.to(torch.float64), var.to(torch.float64)
        count_ = count + self.count
        delta = mean - self.mean
        m = self.var * self.count + var * count + delta ** 2 * self.count * count / count_
        self.mean.copy_(self.mean + delta * count / count_)
        self.var.copy_(m / count_)
        self.count.copy_(count_)

    @torch.no_grad()
    def update(self, x) -> None:
        x = x.reshape(-1, x.size(-1))
        var, mean = torch.var_mean(x, dim=0, unbiased=True)
        self.update_from_moments(mean, var, float(x.size(0)))


def _backbone_init(layer):   # agent.py:98-101
    torch.nn.init.kaiming_normal_(layer.weight, nn.init.calculate_gain("relu"))
    torch.nn.init.constant_(layer.bias, 0.0)
    return layer


def _head_init(layer):       # agent.py:103-106
    torch.nn.init.orthogonal_(layer.weight, gain=0.01)
    torch.nn.init.constant_(layer.bias, 0.0)
    return layer


class Agent(nn.Module):
    def __init__(self, input_dim: int = 128, num_channels: int = 32, num_layers: int = 2, action_buckets=ACTION_BUCKETS):
        """Same positional signature as the reference (agent.py:109); the fused kernels cover num_layers == 2
        (the only value the reference's scripts pass: env.py:107, ppo.py:257)."""
        super().__init__()
        if num_layers != 2:
            raise ValueError("madrona_basketball_b200 supports the reference's 2-layer policy only")
        self.action_buckets = list(action_buckets)
        self.backbone = nn.Sequential(
            _backbone_init(nn.Linear(input_dim, num_channels)), nn.LayerNorm(num_channels), nn.ReLU(),
            _backbone_init(nn.Linear(num_channels, num_channels)), nn.LayerNorm(num_channels), nn.ReLU())
        self.actor = _head_init(nn.Linear(num_channels, sum(self.action_buckets)))
        self.critic = _head_init(nn.Linear(num_channels, 1))
        self.obs_norm = RunningMeanStd(input_dim, clamp=5.0)
        self.value_norm = RunningMeanStd(1, clamp=5.0)

    def _features(self, obs):
        return self.backbone(self.obs_norm(obs))

    def _bucket_log_probs(self, logits):
        out, off = [], 0
        for nb in self.action_buckets:
            sl = logits[:, off:off + nb]
            out.append(sl - sl.logsumexp(dim=-1, keepdim=True))
            off += nb
        return out

    def evaluate(self, obs):
        return self.critic(self._features(obs)).squeeze(-1)

    def get_stats(self, obs, actions):
        """agent.py:171-178: summed log-prob of `actions`, summed entropy, value."""
        x = self._features(obs)
        lsm = self._bucket_log_probs(self.actor(x))
        actions = actions.long()
        logp = sum(l.gather(1, actions[:, i:i + 1]).squeeze(1) for i, l in enumerate(lsm))
        ent = sum(-(l.exp() * l).sum(dim=-1) for l in lsm)
        return logp, ent, self.critic(x).squeeze(-1)
