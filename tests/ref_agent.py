"""Plain-torch fp32 reference of the policy the reference trains (scripts/agent.py:108-154,
scripts/action.py), written for the tests: same layer layout and state_dict key names, so a
checkpoint of the reference loads into it.  Test infrastructure only."""
import torch
import torch.nn as nn

ACTION_BUCKETS = [2, 8, 3, 2, 2, 2]  # scripts/env.py:102


class RefRunningMeanStd(nn.Module):  # agent.py:18-51
    def __init__(self, dim, clamp=0.0):
        super().__init__()
        self.epsilon, self.clamp = 1e-5, clamp
        self.register_buffer("mean", torch.zeros(dim, dtype=torch.float64))
        self.register_buffer("var", torch.ones(dim, dtype=torch.float64))
        self.register_buffer("count", torch.ones((), dtype=torch.float64))

    def forward(self, x, unnorm=False):
        mean = self.mean.to(torch.float32)
        var = self.var.to(torch.float32) + self.epsilon
        if unnorm:
            if self.clamp:
                x = torch.clamp(x, min=-self.clamp, max=self.clamp)
            return mean + torch.sqrt(var) * x
        x = (x - mean) * torch.rsqrt(var)
        return torch.clamp(x, min=-self.clamp, max=self.clamp) if self.clamp else x


class RefAgent(nn.Module):
    def __init__(self, input_dim=128, num_channels=32, buckets=ACTION_BUCKETS):
        super().__init__()
        self.buckets = buckets
        self.backbone = nn.Sequential(nn.Linear(input_dim, num_channels), nn.LayerNorm(num_channels), nn.ReLU(),
                                      nn.Linear(num_channels, num_channels), nn.LayerNorm(num_channels), nn.ReLU())
        self.actor = nn.Linear(num_channels, sum(buckets))
        self.critic = nn.Linear(num_channels, 1)
        self.obs_norm = RefRunningMeanStd(input_dim, clamp=5.0)
        self.value_norm = RefRunningMeanStd(1, clamp=5.0)

    def logits_value(self, obs):
        x = self.backbone(self.obs_norm(obs))
        return self.actor(x), self.critic(x).squeeze(-1)

    def bucket_log_softmax(self, logits):
        out, off = [], 0
        for nb in self.buckets:
            sl = logits[:, off:off + nb]
            out.append(sl - sl.logsumexp(dim=-1, keepdim=True))   # Categorical(logits=...) normalisation
            off += nb
        return out


def randomize(agent: RefAgent, seed: int = 0, scale_heads: float = 1.0):
    """Non-trivial parameters / normaliser statistics (the reference's init makes the heads ~0.01)."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for p in agent.parameters():
            p.copy_(torch.randn(p.shape, generator=g) * (0.3 if p.dim() > 1 else 0.1))
        agent.backbone[1].weight.add_(1.0); agent.backbone[4].weight.add_(1.0)
        agent.actor.weight.mul_(scale_heads)
        agent.obs_norm.mean.copy_(torch.randn(128, generator=g, dtype=torch.float64) * 3.0)
        agent.obs_norm.var.copy_(torch.rand(128, generator=g, dtype=torch.float64) * 20.0 + 0.01)
        agent.value_norm.mean.fill_(1.7); agent.value_norm.var.fill_(9.0)
    return agent
