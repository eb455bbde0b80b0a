"""World_size-2 gloo tests of the multi-GPU PPO plumbing (madrona_basketball_b200/ppo_dist.py): the
reductions must reproduce the single-process arithmetic of scripts/agent.py:41-51 and
scripts/ppo.py:173-214 on the concatenated batch."""
import os
import socket

import pytest
import torch


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _data():
    g = torch.Generator().manual_seed(5)
    x = torch.randn(600, 128, generator=g) * 3 + 1          # rows of "observations"
    adv = torch.randn(600, generator=g) * 2 + 0.3
    act = torch.stack([torch.randint(0, b, (600,), generator=g) for b in [2, 8, 3, 2, 2, 2]], 1).float()
    return x, adv, act


def _worker(rank, ws, port, q):
    import torch.distributed as dist

    from madrona_basketball_b200 import ppo_dist
    from madrona_basketball_b200.policy_torch import Agent

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        x, adv, act = _data()
        lo, hi = (0, 250) if rank == 0 else (250, 600)     # unequal shards
        xs, advs, acts = x[lo:hi], adv[lo:hi].clone(), act[lo:hi]
        mean, var, n = ppo_dist.global_moments(xs)
        torch.manual_seed(100 + rank)                       # replicas start different ...
        agent = Agent()
        ppo_dist.broadcast_module_(agent)                   # ... and are made identical
        ppo_dist.update_normalizer(agent.obs_norm, xs)
        ppo_dist.normalize_advantages_(advs)
        # one gradient step on equal-size shards == the full-batch step
        e_lo, e_hi = (0, 300) if rank == 0 else (300, 600)
        lp, ent, v = agent.get_stats(x[e_lo:e_hi], act[e_lo:e_hi])
        loss = (-(lp * adv[e_lo:e_hi]).mean() + 0.5 * (v ** 2).mean() - 0.01 * ent.mean())
        loss.backward()
        ppo_dist.all_reduce_gradients(agent.parameters())
        grads = torch.cat([p.grad.reshape(-1) for p in agent.parameters()])
        w0 = torch.cat([p.detach().reshape(-1) for p in agent.parameters()])
        np_ = lambda d: {k: (v.detach().cpu().numpy().copy() if torch.is_tensor(v) else
                             ({kk: vv.detach().cpu().numpy().copy() for kk, vv in v.items()} if isinstance(v, dict) else v))
                         for k, v in d.items()}   # plain numpy through the queue (no shared-memory handles)
        if rank == 0:
            q.put(np_({"mean": mean, "var": var, "n": n, "obs_mean": agent.obs_norm.mean.clone(), "obs_var": agent.obs_norm.var.clone(),
                   "obs_count": agent.obs_norm.count.clone(), "adv0": advs.clone(), "grads": grads, "w0": w0,
                   "state": {k: v.clone() for k, v in agent.state_dict().items()}}))
        else:
            q.put(np_({"w0_rank1": w0, "adv1": advs.clone()}))
    finally:
        dist.destroy_process_group()


def test_two_rank_reductions_match_single_process():
    import torch.multiprocessing as mp

    from madrona_basketball_b200 import ppo_dist
    from madrona_basketball_b200.policy_torch import Agent

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(2):
        got.update(q.get(timeout=240))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    T = torch.from_numpy
    got = {k: (T(v) if hasattr(v, "dtype") and hasattr(v, "shape") and getattr(v, "ndim", 0) >= 0 and not isinstance(v, float) and not isinstance(v, dict) else
               ({kk: T(vv) for kk, vv in v.items()} if isinstance(v, dict) else v)) for k, v in got.items()}
    x, adv, act = _data()
    var, mean = torch.var_mean(x.double(), dim=0, unbiased=True)
    assert got["n"] == 600.0
    assert torch.allclose(got["mean"], mean, rtol=1e-12, atol=1e-12) and torch.allclose(got["var"], var, rtol=1e-10)
    # RunningMeanStd.update on the concatenated batch (agent.py:41-51)
    ref = Agent()
    ref.obs_norm.update(x)
    assert torch.allclose(got["obs_mean"], ref.obs_norm.mean, rtol=1e-6, atol=1e-7)
    assert torch.allclose(got["obs_var"], ref.obs_norm.var, rtol=1e-6) and got["obs_count"] == ref.obs_norm.count
    # advantage normalisation (ppo.py:173-174)
    want = (adv - adv.mean()) / (adv.std() + 1e-8)
    assert torch.allclose(torch.cat([got["adv0"], got["adv1"]]), want, rtol=1e-5, atol=1e-6)
    # replicas identical after the broadcast; averaged shard gradients == full-batch gradient
    assert torch.equal(got["w0"], got["w0_rank1"])
    ref.load_state_dict(got["state"])
    ref.obs_norm.mean.copy_(got["obs_mean"]); ref.obs_norm.var.copy_(got["obs_var"])
    ref.zero_grad()
    lp, ent, v = ref.get_stats(x, act)
    ((-(lp * adv).mean() + 0.5 * (v ** 2).mean() - 0.01 * ent.mean())).backward()
    full = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
    assert torch.allclose(got["grads"], full, rtol=1e-4, atol=1e-6)


def test_single_process_fallbacks_and_policy_stats():
    """Without a process group the helpers are the plain formulas; Agent.get_stats agrees with torch's Categorical."""
    from torch.distributions.categorical import Categorical

    from madrona_basketball_b200 import ppo_dist
    from madrona_basketball_b200.policy_torch import ACTION_BUCKETS, Agent

    x, adv, act = _data()
    mean, var, n = ppo_dist.global_moments(x)
    v2, m2 = torch.var_mean(x.double(), dim=0, unbiased=True)
    assert n == 600.0 and torch.allclose(mean, m2) and torch.allclose(var, v2)
    agent = Agent()
    with torch.no_grad():
        agent.actor.weight.mul_(50)
    lp, ent, v = agent.get_stats(x, act)
    logits = agent.actor(agent.backbone(agent.obs_norm(x)))
    off, lp_ref, ent_ref = 0, 0, 0
    for i, nb in enumerate(ACTION_BUCKETS):
        d = Categorical(logits=logits[:, off:off + nb])
        lp_ref = lp_ref + d.log_prob(act[:, i]); ent_ref = ent_ref + d.entropy(); off += nb
    assert torch.allclose(lp, lp_ref, rtol=1e-5, atol=1e-6) and torch.allclose(ent, ent_ref, rtol=1e-5, atol=1e-6)
    sd = agent.state_dict()
    for k in ["backbone.0.weight", "backbone.1.weight", "backbone.3.bias", "backbone.4.bias", "actor.weight", "critic.bias",
              "obs_norm.mean", "obs_norm.var", "obs_norm.count", "value_norm.mean"]:
        assert k in sd, k     # the reference's checkpoint keys (scripts/agent.py)
