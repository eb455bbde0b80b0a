"""Data-parallel PPO iterations on the engine (BASELINE.json configs[3]: scripts/ppo.py at 8192 envs per GPU, 32-step
rollouts, 4 epochs x 4 minibatches; /root/reference/scripts/ppo.py:226-376 made one-process-per-GPU), as library
functions so that tools/train_ppo.py AND bench.py (driver-visible multi-rank evidence) run the same code.

The gradient exchange is the path's only collective (ppo.py:180-216 -> one flat all-reduce per optimiser step): NCCL by
default, or the all-reduce fused into the Adam kernel over NVLink peer memory (bb_ppo_adam_p2p) when asked for.
"""
from __future__ import annotations

import time
from typing import Dict, Optional


def run_iterations(envs_per_gpu: int = 8192, rollout_steps: int = 32, iters: int = 20, warmup: int = 3, trainee_idx: int = 1,
                   seed: int = 321, update: str = "fused", allreduce: str = "nccl", log_every: int = 0,
                   save: Optional[str] = None) -> Dict:
    """Runs warmup + iters PPO iterations on this rank's shard (torch.distributed, if initialised, spans the ranks) and
    returns the timing / statistics record (same on every rank; env_steps_per_s is the whole job's)."""
    import torch
    import torch.distributed as dist

    from . import ppo_dist
    from .policy_torch import Agent
    from .rollout import FusedRollout, PolicyWeights, RolloutBuffers
    from .shard import ShardedSimulator, env_rank, reduce_episode_stats

    rank, local, _ = env_rank()
    ws = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    torch.manual_seed(seed)
    n, T = envs_per_gpu, rollout_steps
    shard = ShardedSimulator(n * ws, rank=rank, world_size=ws, gpu_id=local)
    agent = Agent().to(dev)
    ppo_dist.broadcast_module_(agent)
    cfg = ppo_dist.PPOConfig()
    fused = None
    note = ""
    if update == "fused":   # re-binds the parameters as views of one flat vector: before anything takes their pointers
        from .ppo_fused import FusedPPO
        fused = FusedPPO(agent, lr=cfg.learning_rate, eps=1e-8, clip_coef=cfg.clip_coef, ent_coef=cfg.ent_coef,
                         vf_coef=cfg.vf_coef, max_grad_norm=cfg.max_grad_norm, seed=seed + 7919 * rank)
        if ws > 1 and allreduce == "p2p":
            try:
                fused.enable_p2p()
            except RuntimeError as e:   # raised on every rank together (CUDA IPC / peer access unavailable): NCCL all-reduce instead
                allreduce, note = "nccl", f"p2p unavailable ({e}); fell back to nccl"
    else:
        opt = torch.optim.Adam(agent.parameters(), lr=cfg.learning_rate, eps=1e-8)
    weights = fused.weights if fused is not None else PolicyWeights.from_module(agent)
    buf = RolloutBuffers(T, n, dev)
    roll = FusedRollout(shard.sim, weights, trainee_idx, seed=seed, world_offset=shard.world_offset)
    gen = torch.Generator(device=dev).manual_seed(seed + rank)

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    t_roll = t_adv = t_upd = 0.0
    losses = {}
    curve, curve_prev = [], [0.0, 0.0, 0.0]
    wall0 = time.perf_counter()
    for it in range(warmup + iters):
        if it == warmup:
            if ws > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t_roll = t_adv = t_upd = 0.0
            wall0 = time.perf_counter()
        e = [ev() for _ in range(4)]
        e[0].record()
        agent.eval()
        roll.collect(buf)                                               # ppo.py:60-141
        e[1].record()
        roll.gae(buf, cfg.gamma, cfg.gae_lambda, agent.value_norm.mean, agent.value_norm.var)   # ppo.py:144-166
        if fused is not None:
            fused.finish_advantages(buf)                                # ppo.py:167-176, incl. the float32 normaliser vectors
        else:
            ppo_dist.finish_advantages(agent, buf)
            weights.refresh()                                           # the observation normaliser moved
        e[2].record()
        agent.train()
        if fused is not None:                                           # ppo.py:180-216
            losses = fused.update(buf, cfg.num_minibatches, cfg.update_epochs, generator=gen, refresh_normalizer=False)
        else:
            losses = ppo_dist.update_policy(agent, opt, buf, cfg, generator=gen)
        e[3].record()
        torch.cuda.synchronize()
        t_roll += e[0].elapsed_time(e[1]); t_adv += e[1].elapsed_time(e[2]); t_upd += e[2].elapsed_time(e[3])
        if log_every and (it + 1) % log_every == 0:   # learning curve: episodes finished in this window (this rank)
            cur = buf.ep_sums.tolist()
            d = [c - p for c, p in zip(cur, curve_prev)]
            curve_prev = cur
            if d[2] > 0:
                curve.append({"iter": it + 1, "episodes": d[2], "mean_return": d[0] / d[2], "mean_length": d[1] / d[2]})
    wall = time.perf_counter() - wall0
    t = torch.tensor([wall], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    s = buf.ep_sums.tolist()
    stats = reduce_episode_stats({"return_sum": s[0], "length_sum": s[1], "episodes": s[2]})
    # replicas must hold the same parameters after training: max |theta_r - theta_0| over ranks
    flat = torch.cat([p.detach().reshape(-1).float() for p in agent.parameters()])
    spread = 0.0
    if ws > 1:
        ref = flat.clone()
        dist.broadcast(ref, src=0)
        d = (flat - ref).abs().max().reshape(1)
        dist.all_reduce(d, op=dist.ReduceOp.MAX)
        spread = float(d.item())
    if save and rank == 0:
        torch.save(agent.state_dict(), save)
    steps = n * ws * T * iters
    out = {
        "what": "PPO iterations (rollout + GAE + 4x4 minibatch update), ppo.py defaults, fused rollout path, update=" + update
                + (", allreduce=" + allreduce if (ws > 1 and update == "fused") else ""),
        "n_gpus": ws, "envs_per_gpu": n, "rollout_steps": T, "iters": iters,
        "env_steps_per_s": steps / float(t.item()), "s_per_iter": float(t.item()) / iters,
        "ms_per_iter_rank0": {"rollout": t_roll / iters, "advantages_normalisers": t_adv / iters, "update": t_upd / iters},
        "episodes_finished": stats["episodes"], "mean_return": stats["return_sum"] / max(stats["episodes"], 1.0),
        "mean_episode_length": stats["length_sum"] / max(stats["episodes"], 1.0),
        "last_losses": {k: float(v) for k, v in losses.items()},
        "replica_param_spread": spread,
        "learning_curve_rank0": curve,
    }
    if note:
        out["note"] = note
    if fused is not None and ws > 1:
        fused.disable_p2p()
    del roll, buf, shard
    return out


def check_p2p_against_nccl(steps: int = 64) -> Dict:
    """bb_ppo_adam_p2p against NCCL all-reduce + bb_ppo_adam on the same per-rank random gradients (tools/check_p2p.py's
    parity part, callable inside an initialised process group): the p2p parameters must be bit-identical on every rank and
    within float32 summation-order noise of the NCCL path.  Returns {max_abs_diff, ranks_identical, us_* timings}."""
    import torch
    import torch.distributed as dist

    from .policy_torch import Agent
    from .ppo_fused import NUM_PARAMS, FusedPPO
    from .shard import env_rank

    rank, local, _ = env_rank()
    ws = dist.get_world_size()
    dev = torch.device("cuda", local)

    def make():
        torch.manual_seed(1234)
        return FusedPPO(Agent(128, num_channels=32).to(dev), lr=3e-4, max_grad_norm=1.0)

    a, b = make(), make()
    b.enable_p2p(timeout_ms=2000)
    gen = torch.Generator(device=dev)
    gen.manual_seed(100 + rank)
    grads = torch.randn(32, NUM_PARAMS, device=dev, generator=gen) * 0.05
    for i in range(steps):
        g = grads[i % 32]
        a.grad.copy_(g)
        dist.all_reduce(a.grad, op=dist.ReduceOp.SUM)
        a.apply(1.0 / ws)
        b.grad.copy_(g)
        b.apply_p2p()
    torch.cuda.synchronize()
    timed_out = int(b.p2p_error.item()) != 0
    gathered = [torch.empty_like(b.flat) for _ in range(ws)]
    dist.all_gather(gathered, b.flat)
    identical = all(torch.equal(gathered[r], gathered[0]) for r in range(ws))
    max_diff = float((a.flat - b.flat).abs().max())

    def timed(fn, reps=200):
        for _ in range(20):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps * 1e3], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def nccl_step():
        dist.all_reduce(a.grad, op=dist.ReduceOp.SUM)
        a.apply(1.0 / ws)

    us_nccl = timed(nccl_step)
    us_p2p = timed(b.apply_p2p)
    torch.cuda.synchronize()
    timed_out = timed_out or int(b.p2p_error.item()) != 0
    b.disable_p2p()
    return {"steps": steps, "p2p_max_abs_diff": max_diff, "ranks_identical": bool(identical), "timed_out": bool(timed_out),
            "us_nccl_allreduce_plus_adam": round(us_nccl, 2), "us_p2p_fused": round(us_p2p, 2)}
