#!/usr/bin/env python
"""Rollout throughput at BASELINE.json configs[3] scale (8192 envs x 32 steps, scripts/ppo.py defaults):

  (a) torch loop : what the reference's unmodified ppo.py `rollout()` does per step on top of the drop-in
                   simulator — Agent.forward in torch (~15 kernels), actions[:, idx] = a, sim.step(), three
                   .clone()s (env.py:167-169), PPOStats.update with its `dones.any()` host sync, buffer writes;
  (b) fused      : bb_rollout — one policy kernel + one step kernel per step, no host synchronisation.

Prints one JSON line.  Not the contract bench (bench.py); numbers land in profiles/ and DESIGN.md section 8.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=32)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--agent-idx", type=int, default=1)
    ap.add_argument("--frozen", action="store_true", help="self-play: a frozen opponent policy drives the other agent (fused path only)")
    ap.add_argument("--no-torch", action="store_true", help="skip the torch loop")
    args = ap.parse_args()
    import torch

    from madrona_basketball_b200 import SimpleGridworldSimulator, madrona
    from madrona_basketball_b200.rollout import FusedRollout, PolicyWeights, RolloutBuffers
    from ref_agent import RefAgent, randomize

    n, T, idx = args.envs, args.steps, args.agent_idx
    dev = torch.device("cuda", 0)
    mk = lambda: SimpleGridworldSimulator(discrete_x=32, discrete_y=17, start_x=31.515 / 2.0, start_y=16.764 / 2.0,
                                          max_episode_length=39600, exec_mode=madrona.ExecMode.CUDA, num_worlds=n, gpu_id=0)
    agent = randomize(RefAgent(), seed=1, scale_heads=2.0).to(dev).eval()

    # ---------------------------------------------------------------- (a) torch loop, as ppo.py / env.py do it
    sim = mk()
    obs_t, act_t = sim.observations_tensor().to_torch(), sim.action_tensor().to_torch()
    rew_t, done_t, reset_t = sim.reward_tensor().to_torch(), sim.done_tensor().to_torch(), sim.reset_tensor().to_torch()
    z = lambda *s: torch.zeros(s, device=dev)
    b_obs, b_act, b_val, b_lp, b_rew, b_nd = z(T, n, 128), z(T, n, 6), z(T, n), z(T, n), z(T, n), z(T, n)
    cur, length = z(n), z(n)
    from torch.distributions.categorical import Categorical

    def torch_forward(obs):
        logits, value = agent.logits_value(obs)
        acts, lps, off = [], [], 0
        for nb in agent.buckets:
            d = Categorical(logits=logits[:, off:off + nb], validate_args=False)
            a = d.sample()
            acts.append(a); lps.append(d.log_prob(a)); off += nb
        return torch.stack(acts, 1), torch.stack(lps, 1).sum(-1), value

    def env_step(a):
        act_t[:, idx] = a
        sim.step()
        return obs_t[:, idx].detach().clone(), rew_t[:, idx].detach().clone(), done_t[:, idx].detach().clone()

    @torch.no_grad()
    def torch_rollout():
        nonlocal cur, length
        reset_t.fill_(1)
        obs, _, _ = env_step(torch.zeros((n, 6), dtype=torch.int32, device=dev))
        reset_t.fill_(0)
        for t in range(T):
            a, lp, v = torch_forward(obs)
            obs_, rew, done = env_step(a.to(torch.int32))
            cur += rew; length += 1
            if done.any():                                   # ppo_stats.py:26: one host sync per step
                ind = done.nonzero(as_tuple=False).squeeze(-1)
                _ = cur[ind].mean(), length[ind].mean()
            cur = cur * (1 - done); length = length * (1 - done)
            b_obs[t], b_act[t], b_val[t], b_lp[t], b_nd[t], b_rew[t] = obs, a, v, lp, 1.0 - done, rew
            obs = obs_
        return agent.logits_value(obs)[1]

    t_torch = float("nan")
    if not args.no_torch:
        for _ in range(3):
            torch_rollout()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.iters):
            torch_rollout()
        torch.cuda.synchronize()
        t_torch = (time.perf_counter() - t0) / args.iters

    # ---------------------------------------------------------------- (b) fused
    sim2 = mk()
    buf = RolloutBuffers(T, n, dev)
    frozen_w = PolicyWeights.from_module(randomize(RefAgent(), seed=2, scale_heads=2.0).to(dev).eval()) if args.frozen else None
    fr = FusedRollout(sim2, PolicyWeights.from_module(agent), idx, frozen=frozen_w, seed=321)
    for _ in range(3):
        fr.collect(buf)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.iters):
        fr.collect(buf)
        fr.gae(buf, 0.998, 0.95, agent.value_norm.mean, agent.value_norm.var)
    e1.record()
    torch.cuda.synchronize()
    t_fused = (time.perf_counter() - t0) / args.iters
    t_fused_dev = e0.elapsed_time(e1) * 1e-3 / args.iters

    # policy kernel alone
    from madrona_basketball_b200.rollout import policy_forward
    w = PolicyWeights.from_module(agent)
    view = sim2.observations_tensor().to_torch()[:, idx]
    for _ in range(5):
        policy_forward(w, view)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(50)]
    for a, b in ev:
        a.record(); policy_forward(w, view); b.record()
    torch.cuda.synchronize()
    pol_us = sorted(a.elapsed_time(b) * 1e3 for a, b in ev)[len(ev) // 2]

    steps = n * (T + 1)   # env.reset() is one more simulator step per rollout
    print(json.dumps({
        "what": "rollout of ppo.py defaults on one B200: env.reset() + T x (policy forward + sample, sim step, bookkeeping)",
        "envs": n, "rollout_steps": T, "iters": args.iters, "agent_idx": idx, "frozen_opponent": bool(args.frozen),
        "torch_loop": {"s_per_rollout": t_torch, "env_steps_per_s": steps / t_torch,
                       "what": "torch Agent.forward + Categorical sampling, env.py-style step with 3 clones, PPOStats host sync per step"},
        "fused": {"s_per_rollout": t_fused, "s_per_rollout_device": t_fused_dev, "env_steps_per_s": steps / t_fused,
                  "what": "bb_rollout + bb_gae: 2 kernels per step, no host sync; wall clock incl. launch"},
        "speedup": t_torch / t_fused,
        "policy_kernel_us_median": pol_us,
        "episode_stats": buf.episode_stats(),
    }))


if __name__ == "__main__":
    main()
