#!/usr/bin/env python
"""End-to-end PPO on the B200 engine (BASELINE.json configs[3]: scripts/ppo.py at 8192 envs per GPU, 32-step
rollouts, 4 epochs x 4 minibatches), one process per GPU:

    python tools/train_ppo.py --iters 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/train_ppo.py --iters 20

Rollout and GAE run in libbball.so (bb_rollout, bb_gae: no host sync per step); the clipped-PPO update is
bb_ppo_grad + bb_ppo_adam (--update fused, default) or torch autograd as in the reference (--update torch), with
one flat NCCL all-reduce of the gradient per optimiser step; normaliser updates and advantage normalisation stay
in torch (ppo_dist.py).  Prints one JSON line on rank 0: env-steps/s (SPS) over the timed iterations.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs-per-gpu", type=int, default=8192)
    ap.add_argument("--rollout-steps", type=int, default=32)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--trainee-idx", type=int, default=1)   # ppo.py:34
    ap.add_argument("--seed", type=int, default=321)        # ppo.py:26
    ap.add_argument("--save", default=None, help="write the trained state_dict here (reference checkpoint format)")
    ap.add_argument("--log-every", type=int, default=0,
                    help="every K iterations read the episode statistics (one host sync) and record the window's mean return")
    ap.add_argument("--update", default="fused", choices=["fused", "torch"],
                    help="fused: bb_ppo_grad + bb_ppo_adam (3 kernels per minibatch step); torch: autograd + torch.optim.Adam as in the reference")
    ap.add_argument("--allreduce", default="p2p", choices=["p2p", "nccl"],
                    help="multi-GPU fused update: p2p = gradient all-reduce inside the optimiser kernel over NVLink peer memory "
                         "(bb_ppo_adam_p2p); nccl = one NCCL all-reduce per minibatch step + bb_ppo_adam")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    from madrona_basketball_b200 import ppo_dist
    from madrona_basketball_b200.policy_torch import Agent
    from madrona_basketball_b200.rollout import FusedRollout, PolicyWeights, RolloutBuffers
    from madrona_basketball_b200.shard import ShardedSimulator, env_rank, reduce_episode_stats

    rank, local, ws = env_rank()
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if ws > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(args.seed)
    n, T = args.envs_per_gpu, args.rollout_steps
    shard = ShardedSimulator(n * ws, rank=rank, world_size=ws, gpu_id=local)
    agent = Agent().to(dev)
    ppo_dist.broadcast_module_(agent)
    cfg = ppo_dist.PPOConfig()
    if args.update == "fused":   # re-binds the parameters as views of one flat vector: before anything takes their pointers
        from madrona_basketball_b200.ppo_fused import FusedPPO
        fused = FusedPPO(agent, lr=cfg.learning_rate, eps=1e-8, clip_coef=cfg.clip_coef, ent_coef=cfg.ent_coef,
                         vf_coef=cfg.vf_coef, max_grad_norm=cfg.max_grad_norm, seed=args.seed + 7919 * rank)
        if ws > 1 and args.allreduce == "p2p":
            try:
                fused.enable_p2p()
            except RuntimeError as e:   # raised on every rank together (CUDA IPC / peer access unavailable): NCCL all-reduce instead
                args.allreduce = "nccl"
                if rank == 0:
                    print(f"train_ppo: {e}; using --allreduce nccl", file=sys.stderr)
    else:
        opt = torch.optim.Adam(agent.parameters(), lr=cfg.learning_rate, eps=1e-8)
    weights = fused.weights if args.update == "fused" else PolicyWeights.from_module(agent)
    buf = RolloutBuffers(T, n, dev)
    roll = FusedRollout(shard.sim, weights, args.trainee_idx, seed=args.seed, world_offset=shard.world_offset)
    gen = torch.Generator(device=dev).manual_seed(args.seed + rank)

    ev = lambda: torch.cuda.Event(enable_timing=True)
    t_roll = t_adv = t_upd = 0.0
    losses = {}
    curve, curve_prev = [], [0.0, 0.0, 0.0]
    for it in range(args.warmup + args.iters):
        if it == args.warmup:
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
        if args.update == "fused":
            fused.finish_advantages(buf)                                # ppo.py:167-176, incl. the float32 normaliser vectors
        else:
            ppo_dist.finish_advantages(agent, buf)
            weights.refresh()                                           # the observation normaliser moved
        e[2].record()
        agent.train()
        if args.update == "fused":                                      # ppo.py:180-216
            losses = fused.update(buf, cfg.num_minibatches, cfg.update_epochs, generator=gen, refresh_normalizer=False)
        else:
            losses = ppo_dist.update_policy(agent, opt, buf, cfg, generator=gen)
        e[3].record()
        torch.cuda.synchronize()
        t_roll += e[0].elapsed_time(e[1]); t_adv += e[1].elapsed_time(e[2]); t_upd += e[2].elapsed_time(e[3])
        if args.log_every and (it + 1) % args.log_every == 0:   # learning curve: episodes finished in this window (this rank)
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
    if args.save and rank == 0:
        torch.save(agent.state_dict(), args.save)
    if rank == 0:
        steps = n * ws * T * args.iters
        print(json.dumps({
            "what": "PPO iterations (rollout + GAE + 4x4 minibatch update), ppo.py defaults, fused rollout path, update=" + args.update
                    + (", allreduce=" + args.allreduce if (ws > 1 and args.update == "fused") else ""),
            "n_gpus": ws, "envs_per_gpu": n, "rollout_steps": T, "iters": args.iters,
            "env_steps_per_s": steps / float(t.item()), "s_per_iter": float(t.item()) / args.iters,
            "ms_per_iter_rank0": {"rollout": t_roll / args.iters, "advantages_normalisers": t_adv / args.iters, "update": t_upd / args.iters},
            "episodes_finished": stats["episodes"], "mean_return": stats["return_sum"] / max(stats["episodes"], 1.0),
            "mean_episode_length": stats["length_sum"] / max(stats["episodes"], 1.0),
            "last_losses": {k: float(v) for k, v in losses.items()},
            "learning_curve_rank0": curve,
        }))
    if ws > 1:
        if args.update == "fused":
            fused.disable_p2p()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
