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
    ap.add_argument("--allreduce", default="nccl", choices=["nccl", "p2p"],
                    help="multi-GPU fused update: nccl (default, as north_star names it) = one NCCL all-reduce of the flat gradient per "
                         "minibatch step + bb_ppo_adam; p2p = the all-reduce inside the optimiser kernel over NVLink peer memory "
                         "(bb_ppo_adam_p2p, opt-in)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    from madrona_basketball_b200.ppo_train import run_iterations
    from madrona_basketball_b200.shard import env_rank

    rank, local, ws = env_rank()
    torch.cuda.set_device(torch.device("cuda", local))
    if ws > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = run_iterations(envs_per_gpu=args.envs_per_gpu, rollout_steps=args.rollout_steps, iters=args.iters, warmup=args.warmup,
                         trainee_idx=args.trainee_idx, seed=args.seed, update=args.update, allreduce=args.allreduce,
                         log_every=args.log_every, save=args.save)
    if rank == 0:
        print(json.dumps(out))
    if ws > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
