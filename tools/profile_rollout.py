#!/usr/bin/env python
"""Minimal driver for ncu captures of bb_rollout (one launch of bb_rollout_loop_kernel per rollout by default,
BB_ROLLOUT_LOOP=0: the kernel chain).  Not a benchmark: prints nothing timed."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=32)
    ap.add_argument("--rollouts", type=int, default=3)
    args = ap.parse_args()
    import torch

    from helpers import make_engine
    from madrona_basketball_b200.rollout import FusedRollout, PolicyWeights, RolloutBuffers
    from ref_agent import RefAgent, randomize

    agent = randomize(RefAgent(), seed=1, scale_heads=2.0).cuda()
    sim = make_engine(args.envs)
    buf = RolloutBuffers(args.steps, args.envs, torch.device("cuda"))
    fr = FusedRollout(sim, PolicyWeights.from_module(agent), 1, seed=321)
    for _ in range(args.rollouts):
        fr.collect(buf)
    torch.cuda.synchronize()
    print("ok", args.envs, args.steps, sim.launch_count())


if __name__ == "__main__":
    main()
