#!/usr/bin/env python
"""Minimal driver for ncu captures: N worlds, a few steps of the fused step kernel with the
synthetic action stream (same inputs as bench.py).  Not a benchmark: prints nothing timed."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=1 << 20)
    ap.add_argument("--steps", type=int, default=8)
    args = ap.parse_args()
    import torch
    from madrona_basketball_b200.shard import ShardedSimulator

    sh = ShardedSimulator(args.worlds, rank=0, world_size=1, gpu_id=0)
    act = torch.empty((args.worlds, 2, 6), dtype=torch.int32, device="cuda")
    for s in range(args.steps):
        sh.fill_actions(1234, s, out=act)
        sh.sim.step_with_actions(act)
    torch.cuda.synchronize()
    print("ok", args.worlds, args.steps, sh.sim.launch_count())


if __name__ == "__main__":
    main()
