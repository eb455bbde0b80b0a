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
    ap.add_argument("--times", action="store_true", help="print the event time of every step (to pick an event-heavy step for ncu -s)")
    ap.add_argument("--loop", type=int, default=0, help="after the steps: one bb_step_n launch of this many steps (bb_step_loop_kernel)")
    args = ap.parse_args()
    import torch
    from madrona_basketball_b200.shard import ShardedSimulator

    sh = ShardedSimulator(args.worlds, rank=0, world_size=1, gpu_id=0)
    act = torch.empty((args.worlds, 2, 6), dtype=torch.int32, device="cuda")
    ev = []
    for s in range(args.steps):
        sh.fill_actions(1234, s, out=act)
        if args.times:
            ev.append((torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)))
            ev[-1][0].record()
        sh.sim.step_with_actions(act)
        if args.times:
            ev[-1][1].record()
    if args.loop:
        sh.sim.step_n(args.loop)
    torch.cuda.synchronize()
    if args.times:
        us = [a.elapsed_time(b) * 1e3 for a, b in ev]
        order = sorted(range(len(us)), key=lambda i: -us[i])
        print("slowest steps (index: us):", ", ".join(f"{i}: {us[i]:.0f}" for i in order[:12]))
        print("all:", " ".join(f"{u:.0f}" for u in us))
    print("ok", args.worlds, args.steps, sh.sim.launch_count())


if __name__ == "__main__":
    main()
