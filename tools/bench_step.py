#!/usr/bin/env python
"""Quick kernel timing for tuning (not the contract bench): per-launch microseconds of the fused
step at several world counts, L2 flushed between launches and L2 warm."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, nargs="+", default=[8192, 65536, 1 << 20])
    ap.add_argument("--reps", type=int, default=50)
    args = ap.parse_args()
    import torch
    from madrona_basketball_b200.shard import ShardedSimulator

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for n in args.worlds:
        sh = ShardedSimulator(n, rank=0, world_size=1, gpu_id=0)
        acts = torch.empty((8, n, 2, 6), dtype=torch.int32, device="cuda")
        for s in range(8):
            sh.fill_actions(1234, s, out=acts[s])
        for s in range(5):
            sh.sim.step_with_actions(acts[s % 8])
        out = {}
        for mode in ("flushed", "warm"):
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.reps)]
            torch.cuda.synchronize()
            for s in range(args.reps):
                if mode == "flushed":
                    flush.zero_()
                ev[s][0].record()
                sh.sim.step_with_actions(acts[s % 8])
                ev[s][1].record()
            torch.cuda.synchronize()
            ms = sorted(a.elapsed_time(b) for a, b in ev)
            out[mode] = (sum(ms) / len(ms) * 1e3, ms[len(ms) // 2] * 1e3, ms[0] * 1e3)
        gbs = n * 1872 / (out["flushed"][0] * 1e-6) / 1e9
        print(f"worlds={n:8d} flushed avg/med/min us = {out['flushed'][0]:.2f}/{out['flushed'][1]:.2f}/{out['flushed'][2]:.2f}  "
              f"warm avg/med/min us = {out['warm'][0]:.2f}/{out['warm'][1]:.2f}/{out['warm'][2]:.2f}  "
              f"flushed {n / out['flushed'][0]:.1f} Mws/s  {gbs:.0f} GB/s algorithmic")
        del sh, acts


if __name__ == "__main__":
    main()
