#!/usr/bin/env python
"""Per-step kernel time over a long run at N worlds (episode-phase dependence), with and without the NVML sampler."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from madrona_basketball_b200.shard import ShardedSimulator
import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 400
use_sampler = len(sys.argv) > 3 and sys.argv[3] == "sampler"
sh = ShardedSimulator(n, rank=0, world_size=1, gpu_id=0)
acts = torch.empty((8, n, 2, 6), dtype=torch.int32, device="cuda")
sampler = bench.ClockSampler(0) if use_sampler else None
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
torch.cuda.synchronize()
if sampler: sampler.__enter__()
for s in range(steps):
    sh.fill_actions(1234, s, out=acts[s % 8])
    ev[s][0].record()
    sh.sim.step_with_actions(acts[s % 8])
    ev[s][1].record()
torch.cuda.synchronize()
if sampler: sampler.__exit__()
us = [a.elapsed_time(b) * 1e3 for a, b in ev]
for i in range(0, steps, 20):
    chunk = us[i:i + 20]
    print(f"steps {i:4d}-{i + len(chunk) - 1:4d}: avg {sum(chunk) / len(chunk):7.1f} us  min {min(chunk):7.1f}  max {max(chunk):7.1f}")
if sampler: print(sampler.report())
rc = sh.sim.tensor("rng_count").to_torch()
print("rng_count mean", rc.float().mean().item())
