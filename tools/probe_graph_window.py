#!/usr/bin/env python
"""Where does a short timed window lose time?  bench.py's `value` (one CUDA-graph replay of K cold-row steps behind a spin
kernel) at several K on one box: the fit  window = fixed + K x per_step  separates the replay's start-up cost from the
step time.  Also the same window with the replay split in two graphs of K / 2."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from madrona_basketball_b200.shard import ShardedSimulator

n = 8192
dev = torch.device("cuda", 0)
R = 22
shards = [ShardedSimulator(n * R, rank=i, world_size=R, gpu_id=0) for i in range(R)]
acts = torch.empty((8, n, 2, 6), dtype=torch.int32, device=dev)
for s in range(8):
    shards[s % R].fill_actions(1234, s, out=acts[s])
calls = [[sh.sim.bind_step_with_actions(acts[b]) for b in range(8)] for sh in shards]
for i in range(20000):
    calls[i % R][i % 8]()
    if i % 64 == 0:
        torch.cuda.synchronize()
torch.cuda.synchronize()
side = torch.cuda.Stream(dev)


def capture(first, K):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        for s in range(K):
            shards[(first + s) % R].sim.bind_step_with_actions(acts[s % 8])()
    return g


def window(graphs, spacer_us, reps=7):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    out = []
    for _ in range(reps):
        for i in range(2 * R):
            calls[i % R][i % 8]()
        torch.cuda.synchronize()
        if spacer_us:
            torch.cuda._sleep(int(spacer_us * 1e-6 * 1.9e9))
        e0.record()
        for g in graphs:
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        out.append(e0.elapsed_time(e1) * 1e3)
    out.sort()
    return out[len(out) // 2], out[0]


res = {}
for K in (10, 20, 40, 100, 200):
    g = capture(0, K)
    g.replay(); torch.cuda.synchronize()
    for sp in (300, 0):
        med, mn = window([g], sp)
        res[f"K={K} spacer={sp}"] = {"window_us": round(med, 2), "min_us": round(mn, 2), "us_per_step": round(med / K, 3)}
ga, gb = capture(0, 10), capture(10, 10)
ga.replay(); gb.replay(); torch.cuda.synchronize()
med, mn = window([ga, gb], 300)
res["K=10+10 (two graphs) spacer=300"] = {"window_us": round(med, 2), "min_us": round(mn, 2), "us_per_step": round(med / 20, 3)}
# ordinary launches, same window
for K in (20, 200):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(7):
        torch.cuda.synchronize()
        torch.cuda._sleep(int(300e-6 * 1.9e9))
        e0.record()
        for s in range(K):
            calls[s % R][s % 8]()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    res[f"stream K={K} spacer=300"] = {"window_us": round(ts[3], 2), "us_per_step": round(ts[3] / K, 3)}
print(json.dumps(res, indent=1))
