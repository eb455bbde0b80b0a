#!/usr/bin/env python
"""A/B of two builds of libbball.so on the SAME box (box-to-box spread is ~3 %, more than most kernel changes):
    python tools/ab_step.py libA.so[:ENV=VAL,ENV2=VAL2] libB.so[:...] [worlds ...]
For each library (alternating, 3 rounds): K = 200 steps over rotating cold shards with random actions (bench.py's `value`:
one CUDA-graph replay; and the same as queued stream launches), and bb_step_n graph replay; prints microseconds per step."""
import json, os, subprocess, sys

CHILD = r'''
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(sys.argv[0]))) if False else "/root/repo")
import torch
import bench
from madrona_basketball_b200.shard import ShardedSimulator
out = {}
for n in [int(a) for a in sys.argv[1:]]:
    r = bench.rotating_rate(n, n, 1, 0, 0, 200, 10, 0.3)
    sh = ShardedSimulator(n, rank=0, world_size=1, gpu_id=0)
    sh.sim.step_n(50); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); sh.sim.step_n(200); e1.record(); torch.cuda.synchronize()
    out[n] = {"cold_graph_us": round(r["us_per_step"], 3), "cold_stream_us": round(r["stream_launches"]["us_per_step"], 3),
              "step_n_us": round(e0.elapsed_time(e1) / 200 * 1e3, 3)}
print(json.dumps(out))
'''

def main():
    libs = sys.argv[1:3]
    worlds = sys.argv[3:] or ["8192", "65536", "1048576"]
    for rnd in range(3):
        for spec in libs:
            lib, _, extra = spec.partition(":")
            env = dict(os.environ, BB_LIB_PATH=os.path.abspath(lib))
            env.update(kv.split("=", 1) for kv in extra.split(",") if kv)
            res = subprocess.run([sys.executable, "-c", CHILD] + worlds, env=env, capture_output=True, text=True)
            line = res.stdout.strip().splitlines()[-1] if res.stdout.strip() else res.stderr[-400:]
            print(os.path.basename(spec), line, flush=True)

if __name__ == "__main__":
    main()
