#!/usr/bin/env python
"""Where does the 8192-world step time go?  Event-timed floors on one B200 (tuning aid, prints JSON):

  null        a 1-element torch kernel after an L2 flush: launch + event floor
  step_flush  bb_step_with_actions after an L2 flush (bench.py's `value` condition)
  step_code   same, but a 64-world engine is stepped between flush and timing: the kernel's CODE and constants are back in
              L2 / the instruction caches, the 8192 worlds' data still cold -> cost of cold instruction fetch
  step_warm   no flush at all
  sizes       step_flush at other world counts (latency floor vs bandwidth)
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madrona_basketball_b200 import SimpleGridworldSimulator, madrona  # noqa: E402

ARGS = dict(discrete_x=32, discrete_y=17, start_x=31.515 / 2.0, start_y=16.764 / 2.0, max_episode_length=39600,
            exec_mode=madrona.ExecMode.CUDA, gpu_id=0)
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
one = torch.zeros(1, device=dev)
REPS = int(os.environ.get("REPS", "300"))


def timed(fn, pre=None, reps=REPS):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for i in range(reps):
        if pre is not None:
            pre(i)
        ev[i][0].record()
        fn(i)
        ev[i][1].record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)
    return {"mean_us": 1e3 * sum(ms) / len(ms), "median_us": 1e3 * ms[len(ms) // 2], "p10_us": 1e3 * ms[len(ms) // 10], "p90_us": 1e3 * ms[int(len(ms) * 0.9)]}


def engine(n):
    sim = SimpleGridworldSimulator(num_worlds=n, **ARGS)
    acts = torch.empty((8, n, 2, 6), dtype=torch.int32, device=dev)
    for s in range(8):
        sim.fill_actions(1234, s, out=acts[s])
    for s in range(64):
        sim.step_with_actions(acts[s % 8])
    torch.cuda.synchronize()
    return sim, acts


out = {}
sim, acts = engine(8192)
small, sacts = engine(64)
# pre-warm clocks
for i in range(2000):
    sim.step_with_actions(acts[i % 8])
torch.cuda.synchronize()
out["null"] = timed(lambda i: one.add_(1.0), pre=lambda i: flush.zero_())
out["step_flush"] = timed(lambda i: sim.step_with_actions(acts[i % 8]), pre=lambda i: flush.zero_())


def pre_code(i):
    flush.zero_()
    small.step_with_actions(sacts[i % 8])


out["step_code_warm_data_cold"] = timed(lambda i: sim.step_with_actions(acts[i % 8]), pre=pre_code)
out["step_warm"] = timed(lambda i: sim.step_with_actions(acts[i % 8]))
out["step_n_graph_us_per_step"] = None
sim.step_n(200)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); sim.step_n(200); e1.record(); torch.cuda.synchronize()
out["step_n_graph_us_per_step"] = e0.elapsed_time(e1) * 1e3 / 200
del sim, small
sizes = {}
for n in [int(x) for x in os.environ.get("SIZES", "1024,4096,16384,32768,65536,131072").split(",")]:
    s2, a2 = engine(n)
    sizes[n] = timed(lambda i: s2.step_with_actions(a2[i % 8]), pre=lambda i: flush.zero_(), reps=100)
    sizes[n]["roofline_frac"] = n * 1872 / (sizes[n]["mean_us"] * 1e-6) / 6531.3e9
    del s2, a2
out["sizes_flushed"] = sizes
print(json.dumps(out))
