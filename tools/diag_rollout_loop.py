#!/usr/bin/env python
"""Diagnostic: bb_rollout as one launch (BB_ROLLOUT_LOOP=1) against the kernel chain (=0): where do the buffers differ?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from helpers import make_engine
from madrona_basketball_b200.rollout import FusedRollout, PolicyWeights, RolloutBuffers
from ref_agent import RefAgent, randomize

n, T = int(sys.argv[1]) if len(sys.argv) > 1 else 700, 4
agent = randomize(RefAgent(), seed=1, scale_heads=2.0).cuda()
w = PolicyWeights.from_module(agent)
res = {}
for form in ("0", "1"):
    os.environ["BB_ROLLOUT_LOOP"] = form
    sim = make_engine(n)
    for s in range(25):
        sim.fill_actions(1234, s); sim.step()
    buf = RolloutBuffers(T, n, torch.device("cuda"))
    FusedRollout(sim, w, 1, seed=321).collect(buf)
    torch.cuda.synchronize()
    res[form] = {k: getattr(buf, k).clone() for k in ("obs", "actions", "values", "log_probs", "rewards", "not_dones", "next_value")}
for k in res["0"]:
    a, b = res["0"][k], res["1"][k]
    ne = (a.view(torch.int32) != b.view(torch.int32))
    print(k, "mismatches", int(ne.sum()), "of", ne.numel(), "max abs diff", float((a - b).abs().max()))
    if k in ("values", "log_probs") and ne.any():
        for t in range(T):
            rows = ne[t].nonzero().flatten().tolist()
            print("  t", t, "rows differing:", len(rows), "first", rows[:24], "rows%16 histogram", torch.bincount(torch.tensor(rows) % 16, minlength=16).tolist() if rows else None)
