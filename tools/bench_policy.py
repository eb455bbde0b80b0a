#!/usr/bin/env python
"""bb_policy_forward alone (8192 rows, the rollout's shape: obs[:, idx] view of the engine's Observations column):
microseconds per call for the warp-level kernel (mma.sync 3xTF32) and the tcgen05 / TMEM kernel (BB_POLICY_UMMA_FORCE)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from madrona_basketball_b200.rollout import PolicyWeights, policy_forward
from ref_agent import RefAgent, randomize

dev = torch.device("cuda", 0)
agent = randomize(RefAgent(), seed=1, scale_heads=2.0).to(dev).eval()
w = PolicyWeights.from_module(agent)
out = {}
for n in (8192, 65536):
    obs_all = torch.randn(n, 2, 128, device=dev)
    obs = obs_all[:, 1]
    for flavour in ("mma_sync", "tcgen05"):
        os.environ["BB_POLICY_UMMA_FORCE"] = "1" if flavour == "tcgen05" else "0"
        for _ in range(20):
            policy_forward(w, obs, stochastic=True, seed=3, draw=0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(200):
            policy_forward(w, obs, stochastic=True, seed=3, draw=i)
        e1.record()
        torch.cuda.synchronize()
        out[f"{flavour}@{n}"] = round(e0.elapsed_time(e1) / 200 * 1e3, 2)
print(json.dumps({"us_per_policy_forward": out}))
