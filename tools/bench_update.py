#!/usr/bin/env python
"""PPO minibatch step at the ppo.py defaults (65 536 rows = 8192 envs x 32 steps / 4 minibatches), one B200:
torch autograd + clip_grad_norm_ + Adam as the reference runs it (ppo.py:189-214) vs bb_ppo_grad + bb_ppo_adam.
Prints one JSON line."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from madrona_basketball_b200 import ppo_dist
from madrona_basketball_b200.policy_torch import Agent, ACTION_BUCKETS
from madrona_basketball_b200.ppo_fused import FusedPPO

rows_total, nmb, reps = 8192 * 32, 4, int(sys.argv[1]) if len(sys.argv) > 1 else 20
dev = torch.device("cuda", 0)
torch.manual_seed(0)
g = torch.Generator(device=dev).manual_seed(0)
obs = torch.randn(rows_total, 128, device=dev, generator=g)
act = torch.stack([torch.randint(0, nb, (rows_total,), generator=g, device=dev) for nb in ACTION_BUCKETS], dim=1).float()
adv = torch.randn(rows_total, device=dev, generator=g); ret = torch.randn(rows_total, device=dev, generator=g)
agent_t, agent_f = Agent().to(dev), Agent().to(dev)
agent_f.load_state_dict(agent_t.state_dict())
with torch.no_grad():
    lp, _, v = agent_t.get_stats(obs, act)
lp_old = (lp + 0.1 * torch.randn_like(lp)).contiguous(); v_old = (v + 0.1 * torch.randn_like(v)).contiguous()
cfg = ppo_dist.PPOConfig()
opt = torch.optim.Adam(agent_t.parameters(), lr=cfg.learning_rate, eps=1e-8)
fused = FusedPPO(agent_f)
mb = rows_total // nmb

def torch_step(idx):
    lp, ent, v = agent_t.get_stats(obs[idx], act[idx])
    ratio = torch.exp(lp - lp_old[idx]); a = adv[idx]
    pg = torch.max(-a * ratio, -a * torch.clamp(ratio, 0.8, 1.2)).mean()
    vo, r = v_old[idx], ret[idx]
    vc = vo + (v - vo).clamp(-0.2, 0.2)
    vl = 0.5 * torch.max((v - r) ** 2, (vc - r) ** 2).mean()
    loss = pg + vl - 0.01 * ent.mean()
    opt.zero_grad(set_to_none=False); loss.backward()
    torch.nn.utils.clip_grad_norm_(list(agent_t.parameters()), 1.0); opt.step()

def fused_step(idx):
    fused.gradient(obs, act, lp_old, v_old, adv, ret, idx); fused.apply()

def timeit(fn):
    perm = torch.randperm(rows_total, device=dev, generator=g)
    for s in range(0, rows_total, mb): fn(perm[s:s + mb])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(reps):
        perm = torch.randperm(rows_total, device=dev, generator=g)
        for s in range(0, rows_total, mb): fn(perm[s:s + mb])
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * nmb), (time.perf_counter() - t0) / (reps * nmb) * 1e3

# kernel-only timing of the gradient kernel pair and the Adam kernel
idx = torch.randperm(rows_total, device=dev, generator=g)[:mb].contiguous()
for _ in range(3): fused.gradient(obs, act, lp_old, v_old, adv, ret, idx)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
torch.cuda.synchronize(); ev[0].record()
for _ in range(50): fused.gradient(obs, act, lp_old, v_old, adv, ret, idx)
ev[1].record()
for _ in range(50): fused.apply()
ev[2].record(); torch.cuda.synchronize()
t_dev, t_wall = timeit(torch_step)
f_dev, f_wall = timeit(fused_step)
print(json.dumps({"what": "PPO minibatch step, 65536 rows, Agent(128->32->32->19+1), one B200", "rows": mb,
                  "torch": {"ms_per_step_device": t_dev, "ms_per_step_wall": t_wall},
                  "fused": {"ms_per_step_device": f_dev, "ms_per_step_wall": f_wall,
                            "grad_kernels_us": ev[0].elapsed_time(ev[1]) / 50 * 1e3, "adam_kernel_us": ev[1].elapsed_time(ev[2]) / 50 * 1e3},
                  "speedup_device": t_dev / f_dev, "flop_per_row": 25800,
                  "grad_tflops": mb * 25800 / (ev[0].elapsed_time(ev[1]) / 50 * 1e-3) / 1e12}))
