"""GPU parity of the fused PPO update (include/bb_ppo.h) against plain torch fp32 autograd of the reference's
update_policy (scripts/ppo.py:180-216) on the same minibatch.  Floating-point kernels: the gradient is a sum over
65 536 rows in a different order than torch's, so the tolerance is 1e-4 relative to the largest gradient entry of
each parameter tensor (stated per assert); Adam steps agree to 1e-6 relative."""
import numpy as np
import pytest

from helpers import make_engine
from ref_agent import ACTION_BUCKETS

pytestmark = pytest.mark.gpu


def _agent(seed=0, scale=1.0):
    import torch

    from madrona_basketball_b200.policy_torch import Agent

    torch.manual_seed(seed)
    a = Agent().cuda()
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():   # non-trivial heads and normaliser (the reference's init makes the heads ~0.01)
        for p in a.parameters():
            p.copy_((torch.randn(p.shape, generator=g) * (0.3 if p.dim() > 1 else 0.1) * scale).cuda())
        a.backbone[1].weight.add_(1.0); a.backbone[4].weight.add_(1.0)
        a.obs_norm.mean.copy_((torch.randn(128, generator=g, dtype=torch.float64) * 0.5).cuda())
        a.obs_norm.var.copy_((torch.rand(128, generator=g, dtype=torch.float64) * 4.0 + 0.05).cuda())
    return a


def _batch(rows, seed=3):
    """Real observation rows from the engine + synthetic behaviour statistics."""
    import torch

    n = (rows + 1) // 2
    sim = make_engine(n)
    for s in range(30):
        sim.fill_actions(11, s)
        sim.step()
    torch.cuda.synchronize()
    obs = sim.observations_tensor().to_torch().reshape(-1, 128)[:rows].clone()
    g = torch.Generator(device="cuda").manual_seed(seed)
    act = torch.stack([torch.randint(0, nb, (rows,), generator=g, device="cuda") for nb in ACTION_BUCKETS], dim=1).float()
    adv = torch.randn(rows, generator=g, device="cuda")
    ret = torch.randn(rows, generator=g, device="cuda")
    return obs, act, adv, ret, g


def _torch_loss(agent, obs, act, lp_old, v_old, adv, ret, clip=0.2, ent_coef=0.01, vf_coef=1.0):
    import torch

    lp, ent, v = agent.get_stats(obs, act)
    ratio = torch.exp(lp - lp_old)
    pg = torch.max(-adv * ratio, -adv * torch.clamp(ratio, 1 - clip, 1 + clip))
    v_clip = v_old + (v - v_old).clamp(-clip, clip)
    vl = 0.5 * torch.max((v - ret) ** 2, (v_clip - ret) ** 2)
    loss = pg.mean() + vf_coef * vl.mean() - ent_coef * ent.mean()
    return loss, pg.sum(), vl.sum(), ent.sum(), ((ratio < 1 - clip) | (ratio > 1 + clip)).float().sum()


def _behaviour(agent, obs, act, g, spread):
    """lp_old / v_old around the current policy so that a good share of rows is clipped on either side."""
    import torch

    with torch.no_grad():
        lp, _, v = agent.get_stats(obs, act)
    rows = obs.shape[0]
    return (lp + spread * torch.randn(rows, generator=g, device="cuda")).contiguous(), \
           (v + spread * torch.randn(rows, generator=g, device="cuda")).contiguous()


@pytest.mark.parametrize("rows,use_idx", [(4096, False), (1000, True), (65536, True), (37, False)])
def test_gradient_matches_torch_autograd(rows, use_idx):
    import torch

    from madrona_basketball_b200.ppo_fused import PARAM_ORDER, FusedPPO

    agent = _agent(seed=rows)
    obs, act, adv, ret, g = _batch(rows + 50)
    fused = FusedPPO(agent)
    lp_old, v_old = _behaviour(agent, obs, act, g, 0.25)
    idx = torch.randperm(rows + 50, generator=g, device="cuda")[:rows].contiguous() if use_idx else None
    sel = idx if use_idx else torch.arange(rows, device="cuda")
    o = obs if use_idx else obs[:rows].contiguous()
    a = act if use_idx else act[:rows].contiguous()
    args = (lp_old, v_old, adv, ret) if use_idx else tuple(t[:rows].contiguous() for t in (lp_old, v_old, adv, ret))
    grad = fused.gradient(o, a, *args, indices=idx).clone()
    stats = fused.stats.clone()

    agent.zero_grad()
    loss, pg_sum, vl_sum, ent_sum, nclip = _torch_loss(agent, obs[sel], act[sel], lp_old[sel], v_old[sel], adv[sel], ret[sel])
    loss.backward()
    torch.cuda.synchronize()
    named = dict(agent.named_parameters())
    off = 0
    for k in PARAM_ORDER:
        ref = named[k].grad.reshape(-1)
        got = grad[off:off + ref.numel()]
        off += ref.numel()
        scale = ref.abs().max().item()
        assert scale > 0
        err = (got - ref).abs().max().item()
        assert err <= 1e-4 * scale + 1e-9, f"{k}: max err {err:.3e} vs scale {scale:.3e}"   # 1e-4 of the tensor's largest entry
    ref_stats = torch.stack([pg_sum, vl_sum, ent_sum, nclip]).detach()
    assert torch.allclose(stats[:3], ref_stats[:3], rtol=2e-5, atol=1e-3), (stats, ref_stats)
    assert abs(stats[3].item() - ref_stats[3].item()) <= max(2.0, 1e-3 * rows)   # rows within an ulp of the clip edge
    assert 0.05 * rows < ref_stats[3].item() < 0.95 * rows                          # both regimes are exercised


def test_gradient_is_bitwise_reproducible():
    import torch

    from madrona_basketball_b200.ppo_fused import FusedPPO

    agent = _agent(seed=5)
    obs, act, adv, ret, g = _batch(20000)
    fused = FusedPPO(agent)
    lp_old, v_old = _behaviour(agent, obs, act, g, 0.2)
    g1 = fused.gradient(obs, act, lp_old, v_old, adv, ret).clone()
    g2 = fused.gradient(obs, act, lp_old, v_old, adv, ret).clone()
    torch.cuda.synchronize()
    assert torch.equal(g1, g2)


def test_clip_and_adam_match_torch():
    import torch

    from madrona_basketball_b200.ppo_fused import PARAM_ORDER, FusedPPO

    agent = _agent(seed=9)
    ref = _agent(seed=9)
    fused = FusedPPO(agent, lr=3e-4, max_grad_norm=1.0)
    opt = torch.optim.Adam(ref.parameters(), lr=3e-4, eps=1e-8)
    gen = torch.Generator(device="cuda").manual_seed(1)
    rnamed = dict(ref.named_parameters())
    for step in range(5):
        gflat = torch.randn(5972, generator=gen, device="cuda") * (3.0 if step % 2 == 0 else 0.005)   # clipped / not clipped
        fused.grad.copy_(gflat)
        fused.apply(0.5)
        off = 0
        for k in PARAM_ORDER:
            p = rnamed[k]
            p.grad = (gflat[off:off + p.numel()] * 0.5).view_as(p).clone()
            off += p.numel()
        norm = torch.nn.utils.clip_grad_norm_(list(ref.parameters()), 1.0)
        opt.step()
        torch.cuda.synchronize()
        assert abs(fused.grad_norm.item() - norm.item()) <= 1e-5 * norm.item()
        named = dict(agent.named_parameters())
        for k in PARAM_ORDER:
            assert torch.allclose(named[k], rnamed[k], rtol=1e-6, atol=1e-7), k   # Adam arithmetic: 1e-6 relative
    assert int(fused.step.item()) == 5


def test_fused_update_tracks_torch_update_policy():
    """One epoch (4 minibatches) of ppo.py:180-216 with the same permutation: parameters agree to 1e-4 absolute
    (16 384-row gradients summed in a different order, then Adam's normalisation)."""
    import torch

    from madrona_basketball_b200.ppo_fused import PARAM_ORDER, FusedPPO

    rows, nmb = 16384, 4
    agent, ref = _agent(seed=21), _agent(seed=21)
    obs, act, adv, ret, g = _batch(rows)
    lp_old, v_old = _behaviour(agent, obs, act, g, 0.1)
    fused = FusedPPO(agent)
    opt = torch.optim.Adam(ref.parameters(), lr=3e-4, eps=1e-8)
    perm = torch.randperm(rows, generator=g, device="cuda")
    mb = rows // nmb
    for s in range(0, rows, mb):
        idx = perm[s:s + mb]
        fused.gradient(obs, act, lp_old, v_old, adv, ret, idx)
        fused.apply()
        loss = _torch_loss(ref, obs[idx], act[idx], lp_old[idx], v_old[idx], adv[idx], ret[idx])[0]
        opt.zero_grad()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(list(ref.parameters()), 1.0)
        opt.step()
    torch.cuda.synchronize()
    named, rnamed = dict(agent.named_parameters()), dict(ref.named_parameters())
    for k in PARAM_ORDER:
        assert torch.allclose(named[k], rnamed[k], rtol=0, atol=1e-4), (k, (named[k] - rnamed[k]).abs().max())
    # the module's state_dict still carries the reference's checkpoint keys, now as views of the flat vector
    assert set(PARAM_ORDER) <= set(agent.state_dict().keys())
    assert named["actor.weight"].data_ptr() == fused.flat.data_ptr() + 4 * 5312


def test_finish_advantages_matches_torch():
    """ppo.py:167-176 (normaliser updates, advantage / value normalisation): fused kernels vs the torch float64
    restatement in ppo_dist.finish_advantages.  Normaliser buffers (float64) rtol 1e-9; float32 outputs rtol 1e-5."""
    import copy
    import types

    import torch

    from madrona_basketball_b200 import ppo_dist
    from madrona_basketball_b200.ppo_fused import FusedPPO

    T, n = 8, 1111
    agent = _agent(seed=33)
    with torch.no_grad():
        agent.value_norm.mean.fill_(1.7); agent.value_norm.var.fill_(9.0); agent.value_norm.count.fill_(5000.0)
        agent.obs_norm.count.fill_(12345.0)
    ref = copy.deepcopy(agent)
    obs, _, _, _, g = _batch(T * n)
    mk = lambda scale, shift: (torch.randn(T, n, generator=g, device="cuda") * scale + shift).contiguous()
    buf = types.SimpleNamespace(horizon=T, n_envs=n, obs=obs.view(T, n, 128).contiguous(), values_unnorm=mk(3.0, 2.0),
                                returns=mk(4.0, 1.0), advantages=mk(2.5, -0.7))
    rbuf = types.SimpleNamespace(horizon=T, n_envs=n, obs=buf.obs.clone(), values_unnorm=buf.values_unnorm.clone(),
                                 returns=buf.returns.clone(), advantages=buf.advantages.clone())
    fused = FusedPPO(agent)
    for it in range(2):   # twice: the second pass starts from the updated statistics
        fused.finish_advantages(buf)
        ppo_dist.finish_advantages(ref, rbuf)
        torch.cuda.synchronize()
        for name in ("obs_norm", "value_norm"):
            a, r = getattr(agent, name), getattr(ref, name)
            assert torch.allclose(a.mean, r.mean, rtol=1e-9, atol=1e-12), (name, "mean", it)
            assert torch.allclose(a.var, r.var, rtol=1e-9, atol=1e-12), (name, "var", it)
            assert torch.allclose(a.count, r.count, rtol=0, atol=0), (name, "count", it)
        for f in ("advantages", "values_unnorm", "returns"):
            assert torch.allclose(getattr(buf, f), getattr(rbuf, f), rtol=1e-5, atol=1e-6), (f, it)
        w = fused.weights._t
        assert torch.allclose(w["obs_mean"], ref.obs_norm.mean.float(), rtol=1e-6, atol=1e-7)
        assert torch.allclose(w["obs_inv_std"], torch.rsqrt(ref.obs_norm.var.float() + 1e-5), rtol=1e-6)
        # feed new (already normalised-scale) numbers back in for the second pass
    assert abs(buf.advantages.mean().item()) < 1e-3 and abs(buf.advantages.std().item() - 1.0) < 1e-3


def test_randperm_is_a_permutation_and_looks_uniform():
    """bb_randperm (ppo.py:186 torch.randperm): a permutation of 0..n-1, different per draw, reproducible per (seed,
    draw), position-vs-value uncorrelated, and the minibatch gradient through int32 indices equals the int64 path."""
    import torch

    from madrona_basketball_b200.ppo_fused import FusedPPO

    agent = _agent(seed=2)
    fused, again = FusedPPO(agent, seed=11), FusedPPO(_agent(seed=2), seed=11)
    for n in (262144, 1000, 37):
        p1 = fused.randperm(n).clone()
        p2 = fused.randperm(n).clone()
        q1 = again.randperm(n).clone()
        torch.cuda.synchronize()
        assert p1.dtype == torch.int32 and torch.equal(torch.sort(p1.long())[0], torch.arange(n, device="cuda"))
        assert torch.equal(torch.sort(p2.long())[0], torch.arange(n, device="cuda"))
        assert not torch.equal(p1, p2) and torch.equal(p1, q1)
        again.randperm(n)
        if n >= 1000:
            x = torch.arange(n, device="cuda", dtype=torch.float64)
            r = torch.corrcoef(torch.stack([x, p1.double()]))[0, 1].item()
            assert abs(r) < 5.0 / n ** 0.5                      # 5 sigma of the null distribution
            fixed = (p1.long() == torch.arange(n, device="cuda")).sum().item()
            assert fixed <= 8                                    # expected 1 fixed point
    obs, act, adv, ret, g = _batch(5000)
    lp_old, v_old = _behaviour(agent, obs, act, g, 0.2)
    idx32 = fused.randperm(5000)[:2048].contiguous()
    g32 = fused.gradient(obs, act, lp_old, v_old, adv, ret, idx32).clone()
    g64 = fused.gradient(obs, act, lp_old, v_old, adv, ret, idx32.long().contiguous()).clone()
    torch.cuda.synchronize()
    assert torch.equal(g32, g64)


def test_fused_ppo_iterations_track_the_torch_path():
    """Config 4 in miniature, twice on identical engines / policies / sampling streams: (a) fused rollout + GAE +
    ppo_dist.finish_advantages + torch update_policy, (b) the same with FusedPPO.finish_advantages + FusedPPO.update fed
    the SAME permutations (torch.randperm with equal generators).  After 3 iterations the parameters agree to 2e-3
    absolute (different summation orders through Adam's normalisation, then through three rollouts of sampled actions
    that may flip near a CDF edge) and the normalisers to 1e-6 relative; the kernels read the updated weights in place."""
    import copy

    import torch

    from madrona_basketball_b200 import ppo_dist
    from madrona_basketball_b200.policy_torch import Agent
    from madrona_basketball_b200.ppo_fused import FusedPPO
    from madrona_basketball_b200.rollout import FusedRollout, PolicyWeights, RolloutBuffers, policy_forward

    torch.manual_seed(0)
    n, T = 512, 16
    dev = torch.device("cuda")
    cfg = ppo_dist.PPOConfig(num_minibatches=2, update_epochs=2)
    agent_t = Agent().to(dev)
    agent_f = copy.deepcopy(agent_t)
    sim_t, sim_f = make_engine(n), make_engine(n)
    opt = torch.optim.Adam(agent_t.parameters(), lr=cfg.learning_rate, eps=1e-8)
    fused = FusedPPO(agent_f, lr=cfg.learning_rate, eps=1e-8)
    w_t = PolicyWeights.from_module(agent_t)
    buf_t, buf_f = RolloutBuffers(T, n, dev), RolloutBuffers(T, n, dev)
    roll_t, roll_f = FusedRollout(sim_t, w_t, 1, seed=9), FusedRollout(sim_f, fused.weights, 1, seed=9)
    g_t, g_f = torch.Generator(device=dev).manual_seed(5), torch.Generator(device=dev).manual_seed(5)
    for it in range(3):
        roll_t.collect(buf_t)
        roll_t.gae(buf_t, cfg.gamma, cfg.gae_lambda, agent_t.value_norm.mean, agent_t.value_norm.var)
        ppo_dist.finish_advantages(agent_t, buf_t)
        w_t.refresh()
        ppo_dist.update_policy(agent_t, opt, buf_t, cfg, generator=g_t)

        roll_f.collect(buf_f)
        roll_f.gae(buf_f, cfg.gamma, cfg.gae_lambda, agent_f.value_norm.mean, agent_f.value_norm.var)
        fused.finish_advantages(buf_f)
        losses = fused.update(buf_f, cfg.num_minibatches, cfg.update_epochs, generator=g_f, refresh_normalizer=False,
                              permutation="torch")
        assert all(bool(torch.isfinite(v)) for v in losses.values())
        if it == 0:   # the first rollout is identical, so the first iteration's inputs are too
            assert torch.equal(buf_t.obs, buf_f.obs) and torch.equal(buf_t.actions, buf_f.actions)
            assert torch.allclose(buf_t.advantages, buf_f.advantages, rtol=1e-5, atol=1e-6)
    torch.cuda.synchronize()
    for (k, a), (_, b) in zip(agent_t.named_parameters(), agent_f.named_parameters()):
        assert torch.allclose(a, b, rtol=0, atol=2e-3), (k, float((a - b).abs().max()))
    assert torch.allclose(agent_t.obs_norm.mean, agent_f.obs_norm.mean, rtol=1e-6, atol=1e-6)
    assert float(agent_f.obs_norm.count) == 1.0 + 3 * n * T
    # the fused path's own permutation stream also trains (finite, parameters move)
    before = fused.flat.clone()
    fused.update(buf_f, cfg.num_minibatches, cfg.update_epochs, refresh_normalizer=False)
    torch.cuda.synchronize()
    assert torch.isfinite(fused.flat).all() and not torch.equal(before, fused.flat)
    obs = sim_f.observations_tensor().to_torch()[:, 1]
    _, _, v_kernel, logits = policy_forward(fused.weights, obs, stochastic=False, want_logits=True)
    with torch.no_grad():
        x = agent_f.backbone(agent_f.obs_norm(obs))
        assert torch.allclose(logits, agent_f.actor(x), rtol=1e-5, atol=2e-5)


def test_p2p_adam_with_one_rank_equals_bb_ppo_adam():
    """bb_ppo_adam_p2p at world size 1 (no peers to wait for) must be bb_ppo_adam bit for bit: same clip and Adam arithmetic."""
    import ctypes

    import torch

    from madrona_basketball_b200 import _lib
    from madrona_basketball_b200.ppo_fused import NUM_PARAMS, FusedPPO

    a, b = FusedPPO(_agent(5), max_grad_norm=0.5), FusedPPO(_agent(5), max_grad_norm=0.5)
    assert torch.equal(a.flat, b.flat)
    handle = (ctypes.c_ubyte * 64)()
    ctx = b._lib.bb_p2p_create(0, 1, NUM_PARAMS, handle)
    assert ctx, _lib.last_error()
    assert b._lib.bb_p2p_connect(ctx, handle) == 0
    b._p2p, b._p2p_timeout_ms = ctx, 100
    b.p2p_error = torch.zeros(1, dtype=torch.int32, device="cuda")
    g = torch.Generator(device="cuda").manual_seed(9)
    for _ in range(12):
        grad = torch.randn(NUM_PARAMS, device="cuda", generator=g) * 0.02
        a.grad.copy_(grad); b.grad.copy_(grad)
        a.apply(1.0)
        b.apply_p2p()
    torch.cuda.synchronize()
    assert int(b.p2p_error.item()) == 0 and int(b.step.item()) == 12
    assert torch.equal(a.flat, b.flat) and torch.equal(a.exp_avg_sq, b.exp_avg_sq) and torch.equal(a.grad_norm, b.grad_norm)
    b._lib.bb_p2p_destroy(ctx)
    b._p2p = None
    # bad arguments are refused on the host, not launched
    assert not b._lib.bb_p2p_create(3, 2, NUM_PARAMS, handle)
    assert not b._lib.bb_p2p_create(0, 17, NUM_PARAMS, handle)


def test_p2p_allreduce_adam_on_two_gpus():
    """tools/check_p2p.py under torchrun: rank-identical parameters, agreement with NCCL all-reduce + bb_ppo_adam, and a
    missing peer raises the flag after the timeout instead of hanging.  Needs two GPUs (skipped on a one-GPU box)."""
    import json
    import os
    import subprocess
    import sys

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, CHECK_P2P_STEPS="50")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                          "--master-port", "29541", os.path.join(root, "tools", "check_p2p.py")], capture_output=True, text=True,
                         timeout=300, env=env)
    assert out.returncode == 0, out.stderr[-3000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    r = json.loads(line)
    assert r["timeout_raises_flag"] is True and r["max_param_diff_vs_nccl"] < 2e-5
