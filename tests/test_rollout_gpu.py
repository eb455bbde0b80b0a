"""GPU parity of the fused rollout path (include/bb_rollout.h; SURVEY 8(f) n1 + n2) against a plain
torch fp32 reference of scripts/agent.py / action.py / ppo.py.  These are floating-point kernels:
tolerance 1e-5 relative (+1e-5 absolute on logits of order 1), stated per assert."""
import numpy as np
import pytest

import oracle
from helpers import ENV_ARGS, assert_states_equal, engine_state, make_engine, make_oracle
from ref_agent import ACTION_BUCKETS, RefAgent, randomize

pytestmark = pytest.mark.gpu

OFF = np.concatenate([[0], np.cumsum(ACTION_BUCKETS)[:-1]])


def _uniforms(seed, rows, draw):
    """The sampler's uniforms: Threefry2x32(key = (seed, row), counter = (draw, c)) for c = 0..2 gives the
    words of buckets 2c (x0) and 2c+1 (x1); u = (word >> 8) * 2^-24."""
    u = np.empty((len(rows), 6), np.float32)
    for i, r in enumerate(rows):
        for c in range(3):
            x0, x1 = oracle.threefry2x32(seed, int(r), draw, c)
            u[i, 2 * c] = np.float32((x0 >> 8) * 2.0 ** -24)
            u[i, 2 * c + 1] = np.float32((x1 >> 8) * 2.0 ** -24)
    return u


def _obs_batch(n, seed=0):
    """Real observation rows: a few hundred simulated steps, both agents' rows."""
    import torch

    sim = make_engine(n)
    for s in range(40):
        sim.fill_actions(77, s)
        sim.step()
    torch.cuda.synchronize()
    return sim, sim.observations_tensor().to_torch()


@pytest.fixture(params=["mma_sync", "tcgen05"])
def policy_flavour(request, monkeypatch):
    """bb_policy_forward's two kernels: the default warp-level one (mma.sync 3xTF32) and the tcgen05 / TMEM one
    (BB_POLICY_UMMA, csrc/bb_umma.cuh); both are held to the same torch fp32 reference."""
    monkeypatch.setenv("BB_POLICY_UMMA_FORCE", "1" if request.param == "tcgen05" else "0")
    return request.param


@pytest.mark.parametrize("hidden", [32, 16])
def test_policy_forward_matches_torch(hidden, policy_flavour):
    import torch

    from madrona_basketball_b200.rollout import PolicyWeights, policy_forward

    n = 3000
    sim, obs = _obs_batch(n)
    agent = randomize(RefAgent(num_channels=hidden), seed=hidden).cuda()
    w = PolicyWeights.from_module(agent)
    for idx in (0, 1):
        view = obs[:, idx]                                    # [n,128], row stride 256: read in place
        act, logp, val, logits = policy_forward(w, view, stochastic=True, seed=5, draw=9, want_logits=True)
        with torch.no_grad():
            ref_logits, ref_val = agent.logits_value(view)
            lsm = agent.bucket_log_softmax(ref_logits)
        torch.cuda.synchronize()
        assert torch.allclose(logits, ref_logits, rtol=1e-5, atol=2e-5), (logits - ref_logits).abs().max()
        assert torch.allclose(val, ref_val, rtol=1e-5, atol=2e-5)
        # sampling: same uniforms, inverse CDF over the reference's probabilities
        rows = np.arange(0, n, 37)
        u = _uniforms(5, rows, 9)
        a = act.cpu().numpy()
        ref_logp = np.zeros(len(rows), np.float64)
        for b in range(6):
            p = lsm[b][rows].exp().double().cpu().numpy()
            cdf = np.cumsum(p, axis=1)
            want = np.minimum((u[:, b:b + 1] >= cdf).sum(axis=1), ACTION_BUCKETS[b] - 1)
            margin = np.abs(cdf - u[:, b:b + 1]).min(axis=1)
            bad = (a[rows, b] != want) & (margin > 1e-5)          # only a uniform within 1e-5 of a CDF edge may flip
            assert not bad.any(), (b, np.flatnonzero(bad)[:5])
            ref_logp += lsm[b][rows].double().cpu().numpy()[np.arange(len(rows)), a[rows, b]]
        assert np.allclose(logp.cpu().numpy()[rows], ref_logp, rtol=1e-5, atol=2e-5)
        assert (a >= 0).all() and (a < np.array(ACTION_BUCKETS)).all()
        # deterministic mode = argmax per bucket (action.py:21-23)
        act_d, logp_d, _ = policy_forward(w, view, stochastic=False)
        ad = act_d.cpu().numpy()
        for b in range(6):
            sl = ref_logits[:, OFF[b]:OFF[b] + ACTION_BUCKETS[b]]
            top2 = sl.topk(2, dim=1).values
            clear = ((top2[:, 0] - top2[:, 1]) > 1e-4).cpu().numpy()
            assert np.array_equal(ad[clear, b], sl.argmax(dim=1).cpu().numpy()[clear])


def test_sampler_follows_the_categorical_distribution(policy_flavour):
    """Many draws for a handful of rows: empirical frequencies vs softmax probabilities (chi-square)."""
    import torch

    from madrona_basketball_b200.rollout import PolicyWeights, policy_forward

    agent = randomize(RefAgent(), seed=3).cuda()
    w = PolicyWeights.from_module(agent)
    sim, obs = _obs_batch(8)
    row = obs[3:4, 0].contiguous()
    reps = 20000
    batch = row.expand(reps, 128).contiguous()
    act, _, _ = policy_forward(w, batch, stochastic=True, seed=11, draw=0)     # rows differ only in their RNG key
    with torch.no_grad():
        lsm = agent.bucket_log_softmax(agent.logits_value(row)[0])
    a = act.cpu().numpy()
    for b in range(6):
        p = lsm[b][0].exp().double().cpu().numpy()
        cnt = np.bincount(a[:, b], minlength=ACTION_BUCKETS[b])
        exp = p * reps
        keep = exp > 5
        chi2 = (((cnt - exp) ** 2) / np.maximum(exp, 1e-9))[keep].sum()
        assert chi2 < 40.0, (b, chi2, cnt, exp)      # dof <= 7: p(chi2 > 40) < 1e-6


def _reference_rollout(sim, w, frozen_w, buf_shape, agent_idx, seed, draw_base, stats=None):
    """ppo.py rollout() step by step in Python: bb_policy_forward for the actions (same streams), sim.step(),
    torch bookkeeping as in ppo.py:84-137 / env.py:147,167-185."""
    import torch

    from madrona_basketball_b200.rollout import policy_forward

    T, n = buf_shape
    dev = torch.device("cuda")
    out = {k: torch.zeros((T, n) + s, device=dev) for k, s in
           [("obs", (128,)), ("actions", (6,)), ("values", ()), ("log_probs", ()), ("rewards", ()), ("not_dones", ())]}
    obs_t, act_t = sim.observations_tensor().to_torch(), sim.action_tensor().to_torch()
    rew_t, done_t, reset_t = sim.reward_tensor().to_torch(), sim.done_tensor().to_torch(), sim.reset_tensor().to_torch()
    reset_t.fill_(1); act_t[:, agent_idx] = 0; sim.step(); reset_t.fill_(0)        # env.reset()
    if stats is None:   # PPOStats lives across iterations (ppo.py:279)
        stats = [torch.zeros(n, device=dev), torch.zeros(n, device=dev)]
    cur, length, sums = stats[0], stats[1], torch.zeros(3, dtype=torch.float64, device=dev)
    for t in range(T):
        obs = obs_t[:, agent_idx].clone()
        a, lp, v = policy_forward(w, obs_t[:, agent_idx], True, seed, draw_base + t)
        act_t[:, agent_idx] = a
        if frozen_w is not None:
            fa, _, _ = policy_forward(frozen_w, obs_t[:, 1 - agent_idx], True, seed, 0x80000000 + draw_base + t)
            act_t[:, 1 - agent_idx] = fa
        sim.step()
        rew, done = rew_t[:, agent_idx].clone(), done_t[:, agent_idx].clone()
        out["obs"][t], out["actions"][t], out["values"][t], out["log_probs"][t] = obs, a.float(), v, lp
        out["rewards"][t], out["not_dones"][t] = rew, 1.0 - done
        cur += rew; length += 1                                                   # ppo_stats.py:21-36
        fin = done != 0
        sums += torch.stack([cur[fin].double().sum(), length[fin].double().sum(), fin.double().sum()])
        cur *= (1 - done); length *= (1 - done)
    from madrona_basketball_b200.rollout import policy_forward as pf
    _, _, nv = pf(w, obs_t[:, agent_idx], False)
    out["next_value"], out["cur_reward"], out["ep_length"], out["ep_sums"] = nv, cur, length, sums
    return out


@pytest.mark.parametrize("form", ["one_launch", "one_launch_two_tiles_per_block", "kernel_chain"])
@pytest.mark.parametrize("agent_idx,with_frozen", [(1, False), (0, True)])
def test_fused_rollout_matches_stepwise_loop(agent_idx, with_frozen, form, monkeypatch):
    """bb_rollout in its three forms -- the whole rollout as one launch of bb_rollout_loop_kernel with one or two tiles per block
    (the latter: n = 700 is 44 tiles, so the last block's second tile... is a real one; 22 blocks) and the chain of policy / step
    kernels (BB_ROLLOUT_LOOP=0) -- against ppo.py's rollout() replayed step by step: every buffer bit for bit."""
    import torch

    # 2 / 3: the looping kernel with one / two tiles per block, also when the grid is not resident at once
    monkeypatch.setenv("BB_ROLLOUT_LOOP", {"one_launch": "2", "one_launch_two_tiles_per_block": "3", "kernel_chain": "0"}[form])
    monkeypatch.setenv("BB_POLICY_UMMA_FORCE", "0")     # the stepwise reference below calls bb_policy_forward: same kernel family

    from madrona_basketball_b200.rollout import FusedRollout, PolicyWeights, RolloutBuffers

    n, T = 700, 32
    agent = randomize(RefAgent(), seed=1, scale_heads=2.0).cuda()
    frozen = randomize(RefAgent(), seed=2, scale_heads=2.0).cuda() if with_frozen else None
    w = PolicyWeights.from_module(agent)
    fw = PolicyWeights.from_module(frozen) if frozen is not None else None
    a, b = make_engine(n), make_engine(n)
    for s in range(25):
        for e in (a, b):
            e.fill_actions(1234, s); e.step()
    buf = RolloutBuffers(T, n, torch.device("cuda"))
    fr = FusedRollout(a, w, agent_idx, frozen=fw, seed=321)
    ref_all, stats = [], None
    for it in range(3):                                     # three consecutive rollouts: draws advance, stats persist
        fr.collect(buf)
        ref = _reference_rollout(b, w, fw, (T, n), agent_idx, 321, it * T, stats)
        stats = [ref["cur_reward"], ref["ep_length"]]
        torch.cuda.synchronize()
        for k in ("obs", "actions", "values", "log_probs", "rewards", "not_dones", "next_value"):
            got, want = getattr(buf, k), ref[k]
            assert torch.equal(got.view(torch.int32), want.view(torch.int32)), (it, k)
        ref_all.append(ref)
    assert_states_equal(engine_state(a), engine_state(b), "engine state after fused rollouts")
    assert torch.equal(buf.cur_reward, ref_all[-1]["cur_reward"]) and torch.equal(buf.ep_length, ref_all[-1]["ep_length"])
    sums = sum(r["ep_sums"] for r in ref_all)
    assert torch.allclose(buf.ep_sums, sums, rtol=1e-9) and buf.ep_sums[2] > 0
    st = buf.episode_stats()
    assert st["episodes"] == float(sums[2]) and st["mean_episode_length"] > 1
    # the simulator inside the rollout is the same step path: replay the recorded actions on the oracle
    if not with_frozen:
        orc, c = make_oracle(n), make_engine(n)
        for s in range(25):
            orc.fill_actions(1234, s); orc.step()
            c.fill_actions(1234, s); c.step()
        fr2 = FusedRollout(c, w, agent_idx, seed=321)
        buf2 = RolloutBuffers(T, n, torch.device("cuda"))
        fr2.collect(buf2)
        torch.cuda.synchronize()
        orc.tensor("reset")[...] = 1; orc.tensor("action")[:, agent_idx] = 0; orc.step(); orc.tensor("reset")[...] = 0
        acts = buf2.actions.cpu().numpy().astype(np.int32)
        for t in range(T):
            orc.tensor("action")[:, agent_idx] = acts[t]
            orc.step()
            assert np.array_equal(buf2.rewards[t].cpu().numpy().view(np.int32), orc.tensor("reward")[:, agent_idx].view(np.int32)), t
        assert_states_equal(engine_state(c), orc.state(), "oracle replay of the fused rollout")


def test_gae_kernel_matches_ppo_loop():
    """compute_advantages (ppo.py:144-166) restated in torch vs bb_gae."""
    import torch

    from madrona_basketball_b200.rollout import FusedRollout, PolicyWeights, RolloutBuffers

    n, T = 5000, 32
    dev = torch.device("cuda")
    g = torch.Generator(device="cuda").manual_seed(0)
    buf = RolloutBuffers(T, n, dev)
    buf.values.copy_(torch.randn((T, n), generator=g, device=dev) * 3)          # some beyond the +-5 clamp
    buf.next_value.copy_(torch.randn(n, generator=g, device=dev) * 3)
    buf.rewards.copy_(torch.randn((T, n), generator=g, device=dev) * 2 - 0.5)
    buf.not_dones.copy_((torch.rand((T, n), generator=g, device=dev) > 0.05).float())
    agent = randomize(RefAgent()).cuda()
    sim = make_engine(n)
    fr = FusedRollout(sim, PolicyWeights.from_module(agent), 1)
    gamma, lam = 0.998, 0.95
    fr.gae(buf, gamma, lam, agent.value_norm.mean, agent.value_norm.var)
    values = agent.value_norm(buf.values, unnorm=True)
    next_value = agent.value_norm(buf.next_value, unnorm=True)
    adv = torch.zeros_like(values)
    last = 0.0
    for t in reversed(range(T)):
        if t == T - 1:
            nnt, nv = buf.not_dones[t], next_value
        else:
            nnt, nv = buf.not_dones[t + 1], values[t + 1]
        delta = buf.rewards[t] + gamma * nv * nnt - values[t]
        adv[t] = last = delta + gamma * lam * nnt * last
    ret = adv + values
    torch.cuda.synchronize()
    assert torch.allclose(buf.values_unnorm, values, rtol=1e-6, atol=1e-6)
    assert torch.allclose(buf.advantages, adv, rtol=1e-5, atol=1e-4)
    assert torch.allclose(buf.returns, ret, rtol=1e-5, atol=1e-4)


def test_ppo_iterations_train_end_to_end():
    """Config 4 in miniature: fused rollout + GAE + torch PPO update (ppo_dist, single process) for a few
    iterations: finite losses, parameters move, the kernels see the updated weights without copies."""
    import torch

    from madrona_basketball_b200 import ppo_dist
    from madrona_basketball_b200.policy_torch import Agent
    from madrona_basketball_b200.rollout import FusedRollout, PolicyWeights, RolloutBuffers, policy_forward

    torch.manual_seed(0)
    n, T = 512, 16
    dev = torch.device("cuda")
    sim = make_engine(n)
    agent = Agent().to(dev)
    cfg = ppo_dist.PPOConfig(num_minibatches=2, update_epochs=2)
    opt = torch.optim.Adam(agent.parameters(), lr=1e-3, eps=1e-8)
    w = PolicyWeights.from_module(agent)
    buf = RolloutBuffers(T, n, dev)
    roll = FusedRollout(sim, w, 1, seed=9)
    w0 = agent.backbone[0].weight.detach().clone()
    for it in range(3):
        agent.eval()
        roll.collect(buf)
        roll.gae(buf, cfg.gamma, cfg.gae_lambda, agent.value_norm.mean, agent.value_norm.var)
        ppo_dist.finish_advantages(agent, buf)
        w.refresh()
        assert abs(float(buf.advantages.mean())) < 1e-4 and abs(float(buf.advantages.std()) - 1.0) < 1e-3
        agent.train()
        losses = ppo_dist.update_policy(agent, opt, buf, cfg)
        assert all(torch.isfinite(v) for v in losses.values())
    assert not torch.equal(w0, agent.backbone[0].weight)
    assert float(agent.obs_norm.count) == 1.0 + 3 * n * T
    # zero-copy: the kernel reads the live (updated) parameters and the refreshed normaliser
    obs = sim.observations_tensor().to_torch()[:, 1]
    _, _, v_kernel, logits = policy_forward(w, obs, stochastic=False, want_logits=True)
    with torch.no_grad():
        x = agent.backbone(agent.obs_norm(obs))
        assert torch.allclose(logits, agent.actor(x), rtol=1e-5, atol=2e-5)
        assert torch.allclose(v_kernel, agent.critic(x).squeeze(-1), rtol=1e-5, atol=2e-5)


@pytest.mark.parametrize("k", [8, 32, 128])
def test_umma_selftest_matches_float64_matmul(k):
    """csrc/bb_umma.cuh: tcgen05.mma kind::tf32 (M = 128, N = 32, accumulator in tensor memory) issued as 3xTF32 gives fp32-level
    products: D = A . B^T against the float64 product, on random and on exactly representable inputs."""
    import ctypes

    import torch

    from madrona_basketball_b200 import _lib

    lib = _lib.load()
    lib.bb_umma_selftest.restype = ctypes.c_int
    lib.bb_umma_selftest.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    g = torch.Generator(device="cuda").manual_seed(5 + k)
    for exact in (True, False):
        if exact:   # small integers: every product and sum is exact in tf32 / fp32 -> bitwise equality expected
            a = torch.randint(-8, 9, (128, k), device="cuda", generator=g).float()
            b = torch.randint(-8, 9, (32, k), device="cuda", generator=g).float()
        else:
            a = torch.randn(128, k, device="cuda", generator=g)
            b = torch.randn(32, k, device="cuda", generator=g)
        d = torch.full((128, 32), float("nan"), device="cuda")
        rc = lib.bb_umma_selftest(a.data_ptr(), b.data_ptr(), d.data_ptr(), k, torch.cuda.current_stream().cuda_stream)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        want = a.double() @ b.double().t()
        if exact:
            assert torch.equal(d.double(), want)
        else:
            err = (d.double() - want).abs().max().item()
            assert err < 2e-5 * max(1.0, want.abs().max().item()), err   # plain TF32 would be ~1e-2 here
