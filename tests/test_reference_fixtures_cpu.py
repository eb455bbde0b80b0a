"""tests/golden/reference_ppo.npz was produced by the REFERENCE'S OWN scripts (agent.py, action.py, buffers.py,
ppo.py compute_advantages / update_policy; generator tools/make_golden_ppo_ref.py).  Here, on the CPU: the repo's
torch mirrors of that code (policy_torch.Agent = the multi-GPU trainer's module, tests/ref_agent.RefAgent = the
kernels' torch reference, ppo_dist.finish_advantages / update_policy) reproduce it, so the GPU tests that compare the
kernels with those mirrors are anchored to the reference and not only to this repo's reading of it."""
import os

import numpy as np
import torch

from ref_agent import RefAgent

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_ppo.npz")


def _gold():
    return np.load(GOLD)


def _state_dict(g, prefix="sd0::"):
    return {k[len(prefix):]: torch.from_numpy(g[k]) for k in g.files if k.startswith(prefix)}


def _load(agent, g, prefix="sd0::"):
    sd = _state_dict(g, prefix)
    missing, unexpected = agent.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    return agent


def test_fixture_has_the_reference_checkpoint_keys():
    """The fixture's state_dict is the reference Agent's (agent.py:108-131): the repo's modules load it strictly."""
    from madrona_basketball_b200.policy_torch import Agent

    g = _gold()
    a = Agent(128, 32, 2, [2, 8, 3, 2, 2, 2])        # the reference's positional signature (agent.py:109)
    a.load_state_dict(_state_dict(g), strict=True)
    RefAgent().load_state_dict(_state_dict(g), strict=True)


def test_torch_mirrors_reproduce_the_reference_forward():
    """Agent.get_stats / evaluate / logits (agent.py:134-178, action.py:37-45) on 768 rows: float32, identical
    operation order => agreement to 1e-6 absolute on logits of order 1-10."""
    from madrona_basketball_b200.policy_torch import Agent

    g = _gold()
    obs, act = torch.from_numpy(g["fwd_obs"]), torch.from_numpy(g["fwd_actions"])
    a = _load(Agent(), g)
    r = _load(RefAgent(), g)
    with torch.no_grad():
        lp, ent, v = a.get_stats(obs, act)
        logits, v2 = r.logits_value(obs)
    assert np.allclose(lp.numpy(), g["fwd_logp"], rtol=1e-5, atol=1e-5)
    assert np.allclose(ent.numpy(), g["fwd_entropy"], rtol=1e-5, atol=1e-5)
    assert np.allclose(v.numpy(), g["fwd_value"], rtol=1e-5, atol=1e-5)
    assert np.allclose(logits.numpy(), g["fwd_logits"], rtol=1e-5, atol=1e-5)
    assert np.allclose(v2.numpy(), g["fwd_evaluate"], rtol=1e-5, atol=1e-5)
    # argmax actions (action.py:20-23)
    off = 0
    for b, nb in enumerate([2, 8, 3, 2, 2, 2]):
        assert np.array_equal(logits[:, off:off + nb].argmax(dim=1).numpy(), g["fwd_best"][:, b])
        off += nb


def _gae(values, next_value, rewards, not_dones, gamma, lam):
    """ppo.py:151-165."""
    T = values.shape[0]
    adv = torch.zeros_like(values)
    last = 0.0
    for t in reversed(range(T)):
        nnt, nv = (not_dones[t], next_value) if t == T - 1 else (not_dones[t + 1], values[t + 1])
        delta = rewards[t] + gamma * nv * nnt - values[t]
        adv[t] = last = delta + gamma * lam * nnt * last
    return adv


def test_ppo_dist_reproduces_compute_advantages_and_update_policy():
    """ppo_dist.finish_advantages + update_policy (the multi-GPU trainer's torch path, single process here) against
    the reference's compute_advantages (ppo.py:144-177) and update_policy (ppo.py:180-216) outputs."""
    import types

    from madrona_basketball_b200 import ppo_dist
    from madrona_basketball_b200.policy_torch import Agent

    g = _gold()
    a = _load(Agent(), g)
    t = lambda k: torch.from_numpy(g["gae_in_" + k].copy())
    T, N = g["gae_in_values"].shape
    values = a.value_norm(t("values"), unnorm=True)
    next_value = a.value_norm(t("next_value"), unnorm=True)
    adv = _gae(values, next_value, t("rewards"), t("not_dones"), float(g["gae_gamma"]), float(g["gae_lambda"]))
    buf = types.SimpleNamespace(horizon=T, n_envs=N, obs=t("obs"), actions=t("actions"), log_probs=t("log_probs"), values=t("values"),
                                values_unnorm=values.clone(), advantages=adv, returns=adv + values)
    ppo_dist.finish_advantages(a, buf)
    assert np.allclose(buf.advantages.numpy(), g["gae_advantages"], rtol=1e-4, atol=1e-5)
    assert np.allclose(buf.returns.numpy(), g["gae_returns"], rtol=1e-4, atol=1e-5)
    assert np.array_equal(buf.values.numpy(), g["gae_values_after"])      # ppo.py:175 rebinds a local: the buffer keeps the old values
    for k in ("obs_norm.mean", "obs_norm.var", "obs_norm.count", "value_norm.mean", "value_norm.var", "value_norm.count"):
        # the reference takes the batch moments in float32 (torch.var_mean of the float32 buffer, agent.py:43) before promoting
        # them; this repo accumulates them in float64 (one pass, all ranks): agreement is float32 rounding, not 1e-9
        assert np.allclose(a.state_dict()[k].numpy(), g["sd1::" + k], rtol=2e-6, atol=1e-7), k
    cfg = ppo_dist.PPOConfig(num_minibatches=1, update_epochs=1)
    opt = torch.optim.Adam(a.parameters(), lr=cfg.learning_rate, eps=1e-8)
    ppo_dist.update_policy(a, opt, buf, cfg)
    for k, p in a.named_parameters():
        gr = g["upd1_clipped_grad::" + k]
        assert np.allclose(p.grad.numpy(), gr, rtol=1e-3, atol=1e-5 * max(np.abs(gr).max(), 1e-6)), k
        assert np.allclose(p.detach().numpy(), g["upd1::" + k], rtol=0, atol=2e-6), k     # one Adam step moves each weight by ~lr = 3e-4
    cfg = ppo_dist.PPOConfig(num_minibatches=1, update_epochs=2)
    ppo_dist.update_policy(a, opt, buf, cfg)
    for k, p in a.named_parameters():
        assert np.allclose(p.detach().numpy(), g["upd3::" + k], rtol=0, atol=1e-5), k
