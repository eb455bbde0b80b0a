"""The fused policy / GAE / PPO-update kernels against fixtures produced by the REFERENCE'S OWN scripts
(tests/golden/reference_ppo.npz; generator tools/make_golden_ppo_ref.py imports /root/reference/scripts/agent.py,
action.py, buffers.py and ppo.py's compute_advantages / update_policy).  Floating-point kernels; tolerances per assert."""
import os
import types

import numpy as np
import pytest

from helpers import make_engine

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_ppo.npz")
BUCKETS = [2, 8, 3, 2, 2, 2]


def _agent(g, prefix="sd0::"):
    import torch

    from madrona_basketball_b200.policy_torch import Agent

    a = Agent(128, 32, 2, BUCKETS)
    a.load_state_dict({k[len(prefix):]: torch.from_numpy(g[k]) for k in g.files if k.startswith(prefix)}, strict=True)
    return a.cuda()


@pytest.mark.parametrize("flavour", ["mma_sync", "tcgen05"])
def test_policy_kernel_matches_reference_agent_forward(flavour, monkeypatch):
    """bb_policy_forward vs Agent.norm_obs_backbone + actor / critic (agent.py:134-154) and
    DiscreteActionDistributions.best / action_stats (action.py:20-45): logits / value rtol 1e-5 + atol 2e-5
    (3xTF32 products), argmax actions equal wherever the top two logits of a bucket differ by more than 1e-4."""
    import torch

    from madrona_basketball_b200.rollout import PolicyWeights, policy_forward

    monkeypatch.setenv("BB_POLICY_UMMA_FORCE", "1" if flavour == "tcgen05" else "0")   # csrc/bb_umma.cuh vs the mma.sync kernel
    g = np.load(GOLD)
    agent = _agent(g)
    w = PolicyWeights.from_module(agent)
    obs = torch.from_numpy(g["fwd_obs"]).cuda()
    act, logp, val, logits = policy_forward(w, obs, stochastic=False, want_logits=True)
    torch.cuda.synchronize()
    assert np.allclose(logits.cpu().numpy(), g["fwd_logits"], rtol=1e-5, atol=2e-5), np.abs(logits.cpu().numpy() - g["fwd_logits"]).max()
    assert np.allclose(val.cpu().numpy(), g["fwd_value"], rtol=1e-5, atol=2e-5)
    a = act.cpu().numpy()
    off, clear_all = 0, np.ones(len(a), bool)
    for b, nb in enumerate(BUCKETS):
        sl = np.sort(g["fwd_logits"][:, off:off + nb], axis=1)
        clear = (sl[:, -1] - sl[:, -2]) > 1e-4
        assert np.array_equal(a[clear, b], g["fwd_best"][clear, b]), b
        clear_all &= clear
        off += nb
    assert clear_all.mean() > 0.9
    assert np.allclose(logp.cpu().numpy()[clear_all], g["fwd_best_logp"][clear_all], rtol=1e-5, atol=3e-5)


def _buffers(g, dev):
    import torch

    from madrona_basketball_b200.rollout import RolloutBuffers

    T, N = g["gae_in_values"].shape
    buf = RolloutBuffers(T, N, dev)
    for k in ("obs", "actions", "values", "next_value", "rewards", "not_dones", "log_probs"):
        getattr(buf, k).copy_(torch.from_numpy(g["gae_in_" + k]))
    return buf


def test_gae_and_normaliser_kernels_match_reference_compute_advantages():
    """bb_gae + bb_ppo_moments + bb_ppo_finish_advantages vs ppo.py:144-177 run by the reference itself.
    float32 outputs rtol 1e-4 (the scan accumulates 16 steps), normaliser state to float32 rounding (the reference
    takes the batch moments in float32, the kernels in float64)."""
    import torch

    from madrona_basketball_b200.ppo_fused import FusedPPO
    from madrona_basketball_b200.rollout import FusedRollout

    g = np.load(GOLD)
    dev = torch.device("cuda")
    agent = _agent(g)
    buf = _buffers(g, dev)
    fused = FusedPPO(agent)
    sim = make_engine(buf.n_envs)
    roll = FusedRollout(sim, fused.weights, 1)
    roll.gae(buf, float(g["gae_gamma"]), float(g["gae_lambda"]), agent.value_norm.mean, agent.value_norm.var)
    fused.finish_advantages(buf)
    torch.cuda.synchronize()
    assert np.allclose(buf.advantages.cpu().numpy(), g["gae_advantages"], rtol=1e-4, atol=2e-5)
    assert np.allclose(buf.returns.cpu().numpy(), g["gae_returns"], rtol=1e-4, atol=2e-5)
    assert np.array_equal(buf.values.cpu().numpy(), g["gae_values_after"])        # ppo.py:175 leaves the buffer's values alone
    sd = agent.state_dict()
    for k in ("obs_norm.mean", "obs_norm.var", "obs_norm.count", "value_norm.mean", "value_norm.var", "value_norm.count"):
        assert np.allclose(sd[k].cpu().numpy(), g["sd1::" + k], rtol=2e-6, atol=1e-7), k


def test_fused_update_matches_reference_update_policy():
    """bb_ppo_grad + bb_ppo_adam vs the reference's update_policy (ppo.py:180-216): one minibatch = the whole 1600-row
    buffer, so torch.randperm's stream does not matter.  Clipped gradient within 2e-4 of each tensor's largest entry
    (3xTF32 products, different summation order); parameters after 1 step within 2e-6 and after 3 steps within 1e-5
    absolute (each Adam step moves a weight by about lr = 3e-4)."""
    import torch

    from madrona_basketball_b200.ppo_fused import PARAM_ORDER, FusedPPO

    g = np.load(GOLD)
    dev = torch.device("cuda")
    agent = _agent(g)
    for k in ("obs_norm.mean", "obs_norm.var", "obs_norm.count", "value_norm.mean", "value_norm.var", "value_norm.count"):
        getattr(getattr(agent, k.split(".")[0]), k.split(".")[1]).copy_(torch.from_numpy(g["sd1::" + k]))   # state after compute_advantages
    buf = _buffers(g, dev)
    buf.advantages.copy_(torch.from_numpy(g["gae_advantages"]))
    buf.returns.copy_(torch.from_numpy(g["gae_returns"]))
    fused = FusedPPO(agent, lr=float(g["args_learning_rate"]), clip_coef=float(g["args_clip_coef"]), ent_coef=float(g["args_ent_coef"]),
                     vf_coef=float(g["args_vf_coef"]), max_grad_norm=float(g["args_max_grad_norm"]))
    fused.update(buf, num_minibatches=1, update_epochs=1)
    torch.cuda.synchronize()
    flat_grad = fused.grad.cpu().numpy()
    norm = float(np.sqrt((flat_grad.astype(np.float64) ** 2).sum()))
    clipped = flat_grad * min(1.0, float(g["args_max_grad_norm"]) / (norm + 1e-6))       # clip_grad_norm_ (ppo.py:212)
    named = dict(agent.named_parameters())
    off = 0
    for k in PARAM_ORDER:
        want = g["upd1_clipped_grad::" + k]
        got = clipped[off:off + want.size].reshape(want.shape)
        off += want.size
        assert np.abs(got - want).max() <= 2e-4 * np.abs(want).max() + 1e-9, (k, np.abs(got - want).max(), np.abs(want).max())
        assert np.allclose(named[k].detach().cpu().numpy(), g["upd1::" + k], rtol=0, atol=2e-6), k
    fused.update(buf, num_minibatches=1, update_epochs=2)
    torch.cuda.synchronize()
    for k in PARAM_ORDER:
        assert np.allclose(named[k].detach().cpu().numpy(), g["upd3::" + k], rtol=0, atol=1e-5), k


def test_fused_update_takes_the_remainder_minibatch_like_the_reference():
    """ppo.py:188 `range(0, total_steps, minibatch_size)`: 1600 rows with 3 minibatches = 533 + 533 + 533 + 1 rows, four
    optimiser steps per epoch."""
    import torch

    from madrona_basketball_b200.ppo_fused import FusedPPO

    g = np.load(GOLD)
    agent = _agent(g)
    buf = _buffers(g, torch.device("cuda"))
    buf.advantages.copy_(torch.from_numpy(g["gae_advantages"]))
    buf.returns.copy_(torch.from_numpy(g["gae_returns"]))
    fused = FusedPPO(agent)
    fused.update(buf, num_minibatches=3, update_epochs=1)
    torch.cuda.synchronize()
    assert int(fused.step.item()) == 4
