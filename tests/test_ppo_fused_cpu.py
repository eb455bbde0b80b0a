"""Host logic of the fused PPO update that needs no GPU: the flat parameter layout (include/bb_ppo.h offsets) and the
re-binding of the module's parameters as views of one vector (checkpoints keep the reference's keys)."""
import os
import re

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_flat_layout_matches_header_offsets():
    from madrona_basketball_b200.policy_torch import Agent
    from madrona_basketball_b200.ppo_fused import NUM_PARAMS, PARAM_ORDER

    text = open(os.path.join(ROOT, "include", "bb_ppo.h")).read()
    offs = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define BB_PPO_OFF_(\w+)\s+(\d+)", text)}
    header_order = ["W1", "B1", "G1", "E1", "W2", "B2", "G2", "E2", "WA", "BA", "WC", "BC"]
    named = dict(Agent().named_parameters())
    assert list(named) == PARAM_ORDER
    off = 0
    for key, name in zip(PARAM_ORDER, header_order):
        assert offs[name] == off, (key, name, offs[name], off)
        off += named[key].numel()
    assert off == NUM_PARAMS == int(re.search(r"#define BB_PPO_NUM_PARAMS\s+(\d+)", text).group(1))


def test_flatten_parameters_rebinds_views_and_keeps_checkpoint_keys():
    from madrona_basketball_b200.policy_torch import Agent
    from madrona_basketball_b200.ppo_fused import PARAM_ORDER, flatten_parameters_

    torch.manual_seed(0)
    agent = Agent()
    before = {k: v.clone() for k, v in agent.state_dict().items()}
    obs = torch.randn(7, 128)
    out_before = agent.evaluate(obs)
    flat = flatten_parameters_(agent)
    assert flat.shape == (5972,) and flat.dtype == torch.float32
    after = agent.state_dict()
    assert list(after.keys()) == list(before.keys())                 # the reference's checkpoint keys, same order
    for k in before:
        assert torch.equal(after[k], before[k]), k
    assert torch.equal(agent.evaluate(obs), out_before)
    named = dict(agent.named_parameters())
    off = 0
    for k in PARAM_ORDER:                                            # views of the flat vector, in order
        p = named[k]
        assert p.data_ptr() == flat.data_ptr() + 4 * off and p.is_contiguous(), k
        off += p.numel()
    with torch.no_grad():                                            # a write through the flat vector is seen by the module
        flat.add_(1.0)
    assert torch.allclose(named["critic.bias"], before["critic.bias"] + 1.0)
    loaded = Agent()
    loaded.load_state_dict(agent.state_dict())                       # and round-trips through a plain module
    assert torch.equal(loaded.actor.weight, agent.actor.weight)
