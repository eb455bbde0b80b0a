"""GPU: the CUDA engine against the committed fixtures of tools/make_golden.py (no oracle in the
loop), the device arithmetic spec against the host's, and an EnvWrapper-shaped driving loop
(the call sequence of the reference's scripts/env.py, which must run unmodified)."""
import ctypes
import os

import numpy as np
import pytest

from helpers import ALL_COLUMNS, make_engine

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SEED = 1234


def gold(name):
    return np.load(os.path.join(GOLD, name))


def _checksums(sim):
    return sim.checksum().cpu().numpy().view(np.uint64)


def test_config1_trace_matches_golden():
    """BASELINE.json configs[0]: N=1, random actions, 1000 steps, per-step checksum + per-step columns."""
    import torch

    g = gold("config1_n1_random_1000.npz")
    sim = make_engine(1)
    sums = torch.empty((1000, 1), dtype=torch.int64, device="cuda")
    rew, done, poss = [], [], []
    for s in range(1000):
        sim.fill_actions(SEED, s)
        sim.step()
        sums[s] = sim.checksum()
        rew.append(sim.reward_tensor().to_torch().clone())
        done.append(sim.done_tensor().to_torch().clone())
        poss.append(sim.agent_possession_tensor().to_torch().clone())
    assert np.array_equal(sums.cpu().numpy().view(np.uint64), g["checksum"])
    assert np.array_equal(torch.stack(rew).cpu().numpy().view(np.int32), g["trace_reward"].view(np.int32))
    assert np.array_equal(torch.stack(done).cpu().numpy(), g["trace_done"])
    assert np.array_equal(torch.stack(poss).cpu().numpy(), g["trace_agent_possession"])
    for k in ALL_COLUMNS:
        a = sim.tensor(k).to_torch().cpu().numpy()
        assert np.array_equal(a.view(np.int32), g[f"final_{k}"].view(np.int32)), k


def test_zero_action_and_many_world_traces_match_golden():
    g = gold("n1_zero_action_700.npz")
    sim = make_engine(1)
    act = sim.action_tensor().to_torch()
    sums = []
    for s in range(700):
        act[:, 0] = 0
        sim.step()
        sums.append(sim.checksum())
    import torch

    assert np.array_equal(torch.stack(sums).cpu().numpy().view(np.uint64), g["checksum"])

    g = gold("n64_random_300.npz")
    sim = make_engine(64)
    sums = []
    for s in range(300):
        sim.fill_actions(SEED, s)
        sim.step()
        if s % 10 == 0:
            sums.append(sim.checksum())
    assert np.array_equal(torch.stack(sums).cpu().numpy().view(np.uint64), g["checksum"])


def test_dead_paths_and_collision_stress_match_golden():
    import torch

    g = gold("n32_deadpaths_400.npz")
    sim = make_engine(32, flags=0)
    sim.game_state_tensor().to_torch().view(torch.int32)[::2, 13] = 0
    sums = []
    for s in range(400):
        sim.fill_actions(SEED, s)
        sim.step()
        if s % 10 == 0:
            sums.append(sim.checksum())
    assert np.array_equal(torch.stack(sums).cpu().numpy().view(np.uint64), g["checksum"])

    g = gold("n16_collision_stress_60.npz")
    sim = make_engine(16)
    pe = sim.agent_pos_tensor().to_torch()
    sums = []
    for s in range(60):
        sim.fill_actions(SEED, s)
        pe[:, 1] = pe[:, 0] + pe.new_tensor([0.05, 0.0, 0.0])
        sim.step()
        sums.append(sim.checksum())
    assert np.array_equal(torch.stack(sums).cpu().numpy().view(np.uint64), g["checksum"])


def test_device_math_spec_is_bit_identical_to_host_spec():
    """include/bb_math.h evaluated by sm_100a == evaluated by x86-64 (the claim the parity design rests on):
    committed vectors first, then two million fresh inputs per function against the host build."""
    import torch

    import oracle
    from madrona_basketball_b200 import _lib

    lib = _lib.load()
    g = gold("math_spec.npz")
    rng = np.random.default_rng(11)
    ranges = {"sinf": (-8, 8), "cosf": (-8, 8), "atanf": (-64, 64), "atan2f": (-30, 30), "acosf": (-1, 1),
              "expf": (-20, 2), "erff": (-5, 5)}
    for fn, code in oracle.MATH_FNS.items():
        for big in (False, True):
            if big:
                x = rng.uniform(*ranges[fn], 1 << 21).astype(np.float32)
                y = rng.uniform(-30, 30, 1 << 21).astype(np.float32)
                want = oracle.math_eval(fn, x, y if fn == "atan2f" else None)
            else:
                x, y, want = g[f"{fn}_x"], g["atan2f_y"], g[f"{fn}_out"]
            xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
            out = torch.empty_like(xd)
            rc = lib.bb_math_eval(code, xd.data_ptr(), yd.data_ptr(), out.data_ptr(), x.size,
                                  torch.cuda.current_stream().cuda_stream)
            assert rc == 0, _lib.last_error()
            got = out.cpu().numpy()
            assert np.array_equal(got.view(np.int32), want.view(np.int32)), (fn, big)


def test_env_wrapper_call_sequence():
    """The calls scripts/env.py makes (ctor env.py:25-35, five views 75-79, step 125-170, reset 178-185),
    through the drop-in module name, checked against the oracle."""
    import torch

    import madrona_basketball as mba
    from helpers import ENV_ARGS, make_oracle

    n = 128
    worlds = mba.SimpleGridworldSimulator(discrete_x=32, discrete_y=17, start_x=31.515 / 2.0, start_y=16.764 / 2.0,
                                          max_episode_length=39600, exec_mode=mba.madrona.ExecMode.CUDA,
                                          num_worlds=n, gpu_id=0)
    obs_t = worlds.observations_tensor().to_torch()
    act_t = worlds.action_tensor().to_torch()
    done_t = worlds.done_tensor().to_torch()
    rew_t = worlds.reward_tensor().to_torch()
    reset_t = worlds.reset_tensor().to_torch()
    assert obs_t.shape == (n, 2, 128) and obs_t.dtype == torch.float32 and obs_t.is_cuda
    assert act_t.shape == (n, 2, 6) and act_t.dtype == torch.int32
    assert done_t.shape == (n, 2) and rew_t.shape == (n, 2) and reset_t.shape == (n, 2, 1)
    orc = make_oracle(n)
    gen = torch.Generator(device="cpu").manual_seed(3)
    buckets = torch.tensor([2, 8, 3, 2, 2, 2])

    def env_step(trainee):   # env.py:147,155,167-169
        act_t[:, 0] = trainee.to(act_t.device)
        orc.tensor("action")[:, 0] = trainee.numpy()
        worlds.step()
        orc.step()
        return obs_t[:, 0].detach().clone(), rew_t[:, 0].detach().clone(), done_t[:, 0].detach().clone()

    for it in range(3):
        reset_t.fill_(1); orc.tensor("reset")[...] = 1          # env.py:179-185
        o, r, d = env_step(torch.zeros((n, 6), dtype=torch.int32))
        reset_t.fill_(0); orc.tensor("reset")[...] = 0
        assert (d == 1).all()
        for s in range(32):
            a = (torch.rand((n, 6), generator=gen) * buckets).to(torch.int32)
            o, r, d = env_step(a)
            assert np.array_equal(o.cpu().numpy().view(np.int32), orc.tensor("observations")[:, 0].view(np.int32))
            assert np.array_equal(r.cpu().numpy().view(np.int32), orc.tensor("reward")[:, 0].view(np.int32))
            assert np.array_equal(d.cpu().numpy(), orc.tensor("done")[:, 0])
