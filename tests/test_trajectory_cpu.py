"""The `.npz` trajectory writer against the keys / shapes scripts/viewer.py:1031-1040 loads,
driven here by a CPU stand-in that exposes the oracle's columns through the reference's accessor names."""
import numpy as np
import torch

from helpers import make_oracle
from madrona_basketball_b200.trajectory import TrajectoryRecorder

_ACC = {"agent_pos_tensor": "agent_pos", "basketball_pos_tensor": "basketball_pos", "ball_velocity_tensor": "ball_velocity",
        "orientation_tensor": "orientation", "ball_physics_tensor": "ball_physics", "agent_possession_tensor": "agent_possession",
        "game_state_tensor": "game_state", "reward_tensor": "reward", "action_tensor": "action", "done_tensor": "done",
        "hoop_pos_tensor": "hoop_pos"}


class _View:
    def __init__(self, arr):
        self._t = torch.from_numpy(arr)     # shares memory with the oracle column

    def to_torch(self):
        return self._t


class _OracleSim:
    def __init__(self, orc):
        self.orc = orc

    def __getattr__(self, name):
        if name in _ACC:
            return lambda: _View(self.orc.tensor(_ACC[name]))
        raise AttributeError(name)


def test_episode_log_has_the_viewer_format(tmp_path):
    orc = make_oracle(4)
    rec = TrajectoryRecorder(_OracleSim(orc), worlds=slice(0, 1), agent_idx=0)
    ended = False
    for s in range(400):
        orc.fill_actions(1234, s)
        orc.step()
        ended = rec.record()
        if ended:
            break
    assert ended and len(rec.steps) == s + 1
    path = tmp_path / "iter_1_episode.npz"
    rec.save_episode(str(path))
    d = np.load(path, allow_pickle=True)
    S = s + 1
    want = {"agent_pos": (S, 1, 2, 3), "ball_pos": (S, 1, 1, 3), "ball_vel": (S, 1, 1, 3), "orientation": (S, 1, 2, 4),
            "ball_physics": (S, 1, 1, 7), "agent_possession": (S, 1, 2, 3), "game_state": (S, 1, 14), "rewards": (S, 1, 2),
            "actions": (S, 1, 2, 6), "done": (S, 1), "hoop_pos": (1, 2, 3)}
    assert set(d.files) == set(want)
    for k, shape in want.items():
        assert d[k].shape == shape, (k, d[k].shape)
    assert d["done"][-1, 0] == 1.0 and not d["done"][:-1].any()
    assert np.array_equal(d["hoop_pos"], [[[3.25, 8.5, 0.0], [28.75, 8.5, 0.0]]])     # ppo.py:117
    assert np.array_equal(d["agent_pos"][-1, 0], orc.tensor("agent_pos")[0])
    assert d["actions"].dtype == np.int32 and d["agent_pos"].dtype == np.float32
    assert not rec.steps


def test_run_log_covers_all_worlds(tmp_path):
    orc = make_oracle(3)
    rec = TrajectoryRecorder(_OracleSim(orc), worlds=slice(None), agent_idx=1)
    for s in range(10):
        orc.fill_actions(1234, s)
        orc.step()
        rec.record(done=torch.from_numpy(orc.tensor("done"))[:, 1])
    path = tmp_path / "run.npz"
    rec.save_run(str(path), num_episodes=2)
    d = np.load(path)
    assert d["agent_pos"].shape == (10, 3, 2, 3) and d["done"].shape == (10, 3) and int(d["num_episodes"]) == 2
    assert d["hoop_pos"].shape == (3, 2, 3)
