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


def _reference_viewer():
    """The reference's scripts/viewer.py (staged copy, pygame stubbed); None when the reference is not available."""
    import os
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "tools"))
    import stage_reference

    if not stage_reference.staged() and not stage_reference.stage(verbose=False):
        return None
    saved = list(sys.path)
    sys.path[:0] = stage_reference.script_path()
    try:
        import viewer
        return viewer
    finally:
        sys.path[:] = saved


def test_logs_load_through_the_reference_viewer_loader(tmp_path):
    """files written by TrajectoryRecorder go through the reference's OWN loader,
    ViewerClass.load_and_parse_log (scripts/viewer.py:1028-1100): episode breaks, shot events and counts come out."""
    import pytest

    viewer = _reference_viewer()
    if viewer is None:
        pytest.skip("baseline/_ref/ not staged and /root/reference absent")
    n, steps = 3, 700
    orc = make_oracle(n)
    rec = TrajectoryRecorder(_OracleSim(orc), worlds=slice(None), agent_idx=0)
    dones = []
    for s in range(steps):
        orc.fill_actions(1234, s)
        orc.step()
        rec.record()
        dones.append(orc.tensor("done")[:, 0].copy())
    path = tmp_path / "run.npz"
    rec.save_run(str(path), num_episodes=0)
    parsed = viewer.ViewerClass.load_and_parse_log(object(), str(path), event_to_track="shoot")   # `self` is unused by the loader
    assert parsed is not None, "the reference's loader rejected the file"
    assert (parsed["num_steps"], parsed["num_worlds"], parsed["num_agents"]) == (steps, n, 2)
    dones = np.array(dones)
    for w in range(n):
        ends = [b["end"] for b in parsed["episode_breaks"][w]]
        assert ends == [int(i) for i in np.flatnonzero(dones[1:, w] > 0) + 1]
        starts = [b["start"] for b in parsed["episode_breaks"][w]]
        assert starts == [0] + [e + 1 for e in ends[:-1]]
    assert parsed["total_episodes"] == max(int((dones[1:, w] > 0).sum()) for w in range(n)) > 0
    assert np.array_equal(parsed["episodes_completed_log"], np.cumsum(dones, axis=0))
    assert len(parsed["parsed_events"]) > 0                                   # random actions shoot
    for ev in parsed["parsed_events"]:
        assert ev["pos"].shape == (3,) and bool(ev["outcome"]) in (True, False)     # constants.py:31: shotIsGoingIn of that step

    # the per-episode file of ppo.py:108-121 (world 0 only, ppo.py's fixed hoop positions)
    orc2 = make_oracle(2)
    rec2 = TrajectoryRecorder(_OracleSim(orc2), worlds=slice(0, 1), agent_idx=0)
    for s in range(700):
        orc2.fill_actions(7, s)
        orc2.step()
        if rec2.record():
            break
    ep = tmp_path / "iter_100_episode.npz"
    rec2.save_episode(str(ep))
    parsed2 = viewer.ViewerClass.load_and_parse_log(object(), str(ep))
    assert parsed2 is not None and parsed2["num_worlds"] == 1 and parsed2["num_steps"] == s + 1
