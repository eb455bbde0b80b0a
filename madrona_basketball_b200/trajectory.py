"""Trajectory `.npz` logs in the reference's playback format.

scripts/ppo.py:93-122 (world 0 of a training run, one file per recorded episode) and
scripts/infer.py:116-150 (every world, whole run) write one array per key, stacked over steps:

    agent_pos [S,W,2,3] f32   ball_pos [S,W,1,3] f32   ball_vel [S,W,1,3] f32   orientation [S,W,2,4] f32
    ball_physics [S,W,1,7] i32   agent_possession [S,W,2,3] i32   game_state [S,W,14] f32
    rewards [S,W,2] f32   actions [S,W,2,6] i32   done [S,W] f32   hoop_pos [W or 1,2,3] f32 (static)

which scripts/viewer.py:1031-1040 loads for playback on a machine with a display.  The recorder reads the
simulator's exported tensors (any object with the reference's `*_tensor().to_torch()` accessors), so logs of
B200 runs replay in the unmodified viewer.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np

# key -> accessor of bindings.cpp
_FIELDS = {
    "agent_pos": "agent_pos_tensor", "ball_pos": "basketball_pos_tensor", "ball_vel": "ball_velocity_tensor",
    "orientation": "orientation_tensor", "ball_physics": "ball_physics_tensor",
    "agent_possession": "agent_possession_tensor", "game_state": "game_state_tensor",
    "rewards": "reward_tensor", "actions": "action_tensor",
}
# ppo.py:117 writes these instead of the simulator's own hoop tensor ("corrected" positions for rendering)
PPO_HOOP_POS = np.array([[[3.25, 8.5, 0.0], [28.75, 8.5, 0.0]]])


class TrajectoryRecorder:
    def __init__(self, sim, worlds: Optional[slice] = slice(0, 1), agent_idx: int = 0):
        """worlds: slice of worlds to log (ppo.py logs [:1], infer.py logs all = slice(None))."""
        self.sim, self.worlds, self.agent_idx = sim, worlds, agent_idx
        self._views = {k: getattr(sim, acc)().to_torch() for k, acc in _FIELDS.items()}
        self._done = sim.done_tensor().to_torch()
        self.static = {"hoop_pos": sim.hoop_pos_tensor().to_torch().cpu().numpy().copy()}   # ppo.py:289-290
        self.steps: List[Dict[str, np.ndarray]] = []

    def record(self, done=None) -> bool:
        """Append the current step (call after sim.step()).  `done`: the trainee's done vector as env.step returns
        it (env.py:169); default: done_tensor[:, agent_idx].  Returns True when world 0's episode just ended."""
        w = self.worlds
        entry = {k: v[w].cpu().numpy().copy() for k, v in self._views.items()}
        d = self._done[:, self.agent_idx] if done is None else done
        entry["done"] = d[w].cpu().numpy().copy()
        self.steps.append(entry)
        return bool(entry["done"].reshape(-1)[0] > 0)

    def arrays(self) -> Dict[str, np.ndarray]:
        return {k: np.array([s[k] for s in self.steps]) for k in self.steps[0]} if self.steps else {}

    def save_episode(self, path: str) -> None:
        """ppo.py:108-121: the episode of world 0, with ppo.py's fixed hoop positions."""
        out = self.arrays()
        out["hoop_pos"] = PPO_HOOP_POS.copy()
        np.savez_compressed(path, **out)
        self.steps.clear()

    def save_run(self, path: str, num_episodes: int = 0) -> None:
        """infer.py:139-150: static data + every recorded step."""
        np.savez_compressed(path, **self.static, num_episodes=num_episodes, **self.arrays())
        self.steps.clear()
