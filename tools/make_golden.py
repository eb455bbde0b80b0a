#!/usr/bin/env python
"""Writes the committed fixtures under tests/golden/.

PROVENANCE: the reference cannot be compiled or imported here (its Madrona submodule is
absent) and ships no tests or golden vectors (F2), so these vectors come from
the CPU oracle (oracle/, deterministic math spec build), NOT from the reference.  They pin
the oracle and the CUDA engine against drift and let the GPU box check the engine without
trusting a freshly compiled oracle; they do not pin parity with Madrona itself ("parity
unpinned").

    python tools/make_golden.py            # rewrite tests/golden/*.npz
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
SEED = 1234
ENV_ARGS = dict(discrete_x=32, discrete_y=17, start_x=31.515 / 2.0, start_y=16.764 / 2.0)  # scripts/env.py:22-35

# columns small enough to store per step for a single world
TRACE_COLS = ["reward", "done", "agent_pos", "basketball_pos", "ball_physics", "agent_possession", "game_state",
              "action", "rng_count", "world_clock"]


def trace(n, steps, flags=1, actions="random", mutate=None, every=1):
    orc = oracle.Oracle(n, flags=flags, **ENV_ARGS)
    if mutate:
        mutate(orc, -1)
    sums = np.empty((steps, n), np.uint64)
    cols = {k: [] for k in TRACE_COLS} if n == 1 else {}
    for s in range(steps):
        if actions == "random":
            orc.fill_actions(SEED, s)
        elif actions == "zero0":  # scripts/run.py: agent 0 gets zeros, agent 1 keeps hardCodeDefense's action
            orc.tensor("action")[:, 0] = 0
        if mutate:
            mutate(orc, s)
        orc.step()
        sums[s] = orc.checksum()
        for k in cols:
            cols[k].append(orc.tensor(k).copy())
    out = {"checksum": sums[::every].copy(), "every": np.int64(every)}
    out.update({f"trace_{k}": np.stack(v) for k, v in cols.items()})
    out.update({f"final_{k}": v for k, v in orc.state().items() if n <= 64 or k != "observations"})
    return out


def main():
    os.makedirs(GOLD, exist_ok=True)
    np.savez_compressed(os.path.join(GOLD, "config1_n1_random_1000.npz"), **trace(1, 1000))
    np.savez_compressed(os.path.join(GOLD, "n1_zero_action_700.npz"), **trace(1, 700, actions="zero0"))
    np.savez_compressed(os.path.join(GOLD, "n64_random_300.npz"), **trace(64, 300, every=10))

    def one_v_one_off(orc, s):
        if s == -1:
            orc.tensor("game_state").view(np.int32)[::2, 13] = 0

    np.savez_compressed(os.path.join(GOLD, "n32_deadpaths_400.npz"), **trace(32, 400, flags=0, mutate=one_v_one_off, every=10))

    def collide(orc, s):
        if s >= 0:
            p = orc.tensor("agent_pos")
            p[:, 1] = p[:, 0] + np.array([0.05, 0.0, 0.0], np.float32)

    np.savez_compressed(os.path.join(GOLD, "n16_collision_stress_60.npz"), **trace(16, 60, mutate=collide))

    # libm call sites: spec outputs on fixed inputs, so device and host spec stay pinned
    rng = np.random.default_rng(7)
    x = {
        "sinf": rng.uniform(-8, 8, 4096), "cosf": rng.uniform(-8, 8, 4096), "atanf": rng.uniform(-64, 64, 4096),
        "acosf": rng.uniform(-1, 1, 4096), "expf": rng.uniform(-20, 2, 4096), "erff": rng.uniform(-5, 5, 4096),
        "atan2f": rng.uniform(-30, 30, 4096),
    }
    y = rng.uniform(-30, 30, 4096).astype(np.float32)
    out = {"atan2f_y": y}
    for fn, xs in x.items():
        xs = xs.astype(np.float32)
        out[f"{fn}_x"] = xs
        out[f"{fn}_out"] = oracle.math_eval(fn, xs, y if fn == "atan2f" else None)
    np.savez_compressed(os.path.join(GOLD, "math_spec.npz"), **out)
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


if __name__ == "__main__":
    main()
