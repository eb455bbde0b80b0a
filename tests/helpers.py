"""Shared helpers for the parity tests: run the CUDA engine and the CPU oracle on
the same inputs and compare complete world state bit for bit."""
from __future__ import annotations

import numpy as np

import oracle

ENV_ARGS = dict(discrete_x=32, discrete_y=17, start_x=31.515 / 2.0, start_y=16.764 / 2.0)  # scripts/env.py:22-35

ALL_COLUMNS = list(oracle.SLOTS)
# float32 tolerance stated by BASELINE.json's north_star ("within 1e-5 relative"); the
# deterministic math spec lets the tests demand far more: identical bits.
FLOAT_RTOL = 1e-5


def make_engine(num_worlds: int, flags: int = 1):
    from madrona_basketball_b200 import SimpleGridworldSimulator, madrona

    return SimpleGridworldSimulator(max_episode_length=39600, exec_mode=madrona.ExecMode.CUDA,
                                    num_worlds=num_worlds, gpu_id=0, _flags=flags, **ENV_ARGS)


def make_oracle(num_worlds: int, flags: int = 1, libm: bool = False):
    return oracle.Oracle(num_worlds, flags=flags, libm=libm, **ENV_ARGS)


def engine_state(sim) -> dict:
    return {k: sim.tensor(k).to_torch().cpu().numpy() for k in ALL_COLUMNS}


def assert_states_equal(eng: dict, orc: dict, where: str = "", columns=None):
    """Bit-exact comparison of every column (floats compared as their int32 bit patterns)."""
    for k in (columns or ALL_COLUMNS):
        a = np.ascontiguousarray(eng[k]).view(np.int32)
        b = np.ascontiguousarray(orc[k]).view(np.int32)
        assert a.shape == b.shape, (k, a.shape, b.shape)
        if not np.array_equal(a, b):
            bad = np.argwhere(a != b)
            first = tuple(bad[0])
            fa = np.ascontiguousarray(eng[k])[first]
            fb = np.ascontiguousarray(orc[k])[first]
            raise AssertionError(f"{where}: column '{k}' differs at {len(bad)} elements; first {first}: "
                                 f"engine={fa!r} oracle={fb!r}")


def numpy_fnv_checksum(state: dict) -> np.ndarray:
    """Independent numpy restatement of bb_checksum / orc_checksum (canonical column order)."""
    n = state["world_clock"].shape[0]
    h = np.full(n, 0xcbf29ce484222325, dtype=np.uint64)
    prime = np.uint64(0x100000001b3)
    for k in ALL_COLUMNS:
        arr = np.ascontiguousarray(state[k])
        if k in ("agent_entity_id", "ball_entity_id"):
            raise ValueError("entity columns need the raw 8-byte entities; use raw_words()")
        words = arr.view(np.uint32).reshape(n, -1)
        for j in range(words.shape[1]):
            h = (h ^ words[:, j].astype(np.uint64)) * prime
    return h
