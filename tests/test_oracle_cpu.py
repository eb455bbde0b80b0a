"""CPU tests of the oracle (oracle/): known answers, golden fixtures, reference quirks.

The reference ships no tests and cannot be built here (F1), so the known
answers are (a) the published Random123 Threefry vectors, (b) values the reference's own
scripts pin incidentally (hoop positions ppo.py:117, obs width, action buckets), (c) the
self-derived answers of SURVEY.md App. D, and (d) the committed fixtures of
tools/make_golden.py.  Parity with Madrona itself stays unpinned.
"""
import os

import numpy as np
import pytest

import oracle
from helpers import ENV_ARGS, make_oracle, numpy_fnv_checksum

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SEED = 1234


def gold(name):
    return np.load(os.path.join(GOLD, name))


# ---------------------------------------------------------------- RNG (mgr.cpp:155, sim.cpp:89, helper.cpp:8-11)
@pytest.mark.parametrize("key,ctr,out", [
    ((0x00000000, 0x00000000), (0x00000000, 0x00000000), (0x6b200159, 0x99ba4efe)),
    ((0xffffffff, 0xffffffff), (0xffffffff, 0xffffffff), (0x1cb996fc, 0xbb002be7)),
    ((0x13198a2e, 0x03707344), (0x243f6a88, 0x85a308d3), (0xc4923a9c, 0x483df7a0)),
])
def test_threefry2x32_20_random123_kat(key, ctr, out):
    assert oracle.threefry2x32(key[0], key[1], ctr[0], ctr[1]) == out


def test_world_rng_stream_first_draws():
    """key = split_i(initKey(0), 0, 0); u = (bits >> 8) * 2^-24."""
    want = [0.842314184, 0.477861702, 0.25783962, 0.495449126, 0.779361486, 0.150488496]
    got = [oracle.rng_uniform(0, i) for i in range(6)]
    assert np.allclose(got, want, rtol=0, atol=5e-9)
    assert all(0.0 <= u < 1.0 for u in got)


# ---------------------------------------------------------------- generateWorld (gen.cpp:13-214)
def test_initial_world_matches_reference_constants():
    o = make_oracle(3)
    hoops = o.tensor("hoop_pos")
    # scripts/ppo.py:117 hard-codes these for playback
    assert np.allclose(hoops, [[[3.25, 8.5, 0.0], [28.75, 8.5, 0.0]]] * 3, atol=1e-6)
    gs = o.tensor("game_state")
    gi = gs.view(np.int32)
    assert gi[0, 0] == 0 and gi[0, 1] == 1 and gi[0, 13] == 1      # no inbound, live ball, ONE_ON_ONE
    assert gs[0, 8] == 10.0 and gs[0, 9] == 24.0                     # TIME_PER_PERIOD, shot clock
    assert gi[:, 4].tolist() == [0, 5, 10] and gi[:, 6].tolist() == [1, 6, 11]  # hoop entity ids, 5 entities per world
    assert o.tensor("agent_entity_id").shape == (3, 2) and o.tensor("ball_entity_id").shape == (3, 1)
    poss = o.tensor("agent_possession")
    assert poss[0].tolist() == [[1, 2, 2], [0, 2147483647, 2]]
    assert o.tensor("ball_grabbed")[1].tolist() == [[1, 8]]          # held by agent 0 of world 1 (id 5*1+3)
    # identical RNG stream in every world (sim.cpp:89): same spawn everywhere
    p = o.tensor("agent_pos")
    assert np.array_equal(p[0], p[1]) and np.array_equal(p[0], p[2])
    # spawn: offence = start + (-5 + 10 u0, -5 + 10 u1); defender at radius 8 (helper.cpp:121-136)
    u = [oracle.rng_uniform(0, i) for i in range(3)]
    assert np.allclose(p[0, 0, :2], [15.7575 - 5 + 10 * u[0], 8.382 - 5 + 10 * u[1]], atol=1e-5)
    d = p[0, 1, :2] - p[0, 0, :2]
    want = 8.0 * np.array([np.cos(2 * np.pi * u[2]), np.sin(2 * np.pi * u[2])])
    want_clamped = np.clip(p[0, 0, :2] + want, [0, 0], [32, 17]) - p[0, 0, :2]
    assert np.allclose(d, want_clamped, atol=1e-4)
    assert o.tensor("rng_count").tolist() == [3, 3, 3]
    assert o.tensor("observations").shape == (3, 2, 128) and not o.tensor("observations").any()  # filled by the first step only
    # team colours before the first reset (constants.hpp:18-19), quirk B11
    team = o.tensor("agent_team")
    assert team[0, 1, 1:4].view(np.float32).tolist() == [128.0, 0.0, 128.0]


def test_shapes_and_dtypes_match_manager_accessors():
    """mgr.cpp:317-445 declared shapes / dtypes."""
    n = 5
    o = make_oracle(n)
    want = {
        "reset": ((n, 2, 1), np.int32), "game_state": ((n, 14), np.float32), "action": ((n, 2, 6), np.int32),
        "action_mask": ((n, 2, 4), np.int32), "agent_pos": ((n, 2, 3), np.float32),
        "observations": ((n, 2, 128), np.float32), "orientation": ((n, 2, 4), np.float32),
        "reward": ((n, 2), np.float32), "done": ((n, 2), np.float32), "agent_possession": ((n, 2, 3), np.int32),
        "agent_entity_id": ((n, 2), np.int32), "agent_team": ((n, 2, 5), np.int32), "agent_stats": ((n, 2, 2), np.int32),
        "basketball_pos": ((n, 1, 3), np.float32), "ball_physics": ((n, 1, 7), np.int32),
        "ball_grabbed": ((n, 1, 2), np.int32), "ball_entity_id": ((n, 1), np.int32),
        "ball_velocity": ((n, 1, 3), np.float32), "hoop_pos": ((n, 2, 3), np.float32),
    }
    for k, (shape, dt) in want.items():
        t = o.tensor(k)
        assert t.shape == shape and t.dtype == dt, k


# ---------------------------------------------------------------- step semantics
def test_game_clock_expires_after_620_steps():
    """clockSystem (game.cpp:992-1030): 10 s at 1/62 s per step in float32 -> reset on step 620 (App. D)."""
    o = make_oracle(1)
    # park both agents far apart with the ball held so nothing else ends the episode
    first_reset = None
    for s in range(640):
        o.tensor("action")[...] = 0
        o.tensor("agent_pos")[0, 0] = [10.0, 8.0, 0.0]
        o.tensor("agent_pos")[0, 1] = [25.0, 8.0, 0.0]
        o.step()
        if o.tensor("done")[0, 0] == 1.0 and first_reset is None:
            first_reset = s + 1
            assert o.tensor("reward")[0, 0] >= 10.0   # offence gets +10 at expiry (game.cpp:1009-1021)
    assert first_reset == 620


def test_python_reset_only_forces_done(tmp_path):
    """Quirk B1 (game.cpp:978-984): the Reset flag sets done=1 / CurStep=0 but does not regenerate the world."""
    o = make_oracle(2)
    for s in range(5):
        o.fill_actions(SEED, s)
        o.step()
    before_rng = o.tensor("rng_count").copy()
    o.tensor("reset")[...] = 1
    o.tensor("action")[...] = 0
    o.step()
    assert (o.tensor("done") == 1.0).all()
    assert (o.tensor("cur_step") == 0).all()
    assert np.array_equal(o.tensor("rng_count"), before_rng)       # no resetWorld -> no new draws
    o.tensor("reset")[...] = 0
    o.step()
    assert (o.tensor("done") == 0.0).all() and (o.tensor("cur_step") == 1).all()


def test_action_mask_tag_override_and_defense_writes_actions():
    """a2: can_grab = can_pass = 0 (game.cpp:525-528).  a17 / quirk B12: the defender's Action row is rewritten."""
    o = make_oracle(4)
    o.fill_actions(SEED, 0)
    o.step()
    m = o.tensor("action_mask")
    assert (m[:, :, 1] == 0).all() and (m[:, :, 2] == 0).all()
    a = o.tensor("action")
    assert (a[:, 1, 3] == 1).all()                                   # defender grab = 1
    assert (a[:, 0, 0] == 0).all()                                   # team in possession: move = 0


def test_observation_layout_spot_checks():
    """SURVEY App. A indices (game.cpp:1256-1460)."""
    o = make_oracle(1)
    o.fill_actions(SEED, 0)
    o.tensor("action")[0, 0, 5] = 0   # do not shoot on the first step
    o.step()
    obs = o.tensor("observations")[0]
    gs = o.tensor("game_state")[0]
    pos = o.tensor("agent_pos")[0]
    assert obs[0, 0] == gs[8] and obs[1, 0] == gs[8]                 # game clock
    assert np.array_equal(obs[0, 23:26], pos[0]) and np.array_equal(obs[1, 23:26], pos[1])
    assert np.array_equal(obs[0, 61:64], pos[1]) and np.array_equal(obs[1, 61:64], pos[0])
    assert np.array_equal(obs[0, 17:20], o.tensor("hoop_pos")[0, 1])  # agent 0 attacks hoop 1
    assert np.array_equal(obs[1, 17:20], o.tensor("hoop_pos")[0, 0])
    assert np.array_equal(obs[0, 30:34], o.tensor("orientation")[0, 0])
    assert obs[0, 60] == 1.0 and obs[1, 60] == 0.0                   # hasBall
    assert obs[0, 99] == 1.0 and obs[0, 100] == 0.0                  # ball-holder one-hot
    assert obs[0, 16] == np.float32(o.tensor("ball_physics")[0, 0, 2]) == 0.0  # lastTouchedByTeamID: holder's team (a11)
    assert not obs[:, 103:].any() and not obs[:, 26:30].any()
    assert abs(np.linalg.norm(obs[0, 64:67]) - 1.0) < 1e-6 and abs(obs[0, 67] - np.linalg.norm(pos[1] - pos[0])) < 1e-5


def test_collision_triggers_reset_one_step_late():
    """Quirk B2: a16 runs after a14, so the reset (done=1) lands on the step after the +-10 rewards."""
    o = make_oracle(1)
    o.tensor("action")[...] = 0
    p = o.tensor("agent_pos")
    p[0, 1] = p[0, 0] + np.array([0.05, 0, 0], np.float32)
    o.step()
    r = o.tensor("reward")[0].copy()
    assert r[0] <= -10.0 + 1.5 and r[1] >= 10.0 - 1.5                # offence -10, defence +10 (plus shaping terms)
    assert o.tensor("done")[0].tolist() == [0.0, 0.0] and o.tensor("world_clock")[0] == 1
    o.tensor("action")[...] = 0
    o.step()
    assert o.tensor("done")[0].tolist() == [1.0, 1.0] and o.tensor("rng_count")[0] == 6
    # quirk B11: team 1 colour changes at the first reset (gen.cpp:258,283)
    assert o.tensor("agent_team")[0, 1, 1:4].view(np.float32).tolist() == [255.0, 0.0, 100.0]


# ---------------------------------------------------------------- determinism / fixtures
def test_thread_count_does_not_change_results():
    a, b = make_oracle(257), make_oracle(257)
    for s in range(60):
        a.fill_actions(SEED, s); b.fill_actions(SEED, s)
        a.step(1); b.step(7)
    assert np.array_equal(a.checksum(), b.checksum())
    c = make_oracle(257)
    c.run_synthetic(SEED, 0, 60, 0, 5)
    assert np.array_equal(a.checksum(), c.checksum())


def test_checksum_is_fnv1a_over_canonical_columns():
    o = make_oracle(9)
    for s in range(12):
        o.fill_actions(SEED, s)
        o.step()
    st = o.state()
    # entity columns are 8-byte {gen, id} underneath; the exported view shows only half (quirk B10)
    st = dict(st)
    n = 9
    ae = np.zeros((n, 4), np.int32); ae[:, 1] = 5 * np.arange(n) + 3; ae[:, 3] = 5 * np.arange(n) + 4
    be = np.zeros((n, 2), np.int32); be[:, 1] = 5 * np.arange(n) + 2
    st["agent_entity_id"], st["ball_entity_id"] = ae, be
    h = np.full(n, 0xcbf29ce484222325, dtype=np.uint64)
    with np.errstate(over="ignore"):
        for k in oracle.SLOTS:
            words = np.ascontiguousarray(st[k]).view(np.uint32).reshape(n, -1)
            for j in range(words.shape[1]):
                h = (h ^ words[:, j].astype(np.uint64)) * np.uint64(0x100000001b3)
    assert np.array_equal(h, o.checksum())


def _replay(n, steps, flags=1, actions="random", mutate=None):
    o = make_oracle(n, flags)
    if mutate:
        mutate(o, -1)
    sums = []
    for s in range(steps):
        if actions == "random":
            o.fill_actions(SEED, s)
        else:
            o.tensor("action")[:, 0] = 0
        if mutate:
            mutate(o, s)
        o.step()
        sums.append(o.checksum())
    return o, np.stack(sums)


def test_golden_config1_single_world_1000_steps():
    """BASELINE.json configs[0]: per-step state checksum trace."""
    g = gold("config1_n1_random_1000.npz")
    o, sums = _replay(1, 1000)
    assert np.array_equal(sums, g["checksum"])
    for k, v in o.state().items():
        assert np.array_equal(v.view(np.int32), g[f"final_{k}"].view(np.int32)), k
    # the trace exercises the path: shots, resets, rewards of every kind
    assert g["trace_done"].sum() > 10 and g["trace_rng_count"][-1, 0] > 30
    assert (g["trace_reward"] <= -100).any()            # out of bounds
    assert (g["trace_ball_physics"][:, 0, 0, 0] == 1).any()  # ball in flight


def test_golden_zero_action_run():
    g = gold("n1_zero_action_700.npz")
    _, sums = _replay(1, 700, actions="zero0")
    assert np.array_equal(sums, g["checksum"])


def test_golden_many_worlds_and_dead_paths():
    g = gold("n64_random_300.npz")
    _, sums = _replay(64, 300)
    assert np.array_equal(sums[::10], g["checksum"])
    assert len(np.unique(g["checksum"][-1])) > 32       # worlds diverged through their actions

    def off(o, s):
        if s == -1:
            o.tensor("game_state").view(np.int32)[::2, 13] = 0
    g = gold("n32_deadpaths_400.npz")
    o, sums = _replay(32, 400, flags=0, mutate=off)
    assert np.array_equal(sums[::10], g["checksum"])
    assert (o.tensor("agent_possession")[:, 1, 0] == 1).any() or (g["final_game_state"][::2, 11] > 0).any()


def test_golden_collision_stress():
    def collide(o, s):
        if s >= 0:
            p = o.tensor("agent_pos")
            p[:, 1] = p[:, 0] + np.array([0.05, 0.0, 0.0], np.float32)
    g = gold("n16_collision_stress_60.npz")
    o, sums = _replay(16, 60, mutate=collide)
    assert np.array_equal(sums, g["checksum"])
    assert o.tensor("rng_count").min() >= 3 + 3 * 25


# ---------------------------------------------------------------- arithmetic spec vs glibc
def test_math_spec_golden_and_glibc_deviation():
    """include/bb_math.h on the host reproduces the committed outputs bit for bit and stays within
    1 ulp of glibc's float functions (the arithmetic the reference's CPU build would use)."""
    g = gold("math_spec.npz")
    for fn in oracle.MATH_FNS:
        x = g[f"{fn}_x"]
        y = g["atan2f_y"] if fn == "atan2f" else None
        spec = oracle.math_eval(fn, x, y)
        assert np.array_equal(spec.view(np.int32), g[f"{fn}_out"].view(np.int32)), fn
        libm = oracle.math_eval(fn, x, y, libm=True)
        ulp = np.abs(spec.view(np.int32).astype(np.int64) - libm.view(np.int32).astype(np.int64))
        assert ulp.max() <= 1, (fn, int(ulp.max()))
        # the spec is the correctly rounded result (tools/check_math.c: 0 mismatches vs (float)f((double)x));
        # glibc's float functions are not: up to 25 % of atan2f results with x < 0 are 1 ulp off
        assert (ulp > 0).mean() < 0.30, fn


def test_libm_build_agrees_on_discrete_state_for_a_while():
    """The glibc build and the spec build differ only in last-ulp libm results; over a short horizon
    all discrete state agrees and floats stay within the 1e-5 relative budget of north_star."""
    a, b = make_oracle(64), make_oracle(64, libm=True)
    for s in range(100):
        a.fill_actions(SEED, s); b.fill_actions(SEED, s)
        a.step(); b.step()
    for k in ["done", "agent_possession", "ball_physics", "ball_grabbed", "world_clock", "rng_count", "action_mask"]:
        assert np.array_equal(a.tensor(k), b.tensor(k)), k
    for k in ["agent_pos", "basketball_pos", "orientation", "agent_velocity"]:
        assert np.allclose(a.tensor(k), b.tensor(k), rtol=1e-5, atol=1e-6), k


# ---------------------------------------------------------------- vectors generated by the reference's own Python
def test_oracle_against_reference_python_fixtures():
    """tools/make_golden_ref.py ran the reference's scripts/viewer.py `rotate_vec` (its Python mirror of Madrona's
    Quat::rotateVec) and read src/constants.py; the oracle's primitive / constants must agree.  This is the one
    piece of the Madrona boundary the reference itself pins."""
    g = gold("reference_python.npz")
    q, v, want = g["rotate_q"], g["rotate_v"], g["rotate_out"]
    for i in range(len(q)):
        got = oracle.rotate_vec(q[i], v[i])
        scale = max(1.0, float(np.abs(want[i]).max()), float(np.dot(q[i], q[i])) * float(np.abs(v[i]).max()))
        assert np.allclose(got, want[i], rtol=0, atol=4e-6 * scale), (i, got, want[i])   # float32 vs the float64 formula
    assert np.allclose(oracle.rotate_vec([1, 0, 0, 0], [0, 1, 0]), [0, 1, 0])
    c = oracle.court_constants()
    for name, val in zip(g["const_names"], g["const_values"]):
        assert abs(c[str(name)] - float(val)) <= 1e-6 * max(1.0, abs(float(val))), (name, c[str(name)], val)
    assert int(g["scalar_NUM_AGENTS"]) == 2 and int(g["scalar_NUM_HOOPS"]) == 2 and float(g["scalar_SIMULATION_HZ"]) == 62.0
    assert abs(float(g["scalar_AGENT_SHOULDER_WIDTH"]) - 0.429) < 1e-12 and float(g["scalar_HOOP_SCORE_ZONE_SIZE"]) == 0.1
