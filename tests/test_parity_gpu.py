"""GPU parity: CUDA engine (through the C ABI / Python mirror) vs the CPU oracle.

Bar (BASELINE.json north_star): integer / discrete state bit-exact, float32 within 1e-5
relative over 1000 steps.  Because both sides share the arithmetic spec include/bb_math.h
and the kernel is built with -fmad=false, these tests demand identical bits for floats too.
"""
import numpy as np
import pytest

from helpers import (ALL_COLUMNS, assert_states_equal, engine_state, make_engine, make_oracle)

pytestmark = pytest.mark.gpu

SEED = 1234


def _run_lockstep(n, steps, flags=1, check_every=1, mutate=None):
    import torch

    sim = make_engine(n, flags)
    orc = make_oracle(n, flags)
    assert_states_equal(engine_state(sim), orc.state(), f"N={n} after construction")
    for s in range(steps):
        sim.fill_actions(SEED, s)
        orc.fill_actions(SEED, s)
        if mutate is not None:
            mutate(sim, orc, s)
        sim.step()
        orc.step(num_threads=8 if n > 256 else 1)
        if s % check_every == 0 or s == steps - 1:
            torch.cuda.synchronize()
            assert_states_equal(engine_state(sim), orc.state(), f"N={n} step {s}")
    return sim, orc


def test_config1_single_world_1000_steps_bit_exact():
    """BASELINE.json configs[0]: num_worlds=1, random actions (fixed seed), 1000 steps, every step compared."""
    sim, orc = _run_lockstep(1, 1000)
    # the episode must actually have exercised resets (RNG draws beyond the 3 of generateWorld)
    assert int(orc.tensor("rng_count")[0]) > 3


def test_many_worlds_random_actions():
    sim, orc = _run_lockstep(4096, 300, check_every=25)
    rng = orc.tensor("rng_count")
    assert rng.max() > 3 and len(np.unique(rng)) > 1  # worlds diverged and reset


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 127, 129, 1000])
def test_ragged_world_counts(n):
    """Tail handling: world counts that are not multiples of the warp / block size."""
    _run_lockstep(n, 40, check_every=13)


@pytest.mark.parametrize("shape", [dict(BB_BLOCK_THREADS="128", BB_PHASE_SYNC="1", BB_OBS_ROUND="8"),
                                   dict(BB_BLOCK_THREADS="128", BB_PHASE_SYNC="0", BB_OBS_ROUND="16"),
                                   dict(BB_BLOCK_THREADS="64", BB_OBS_ROUND="8"),
                                   dict(BB_BLOCK_THREADS="256", BB_PHASE_SYNC="1"),
                                   dict(BB_PDL="0"),
                                   dict(BB_LONE="1"),
                                   dict(BB_PDL_EARLY="0", BB_PREFETCH_SELF="0"),
                                   dict(BB_BLOCK_THREADS="128", BB_PHASE_SYNC="1", BB_OBS_ROUND="8", BB_PDL_SYNC="1")])
def test_every_launch_shape_is_bit_exact(monkeypatch, shape):
    """The launch shapes the engine picks by world count (one-warp blocks / 64 threads / 128 threads with phase barriers and
    two observation rounds / the 512-thread instantiation) and its tuning knobs (no programmatic launches, the lone-warps
    shape, no early release / self-prefetch, all three with the phase-barrier shape), forced here at a size the oracle checks
    every column of (observations included): ragged count, dead-path flags off, 60 steps."""
    for k, v in shape.items():
        monkeypatch.setenv(k, v)
    _run_lockstep(1000 + 37, 60, check_every=7)
    _run_lockstep(333, 40, flags=0, check_every=5)


def test_zero_actions_like_run_py():
    """scripts/run.py drives zero actions for agent 0 only; the hard-coded defender chases (a17)."""
    import torch

    n = 64
    sim, orc = make_engine(n), make_oracle(n)
    act_e = sim.action_tensor().to_torch()
    act_o = orc.tensor("action")
    for s in range(700):  # crosses the 620-step game-clock expiry and several collisions
        act_e[:, 0] = 0
        act_o[:, 0] = 0
        sim.step()
        orc.step()
        if s % 50 == 0 or s > 600:
            torch.cuda.synchronize()
            assert_states_equal(engine_state(sim), orc.state(), f"zero-action step {s}")
    assert orc.tensor("rng_count").max() > 3


def test_env_reset_semantics():
    """EnvWrapper.reset (scripts/env.py:178-185): resets.fill_(1); step; resets.fill_(0) only forces done=1 (quirk B1)."""
    import torch

    n = 33
    sim, orc = make_engine(n), make_oracle(n)
    for s in range(100):
        if s % 33 == 0:
            sim.reset_tensor().to_torch().fill_(1)
            orc.tensor("reset")[...] = 1
            sim.action_tensor().to_torch()[:, 0] = 0
            orc.tensor("action")[:, 0] = 0
        else:
            sim.fill_actions(SEED, s)
            orc.fill_actions(SEED, s)
        sim.step()
        orc.step()
        torch.cuda.synchronize()
        if s % 33 == 0:
            done = sim.done_tensor().to_torch().cpu().numpy()
            assert (done == 1.0).all()
            sim.reset_tensor().to_torch().fill_(0)
            orc.tensor("reset")[...] = 0
        assert_states_equal(engine_state(sim), orc.state(), f"reset-semantics step {s}")


def test_set_action_and_trigger_reset():
    import torch

    n = 8
    sim, orc = make_engine(n), make_oracle(n)
    sim.set_action(3, 1, 1, 5, 2, 0, 0, 1)
    orc.tensor("action")[3, 1] = [1, 5, 2, 0, 0, 1]
    sim.trigger_reset(5)
    orc.tensor("reset")[5] = 1
    sim.set_action(99, 0, 1, 1, 1, 1, 1, 1)  # invalid: prints the reference's error line, ignored
    sim.trigger_reset(-1)                     # invalid: silently ignored
    sim.step()
    orc.step()
    torch.cuda.synchronize()
    assert_states_equal(engine_state(sim), orc.state(), "set_action/trigger_reset")
    assert sim.done_tensor().to_torch()[5].tolist() == [1.0, 1.0]


def test_stress_forced_collisions():
    """BASELINE.json configs[4] (small): agent 1 teleported next to agent 0 every step through the
    writable exported view, so agentCollision fires each step and the world resets every 2nd step."""
    def mutate(sim, orc, s):
        pe = sim.agent_pos_tensor().to_torch()
        pe[:, 1] = pe[:, 0] + pe.new_tensor([0.05, 0.0, 0.0])
        po = orc.tensor("agent_pos")
        po[:, 1] = po[:, 0] + np.array([0.05, 0.0, 0.0], dtype=np.float32)

    sim, orc = _run_lockstep(512, 60, check_every=5, mutate=mutate)
    assert orc.tensor("rng_count").min() >= 3 + 3 * 25  # >= 25 resets in 60 steps


def test_dead_paths_grab_pass_inbound():
    """Test hook: with the FOR-TAG mask override off and 1v1 disabled in some worlds the grab /
    pass / steal / out-of-bounds-turnover / inbound-violation systems (a4, a5, a9, a10, a13) run."""
    import torch

    n = 2048
    sim, orc = make_engine(n, flags=0), make_oracle(n, flags=0)
    gs_e = sim.game_state_tensor().to_torch().view(torch.int32)
    gs_o = orc.tensor("game_state").view(np.int32)
    gs_e[::2, 13] = 0   # isOneOnOne = 0 in even worlds
    gs_o[::2, 13] = 0
    for s in range(600):
        sim.fill_actions(SEED, s)
        orc.fill_actions(SEED, s)
        sim.step()
        orc.step(num_threads=8)
        if s % 20 == 0 or s == 599:
            torch.cuda.synchronize()
            assert_states_equal(engine_state(sim), orc.state(), f"dead-path step {s}")
    st = orc.state()
    assert (st["game_state"].view(np.int32)[::2, 0] == 1).any() or st["game_state"][::2, 11].max() > 0  # inbounds happened


def test_checksum_matches_oracle():
    import torch

    n = 777
    sim, orc = make_engine(n), make_oracle(n)
    for s in range(30):
        sim.fill_actions(SEED, s)
        orc.fill_actions(SEED, s)
        sim.step()
        orc.step()
    ce = sim.checksum().cpu().numpy().view(np.uint64)
    co = orc.checksum()
    assert np.array_equal(ce, co)


def test_step_with_actions_and_host_step_equivalent():
    import torch

    n = 300
    a, b, c = make_engine(n), make_engine(n), make_engine(n)
    buf = torch.empty((n, 2, 6), dtype=torch.int32, device="cuda")
    rew = np.empty((n, 2), np.float32); done = np.empty((n, 2), np.float32); obs = np.empty((n, 2, 128), np.float32)
    for s in range(50):
        a.fill_actions(SEED, s); a.step()
        b.fill_actions(SEED, s, out=buf); b.step_with_actions(buf)
        torch.cuda.synchronize()
        c.step_host(buf.cpu().numpy(), rew, done, obs)
    torch.cuda.synchronize()
    sa, sb, sc = engine_state(a), engine_state(b), engine_state(c)
    assert_states_equal(sb, sa, "step_with_actions")
    assert_states_equal(sc, sa, "step_host")
    assert np.array_equal(obs.view(np.int32), sa["observations"].view(np.int32))
    assert np.array_equal(rew.view(np.int32), sa["reward"].view(np.int32))
    assert np.array_equal(done.view(np.int32), sa["done"].view(np.int32))


def test_world_partition_invariance():
    """Sharding property: a shard of worlds [k, k+m) evolves exactly like those worlds
    inside the full simulation, because the action stream is keyed by global world index."""
    import torch

    n, k, m = 256, 96, 64
    full, part = make_engine(n), make_engine(m)
    for s in range(80):
        full.fill_actions(SEED, s)
        part.fill_actions(SEED, s, world_offset=k)
        full.step(); part.step()
    torch.cuda.synchronize()
    sf, sp = engine_state(full), engine_state(part)
    ids = {"agent_entity_id", "ball_entity_id", "agent_possession", "agent_team", "ball_physics", "ball_grabbed", "game_state"}
    for col in ALL_COLUMNS:
        if col in ids:  # entity ids are instance-local (as in the reference): compare the id-free columns
            continue
        assert np.array_equal(sf[col][k:k + m].view(np.int32), sp[col].view(np.int32)), col


@pytest.mark.parametrize("n", [1, 50, 300, 8192, 40000])
def test_persistent_server_steps_match_launch_per_step(n):
    """bb_serve_*: the resident kernel fed through the host doorbell evolves the worlds exactly like
    bb_step launches, mirrors reward / done / observations into pinned host memory, survives its own
    idle exit, and leaves the device columns consistent for ordinary steps afterwards."""
    import time

    import torch

    a, b = make_engine(n), make_engine(n)
    act = torch.empty((n, 2, 6), dtype=torch.int32, device="cuda")
    hact = torch.empty((n, 2, 6), dtype=torch.int32).pin_memory()
    hrew = torch.empty((n, 2)).pin_memory(); hdone = torch.empty((n, 2)).pin_memory()
    hobs = torch.empty((n, 2, 128)).pin_memory() if n <= 8192 else None
    b.serve_begin()
    steps = 60 if n <= 8192 else 12
    for s in range(steps):
        a.fill_actions(SEED, s, out=act); a.step_with_actions(act)
        torch.cuda.current_stream().synchronize()
        hact.copy_(act)
        b.serve_step(hact.numpy(), hrew.numpy(), hdone.numpy(), hobs.numpy() if hobs is not None else None)
        assert np.array_equal(hrew.numpy().view(np.int32), a.reward_tensor().to_torch().cpu().numpy().view(np.int32)), s
        assert np.array_equal(hdone.numpy(), a.done_tensor().to_torch().cpu().numpy()), s
        if hobs is not None and s % 7 == 0:
            assert np.array_equal(hobs.numpy().view(np.int32), a.observations_tensor().to_torch().cpu().numpy().view(np.int32)), s
    b.serve_end()
    torch.cuda.synchronize()
    assert_states_equal(engine_state(b), engine_state(a), f"server N={n}")
    # ordinary launches continue from the served state
    a.fill_actions(SEED, 99, out=act); a.step_with_actions(act); b.step_with_actions(act)
    torch.cuda.synchronize()
    assert_states_equal(engine_state(b), engine_state(a), f"after server N={n}")


@pytest.mark.parametrize("n", [37, 8192])
def test_packed_u8_action_rows_match_int32_rows(n):
    """bb_step_host_u8 / bb_serve_step_u8 (declared extension): the same actions as bytes evolve the worlds exactly like the
    int32 rows; the Action column ends up as the int32 rows the reference exports; ordinary steps refuse to run while the
    server is resident."""
    import torch

    a, b, c = make_engine(n), make_engine(n), make_engine(n)
    act = torch.empty((n, 2, 6), dtype=torch.int32, device="cuda")
    hu8 = torch.empty((n, 2, 6), dtype=torch.uint8).pin_memory()
    hrew = torch.empty((n, 2)).pin_memory(); hdone = torch.empty((n, 2)).pin_memory()
    hrew2 = torch.empty((n, 2)).pin_memory(); hdone2 = torch.empty((n, 2)).pin_memory()
    c.serve_begin()
    with pytest.raises(RuntimeError, match="persistent server"):
        c.step()
    for s in range(40):
        a.fill_actions(SEED, s, out=act); a.step_with_actions(act)
        torch.cuda.current_stream().synchronize()
        hu8.copy_(act.to(torch.uint8))
        b.step_host(hu8.numpy(), hrew.numpy(), hdone.numpy())
        c.serve_step(hu8.numpy(), hrew2.numpy(), hdone2.numpy())
        ref = a.reward_tensor().to_torch().cpu().numpy().view(np.int32)
        assert np.array_equal(hrew.numpy().view(np.int32), ref), s
        assert np.array_equal(hrew2.numpy().view(np.int32), ref), s
        assert np.array_equal(hdone.numpy(), a.done_tensor().to_torch().cpu().numpy()), s
    c.serve_end()
    torch.cuda.synchronize()
    assert_states_equal(engine_state(b), engine_state(a), "step_host_u8")
    assert_states_equal(engine_state(c), engine_state(a), "serve_step_u8")
    with pytest.raises(RuntimeError, match="pinned"):
        b.step_host(np.zeros((n, 2, 6), np.uint8))


@pytest.mark.parametrize("serve_threads,binding", [("32", "fastcall"), ("128", "fastcall"), ("32", "ctypes")])
def test_server_buffer_combinations_and_bound_calls(monkeypatch, serve_threads, binding):
    """The doorbell is one word carrying the slot of a device-side table of buffer combinations {actions, reward, done, obs}:
    a ring of action buffers LARGER than the 64-slot table (slots are re-used), int32 and packed rows in turn, two sets of
    result buffers, steps without result buffers, every call through bind_serve_step (the bound C-ABI call of
    csrc/bb_fastcall.c, and the ctypes closure it falls back to) == the same steps as ordinary launches.  Both block sizes
    of the server."""
    import torch

    monkeypatch.setenv("BB_SERVE_THREADS", serve_threads)
    from madrona_basketball_b200 import manager

    if binding == "ctypes":       # the closure over the ctypes function that bind_serve_step falls back to without the glue module
        monkeypatch.setattr(manager, "_FASTCALL", None)
    else:
        assert manager._fastcall_module() is not None, "csrc/bb_fastcall.c is not built (python -m madrona_basketball_b200.build)"
    n, ring, steps = 300, 70, 160
    a, b = make_engine(n), make_engine(n)
    act = torch.empty((n, 2, 6), dtype=torch.int32, device="cuda")
    h32 = [torch.empty((n, 2, 6), dtype=torch.int32).pin_memory() for _ in range(ring)]
    h8 = [torch.empty((n, 2, 6), dtype=torch.uint8).pin_memory() for _ in range(ring)]
    res = [(torch.empty((n, 2)).pin_memory(), torch.empty((n, 2)).pin_memory()) for _ in range(2)]
    b.serve_begin()
    for s in range(steps):
        a.fill_actions(SEED, s, out=act); a.step_with_actions(act)
        torch.cuda.current_stream().synchronize()
        i, packed, (rew, done) = s % ring, (s // 3) % 2 == 1, res[s % 2]
        buf = h8[i] if packed else h32[i]
        buf.copy_(act.to(buf.dtype))
        if s % 5 == 4:
            b.bind_serve_step(buf.numpy())()                       # no result buffers: another combination
        else:
            b.bind_serve_step(buf.numpy(), rew.numpy(), done.numpy())()
            assert np.array_equal(rew.numpy().view(np.int32), a.reward_tensor().to_torch().cpu().numpy().view(np.int32)), s
            assert np.array_equal(done.numpy(), a.done_tensor().to_torch().cpu().numpy()), s
    b.serve_end()
    torch.cuda.synchronize()
    assert_states_equal(engine_state(b), engine_state(a), f"server combinations, {serve_threads}-thread blocks")


def test_persistent_server_idle_exit_and_restart(monkeypatch):
    import time

    import torch

    monkeypatch.setenv("BB_SERVE_IDLE_MS", "50")
    n = 64
    a, b = make_engine(n), make_engine(n)
    act = torch.empty((n, 2, 6), dtype=torch.int32, device="cuda")
    hact = torch.empty((n, 2, 6), dtype=torch.int32).pin_memory()
    hrew = torch.empty((n, 2)).pin_memory(); hdone = torch.empty((n, 2)).pin_memory()
    b.serve_begin()
    for s in range(6):
        a.fill_actions(SEED, s, out=act); a.step_with_actions(act)
        torch.cuda.current_stream().synchronize()
        hact.copy_(act)
        if s == 3:
            time.sleep(0.3)                 # the server leaves after 50 ms without a doorbell ...
            torch.cuda.synchronize()        # ... so a device-wide sync returns
        b.serve_step(hact.numpy(), hrew.numpy(), hdone.numpy())   # ... and the next step restarts it
        assert np.array_equal(hrew.numpy().view(np.int32), a.reward_tensor().to_torch().cpu().numpy().view(np.int32)), s
    b.serve_end()
    with pytest.raises(RuntimeError, match="pinned"):
        b.serve_begin()
        try:
            b.serve_step(np.zeros((n, 2, 6), np.int32))
        finally:
            b.serve_end()


def test_full_size_65536_worlds_checksums():
    """BASELINE.json configs[2] size (65 536 worlds, here on one GPU): engine vs oracle per-world checksums."""
    import torch

    n = 65536
    sim, orc = make_engine(n), make_oracle(n)
    for s in range(150):
        sim.fill_actions(SEED, s)
        sim.step()
    orc.run_synthetic(SEED, 0, 150, 0, 16)
    ce = sim.checksum().cpu().numpy().view(np.uint64)
    assert np.array_equal(ce, orc.checksum())
    assert orc.tensor("rng_count").max() > 3          # resets happened


def test_full_size_million_world_stress_checksum_of_checksums():
    """BASELINE.json configs[4] (2^20 worlds, forced collisions -> a reset every second step), one GPU:
    per-world checksums against the oracle, plus size-independent properties."""
    import torch

    n = 1 << 20
    sim, orc = make_engine(n), make_oracle(n)
    pe = sim.agent_pos_tensor().to_torch()
    po = orc.tensor("agent_pos")
    off_e, off_o = pe.new_tensor([0.05, 0.0, 0.0]), np.array([0.05, 0.0, 0.0], np.float32)
    for s in range(8):
        sim.fill_actions(SEED, s); orc.fill_actions(SEED, s)
        pe[:, 1] = pe[:, 0] + off_e
        po[:, 1] = po[:, 0] + off_o
        sim.step(); orc.step(16)
    ce = sim.checksum().cpu().numpy().view(np.uint64)
    co = orc.checksum()
    assert np.array_equal(ce, co)
    rc = sim.tensor("rng_count").to_torch()
    assert int(rc.min()) >= 3 + 3 * 4                             # every world reset on every second step (shots add draws)
    obs = sim.observations_tensor().to_torch()
    assert torch.isfinite(obs).all() and not obs[:, :, 103:].any()
    done = sim.done_tensor().to_torch()
    assert ((done == 0) | (done == 1)).all()


def test_north_star_tolerance_against_glibc_arithmetic():
    """north_star's own bar, against the arithmetic the reference's CPU executor would run (glibc libm instead of the
    correctly rounded spec): 1000 steps of random actions, discrete state identical, agent position / velocity /
    orientation within 1e-5 relative.  The ball integrates a velocity that differs by glibc's 1-ulp sinf / cosf errors
    over flights of up to ~300 steps, so its position gets 1e-4 (measured worst: 1.5e-5)."""
    import torch

    n, steps = 64, 1000
    sim, orc = make_engine(n), make_oracle(n, libm=True)
    discrete = ["done", "agent_possession", "ball_physics", "ball_grabbed", "world_clock", "rng_count", "action_mask",
                "action", "reset", "cur_step", "inbounding"]
    for s in range(steps):
        sim.fill_actions(SEED, s)
        orc.fill_actions(SEED, s)
        sim.step()
        orc.step(num_threads=8)
        if s % 50 == 49:
            torch.cuda.synchronize()
            e, o = engine_state(sim), orc.state()
            for k in discrete:
                assert np.array_equal(e[k], o[k]), (k, s)
            for k, rtol in (("agent_pos", 1e-5), ("agent_velocity", 1e-5), ("orientation", 1e-5), ("basketball_pos", 1e-4),
                            ("ball_velocity", 1e-5)):
                err = np.abs(e[k] - o[k]) / np.maximum(np.abs(o[k]), 1.0)
                assert err.max() <= rtol, (k, s, float(err.max()))
    assert int(orc.tensor("rng_count").max()) > 3   # shots / resets happened


def test_config2_8192_worlds_10000_steps_checksums():
    """8192 worlds x 10 000 steps of random actions on one GPU (16 full 620-step episodes per
    world, thousands of shots / resets): per-world checksums of the complete state against the oracle every 1000 steps."""
    import torch

    n, steps, every = 8192, 10000, 1000
    sim, orc = make_engine(n), make_oracle(n)
    for s0 in range(0, steps, every):
        for s in range(s0, s0 + every):
            sim.fill_actions(SEED, s)
            sim.step()
        orc.run_synthetic(SEED, s0, every, 0, 16)
        ce = sim.checksum().cpu().numpy().view(np.uint64)
        co = orc.checksum()
        assert np.array_equal(ce, co), (s0 + every, int((ce != co).sum()))
    assert int(orc.tensor("rng_count").min()) >= 3 + 3 * 15      # every world went through >= 15 resets


@pytest.mark.parametrize("n", [300, 8192, 200000])
def test_step_n_graph_replay_equals_separate_steps(n):
    """bb_step_n: k steps as ONE launch of the looping kernel (one-warp blocks, 64-thread blocks and the large-batch shape
    with phase barriers) == k separate step() calls, bit for bit; another k included."""
    import torch

    a, b = make_engine(n), make_engine(n)
    for s in range(5):                       # some non-trivial state and actions first
        for sim in (a, b):
            sim.fill_actions(SEED, s)
            sim.step()
    for k in (40, 40, 7, 1):
        for _ in range(k):
            a.step()
        b.step_n(k)
        torch.cuda.synchronize()
        assert_states_equal(engine_state(b), engine_state(a), f"step_n N={n} k={k}")
    assert a.launch_count() == b.launch_count()


@pytest.mark.parametrize("n", [300, 8192, 200000])
def test_step_n_with_actions_equals_separate_steps(n):
    """bb_step_n_with_actions: k steps, each with its own action rows, as ONE launch of the looping kernel (every warp takes
    its 16 worlds through the k steps in a row) == k step_with_actions launches, bit for bit on every column; a second call
    continues from the state the first one left."""
    import torch

    k = 24
    a, b = make_engine(n), make_engine(n)
    acts = torch.empty((2 * k, n, 2, 6), dtype=torch.int32, device="cuda")
    for s in range(2 * k):
        a.fill_actions(SEED, s, out=acts[s])
    for half in range(2):
        for s in range(half * k, (half + 1) * k):
            a.step_with_actions(acts[s])
        b.step_n_with_actions(acts[half * k:(half + 1) * k])
        torch.cuda.synchronize()
        assert_states_equal(engine_state(b), engine_state(a), f"step_n_with_actions N={n} half={half}")
    assert b.launch_count() == 1 + 2 * k          # world generation + one pass of the step code per step


def test_step_n_graph_form_still_matches(monkeypatch):
    """BB_STEP_N_GRAPH=1: bb_step_n as round 1's CUDA graph of n step kernels == the looping kernel == separate steps."""
    import torch

    n = 8192
    a, b = make_engine(n), make_engine(n)
    monkeypatch.setenv("BB_STEP_N_GRAPH", "1")
    c = make_engine(n)
    monkeypatch.delenv("BB_STEP_N_GRAPH")
    for s in range(5):
        for sim in (a, b, c):
            sim.fill_actions(SEED, s)
            sim.step()
    for _ in range(33):
        a.step()
    b.step_n(33)
    c.step_n(33)
    torch.cuda.synchronize()
    assert_states_equal(engine_state(b), engine_state(a), "looping kernel")
    assert_states_equal(engine_state(c), engine_state(a), "graph of step kernels")


@pytest.mark.parametrize("n", [300, 8192])
def test_caller_captured_graph_of_steps_equals_separate_steps(n):
    """A caller's own stream capture around ordinary step calls (what bench.py's `value` replays): step_with_actions of two
    engines interleaved inside torch.cuda.graph -- the second and later kernel nodes take programmatic-dependent-launch
    edges, the first a plain one -- replayed twice == the same steps launched one by one, bit for bit."""
    import torch

    k = 6
    eng = [make_engine(n) for _ in range(4)]      # 0, 1: stepped by the graph; 2, 3: their twins, ordinary launches
    acts = torch.empty((k, n, 2, 6), dtype=torch.int32, device="cuda")
    for s in range(k):
        eng[0].fill_actions(SEED, s, out=acts[s])
    for sim in eng:
        for s in range(3):
            sim.step_with_actions(acts[s])
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        for s in range(k):
            eng[s % 2].bind_step_with_actions(acts[s])()
    for _ in range(2):
        graph.replay()
        for s in range(k):
            eng[2 + s % 2].step_with_actions(acts[s])
    torch.cuda.synchronize()
    for i in (0, 1):
        assert_states_equal(engine_state(eng[i]), engine_state(eng[2 + i]), f"captured steps N={n} engine {i}")
