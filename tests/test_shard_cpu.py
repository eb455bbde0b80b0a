"""World sharding: partition arithmetic, shard invariance of the world-keyed
action stream (checked on the oracle), and the only collective near the path — episode-stat
reduction — over gloo with world_size 2."""
import os
import socket

import numpy as np
import pytest

from helpers import make_oracle
from madrona_basketball_b200.shard import shard_worlds

SEED = 1234


@pytest.mark.parametrize("total,ws", [(1, 1), (8192, 1), (65536, 8), (10, 3), (7, 8), (1 << 20, 8), (1000, 7)])
def test_shard_worlds_is_a_contiguous_partition(total, ws):
    first, covered = 0, 0
    sizes = []
    for r in range(ws):
        f, c = shard_worlds(total, ws, r)
        assert f == first
        first += c
        covered += c
        sizes.append(c)
    assert covered == total and max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_worlds(total, ws, ws)


def test_shard_evolves_like_the_same_worlds_of_the_full_simulation():
    """A shard [k, k+m) fed the global-world-keyed action stream reproduces those worlds of the
    unsharded run (entity ids are instance-local, as in the reference, so id-free columns only)."""
    n, ws = 96, 3
    full = make_oracle(n)
    parts = []
    for r in range(ws):
        f, c = shard_worlds(n, ws, r)
        parts.append((f, c, make_oracle(c)))
    for s in range(120):
        full.fill_actions(SEED, s)
        full.step()
        for f, c, o in parts:
            o.fill_actions(SEED, s, world_offset=f)
            o.step()
    id_free = ["observations", "reward", "done", "agent_pos", "orientation", "basketball_pos", "ball_velocity",
               "action", "action_mask", "agent_velocity", "attributes", "rng_count", "world_clock", "cur_step"]
    for f, c, o in parts:
        for k in id_free:
            assert np.array_equal(full.tensor(k)[f:f + c].view(np.int32), o.tensor(k).view(np.int32)), (f, k)
    assert len(np.unique(full.tensor("rng_count"))) > 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world_size, port, total, out_q):
    import torch.distributed as dist

    from madrona_basketball_b200.shard import reduce_episode_stats, shard_worlds as sw

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world_size))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        import oracle
        first, count = sw(total, world_size, rank)
        o = oracle.Oracle(count)
        episodes, reward_sum = 0.0, 0.0
        for s in range(220):
            o.fill_actions(SEED, s, world_offset=first)
            o.step()
            done = o.tensor("done")[:, 0]
            episodes += float(done.sum())
            reward_sum += float(o.tensor("reward").astype(np.float64).sum())
        stats = reduce_episode_stats({"episodes": episodes, "reward_sum": reward_sum, "world_steps": 220.0 * count})
        if rank == 0:
            out_q.put(stats)
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_stat_reduction_matches_single_process():
    import torch.multiprocessing as mp

    total, ws = 50, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, ws, port, total, q)) for r in range(ws)]
    for p in procs:
        p.start()
    stats = q.get(timeout=180)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    o = make_oracle(total)
    episodes, reward_sum = 0.0, 0.0
    for s in range(220):
        o.fill_actions(SEED, s)
        o.step()
        episodes += float(o.tensor("done")[:, 0].sum())
        reward_sum += float(o.tensor("reward").astype(np.float64).sum())
    assert stats["episodes"] == episodes and stats["world_steps"] == 220.0 * total
    assert abs(stats["reward_sum"] - reward_sum) <= 1e-6 * max(1.0, abs(reward_sum))
    assert episodes > 0


def test_reduce_episode_stats_without_process_group():
    from madrona_basketball_b200.shard import reduce_episode_stats

    assert reduce_episode_stats({"b": 2, "a": 1.5}) == {"a": 1.5, "b": 2.0}
