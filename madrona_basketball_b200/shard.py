"""World sharding across GPUs: one process + one engine per GPU, contiguous world
ranges, no collective in the step (every ctx.get in the reference
targets the caller's own world, and the RNG key is identical in every world —
src/sim.cpp:89 — so shards need no seed offset and never communicate).

torch.distributed is used only outside the step: episode-statistics reduction and
the max-over-ranks timing of bench.py.  Works with the nccl backend on GPUs and
with gloo on CPU (tests/test_shard.py).
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Tuple


def shard_worlds(total_worlds: int, world_size: int, rank: int) -> Tuple[int, int]:
    """(first_world, num_worlds) of `rank`: contiguous, sizes differ by at most one."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world_size {world_size}")
    base, extra = divmod(total_worlds, world_size)
    count = base + (1 if rank < extra else 0)
    first = rank * base + min(rank, extra)
    return first, count


def env_rank() -> Tuple[int, int, int]:
    """(rank, local_rank, world_size) from the torchrun environment (defaults: single process)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


class ShardedSimulator:
    """This rank's shard of a `total_worlds`-world simulation.

    `.sim` is a plain SimpleGridworldSimulator over the local worlds (entity ids are
    instance-local, as in the reference).  `world_offset` is the global index of local
    world 0 — the synthetic action stream is keyed by global world index so that any
    sharding replays the same per-world inputs.
    """

    def __init__(self, total_worlds: int, rank: Optional[int] = None, world_size: Optional[int] = None,
                 gpu_id: Optional[int] = None, **sim_kwargs):
        from .manager import SimpleGridworldSimulator
        from .madrona import ExecMode

        r, local, ws = env_rank()
        self.rank = r if rank is None else rank
        self.world_size = ws if world_size is None else world_size
        self.total_worlds = total_worlds
        self.world_offset, self.num_worlds = shard_worlds(total_worlds, self.world_size, self.rank)
        kwargs = dict(discrete_x=32, discrete_y=17, start_x=31.515 / 2.0, start_y=16.764 / 2.0,
                      max_episode_length=39600, exec_mode=ExecMode.CUDA)
        kwargs.update(sim_kwargs)
        self.sim = SimpleGridworldSimulator(num_worlds=self.num_worlds, gpu_id=local if gpu_id is None else gpu_id,
                                            **kwargs)

    def step(self) -> None:
        self.sim.step()

    def fill_actions(self, seed: int, step: int, out=None) -> None:
        self.sim.fill_actions(seed, step, world_offset=self.world_offset, out=out)


def reduce_episode_stats(local: Dict[str, float], group=None) -> Dict[str, float]:
    """Sum per-rank episode statistics (finished episodes, reward sums, step counts).

    The only collective near the step path; one flat all-reduce per call."""
    import torch
    import torch.distributed as dist

    keys = sorted(local)
    if not (dist.is_available() and dist.is_initialized()):
        return {k: float(local[k]) for k in keys}
    device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    buf = torch.tensor([float(local[k]) for k in keys], dtype=torch.float64, device=device)
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return {k: float(v) for k, v in zip(keys, buf.tolist())}
