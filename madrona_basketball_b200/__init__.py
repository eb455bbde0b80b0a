"""B200-native step engine for the madrona_basketball environment.

Only the per-world simulation step (the systems of Sim::setupTasks in the
reference) lives here, behind the reference's own Python surface:
`SimpleGridworldSimulator` mirrors /root/reference/src/bindings.cpp.
"""
from . import madrona
from .manager import SimpleGridworldSimulator, Tensor, FLAG_TAG_MASK
from .shard import shard_worlds, ShardedSimulator

__all__ = ["SimpleGridworldSimulator", "Tensor", "madrona", "FLAG_TAG_MASK", "shard_worlds", "ShardedSimulator"]
