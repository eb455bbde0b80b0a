"""`from madrona_basketball.madrona import ExecMode` (scripts/viewer.py:36)."""
from madrona_basketball_b200.madrona import ExecMode  # noqa: F401
