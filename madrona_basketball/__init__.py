"""Drop-in import name of the reference module (`import madrona_basketball as mba`,
scripts/env.py:1).  Re-exports the B200 engine's mirror of src/bindings.cpp."""
from madrona_basketball_b200.manager import SimpleGridworldSimulator  # noqa: F401
from madrona_basketball_b200 import madrona  # noqa: F401
