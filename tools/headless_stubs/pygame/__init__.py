"""Import-only stand-in for pygame on headless boxes.

The reference's scripts/env.py:12 imports viewer.py, which imports pygame at module top
(viewer.py:3, controllers.py:3) even when the viewer is disabled (--no-viewer).  The viewer
(pygame 2-D drawing) is out of scope for this repository; this stub only lets those imports
succeed.  Any attempt to actually *use* pygame raises.

    PYTHONPATH=<repo>:<repo>/tools/headless_stubs python scripts/ppo.py --no-viewer ...
"""


class _Missing:
    def __init__(self, name):
        self._name = name

    def __getattr__(self, item):
        return _Missing(f"{self._name}.{item}")

    def __call__(self, *a, **k):
        raise RuntimeError(f"{self._name}: pygame is not installed (headless stub); run with --no-viewer")


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    if name.isupper() or name.startswith("K_"):
        return 0  # key / event constants referenced at import time
    return _Missing(f"pygame.{name}")
