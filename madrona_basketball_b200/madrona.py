"""Stand-in for the `madrona` submodule nanobind attaches to the reference module
(madrona::py::setupMadronaSubmodule, /root/reference/src/bindings.cpp:16): callers
only use `madrona.ExecMode.{CPU,CUDA}` (scripts/env.py:31-32, scripts/viewer.py:36)."""
import enum


class ExecMode(enum.Enum):
    CPU = 0
    CUDA = 1
