#!/usr/bin/env python
"""Stages the reference's own Python (scripts/*.py + src/constants.py) under the git-ignored
baseline/_ref/ so that it travels to the GPU box with the gpurun snapshot.

The reference's C++ cannot be built or pip-installed here (its build backend and every source file
need the absent Madrona submodule), but its scripts are plain Python on top of
`import madrona_basketball`: EnvWrapper (scripts/env.py:14-185), run.py, ppo.py, agent.py, action.py,
buffers.py, viewer.py's log loader.  The GPU tests run THOSE files, unmodified, on this repo's drop-in
module (tests/test_reference_scripts_gpu.py); nothing of them is committed.

    python tools/stage_reference.py            # no-op with a message when /root/reference is absent
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
DST = os.path.join(ROOT, "baseline", "_ref")
STUBS = os.path.join(ROOT, "tools", "headless_stubs")


def staged() -> bool:
    return os.path.exists(os.path.join(DST, "scripts", "env.py")) and os.path.exists(os.path.join(DST, "src", "constants.py"))


def stage(verbose: bool = True) -> bool:
    if not os.path.isdir(os.path.join(REF, "scripts")):
        if verbose:
            print(f"[stage_reference] {REF} not present; keeping whatever is under {DST}")
        return staged()
    manifest = {}
    for sub, names in (("scripts", sorted(f for f in os.listdir(os.path.join(REF, "scripts")) if f.endswith(".py"))),
                       ("src", ["constants.py"])):
        os.makedirs(os.path.join(DST, sub), exist_ok=True)
        for name in names:
            src, dst = os.path.join(REF, sub, name), os.path.join(DST, sub, name)
            shutil.copyfile(src, dst)
            with open(dst, "rb") as f:
                manifest[f"{sub}/{name}"] = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(DST, "MANIFEST.json"), "w") as f:
        json.dump({"source": REF, "files": manifest}, f, indent=1, sort_keys=True)
    if verbose:
        print(f"[stage_reference] staged {len(manifest)} files under {DST}")
    return True


def script_path() -> list:
    """sys.path entries (in order) under which the staged scripts import: the drop-in `madrona_basketball`
    (repo root), the pygame import stub, the scripts themselves, and the parent of `src/` (env.py:8-10)."""
    return [ROOT, STUBS, os.path.join(DST, "scripts"), DST]


if __name__ == "__main__":
    sys.exit(0 if stage() else 1)
