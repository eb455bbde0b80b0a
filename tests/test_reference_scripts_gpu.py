"""The reference's OWN scripts, unmodified, on the drop-in module (north_star: "scripts/env.py, ppo.py and
self_play.py run unmodified").

tools/stage_reference.py copies /root/reference/scripts/*.py + src/constants.py into the git-ignored
baseline/_ref/ in the build container; the copy travels to the GPU box with the snapshot.  These tests import
THOSE files (EnvWrapper: env.py:14-185; run.py:6-19's loop; ppo.py:226-376 as a subprocess) with
`import madrona_basketball` resolving to this repository's alias package, and compare what EnvWrapper returns with
the CPU oracle step by step.  Skipped with a clear reason when the staged copy is absent."""
import json
import os
import re
import subprocess
import sys
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import stage_reference  # noqa: E402

pytestmark = pytest.mark.gpu

needs_ref = pytest.mark.skipif(not stage_reference.staged(),
                               reason="baseline/_ref/ not staged (run tools/stage_reference.py where /root/reference is mounted)")


@pytest.fixture()
def ref_env_module():
    """`import env` = the staged scripts/env.py, under the sys.path the reference's own layout implies."""
    saved_path, saved_mods = list(sys.path), set(sys.modules)
    sys.path[:0] = stage_reference.script_path()
    try:
        import env as ref_env

        assert os.path.realpath(ref_env.__file__).startswith(os.path.realpath(stage_reference.DST)), ref_env.__file__
        yield ref_env
    finally:
        sys.path[:] = saved_path
        for m in set(sys.modules) - saved_mods:
            if m in ("env", "viewer", "agent", "action", "moving_avg", "controllers") or m.startswith("src"):
                sys.modules.pop(m, None)


def _assert_step(o, r, d, orc, idx, where):
    assert np.array_equal(o.cpu().numpy().view(np.int32), orc.tensor("observations")[:, idx].view(np.int32)), where
    assert np.array_equal(r.cpu().numpy().view(np.int32), orc.tensor("reward")[:, idx].view(np.int32)), where
    assert np.array_equal(d.cpu().numpy().view(np.int32), orc.tensor("done")[:, idx].view(np.int32)), where


@needs_ref
def test_real_env_wrapper_run_py_loop_matches_oracle(ref_env_module):
    """scripts/run.py:6-19: EnvWrapper(num_worlds, use_gpu=True, gpu_id=0), set_agent_idx(0), 1000 steps of
    get_blank_actions() — the reference's own benchmark loop — obs / reward / done equal to the oracle at every step."""
    from helpers import make_oracle

    n = 64
    env = ref_env_module.EnvWrapper(n, use_gpu=True, gpu_id=0)       # viewer=True by default: ViewerClass fails on the pygame stub,
    env.set_agent_idx(0)                                             # env.py:66-72 catches it and continues without a viewer
    assert env.viewer is None
    assert type(env.worlds).__module__.startswith("madrona_basketball_b200")
    orc = make_oracle(n)
    for i in range(1000):
        actions = env.get_blank_actions()
        obs, rew, done = env.step(actions)
        orc.tensor("action")[:, 0] = 0                                # env.py:147 writes the trainee's rows only
        orc.step()
        _assert_step(obs, rew, done, orc, 0, f"run.py step {i}")
    assert env.get_input_dim() == 128 and env.get_action_space_size() == 6 and env.get_action_buckets() == [2, 8, 3, 2, 2, 2]


@needs_ref
@pytest.mark.parametrize("trainee", [0, 1])
def test_real_env_wrapper_reset_and_random_steps_match_oracle(ref_env_module, trainee):
    """ppo.py's use of the wrapper (ppo.py:63-89): 3 x (env.reset() + 32 random trainee actions)."""
    import torch

    from helpers import make_oracle

    n = 128
    env = ref_env_module.EnvWrapper(n, use_gpu=True, gpu_id=0, viewer=False, trainee_agent_idx=trainee)
    orc = make_oracle(n)
    gen = torch.Generator(device="cpu").manual_seed(5 + trainee)
    buckets = torch.tensor(env.get_action_buckets())

    def orc_step(a):
        orc.tensor("action")[:, trainee] = a.numpy()
        orc.step()

    for it in range(3):
        obs, rew, done = env.reset()                                  # env.py:178-185
        orc.tensor("reset")[...] = 1
        orc_step(torch.zeros((n, 6), dtype=torch.int32))
        orc.tensor("reset")[...] = 0
        _assert_step(obs, rew, done, orc, trainee, f"reset {it}")
        assert (done == 1).all()
        for s in range(32):
            a = (torch.rand((n, 6), generator=gen) * buckets).to(torch.int32)
            obs, rew, done = env.step(a.to(env.actions.device))
            orc_step(a)
            _assert_step(obs, rew, done, orc, trainee, f"iteration {it} step {s}")


@needs_ref
def test_real_ppo_py_runs_on_the_drop_in(tmp_path):
    """`python scripts/ppo.py --no-viewer --num-envs 8192 --num-iterations 3` (BASELINE configs[3]'s script, README
    default size), unmodified, as a subprocess; its own timer prints the FPS, recorded under gpurun_out/."""
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join(stage_reference.script_path() + [env.get("PYTHONPATH", "")])
    env["WANDB_MODE"] = "disabled"
    cmd = [sys.executable, os.path.join(stage_reference.DST, "scripts", "ppo.py"), "--no-viewer", "--num-envs", "8192",
           "--num-iterations", "3", "--log-every-n-iterations", "1", "--save-model-every-n-iterations", "3"]
    t0 = time.time()
    p = subprocess.run(cmd, cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=900)
    wall = time.time() - t0
    tail = (p.stdout[-3000:] + "\n--- stderr ---\n" + p.stderr[-3000:])
    assert p.returncode == 0, tail
    fps = [int(x) for x in re.findall(r"FPS: (\d+)", p.stdout)]
    assert len(fps) == 3, tail
    assert "Model Model saved at iteration 3" in p.stdout, tail
    assert os.path.exists(tmp_path / "checkpoints" / "Model" / "Model_3.pth")
    import torch

    sd = torch.load(tmp_path / "checkpoints" / "Model" / "Model_3.pth", weights_only=True, map_location="cpu")
    assert all(torch.isfinite(v).all() for v in sd.values())
    assert float(sd["obs_norm.count"]) == 1.0 + 3 * 8192 * 32             # compute_advantages ran on three full rollouts
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "ref_ppo_py_sps.json"), "w") as f:
        json.dump({"what": "the reference's scripts/ppo.py, unmodified, on the drop-in simulator: --no-viewer --num-envs 8192 "
                           "--num-iterations 3 (rollout 32 steps, 4 epochs x 4 minibatches, torch policy + torch update)",
                   "fps_per_iteration": fps, "wall_s_whole_process": wall}, f)
