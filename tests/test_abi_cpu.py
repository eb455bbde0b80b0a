"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/bb_engine.h declares, the Python mirror has the reference's surface
(src/bindings.cpp:14-101), and nothing computes without a GPU (no fallback)."""
import ctypes
import inspect
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols(header="bb_engine.h"):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bb_[a-z_0-9]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    from madrona_basketball_b200 import build, _lib

    build.build_engine()      # no-op when libbball.so is newer than its sources
    return _lib.load()


def test_library_exports_every_declared_symbol(lib):
    from madrona_basketball_b200 import _lib

    declared = _declared_symbols()
    assert len(declared) >= 13
    for sym in declared:
        assert hasattr(lib, sym), f"libbball.so lacks {sym}"
    assert sorted(_lib.SYMBOLS) == declared   # the ctypes table covers the whole header
    rollout = _declared_symbols("bb_rollout.h")
    assert rollout == ["bb_gae", "bb_policy_forward", "bb_rollout", "bb_umma_selftest"]
    for sym in rollout:
        assert hasattr(lib, sym), f"libbball.so lacks {sym}"
    ppo = _declared_symbols("bb_ppo.h")
    assert ppo == ["bb_p2p_connect", "bb_p2p_create", "bb_p2p_destroy", "bb_ppo_adam", "bb_ppo_adam_p2p", "bb_ppo_finish_advantages",
                   "bb_ppo_grad", "bb_ppo_moments", "bb_ppo_moments_static", "bb_ppo_moments_workspace_bytes", "bb_ppo_workspace_bytes", "bb_randperm",
                   "bb_randperm_workspace_bytes"]
    for sym in ppo:
        assert hasattr(lib, sym), f"libbball.so lacks {sym}"


def test_library_is_sm100a_native_code():
    """The .so must carry sm_100a SASS for the step kernel (no PTX-only / other-arch build)."""
    import shutil
    import subprocess

    from madrona_basketball_b200 import _lib

    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out, out


def test_slot_numbers_follow_export_id_enum():
    """types.hpp:10-42 ExportID order == bb_slot == the Python table."""
    from madrona_basketball_b200.manager import SLOTS

    text = open(os.path.join(ROOT, "include", "bb_engine.h")).read()
    enum = dict((m.group(1), int(m.group(2))) for m in re.finditer(r"BB_SLOT_([A-Z_]+) = (\d+)", text))
    order = ["RESET", "GAME_STATE", "ACTION", "ACTION_MASK", "AGENT_POS", "OBSERVATIONS", "REWARD", "DONE",
             "AGENT_ENTITY_ID", "AGENT_POSSESSION", "ORIENTATION", "TEAM_DATA", "AGENT_STATS", "BASKETBALL_POS",
             "BALL_PHYSICS", "BALL_ENTITY_ID", "BALL_GRABBED", "BALL_VELOCITY", "HOOP_POS"]
    assert [enum[k] for k in order] == list(range(19))
    assert SLOTS["agent_team"] == enum["TEAM_DATA"] and SLOTS["observations"] == enum["OBSERVATIONS"]


def test_python_surface_matches_bindings_cpp():
    """Class name, constructor keywords and the 22 methods of src/bindings.cpp:17-99."""
    import madrona_basketball as mba
    from madrona_basketball.madrona import ExecMode

    assert ExecMode.CPU is not ExecMode.CUDA and mba.madrona.ExecMode is ExecMode
    cls = mba.SimpleGridworldSimulator
    params = list(inspect.signature(cls.__init__).parameters)
    assert params[1:9] == ["discrete_x", "discrete_y", "start_x", "start_y", "max_episode_length", "exec_mode",
                           "num_worlds", "gpu_id"]
    assert inspect.signature(cls.__init__).parameters["gpu_id"].default == -1
    methods = ["step", "set_action", "trigger_reset", "reset_tensor", "action_tensor", "observations_tensor",
               "agent_pos_tensor", "reward_tensor", "done_tensor", "basketball_pos_tensor", "ball_physics_tensor",
               "ball_grabbed_tensor", "agent_possession_tensor", "agent_entity_id_tensor", "ball_entity_id_tensor",
               "agent_team_tensor", "game_state_tensor", "orientation_tensor", "hoop_pos_tensor",
               "agent_stats_tensor", "action_mask_tensor", "ball_velocity_tensor"]
    for m in methods:
        assert callable(getattr(cls, m)), m
    # nanobind exposes set_action(world_idx, agent_idx, move_speed, move_angle, rotate, grab, pass, shoot)
    sa = list(inspect.signature(cls.set_action).parameters)
    assert sa[1:7] == ["world_idx", "agent_idx", "move_speed", "move_angle", "rotate", "grab"]


def test_no_cpu_path_behind_the_boundary(lib):
    """ExecMode.CPU is refused, and on a box without a GPU bb_create fails loudly instead of falling back."""
    import torch

    import madrona_basketball as mba
    from madrona_basketball_b200 import _lib

    args = dict(discrete_x=32, discrete_y=17, start_x=15.7575, start_y=8.382, max_episode_length=39600, num_worlds=4)
    with pytest.raises(RuntimeError, match="CUDA-only"):
        mba.SimpleGridworldSimulator(exec_mode=mba.madrona.ExecMode.CPU, **args)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CUDA device|bb_create"):
            mba.SimpleGridworldSimulator(exec_mode=mba.madrona.ExecMode.CUDA, gpu_id=0, **args)
        assert "no CPU path" in _lib.last_error() or "CUDA" in _lib.last_error()
    cfg = _lib.BBConfig(0, 32, 17, 15.7575, 8.382, 39600, 0, 0, 1)
    assert not lib.bb_create(ctypes.byref(cfg))       # num_worlds must be positive
    assert "num_worlds" in _lib.last_error()


def test_product_package_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under madrona_basketball_b200/ (or the drop-in
    alias package) may reference it."""
    for pkg in ("madrona_basketball_b200", "madrona_basketball"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, pkg)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                    text = open(os.path.join(dirpath, f)).read()
                    assert not re.search(r"^\s*(import|from)\s+oracle\b", text, flags=re.M), (dirpath, f)
                    assert "liboracle" not in text and "bb_oracle" not in text, (dirpath, f)


def test_library_builds_from_clean_sources(tmp_path):
    """A from-clean nvcc compile of csrc/bb_engine.cu into a scratch path (not the shipped libbball.so): every symbol
    of include/*.h is exported and the ELF is sm_100a.  ~45 s of nvcc."""
    import glob
    import shutil
    import subprocess

    from madrona_basketball_b200 import build

    try:
        build.find_nvcc()
    except RuntimeError:
        pytest.skip("nvcc not available")
    out = str(tmp_path / "libbball_clean.so")
    assert build.build_engine(force=True, out=out) == out and os.path.getsize(out) > 100_000
    fresh = ctypes.CDLL(out)
    declared = []
    for header in sorted(glob.glob(os.path.join(ROOT, "include", "*.h"))):
        if os.path.basename(header) != "bb_math.h":      # the arithmetic spec: static inline functions, not exports
            declared += _declared_symbols(os.path.basename(header))
    assert len(declared) >= 34
    for sym in declared:
        assert hasattr(fresh, sym), f"clean build lacks {sym}"
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if os.path.exists(cuobjdump):
        elf = subprocess.run([cuobjdump, "-lelf", out], capture_output=True, text=True).stdout
        assert "sm_100a" in elf and "sm_90" not in elf, elf
    # the shipped library is the one these sources produce (content hash written by the build)
    assert not build.needs_build(), "madrona_basketball_b200/libbball.so is stale w.r.t. the sources in the tree"


def test_squared_length_thresholds_equal_the_sqrt_tests():
    """bb_world.cuh replaces `length() <= c` / `>= c` by a test on the squared length (kLen2Le* / kLen2Ge*): for every float
    around the threshold the two verdicts must agree (numpy's float32 sqrt is correctly rounded, like the device's
    __fsqrt_rn and the oracle's sqrtf)."""
    import re

    import numpy as np

    src = open(os.path.join(ROOT, "madrona_basketball_b200", "csrc", "bb_world.cuh")).read()
    consts = {m.group(1): float.fromhex(m.group(2)) for m in re.finditer(r"constexpr float (kLen2\w+) = (0x[0-9a-fp.+-]+)f;", src)}
    cases = [("kLen2LeHoopZone", np.float32(0.1), "le"), ("kLen2LeAgentSize", np.float32(0.2), "le"), ("kLen2GeArc", np.float32(7.24), "ge")]
    assert set(consts) == {c[0] for c in cases}
    for name, c, kind in cases:
        T = np.float32(consts[name])
        assert float(T) == consts[name]                      # the literal is a float32
        bits = np.array([T]).view(np.uint32)[0]
        s = np.arange(int(bits) - 100000, int(bits) + 100000, dtype=np.uint32).view(np.float32)
        root = np.sqrt(s)
        assert root.dtype == np.float32
        if kind == "le":
            assert np.array_equal(root <= c, s <= T), name
        else:
            assert np.array_equal(root >= c, s >= T), name


def test_fastcall_glue_calls_the_bound_entry_point(lib):
    """csrc/bb_fastcall.c: bind5(fn, last_error, a0..a4)() is `fn(a0..a4)` -- the arguments arrive unchanged, a zero return
    yields None, a non-zero return raises with bb_last_error()'s text.  No compute: the bound function here is a ctypes
    callback standing in for bb_serve_step."""
    from madrona_basketball_b200 import build

    build.build_fastcall()
    from madrona_basketball_b200 import _fastcall

    seen = []
    proto = ctypes.CFUNCTYPE(ctypes.c_int, *([ctypes.c_void_p] * 5))

    def target(a, b, c, d, e):
        seen.append((a, b, c, d, e))
        return 0 if len(seen) == 1 else 7

    cb = proto(target)
    addr = lambda f: ctypes.cast(f, ctypes.c_void_p).value
    call = _fastcall.bind5(addr(cb), addr(lib.bb_last_error), 0x1000, 0x2000, 0x3000, 0, 0x5000)
    assert call() is None
    assert seen == [(0x1000, 0x2000, 0x3000, None, 0x5000)]
    with pytest.raises(RuntimeError, match="rc 7"):
        call()
    with pytest.raises(TypeError):
        call(1)
    with pytest.raises(ValueError):
        _fastcall.bind5(0, 0, 0, 0, 0, 0, 0)
