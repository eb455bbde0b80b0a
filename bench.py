#!/usr/bin/env python
"""Benchmark of the basketball step path (BASELINE.json metric: env steps/sec = worlds x steps / s).

    python bench.py --gpus N --steps K --warmup W            # B200 engine (this repo)
    python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the oracle restatement of the
                                                             # reference's CPU path on all host threads
                                                             # (the Madrona CPU executor itself cannot be
                                                             # built here: its submodule is absent)

A "step" is one Manager::step() over one batch of synthetic random actions for
this rank's worlds.  Workload at N GPUs: 8192 worlds per GPU (BASELINE.json configs[1]; 8 GPUs x 8192
= configs[2]'s 65 536 worlds), sharded by world with no collective in the step => weak scaling.

Prints ONE JSON line on rank 0.  Timing: CUDA events on the launching stream, max over ranks; L2 is
flushed (256 MiB memset) before every timed step and the flush is outside the per-step events.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "env_steps_per_sec"
UNIT = "world-steps/s"
WORLDS_PER_GPU = 8192
BYTES_PER_WORLD_STEP = 1872  # 56 in + 348 state read + 348 state write + 1120 outputs
SEED = 1234
LARGE_WORLDS = 1 << 20       # HBM-bound regime (working set 1.9 GB >> 126 MB L2)
PREWARM_S = 0.5              # declared device pre-warm before the timed window (config.prewarm)
DEPHASE_SPAN = 150           # steps: the shards of a rank are spread over this much of the episode cycle before anything is timed
SPACER_US = 300              # a spin kernel of this length is queued ahead of the timed window while the host enqueues its K launches


def enqueue_spacer(torch, flush):
    """Keeps the stream busy for ~SPACER_US without touching memory (torch.cuda._sleep: one thread spinning on the clock), so
    that the K launches enqueued behind it are all queued when the window opens and the kernel's code stays cached as in any
    steady-state loop; the rows are cold by the shard rotation, not by a flush.  Falls back to 256 MiB memsets."""
    try:
        torch.cuda._sleep(int(SPACER_US * 1e-6 * 1.9e9))
    except Exception:
        if flush is None:
            flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        for _ in range(6):
            flush.zero_()


def dephase_steps(R: int) -> int:
    """Extra untimed steps per shard index.  Every world of every shard starts from the same spawn and the game's clocks run
    in lock step, so shots, rebounds and resets come in waves (a step costs 1.0x on a quiet step, 1.6x in a wave, waves
    48-96 steps apart: profiles/r02b_step_times_1m_200_steps.txt).  With all shards at the same phase a 20-launch window
    lands on ONE phase of that cycle and reads 8.1 or 10.3 us per step by chance; with shard i advanced i x dephase_steps(R)
    steps the window's launches sample the whole cycle -- the long-run mean, run after run."""
    return max(7, -(-DEPHASE_SPAN // R))


def measured_peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples SM clock + throttle reasons through NVML while timed regions run."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._active = threading.Event()
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        except Exception as e:  # NVML missing: report that instead of inventing numbers
            self._nv = None
            self.error = repr(e)

    def _run(self):
        nv = self._nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self._stop.is_set():
            if not self._active.wait(timeout=0.5):
                continue
            if self._active.is_set():
                try:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                    for bit, name in names.items():
                        if mask & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
            time.sleep(0.002)

    def __enter__(self):
        self._active.set()
        return self

    def __exit__(self, *a):
        self._active.clear()

    def pause(self):
        """Parks the polling thread (it blocks on an event instead of waking every 2 ms)."""
        self._active.clear()

    def report(self):
        self._stop.set()
        if self._nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": getattr(self, "error", "nvml unavailable")}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def workload_string(worlds_per_gpu: int, n_gpus: int) -> str:
    """config.workload, the same text for both arms (the driver compares them)."""
    return (f"{worlds_per_gpu} worlds per GPU x {n_gpus} GPU = {worlds_per_gpu * n_gpus} worlds, 2 agents + ball per world, "
            f"random actions for both agents (Threefry, seed {SEED}), sim seed 0")


_CPU_BUILD = {}


def pick_cpu_build(worlds: int, threads: int):
    """The CPU arm times the FASTER of the oracle's two glibc-libm builds on this host: the generic -O2 one
    (oracle/Makefile) and a -O3 -march=native one compiled on this machine (oracle.build_native).  Same results."""
    if "kind" in _CPU_BUILD:
        return _CPU_BUILD["kind"], _CPU_BUILD["note"]
    import oracle

    oracle.build()
    rates = {}
    for kind in (True, "native"):
        try:
            orc = oracle.Oracle(worlds, start_x=31.515 / 2.0, start_y=16.764 / 2.0, libm=kind)
        except Exception as e:   # no compiler on the box: the generic build stands
            rates[kind] = 0.0
            continue
        orc.run_synthetic(SEED, 0, 20, 0, threads)
        t0 = time.perf_counter()
        orc.run_synthetic(SEED, 20, 100, 0, threads)
        rates[kind] = worlds * 100 / (time.perf_counter() - t0)
        orc.close()
    kind = "native" if rates["native"] > rates[True] else True
    _CPU_BUILD["kind"] = kind
    _CPU_BUILD["note"] = (f"oracle/ restatement on glibc libm, {'-O3 -march=native' if kind == 'native' else '-O2 generic'} build "
                          f"(probe: generic {rates[True] / 1e6:.1f} M/s, native {rates['native'] / 1e6:.1f} M/s)")
    return kind, _CPU_BUILD["note"]


def time_cpu_arm(worlds: int, steps: int, warmup: int, threads: int, kind=True):
    """Times the oracle (CPU restatement of the reference's step) on `threads` host threads: ONE call, every
    thread steps its own block of worlds through all `steps` steps (no thread start-up or barrier per step: the
    steady-state rate a CPU executor with a persistent thread pool delivers)."""
    import oracle

    orc = oracle.Oracle(worlds, start_x=31.515 / 2.0, start_y=16.764 / 2.0, libm=kind)
    orc.run_synthetic(SEED, 0, warmup, 0, threads)
    t0 = time.perf_counter()
    orc.run_synthetic(SEED, warmup, steps, 0, threads)
    dt = time.perf_counter() - t0
    orc.close()
    return worlds * steps / dt, dt


def time_cpu_arm_for(worlds: int, seconds: float, threads: int, multiple_of: int = 1):
    """Bounded sample: a short probe sizes ONE run to about `seconds` of CPU work."""
    kind, note = pick_cpu_build(worlds, threads)
    rate, _ = time_cpu_arm(worlds, 100, 10, threads, kind)
    steps = max(200, int(rate * seconds / worlds))
    steps = ((steps + multiple_of - 1) // multiple_of) * multiple_of
    value, dt = time_cpu_arm(worlds, steps, 50, threads, kind)
    return value, dt, steps, note


def ncu_traffic(key: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of bb_step_kernel from the committed
    `ncu --set full` capture (profiles/traffic.json; written by tools/ncu_summary.py runs)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(key)
    except Exception:
        return None


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    worlds = args.worlds_per_gpu * args.gpus
    steps, warmup = args.steps, args.warmup
    # A K-step block at 8192 worlds lasts only milliseconds, so the arm runs B consecutive K-step blocks inside ONE
    # oracle call of about args.cpu_seconds (threads started once, no barrier between steps) and reports the mean block:
    # the steady-state rate, the same number the engine arm's cpu_baseline prints on this box.
    value, dt, total_steps, note = time_cpu_arm_for(worlds, args.cpu_seconds, threads, multiple_of=steps)
    blocks = total_steps // steps
    sample = (f"{worlds} worlds x {total_steps} steps = {blocks} consecutive blocks of {steps} steps in one call ({dt:.1f} s of CPU work) "
              f"after 50 warm-up steps; {note}; {threads} threads, synthetic random actions")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": dt / total_steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(args.worlds_per_gpu, args.gpus),
                   "what": "CPU restatement of the reference step (the Madrona CPU executor cannot be built on this box)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


def rotating_rate(n, total_worlds, n_gpus, rank, local_rank, K, Wm, prewarm_s, seed_base=0):
    """World-steps/s of K back-to-back launches over R independent n-world shards on this GPU (inputs larger than L2),
    max over ranks.  The measurement `value` uses, as a function (strong-scaling points)."""
    import torch
    import torch.distributed as dist

    from madrona_basketball_b200.shard import ShardedSimulator

    dev = torch.device("cuda", local_rank)
    l2_bytes = torch.cuda.get_device_properties(dev).L2_cache_size
    R = max(2, -(-int(2.5 * l2_bytes) // (n * BYTES_PER_WORLD_STEP)))
    shards = [ShardedSimulator(total_worlds * R, rank=rank * R + i, world_size=n_gpus * R, gpu_id=local_rank) for i in range(R)]
    acts = torch.empty((min(K + Wm, 8), n, 2, 6), dtype=torch.int32, device=dev)
    for s in range(acts.shape[0]):
        shards[s % R].fill_actions(SEED, seed_base + s, out=acts[s])
    calls = [[sh.sim.bind_step_with_actions(acts[b]) for b in range(acts.shape[0])] for sh in shards]
    for j in range(R):                       # shards at different phases of the episode cycle (dephase_steps)
        for s in range(j * dephase_steps(R)):
            calls[j][s % acts.shape[0]]()
        torch.cuda.synchronize()
    t0, i = time.perf_counter(), 0
    while i < 2 * R or time.perf_counter() - t0 < prewarm_s:
        calls[i % R][i % acts.shape[0]]()
        i += 1
        if i % 64 == 0:
            torch.cuda.synchronize()
    for s in range(Wm):
        calls[(i + s) % R][s % acts.shape[0]]()
    first = (i + Wm) % R
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if n_gpus > 1:
        dist.barrier()
    torch.cuda.synchronize()
    enqueue_spacer(torch, None)   # see run_engine: the K launches queue up behind it
    e0.record()
    for s in range(K):
        calls[(first + s) % R][s % acts.shape[0]]()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if n_gpus > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    stream_ms = float(t.item())
    # the same K steps as one CUDA-graph replay (run_engine's `value`): capture, one untimed replay, complete the shard cycle
    first = (first + K) % R
    g, side = torch.cuda.CUDAGraph(), torch.cuda.Stream(dev)
    with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
        for s in range(K):
            shards[(first + s) % R].sim.bind_step_with_actions(acts[s % acts.shape[0]])()
    g.replay()
    n_fill = (R - K % R) % R
    while n_fill < Wm:
        n_fill += R
    for s in range(n_fill):
        calls[(first + K + s) % R][s % acts.shape[0]]()
    if n_gpus > 1:
        dist.barrier()
    torch.cuda.synchronize()
    enqueue_spacer(torch, None)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if n_gpus > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    del calls, shards, g
    return {"value": total_worlds * K / (ms * 1e-3), "unit": UNIT, "worlds_per_gpu": n, "us_per_step": ms / K * 1e3, "shards_per_gpu": R,
            "hbm_frac": n * BYTES_PER_WORLD_STEP / (ms / K * 1e-3) / 1e9 / measured_peak_gbs()[0],
            "stream_launches": {"value": total_worlds * K / (stream_ms * 1e-3), "us_per_step": stream_ms / K * 1e3}}


def stress_rate(total_worlds, n_gpus, rank, local_rank, steps=100, warm=20):
    """BASELINE.json configs[4]: reset / collision stress: every step agent 1 is teleported next to agent 0
    through the writable exported agent_pos view, so agentCollisionSystem fires each step and the world resets every
    second step.  Per-step events (the teleport write and the action fill are outside them), max over ranks of the sum."""
    import torch
    import torch.distributed as dist

    from madrona_basketball_b200.shard import ShardedSimulator

    dev = torch.device("cuda", local_rank)
    sh = ShardedSimulator(total_worlds, rank=rank, world_size=n_gpus, gpu_id=local_rank)
    sim, n = sh.sim, sh.num_worlds
    acts = torch.empty((4, n, 2, 6), dtype=torch.int32, device=dev)
    pos = sim.agent_pos_tensor().to_torch()
    off = pos.new_tensor([0.05, 0.0, 0.0])
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for s in range(warm + steps):
        sh.fill_actions(SEED, s, out=acts[s % 4])
        pos[:, 1] = pos[:, 0] + off
        if s >= warm:
            ev[s - warm][0].record()
        sim.step_with_actions(acts[s % 4])
        if s >= warm:
            ev[s - warm][1].record()
    torch.cuda.synchronize()
    t = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device=dev)
    if n_gpus > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    draws = int(sim.tensor("rng_count").to_torch().min().item())
    del sim, sh
    return {"value": total_worlds * steps / (ms * 1e-3), "unit": UNIT, "worlds": total_worlds, "worlds_per_gpu": n,
            "us_per_step": ms / steps * 1e3, "hbm_frac": n * BYTES_PER_WORLD_STEP / (ms / steps * 1e-3) / 1e9 / measured_peak_gbs()[0],
            "min_rng_draws_per_world": draws,
            "what": "configs[4]: agent 1 teleported onto agent 0 before every step (collision each step, reset every second step); "
                    "working set > L2 at this size, per-step CUDA events"}


def run_engine(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from madrona_basketball_b200.shard import ShardedSimulator, env_rank

    rank, local_rank, world_size = env_rank()
    if world_size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n_gpus = max(world_size, 1)
    if args.gpus != n_gpus and rank == 0:
        print(f"[bench] --gpus {args.gpus} but WORLD_SIZE={world_size}; using {n_gpus}", file=sys.stderr)
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)

    K, Wm = args.steps, args.warmup
    total_worlds = args.worlds_per_gpu * n_gpus
    n = args.worlds_per_gpu
    # `value` times K back-to-back step launches between ONE pair of CUDA events, as the bench contract words it.  So that
    # no launch finds its rows in the 126 MB L2 ("inputs larger than L2"), the rank owns R independent shards of `n` worlds
    # (R x n x 1872 B >= 2.5 x L2) and every launch advances the next shard, round-robin: by the time a shard comes round
    # again 2.5 L2 sizes of other rows have passed through the cache.  One launch = one Manager::step() over n worlds.
    l2_bytes = torch.cuda.get_device_properties(dev).L2_cache_size
    R = max(2, -(-int(2.5 * l2_bytes) // (n * BYTES_PER_WORLD_STEP)))
    shards = [ShardedSimulator(total_worlds * R, rank=rank * R + i, world_size=n_gpus * R, gpu_id=local_rank) for i in range(R)]
    shard = shards[0]
    sim = shard.sim
    assert all(sh.num_worlds == n for sh in shards)
    sampler = ClockSampler(local_rank)

    # inputs resident in HBM before the timed region: one action batch per step and shard slot
    actions = torch.empty((K + Wm, n, 2, 6), dtype=torch.int32, device=dev)
    for s in range(K + Wm):
        shards[s % R].fill_actions(SEED, s, out=actions[s])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------------------------------------------------------- value: device-resident, cold rows, K launches in one event pair
    # Declared device pre-warm (stated in config.prewarm): an idle GPU needs a few hundred milliseconds of work before its
    # clocks and the driver's launch path reach steady state; with --steps 20 --warmup 5 the timed window would otherwise
    # sit on that ramp.  Same launches as the timed loop, at least PREWARM_S seconds, then the W warm-up steps as passed.
    # launches pre-bound (pointer / stream look-ups done once): a Python-level step_with_actions call costs ~10 us of host
    # time, more than the kernel lasts, and the timed loop must not be host-bound
    bound = {}

    def launch(shard_idx, batch):
        key = (shard_idx, batch)
        if key not in bound:
            bound[key] = shards[shard_idx].sim.bind_step_with_actions(actions[batch])
        bound[key]()

    dephase = dephase_steps(R)
    for i in range(R):                       # untimed: shard i sits i x dephase steps further into the episode cycle
        for s in range(i * dephase):
            launch(i, s % (K + Wm))
        torch.cuda.synchronize()
    t_pw, n_pw = time.perf_counter(), 0
    while n_pw < 4 * R or time.perf_counter() - t_pw < PREWARM_S:
        launch(n_pw % R, n_pw % (K + Wm))
        n_pw += 1
        if n_pw % 64 == 0:
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    for s in range(Wm):
        launch((n_pw + s) % R, s)
    first = (n_pw + Wm) % R
    timed = [shards[(first + s) % R].sim.bind_step_with_actions(actions[Wm + s]) for s in range(K)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = sum(sh.sim.launch_count() for sh in shards)
    barrier()
    with sampler:
        wall0 = time.perf_counter()
        # The K launches are enqueued while the GPU is still busy with a spacer (a ~300 us spin kernel) issued just before
        # the first event, so the timed window holds K step kernels the stream already has in its queue: device time, not
        # the host's enqueue pace or its scheduling jitter (5 us per launch from Python, half a step).
        enqueue_spacer(torch, flush)
        e0.record()
        for call in timed:
            call()
        e1.record()
        barrier()
        wall = time.perf_counter() - wall0
    stream_launches = sum(sh.sim.launch_count() for sh in shards) - launches0
    stream_dev_ms = float(e0.elapsed_time(e1))
    t = torch.tensor([stream_dev_ms], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    stream_value = total_worlds * K / (float(t.item()) * 1e-3)

    # ---------------------------------------------------------------- value: the same K cold-row steps as ONE CUDA graph launch
    # A step kernel lasts ~8 us, so a loop of ordinary launches is bound by the stream's per-launch front-end cost
    # (stream_launches above); a launch-bound inner loop belongs in a CUDA graph.  The K steps over the rotating shards are
    # captured once (the engine's launches carry their programmatic-dependent-launch attribute into the graph's edges) and the
    # timed window is ONE replay: K step kernels, each on the next shard's rows, which are as cold as above -- between two
    # touches of a shard the other R - 1 shards (2.4 x L2) pass through the cache, in the warm-up below and across replays.
    first = (first + K) % R
    side = torch.cuda.Stream(dev)
    gK = torch.cuda.CUDAGraph()
    launches0 = sum(sh.sim.launch_count() for sh in shards)
    with torch.cuda.graph(gK, stream=side, capture_error_mode="thread_local"):
        for s in range(K):     # bound inside the capture: the closure keeps the capturing stream
            shards[(first + s) % R].sim.bind_step_with_actions(actions[Wm + s])()
    gpu_launches = sum(sh.sim.launch_count() for sh in shards) - launches0     # kernel nodes in the graph = launches per replay
    assert gpu_launches == K, (gpu_launches, K)
    gK.replay()                # untimed: instantiation / upload of the graph, and K more warm-up steps
    n_fill = (R - K % R) % R   # ordinary launches over the shards the graph leaves out of this cycle, then whole cycles up to >= W
    while n_fill < Wm:
        n_fill += R
    for s in range(n_fill):
        launch((first + K + s) % R, s % (K + Wm))
    barrier()
    with sampler:
        wall0 = time.perf_counter()
        enqueue_spacer(torch, flush)
        e0.record()
        gK.replay()
        e1.record()
        barrier()
        wall = time.perf_counter() - wall0
    dev_ms = float(e0.elapsed_time(e1))
    t = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    max_ms = float(t.item())
    value = total_worlds * K / (max_ms * 1e-3)

    # ---------------------------------------------------------------- latency of ONE isolated step, L2 flushed (informational)
    # Round 1's `value`: a 256 MiB memset before every step and one event pair around each launch.  It carries the
    # launch + event floor of an idle stream, which `null_floor_us` measures with an empty-handed 1-element kernel in the
    # same place; kept so that the two rounds can be compared.
    one = torch.zeros(1, device=dev)
    lat = {}
    for name, fn in (("null_floor_us", lambda s: one.add_(1.0)), ("step_us", lambda s: sim.step_with_actions(actions[s % (K + Wm)]))):
        reps = max(K, 50)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for s in range(3):
            flush.zero_()
            fn(s)
        torch.cuda.synchronize()
        for s in range(reps):
            flush.zero_()
            ev[s][0].record()
            fn(s)
            ev[s][1].record()
        torch.cuda.synchronize()
        ms = sorted(a.elapsed_time(b) for a, b in ev)
        lat[name] = {"mean": 1e3 * sum(ms) / len(ms), "median": 1e3 * ms[len(ms) // 2]}
    tl = torch.tensor([lat["step_us"]["mean"]], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(tl, op=dist.ReduceOp.MAX)
    flushed_value = total_worlds / (float(tl.item()) * 1e-6)

    # ---------------------------------------------------------------- L2-warm, graph-launched (informational)
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream(dev)
    with torch.cuda.stream(side):
        for s in range(3):
            sim.step_with_actions(actions[s])
    torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=side):
        for s in range(K):
            sim.step_with_actions(actions[Wm + s])
    g.replay()
    barrier()
    with sampler:
        e0.record()
        g.replay()
        e1.record()
        barrier()
    warm_ms = e0.elapsed_time(e1)
    t = torch.tensor([warm_ms], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    warm_value = total_worlds * K / (float(t.item()) * 1e-3)

    # ---------------------------------------------------------------- bb_step_n: the engine's own graph replay (informational)
    # K steps as ONE CUDA-graph launch with programmatic-dependent-launch edges between the step kernels; the Action column is
    # whatever hardCodeDefenseSystem leaves in it (the scripts/run.py loop: no host work between steps), L2 warm.
    sim.step_n(K)
    barrier()
    with sampler:
        e0.record()
        sim.step_n(K)
        e1.record()
        barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    step_n_value = total_worlds * K / (float(t.item()) * 1e-3)

    # ---------------------------------------------------------------- fused K steps with K action batches (informational)
    # bb_step_n_with_actions: the same K random-action batches as `value`, but taken by ONE launch of the looping kernel
    # on one 8192-world shard (BASELINE.md section 3's "fused-K-steps kernel"); the shard's rows stay in L1 / L2 between steps.
    fused_acts = actions[Wm:Wm + K].contiguous()
    sim.step_n_with_actions(fused_acts)
    barrier()
    with sampler:
        e0.record()
        sim.step_n_with_actions(fused_acts)
        e1.record()
        barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    fused_k_value = total_worlds * K / (float(t.item()) * 1e-3)

    # ---------------------------------------------------------------- e2e: host buffers through the C ABI
    # Every step takes its own action batch from pinned host memory (a ring of pre-filled pinned
    # buffers: the producer is not what is measured) and returns reward + done into pinned host memory.
    ring = min(K + Wm, 32)
    h_act = [actions[s].cpu().pin_memory() for s in range(ring)]
    h_rew = torch.empty((n, 2), dtype=torch.float32).pin_memory()
    h_done = torch.empty((n, 2), dtype=torch.float32).pin_memory()
    np_act = [a.numpy() for a in h_act]
    h_act8 = [actions[s].to(torch.uint8).cpu().pin_memory() for s in range(ring)]   # packed rows (declared extension)
    np_act8 = [a.numpy() for a in h_act8]
    np_rew, np_done = h_rew.numpy(), h_done.numpy()
    h_obs = torch.empty((n, 2, 128), dtype=torch.float32).pin_memory()
    np_obs = h_obs.numpy()
    sampler.pause()   # no NVML polling thread next to the latency-bound host loops (clocks were sampled in the regions above)

    def timed_host_loop(step_fn, with_obs=False, acts=None):
        acts = np_act if acts is None else acts
        extra = (np_obs,) if with_obs else ()
        Ke = max(K, 200)  # host loops are wall-clock timed: at least 200 steps so that one scheduler hiccup does not decide the number
        if step_fn == sim.serve_step:   # argument checks / pointer look-ups once per ring slot, not once per step
            calls = [sim.bind_serve_step(a, np_rew, np_done, *extra) for a in acts]
        else:
            calls = [(lambda a=a: step_fn(a, np_rew, np_done, *extra)) for a in acts]
        for s in range(max(Wm, 20)):
            calls[s % ring]()
        if world_size > 1:
            dist.barrier()
        windows = []              # five windows of Ke steps, the MEDIAN window counts: one host scheduler hiccup does not decide the number
        seq = [calls[(Wm + s) % ring] for s in range(Ke)]   # the ring order, resolved outside the timed loop
        for _ in range(5):
            t0 = time.perf_counter()
            for call in seq:
                call()                    # returns when reward + done (+ obs) are in host memory
            windows.append(time.perf_counter() - t0)
        dt = sorted(windows)[len(windows) // 2]
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world_size > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return total_worlds * Ke / float(tt.item())

    # (a) launch per step: bb_step_host = one kernel launch + one stream wait per step
    e2e_launch_value = timed_host_loop(sim.step_host)
    barrier()
    # (b) persistent server: bb_serve_step = doorbell in mapped pinned memory, no launch / stream wait per step
    e2e_obs_launch_value = timed_host_loop(sim.step_host, with_obs=True)
    barrier()
    e2e_launch8_value = timed_host_loop(sim.step_host, acts=np_act8)
    barrier()
    if args.no_serve:   # profiler runs: a resident kernel cannot live under ncu's serialised replay
        e2e_serve_value, e2e_obs_serve_value, e2e_serve8_value = None, None, None
    else:
        sim.serve_begin()
        e2e_serve_value = timed_host_loop(sim.serve_step)
        e2e_serve8_value = timed_host_loop(sim.serve_step, acts=np_act8)
        e2e_obs_serve_value = timed_host_loop(sim.serve_step, with_obs=True)
        sim.serve_end()
        barrier()
    # the headline is whichever host path is fastest on this box at this N (the server spins one host thread per rank;
    # with many ranks on few cores launch-per-step can win): all are reported
    paths = {"bb_serve_step_u8": e2e_serve8_value, "bb_step_host_u8": e2e_launch8_value,
             "bb_serve_step": e2e_serve_value, "bb_step_host": e2e_launch_value}
    e2e_path = max((k for k in paths if paths[k] is not None), key=lambda k: paths[k])
    e2e_value = paths[e2e_path]
    e2e_i32_value = max(v for v in (e2e_launch_value, e2e_serve_value) if v is not None)
    e2e_obs_value = max(v for v in (e2e_obs_launch_value, e2e_obs_serve_value) if v is not None)
    h2d, d2h = n * (12 if e2e_path.endswith("_u8") else 48), n * 16

    # ---------------------------------------------------------------- roofline
    peak, peak_src = measured_peak_gbs()
    avg_kernel_s = (dev_ms / K) * 1e-3
    achieved = n * BYTES_PER_WORLD_STEP / avg_kernel_s / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic(f"bb_step_kernel@{n}"), "kernel": "bb_step_kernel", "worlds_per_launch": n,
                "algorithmic_bytes_per_launch": n * BYTES_PER_WORLD_STEP, "peak_source": peak_src,
                "traffic_note": "ncu --set full capture with ncu's default cache control (caches flushed before the launch, like the "
                                "timed launches): DRAM reads = the state + action rows; the 11.7 MB of stores stay in the L2 for the "
                                "lifetime of the single profiled launch, so ncu's DRAM write bytes read 0 (in the timed cycle over the "
                                "shards they are written back while later launches run)",
                "note": "8192 worlds (15 MB) is latency/launch-bound by construction; see roofline_hbm_bound"}

    roofline_large = None
    cpu_baseline = None
    if rank == 0 and n_gpus == 1:
        # HBM-bound regime: 2^20 worlds on one GPU, same kernel
        del sim, shard, shards
        big = ShardedSimulator(LARGE_WORLDS, rank=0, world_size=1, gpu_id=local_rank)
        bsim = big.sim
        # Step time depends on the episode phase (every world shares one RNG stream and spawn, so shots and
        # resets cluster on the same steps): time a long window with a fresh action batch per step and
        # report the MEAN; median / p90 show the quiet and the event-heavy steps.
        bact = torch.empty((4, LARGE_WORLDS, 2, 6), dtype=torch.int32, device=dev)
        big_warm, reps = 64, 256
        for s in range(big_warm):
            big.fill_actions(SEED, s, out=bact[s % 4])
            bsim.step_with_actions(bact[s % 4])
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        torch.cuda.synchronize()
        with sampler:
            for s in range(reps):
                big.fill_actions(SEED, big_warm + s, out=bact[s % 4])   # untimed: outside the events
                ev[s][0].record()
                bsim.step_with_actions(bact[s % 4])
                ev[s][1].record()
            torch.cuda.synchronize()
        ms = sorted(a.elapsed_time(b) for a, b in ev)
        avg = sum(ms) / len(ms) * 1e-3
        ach = LARGE_WORLDS * BYTES_PER_WORLD_STEP / avg / 1e9
        roofline_large = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                          "traffic": ncu_traffic(f"bb_step_kernel@{LARGE_WORLDS}"), "worlds_per_launch": LARGE_WORLDS,
                          "ms_per_launch": avg * 1e3, "ms_per_launch_median": ms[len(ms) // 2], "ms_per_launch_p90": ms[int(len(ms) * 0.9)],
                          "launches_timed": reps, "world_steps_per_s": LARGE_WORLDS / avg,
                          "note": "same kernel at 2^20 worlds (1.96 GB per launch > L2, no flush needed); mean over 256 consecutive steps "
                                  "after 64 warm-up steps, fresh random actions each step; median = steps without shots/resets, "
                                  "p90 = steps where most warps also run the shot / reset paths"}
        del bsim, big, bact
        threads = os.cpu_count() or 1
        cv, cdt, cpu_steps, cnote = time_cpu_arm_for(args.worlds_per_gpu, args.cpu_seconds, threads)
        cpu_baseline = {"value": cv, "unit": UNIT, "cores": threads, "kind": "port",
                        "sample": f"{args.worlds_per_gpu} worlds x {cpu_steps} steps in one call, {cdt:.1f} s of CPU work on {threads} threads; "
                                  f"{cnote} (the reference's Madrona CPU executor cannot be built: submodule absent)"}

    # ---------------------------------------------------------------- informational keys for the other BASELINE.json configs
    extras = {}
    if not args.no_extras:
        if n_gpus > 1 and 65536 % n_gpus == 0:
            # configs[2] as STRONG scaling: 65 536 worlds / N per GPU (the weak headline above matches it only at N = 8)
            extras["strong_65536"] = rotating_rate(65536 // n_gpus, 65536, n_gpus, rank, local_rank, K, Wm, 0.2, seed_base=1000)
            extras["strong_65536"]["what"] = "BASELINE configs[2]: 65 536 worlds sharded over the N GPUs, same measurement as `value`"
        if n_gpus == 1:
            # the metric's second size ("at 8192/64K worlds") on ONE GPU, measured exactly like `value`: 65 536 worlds per launch,
            # K launches as one graph replay over rotating shards (rows cold), with its HBM fraction
            extras["worlds_65536"] = rotating_rate(65536, 65536, 1, rank, local_rank, K, Wm, 0.2, seed_base=1000)
            extras["worlds_65536"]["what"] = ("65 536 worlds on one GPU, same kernel and same measurement as `value` (K launches over "
                                              "rotating shards larger than L2); hbm_frac = 1872 B x 65 536 / time per launch / measured peak")
        if n_gpus == 8:
            extras["stress_2p20"] = stress_rate(1 << 20, n_gpus, rank, local_rank)
        # configs[3]: data-parallel PPO (scripts/ppo.py:226-376 one process per GPU), gradient all-reduce with NCCL (default) and
        # fused into the Adam kernel over NVLink peer memory; the parity of the two exchange paths is checked in-process
        from madrona_basketball_b200 import ppo_train

        ppo = {}
        r_nccl = ppo_train.run_iterations(iters=args.ppo_iters, warmup=3, allreduce="nccl")
        ppo["env_steps_per_s_nccl"] = r_nccl["env_steps_per_s"]
        ppo["ms_per_iter_rank0"] = r_nccl["ms_per_iter_rank0"]
        ppo["replica_param_spread_nccl"] = r_nccl["replica_param_spread"]
        if n_gpus > 1:
            try:
                ppo.update(ppo_train.check_p2p_against_nccl(64))
                r_p2p = ppo_train.run_iterations(iters=args.ppo_iters, warmup=3, allreduce="p2p")
                ppo["env_steps_per_s_p2p"] = r_p2p["env_steps_per_s"]
                ppo["replica_param_spread_p2p"] = r_p2p["replica_param_spread"]
                if "note" in r_p2p:
                    ppo["note"] = r_p2p["note"]
            except RuntimeError as e:   # CUDA IPC unavailable on this box: the NCCL path is the product default
                ppo["p2p_unavailable"] = str(e)[:200]
        ppo["what"] = (f"{args.ppo_iters} PPO iterations of 8192 envs x 32 steps per GPU (rollout + GAE + 4 epochs x 4 minibatches, ppo.py "
                       "defaults) through bb_rollout / bb_gae / bb_ppo_grad / bb_ppo_adam; whole-job env-steps/s, wall clock, max over ranks")
        extras["ppo"] = ppo

    clocks = sampler.report()
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": K, "warmup": Wm,
            "ms_per_step": max_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_string(args.worlds_per_gpu, n_gpus),
                       "l2": f"inputs larger than L2: every launch advances the next of {R} independent {n}-world shards on this GPU "
                             f"({R * n * BYTES_PER_WORLD_STEP / 1e6:.0f} MB of rows per cycle against a {l2_bytes / 1e6:.0f} MB L2), so each of the K timed "
                             "launches reads its rows from HBM; the K step kernels are one CUDA-graph replay between one pair of CUDA "
                             f"events, enqueued behind a {SPACER_US} us spin kernel so that the window is device time",
                       "launch": f"one CUDA graph of the K = {K} step kernels (captured from the engine's ordinary bb_step_with_actions calls, "
                                 "programmatic dependent launch edges), replayed once in the timed window; stream_launches = the same K "
                                 "steps as K ordinary stream launches",
                       "dephase": f"before anything is timed shard i is advanced i x {dephase} extra steps: the worlds' game clocks run in lock "
                                  "step, so shots / resets come in waves (a step costs 1.0x to 1.6x depending on the phase); with the "
                                  "shards spread over the cycle the K timed launches sample all of it instead of one phase by chance",
                       "prewarm": f"{n_pw} untimed launches of the same cycle ({PREWARM_S} s), one untimed replay of the graph and "
                                  f"{n_fill} ordinary launches (>= W, completing the shard cycle) before the timed window, so that it does "
                                  "not sit on the clock / launch-path ramp of an idle GPU",
                       "parallelism": f"world-sharded x{n_gpus}, no collective in the step"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "path": e2e_path,
                    "what": "per step: action rows read from pinned host memory over PCIe, fused step, reward + done written back "
                            "to pinned host memory, the call returns when they are there; synchronous, one step in flight; wall clock, median of "
                            "five 200-step windows.  The call is SimpleGridworldSimulator.bind_serve_step(...)() / step_host(...): the "
                            "C-ABI entry point with its pointers looked up once per buffer (the served paths go through the bound-call "
                            "glue csrc/bb_fastcall.c instead of ctypes, which costs ~1.5 us per call).  value = "
                            "the fastest of the four host paths below (named in `path`); the *_u8 paths take the rows PACKED, u8 "
                            "[N,2,6] (12 B/world, a declared extension of the C ABI: every field the policy emits is < 8), the "
                            "others the reference's int32 rows (48 B/world; int32_rows.value).  The 1024 B/world observation rows "
                            "STAY ON THE DEVICE: the reference's consumer of them is a torch policy on the same device "
                            "(env.py:75-79, ppo.py:66); see e2e_with_obs for the host-consumer case",
                    "int32_rows": {"value": e2e_i32_value, "unit": UNIT, "h2d_bytes_per_step": n * 48,
                                   "serve": e2e_serve_value, "launch_per_step": e2e_launch_value,
                                   "what": "the reference's Action dtype: bb_serve_step (persistent server kernel, doorbell in mapped pinned "
                                           "memory, no launch / stream wait per step) and bb_step_host (one launch + one stream wait per step)"},
                    "u8_rows": {"serve": e2e_serve8_value, "launch_per_step": e2e_launch8_value, "unit": UNIT, "h2d_bytes_per_step": n * 12,
                                "what": "bb_serve_step_u8 / bb_step_host_u8"}},
            "e2e_with_obs": {"value": e2e_obs_value, "unit": UNIT, "h2d_bytes_per_step": n * 48, "d2h_bytes_per_step": d2h + n * 1024,
                             "serve": e2e_obs_serve_value, "launch_per_step": e2e_obs_launch_value,
                             "what": "same, with both agents' observation rows mirrored to pinned host memory too (a policy on the host)"},
            "gpu_launches": int(gpu_launches),
            "clocks": clocks,
            "roofline": roofline,
            "roofline_hbm_bound": roofline_large,
            "cpu_baseline": cpu_baseline,
            "step_latency_flushed": {"value": flushed_value, "unit": UNIT, "step_us": lat["step_us"], "null_floor_us": lat["null_floor_us"],
                                     "what": "round 1's `value`: ONE isolated launch after a 256 MiB L2-flush memset, one event pair per step "
                                             "(max over ranks of the mean); null_floor_us = a 1-element torch kernel timed the same way on "
                                             "this box, i.e. the launch + event floor that number contains (informational)"},
            "stream_launches": {"value": stream_value, "unit": UNIT, "us_per_step": stream_dev_ms / K * 1e3, "gpu_launches": int(stream_launches),
                                "what": "the same K cold-row steps as K ordinary launches queued on the stream (round 2's earlier `value`): "
                                        "carries the stream's per-launch cost that the graph replay avoids (informational)"},
            "l2_warm_graph": {"value": warm_value, "unit": UNIT,
                              "what": "same K steps replayed as one CUDA graph, no L2 flush (informational)"},
            "graph_step_n": {"value": step_n_value, "unit": UNIT,
                             "what": "K steps through bb_step_n: ONE launch of the looping kernel (every warp takes its 16 worlds through "
                                     "the K steps in a row; round 1 / BB_STEP_N_GRAPH=1: a CUDA graph of K step kernels), device-resident "
                                     "actions (scripts/run.py loop), no L2 flush (informational)"},
            "fused_k_steps": {"value": fused_k_value, "unit": UNIT,
                              "what": "K steps with K random-action batches through bb_step_n_with_actions: one launch of the looping "
                                      "kernel on one shard, rows L1 / L2 resident between steps (informational)"},
            "wall_s_timed_region": wall,
        }
        line.update(extras)
        emit(line)
    if world_size > 1:
        dist.destroy_process_group()
    return 0


def _claim_stdout():
    """stdout carries exactly ONE JSON line: libraries that write to fd 1 on their own (NCCL prints its
    version banner there) are sent to stderr; emit() writes to the saved descriptor."""
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


_JSON_OUT = None


def emit(line: dict) -> None:
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--worlds-per-gpu", type=int, default=WORLDS_PER_GPU)
    ap.add_argument("--no-serve", action="store_true", help="e2e through bb_step_host only (use under ncu)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="CPU work of the cpu_baseline sample")
    ap.add_argument("--no-extras", action="store_true", help="skip strong_65536 / stress_2p20 / ppo (profiler runs)")
    ap.add_argument("--ppo-iters", type=int, default=20)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_engine(args)


if __name__ == "__main__":
    sys.exit(main())
