#!/usr/bin/env python
"""Step-kernel throughput for the BASELINE.json configurations bench.py does not print (one process per GPU,
this rank's share; no collective in the step):

  configs[2]  65 536 worlds over N GPUs, random actions                       -> worlds/GPU = 65536 / N
  configs[4]  2^20 worlds over N GPUs, reset/collision stress: agent 1 is teleported next to agent 0 through the
              writable exported view every step, so agentCollision fires each step and the world resets every 2nd step

    python tools/bench_configs.py                                   # N = 1: full sizes on one GPU
    python -m torch.distributed.run --nproc-per-node 8 ... tools/bench_configs.py

CUDA events around every launch (the teleport write and the action fill are outside them), max over ranks of the
summed step time.  Prints one JSON line on rank 0."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from madrona_basketball_b200.shard import ShardedSimulator, env_rank

BYTES = 1872


def run(total, rank, ws, local, stress, steps=200, warm=20):
    dev = torch.device("cuda", local)
    sh = ShardedSimulator(total, rank=rank, world_size=ws, gpu_id=local)
    sim, n = sh.sim, sh.num_worlds
    acts = torch.empty((4, n, 2, 6), dtype=torch.int32, device=dev)
    pos = sim.agent_pos_tensor().to_torch()
    off = pos.new_tensor([0.05, 0.0, 0.0])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if n * BYTES < (512 << 20) else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for s in range(warm + steps):
        sh.fill_actions(1234, s, out=acts[s % 4])
        if stress:
            pos[:, 1] = pos[:, 0] + off
        if flush is not None:
            flush.zero_()
        if s >= warm:
            ev[s - warm][0].record()
        sim.step_with_actions(acts[s % 4])
        if s >= warm:
            ev[s - warm][1].record()
    torch.cuda.synchronize()
    ms = [a.elapsed_time(b) for a, b in ev]
    t = torch.tensor([sum(ms)], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot_ms = float(t.item())
    resets = int(sim.tensor("rng_count").to_torch().min().item())
    del sim, sh
    return {"worlds": total, "worlds_per_gpu": n, "steps": steps, "us_per_step": tot_ms / steps * 1e3,
            "world_steps_per_s": total * steps / (tot_ms * 1e-3), "hbm_gbs_per_gpu": n * BYTES * steps / (tot_ms * 1e-3) / 1e9,
            "l2": "flushed before every step" if flush is not None else "working set > L2",
            "min_rng_draws_per_world": resets}


def main():
    rank, local, ws = env_rank()
    torch.cuda.set_device(local)
    if ws > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    try:
        peak = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    out = {"n_gpus": ws, "hbm_peak_gbs": peak,
           "config2_65536_worlds_random": run(65536, rank, ws, local, False),
           "config4_2p20_worlds_reset_collision_stress": run(1 << 20, rank, ws, local, True),
           "config4_2p20_worlds_random_for_comparison": run(1 << 20, rank, ws, local, False)}
    for v in out.values():
        if isinstance(v, dict):
            v["hbm_frac"] = v["hbm_gbs_per_gpu"] / peak
    if rank == 0:
        print(json.dumps(out))
    if ws > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
