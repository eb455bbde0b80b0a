"""Multi-GPU check + timing of the gradient all-reduce fused into the optimiser step (bb_ppo_adam_p2p, include/bb_ppo.h).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/check_p2p.py

Two optimisers per rank start from the same parameters and see the same per-rank random gradients: one goes through
NCCL all-reduce + bb_ppo_adam, the other through bb_ppo_adam_p2p.  Checks: (1) the p2p parameters are bit-identical on
every rank (rank-order sum), (2) they track the NCCL path within float32 summation-order noise, (3) a rank that does
not show up makes the peers give up after the timeout and raise the error flag WITHOUT applying the step (no hang).
Rank 0 prints one JSON line with the timings.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from madrona_basketball_b200.policy_torch import Agent  # noqa: E402
from madrona_basketball_b200.ppo_fused import NUM_PARAMS, FusedPPO  # noqa: E402


def main() -> None:
    rank, ws = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    steps = int(os.environ.get("CHECK_P2P_STEPS", 200))

    def make():
        torch.manual_seed(1234)
        return FusedPPO(Agent(128, num_channels=32).to(dev), lr=3e-4, max_grad_norm=1.0)

    a, b = make(), make()
    b.enable_p2p(timeout_ms=2000)
    assert torch.equal(a.flat, b.flat)
    gen = torch.Generator(device=dev)
    gen.manual_seed(100 + rank)
    grads = torch.randn(32, NUM_PARAMS, device=dev, generator=gen) * 0.05

    # ---- 1, 2: parity over 64 steps ----
    for i in range(64):
        g = grads[i % 32]
        a.grad.copy_(g)
        dist.all_reduce(a.grad, op=dist.ReduceOp.SUM)
        a.apply(1.0 / ws)
        b.grad.copy_(g)
        b.apply_p2p()
    torch.cuda.synchronize()
    assert int(b.p2p_error.item()) == 0, "p2p timed out"
    assert int(b.step.item()) == 64
    gathered = [torch.empty_like(b.flat) for _ in range(ws)]
    dist.all_gather(gathered, b.flat)
    for r in range(ws):
        assert torch.equal(gathered[r], gathered[0]), f"p2p parameters of rank {r} differ from rank 0"
    max_diff = float((a.flat - b.flat).abs().max())
    norm_diff = float((a.grad_norm - b.grad_norm).abs().max())
    assert max_diff < 2e-5, max_diff       # 64 Adam steps of lr 3e-4: a sign flip of a ~0 gradient moves a weight by 2 * lr at most
    assert norm_diff < 1e-5, norm_diff

    # ---- timing: gradient -> averaged + applied, back to back on the stream ----
    def timed(fn):
        for _ in range(20):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / steps * 1e3], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def nccl_step():
        dist.all_reduce(a.grad, op=dist.ReduceOp.SUM)
        a.apply(1.0 / ws)

    us_nccl = timed(nccl_step)
    us_p2p = timed(b.apply_p2p)
    us_local = timed(lambda: a.apply(1.0))
    torch.cuda.synchronize()
    assert int(b.p2p_error.item()) == 0

    # ---- 3: a missing peer must not hang anybody ----
    c = make()
    c.enable_p2p(timeout_ms=100)
    before = c.flat.clone()
    gave_up = None
    if ws > 1:
        if rank != ws - 1:          # the last rank stays away
            c.grad.copy_(grads[0])
            c.apply_p2p()
            torch.cuda.synchronize()
            gave_up = int(c.p2p_error.item())
            assert gave_up == 1, "the kernel should have given up"
            assert torch.equal(c.flat, before), "a timed-out step must not be applied"
            assert int(c.step.item()) == 0
        torch.cuda.synchronize()
        dist.barrier()
        c._p2p_error_event = None    # the flag is expected here
    c.disable_p2p()
    b.disable_p2p()
    if rank == 0:
        print(json.dumps({"check": "bb_ppo_adam_p2p", "n_gpus": ws, "steps": steps, "max_param_diff_vs_nccl": max_diff,
                          "us_nccl_allreduce_plus_adam": round(us_nccl, 2), "us_p2p_fused": round(us_p2p, 2),
                          "us_adam_alone": round(us_local, 2), "timeout_raises_flag": gave_up == 1 if ws > 1 else None}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
