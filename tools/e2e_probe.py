#!/usr/bin/env python
"""Where the end-to-end step time goes (tuning aid): raw C-ABI calls in a tight loop."""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from madrona_basketball_b200.shard import ShardedSimulator

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
sh = ShardedSimulator(n, rank=0, world_size=1, gpu_id=0)
sim = sh.sim
lib, h = sim._lib, sim._h
dact = torch.empty((n, 2, 6), dtype=torch.int32, device="cuda")
sh.fill_actions(1234, 0, out=dact)
pact = dact.cpu().pin_memory(); prew = torch.empty((n, 2)).pin_memory(); pdone = torch.empty((n, 2)).pin_memory()
uact = dact.cpu().numpy().copy(); urew = np.empty((n, 2), np.float32); udone = np.empty((n, 2), np.float32)
stream = torch.cuda.current_stream().cuda_stream
R = 2000

def loop(name, fn, sync=True):
    for _ in range(50): fn()
    if sync: torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(R): fn()
    if sync: torch.cuda.synchronize()
    print(f"{name:55s} {(time.perf_counter() - t) / R * 1e6:7.2f} us/step")

def dev_async(): lib.bb_step_with_actions(h, dact.data_ptr(), stream)
loop("device actions, async launches (no wait)", dev_async)
def dev_sync():
    lib.bb_step_with_actions(h, dact.data_ptr(), stream); torch.cuda.current_stream().synchronize()
loop("device actions, launch + stream sync", dev_sync)
pa, pr, pd = pact.data_ptr(), prew.data_ptr(), pdone.data_ptr()
loop("bb_step_host pinned (zero-copy), raw pointers", lambda: lib.bb_step_host(h, pa, pr, pd, None))
loop("bb_step_host pinned actions only", lambda: lib.bb_step_host(h, pa, None, None, None))
loop("bb_step_host pinned reward+done only", lambda: lib.bb_step_host(h, None, pr, pd, None))
loop("bb_step_host nothing (launch + wait)", lambda: lib.bb_step_host(h, None, None, None, None))
ua, ur, ud = uact.ctypes.data, urew.ctypes.data, udone.ctypes.data
loop("bb_step_host pageable (staged copies)", lambda: lib.bb_step_host(h, ua, ur, ud, None))
na, nr, nd = pact.numpy(), prew.numpy(), pdone.numpy()
loop("sim.step_host(numpy pinned) python wrapper", lambda: sim.step_host(na, nr, nd))

lib.bb_serve_begin(h)
loop("bb_serve_step pinned (persistent server)", lambda: lib.bb_serve_step(h, pa, pr, pd, None), sync=False)
loop("bb_serve_step actions only", lambda: lib.bb_serve_step(h, pa, None, None, None), sync=False)
loop("bb_serve_step nothing (doorbell + step + completion)", lambda: lib.bb_serve_step(h, None, None, None, None), sync=False)
lib.bb_serve_last_gpu_ns.restype = ctypes.c_longlong; lib.bb_serve_last_gpu_ns.argtypes = [ctypes.c_void_p]
acc = []
for _ in range(200):
    lib.bb_serve_step(h, pa, pr, pd, None); acc.append(lib.bb_serve_last_gpu_ns(h))
print("GPU-side doorbell-seen -> last-block-done, pinned io: avg %.2f us min %.2f" % (sum(acc) / len(acc) / 1e3, min(acc) / 1e3))
acc = []
for _ in range(200):
    lib.bb_serve_step(h, None, None, None, None); acc.append(lib.bb_serve_last_gpu_ns(h))
print("GPU-side doorbell-seen -> last-block-done, no host io: avg %.2f us min %.2f" % (sum(acc) / len(acc) / 1e3, min(acc) / 1e3))
lib.bb_serve_end(h)
