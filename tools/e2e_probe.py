#!/usr/bin/env python
"""Where the end-to-end step time goes (tuning aid): raw C-ABI calls in a tight loop."""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from madrona_basketball_b200.shard import ShardedSimulator

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
MARKS = (int(os.environ.get("BB_SERVE_FLAGS", "0")) & 4) != 0
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
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(400): dev_async()
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"400 async launches: host enqueue {(t1 - t) / 400 * 1e6:.2f} us each, until drained {(t2 - t) / 400 * 1e6:.2f} us each")
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

u8 = dact.to(torch.uint8).cpu().pin_memory(); p8 = u8.data_ptr()
loop("bb_step_host_u8 pinned packed rows", lambda: lib.bb_step_host_u8(h, p8, pr, pd, None))
lib.bb_serve_begin(h)
loop("bb_serve_step_u8 pinned packed rows", lambda: lib.bb_serve_step_u8(h, p8, pr, pd, None), sync=False)
loop("bb_serve_step_u8 packed rows only (no reward/done)", lambda: lib.bb_serve_step_u8(h, p8, None, None, None), sync=False)
# bench.py's e2e conditions: a fresh random batch every step (some tile always runs a shot / reset path), ring of 32 buffers
ring8 = []
for s_ in range(32):
    sh.fill_actions(1234, 100 + s_, out=dact)
    ring8.append(dact.to(torch.uint8).cpu().pin_memory())
ring_ptrs = [t_.data_ptr() for t_ in ring8]
ring_i = [0]
def serve_ring():
    ring_i[0] = (ring_i[0] + 1) & 31
    lib.bb_serve_step_u8(h, ring_ptrs[ring_i[0]], pr, pd, None)
loop("bb_serve_step_u8 ring of 32 random batches (bench e2e)", serve_ring, sync=False)
n8, = (u8.numpy(),)
loop("sim.serve_step(numpy u8) python wrapper", lambda: sim.serve_step(n8, nr, nd), sync=False)
def marks(name, fn):
    m = (ctypes.c_longlong * 8)()
    acc = np.zeros(8); wall = []
    for _ in range(300):
        t0 = time.perf_counter_ns(); fn(); wall.append(time.perf_counter_ns() - t0)
        lib.bb_serve_last_marks(h, m); acc += np.array(list(m), dtype=np.float64)
    acc /= 300
    print(f"{name}: host call {np.mean(wall) / 1e3:.2f} us (median {np.median(wall) / 1e3:.2f}); GPU marks (us after doorbell seen): "
          f"fence {acc[1] / 1e3:.2f}, released {acc[2] / 1e3:.2f}, blk0 tile start {acc[3] / 1e3:.2f} end {acc[4] / 1e3:.2f}, "
          f"last blk tile end {acc[5] / 1e3:.2f}, counted {acc[6] / 1e3:.2f}, before flag {acc[7] / 1e3:.2f}")
if MARKS: marks("serve u8 + reward/done", lambda: lib.bb_serve_step_u8(h, p8, pr, pd, None))
if MARKS: marks("serve no io", lambda: lib.bb_serve_step_u8(h, None, None, None, None))
# host-side enqueue cost of an asynchronous launch vs GPU time per launch
loop("bb_serve_step pinned (persistent server)", lambda: lib.bb_serve_step(h, pa, pr, pd, None), sync=False)
loop("bb_serve_step actions only", lambda: lib.bb_serve_step(h, pa, None, None, None), sync=False)
loop("bb_serve_step nothing (doorbell + step + completion)", lambda: lib.bb_serve_step(h, None, None, None, None), sync=False)
acc = []
for _ in range(200):
    lib.bb_serve_step(h, pa, pr, pd, None); acc.append(lib.bb_serve_last_gpu_ns(h))
print("GPU-side doorbell-seen -> last-block-done, pinned io: avg %.2f us min %.2f" % (sum(acc) / len(acc) / 1e3, min(acc) / 1e3))
acc = []
for _ in range(200):
    lib.bb_serve_step(h, None, None, None, None); acc.append(lib.bb_serve_last_gpu_ns(h))
print("GPU-side doorbell-seen -> last-block-done, no host io: avg %.2f us min %.2f" % (sum(acc) / len(acc) / 1e3, min(acc) / 1e3))
lib.bb_serve_end(h)
