"""Where the end-to-end step of bench.py spends its time: each C-ABI call of the step timed on the host with a device
synchronise after it (serialised), next to the pipelined step as the bench runs it."""
import os, sys, time
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200"))
sys.path.insert(0, ROOT)
import osc_abi as abi, osc_b200
import bench

chain = bench.load_chain()
n, B = chain["n"], 65536
batch, w, t0, t1 = bench.build_batch(osc_b200, abi, chain, B, 0, 2, 64)
if "--no-otg" in sys.argv:
    pass
batch.tick(True, fetch=False)
real = torch.float64
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a.T)).to(real).pin_memory()
h_in = [(abi.F_Q, -1, pin(w["q"])), (abi.F_DQ, -1, pin(w["dq"])), (abi.F_GOAL_POSITION, t0, pin(w["goal_pos"])),
        (abi.F_GOAL_ORIENTATION, t0, pin(w["goal_ori"])), (abi.F_JT_GOAL_POSITION, t1, pin(w["jt_goal"]))]
h_tau = torch.empty((n, B), dtype=real).pin_memory()
L, h = batch.L, batch.h
sync = torch.cuda.synchronize


def timed(fn, reps=30):
    for _ in range(3):
        fn(); sync()
    ts = []
    for _ in range(reps):
        sync(); t = time.perf_counter(); fn(); sync(); ts.append(time.perf_counter() - t)
    return 1e3 * float(np.median(ts))


for f, task, t in h_in:
    ms = timed(lambda: L.osc_set_field(h, task, f, t.data_ptr(), abi.LAYOUT_SOA, abi.MEM_HOST))
    print("set_field %3d  %5.2f MB  %.3f ms" % (f, t.numel() * 8e-6, ms))
print("all 5 set_field      %.3f ms" % timed(lambda: [L.osc_set_field(h, task, f, t.data_ptr(), abi.LAYOUT_SOA, abi.MEM_HOST) for f, task, t in h_in]))
print("tick, no fetch       %.3f ms" % timed(lambda: batch.tick(True, fetch=False)))
print("tick + tau to host   %.3f ms" % timed(lambda: L.osc_tick(h, 1, h_tau.data_ptr(), abi.LAYOUT_SOA, abi.MEM_HOST)))
d_tau = torch.empty((n, B), dtype=real, device="cuda")
print("torch d2h 3.67 MB    %.3f ms" % timed(lambda: h_tau.copy_(d_tau, non_blocking=True)))
big = torch.empty(int(17.3e6) // 8, dtype=real).pin_memory(); dbig = torch.empty_like(big, device="cuda")
print("torch h2d 17.3 MB    %.3f ms" % timed(lambda: dbig.copy_(big, non_blocking=True)))


def step():
    for f, task, t in h_in:
        L.osc_set_field(h, task, f, t.data_ptr(), abi.LAYOUT_SOA, abi.MEM_HOST)
    L.osc_tick(h, 1, h_tau.data_ptr(), abi.LAYOUT_SOA, abi.MEM_HOST)


for _ in range(5):
    step()
sync(); t = time.perf_counter()
for _ in range(50):
    step()
sync(); print("whole step           %.3f ms" % ((time.perf_counter() - t) / 50 * 1e3))
