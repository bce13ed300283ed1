#!/usr/bin/env python3
"""BASELINE config 5 as stated: the Panda two-task hierarchy (config 2) at B_total in {1e4 ... 1e7} robots on 1 / 2 / 4 / 8 B200s,
next to the CPU port on all host cores at 1e4 and 1e5 robots.  One JSON line per (B_total, GPUs) point.

    python tools/sweep_config5.py                                       # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/sweep_config5.py
    python tools/sweep_config5.py --one-process N                       # the same sweep through ONE process: osc_batch_create_multi

Device-resident timing: CUDA events around each of 20 ticks after 5 warm-up ticks, no L2 flush (every batch >= 3e4 robots is larger
than L2; the small ones are the latency floor the sweep is meant to show), max over ranks of the summed tick time.  The
one-process mode has one stream per device inside the library, so it is timed by the host clock around [K ticks + synchronize]."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import osc_abi as abi  # noqa: E402
import osc_b200  # noqa: E402
import workloads  # noqa: E402

SIZES = (10000, 30000, 100000, 300000, 1000000, 3000000, 10000000)


def chain_of(name="panda"):
    with open(os.path.join(ROOT, "fixtures", name + ".chain.json")) as f:
        return json.load(f)


def build(chain, B, base, **kw):
    rep = (B + len(base["q"]) - 1) // len(base["q"])
    tile = lambda a: np.tile(a, (rep, 1))[:B]
    b = osc_b200.Batch(chain, B, q0=[j["q0"] for j in chain["joints"]], **kw)
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = b.add_joint_task(abi.joint_task_desc(chain["n"]))
    b.set(abi.F_Q, tile(base["q"])); b.set(abi.F_DQ, tile(base["dq"])); b.update_model()
    b.set(abi.F_GOAL_POSITION, tile(base["goal_pos"]), t0); b.set(abi.F_GOAL_ORIENTATION, tile(base["goal_ori"]), t0)
    b.set(abi.F_JT_GOAL_POSITION, tile(base["jt_goal"]), t1)
    return b


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--one-process", type=int, default=0, help="number of devices driven by ONE process through osc_batch_create_multi")
    ap.add_argument("--max-total", type=int, default=10000000)
    ap.add_argument("--ticks", type=int, default=20)
    args = ap.parse_args()
    import torch
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    chain = chain_of()
    base = workloads.config2(chain, 65536, seed=5)
    if args.one_process:
        G = args.one_process
        for Bt in SIZES:
            if Bt > args.max_total:
                continue
            b = build(chain, Bt, base, devices=list(range(G)))
            for _ in range(5):
                b.tick(True, fetch=False)
            b.synchronize()
            t = time.perf_counter()
            for _ in range(args.ticks):
                b.tick(True, fetch=False)
            b.synchronize()
            el = time.perf_counter() - t
            print(json.dumps({"config": 5, "robots_total": Bt, "gpus": G, "path": "one process, osc_batch_create_multi (one host thread + stream per device)",
                              "ms_per_tick": el / args.ticks * 1e3, "robot_ticks_per_s": Bt * args.ticks / el, "timing": "host clock around %d ticks + synchronize" % args.ticks}), flush=True)
            b.close()
        return
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    for Bt in SIZES:
        if Bt > args.max_total:
            continue
        lo, hi = Bt * rank // world, Bt * (rank + 1) // world
        b = build(chain, hi - lo, base, device=local)
        b.set_stream(torch.cuda.current_stream().cuda_stream)
        for _ in range(5):
            b.tick(True, fetch=False)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.ticks)]
        for s, e in ev:
            s.record(); b.tick(True, fetch=False); e.record()
        torch.cuda.synchronize()
        ms = np.array([s.elapsed_time(e) for s, e in ev])
        tot = torch.tensor([float(ms.sum())], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        fb = b.tick_fallback_count()
        if rank == 0:
            print(json.dumps({"config": 5, "robots_total": Bt, "gpus": world, "path": "torchrun, one batch per rank", "ms_per_tick": float(tot.item()) / args.ticks,
                              "p50_us_rank0": float(np.median(ms) * 1e3), "robot_ticks_per_s": Bt * args.ticks / (float(tot.item()) * 1e-3),
                              "fallback_robots_rank0": int(fb)}), flush=True)
        b.close()
        torch.cuda.empty_cache()
    if rank == 0 and world == 1:   # the CPU port on all host cores at 1e4 and 1e5 robots (linear beyond)
        import orc
        nthreads = os.cpu_count() or 1
        for Bc in (10000, 100000):
            rep = (Bc + 65535) // 65536
            tile = lambda a: np.tile(a, (rep, 1))[:Bc]
            o = orc.OracleBatch(chain, Bc, q0=[j["q0"] for j in chain["joints"]], nthreads=nthreads)
            t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07))); t1 = o.add_joint_task(abi.joint_task_desc(7))
            o.set(abi.F_Q, tile(base["q"])); o.set(abi.F_DQ, tile(base["dq"])); o.update_model()
            o.set(abi.F_GOAL_POSITION, tile(base["goal_pos"]), t0); o.set(abi.F_GOAL_ORIENTATION, tile(base["goal_ori"]), t0)
            o.set(abi.F_JT_GOAL_POSITION, tile(base["jt_goal"]), t1)
            o.L.orc_tick(o.h, 1, nthreads)
            ticks = 20 if Bc == 10000 else 4
            t = time.perf_counter()
            for _ in range(ticks):
                o.L.orc_tick(o.h, 1, nthreads)
            el = time.perf_counter() - t
            print(json.dumps({"config": 5, "robots_total": Bc, "gpus": 0, "path": "CPU port (oracle, g++ -O2), %d host threads" % nthreads,
                              "ms_per_tick": el / ticks * 1e3, "robot_ticks_per_s": Bc * ticks / el}), flush=True)
            del o
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
