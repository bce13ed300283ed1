#!/usr/bin/env python3
"""Throughput of BASELINE configs 3, 4 and 5 on one B200 (config 2 is bench.py).  One JSON line per measurement.

    python tools/bench_configs.py [--max-batch 3000000]

Device-resident timing with CUDA events around the control calls only (inputs already on the device), after warm-up ticks.
These are companion numbers for BASELINE.md section 3; the headline contract lives in bench.py.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import osc_abi as abi  # noqa: E402
import osc_b200  # noqa: E402
import workloads  # noqa: E402


def chain_of(name):
    with open(os.path.join(ROOT, "fixtures", name + ".chain.json")) as f:
        return json.load(f)


def timed(torch, fn, ticks, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(ticks)]
    for s, e in ev:
        s.record(); fn(); e.record()
    torch.cuda.synchronize()
    return np.array([s.elapsed_time(e) for s, e in ev])


def config5(torch, chain, B):
    w = workloads.config2(chain, min(B, 65536), seed=5)
    rep = (B + len(w["q"]) - 1) // len(w["q"])
    tile = lambda a: np.tile(a, (rep, 1))[:B]
    b = osc_b200.Batch(chain, B, q0=[j["q0"] for j in chain["joints"]])
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = b.add_joint_task(abi.joint_task_desc(chain["n"]))
    b.set_stream(torch.cuda.current_stream().cuda_stream)
    b.set(abi.F_Q, tile(w["q"])); b.set(abi.F_DQ, tile(w["dq"])); b.update_model()
    b.set(abi.F_GOAL_POSITION, tile(w["goal_pos"]), t0); b.set(abi.F_GOAL_ORIENTATION, tile(w["goal_ori"]), t0)
    b.set(abi.F_JT_GOAL_POSITION, tile(w["jt_goal"]), t1)
    ms = timed(torch, lambda: b.tick(True, fetch=False), 20)
    fb = b.tick_fallback_count()
    del b
    return {"config": 5, "robots": B, "ms_per_tick": float(ms.mean()), "p50_us": float(np.median(ms) * 1e3),
            "robot_ticks_per_s": B / (ms.mean() * 1e-3), "fallback_robots": int(fb)}


def config3(torch, chain, B=262144, ticks=100):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_config3_force_popc_gpu as T
    from common import q0_of, random_q
    rng = np.random.default_rng(3)
    q = random_q(chain, B, rng, 0.3); dq = rng.uniform(-0.5, 0.5, (B, 7)); gq = q + rng.normal(0, 0.1, (B, 7))
    b = osc_b200.Batch(chain, B, q0=q0_of(chain))
    b.set_stream(torch.cuda.current_stream().cuda_stream)
    t0, _ = T._configure(b, chain, q, dq, 3, gq)
    f, _ = workloads.config3_wrench(B, 0)
    fd = torch.from_numpy(np.ascontiguousarray(f.T)).cuda(); md = torch.zeros_like(fd)
    L, h = b.L, b.h

    def step_per_call():
        L.osc_update_model(h); L.osc_update_task_models(h)
        L.osc_mft_update_sensed_force_and_moment(h, t0, fd.data_ptr(), md.data_ptr(), abi.LAYOUT_SOA, abi.MEM_DEVICE)
        L.osc_compute_torques(h, 1, None, abi.LAYOUT_SOA, abi.MEM_DEVICE)

    def step_tick():   # the control cycle around osc_tick: sensed wrench in, fused tick (k_mft_sensed + k_tick_fused3 + k_tick_tail)
        L.osc_mft_update_sensed_force_and_moment(h, t0, fd.data_ptr(), md.data_ptr(), abi.LAYOUT_SOA, abi.MEM_DEVICE)
        L.osc_tick(h, 1, None, abi.LAYOUT_SOA, abi.MEM_DEVICE)
    out = []
    for name, fn in (("per-call kernels (API-faithful route)", step_per_call), ("osc_tick: k_mft_sensed + k_tick_fused3 + k_tick_tail", step_tick)):
        ms = timed(torch, fn, ticks)
        out.append({"config": 3, "robots": B, "ms_per_tick": float(ms.mean()), "p50_us": float(np.median(ms) * 1e3),
                    "robot_ticks_per_s": B / (ms.mean() * 1e-3), "route": name, "fused": bool(b.tick_is_fused()),
                    "fallback_robots": int(b.tick_fallback_count())})
    return out


def config4(torch, name, B=65536, ticks=50):
    import test_singular_gpu as S
    from common import q0_of
    chain = chain_of(name)
    rng = np.random.default_rng(4)
    base = S._panda_q(rng) if name == "panda" else S._puma_q(rng)
    q = np.tile(base, ((B + len(base) - 1) // len(base), 1))[:B]
    n = chain["n"]
    b = osc_b200.Batch(chain, B, q0=q0_of(chain))
    b.set_stream(torch.cuda.current_stream().cuda_stream)
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0)))
    if n > 6:
        b.add_joint_task(abi.joint_task_desc(n))
    b.set(abi.F_Q, q); b.set(abi.F_DQ, np.zeros((B, n))); b.update_model(); b.reinitialize_tasks()
    ms = timed(torch, lambda: b.tick(n > 6, fetch=False), ticks)
    return {"config": 4, "chain": name, "robots": B, "ms_per_tick": float(ms.mean()), "p50_us": float(np.median(ms) * 1e3),
            "robot_ticks_per_s": B / (ms.mean() * 1e-3), "fallback_robots": int(b.tick_fallback_count()),
            "route": "fused tick; robots in the blend branch go through k_tick_tail"}


def config1(torch, chain, ticks=10000):
    """BASELINE config 1: ONE Panda, the control cycle of examples/04-task_and_redundancy (…cpp:138-199) with the host in the loop
    every tick: q, dq from host memory -> osc_tick -> tau back in host memory, wall-clock per tick (what a 1 kHz control thread
    would see).  The goal moves on the circle of :163-172.  Not a throughput configuration: a single robot uses one thread of one SM."""
    import time
    from common import q0_of
    b = osc_b200.Batch(chain, 1, q0=q0_of(chain))
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = b.add_joint_task(abi.joint_task_desc(chain["n"]))
    q = np.array([q0_of(chain)], float); dq = np.zeros((1, chain["n"]))
    b.set(abi.F_Q, q); b.set(abi.F_DQ, dq); b.update_model(); b.reinitialize_tasks()
    x0 = np.asarray(b.get(abi.F_CURRENT_POSITION, t0)).copy()
    lat = np.empty(ticks)
    for k in range(ticks + 200):
        t = 1e-3 * k
        goal = x0 + 0.1 * np.array([[np.sin(np.pi * t), 1 - np.cos(np.pi * t), 0.0]])
        b.set(abi.F_GOAL_POSITION, goal, t0)
        a = time.perf_counter()
        b.set(abi.F_Q, q); b.set(abi.F_DQ, dq)
        tau = b.tick(True)          # returns the torques in host memory (synchronises)
        e = time.perf_counter()
        if k >= 200:
            lat[k - 200] = (e - a) * 1e6
        q = q + 1e-6 * np.asarray(tau)  # keep the inputs changing without leaving the neighbourhood
    return {"config": 1, "robots": 1, "ticks": ticks, "p50_us": float(np.percentile(lat, 50)), "p99_us": float(np.percentile(lat, 99)),
            "what": "host in the loop: 2 x osc_set_field(q, dq) + osc_tick returning tau to host memory, wall clock through the ctypes binding",
            "fused": bool(b.tick_is_fused())}


def single_task(torch, name="puma560_synthetic", B=65536, ticks=50):
    """A 6-DoF arm under one full MotionForceTask at random (mostly non-singular) postures: the single-task fused tick."""
    from common import q0_of, random_q
    chain = chain_of(name)
    rng = np.random.default_rng(6)
    q = random_q(chain, B, rng, 0.3)
    n = chain["n"]
    out = []
    for outputs, tag in ((None, "fused tick (k_tick_fused2<6> + k_tick_tail)"), (abi.TICK_OUTPUTS_ALL, "per-call kernels (API-faithful route)")):
        b = osc_b200.Batch(chain, B, q0=q0_of(chain))
        b.set_stream(torch.cuda.current_stream().cuda_stream)
        b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0)))
        if outputs is not None:
            b.set_tick_outputs(outputs)
        b.set(abi.F_Q, q); b.set(abi.F_DQ, rng.uniform(-0.3, 0.3, (B, n))); b.update_model(); b.reinitialize_tasks()
        ms = timed(torch, lambda: b.tick(False, fetch=False), ticks)
        out.append({"config": "single 6-DoF task", "chain": name, "robots": B, "ms_per_tick": float(ms.mean()), "p50_us": float(np.median(ms) * 1e3),
                    "robot_ticks_per_s": B / (ms.mean() * 1e-3), "fallback_robots": int(b.tick_fallback_count()) if outputs is None else 0, "route": tag})
        del b
    return out


def single_task7(torch, chain, B=65536, ticks=50):
    """examples/07's hierarchy: a Panda whose RobotController holds ONE MotionForceTask, config 2's postures.  Plain motion task (the SIMPLE
    SINGLE instantiation) and force along z + two moment axes in closed loop with the passivity observer (the general one), each through the
    fused tick and through the per-call kernels."""
    w = workloads.config2(chain, B, seed=5)
    f = np.tile([0.0, 0.0, 5.0], (B, 1)); m = np.zeros((B, 3))
    out = []
    for law in ("motion", "surface contact (force 1 + moment 2, closed loop, POPC)"):
        for outputs, tag in ((None, "fused tick (k_tick_fused2<7, SINGLE> + k_tick_tail)"), (abi.TICK_OUTPUTS_ALL, "per-call kernels (API-faithful route)")):
            b = osc_b200.Batch(chain, B, q0=[j["q0"] for j in chain["joints"]])
            b.set_stream(torch.cuda.current_stream().cuda_stream)
            t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
            if outputs is not None:
                b.set_tick_outputs(outputs)
            b.set(abi.F_Q, w["q"]); b.set(abi.F_DQ, w["dq"]); b.update_model()
            if law != "motion":
                p = b.mft_params(t0)
                p.force_space_dimension, p.moment_space_dimension, p.closed_loop_force_control, p.closed_loop_moment_control, p.passivity_enabled = 1, 2, 1, 1, 1
                for i, x in enumerate((0.0, 0.0, 1.0)):
                    p.force_or_motion_axis[i] = x; p.moment_or_rotmotion_axis[i] = x
                b.set_mft_params(t0, p)
                b.set(abi.F_GOAL_FORCE, np.tile([0.0, 0.0, -5.0], (B, 1)), t0)
                b.update_sensed_force_and_moment(t0, f, m)
            b.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); b.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0)
            ms = timed(torch, lambda: b.tick(True, fetch=False), ticks)
            out.append({"config": "single 6-DoF task on the 7-DoF Panda", "law": law, "robots": B, "ms_per_tick": float(ms.mean()), "p50_us": float(np.median(ms) * 1e3),
                        "robot_ticks_per_s": B / (ms.mean() * 1e-3), "fallback_robots": int(b.tick_fallback_count()) if outputs is None else 0, "route": tag,
                        "fused": bool(b.tick_is_fused())})
            del b
    return out


def joint_only(torch, name="puma560_synthetic", B=65536, ticks=50):
    """examples/01: a full JointTask alone (k_tick_joint) against the per-call kernels"""
    from common import q0_of, random_q
    chain = chain_of(name)
    rng = np.random.default_rng(8)
    n = chain["n"]
    q = random_q(chain, B, rng, 0.3)
    out = []
    for outputs, tag in ((None, "fused tick (k_tick_joint)"), (abi.TICK_OUTPUTS_ALL, "per-call kernels (API-faithful route)")):
        b = osc_b200.Batch(chain, B, q0=q0_of(chain))
        b.set_stream(torch.cuda.current_stream().cuda_stream)
        t0 = b.add_joint_task(abi.joint_task_desc(n))
        if outputs is not None:
            b.set_tick_outputs(outputs)
        b.set(abi.F_Q, q); b.set(abi.F_DQ, rng.uniform(-0.3, 0.3, (B, n))); b.update_model()
        b.set(abi.F_JT_GOAL_POSITION, q + 0.1, t0)
        ms = timed(torch, lambda: b.tick(True, fetch=False), ticks)
        out.append({"config": "full JointTask alone", "chain": name, "robots": B, "ms_per_tick": float(ms.mean()), "p50_us": float(np.median(ms) * 1e3),
                    "robot_ticks_per_s": B / (ms.mean() * 1e-3), "route": tag, "fused": bool(b.tick_is_fused())})
        del b
    return out


def config4_stages(torch, name="panda", B=65536, ticks=10):
    """Where the per-call route spends its time when EVERY robot is in the singular blend branch (what k_tick_tail runs per robot)."""
    import test_singular_gpu as S
    from common import q0_of
    chain = chain_of(name)
    rng = np.random.default_rng(4)
    base = S._panda_q(rng) if name == "panda" else S._puma_q(rng)
    q = np.tile(base, ((B + len(base) - 1) // len(base), 1))[:B]
    n = chain["n"]
    b = osc_b200.Batch(chain, B, q0=q0_of(chain))
    b.set_stream(torch.cuda.current_stream().cuda_stream)
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0)))
    t1 = b.add_joint_task(abi.joint_task_desc(n)) if n > 6 else None
    b.set(abi.F_Q, q); b.set(abi.F_DQ, np.zeros((B, n))); b.update_model(); b.reinitialize_tasks()
    calls = [("k_model", lambda: b.update_model()), ("k_mft_update", lambda: b.task_update_model(t0, -1)),
             ("k_mft_compute", lambda: b.task_compute_torques(t0, False))]
    if t1 is not None:
        calls += [("k_jt_update", lambda: b.task_update_model(t1, t0)), ("k_jt_compute", lambda: b.task_compute_torques(t1, True))]
    out = {}
    for nm, fn in calls:
        out[nm] = float(np.median(timed(torch, fn, ticks, warm=2)) * 1e3)
    return {"config": 4, "chain": name, "robots": B, "route": "per-call kernels, every robot in the blend branch", "stage_us": out}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-batch", type=int, default=3000000)
    ap.add_argument("--only", default="", help="comma list of: 1, 3, 4, 4s, 5, st, st7, jt")
    args = ap.parse_args()
    import torch
    panda = chain_of("panda")
    only = set(args.only.split(",")) if args.only else None
    if only is not None:
        if "1" in only:
            print(json.dumps(config1(torch, panda)), flush=True)
        if "5" in only:
            for B in (10000, 30000, 100000, 300000, 1000000, 3000000, 10000000):
                if B <= args.max_batch:
                    print(json.dumps(config5(torch, panda, B)), flush=True)
        if "3" in only:
            for rec in config3(torch, panda):
                print(json.dumps(rec), flush=True)
        if "4" in only:
            print(json.dumps(config4(torch, "panda")), flush=True)
            print(json.dumps(config4(torch, "puma560_synthetic")), flush=True)
        if "st" in only:
            for rec in single_task(torch):
                print(json.dumps(rec), flush=True)
        if "st7" in only:
            for rec in single_task7(torch, panda):
                print(json.dumps(rec), flush=True)
        if "jt" in only:
            for nm in ("puma560_synthetic", "panda"):
                for rec in joint_only(torch, nm):
                    print(json.dumps(rec), flush=True)
        if "4s" in only:
            print(json.dumps(config4_stages(torch, "panda")), flush=True)
            print(json.dumps(config4_stages(torch, "puma560_synthetic")), flush=True)
        return
    print(json.dumps(config1(torch, panda)), flush=True)
    for B in (10000, 30000, 100000, 300000, 1000000, 3000000, 10000000):
        if B <= args.max_batch:
            print(json.dumps(config5(torch, panda, B)), flush=True)
            torch.cuda.empty_cache()
    for rec in config3(torch, panda):
        print(json.dumps(rec), flush=True)
    print(json.dumps(config4(torch, "panda")), flush=True)
    print(json.dumps(config4(torch, "puma560_synthetic")), flush=True)
    for rec in single_task(torch):
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
