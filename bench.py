#!/usr/bin/env python
"""bench.py -- OSC control updates/s (robots x ticks / s) on N B200s, BASELINE.json config 2:
Panda 7-DoF, 6-DoF MotionForceTask + full JointTask in its nullspace through RobotController semantics, 65,536 robots
per GPU, FP64.  One "step" = one control tick of every robot of the batch (updateModel + updateControllerTaskModels +
computeControlTorques).  Robots are independent: the batch shards over GPUs with no collective ("scaling": "weak").

    python bench.py --gpus 1 --steps 50 --warmup 5            # this repo's CUDA path (through the C ABI)
    python bench.py --impl reference --gpus 1 --steps 3 --warmup 1   # the reference algorithm on the host cores (oracle port)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200"))

METRIC = "OSC control updates/sec (robots x ticks/s)"
UNIT = "robot-ticks/s"
WORKLOAD = "Panda 7-DoF 6-DoF OSC + joint nullspace task, batch 65,536 robots per B200, FP64 (BASELINE config 2)"
# algorithmic work per robot-tick, frozen from SURVEY.md 8(d) (the reference's own route: two Jacobi SVDs included)
ALG_FLOP_PER_TICK = 3.6e4
ALG_BYTES_PER_TICK = 416.0
# internal OTG state advance, both tasks: per DoF the current segment (p, v, a, end time) read and the
# desired (p, v, a) written, plus the trajectory clock; the cartesian OTG also writes the 9 entries of the desired orientation
# joints (n = 7): (4*7 + 1 + 3*7) * 8 = 400 B; cartesian: (4*6 + 1 + 24) * 8 = 392 B
OTG_BYTES_PER_TICK = 792.0
# rollout step, n = 7: FK 504 + CRBA 1,680 + recursive Newton-Euler ~ 150 n = 1,050
# + Cholesky factor n^3/3 and two triangular solves 2 n^2 (FMA = 2) = 327 + integrator 4 n = 28
ROLLOUT_FLOP_PER_STEP = 3589.0


def load_chain(name="panda"):
    with open(os.path.join(ROOT, "fixtures", name + ".chain.json")) as f:
        return json.load(f)


class NearGpu:
    """Pinned host buffers of the end-to-end leg are allocated (and first touched) by a thread bound to the CPUs of the
    GPU's own NUMA node when the box shows more than one: a buffer on the other socket crosses the inter-socket link on
    every DMA.  (The boxes of this pool are VMs with a single visible node, so this is a no-op there; see the candidate
    selection at the end-to-end leg.)  The affinity is restored on exit, so the CPU baseline and the clock sampler keep
    all the cores."""

    def __init__(self, device_index):
        self.cpus, self.saved = None, None
        try:
            import torch
            p = torch.cuda.get_device_properties(device_index)
            bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
            with open("/sys/bus/pci/devices/%s/local_cpulist" % bdf) as f:
                cpus = set()
                for part in f.read().strip().split(","):
                    lo, _, hi = part.partition("-")
                    cpus.update(range(int(lo), int(hi or lo) + 1))
            allowed = os.sched_getaffinity(0)
            if cpus & allowed and (cpus & allowed) != allowed:
                self.cpus = cpus & allowed
        except Exception:
            self.cpus = None

    def __enter__(self):
        if self.cpus:
            try:
                self.saved = os.sched_getaffinity(0)
                os.sched_setaffinity(0, self.cpus)
            except OSError:  # not permitted in this container: carry on unbound
                self.saved, self.cpus = None, None
        return self

    def __exit__(self, *a):
        if self.saved:
            try:
                os.sched_setaffinity(0, self.saved)
            except OSError:
                pass
        return False


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed regions (B200_PROFILING.md clocks line).  A timed region of
    this bench is a few milliseconds, far below nvidia-smi's sampling period, so NVML is polled directly from a thread
    (about 10 kHz); nvidia-smi is the fallback when the NVML binding is missing."""
    BITS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, device):
        self.device, self.sm, self.reasons, self.mx, self.power = device, [], set(), None, []
        self.stop_flag, self.t, self.h, self.nv = threading.Event(), None, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            idx = self.device
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                try:
                    idx = int(vis.split(",")[self.device])
                except ValueError:
                    pass
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.nv = pynvml
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self._poll if self.nv else self._smi, daemon=True)
        self.t.start()

    def _poll(self):
        nv = self.nv
        while not self.stop_flag.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for name, bit in self.BITS:
                    if r & bit:
                        self.reasons.add(name)
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception:
                break
            time.sleep(0.0005)  # ~2 kHz: leaves the interpreter lock to the launching thread

    def _smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=10).stdout
                f = [x.strip() for x in out.strip().split(",")]
                self.sm.append(float(f[0])); self.mx = float(f[1])
                for (name, _), v in zip(self.BITS, f[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(name)
            except Exception:
                break

    def stop(self):
        self.stop_flag.set()
        if self.t:
            self.t.join(timeout=15)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx, "reasons": sorted(self.reasons),
                "samples": len(self.sm), "power_w_max": max(self.power) if self.power else None,
                "how": "NVML polled from a thread during the device-resident and end-to-end timed regions" if self.nv else "nvidia-smi loop"}


def build_batch(osc_b200, abi, chain, B, device, seed, precision):
    import workloads
    w = workloads.config2(chain, B, seed)
    b = osc_b200.Batch(chain, B, precision=precision, device=device, q0=[j["q0"] for j in chain["joints"]])
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = b.add_joint_task(abi.joint_task_desc(chain["n"]))
    b.set(abi.F_Q, w["q"]); b.set(abi.F_DQ, w["dq"])
    b.update_model()
    b.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); b.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0)
    b.set(abi.F_GOAL_LINEAR_VELOCITY, w["goal_linvel"], t0); b.set(abi.F_GOAL_ANGULAR_VELOCITY, w["goal_angvel"], t0)
    b.set(abi.F_JT_GOAL_POSITION, w["jt_goal"], t1)
    return b, w, t0, t1


class CpuRef:
    """The reference algorithm (oracle port of the reference path, robots statically partitioned over the host threads, one
    robot instance per thread at a time, each with its own model copy) on a bounded sample of config 2.
    variant: "" = g++ -O2 (the reference's Release default, CMakeLists.txt:13-26), "O3" = -O3 -march=native,
    "heap" = malloc'ed matrix temporaries like Eigen's dynamic MatrixXd (oracle/Makefile)."""

    def __init__(self, chain, nthreads, Bc=4096, variant=""):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import orc
        import osc_abi as abi
        import workloads
        self.Bc, self.nthreads, self.variant = Bc, nthreads, variant
        w = workloads.config2(chain, Bc, seed=2)
        o = orc.OracleBatch(chain, Bc, q0=[j["q0"] for j in chain["joints"]], nthreads=nthreads, variant=variant)
        t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
        t1 = o.add_joint_task(abi.joint_task_desc(chain["n"]))
        o.set(abi.F_Q, w["q"]); o.set(abi.F_DQ, w["dq"]); o.update_model()
        o.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); o.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0)
        o.set(abi.F_GOAL_LINEAR_VELOCITY, w["goal_linvel"], t0); o.set(abi.F_GOAL_ANGULAR_VELOCITY, w["goal_angvel"], t0)
        o.set(abi.F_JT_GOAL_POSITION, w["jt_goal"], t1)
        self.o = o

    def run(self, ticks):
        L, h = self.o.L, self.o.h
        t = time.perf_counter()
        for _ in range(ticks):
            L.orc_tick(h, 1, self.nthreads)
        return time.perf_counter() - t

    def sample(self, ticks):
        flags = {"": "g++ -O2", "O3": "g++ -O3 -march=native", "heap": "g++ -O2, heap-allocated matrix temporaries"}[self.variant]
        return "%d robots x %d ticks of config 2 (seed 2), %d threads, %s" % (self.Bc, ticks, self.nthreads, flags)


def cpu_baseline(chain, seconds=10.0):
    """bounded CPU legs, rank 0 only: config 2 on all host threads (-O2 = the reported baseline; -O3 -march=native and the
    heap-temporaries proxy beside it), BASELINE config 1 (ONE robot, 10,000 cycles of the 04-task_and_redundancy loop, per-cycle
    latency), and the operation count of one config-2 tick that freezes the algorithmic FLOP figure of SURVEY.md 8(d)."""
    nthreads = os.cpu_count() or 1
    out = None
    for variant, secs in (("", seconds), ("O3", 3.0), ("heap", 3.0)):
        ref = CpuRef(chain, nthreads, variant=variant)
        ref.run(1)
        one = ref.run(1)
        ticks = int(max(2, min(5000, secs / max(one, 1e-6))))
        el = ref.run(ticks)
        rec = {"value": ref.Bc * ticks / el, "unit": UNIT, "cores": nthreads, "kind": "port", "sample": ref.sample(ticks)}
        if out is None:
            out = rec
            out["variants"] = {}
        else:
            out["variants"]["O3_march_native" if variant == "O3" else "heap_temporaries"] = {"value": rec["value"], "sample": rec["sample"]}
    import orc
    import osc_abi as abi
    q0 = [j["q0"] for j in chain["joints"]]
    one = orc.OracleBatch(chain, 1, q0=q0)
    one.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07))); one.add_joint_task(abi.joint_task_desc(chain["n"]))
    us = one.config1_run(10000)
    out["config1_single_robot"] = {"p50_us_per_tick": float(np.median(us)), "p99_us_per_tick": float(np.percentile(us, 99)),
                                   "ticks": 10000, "what": "BASELINE config 1: one Panda, the 04-task_and_redundancy control cycle "
                                   "(setQ/setDq, updateModel, 2 x updateTaskModel, goal generator, 2 x computeTorques), one host thread, g++ -O2"}
    cnt = CpuRef(chain, 1, Bc=64, variant="count")
    flops = [cnt.o.count_tick_flops(True) for _ in range(2)][-1]
    out["counted_flop_per_robot_tick"] = {"value": flops, "what": "adds + multiplies of one config-2 tick of the oracle (reference route, two Jacobi "
                                          "SVDs), counted by its instrumented primitives (oracle/Makefile target count); SURVEY.md 8(d) estimated 3.55e4"}
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU algorithm on this box's host cores.  The reference library itself cannot be
    built here (needs Eigen3 + un-vendored sai2-model), so this is the oracle port ("kind": "port").
    One step = 10 ticks of a 4096-robot sample."""
    if rank != 0:
        return
    chain = load_chain()
    nthreads = os.cpu_count() or 1
    ref = CpuRef(chain, nthreads)
    TICKS = 10
    for _ in range(max(1, args.warmup)):
        ref.run(1)
    el = [ref.run(TICKS) for _ in range(args.steps)]
    value = float(ref.Bc * TICKS * len(el) / sum(el))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": float(np.mean(el) * 1e3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "step_sample": "each step = %d ticks of a %d-robot sample on the host cores" % (TICKS, ref.Bc)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": nthreads, "kind": "port", "sample": ref.sample(TICKS * len(el))},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=65536, help="robots per GPU")
    ap.add_argument("--precision", type=int, default=64, choices=[64, 32])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-otg", action="store_true", help="skip the extra measurement with the internal OTGs enabled")
    ap.add_argument("--tick-outputs", default="lazy", choices=["lazy", "tau", "all"],
                    help="lazy (default): the tick also writes every vector-valued getter of the reference; tau: torques only")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    run_b200(args, rank, world, local_rank)


def run_b200(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import osc_abi as abi
    import osc_b200
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the batched controller has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    chain = load_chain()
    n, B = chain["n"], args.batch
    real = torch.float64 if args.precision == 64 else torch.float32
    batch, w, t0, t1 = build_batch(osc_b200, abi, chain, B, local_rank, seed=2 + rank, precision=args.precision)
    batch.set_stream(torch.cuda.current_stream().cuda_stream)
    if args.tick_outputs == "tau":
        batch.set_tick_outputs(abi.TICK_OUTPUTS_TAU)
    elif args.tick_outputs == "all":  # the per-call (API-faithful) kernel sequence: every getter written every tick
        batch.set_tick_outputs(abi.TICK_OUTPUTS_ALL)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident tick: W warm-up + K timed steps, CUDA events on the launching stream, L2 flushed between steps
    for _ in range(args.warmup):
        batch.tick(True, fetch=False)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = batch.launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    noflush = bool(os.environ.get("OSC_BENCH_NOFLUSH"))  # diagnosis only: a line measured this way is not a bench value
    for s, e in ev:
        if not noflush:
            flush.zero_()
        s.record()
        batch.tick(True, fetch=False)
        e.record()
    barrier()
    launches = batch.launch_count() - l0
    step_ms = np.array([s.elapsed_time(e) for s, e in ev])
    total_ms = max_over_ranks(float(step_ms.sum()))
    value = world * B * args.steps / (total_ms * 1e-3)

    # ---- parity of the benchmarked state: 512 robots of this very batch against the oracle (outside every timed region)
    parity = None
    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import orc
        pick = np.sort(np.random.default_rng(7).choice(B, min(512, B), replace=False))
        tau_gpu = np.asarray(batch.tick(True), dtype=np.float64)[pick]
        o = orc.OracleBatch(chain, len(pick), q0=[j["q0"] for j in chain["joints"]], nthreads=os.cpu_count() or 1)
        o0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07))); o1 = o.add_joint_task(abi.joint_task_desc(n))
        o.set(abi.F_Q, w["q"][pick]); o.set(abi.F_DQ, w["dq"][pick]); o.update_model()
        o.set(abi.F_GOAL_POSITION, w["goal_pos"][pick], o0); o.set(abi.F_GOAL_ORIENTATION, w["goal_ori"][pick], o0)
        o.set(abi.F_GOAL_LINEAR_VELOCITY, w["goal_linvel"][pick], o0); o.set(abi.F_GOAL_ANGULAR_VELOCITY, w["goal_angvel"][pick], o0)
        o.set(abi.F_JT_GOAL_POSITION, w["jt_goal"][pick], o1)
        tau_ref = o.tick(True)
        err = np.abs(tau_gpu - tau_ref).max(axis=1) / np.maximum(np.abs(tau_ref).max(axis=1), 1e-3 * np.abs(tau_ref).max())
        tol = 1e-9 if args.precision == 64 else 1e-4
        parity = {"robots": int(len(pick)), "max_rel_err": float(err.max()), "tolerance": tol, "ok": bool(err.max() <= tol),
                  "what": "torques of %d random robots of the benchmarked batch after the timed ticks vs the CPU oracle" % len(pick)}
        del o

    # ---- the same tick captured in a CUDA graph (osc_tick_noalloc: no allocation under capture), replayed: launch-gap-free latency
    graph_tick = None
    if batch.tick_is_fused() and B <= abi.NOALLOC_MAX_ROBOTS:
        cap = torch.cuda.Stream()
        g = torch.cuda.CUDAGraph()
        batch.set_stream(cap.cuda_stream)
        with torch.cuda.stream(cap):
            batch.tick_noalloc(None, True)  # warm (function attributes, module load) outside the capture
            cap.synchronize()
            g.capture_begin()
            batch.tick_noalloc(None, True)
            g.capture_end()
            cap.synchronize()
            for _ in range(args.warmup):
                g.replay()
            evg = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
            for s_, e_ in evg:
                if not noflush:
                    flush.zero_()
                s_.record(cap); g.replay(); e_.record(cap)
            cap.synchronize()
        gms = np.array([s_.elapsed_time(e_) for s_, e_ in evg])
        graph_tick = {"p50_us": float(np.median(gms) * 1e3), "mean_us": float(gms.mean() * 1e3),
                      "what": "osc_tick_noalloc captured once in a CUDA graph and replayed (k_tick_fused2 + k_tick_tail, L2 flushed between replays)"}
        batch.set_stream(torch.cuda.current_stream().cuda_stream)
        del g
    barrier()

    # ---- end to end through the C ABI with HOST buffers: H2D of this step's state + goals, tick, D2H of the torques
    # The DMA rate of a pinned buffer depends on where its pages landed (measured on this pool, same process, same minute:
    # 47 GB/s for one cudaHostAlloc, 11 GB/s for the next; the boxes are VMs that show one NUMA node, so there is nothing to
    # bind to).  Like an application that keeps its I/O buffers for its lifetime, each host buffer is picked as the fastest
    # of a few pinned candidates, by a timed H2D copy; this happens once, outside every timed region.
    keep, rates = [], []

    def pinned(a, candidates=4):
        src = torch.from_numpy(np.ascontiguousarray(a.T)).to(real)  # SoA [k][B]
        dev = torch.empty_like(src, device="cuda")
        best, best_ms, first = None, None, None
        for _ in range(candidates):
            t = src.pin_memory()
            keep.append(t)  # held until all buffers are chosen, so that a rejected block is not handed out again at once
            dev.copy_(t, non_blocking=True)
            ms = []
            for _ in range(3):
                s_, e_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s_.record(); dev.copy_(t, non_blocking=True); e_.record(); torch.cuda.synchronize()
                ms.append(s_.elapsed_time(e_))
            if first is None:
                first = t
            if best is None or min(ms) < best_ms:
                best, best_ms = t, min(ms)
            rates.append(src.numel() * src.element_size() / (min(ms) * 1e-3) / 1e9)
        return first, best
    # every per-tick input of SURVEY.md 8(d): q, dq (14) + the six MotionForceTask goals (24) + the joint goal (7) = 45 scalars
    zeros3 = np.zeros((B, 3))
    fields = [(abi.F_Q, -1, w["q"]), (abi.F_DQ, -1, w["dq"]), (abi.F_GOAL_POSITION, t0, w["goal_pos"]), (abi.F_GOAL_ORIENTATION, t0, w["goal_ori"]),
              (abi.F_GOAL_LINEAR_VELOCITY, t0, w["goal_linvel"]), (abi.F_GOAL_ANGULAR_VELOCITY, t0, w["goal_angvel"]),
              (abi.F_GOAL_LINEAR_ACCELERATION, t0, zeros3), (abi.F_GOAL_ANGULAR_ACCELERATION, t0, zeros3), (abi.F_JT_GOAL_POSITION, t1, w["jt_goal"])]
    near = NearGpu(local_rank)
    with near:
        bufs = [(f, task, pinned(a)) for f, task, a in fields]
        tau_bufs = pinned(np.zeros((B, n)))
    h2d = sum(t[0].numel() * t[0].element_size() for _, _, t in bufs)
    d2h = tau_bufs[0].numel() * tau_bufs[0].element_size()
    L, h = batch.L, batch.h

    def e2e_run(which):
        h_in = [(f, task, t[which]) for f, task, t in bufs]
        h_tau = tau_bufs[which]

        def e2e_step():
            for f, task, t in h_in:
                L.osc_set_field(h, task, f, t.data_ptr(), abi.LAYOUT_SOA, abi.MEM_HOST)
            rc = L.osc_tick(h, 1, h_tau.data_ptr(), abi.LAYOUT_SOA, abi.MEM_HOST)  # returns after the torques landed
            assert rc == 0
        for _ in range(args.warmup):
            e2e_step()
        barrier()
        t_e2e = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        torch.cuda.synchronize()
        el = max_over_ranks(time.perf_counter() - t_e2e)
        barrier()
        return el
    e2e_first_s = e2e_run(0)   # the buffers as cudaHostAlloc first handed them out: the typical case
    e2e_s = e2e_run(1)         # each buffer the fastest of 4 candidates: what an application that keeps its I/O buffers can reach
    keep.clear()
    host_buffers = "pinned; value = each buffer the fastest of 4 candidates by a timed H2D copy (candidates ran %.0f-%.0f GB/s); first_allocation = the first candidate of each" % (min(rates), max(rates))
    if near.cpus:
        host_buffers += "; allocated on the GPU's NUMA node (%d local CPUs)" % len(near.cpus)
    clocks = sampler.stop() if rank == 0 else None
    e2e_value = world * B * args.steps / e2e_s
    e2e_first_value = world * B * args.steps / e2e_first_s

    # ---- roofline of the dominant kernel (its own launches timed with CUDA events, L2 flushed)
    fused = batch.tick_is_fused()
    stages = {}
    FUSED = "k_tick_fused2"
    if fused:
        stages[FUSED] = float(np.mean(step_ms))  # the step is this kernel + the (empty) k_tick_tail launch
    else:
        calls = [("k_model", lambda: batch.update_model()), ("k_mft_update", lambda: batch.task_update_model(0, -1)),
                 ("k_jt_update", lambda: batch.task_update_model(1, 0)), ("k_mft_compute", lambda: batch.task_compute_torques(0, False)),
                 ("k_jt_compute", lambda: batch.task_compute_torques(1, True))]
        for name, fn in calls:
            ts = []
            for _ in range(min(args.steps, 10)):
                flush.zero_()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record(); fn(); e.record(); torch.cuda.synchronize()
                ts.append(s.elapsed_time(e))
            stages[name] = float(np.mean(ts))
    # ---- the same tick with the reference's default internal OTGs on: k_otg_tick x2 ahead of the fused
    # kernel.  First tick after a goal change = every robot recalculates its trajectory; then the steady state-advance.
    otg = None
    if not args.no_otg and fused:
        batch.enable_internal_otg(t0, [0.3, np.pi / 3], [2.0, 2 * np.pi])   # MotionForceTask.h:68-71
        batch.enable_internal_otg(t1, np.pi / 3, 2 * np.pi)                  # JointTask.h:40-41
        rng = np.random.default_rng(99)
        batch.set(abi.F_GOAL_POSITION, w["goal_pos"] + rng.uniform(-0.2, 0.2, (B, 3)), t0)
        batch.set(abi.F_JT_GOAL_POSITION, w["jt_goal"] + rng.uniform(-1.0, 1.0, (B, n)), t1)
        batch.tick(True, fetch=False)  # untimed: the first launch of the OTG kernels also loads their code (lazy module loading)
        for _ in range(args.warmup):
            batch.tick(True, fetch=False)
        # new goals for every robot in mid-trajectory: the tick that follows recalculates every trajectory (MPC-style goal updates)
        batch.set(abi.F_GOAL_POSITION, w["goal_pos"] + rng.uniform(-0.2, 0.2, (B, 3)), t0)
        batch.set(abi.F_JT_GOAL_POSITION, w["jt_goal"] + rng.uniform(-1.0, 1.0, (B, n)), t1)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        flush.zero_(); s.record(); batch.tick(True, fetch=False); e.record(); torch.cuda.synchronize()
        recalc_ms = s.elapsed_time(e)
        for _ in range(args.warmup):
            batch.tick(True, fetch=False)
        ts = []
        for _ in range(args.steps):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); batch.tick(True, fetch=False); e.record(); torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        reached = (np.asarray(batch.get(abi.F_OTG_FLAGS, t1))[:, 0] & 1).mean()
        otg_ms = float(np.mean(ts))
        extra_ms = otg_ms - float(np.mean(step_ms))
        otg = {"ms_per_step": otg_ms, "value": B / (otg_ms * 1e-3), "unit": UNIT, "tick_all_robots_recalculate_ms": recalc_ms,
               "k_otg_tick_x2_ms": extra_ms, "alg_bytes_per_robot_tick": OTG_BYTES_PER_TICK,
               "hbm_frac_of_otg_kernels": OTG_BYTES_PER_TICK * B / max(extra_ms * 1e-3, 1e-9) / 1e9 / (
                   json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0),
               "robots_at_goal_after": float(reached), "limits": "reference defaults (MotionForceTask.h:68-71, JointTask.h:40-41)"}
        batch.disable_internal_otg(t0); batch.disable_internal_otg(t1)
    # ---- device-side rollout: [tick ; forward dynamics + symplectic Euler] with q, dq never leaving the GPU
    rollout = None
    if not args.no_otg and fused:
        snap = batch.snapshot()
        batch.enable_gravity_compensation(True)
        for _ in range(args.warmup):
            batch.tick(True, fetch=False); batch.rollout_step(1e-3, 1, 0.0)
        ts, ts_step = [], []
        for _ in range(min(args.steps, 20)):
            flush.zero_()
            s, m, e = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            s.record(); batch.tick(True, fetch=False); m.record(); batch.rollout_step(1e-3, 1, 0.0); e.record(); torch.cuda.synchronize()
            ts.append(s.elapsed_time(e)); ts_step.append(m.elapsed_time(e))
        finite = bool(np.isfinite(np.asarray(batch.get(abi.F_Q))).all())
        rollout = {"ms_per_tick_and_step": float(np.mean(ts)), "value": B / (float(np.mean(ts)) * 1e-3), "unit": "robot-(tick+sim step)s/s",
                   "k_rollout_step_ms": float(np.mean(ts_step)), "sim_dt": 1e-3, "substeps": 1, "state_finite_after": finite,
                   "alg_flop_per_robot_step": ROLLOUT_FLOP_PER_STEP,
                   "note": "k_rollout_step2 = sin/cos + link-frame recursive Newton-Euler + body-frame CRBA + Cholesky solve + integrator per robot; q, dq, tau stay in HBM"}
        batch.enable_gravity_compensation(False)
        batch.restore(snap)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    fp64_peak = osc_b200.measure_fp64_peak(local_rank)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    top = max(stages, key=stages.get)
    # algorithmic FLOP of each kernel per robot (whole tick = 3.6e4)
    alg = {"k_tick_fused": ALG_FLOP_PER_TICK, "k_tick_fused2": ALG_FLOP_PER_TICK, "k_model": 504 + 84 + 1680 + 686, "k_mft_update": 10800 + 504 + 1314 + 1092 + 2000,
           "k_jt_update": 15100 + 240, "k_mft_compute": 300 + 162 + 84 + 504, "k_jt_compute": 364}
    traffic, executed, capture = None, None, None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):  # per-launch figures of the committed `ncu --set full` capture of the same kernel (profiles/README.md)
        rec = json.load(open(tpath)).get(top) or {}
        traffic, executed, capture = rec.get("dram_bytes_per_launch"), rec.get("executed_fp64_flop_per_robot"), rec.get("capture")
        if traffic is not None and rec.get("robots") and rec["robots"] != B:
            traffic = traffic * B / rec["robots"]
    t_top = stages[top] * 1e-3
    work = executed if executed else alg[top]  # FLOP the kernel really executes per robot (ncu SASS count); algorithmic if no capture
    roof = {"bound": "fp64", "kernel": top, "achieved": work * B / t_top / 1e12, "peak": fp64_peak, "unit": "TFLOP/s",
            "frac": work * B / t_top / 1e12 / fp64_peak, "traffic": traffic,
            "peak_source": "DFMA microbenchmark in this run (osc_measure_fp64_peak); nominal 148 SM x 64 DFMA/clk x 2 x 1.965 GHz = 37.2; "
                           "MEASURED_PEAKS.json has no FP64 entry",
            "flop_per_robot": work,
            "flop_source": ("FP64 add/mul/fma thread instructions of the committed ncu capture %s (profiles/traffic.json)" % capture) if executed
                           else "algorithmic count of SURVEY.md 8(d) (no ncu capture of this kernel committed)",
            "alg_flop_per_robot_reference_route": alg[top],
            "alg_frac_reference_route": alg[top] * B / t_top / 1e12 / fp64_peak,
            "note": "frac = EXECUTED FP64 work / measured DFMA peak over the bench step.  alg_frac_reference_route uses the frozen algorithmic "
                    "count of SURVEY.md 8(d) (the reference's route with two Jacobi SVDs, 3.6e4 FLOP; the oracle's counter measures "
                    "3.4e4) and exceeds 1 because the fused kernel does not execute that work (closed forms + a certificate instead of the SVDs)",
            "stage_ms": stages}
    roof_hbm = {"bound": "hbm", "achieved": ALG_BYTES_PER_TICK * B / (float(np.mean(step_ms)) * 1e-3) / 1e9, "peak": hbm_peak,
                "unit": "GB/s", "peak_source": hbm_src}
    roof_hbm["frac"] = roof_hbm["achieved"] / hbm_peak
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64" if args.precision == 64 else "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "robots_per_gpu": B, "hierarchy": "MotionForceTask(6-DoF, end-effector + 0.07 m) > JointTask(full)",
                       "stacked": True, "tick_outputs": args.tick_outputs, "l2": "NOT FLUSHED (diagnostic run)" if noflush else "256 MB buffer written between timed steps (L2 flush)", "fused_tick": fused},
            "p50_tick_us": float(np.median(step_ms) * 1e3), "p99_tick_us": float(np.percentile(step_ms, 99) * 1e3),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "path": "osc_set_field x9 (q, dq, six MotionForceTask goals, joint goal: all 45 input scalars of SURVEY.md 8d) + osc_tick "
                            "with pinned HOST SoA buffers (C ABI), strictly serial on one stream",
                    "ms_per_step": e2e_s / args.steps * 1e3,
                    "first_allocation": {"value": e2e_first_value, "ms_per_step": e2e_first_s / args.steps * 1e3},
                    "pcie_gb_per_s": (h2d + d2h) / (e2e_s / args.steps) / 1e9,
                    "host_buffers": host_buffers},
            "parity_sample": parity, "graph_tick": graph_tick,
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "roofline_hbm": roof_hbm}
    if otg:
        line["otg_on"] = otg
    if rollout:
        rollout["fp64_frac"] = ROLLOUT_FLOP_PER_STEP * B / (rollout["k_rollout_step_ms"] * 1e-3) / 1e12 / fp64_peak
        line["rollout"] = rollout
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(chain)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
