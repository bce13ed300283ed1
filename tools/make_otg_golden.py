"""Generates tests/golden/otg_*.npz with the REAL Ruckig library of the reference underneath (oracle/_ref/libruckig_ref.so,
built by `make -C oracle ref` from /root/reference/ruckig).  Run in the build container (needs /root/reference):

    python tools/make_otg_golden.py

  otg_ruckig.npz     4000 single Ruckig::update calls (inputs in the regimes of tests/test_otg_oracle.py::random_case)
  otg_joints.npz     OTG_joints roll-outs: goal sequences (steps, moving goals, goal changes mid-trajectory, invalid goals)
  otg_cartesian.npz  OTG_6dof_cartesian roll-outs (position + orientation goals, re-referencing)
  otg_tasks.npz      task-level: config-2-style Panda batch with both internal OTGs on, desired states and torques of the
                     C++ oracle running on the real Ruckig (consumed by tests/test_otg_gpu.py)
The wrappers (OTG_joints.cpp / OTG_6dof_cartesian.cpp) need Eigen and stay restatements (oracle/otg_oracle.hpp); what these
files pin is everything below them.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200"))
import orc  # noqa: E402
from test_otg_oracle import random_case  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
DT = 0.001


def rot(axis, ang):
    axis = np.asarray(axis, float) / np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K


def joints_script(rng, n, steps):
    """goal sequence [steps][2n]: piecewise-constant goals with occasional changes, a moving goal stretch, one invalid goal"""
    p = rng.uniform(-1, 1, n)
    goals = np.zeros((steps, 2 * n))
    g, gv = p.copy(), np.zeros(n)
    t = 0
    while t < steps:
        seg = int(rng.integers(30, 400))
        kind = rng.integers(0, 5)
        if kind == 0:
            g = g + rng.uniform(-0.8, 0.8, n)
            gv = np.zeros(n)
        elif kind == 1:  # collinear step (phase sync)
            g = g + rng.uniform(0.1, 1.0) * rng.uniform(-1, 1, n)
            gv = np.zeros(n)
        elif kind == 2:  # goal with a velocity (finishes moving: Q14 branch)
            g = g + rng.uniform(-0.5, 0.5, n)
            gv = rng.uniform(-0.2, 0.2, n)
        elif kind == 3:  # tiny change (below nothing: exact isApprox 1e-12) and a real one
            g = g * (1 + 1e-14)
        else:  # target velocity above the limit on one joint: ErrorInvalidInput, output frozen
            g = g + rng.uniform(-0.3, 0.3, n)
            gv = np.zeros(n)
            gv[rng.integers(0, n)] = 50.0
            seg = int(rng.integers(3, 10))
        for s in range(seg):
            if t >= steps:
                break
            goals[t, :n] = g
            goals[t, n:] = gv
            t += 1
    return p, goals


def main():
    assert orc.have_reference_ruckig(), "build oracle/_ref first: make -C oracle ref"
    os.makedirs(GOLDEN, exist_ok=True)
    rng = np.random.default_rng(7)
    orc.otg_use_reference(True)

    # ---- single updates
    N = 4000
    io = np.zeros((N, 49))
    out = np.zeros((N, 21))
    info = np.zeros((N, 3))
    res = np.zeros(N, dtype=np.int32)
    ns = np.zeros(N, dtype=np.int32)
    for i in range(N):
        n = int(rng.integers(1, 8))
        case = random_case(rng, n)
        rk = orc.Ruckig2(n, DT)
        r, o, inf = rk.update(*case)
        io[i, :7 * n] = np.concatenate(case)
        ns[i], res[i], info[i] = n, r, inf
        if r >= 0:
            out[i, :3 * n] = o.ravel()
    np.savez_compressed(os.path.join(GOLDEN, "otg_ruckig.npz"), io=io, out=out, info=info, result=res, n=ns, dt=DT)
    print("otg_ruckig.npz: %d cases, results:" % N, dict(zip(*np.unique(res, return_counts=True))))

    # ---- OTG_joints roll-outs
    d = {"dt": DT}
    cases = 12
    for c in range(cases):
        n = int(rng.integers(1, 8))
        steps = 1500
        p0, goals = joints_script(rng, n, steps)
        lim = np.stack([rng.uniform(0.5, 2.0, n), rng.uniform(1.0, 8.0, n)])
        o = orc.OtgJoints(p0, DT, lim[0], lim[1])
        outs, results, reached = np.zeros((steps, 3 * n)), np.zeros(steps, np.int32), np.zeros(steps, np.int32)
        for t in range(steps):
            r, y, fl = o.step(goals[t, :n], goals[t, n:])
            outs[t], results[t], reached[t] = y.ravel(), r, fl
        d.update({"p0_%d" % c: p0, "goals_%d" % c: goals, "outs_%d" % c: outs, "results_%d" % c: results, "reached_%d" % c: reached,
                  "limits_%d" % c: lim})
        print("joints case %d: n=%d results" % (c, n), dict(zip(*np.unique(results, return_counts=True))))
    d["cases"] = cases
    np.savez_compressed(os.path.join(GOLDEN, "otg_joints.npz"), **d)

    # ---- OTG_6dof_cartesian roll-outs
    d = {"dt": DT}
    cases = 8
    for c in range(cases):
        steps = 1500
        p0 = rng.uniform(-0.5, 0.5, 3)
        R0 = rot(rng.normal(size=3), rng.uniform(0, 2.5))
        lim = np.array([0.3, 2.0, np.pi / 3, 2 * np.pi]) * rng.uniform(0.5, 2.0)  # MotionForceTask.h:68-71 scaled
        goals = np.zeros((steps, 18))
        gp, gR, gv, gw = p0.copy(), R0.copy(), np.zeros(3), np.zeros(3)
        t = 0
        while t < steps:
            seg = int(rng.integers(40, 400))
            kind = rng.integers(0, 5)
            if kind == 0:
                gp = gp + rng.uniform(-0.2, 0.2, 3)
            elif kind == 1:
                gR = gR @ rot(rng.normal(size=3), rng.uniform(0.05, 1.5))
            elif kind == 2:
                gp = gp + rng.uniform(-0.2, 0.2, 3)
                gR = gR @ rot(rng.normal(size=3), rng.uniform(0.05, 1.0))
            elif kind == 3:
                gp = gp + rng.uniform(-0.1, 0.1, 3)
                gv = rng.uniform(-0.05, 0.05, 3)
                gw = rng.uniform(-0.1, 0.1, 3) if rng.random() < 0.5 else np.zeros(3)
            else:
                gp = gp * (1 + 1e-5)  # below the wrapper's 1e-3 isApprox threshold: ignored
            for s in range(seg):
                if t >= steps:
                    break
                goals[t] = np.concatenate([gp, gR.ravel(), gv, gw])
                t += 1
            if kind == 3:
                gv, gw = np.zeros(3), np.zeros(3)
        lin_v, lin_a, ang_v, ang_a = lim
        o = orc.OtgCartesian(p0, R0, DT, lin_v, lin_a, ang_v, ang_a)
        outs, results, reached = np.zeros((steps, 24)), np.zeros(steps, np.int32), np.zeros(steps, np.int32)
        for t in range(steps):
            r, y, fl = o.step(goals[t, :3], goals[t, 3:12], goals[t, 12:15], goals[t, 15:18])
            outs[t], results[t], reached[t] = y, r, fl
        d.update({"p0_%d" % c: p0, "R0_%d" % c: R0, "goals_%d" % c: goals, "outs_%d" % c: outs, "results_%d" % c: results,
                  "reached_%d" % c: reached, "limits_%d" % c: lim})
        print("cartesian case %d: results" % c, dict(zip(*np.unique(results, return_counts=True))))
    d["cases"] = cases
    np.savez_compressed(os.path.join(GOLDEN, "otg_cartesian.npz"), **d)

    # ---- task level (Panda, config-2 style, both OTGs on, reference defaults MotionForceTask.h:67-72 / JointTask.h:38-41)
    import osc_abi as abi
    import workloads
    chain = json.load(open(os.path.join(ROOT, "fixtures", "panda.chain.json")))
    B, T = 48, 1000
    w = workloads.config2(chain, B, seed=31)
    q0 = [j["q0"] for j in chain["joints"]]
    o = orc.OracleBatch(chain, B, q0=q0)
    t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = o.add_joint_task(abi.joint_task_desc(chain["n"]))
    o.set(abi.F_Q, w["q"]); o.set(abi.F_DQ, w["dq"]); o.update_model()
    o.reinitialize_tasks()
    o.enable_internal_otg(t0, [0.3, np.pi / 3], [2.0, 2 * np.pi])
    o.enable_internal_otg(t1, np.pi / 3, 2 * np.pi)
    rec = workloads.otg_rollout(o, w, t0, t1, T, abi)
    np.savez_compressed(os.path.join(GOLDEN, "otg_tasks.npz"), B=B, T=T, seed=31, **rec)
    print("otg_tasks.npz: tau range", np.abs(rec["tau"]).max())
    orc.otg_use_reference(False)


if __name__ == "__main__":
    main()
