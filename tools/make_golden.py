"""Generate tests/golden/*.npz: frozen input/output vectors of the control tick, computed by the numpy/LAPACK oracle
(oracle/np_oracle.py).  PARITY UNPINNED: the reference has no golden vectors of its own and cannot be built here
(DESIGN.md 1), so these freeze OUR restatement; tests/test_golden.py replays them on the C++ oracle (CPU) and on the CUDA
path (GPU).  Run in the build container:   python tools/make_golden.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("sai2-primitives-perso_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import np_oracle as npo  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def chain(name):
    return json.load(open(os.path.join(ROOT, "fixtures", name + ".chain.json")))


def rq(c, B, rng, margin=0.1):
    lo = np.array([j["lower"] for j in c["joints"]]) + margin
    hi = np.array([j["upper"] for j in c["joints"]]) - margin
    return rng.uniform(lo, hi, (B, c["n"]))


def two_task(name="panda", B=16, seed=101, ticks=2, link="end-effector", p_c=(0, 0, 0.07)):
    """examples/05: 6-DoF MotionForceTask + full JointTask, RobotController (stacked)"""
    c = chain(name); n = c["n"]; rng = np.random.default_rng(seed)
    q, dq = rq(c, B, rng), rng.uniform(-0.5, 0.5, (B, n))
    dpos, gv, gw = rng.normal(0, 0.05, (B, 3)), rng.normal(0, 0.05, (B, 3)), rng.normal(0, 0.05, (B, 3))
    gq = q + rng.normal(0, 0.1, (B, n))
    dqt = 0.01 * rng.normal(0, 1, (ticks, B, n))
    out = dict(tau=[], M=[], J=[], lam=[], Ntot=[], alpha=[], x=[])
    for r in range(B):
        m = npo.Model(c); m.setQ(q[r]); m.setDq(dq[r]); m.updateModel()
        t0 = npo.MotionForceTask(m, link, np.array(p_c, float)); t1 = npo.JointTask(m)
        t0.goal_pos = t0.x + dpos[r]; t0.goal_linvel = gv[r]; t0.goal_angvel = gw[r]; t1.goal_pos = gq[r]
        ctl = npo.RobotController(m, [t0, t1])
        for t in range(ticks):
            m.setQ(q[r] + dqt[t, r]); m.updateModel(); ctl.updateControllerTaskModels(); tau = ctl.computeControlTorques()
        out["tau"].append(tau); out["M"].append(m.M().ravel()); out["J"].append(t0.J.ravel()); out["lam"].append(t0.handler.lambda_world().ravel())
        out["Ntot"].append(ctl.N_total.ravel()); out["alpha"].append([t0.handler.alpha]); out["x"].append(t0.x)
    return dict(q=q, dq=dq, dpos=dpos, gv=gv, gw=gw, gq=gq, dqt=dqt, p_c=np.array(p_c, float), **{k: np.array(v) for k, v in out.items()})


def force_popc(B=4, seed=103, ticks=120):
    """examples/09: translation-only partial task, force axis Z, closed loop + passivity, JointTask below"""
    c = chain("panda"); rng = np.random.default_rng(seed)
    q, dq = rq(c, B, rng, 0.3), rng.uniform(-0.3, 0.3, (B, 7))
    dirs = [(1, 0, 0), (0, 1, 0), (0, 0, 1)]
    f = np.array([0, 0, -5.0])[None, None] * (1 + 0.3 * np.sin(2 * np.pi * 3e-3 * np.arange(ticks)[:, None, None] + np.arange(B)[None, :, None])) \
        + rng.normal(0, 0.2, (ticks, B, 3))
    taus, rc, integ = np.zeros((ticks, B, 7)), np.zeros(B), np.zeros((B, 3))
    for r in range(B):
        m = npo.Model(c); m.setQ(q[r]); m.setDq(dq[r]); m.updateModel()
        t0 = npo.MotionForceTask(m, "end-effector", np.zeros(3), dirs_t=dirs, dirs_r=[])
        t0.parametrizeForceMotionSpaces(1, np.array([0, 0, 1.0])); t0.setClosedLoopForceControl(True); t0.popc.enable()
        t0.goal_force = np.array([0, 0, -5.0])
        ctl = npo.RobotController(m, [t0, npo.JointTask(m)])
        for t in range(ticks):
            m.setQ(q[r] + 0.02 * np.sin(0.01 * t)); m.updateModel(); ctl.updateControllerTaskModels()
            t0.updateSensedForceAndMoment(f[t, r], np.zeros(3))
            taus[t, r] = ctl.computeControlTorques()
        rc[r] = t0.popc.Rc; integ[r] = t0.integ_force
    return dict(q=q, dq=dq, f=f, tau=taus[[0, ticks // 2, ticks - 1]], tick_ids=np.array([0, ticks // 2, ticks - 1]), rc=rc, integ=integ)


def puma_singular(B=8, seed=119, ticks=3):
    """examples/19: wrist-lock posture +- eps, handler gains (50, 20, 20), manual single task"""
    c = chain("puma560_synthetic"); rng = np.random.default_rng(seed)
    base = np.array([0.3, -np.pi / 4, 3 * np.pi / 4, np.pi / 2, 0.0, np.pi])
    q = base[None] + rng.normal(0, 1, (B, 6)) * np.repeat([1e-1, 1e-2, 1e-3, 1e-4], 2)[:, None]
    dq, off = rng.uniform(-0.2, 0.2, (B, 6)), rng.normal(0, 0.03, (B, 3))
    tau, alpha = np.zeros((B, 6)), np.zeros(B)
    for r in range(B):
        m = npo.Model(c); m.setQ(q[r]); m.setDq(dq[r]); m.updateModel()
        t0 = npo.MotionForceTask(m, "end-effector", np.zeros(3))
        t0.handler.kp1, t0.handler.kv1, t0.handler.kv2 = 50.0, 20.0, 20.0
        t0.goal_pos = t0.x + off[r]
        for t in range(ticks):
            m.setQ(q[r] + 1e-3 * t); m.updateModel(); t0.updateTaskModel(np.eye(6)); tau[r] = t0.computeTorques()
        alpha[r] = t0.handler.alpha
    return dict(q=q, dq=dq, off=off, ticks=np.array(ticks), tau=tau, alpha=alpha)


def planar_partial(B=8, seed=111):
    """examples/11: planar 4R, X, Y translation + Z rotation, JointTask below"""
    c = chain("rrrrbot"); rng = np.random.default_rng(seed)
    q, dq, off = rng.uniform(-1.5, 1.5, (B, 4)), rng.uniform(-0.5, 0.5, (B, 4)), rng.normal(0, 0.05, (B, 3))
    tau = np.zeros((B, 4)); M = np.zeros((B, 16))
    for r in range(B):
        m = npo.Model(c); m.setQ(q[r]); m.setDq(dq[r]); m.updateModel()
        t0 = npo.MotionForceTask(m, "link4", np.array([0.5, 0, 0]), dirs_t=[(1, 0, 0), (0, 1, 0)], dirs_r=[(0, 0, 1)])
        t0.goal_pos = t0.x + off[r]
        ctl = npo.RobotController(m, [t0, npo.JointTask(m)]); ctl.updateControllerTaskModels()
        tau[r] = ctl.computeControlTorques(); M[r] = m.M().ravel()
    return dict(q=q, dq=dq, off=off, tau=tau, M=M)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for name, fn in (("panda_two_task", two_task), ("panda_force_popc", force_popc), ("puma_singular", puma_singular), ("planar_partial", planar_partial)):
        d = fn()
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, {k: v.shape for k, v in d.items()})
