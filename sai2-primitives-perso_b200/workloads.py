"""Synthetic inputs of the BASELINE.json configs, generated on the host with numpy from a seed.

Used by bench.py and by the full-size GPU tests.  This is input generation only (vectorised forward kinematics to
place goals and to reject near-singular draws); it is not part of the controller and not the oracle.
"""
import numpy as np


def _rot_axis_batch(axis, q):
    """Rodrigues for a fixed unit axis and a batch of angles -> [B,3,3]"""
    a = np.asarray(axis, float)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    s, c = np.sin(q)[:, None, None], np.cos(q)[:, None, None]
    return np.eye(3)[None] + s * K[None] + (1 - c) * (K @ K)[None]


def batched_fk(chain, q, link, p_in_link=(0, 0, 0)):
    """world pose of a point of `link` and the 6xn world Jacobian (linear rows first) for q [B,n]"""
    B, n = q.shape
    R = np.broadcast_to(np.array(chain["world_rot"], float), (B, 3, 3)).copy()
    p = np.broadcast_to(np.array(chain["world_xyz"], float), (B, 3)).copy()
    zs, os_ = [], []
    L = chain["links"][link]
    body = L["body"]
    for i in range(body + 1):
        j = chain["joints"][i]
        p = p + np.einsum("bij,j->bi", R, np.array(j["xyz"], float))
        R = R @ np.array(j["rot"], float)[None]
        ax = np.array(j["axis"], float)
        if j["type"] == "prismatic":
            p = p + np.einsum("bij,j->bi", R, ax) * q[:, i:i + 1]
        else:
            R = R @ _rot_axis_batch(ax, q[:, i])
        zs.append(np.einsum("bij,j->bi", R, ax))
        os_.append(p.copy())
    pl = np.array(L["xyz"], float) + np.array(L["rot"], float) @ np.array(p_in_link, float)
    x = p + np.einsum("bij,j->bi", R, pl)
    Rl = R @ np.array(L["rot"], float)[None]
    J = np.zeros((B, 6, n))
    for i in range(body + 1):
        if chain["joints"][i]["type"] == "prismatic":
            J[:, :3, i] = zs[i]
        else:
            J[:, :3, i] = np.cross(zs[i], x - os_[i])
            J[:, 3:, i] = zs[i]
    return x, Rl, J


def config2(chain, B, seed=2, link="end-effector", p_in_link=(0, 0, 0.07), min_inv_cond=0.1):
    """BASELINE config 2: q ~ U(lower+0.1, upper-0.1) redrawn while sigma_min/sigma_max of J < 0.1 (keeps the batch in
    the non-singular branch), dq ~ U(-0.5, 0.5), goal pose = FK(q + N(0, 0.05^2)), goal twist ~ N(0, 0.05^2), goal
    acceleration 0, joint goal q + N(0, 0.1^2)."""
    rng = np.random.default_rng(seed)
    n = chain["n"]
    lo = np.array([j["lower"] for j in chain["joints"]]) + 0.1
    hi = np.array([j["upper"] for j in chain["joints"]]) - 0.1
    q = rng.uniform(lo, hi, (B, n))
    for _ in range(200):
        _, _, J = batched_fk(chain, q, link, p_in_link)
        s = np.linalg.svd(J, compute_uv=False)
        bad = s[:, min(5, n - 1)] / s[:, 0] < min_inv_cond
        if not bad.any():
            break
        q[bad] = rng.uniform(lo, hi, (int(bad.sum()), n))
    dq = rng.uniform(-0.5, 0.5, (B, n))
    xg, Rg, _ = batched_fk(chain, q + rng.normal(0, 0.05, (B, n)), link, p_in_link)
    return dict(q=q, dq=dq, goal_pos=xg, goal_ori=Rg.reshape(B, 9), goal_linvel=rng.normal(0, 0.05, (B, 3)),
                goal_angvel=rng.normal(0, 0.05, (B, 3)), jt_goal=q + rng.normal(0, 0.1, (B, n)))


def config3_wrench(B, tick, seed=3, goal_force=(0.0, 0.0, -5.0), dt=0.001):
    """BASELINE config 3 surrogate contact signal: F_s(t) = F_goal (1 + 0.3 sin(2 pi 3 t + phi_r)) + N(0, 0.2^2)"""
    rng = np.random.default_rng([seed, tick])
    phase = np.random.default_rng(seed).uniform(0, 2 * np.pi, B)
    g = np.asarray(goal_force, float)
    f = g[None] * (1 + 0.3 * np.sin(2 * np.pi * 3 * tick * dt + phase))[:, None] + rng.normal(0, 0.2, (B, 3))
    return f, np.zeros((B, 3))


def otg_rollout(batch, w, t0, t1, ticks, abi, every=10, per_call=True):
    """Roll-out with the internal OTGs on, shared by tools/make_otg_golden.py (CPU oracle on the
    real Ruckig) and tests/test_otg_gpu.py (CUDA path): the robots stay at the config-2 state (q, dq of `w`: no simulator
    here, and the OTG does not read the robot state), the goals change at ticks 0, 120 (new pose + joint goal), 200
    (goal velocity on half of the robots) and 300 (back to the first goals).  `batch` is an osc_b200.Batch or an
    orc.OracleBatch (same method names).  Records, every `every` ticks: torques, desired pose / joint positions, OTG flags."""
    B, n = w["q"].shape
    rng = np.random.default_rng(1234)
    goals2 = dict(pos=w["goal_pos"] + rng.uniform(-0.08, 0.08, (B, 3)), jt=w["jt_goal"] + rng.uniform(-0.4, 0.4, (B, n)))
    half = (np.arange(B) % 2 == 0)[:, None]
    rec = dict(tau=[], des_pos=[], des_ori=[], des_linvel=[], des_angvel=[], jt_des_pos=[], jt_des_vel=[], jt_des_acc=[], flags0=[], flags1=[])
    for t in range(ticks):
        if t == 0:
            batch.set(abi.F_GOAL_POSITION, w["goal_pos"], t0)
            batch.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0)
            batch.set(abi.F_JT_GOAL_POSITION, w["jt_goal"], t1)
        elif t == 120:
            batch.set(abi.F_GOAL_POSITION, goals2["pos"], t0)
            batch.set(abi.F_JT_GOAL_POSITION, goals2["jt"], t1)
        elif t == 200:
            batch.set(abi.F_GOAL_LINEAR_VELOCITY, np.where(half, w["goal_linvel"], 0.0), t0)
            batch.set(abi.F_GOAL_ANGULAR_VELOCITY, np.where(half, w["goal_angvel"], 0.0), t0)
            batch.set(abi.F_JT_GOAL_VELOCITY, np.where(half, 0.05, 0.0) * np.ones((B, n)), t1)
        elif t == 300:
            batch.set(abi.F_GOAL_LINEAR_VELOCITY, np.zeros((B, 3)), t0)
            batch.set(abi.F_GOAL_ANGULAR_VELOCITY, np.zeros((B, 3)), t0)
            batch.set(abi.F_JT_GOAL_VELOCITY, np.zeros((B, n)), t1)
            batch.set(abi.F_GOAL_POSITION, w["goal_pos"], t0)
            batch.set(abi.F_JT_GOAL_POSITION, w["jt_goal"], t1)
        if per_call:
            batch.update_model()
            batch.update_task_models()
            tau = batch.compute_torques(True)
        else:
            tau = batch.tick(True)
        if t % every == 0 or t == ticks - 1:
            rec["tau"].append(np.asarray(tau, dtype=np.float64))
            rec["des_pos"].append(np.asarray(batch.get(abi.F_DESIRED_POSITION, t0), dtype=np.float64))
            rec["des_ori"].append(np.asarray(batch.get(abi.F_DESIRED_ORIENTATION, t0), dtype=np.float64))
            rec["des_linvel"].append(np.asarray(batch.get(abi.F_DESIRED_LINEAR_VELOCITY, t0), dtype=np.float64))
            rec["des_angvel"].append(np.asarray(batch.get(abi.F_DESIRED_ANGULAR_VELOCITY, t0), dtype=np.float64))
            rec["jt_des_pos"].append(np.asarray(batch.get(abi.F_JT_DESIRED_POSITION, t1), dtype=np.float64))
            rec["jt_des_vel"].append(np.asarray(batch.get(abi.F_JT_DESIRED_VELOCITY, t1), dtype=np.float64))
            rec["jt_des_acc"].append(np.asarray(batch.get(abi.F_JT_DESIRED_ACCELERATION, t1), dtype=np.float64))
            rec["flags0"].append(np.asarray(batch.get(abi.F_OTG_FLAGS, t0)).astype(np.int32) & 1)
            rec["flags1"].append(np.asarray(batch.get(abi.F_OTG_FLAGS, t1)).astype(np.int32) & 1)
    return {k: np.stack(v) for k, v in rec.items()}
