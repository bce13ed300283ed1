import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import osc_abi as abi
import orc
orc.build()
import osc_b200, workloads
from common import load_chain, q0_of, rel_err
chain = load_chain("panda")
B = 256
w = workloads.config2(chain, B, 46)
def setup(b):
    b.set(abi.F_Q, w["q"]); b.set(abi.F_DQ, w["dq"]); b.update_model()
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = b.add_joint_task(abi.joint_task_desc(7))
    p = b.mft_params(t0)
    for a in range(3):
        p.kp_pos[a], p.kv_pos[a], p.kp_ori[a], p.kv_ori[a] = 120.0, 22.0, 210.0, 29.0
    b.set_mft_params(t0, p)
    jp = b.joint_task_params(t1)
    for a in range(7):
        jp.kp[a], jp.kv[a] = 55.0, 15.0
    b.set_joint_task_params(t1, jp)
    b.set(abi.F_GOAL_LINEAR_VELOCITY, np.array([0.01, -0.02, 0.03]), t0)
    p = b.mft_params(t0)
    p.force_space_dimension = 1
    p.force_or_motion_axis[0], p.force_or_motion_axis[1], p.force_or_motion_axis[2] = 0.0, 0.0, 1.0
    p.closed_loop_force_control = 1
    b.set_mft_params(t0, p)
    b.set(abi.F_GOAL_FORCE, np.array([0.0, 0.0, 5.0]), t0)
    b.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); b.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0); b.set(abi.F_JT_GOAL_POSITION, w["jt_goal"], t1)
    return t0, t1
for variant in ("update_model", "defer"):
    g = osc_b200.Batch(chain, B, q0=q0_of(chain)); o = orc.OracleBatch(chain, B, q0=q0_of(chain))
    t0, t1 = setup(g); setup(o)
    q = np.array(w["q"], float)
    for t in range(4):
        q = q + 1e-3 * (np.arange(7) - 3.0)
        f, m = np.array([0.3, -0.2, 4.0 + 0.1 * t]), np.array([0.01, 0.02, -0.03])
        o.set(abi.F_Q, q); o.update_model(); o.update_sensed_force_and_moment(t0, f, m); ref = o.tick(True)
        g.set(abi.F_Q, q)
        if variant == "update_model": g.update_model()
        elif variant == "defer": g._chk(g.L.osc_defer_model_update(g.h))
        g.update_sensed_force_and_moment(t0, f, m)
        tau = g.tick(True)
        print(variant, t, "err %.3e" % rel_err(tau, ref)[0], "fused", g.tick_is_fused())
