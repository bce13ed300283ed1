import os, subprocess, sys, tempfile, pathlib
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import osc_abi as abi
import orc
orc.build()
import osc_b200
import test_facade_cpp as T
from common import load_chain, q0_of, rel_err
chain = load_chain("panda")
B = 256
tmp = pathlib.Path(tempfile.mkdtemp())
exe = T._build()
for ticks in (1,):
    w = T._inputs(tmp, chain, B, 46)
    r = subprocess.run([exe, str(tmp / "chain.bin"), str(tmp / "in.bin"), str(tmp / "out.bin"), str(B), str(ticks), "7"], capture_output=True, text=True, env=dict(os.environ, FACADE_DEMO_DEBUG="1"))
    print("demo rc", r.returncode, r.stderr[:900])
    out = np.fromfile(tmp / "out.bin", dtype=np.float64)
    tau = out[:B * 7].reshape(B, 7)
    res = {}
    for name, b in (("oracle", orc.OracleBatch(chain, B, q0=q0_of(chain))), ("pygpu", osc_b200.Batch(chain, B, q0=q0_of(chain)))):
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
        b.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); b.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0)
        b.set(abi.F_GOAL_LINEAR_VELOCITY, np.array([0.01, -0.02, 0.03]), t0); b.set(abi.F_JT_GOAL_POSITION, w["jt_goal"], t1)
        p = b.mft_params(t0)
        p.force_space_dimension = 1
        p.force_or_motion_axis[0], p.force_or_motion_axis[1], p.force_or_motion_axis[2] = 0.0, 0.0, 1.0
        p.closed_loop_force_control = 1
        b.set_mft_params(t0, p)
        b.set(abi.F_GOAL_FORCE, np.array([0.0, 0.0, 5.0]), t0)
        q = np.array(w["q"], float)
        for t in range(ticks):
            q = q + 1e-3 * (np.arange(7) - 3.0)
            b.set(abi.F_Q, q); b.update_model()
            b.update_sensed_force_and_moment(t0, np.array([0.3, -0.2, 4.0 + 0.1 * t]), np.array([0.01, 0.02, -0.03]))
            res[name] = np.asarray(b.tick(True), float)
    print(ticks, "demo vs oracle %.3e  pygpu vs oracle %.3e  demo vs pygpu %.3e" % (rel_err(tau, res["oracle"])[0], rel_err(res["pygpu"], res["oracle"])[0], rel_err(tau, res["pygpu"])[0]))
    print("  robot 0 demo", tau[0][:4], "oracle", res["oracle"][0][:4])
    pp = b.mft_params(t0)
    print("  pygpu force_dim", pp.force_space_dimension, list(pp.force_or_motion_axis), pp.closed_loop_force_control, pp.kp_force, pp.kv_force, pp.ki_force,
          "sensed", b.get(abi.F_SENSED_FORCE_WORLD, t0)[0], "goal", b.get(abi.F_GOAL_FORCE, t0)[0], "integ", b.get(abi.F_INTEGRATED_FORCE_ERROR, t0)[0])
    print("  pygpu umf", b.get(abi.F_UNIT_MASS_FORCE, t0)[0], "cur", b.get(abi.F_CURRENT_POSITION, t0)[0], "goal", b.get(abi.F_GOAL_POSITION, t0)[0], "q", b.get(abi.F_Q)[0][:3])
