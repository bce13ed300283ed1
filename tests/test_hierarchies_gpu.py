"""GPU parity of the hierarchies and inputs round 1 left untested on the device:

* examples/06-partial_joint_task/06-partial_joint_task.cpp:105-160 -- n = 8 sliding-base Panda (prismatic joint), PARTIAL
  JointTask first (slider + last joint, internal OTG on: the reference default), 6-DoF MotionForceTask second whose goal
  follows the joint task's desired state, RobotController stacking;
* hierarchies with three tasks (RobotController.cpp:18-60): partial JointTask > MotionForceTask > full JointTask on the
  sliding base, and partial MotionForceTask > partial JointTask > full JointTask on the Panda;
* BASELINE config 1: ONE Panda, the 04-task_and_redundancy setup (04...cpp:108-201), 10,000 ticks of its goal generator
  (circle r = 0.05 m at 0.33 Hz in y-z, orientation +-pi/8 about Y at 0.2 Hz, joint goal +1.5 rad on joint 1 at tick 5000),
  stack C (plain sum, manual chaining as the example does) and stack B (RobotController), tau against the oracle EVERY tick;
* a non-identity force-sensor frame (MotionForceTask::setForceSensorFrame / updateSensedForceAndMoment,
  MotionForceTask.cpp:794-828).
All FP64 at 1e-9 against the C++ oracle."""
import numpy as np
import pytest

import osc_abi as abi
from common import FP64_TOL, Pair, load_chain, q0_of, random_q

pytestmark = pytest.mark.gpu


def _tau_err(tg, tr):
    tg, tr = np.asarray(tg, float), np.asarray(tr, float)
    return float((np.abs(tg - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-3 * max(np.abs(tr).max(), 1e-300))).max())


def test_example06_partial_joint_task_first(oracle_built):
    chain = load_chain("panda_sliding_base")
    n, B, T = chain["n"], 48, 1400
    assert n == 8 and chain["joints"][0]["type"] == "prismatic"
    rng = np.random.default_rng(606)
    q0 = q0_of(chain)
    q = np.tile(q0, (B, 1)) + rng.normal(0, 0.1, (B, n))
    amp = rng.normal(0, 0.05, (B, n))
    S = np.zeros((2, n)); S[0, 0] = 1; S[1, 7] = 1      # 06...cpp:107-109
    p = Pair(chain, B, q0=q0)
    p.set(abi.F_Q, q); p.set(abi.F_DQ, np.zeros((B, n))); p.both("update_model")
    ta = p.add_jt(S)
    tb = p.add_mft("end-effector", (0, 0, 0.07))
    p.both("enable_internal_otg", ta, np.pi / 3, 2 * np.pi)   # JointTask default (JointTask.h:38-41); the MotionForceTask OTG stays off (:119)
    x0 = p.ref.get(abi.F_CURRENT_POSITION, tb)
    jgoal = p.ref.get(abi.F_JT_CURRENT_POSITION, ta).copy()
    w = 2 * np.pi * 0.3
    worst = 0.0
    for tick in range(T):
        t = tick * 0.001
        qt = q + amp * np.sin(2 * np.pi * 0.4 * t)
        p.set(abi.F_Q, qt); p.set(abi.F_DQ, amp * 2 * np.pi * 0.4 * np.cos(2 * np.pi * 0.4 * t))
        p.both("update_model"); p.both("update_task_models")
        if tick % 4000 == 1000:
            jgoal[:, 0] -= 1.0
        p.set(abi.F_JT_GOAL_POSITION, jgoal, ta)
        # the example reads the joint task's desired state back and feeds it to the motion task (06...cpp:155-163): done
        # separately on each side so that a divergence of the getters would show up in the torques
        for b in (p.gpu, p.ref):
            gp = x0 + 0.1 * np.array([np.sin(w * t), 0.0, 1 - np.cos(w * t)])
            gv = np.tile(0.1 * np.array([w * np.cos(w * t), 0.0, w * np.sin(w * t)]), (B, 1))
            ga = np.tile(0.1 * np.array([-w * w * np.sin(w * t), 0.0, w * w * np.cos(w * t)]), (B, 1))
            gp[:, 1] = b.get(abi.F_JT_DESIRED_POSITION, ta)[:, 0]
            gv[:, 1] = b.get(abi.F_JT_DESIRED_VELOCITY, ta)[:, 0]
            ga[:, 1] = b.get(abi.F_JT_DESIRED_ACCELERATION, ta)[:, 0]
            b.set(abi.F_GOAL_POSITION, gp, tb); b.set(abi.F_GOAL_LINEAR_VELOCITY, gv, tb); b.set(abi.F_GOAL_LINEAR_ACCELERATION, ga, tb)
        tg, tr = p.both("compute_torques", True)
        worst = max(worst, _tau_err(tg, tr))
        if tick in (0, 999, 1001, T - 1):
            p.compare(FP64_TOL)
    assert worst <= FP64_TOL, worst
    moved = p.ref.get(abi.F_JT_DESIRED_POSITION, ta)[:, 0] - jgoal[:, 0]
    assert np.abs(moved).max() > 0.1, "the OTG of the partial joint task should still be under way after the goal step"
    assert not p.gpu.tick_is_fused()


def test_three_task_hierarchy_sliding_base(oracle_built):
    chain = load_chain("panda_sliding_base")
    n, B = chain["n"], 96
    rng = np.random.default_rng(63)
    q = random_q(chain, B, rng, 0.2); dq = rng.uniform(-0.3, 0.3, (B, n))
    S = np.zeros((1, n)); S[0, 0] = 1.0
    p = Pair(chain, B, q0=q0_of(chain))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.both("update_model")
    ta = p.add_jt(S); tb = p.add_mft("end-effector", (0, 0, 0.07)); tc = p.add_jt()
    p.set(abi.F_JT_GOAL_POSITION, q[:, :1] + 0.1, ta)
    p.set(abi.F_GOAL_POSITION, p.ref.get(abi.F_CURRENT_POSITION, tb) + rng.normal(0, 0.05, (B, 3)), tb)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, (B, n)), tc)
    for tick in range(3):
        p.set(abi.F_Q, q + 0.01 * tick)
        for stacked in (True, False):
            tg, tr = p.both("tick", stacked)
            assert _tau_err(tg, tr) <= FP64_TOL
        p.compare(FP64_TOL)


def test_three_task_hierarchy_panda_partial_first(oracle_built):
    """partial MotionForceTask (translation only, examples/09:114-121) > partial JointTask (joints 1-3) > full JointTask"""
    chain = load_chain("panda")
    n, B = 7, 128
    rng = np.random.default_rng(64)
    q = random_q(chain, B, rng, 0.3); dq = rng.uniform(-0.5, 0.5, (B, n))
    p = Pair(chain, B, q0=q0_of(chain))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.both("update_model")
    ta = p.add_mft("end-effector", (0, 0, 0.07), dirs_translation=[(1, 0, 0), (0, 1, 0), (0, 0, 1)], dirs_rotation=[])
    tb = p.add_jt(np.eye(n)[:3]); tc = p.add_jt()
    p.set(abi.F_GOAL_POSITION, p.ref.get(abi.F_CURRENT_POSITION, ta) + rng.normal(0, 0.05, (B, 3)), ta)
    p.set(abi.F_JT_GOAL_POSITION, q[:, :3] + rng.normal(0, 0.1, (B, 3)), tb)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, (B, n)), tc)
    for tick in range(3):
        p.set(abi.F_Q, q + 0.01 * tick)
        tg, tr = p.both("tick", True)
        assert _tau_err(tg, tr) <= FP64_TOL
        p.compare(FP64_TOL)
    assert (p.ref.get(abi.F_JT_RANGE_RANK, tc) == 1).all() and (p.ref.get(abi.F_JT_RANGE_RANK, tb) == 3).all()


def _config1_goals(t, x0, R0):
    """04-task_and_redundancy.cpp:155-185"""
    w_o, a_o = 2 * np.pi * 0.2, np.pi / 8
    ang, angv, anga = a_o * np.sin(w_o * t), a_o * w_o * np.cos(w_o * t), -a_o * w_o * w_o * np.sin(w_o * t)
    c, s = np.cos(ang), np.sin(ang)
    Ry = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
    r, w = 0.05, 2 * np.pi * 0.33
    return dict(ori=(Ry.T @ R0).reshape(1, 9), angvel=np.array([[0, angv, 0]]), angacc=np.array([[0, anga, 0]]),
                pos=x0 + r * np.array([[0.0, np.sin(w * t), 1 - np.cos(w * t)]]),
                linvel=r * w * np.array([[0.0, np.cos(w * t), np.sin(w * t)]]),
                linacc=r * w * w * np.array([[0.0, -np.sin(w * t), np.cos(w * t)]]))


@pytest.mark.parametrize("stack", ["C_manual_sum", "B_robot_controller"])
def test_config1_single_panda_10000_ticks(oracle_built, stack):
    """BASELINE config 1 replayed on the GPU: B = 1, kinematic playback q(t) = q0 + 0.3 sin(w_j t) (no simulator
    here), joint-task OTG on as in the example, tau compared with the oracle at every one of the 10,000 ticks."""
    chain = load_chain("panda")
    n, T = 7, 10000
    q0 = q0_of(chain)
    p = Pair(chain, 1, q0=q0)
    p.set(abi.F_Q, q0[None]); p.set(abi.F_DQ, np.zeros((1, n))); p.both("update_model")
    t0 = p.add_mft("end-effector", (0, 0, 0.07))
    t1 = p.add_jt()
    p.both("enable_internal_otg", t1, np.pi / 3, 2 * np.pi)     # JointTask default; motion_force_task->disableInternalOtg() (:115)
    x0 = p.ref.get(abi.F_CURRENT_POSITION, t0).copy()
    R0 = p.ref.get(abi.F_CURRENT_ORIENTATION, t0).reshape(3, 3).copy()
    wj = 2 * np.pi * np.array([0.11, 0.13, 0.17, 0.19, 0.23, 0.29, 0.31])
    worst, worst_tick = 0.0, -1
    for tick in range(T):
        t = tick * 0.001
        qt = q0 + 0.3 * np.sin(wj * t) * np.array([1, 0.5, 1, 0.3, 1, 0.5, 1])   # stays inside the joint limits
        dqt = 0.3 * wj * np.cos(wj * t) * np.array([1, 0.5, 1, 0.3, 1, 0.5, 1])
        p.set(abi.F_Q, qt[None]); p.set(abi.F_DQ, dqt[None])
        g = _config1_goals(t, x0, R0)
        if stack == "C_manual_sum":
            p.both("update_model")
            p.both("task_update_model", t0, -1); p.both("task_update_model", t1, t0)
        for f, k in ((abi.F_GOAL_ORIENTATION, "ori"), (abi.F_GOAL_ANGULAR_VELOCITY, "angvel"), (abi.F_GOAL_ANGULAR_ACCELERATION, "angacc"),
                     (abi.F_GOAL_POSITION, "pos"), (abi.F_GOAL_LINEAR_VELOCITY, "linvel"), (abi.F_GOAL_LINEAR_ACCELERATION, "linacc")):
            p.set(f, g[k], t0)
        if tick == 5000:     # 04...cpp:196-201
            p.both("reinitialize_tasks", t1)
            gq = q0.copy(); gq[0] += 1.5
            p.set(abi.F_JT_GOAL_POSITION, gq[None], t1)
        if stack == "C_manual_sum":
            p.both("task_compute_torques", t0, False); p.both("task_compute_torques", t1, False)
            taus = [b.get(abi.F_TASK_TORQUES, t0) + (b.get(abi.F_TASK_TORQUES, t1) if tick >= 5000 else 0.0) for b in (p.gpu, p.ref)]
        else:
            taus = p.both("tick", True)
        e = _tau_err(taus[0], taus[1])
        if e > worst:
            worst, worst_tick = e, tick
        if tick % 2500 == 2499:
            # N_TOTAL is RobotController state (RobotController.cpp:58): with manual chaining the oracle never sets it
            p.compare(FP64_TOL, skip=(abi.F_N_TOTAL,) if stack == "C_manual_sum" else ())
    assert worst <= FP64_TOL, (worst, worst_tick)
    if stack == "B_robot_controller":
        assert p.gpu.tick_is_fused()


def test_force_sensor_frame_not_identity(oracle_built):
    """updateSensedForceAndMoment with _T_control_to_sensor != identity (MotionForceTask.cpp:794-828): rotated and offset
    sensor, the moment picks up r x F; closed-loop force + moment control so the sensed wrench reaches the torques"""
    chain = load_chain("panda")
    n, B = 7, 256
    rng = np.random.default_rng(815)
    q = random_q(chain, B, rng, 0.3); dq = rng.uniform(-0.5, 0.5, (B, n))
    a = 0.7
    Rz = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1.0]])
    Rx = np.array([[1, 0, 0], [0, np.cos(0.3), -np.sin(0.3)], [0, np.sin(0.3), np.cos(0.3)]])
    p = Pair(chain, B, q0=q0_of(chain))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.both("update_model")
    t0 = p.add_mft("end-effector", (0, 0, 0.07)); t1 = p.add_jt()
    p.set_mft_params(t0, sensor_rot=(Rz @ Rx).reshape(9), sensor_xyz=[0.02, -0.03, 0.05], force_space_dimension=1,
                     force_or_motion_axis=[0, 0, 1.0], moment_space_dimension=2, moment_or_rotmotion_axis=[0, 0, 1.0],
                     closed_loop_force_control=1, closed_loop_moment_control=1)
    p.set(abi.F_GOAL_FORCE, np.array([0, 0, -5.0]), t0); p.set(abi.F_GOAL_MOMENT, np.array([0.1, -0.2, 0.0]), t0)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, (B, n)), t1)
    for tick in range(4):
        p.set(abi.F_Q, q + 0.01 * tick)
        f, m = rng.normal(0, 3, (B, 3)), rng.normal(0, 0.5, (B, 3))
        for route in ("per_call", "tick"):
            if route == "per_call":
                p.both("update_model"); p.both("update_task_models")
                p.both("update_sensed_force_and_moment", t0, f, m)
                tg, tr = p.both("compute_torques", True)
            else:
                p.both("update_sensed_force_and_moment", t0, f, m)
                tg, tr = p.both("tick", True)
            assert _tau_err(tg, tr) <= FP64_TOL, (tick, route)
        p.compare(FP64_TOL)
    fw, mw = p.ref.get(abi.F_SENSED_FORCE_WORLD, t0), p.ref.get(abi.F_SENSED_MOMENT_WORLD, t0)
    assert np.abs(np.linalg.norm(fw, axis=1) - np.linalg.norm(f, axis=1)).max() < 1e-9   # a rotation of the force ...
    assert np.abs(np.linalg.norm(mw, axis=1) - np.linalg.norm(m, axis=1)).max() > 1e-3   # ... and r x F in the moment
