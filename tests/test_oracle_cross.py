"""CPU: the C++ oracle (oracle/osc_oracle.hpp, the checker of the GPU parity tests and the timed CPU baseline) against
the independent numpy/LAPACK oracle (oracle/np_oracle.py) on the reference's example setups.

PARITY UNPINNED: the reference has no tests / golden vectors for this path and cannot be compiled here
(Eigen3 + sai2-model absent), so the two restatements -- written on different numerical routes (world-frame composite
rigid body + Cholesky + one-sided Jacobi   vs   per-body kinetic-energy Hessians + LAPACK inv/svd) -- pin each other.
"""
import copy

import numpy as np
import pytest

import np_oracle as npo
import osc_abi as abi
from common import load_chain, q0_of, random_q

TOL = 1e-9


def _rel(a, b, floor=1e-6):
    a, b = np.asarray(a, float).ravel(), np.asarray(b, float).ravel()
    return np.abs(a - b).max() / max(np.abs(b).max(), floor)


class NpRobot:
    """one numpy-oracle robot with the same hierarchy as an OracleBatch"""

    def __init__(self, chain):
        self.model = npo.Model(chain)
        self.tasks = []

    def controller(self):
        return npo.RobotController(self.model, self.tasks)


def _two_task_np(chain, q, dq, link="end-effector", p_c=(0, 0, 0.07)):
    r = NpRobot(chain)
    r.model.setQ(q); r.model.setDq(dq); r.model.updateModel()
    r.tasks.append(npo.MotionForceTask(r.model, link, np.array(p_c, float)))
    r.tasks.append(npo.JointTask(r.model))
    return r


def test_panda_two_task_stacked(oracle_built):
    """examples/05 setup (BASELINE config 2): 6-DoF MotionForceTask + full JointTask, RobotController semantics"""
    orc = oracle_built
    chain = load_chain("panda")
    B = 24
    rng = np.random.default_rng(5)
    q = random_q(chain, B, rng)
    dq = rng.uniform(-0.5, 0.5, (B, 7))
    o = orc.OracleBatch(chain, B, q0=q0_of(chain))
    o.set(abi.F_Q, q); o.set(abi.F_DQ, dq); o.update_model()
    t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = o.add_joint_task(abi.joint_task_desc(7))
    robots = [_two_task_np(chain, q[r], dq[r]) for r in range(B)]
    gp = rng.normal(0, 0.05, (B, 3)); gv = rng.normal(0, 0.05, (B, 3)); gw = rng.normal(0, 0.05, (B, 3))
    gq = q + rng.normal(0, 0.1, (B, 7))
    x0 = o.get(abi.F_CURRENT_POSITION, t0)
    o.set(abi.F_GOAL_POSITION, x0 + gp, t0); o.set(abi.F_GOAL_LINEAR_VELOCITY, gv, t0)
    o.set(abi.F_GOAL_ANGULAR_VELOCITY, gw, t0); o.set(abi.F_JT_GOAL_POSITION, gq, t1)
    for tick in range(3):
        tau = o.tick(True)
        for r, rb in enumerate(robots):
            m, j = rb.tasks
            if tick == 0:
                assert _rel(m.x, x0[r]) < 1e-12
                m.goal_pos = m.x + gp[r]; m.goal_linvel = gv[r]; m.goal_angvel = gw[r]; j.goal_pos = gq[r]
            c = rb.controller()
            rb.model.updateModel()
            c.updateControllerTaskModels()
            t_np = c.computeControlTorques()
            assert _rel(tau[r], t_np) < TOL, (tick, r)
            assert _rel(o.get(abi.F_M)[r], rb.model.M()) < TOL
            assert _rel(o.get(abi.F_N_TOTAL)[r], c.N_total, 1.0) < TOL
            assert _rel(o.get(abi.F_JACOBIAN, t0)[r], m.J) < TOL
            assert _rel(o.get(abi.F_LAMBDA_WORLD, t0)[r], m.handler.lambda_world()) < TOL
            assert _rel(o.get(abi.F_ALPHA, t0)[r], m.handler.alpha, 1.0) < TOL


def test_panda_force_control_popc(oracle_built):
    """examples/09 setup (BASELINE config 3): translation-only partial task, one force axis (Z), closed loop force
    control with the passivity observer, full JointTask below; 130 ticks so the 50-tick PC period fires twice."""
    orc = oracle_built
    chain = load_chain("panda")
    B = 4
    rng = np.random.default_rng(9)
    q = random_q(chain, B, rng, 0.3)
    dq = rng.uniform(-0.3, 0.3, (B, 7))
    dirs = [(1, 0, 0), (0, 1, 0), (0, 0, 1)]
    o = orc.OracleBatch(chain, B, q0=q0_of(chain))
    o.set(abi.F_Q, q); o.set(abi.F_DQ, dq); o.update_model()
    t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0), dirs_translation=dirs, dirs_rotation=[]))
    t1 = o.add_joint_task(abi.joint_task_desc(7))
    p = o.mft_params(t0)
    p.force_space_dimension = 1
    p.force_or_motion_axis[0], p.force_or_motion_axis[1], p.force_or_motion_axis[2] = 0.0, 0.0, 1.0
    p.closed_loop_force_control = 1
    p.passivity_enabled = 1
    o.set_mft_params(t0, p)
    o.set(abi.F_GOAL_FORCE, np.array([0, 0, -5.0]), t0)
    robots = []
    for r in range(B):
        rb = NpRobot(chain)
        rb.model.setQ(q[r]); rb.model.setDq(dq[r]); rb.model.updateModel()
        m = npo.MotionForceTask(rb.model, "end-effector", np.zeros(3), dirs_t=dirs, dirs_r=[])
        m.parametrizeForceMotionSpaces(1, np.array([0, 0, 1.0]))
        m.setClosedLoopForceControl(True)
        m.popc.enable()
        m.goal_force = np.array([0, 0, -5.0])
        rb.tasks = [m, npo.JointTask(rb.model)]
        robots.append(rb)
    worst = 0.0
    for tick in range(130):
        f = np.array([0, 0, -5.0]) * (1 + 0.3 * np.sin(2 * np.pi * 3 * tick * 1e-3 + np.arange(B)))[:, None] + rng.normal(0, 0.2, (B, 3))
        qq = q + 0.02 * np.sin(0.01 * tick)
        o.set(abi.F_Q, qq); o.set(abi.F_DQ, dq)
        o.update_model()
        o.update_task_models()
        o.update_sensed_force_and_moment(t0, f, np.zeros((B, 3)))
        tau = o.compute_torques(True)
        for r, rb in enumerate(robots):
            rb.model.setQ(qq[r]); rb.model.setDq(dq[r]); rb.model.updateModel()
            c = rb.controller()
            c.updateControllerTaskModels()
            rb.tasks[0].updateSensedForceAndMoment(f[r], np.zeros(3))
            worst = max(worst, _rel(tau[r], c.computeControlTorques()))
    assert worst < TOL, worst
    rc = o.get(abi.F_POPC_RC, t0)[:, 0]
    assert np.allclose(rc, [rb.tasks[0].popc.Rc for rb in robots], atol=1e-12)
    assert _rel(o.get(abi.F_INTEGRATED_FORCE_ERROR, t0), [rb.tasks[0].integ_force for rb in robots]) < TOL


def test_planar_partial_task(oracle_built):
    """examples/11: planar 4R, partial task X,Y translation + Z rotation (m = 3, n = 4)"""
    orc = oracle_built
    chain = load_chain("rrrrbot")
    B = 16
    rng = np.random.default_rng(11)
    q = rng.uniform(-1.5, 1.5, (B, 4)); dq = rng.uniform(-0.5, 0.5, (B, 4))
    o = orc.OracleBatch(chain, B, q0=q0_of(chain))
    o.set(abi.F_Q, q); o.set(abi.F_DQ, dq); o.update_model()
    dt_, dr_ = [(1, 0, 0), (0, 1, 0)], [(0, 0, 1)]
    t0 = o.add_motion_force_task(abi.mft_desc(chain, "link4", (0.5, 0, 0), dirs_translation=dt_, dirs_rotation=dr_))
    t1 = o.add_joint_task(abi.joint_task_desc(4))
    assert o.task_dof(t0) == 3
    off = rng.normal(0, 0.05, (B, 3))
    o.set(abi.F_GOAL_POSITION, o.get(abi.F_CURRENT_POSITION, t0) + off, t0)
    tau = o.tick(True)
    for r in range(B):
        rb = NpRobot(chain)
        rb.model.setQ(q[r]); rb.model.setDq(dq[r]); rb.model.updateModel()
        m = npo.MotionForceTask(rb.model, "link4", np.array([0.5, 0, 0]), dirs_t=dt_, dirs_r=dr_)
        m.goal_pos = m.x + off[r]
        rb.tasks = [m, npo.JointTask(rb.model)]
        c = rb.controller(); c.updateControllerTaskModels()
        assert _rel(tau[r], c.computeControlTorques()) < TOL, r


def test_puma_wrist_singularity(oracle_built):
    """examples/19: 6R PUMA-geometry chain around the wrist-lock posture, singularity handling gains (50, 20, 20),
    several ticks so the classification history and the q_prior latch are exercised."""
    orc = oracle_built
    chain = load_chain("puma560_synthetic")
    B = 12
    rng = np.random.default_rng(19)
    base = np.array([0.3, -np.pi / 4, 3 * np.pi / 4, np.pi / 2, 0.0, np.pi])
    eps = np.repeat([1e-1, 1e-2, 1e-3, 1e-4], 3)
    q = base[None] + rng.normal(0, 1, (B, 6)) * eps[:, None]
    dq = rng.uniform(-0.2, 0.2, (B, 6))
    o = orc.OracleBatch(chain, B, q0=q0_of(chain))
    o.set(abi.F_Q, q); o.set(abi.F_DQ, dq); o.update_model()
    t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0)))
    p = o.mft_params(t0)
    p.kp_type_1, p.kv_type_1, p.kv_type_2 = 50.0, 20.0, 20.0
    o.set_mft_params(t0, p)
    off = rng.normal(0, 0.03, (B, 3))
    o.set(abi.F_GOAL_POSITION, o.get(abi.F_CURRENT_POSITION, t0) + off, t0)
    robots = []
    for r in range(B):
        rb = NpRobot(chain)
        rb.model.setQ(q[r]); rb.model.setDq(dq[r]); rb.model.updateModel()
        m = npo.MotionForceTask(rb.model, "end-effector", np.zeros(3))
        m.handler.kp1, m.handler.kv1, m.handler.kv2 = 50.0, 20.0, 20.0
        m.goal_pos = m.x + off[r]
        rb.tasks = [m]
        robots.append(rb)
    seen_blend = False
    for tick in range(4):
        qq = q + 1e-3 * tick
        o.set(abi.F_Q, qq)
        tau = o.tick(False)
        alpha = o.get(abi.F_ALPHA, t0)[:, 0]
        seen_blend |= bool(((alpha > 0) & (alpha < 1)).any()) or bool((alpha == 0).any())
        for r, rb in enumerate(robots):
            rb.model.setQ(qq[r]); rb.model.updateModel()
            m = rb.tasks[0]
            m.updateTaskModel(np.eye(6))
            t_np = m.computeTorques()
            assert _rel(tau[r], t_np, 1e-3) < 1e-8, (tick, r, alpha[r])
            assert abs(alpha[r] - m.handler.alpha) < 1e-9
    assert seen_blend, "the batch should reach the blend / singular branch"


def test_sliding_base_partial_joint_first(oracle_built):
    """examples/06: n = 8 (prismatic base + Panda), partial JointTask on the prismatic joint first, MotionForceTask
    second, full JointTask third."""
    orc = oracle_built
    chain = load_chain("panda_sliding_base")
    n = chain["n"]
    B = 6
    rng = np.random.default_rng(6)
    q = random_q(chain, B, rng, 0.2); dq = rng.uniform(-0.3, 0.3, (B, n))
    S = np.zeros((1, n)); S[0, 0] = 1.0
    o = orc.OracleBatch(chain, B, q0=q0_of(chain))
    o.set(abi.F_Q, q); o.set(abi.F_DQ, dq); o.update_model()
    ta = o.add_joint_task(abi.joint_task_desc(n, S))
    tb = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    tc = o.add_joint_task(abi.joint_task_desc(n))
    o.set(abi.F_JT_GOAL_POSITION, q[:, :1] + 0.1, ta)
    tau = o.tick(True)
    for r in range(B):
        rb = NpRobot(chain)
        rb.model.setQ(q[r]); rb.model.setDq(dq[r]); rb.model.updateModel()
        a = npo.JointTask(rb.model, S, name="a"); a.goal_pos = q[r, :1] + 0.1
        rb.tasks = [a, npo.MotionForceTask(rb.model, "end-effector", np.array([0, 0, 0.07])), npo.JointTask(rb.model, name="c")]
        c = rb.controller(); c.updateControllerTaskModels()
        assert _rel(tau[r], c.computeControlTorques()) < TOL, r


def test_controller_rules():
    """RobotController.cpp:8-51: nothing may follow a full joint task; every task has the same loop timestep; unique
    names (numpy oracle; the product enforces the same rules in osc_batch_add_*: tests/test_api_gpu.py)"""
    chain = load_chain("panda")
    m = npo.Model(chain)
    with pytest.raises(ValueError):
        npo.RobotController(m, [npo.JointTask(m), npo.MotionForceTask(m, "end-effector")])
    with pytest.raises(ValueError):
        npo.RobotController(m, [npo.MotionForceTask(m, "end-effector", dt=0.001), npo.JointTask(m, dt=0.002)])
    with pytest.raises(ValueError):
        npo.RobotController(m, [npo.JointTask(m, np.eye(7)[:1], name="a"), npo.JointTask(m, name="a")])
    with pytest.raises(ValueError):
        npo.RobotController(m, [])


def test_tilted_base_full_inertias_gravity(oracle_built):
    """the two restatements on a chain the fixtures do not offer: base rotated and offset in the world, gravity off the world
    axes, inertia tensors with products of inertia, gravity compensation on (the GPU tests use the same chain:
    tests/test_fused_tick.py::test_tilted_base_full_inertias_gravity_compensation, tests/test_rollout_gpu.py)"""
    orc = oracle_built
    chain = copy.deepcopy(load_chain("panda"))
    cz, sz, cy, sy = np.cos(0.4), np.sin(0.4), np.cos(0.7), np.sin(0.7)
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]]); Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    chain["world_rot"] = (Rz @ Ry).tolist()
    chain["world_xyz"] = [0.1, -0.2, 0.3]
    chain["gravity"] = [1.0, -2.0, -9.0]
    rng = np.random.default_rng(78)
    for b in chain["bodies"]:
        Q = np.linalg.qr(np.eye(3) + rng.uniform(-0.3, 0.3, (3, 3)))[0]
        b["inertia"] = (Q @ np.array(b["inertia"], float) @ Q.T).tolist()
    B = 12
    q = random_q(chain, B, rng)
    dq = rng.uniform(-0.5, 0.5, (B, 7))
    o = orc.OracleBatch(chain, B, q0=q0_of(chain))
    o.set(abi.F_Q, q); o.set(abi.F_DQ, dq); o.update_model()
    t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = o.add_joint_task(abi.joint_task_desc(7))
    o.enable_gravity_compensation(True)
    x0 = o.get(abi.F_CURRENT_POSITION, t0)
    gp = rng.normal(0, 0.05, (B, 3)); gq = q + rng.normal(0, 0.1, (B, 7))
    o.set(abi.F_GOAL_POSITION, x0 + gp, t0); o.set(abi.F_JT_GOAL_POSITION, gq, t1)
    tau = o.tick(True)
    g = o.get(abi.F_GRAVITY)
    assert np.abs(g).max() > 1.0
    for r in range(B):
        rb = _two_task_np(chain, q[r], dq[r])
        m, j = rb.tasks
        assert _rel(m.x, x0[r]) < 1e-12
        m.goal_pos = m.x + gp[r]; j.goal_pos = gq[r]
        c = rb.controller()
        c.gravity_comp = True
        rb.model.updateModel()
        c.updateControllerTaskModels()
        assert _rel(tau[r], c.computeControlTorques()) < TOL, r
        assert _rel(g[r], rb.model.jointGravityVector()) < TOL
        assert _rel(o.get(abi.F_M)[r], rb.model.M()) < TOL
        assert _rel(o.get(abi.F_JACOBIAN, t0)[r], m.J) < TOL
        assert _rel(o.get(abi.F_LAMBDA_WORLD, t0)[r], m.handler.lambda_world()) < TOL
