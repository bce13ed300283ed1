"""[full JointTask] alone on the fused tick (examples/01-joint_control on the PUMA, 02-joint_control_internal_otg): with N_prec = I the
task's range is the whole joint space, Lambda = M and Lambda_mod = M_BIE | M | I (JointTask.cpp:218-283), tau = M a_des + Lambda_mod u
(:342-355) -- k_tick_joint2 (all-revolute chains: body-frame composite-rigid-body recursion, shared-memory stash) and k_tick_joint
(any chain) form the mass matrix and two products, no factorisation.  Against the C++ oracle (which, like the reference,
goes through M^-1, the SVD range basis and the projected inertias), FP64 1e-9 / FP32 1e-4."""
import numpy as np
import pytest

import osc_abi as abi
from common import FP32_TOL, FP64_TOL, Pair, load_chain, q0_of, random_q, rel_err

pytestmark = pytest.mark.gpu


def _pair(name, B, seed, precision=abi.PRECISION_F64, **params):
    chain = load_chain(name)
    n = chain["n"]
    rng = np.random.default_rng(seed)
    q = random_q(chain, B, rng, 0.3)
    p = Pair(chain, B, precision, q0=q0_of(chain))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, rng.uniform(-0.5, 0.5, (B, n))); p.both("update_model")
    t0 = p.add_jt()
    if params:
        p.set_jt_params(t0, **params)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.2, (B, n)), t0)
    p.set(abi.F_JT_GOAL_VELOCITY, rng.normal(0, 0.1, (B, n)), t0)
    p.set(abi.F_JT_GOAL_ACCELERATION, rng.normal(0, 0.1, (B, n)), t0)
    return p, t0, q, rng


@pytest.mark.parametrize("name", ["puma560_synthetic", "panda", "panda_sliding_base", "rrrrbot"])
@pytest.mark.parametrize("decoupling", [abi.BOUNDED_INERTIA_ESTIMATES, abi.FULL_DYNAMIC_DECOUPLING, abi.IMPEDANCE])
def test_joint_task_alone_is_fused(oracle_built, name, decoupling):
    """every fixture chain (n = 4, 6, 7, 8; the sliding base has a prismatic joint), three decoupling modes, ki != 0 over three ticks"""
    p, t0, q, rng = _pair(name, 512, 91 + decoupling, dynamic_decoupling_type=decoupling, ki=2.5)
    assert p.gpu.tick_is_fused(), "a full JointTask alone must take the fused tick"
    l0 = p.gpu.launch_count()
    for tick in range(3):
        p.set(abi.F_Q, q + 0.01 * tick * rng.normal(0, 1, q.shape))
        tg, tr = p.both("tick", True)
        e, where = rel_err(tg, tr)
        assert e <= FP64_TOL, "%.3e at robot %d" % (e, where)
    p.compare(FP64_TOL)   # every getter afterwards (lazy materialisation): Lambda = M, N = 0, range = I, integrated error ...
    assert np.abs(np.asarray(p.gpu.get(abi.F_N_TOTAL))).max() <= 1e-9


def test_joint_task_alone_velocity_saturation_gravity_and_otg(oracle_built):
    """examples/02: the task's internal OTG (acceleration-limited, the reference default) feeds the desired state; velocity saturation
    (JointTask.cpp:327-340) and gravity compensation (RobotController.cpp:70-72) on top; 400 ticks with a goal step"""
    p, t0, q, rng = _pair("puma560_synthetic", 256, 97, use_velocity_saturation=1, saturation_velocity=0.4)
    p.both("enable_gravity_compensation", True)
    p.both("enable_internal_otg", t0, np.pi / 3, 2 * np.pi)
    goal = p.ref.get(abi.F_JT_GOAL_POSITION, t0).copy()
    p.set(abi.F_JT_GOAL_VELOCITY, np.zeros_like(q), t0); p.set(abi.F_JT_GOAL_ACCELERATION, np.zeros_like(q), t0)
    assert p.gpu.tick_is_fused()
    worst = 0.0
    for tick in range(400):
        if tick == 150:
            goal[:, 1] += 0.5
            p.set(abi.F_JT_GOAL_POSITION, goal, t0)
        p.set(abi.F_Q, q + 0.05 * np.sin(2 * np.pi * 1.0 * tick * 1e-3 + np.arange(q.shape[1]))[None, :])
        tg, tr = p.both("tick", True)
        worst = max(worst, rel_err(tg, tr)[0])
    assert worst <= FP64_TOL, worst
    assert rel_err(p.gpu.get(abi.F_JT_DESIRED_POSITION, t0), p.ref.get(abi.F_JT_DESIRED_POSITION, t0))[0] <= 1e-9


def test_joint_task_alone_fp32_and_large_batch(oracle_built):
    p, t0, q, rng = _pair("panda", 40000, 99, precision=abi.PRECISION_F32)
    assert p.gpu.tick_is_fused()
    tg, tr = p.both("tick", True)
    assert rel_err(np.asarray(tg, float), tr)[0] <= FP32_TOL


@pytest.mark.parametrize("name", ["puma560_synthetic", "panda"])
def test_joint_task_alone_large_batch_block_shape(oracle_built, name):
    """224-thread blocks of k_tick_joint2 (from 33,152 robots on), ragged last block, gravity compensation through k_gravity"""
    p, t0, q, rng = _pair(name, 224 * 148 + 5, 101)
    p.both("enable_gravity_compensation", True)
    assert p.gpu.tick_is_fused()
    tg, tr = p.both("tick", True)
    e, where = rel_err(tg, tr)
    assert e <= FP64_TOL, "%.3e at robot %d" % (e, where)
    assert rel_err(p.gpu.get(abi.F_TASK_TORQUES, t0), p.ref.get(abi.F_TASK_TORQUES, t0))[0] <= FP64_TOL
