"""A 6-DoF arm under ONE full MotionForceTask (examples/19-puma_singularity: the PUMA; the most common OSC set-up) on the fused tick:
hierarchy [MotionForceTask] with n = 6 has no nullspace, k_tick_fused2<.., 6> serves it with L_BIE = chol(M_BIE) in place of L for any
bounded-inertia clamp pattern (the PUMA clamps its three wrist joints), and the robots it cannot certify go through the closed-form
blend routine (osc_blend.cuh, n = 6: L_BIE kept beside L) -- against the C++ oracle."""
import numpy as np
import pytest

import osc_abi as abi
from common import FP32_TOL, FP64_TOL, Pair, load_chain, q0_of, random_q
from test_singular_gpu import _puma_q

pytestmark = pytest.mark.gpu


def _tau_err(tg, tr):
    tg, tr = np.asarray(tg, float), np.asarray(tr, float)
    return np.abs(tg - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-3 * max(np.abs(tr).max(), 1e-300))


def _pair(q, rng, precision=abi.PRECISION_F64, with_jt=False, **params):
    chain = load_chain("puma560_synthetic")
    B = len(q)
    p = Pair(chain, B, precision, q0=q0_of(chain))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, rng.uniform(-0.3, 0.3, q.shape)); p.both("update_model")
    t0 = p.add_mft("end-effector", (0, 0, 0))
    t1 = p.add_jt() if with_jt else None
    if params:
        p.set_mft_params(t0, **params)
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    p.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_LINEAR_VELOCITY, rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_ANGULAR_VELOCITY, rng.normal(0, 0.05, (B, 3)), t0)
    return p, t0, t1


@pytest.mark.parametrize("decoupling", [abi.BOUNDED_INERTIA_ESTIMATES, abi.FULL_DYNAMIC_DECOUPLING, abi.IMPEDANCE])
def test_single_task_six_dof_arm_is_fused(oracle_built, decoupling):
    chain = load_chain("puma560_synthetic")
    rng = np.random.default_rng(61 + decoupling)
    q = random_q(chain, 1024, rng, 0.3)
    p, t0, _ = _pair(q, rng, dynamic_decoupling_type=decoupling)
    assert p.gpu.tick_is_fused(), "one full MotionForceTask on a 6-DoF revolute arm must take the fused tick"
    for tick in range(3):
        p.set(abi.F_Q, q + 0.01 * tick * rng.normal(0, 1, q.shape))
        tg, tr = p.both("tick", True)
        alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
        e = _tau_err(tg, tr)
        assert e[alpha == 1].max() <= FP64_TOL, "non-singular robots: %.3e" % e[alpha == 1].max()
        assert e.max() <= 1e-7, "blend robots: %.3e" % e.max()   # IMPEDANCE takes the API-faithful chain: kappa-scaled
    alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
    fb = p.gpu.tick_fallback_count()
    assert int((alpha < 1).sum()) <= fb <= int((alpha < 1).sum()) + 1024 // 8, (fb, int((alpha < 1).sum()))   # the clamp pattern no longer sends every robot to the tail
    p.compare(1e-6)   # every getter afterwards (lazy materialisation)


def test_single_task_in_the_singular_set_flat_tolerance(oracle_built):
    """The wrist-lock set of BASELINE config 4 (examples/19:76-80 +- eps): through the closed-form blend routine, bounded inertia
    estimates with the PUMA's three clamped wrist joints."""
    rng = np.random.default_rng(67)
    q = _puma_q(rng)
    p, t0, _ = _pair(q, rng, kp_type_1=50.0, kv_type_1=20.0, kv_type_2=20.0)   # setSingularityHandlingGains(50, 20, 20), examples/19
    assert p.gpu.tick_is_fused()
    for tick in range(6):
        p.set(abi.F_Q, q + 2e-4 * tick * rng.normal(0, 1, q.shape))
        tg, tr = p.both("tick", True)
        assert _tau_err(tg, tr).max() <= FP64_TOL
        for f in (abi.F_SINGULARITY_TYPES, abi.F_SINGULARITY_COUNTERS, abi.F_NS_RANK):
            assert (np.asarray(p.gpu.get(f, t0)) == p.ref.get(f, t0)).all()
    alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
    assert (alpha < 1).sum() >= len(q) // 2


def test_six_dof_arm_with_a_joint_task_below(oracle_built):
    """[MotionForceTask, JointTask] on n = 6: the joint task's range is empty (JointTask.cpp:234-239), also for blend robots."""
    rng = np.random.default_rng(71)
    q = _puma_q(rng)
    p, t0, t1 = _pair(q, rng, with_jt=True)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, q.shape), t1)
    assert p.gpu.tick_is_fused()
    for tick in range(3):
        p.set(abi.F_Q, q + 2e-4 * tick * rng.normal(0, 1, q.shape))
        for stacked in (True, False):
            tg, tr = p.both("tick", stacked)
            assert _tau_err(tg, tr).max() <= FP64_TOL
    assert np.abs(np.asarray(p.gpu.get(abi.F_TASK_TORQUES, t1))).max() == 0.0


def test_single_task_fp32(oracle_built):
    chain = load_chain("puma560_synthetic")
    rng = np.random.default_rng(73)
    q = random_q(chain, 512, rng, 0.3)
    p, t0, _ = _pair(q, rng, precision=abi.PRECISION_F32)
    assert p.gpu.tick_is_fused()
    tg, tr = p.both("tick", True)
    alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
    assert _tau_err(tg, tr)[alpha == 1].max() <= FP32_TOL


# ---- n = 7 with the motion task as the whole hierarchy (examples/07-surface_surface_contact: RobotController over one task) ----
def _panda_single(q, rng, precision=abi.PRECISION_F64, **params):
    chain = load_chain("panda")
    B = len(q)
    p = Pair(chain, B, precision, q0=q0_of(chain))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, rng.uniform(-0.3, 0.3, q.shape)); p.both("update_model")
    t0 = p.add_mft("end-effector", (0, 0, 0.07))
    if params:
        p.set_mft_params(t0, **params)   # parametrize* re-initialises the goal pose: goals are set afterwards
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    p.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_LINEAR_VELOCITY, rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_ANGULAR_VELOCITY, rng.normal(0, 0.05, (B, 3)), t0)
    return p, t0


@pytest.mark.parametrize("decoupling", [abi.BOUNDED_INERTIA_ESTIMATES, abi.FULL_DYNAMIC_DECOUPLING, abi.IMPEDANCE])
@pytest.mark.parametrize("B", [1024, 224 * 148 + 5])
def test_single_task_seven_dof_arm_is_fused(oracle_built, decoupling, B):
    """Both block shapes of k_tick_fused2<.., SINGLE> (64 threads; 224 with phase barriers from 33,152 robots on), plain motion task
    (the SIMPLE instantiation), stacked and plain sums, gravity compensation on the larger batch."""
    chain = load_chain("panda")
    rng = np.random.default_rng(81 + decoupling)
    q = random_q(chain, B, rng, 0.3)
    p, t0 = _panda_single(q, rng, dynamic_decoupling_type=decoupling)
    assert p.gpu.tick_is_fused(), "one full MotionForceTask on a 7-DoF revolute arm must take the fused tick"
    if B > 2000:
        p.both("enable_gravity_compensation", True)
    for tick in range(2):
        p.set(abi.F_Q, q + 0.01 * tick * rng.normal(0, 1, q.shape))
        for stacked in (True, False):
            tg, tr = p.both("tick", stacked)
            alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
            e = _tau_err(tg, tr)
            assert e[alpha == 1].max() <= FP64_TOL, "non-singular robots: %.3e" % e[alpha == 1].max()
            assert e.max() <= 1e-7, "blend robots: %.3e" % e.max()
    alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
    fb = p.gpu.tick_fallback_count()
    print("decoupling %d: %d of %d robots through k_tick_tail, %d of them in the blend band" % (decoupling, fb, B, int((alpha < 1).sum())))
    # over 40 % of the joint range more than half of the postures sit in the reference's blend band (a 6 x 7 Jacobian in mixed units); the
    # certificate of the fused kernel adds its margin (a few per cent of the batch)
    assert int((alpha < 1).sum()) <= fb <= int((alpha < 1).sum()) + B // 8
    if B < 2000:
        p.compare(1e-6)   # every getter afterwards (lazy materialisation)


def test_single_task_seven_dof_in_the_singular_set(oracle_built):
    """examples/18's postures (elbow extended, wrist aligned) under the single task: the closed-form blend routine without a joint
    task below it, flat 1e-9, discrete classification state equal to the oracle's."""
    from test_singular_gpu import _panda_q
    rng = np.random.default_rng(83)
    q = _panda_q(rng)
    p, t0 = _panda_single(q, rng)
    assert p.gpu.tick_is_fused()
    for tick in range(5):
        p.set(abi.F_Q, q + 2e-4 * tick * rng.normal(0, 1, q.shape))
        tg, tr = p.both("tick", True)
        e = _tau_err(tg, tr)
        assert e.max() <= FP64_TOL, "%.3e at robot %d" % (e.max(), int(e.argmax()))
        for f in (abi.F_SINGULARITY_TYPES, abi.F_SINGULARITY_COUNTERS, abi.F_NS_RANK):
            assert (np.asarray(p.gpu.get(f, t0)) == p.ref.get(f, t0)).all()
    alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
    assert (alpha < 1).sum() >= len(q) // 2 and p.gpu.tick_fallback_count() >= int((alpha < 1).sum())


def test_single_task_seven_dof_surface_contact_control(oracle_built):
    """examples/07: force control along the tool axis and moment control about the two others, closed loop with the passivity
    observer, sensed wrench refreshed every tick -- the general (non-SIMPLE) SINGLE instantiation, 300 ticks so that the POPC
    controller period fires and the integrators matter."""
    chain = load_chain("panda")
    rng = np.random.default_rng(87)
    B = 192
    q = random_q(chain, B, rng, 0.25)
    p, t0 = _panda_single(q, rng, force_space_dimension=1, force_or_motion_axis=(0, 0, 1), moment_space_dimension=2, moment_or_rotmotion_axis=(0, 0, 1),
                          closed_loop_force_control=1, closed_loop_moment_control=1, passivity_enabled=1)
    assert p.gpu.tick_is_fused()
    p.set(abi.F_GOAL_FORCE, np.tile([0, 0, -5.0], (B, 1)), t0)
    amp = rng.normal(0, 0.02, q.shape)
    worst = 0.0
    for tick in range(300):
        t = tick * 1e-3
        p.set(abi.F_Q, q + amp * np.sin(2 * np.pi * 2.0 * t)); p.set(abi.F_DQ, amp * 2 * np.pi * 2.0 * np.cos(2 * np.pi * 2.0 * t))
        f = np.tile([0, 0, 5.0], (B, 1)) + rng.normal(0, 1.0, (B, 3)); m = rng.normal(0, 0.1, (B, 3))
        # the reference's order robot->updateModel(); task->updateSensedForceAndMoment() (examples/07...cpp:150-175): the product evaluates
        # the compliant frame at the batch's current q once a fused tick has run (stale rows), see tests/test_fused3_gpu.py
        p.ref.update_model()
        if tick == 0:
            p.gpu.update_model()
        p.both("update_sensed_force_and_moment", t0, f, m)
        tg, tr = p.both("tick", True)
        worst = max(worst, _tau_err(tg, tr).max())
    assert worst <= FP64_TOL, worst
    rc = p.ref.get(abi.F_POPC_RC, t0)[:, 0]
    assert np.abs(np.asarray(p.gpu.get(abi.F_POPC_RC, t0))[:, 0] - rc).max() <= 1e-9
    p.compare(1e-6)


def test_single_task_seven_dof_fp32(oracle_built):
    chain = load_chain("panda")
    rng = np.random.default_rng(89)
    q = random_q(chain, 512, rng, 0.3)
    p, t0 = _panda_single(q, rng, precision=abi.PRECISION_F32)
    assert p.gpu.tick_is_fused()
    tg, tr = p.both("tick", True)
    alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
    assert _tau_err(tg, tr)[alpha == 1].max() <= FP32_TOL
