"""The singular-blend branch in closed form (csrc/osc_blend.cuh, tried by k_tick_tail before the API-faithful chain) against the
C++ oracle: SingularityHandler.cpp:75-368 for [MotionForceTask of rank M, full JointTask], n = 7.

What is asserted here that tests/test_singular_gpu.py (kappa-scaled tolerances, written for the chain) does not:
  * a FLAT 1e-9 on tau, on the task torques of both tasks and on alpha for robots deep inside the blend band -- the routine never
    inverts a matrix that contains the 1 / sigma^2 of the singular block, so it has no kappa term to concede;
  * the discrete classification state (types, counters, q_prior) equal to the oracle's from the first tick: the type of a singular
    direction comes from a pose change that is SECOND order in the perturbation step, it depends on the sign convention of the
    singular vector pair;
  * both joint-space strategies (type 1 enforced / type 2 by history), bounded-inertia and full decoupling, stacked and plain sums;
  * the configurations the routine DECLINES (IMPEDANCE decoupling, singularity handling off) still give the oracle's torques through
    the chain behind it;
  * partial tasks (ranks 3 and 4) driven into their own singular set."""
import numpy as np
import pytest

import osc_abi as abi
from common import FP64_TOL, Pair, load_chain, q0_of, rel_err
from test_singular_gpu import _panda_q

pytestmark = pytest.mark.gpu


def _tau_err(tg, tr):
    tg, tr = np.asarray(tg, float), np.asarray(tr, float)
    return np.abs(tg - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-3 * max(np.abs(tr).max(), 1e-300))


def _panda_pair(seed, **mft_params):
    chain = load_chain("panda")
    rng = np.random.default_rng(seed)
    q = _panda_q(rng)
    B = len(q)
    p = Pair(chain, B, abi.PRECISION_F64, q0=q0_of(chain))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, rng.uniform(-0.3, 0.3, q.shape)); p.both("update_model")
    t0 = p.add_mft("end-effector", (0, 0, 0.07)); t1 = p.add_jt()
    if mft_params:
        p.set_mft_params(t0, **mft_params)
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    p.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_LINEAR_VELOCITY, rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, q.shape), t1)
    p.set(abi.F_JT_GOAL_VELOCITY, rng.normal(0, 0.1, q.shape), t1)
    return p, t0, t1, q, rng


def _check_tick(p, t0, t1, stacked, tol=FP64_TOL):
    tg, tr = p.both("tick", stacked)
    e = _tau_err(tg, tr)
    assert e.max() <= tol, "tau: %.3e at robot %d" % (e.max(), int(e.argmax()))
    for t in (t0, t1):
        e = _tau_err(p.gpu.get(abi.F_TASK_TORQUES, t), p.ref.get(abi.F_TASK_TORQUES, t))
        assert e.max() <= tol, "task %d torques: %.3e at robot %d" % (t, e.max(), int(e.argmax()))
    assert np.abs(np.asarray(p.gpu.get(abi.F_ALPHA, t0)) - p.ref.get(abi.F_ALPHA, t0)).max() <= 1e-9
    for f in (abi.F_SINGULARITY_TYPES, abi.F_SINGULARITY_COUNTERS, abi.F_NS_RANK):
        assert (np.asarray(p.gpu.get(f, t0)) == p.ref.get(f, t0)).all(), "discrete state %d differs" % f
    assert rel_err(p.gpu.get(abi.F_Q_PRIOR, t0), p.ref.get(abi.F_Q_PRIOR, t0))[0] <= 1e-12


@pytest.mark.parametrize("decoupling", [abi.BOUNDED_INERTIA_ESTIMATES, abi.FULL_DYNAMIC_DECOUPLING])
@pytest.mark.parametrize("enforce_type_1", [0, 1])
def test_blend_routine_flat_tolerance_and_discrete_state(oracle_built, decoupling, enforce_type_1):
    p, t0, t1, q, rng = _panda_pair(11 + decoupling, dynamic_decoupling_type=decoupling, enforce_type_1_strategy=enforce_type_1)
    p.set_jt_params(t1, dynamic_decoupling_type=decoupling)
    assert p.gpu.tick_is_fused()
    alpha = None
    for tick in range(6):
        p.set(abi.F_Q, q + 2e-4 * tick * rng.normal(0, 1, q.shape))   # small moves: robots cross the band edges and change type
        p.set(abi.F_DQ, rng.uniform(-0.3, 0.3, q.shape))
        for stacked in (True, False):
            _check_tick(p, t0, t1, stacked)
        alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
    assert (alpha < 1).sum() >= len(q) // 2 and (alpha == 0).sum() >= 5      # the set really sits in the blend band, some at alpha = 0
    assert p.gpu.tick_fallback_count() >= int((alpha < 1).sum())            # ... and those robots went through k_tick_tail
    types = p.ref.get(abi.F_SINGULARITY_TYPES, t0)
    if not enforce_type_1:
        assert (types == 1).any() and (types == 2).any()                    # both classifications occur
    p.compare(1e-6)   # every getter of the reference after the tick (lazy materialisation through the API-faithful kernels)


def test_blend_routine_with_gravity_compensation_velocity_saturation_and_fp32(oracle_built):
    """Gravity compensation (RobotController.h:31-33) is added by the routine itself; velocity saturation of both tasks (examples/18:
    130-131) changes the control law the fused kernel hands over; an FP32 batch runs the same routine in double over float rows."""
    p, t0, t1, q, rng = _panda_pair(31, use_velocity_saturation=1, linear_saturation_velocity=0.3, angular_saturation_velocity=0.5)
    p.both("enable_gravity_compensation", True)
    for tick in range(3):
        p.set(abi.F_Q, q + 2e-4 * tick * rng.normal(0, 1, q.shape))
        for stacked in (True, False):
            _check_tick(p, t0, t1, stacked)
    chain = load_chain("panda")
    import osc_b200
    f = osc_b200.Batch(chain, len(q), precision=abi.PRECISION_F32, q0=q0_of(chain))
    f.set(abi.F_Q, q); f.set(abi.F_DQ, np.zeros_like(q)); f.update_model()
    f0 = f.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07))); f1 = f.add_joint_task(abi.joint_task_desc(7))
    p2, s0, s1, _, _ = _panda_pair(31)
    p2.set(abi.F_DQ, np.zeros_like(q))
    f.set(abi.F_GOAL_POSITION, p2.ref.get(abi.F_GOAL_POSITION, s0), f0); f.set(abi.F_GOAL_LINEAR_VELOCITY, p2.ref.get(abi.F_GOAL_LINEAR_VELOCITY, s0), f0)
    f.set(abi.F_JT_GOAL_POSITION, p2.ref.get(abi.F_JT_GOAL_POSITION, s1), f1); f.set(abi.F_JT_GOAL_VELOCITY, p2.ref.get(abi.F_JT_GOAL_VELOCITY, s1), f1)
    tr = p2.ref.tick(True)
    tf = np.asarray(f.tick(True), float)
    same_branch = np.asarray(f.get(abi.F_NS_RANK, f0))[:, 0] == p2.ref.get(abi.F_NS_RANK, s0)[:, 0]
    assert same_branch.mean() > 0.9       # float inputs may put a robot on the other side of a band edge
    assert _tau_err(tf[same_branch], tr[same_branch]).max() <= 1e-3   # q itself is rounded to float: kappa * 6e-8 inside the band


@pytest.mark.parametrize("declined_by", ["impedance", "handling_off"])
def test_configurations_the_routine_declines_take_the_chain(oracle_built, declined_by):
    kw = {"dynamic_decoupling_type": abi.IMPEDANCE} if declined_by == "impedance" else {"singularity_handling_enabled": 0}
    p, t0, t1, q, rng = _panda_pair(23, **kw)
    for tick in range(3):
        p.set(abi.F_Q, q + 2e-4 * tick * rng.normal(0, 1, q.shape))
        tg, tr = p.both("tick", True)
        assert _tau_err(tg, tr).max() <= 1e-6   # the chain's tolerance inside the band (kappa up to 1e7, tests/test_singular_gpu.py)
    assert p.gpu.tick_fallback_count() > 0


def _blend_candidates(chain, rank, dirs_r, rng, want):
    """Random postures of which the oracle says that the partial task sits in its blend band (about 5 % of the joint range)."""
    import orc
    from common import random_q
    N = 8192
    q = random_q(chain, N, rng, 0.3)
    o = orc.OracleBatch(chain, N, q0=q0_of(chain))
    o.set(abi.F_Q, q); o.set(abi.F_DQ, np.zeros_like(q)); o.update_model()
    t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07), dirs_translation=[(1, 0, 0), (0, 1, 0), (0, 0, 1)], dirs_rotation=dirs_r))
    o.add_joint_task(abi.joint_task_desc(chain["n"]))
    o.tick(True)
    alpha = o.get(abi.F_ALPHA, t0)[:, 0]
    blend = np.nonzero(alpha < 1)[0]
    clean = np.nonzero(alpha == 1)[0]
    assert len(blend) >= want // 2, "only %d blend postures among %d" % (len(blend), N)
    pick = np.concatenate([blend[:want - want // 4], clean[:want // 4]])
    return q[pick]


@pytest.mark.parametrize("rank", [3, 4])
def test_partial_task_in_its_singular_set(oracle_built, rank):
    """Translation-only (rank 3) and translation + one rotation (rank 4) tasks at postures where the oracle finds the task in its
    blend band (arm towards full extension), a quarter of the batch outside it."""
    chain = load_chain("panda")
    rng = np.random.default_rng(5 + rank)
    dirs_r = [] if rank == 3 else [(0, 0, 1)]
    q = _blend_candidates(chain, rank, dirs_r, rng, 256)
    B = len(q)
    p = Pair(chain, B, abi.PRECISION_F64, q0=q0_of(chain))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, rng.uniform(-0.3, 0.3, q.shape)); p.both("update_model")
    t0 = p.add_mft("end-effector", (0, 0, 0.07), dirs_translation=[(1, 0, 0), (0, 1, 0), (0, 0, 1)], dirs_rotation=dirs_r)
    t1 = p.add_jt()
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    p.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, q.shape), t1)
    assert p.gpu.tick_is_fused() and p.gpu.task_dof(t0) == rank
    for tick in range(4):
        p.set(abi.F_Q, q + 2e-4 * tick * rng.normal(0, 1, q.shape))
        for stacked in (True, False):
            _check_tick(p, t0, t1, stacked)
    alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
    assert (alpha < 1).sum() >= B // 3, "the set does not sit in the blend band: %d of %d" % (int((alpha < 1).sum()), B)
