"""BASELINE config 3 on the GPU: Panda, PARTIAL MotionForceTask (translation only, examples/09:114-121), force axes from
parametrizeForceMotionSpaces(3) and (1, UnitZ) (examples/09:178-182), closed-loop force control with the passivity
observer / controller (MotionForceTask.cpp:511-569), full JointTask in the nullspace, RobotController stacking.

Small batches run tick by tick beside the C++ oracle (FP64, 1e-9); the full 262,144-robot batch is checked through
size-independent properties and through oracle parity of a random subset of its robots (robots are independent, so a
robot's trajectory does not depend on the batch it sits in)."""
import numpy as np
import pytest

import osc_abi as abi
import workloads
from common import FP64_TOL, Pair, load_chain, q0_of, random_q

pytestmark = pytest.mark.gpu
DIRS = [(1, 0, 0), (0, 1, 0), (0, 0, 1)]


def _tau_err(tg, tr):
    return float((np.abs(tg - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-3 * np.abs(tr).max())).max())


def _configure(b, chain, q, dq, force_dim, jt_goal):
    """the task setup of config 3 on one batch object (product or oracle: same method names)"""
    b.set(abi.F_Q, q); b.set(abi.F_DQ, dq); b.update_model()
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07), dirs_translation=DIRS, dirs_rotation=[]))
    t1 = b.add_joint_task(abi.joint_task_desc(chain["n"]))
    p = b.mft_params(t0)
    p.force_space_dimension = force_dim
    p.force_or_motion_axis[0], p.force_or_motion_axis[1], p.force_or_motion_axis[2] = 0.0, 0.0, 1.0
    p.closed_loop_force_control = 1
    p.passivity_enabled = 1
    b.set_mft_params(t0, p)
    b.set(abi.F_GOAL_FORCE, np.array([0, 0, -5.0]), t0)
    b.set(abi.F_JT_GOAL_POSITION, jt_goal, t1)
    return t0, t1


def _drive(b, t0, q, tick, f, dt=0.001):
    """one control cycle: kinematic playback of q (no simulator), sensed wrench, the three calls of the loop"""
    b.set(abi.F_Q, q + 0.05 * np.sin(2 * np.pi * 0.5 * tick * dt + np.arange(q.shape[1]))[None, :])
    b.update_model()
    b.update_task_models()
    b.update_sensed_force_and_moment(t0, f, np.zeros_like(f))
    return b.compute_torques(True)


@pytest.mark.parametrize("force_dim", [3, 1])
def test_config3_parity_600_ticks(oracle_built, force_dim):
    """T = 600 > POPC window (250) and > the 50-tick controller period: the ring buffer wraps, R_c moves"""
    import orc
    import osc_b200
    chain = load_chain("panda")
    B, T = 96, 600
    rng = np.random.default_rng(30 + force_dim)
    q = random_q(chain, B, rng, 0.3)
    dq = rng.uniform(-0.5, 0.5, (B, 7))
    gq = q + rng.normal(0, 0.1, (B, 7))
    g = osc_b200.Batch(chain, B, q0=q0_of(chain))
    r = orc.OracleBatch(chain, B, q0=q0_of(chain))
    t0, _ = _configure(g, chain, q, dq, force_dim, gq)
    _configure(r, chain, q, dq, force_dim, gq)
    worst, rc_moved = 0.0, False
    for tick in range(T):
        f, _ = workloads.config3_wrench(B, tick)
        f = f * 4.0 if tick % 100 > 60 else f  # bursts of active behaviour so that the passivity controller engages
        tg, tr = _drive(g, t0, q, tick, f), _drive(r, t0, q, tick, f)
        worst = max(worst, _tau_err(np.asarray(tg), np.asarray(tr)))
        if tick % 50 == 49:
            rcg, rcr = g.get(abi.F_POPC_RC, t0), r.get(abi.F_POPC_RC, t0)
            assert np.abs(rcg - rcr).max() <= 1e-9
            rc_moved = rc_moved or bool((rcr < 1).any())
    assert worst <= FP64_TOL, worst
    assert rc_moved, "the drive never made the passivity controller reduce R_c: the test does not exercise it"
    for fld in (abi.F_INTEGRATED_FORCE_ERROR, abi.F_POPC_OBSERVER, abi.F_N_TOTAL, abi.F_LAMBDA_WORLD, abi.F_JACOBIAN):
        a, b_ = g.get(fld, t0 if fld != abi.F_N_TOTAL else -1), r.get(fld, t0 if fld != abi.F_N_TOTAL else -1)
        assert np.abs(a - b_).max() <= FP64_TOL * max(1.0, np.abs(b_).max()), fld


def test_config3_full_batch_properties(oracle_built):
    """262,144 robots x 60 ticks on the device: invariants of every robot + oracle parity of 64 robots picked at random"""
    import orc
    import osc_b200
    chain = load_chain("panda")
    B, T, S = 262144, 60, 64
    rng = np.random.default_rng(3)
    q = random_q(chain, B, rng, 0.3)
    dq = rng.uniform(-0.5, 0.5, (B, 7))
    gq = q + rng.normal(0, 0.1, (B, 7))
    pick = np.sort(rng.choice(B, S, replace=False))
    g = osc_b200.Batch(chain, B, q0=q0_of(chain))
    r = orc.OracleBatch(chain, S, q0=q0_of(chain))
    t0, t1 = _configure(g, chain, q, dq, 3, gq)
    _configure(r, chain, q[pick], dq[pick], 3, gq[pick])
    for tick in range(T):
        f, _ = workloads.config3_wrench(B, tick)
        tg, tr = np.asarray(_drive(g, t0, q, tick, f)), np.asarray(_drive(r, t0, q[pick], tick, f[pick]))
        assert _tau_err(tg[pick], tr) <= FP64_TOL
    assert np.isfinite(tg).all()
    rc = g.get(abi.F_POPC_RC, t0)
    assert ((rc >= 0) & (rc <= 1)).all()  # R_c stays in [0, 1] (MotionForceTask.cpp:552-563)
    # the projector handed down the hierarchy is idempotent and annihilates the task Jacobian: N N = N, J N = 0
    N = g.get(abi.F_N_TASK_AND_PREC, t0).reshape(B, 7, 7)
    J = g.get(abi.F_PROJECTED_JACOBIAN, t0).reshape(B, -1, 7)
    assert np.abs(N @ N - N).max() <= 1e-8
    assert np.abs(J @ N).max() <= 1e-8 * max(1.0, np.abs(J).max())
    # no error bit on any robot (bit 5 is informational: the robot sits in the blend branch of the SingularityHandler)
    assert ((g.get(abi.F_STATUS) & ~32) == 0).all()
