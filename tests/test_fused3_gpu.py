"""k_tick_fused3 (csrc/osc_fused3.cuh): the single-launch tick for a PARTIAL MotionForceTask of rank M over a full JointTask --
BASELINE config 3 (translation-only task, 3 force axes / 1 force axis, closed-loop force control with the passivity observer,
examples/09:113-121,176-182) and general direction sets of rank 2, 4, 5 -- against the C++ oracle at 1e-9, tick by tick, with the
getters of the reference answering afterwards (lazy materialisation)."""
import numpy as np
import pytest

import osc_abi as abi
import workloads
from common import FP32_TOL, FP64_TOL, Pair, load_chain, q0_of, random_q, rel_err

pytestmark = pytest.mark.gpu


def _rot(axis, ang):
    a = np.asarray(axis, float) / np.linalg.norm(axis)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K


Q = _rot((1, 2, 3), 0.7)   # a frame that is not aligned with the world axes
CASES = {   # name: (translation directions, rotation directions) -> task rank
    "config3_translation_only": ([(1, 0, 0), (0, 1, 0), (0, 0, 1)], []),
    "rank2_skew_plane": ([Q[:, 0], Q[:, 1]], []),
    "rank4_plane_plus_two_rotations": ([Q[:, 0], Q[:, 2]], [Q[:, 1], Q[:, 2]]),
    "rank5_all_but_one_rotation": ([(1, 0, 0), (0, 1, 0), (0, 0, 1)], [Q[:, 0], Q[:, 1]]),
    "rank3_redundant_direction_list": ([(1, 0, 0), (2, 0, 0), (0, 0, 1)], [(0, 1, 0)]),
}


def _setup(chain, B, seed, dirs_t, dirs_r, precision=abi.PRECISION_F64, **params):
    rng = np.random.default_rng(seed)
    n = chain["n"]
    p = Pair(chain, B, precision, q0=q0_of(chain))
    q = random_q(chain, B, rng, 0.3); dq = rng.uniform(-0.5, 0.5, (B, n))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.both("update_model")
    t0 = p.add_mft("end-effector", (0, 0, 0.07), dirs_translation=dirs_t, dirs_rotation=dirs_r)
    t1 = p.add_jt()
    if params:
        p.set_mft_params(t0, **params)
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    p.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_LINEAR_VELOCITY, rng.normal(0, 0.05, (B, 3)), t0); p.set(abi.F_GOAL_ANGULAR_VELOCITY, rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_LINEAR_ACCELERATION, rng.normal(0, 0.1, (B, 3)), t0)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, (B, n)), t1); p.set(abi.F_JT_GOAL_VELOCITY, rng.normal(0, 0.1, (B, n)), t1)
    p.set(abi.F_JT_GOAL_ACCELERATION, rng.normal(0, 0.1, (B, n)), t1)
    return p, t0, t1, q, rng


@pytest.mark.parametrize("case", sorted(CASES))
@pytest.mark.parametrize("decoupling", [abi.BOUNDED_INERTIA_ESTIMATES, abi.FULL_DYNAMIC_DECOUPLING, abi.IMPEDANCE])
def test_partial_task_fused_tick_matches_oracle(oracle_built, case, decoupling):
    chain = load_chain("panda")
    B = 1024
    dirs_t, dirs_r = CASES[case]
    p, t0, t1, q, rng = _setup(chain, B, 7 + len(case), dirs_t, dirs_r, dynamic_decoupling_type=decoupling)
    p.set_jt_params(t1, dynamic_decoupling_type=decoupling)
    assert p.gpu.tick_is_fused(), "a partial task over a full joint task must take k_tick_fused3"
    rank = p.gpu.task_dof(t0)
    for tick in range(3):
        p.set(abi.F_Q, q + 0.01 * tick * rng.normal(0, 1, q.shape))
        for stacked in (True, False):
            tg, tr = p.both("tick", stacked)
            assert rel_err(tg, tr)[0] <= FP64_TOL, (tick, stacked)
    fb = p.gpu.tick_fallback_count()
    blend = int((p.ref.get(abi.F_ALPHA, t0)[:, 0] < 1).sum())
    assert blend <= fb <= blend + B // 10, (fb, blend)       # the fast path really served the non-singular robots
    assert fb < B // 2
    assert (p.ref.get(abi.F_JT_RANGE_RANK, t1)[p.ref.get(abi.F_ALPHA, t0)[:, 0] == 1] == 7 - rank).all()
    p.compare(FP64_TOL)


@pytest.mark.parametrize("force_dim", [3, 1])
def test_config3_through_the_fused_tick_600_ticks(oracle_built, force_dim):
    """BASELINE config 3 with osc_tick: sensed wrench in, tick out, 600 ticks (POPC window 250, controller period 50)"""
    chain = load_chain("panda")
    B, T = 96, 600
    p, t0, t1, q, rng = _setup(chain, B, 30 + force_dim, *CASES["config3_translation_only"], force_space_dimension=force_dim,
                               force_or_motion_axis=[0, 0, 1.0], closed_loop_force_control=1, passivity_enabled=1)
    p.set(abi.F_GOAL_FORCE, np.array([0, 0, -5.0]), t0)
    assert p.gpu.tick_is_fused()
    worst, rc_moved, launches0 = 0.0, False, p.gpu.launch_count()
    for tick in range(T):
        f, _ = workloads.config3_wrench(B, tick)
        f = f * 4.0 if tick % 100 > 60 else f
        p.set(abi.F_Q, q + 0.05 * np.sin(2 * np.pi * 0.5 * tick * 0.001 + np.arange(7))[None, :])
        # osc_mft_update_sensed_force_and_moment around a fused tick evaluates the compliant frame at the batch's CURRENT q, i.e.
        # the reference's order robot->updateModel(); task->updateSensedForceAndMoment() (examples/07...cpp:150-175): mirrored on the oracle
        p.ref.update_model()
        if tick == 0:
            p.gpu.update_model()   # no fused tick has run yet: the product, like the reference, still holds the kinematics of the setup posture
        p.both("update_sensed_force_and_moment", t0, f, np.zeros_like(f))
        tg, tr = p.both("tick", True)
        worst = max(worst, rel_err(tg, tr)[0])
        if tick % 50 == 49:
            rcg, rcr = p.gpu.get(abi.F_POPC_RC, t0), p.ref.get(abi.F_POPC_RC, t0)
            assert np.abs(rcg - rcr).max() <= 1e-9
            rc_moved = rc_moved or bool((rcr < 1).any())
    assert worst <= FP64_TOL, worst
    assert rc_moved
    p.compare(FP64_TOL)


def test_partial_task_fp32(oracle_built):
    chain = load_chain("panda")
    p, t0, t1, q, rng = _setup(chain, 512, 44, *CASES["rank4_plane_plus_two_rotations"], precision=abi.PRECISION_F32)
    assert p.gpu.tick_is_fused()
    tg, tr = p.both("tick", True)
    alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
    err = np.abs(np.asarray(tg, float) - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-2)
    assert err[alpha == 1].max() <= FP32_TOL, err[alpha == 1].max()
