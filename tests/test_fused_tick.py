"""GPU parity of the single-launch tick (csrc/osc_fused.cuh): k_tick_fused for the robots it can certify non-singular,
k_tick_tail (API-faithful route) for the rest, against the C++ oracle, which follows the reference's route (two Jacobi
SVDs per tick) line by line.  Tolerance: 1e-9 relative in FP64 (BASELINE north_star)."""
import copy

import numpy as np
import pytest

import osc_abi as abi
from common import FP32_TOL, FP64_TOL, Pair, load_chain, random_q, rel_err, task_kappa

pytestmark = pytest.mark.gpu


def _two_task(chain, B, seed, precision=abi.PRECISION_F64, link="end-effector", p_c=(0, 0, 0.07), margin=0.1):
    rng = np.random.default_rng(seed)
    n = chain["n"]
    p = Pair(chain, B, precision)
    t0 = p.add_mft(link, p_c)
    t1 = p.add_jt()
    q = random_q(chain, B, rng, margin)
    dq = rng.uniform(-0.5, 0.5, (B, n))
    p.set(abi.F_Q, q)
    p.set(abi.F_DQ, dq)
    p.both("update_model")
    p.both("reinitialize_tasks")
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    p.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_LINEAR_VELOCITY, rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_ANGULAR_VELOCITY, rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_LINEAR_ACCELERATION, rng.normal(0, 0.1, (B, 3)), t0)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, (B, n)), t1)
    p.set(abi.F_JT_GOAL_VELOCITY, rng.normal(0, 0.1, (B, n)), t1)
    p.set(abi.F_JT_GOAL_ACCELERATION, rng.normal(0, 0.1, (B, n)), t1)
    return p, q, dq, rng


def _tau_err(tg, tr):
    return rel_err(tg, tr)[0]


def test_fused_is_selected_and_matches_oracle_over_ticks(oracle_built):
    chain = load_chain("panda")
    B = 2048
    p, q, dq, rng = _two_task(chain, B, 21)
    assert p.gpu.tick_is_fused()
    for tick in range(5):
        qq = q + 0.01 * tick * rng.normal(0, 1, q.shape)
        p.set(abi.F_Q, qq)
        tg, tr = p.both("tick", True)
        assert _tau_err(tg, tr) <= FP64_TOL
        fb = p.gpu.tick_fallback_count()
        blend = int((p.ref.get(abi.F_ALPHA, 0)[:, 0] < 1).sum())  # robots the reference puts in the blend / singular branch
        assert 0 < blend < B, "random Panda postures cover both branches"
        assert blend <= fb <= blend + B // 20, "fallback = blend-branch robots + the few the certificate cannot prove (%d vs %d)" % (fb, blend)
    # every getter of the reference still answers (lazy materialisation of the matrices)
    p.compare(FP64_TOL)


def test_fused_state_carries_like_the_per_call_route(oracle_built):
    """integrators (ki != 0) and the classification history advance identically through fused and per-call ticks"""
    chain = load_chain("panda")
    B = 512
    p, q, dq, rng = _two_task(chain, B, 22)
    p.set_mft_params(0, ki_pos=[3.0, 4.0, 5.0], ki_ori=[1.0, 2.0, 3.0])
    p.set_jt_params(1, ki=2.5)
    for tick in range(6):
        p.set(abi.F_Q, q + 0.02 * np.sin(0.3 * tick))
        if tick % 2 == 0:
            tg, tr = p.both("tick", True)
        else:  # per-call route on the same batch: shares all persistent state with the fused route
            p.both("update_model"); p.both("update_task_models")
            tg, tr = p.both("compute_torques", True)
        assert _tau_err(tg, tr) <= FP64_TOL, tick
    p.compare(FP64_TOL)


@pytest.mark.parametrize("decoupling", [abi.FULL_DYNAMIC_DECOUPLING, abi.BOUNDED_INERTIA_ESTIMATES, abi.IMPEDANCE])
@pytest.mark.parametrize("stacked", [True, False])
def test_fused_decoupling_modes(oracle_built, decoupling, stacked):
    chain = load_chain("panda")
    p, q, dq, rng = _two_task(chain, 512, 23 + decoupling)
    p.set_mft_params(0, dynamic_decoupling_type=decoupling)
    p.set_jt_params(1, dynamic_decoupling_type=decoupling)
    assert p.gpu.tick_is_fused()
    for _ in range(2):
        tg, tr = p.both("tick", stacked)
        assert _tau_err(tg, tr) <= FP64_TOL
    p.compare(FP64_TOL)


def test_fused_velocity_saturation_gravity_and_bie_threshold(oracle_built):
    chain = load_chain("panda_heavy_ee")
    p, q, dq, rng = _two_task(chain, 512, 27)
    p.set_mft_params(0, use_velocity_saturation=1, bie_threshold=0.25)
    p.set_jt_params(1, use_velocity_saturation=1, bie_threshold=0.25)
    p.both("enable_gravity_compensation", True)
    assert p.gpu.tick_is_fused()
    tg, tr = p.both("tick", True)
    assert _tau_err(tg, tr) <= FP64_TOL
    p.compare(FP64_TOL)


def test_fused_force_axis_closed_loop(oracle_built):
    """6-DoF task with one force axis in the compliant frame, closed loop + passivity (examples/07 style), 60 ticks"""
    chain = load_chain("panda")
    B = 256
    rng = np.random.default_rng(29)
    p = Pair(chain, B)
    t0 = p.add_mft("end-effector", (0, 0, 0.04), in_compliant_frame=True)
    t1 = p.add_jt()
    q = random_q(chain, B, rng, 0.3)
    dq = rng.uniform(-0.2, 0.2, (B, 7))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq)
    p.both("update_model"); p.both("reinitialize_tasks")
    p.set_mft_params(t0, force_space_dimension=1, force_or_motion_axis=[0, 0, 1.0], closed_loop_force_control=1, passivity_enabled=1)
    p.set(abi.F_GOAL_FORCE, np.array([0, 0, 5.0]), t0)
    assert p.gpu.tick_is_fused()
    worst = 0.0
    for tick in range(60):
        f = np.array([0, 0, 5.0]) * (1 + 0.3 * np.sin(0.02 * tick + np.arange(B)))[:, None] + rng.normal(0, 0.2, (B, 3))
        p.both("update_sensed_force_and_moment", t0, f, np.zeros((B, 3)))
        tg, tr = p.both("tick", True)
        worst = max(worst, _tau_err(tg, tr))
    assert worst <= FP64_TOL, worst
    p.compare(FP64_TOL)


@pytest.mark.parametrize("decoupling", [abi.FULL_DYNAMIC_DECOUPLING, abi.BOUNDED_INERTIA_ESTIMATES])
def test_fused_puma_six_dof(oracle_built, decoupling):
    """n = 6: the joint task has an empty range below a 6-DoF task (JointTask.cpp:234-239).  With bounded inertia estimates
    the PUMA clamps its three wrist joints: k_tick_fused2 serves only the last-joint clamp pattern, so every robot takes
    the API-faithful route; with full decoupling the certified ones stay in the fused kernel."""
    chain = load_chain("puma560_synthetic")
    p, q, dq, rng = _two_task(chain, 512, 31, p_c=(0, 0, 0))
    p.set_mft_params(0, dynamic_decoupling_type=decoupling)
    p.set_jt_params(1, dynamic_decoupling_type=decoupling)
    assert p.gpu.tick_is_fused()
    for _ in range(2):
        tg, tr = p.both("tick", True)
        assert _tau_err(tg, tr) <= FP64_TOL
    blend = int((p.ref.get(abi.F_ALPHA, 0)[:, 0] < 1).sum())
    fb = p.gpu.tick_fallback_count()
    if decoupling == abi.FULL_DYNAMIC_DECOUPLING:
        assert blend <= fb <= blend + 512 // 20
    else:
        assert blend <= fb <= 512
    # nullspace projectors: N = I - Jbar J is a cancellation (for this 6-DoF robot the exact N is 0 outside the blend: what the
    # getter returns is rounding noise in EVERY implementation, the oracle included).  Measured oracle-vs-LAPACK
    # (tools/scratch, r2): the noise is 0.2-0.6 eps kappa(M) outside the blend (kappa(M) ~ 5e4 for the PUMA's light wrist) and
    # grows with the conditioning of the inverted projected inertias inside it.  Per-robot bound:
    # max(1e-9, 64 eps kappa_task sqrt(kappa(M))), instead of the flat 1e-6 of round 1.
    kM = np.linalg.cond(p.ref.get(abi.F_M).reshape(512, 6, 6))
    nb = np.maximum(FP64_TOL, 64 * 2.2e-16 * task_kappa(p.ref, 0) * np.sqrt(kM))
    assert np.median(nb) <= 1e-6
    p.compare(FP64_TOL, loose={abi.F_N_TOTAL: nb, abi.F_N_TASK: nb, abi.F_N_TASK_AND_PREC: nb, abi.F_N_PREC: nb})


def test_outputs_modes_agree(oracle_built):
    """OSC_TICK_OUTPUTS_ALL (per-call kernel sequence), _LAZY and _TAU give the same torques and persistent state"""
    import osc_b200
    chain = load_chain("panda")
    B = 1024
    outs = []
    for mode in (abi.TICK_OUTPUTS_ALL, abi.TICK_OUTPUTS_LAZY, abi.TICK_OUTPUTS_TAU):
        p, q, dq, rng = _two_task(chain, B, 33)
        p.gpu.set_tick_outputs(mode)
        assert p.gpu.tick_is_fused() == (mode != abi.TICK_OUTPUTS_ALL)
        for _ in range(3):
            tau = p.gpu.tick(True)
        outs.append((tau, p.gpu.get(abi.F_INTEGRATED_POSITION_ERROR, 0), p.gpu.get(abi.F_JT_INTEGRATED_ERROR, 1)))
    for other in outs[1:]:
        for a, b in zip(outs[0], other):
            assert rel_err(a, b)[0] <= FP64_TOL


def test_fused_fp32(oracle_built):
    chain = load_chain("panda")
    p, q, dq, rng = _two_task(chain, 512, 35, abi.PRECISION_F32)
    tg, tr = p.both("tick", True)
    sv = p.ref.get(abi.F_SINGULAR_VALUES, 0)
    well = sv[:, 5] / sv[:, 0] > 0.09
    assert well.sum() > 30
    err = np.abs(tg - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-2)
    assert err[well].max() <= FP32_TOL, err[well].max()                 # north star: 1e-4 in FP32, on the MAX
    alpha = p.ref.get(abi.F_ALPHA, 0)[:, 0]
    assert err[alpha == 1].max() <= FP32_TOL, err[alpha == 1].max()      # every robot outside the blend, certified or not


def test_small_then_large_batch_in_one_process(oracle_built):
    """The fused tick switches to 224-thread blocks (100 KB of stash per block) once a batch fills the GPU; that kernel needs
    its own shared-memory attribute even when the 64-thread one already ran in this process (regression: the per-kernel flag
    was once shared between the two block shapes, and the large launch failed with "invalid argument")."""
    chain = load_chain("panda")
    for B, seed in ((512, 5), (224 * 148 + 11, 6), (512, 7)):
        p, q, dq, rng = _two_task(chain, B, seed)
        assert p.gpu.tick_is_fused()
        tg, tr = p.both("tick", True)
        assert _tau_err(tg, tr) <= FP64_TOL, B
        p.both("rollout_step", 1e-3, 1, 0.0)
        assert rel_err(p.gpu.get(abi.F_Q), p.ref.get(abi.F_Q))[0] <= FP64_TOL, B


def test_tilted_base_full_inertias_gravity_compensation(oracle_built):
    """Nothing in the fixtures rotates the base, tilts gravity or has products of inertia; the body-frame mass matrix of the fused
    tick, the world-frame one of the per-call route and the gravity vector must all agree with the oracle when they do."""
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
    p, q, dq, rng = _two_task(chain, 1024, 79)
    p.both("enable_gravity_compensation", True)
    assert p.gpu.tick_is_fused()
    tg, tr = p.both("tick", True)
    assert _tau_err(tg, tr) <= FP64_TOL
    assert np.abs(p.ref.get(abi.F_GRAVITY)).max() > 1.0
    p.compare(FP64_TOL)  # every getter, through the per-call kernels (lazy materialisation)


@pytest.mark.parametrize("precision", [abi.PRECISION_F64, abi.PRECISION_F32])
def test_first_generation_kernel_control_link_not_on_last_body(oracle_built, precision):
    """k_tick_fused2 needs the control link on the last body; a MotionForceTask on `link6` (body 5 of 7) makes osc_tick take
    the first-generation fused kernel (csrc/osc_fused.cuh k_tick_fused + k_tick_tail): the last joint does not move the
    control point, so J has a zero column and the joint task's range is wider than in config 2."""
    chain = load_chain("panda")
    B = 1024
    p, q, dq, rng = _two_task(chain, B, 37, precision, link="link6", p_c=(0.05, 0.0, 0.1), margin=0.3)
    assert p.gpu.tick_is_fused()
    for tick in range(3):
        p.set(abi.F_Q, q + 0.01 * tick)
        tg, tr = p.both("tick", True)
        if precision == abi.PRECISION_F64:
            assert _tau_err(tg, tr) <= FP64_TOL
    if precision == abi.PRECISION_F64:
        fb = p.gpu.tick_fallback_count()
        blend = int((p.ref.get(abi.F_ALPHA, 0)[:, 0] < 1).sum())
        assert blend <= fb < B
        p.compare(FP64_TOL)
    else:
        sv = p.ref.get(abi.F_SINGULAR_VALUES, 0)
        well = sv[:, 5] / sv[:, 0] > 0.09
        err = np.abs(tg - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-2)
        assert well.sum() > 30 and err[well].max() <= FP32_TOL, err[well].max()
