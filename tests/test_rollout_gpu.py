"""GPU parity of the device-side rollout step against the C++ oracle: b(q,dq) + g(q), one and many
symplectic-Euler steps, and the closed loop controller + dynamics that an MPC rollout runs without leaving the device."""
import numpy as np
import pytest

import osc_abi as abi
import workloads
from common import FP32_TOL, FP64_TOL, Pair, load_chain, random_q, rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["panda", "panda_sliding_base", "puma560_synthetic", "rrrrbot"])
def test_step_and_nonlinear_effects(oracle_built, name):
    chain = load_chain(name)
    n, B = chain["n"], 256
    rng = np.random.default_rng(7)
    p = Pair(chain, B)
    q = random_q(chain, B, rng)
    dq = rng.uniform(-1, 1, (B, n))
    tau = rng.uniform(-5, 5, (B, n))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.set(abi.F_TAU, tau)
    p.both("rollout_step", 1e-3, 1, 0.3)
    h_ref = p.ref.get(abi.F_NONLINEAR_EFFECTS)
    assert rel_err(p.gpu.get(abi.F_NONLINEAR_EFFECTS), h_ref)[0] <= FP64_TOL
    assert rel_err(p.gpu.get(abi.F_DQ), p.ref.get(abi.F_DQ))[0] <= FP64_TOL
    assert rel_err(p.gpu.get(abi.F_Q), p.ref.get(abi.F_Q))[0] <= FP64_TOL
    p.both("rollout_step", 2.5e-4, 40, 0.0)  # substeps with the torque held
    assert rel_err(p.gpu.get(abi.F_DQ), p.ref.get(abi.F_DQ))[0] <= 1e-8
    assert rel_err(p.gpu.get(abi.F_Q), p.ref.get(abi.F_Q))[0] <= 1e-8


def test_closed_loop_rollout(oracle_built):
    """300 ticks of [tick; rollout step] with gravity compensation on: the robots converge towards their goals, GPU and oracle
    stay together, and osc_rollout (one call, no host synchronisation in between) equals the explicit loop."""
    chain = load_chain("panda")
    B, ticks = 128, 300
    w = workloads.config2(chain, B, seed=9)
    p = Pair(chain, B)
    t0 = p.add_mft("end-effector", (0, 0, 0.07))
    t1 = p.add_jt()
    p.both("enable_gravity_compensation", True)
    p.set(abi.F_Q, w["q"]); p.set(abi.F_DQ, np.zeros_like(w["dq"]))
    p.both("update_model"); p.both("reinitialize_tasks")
    p.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); p.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0)
    p.set(abi.F_JT_GOAL_POSITION, w["q"], t1)
    snap = p.gpu.snapshot()
    e0 = None
    for t in range(ticks):
        p.both("tick", True)
        if t == 0:
            e0 = np.linalg.norm(p.ref.get(abi.F_CURRENT_POSITION, t0) - w["goal_pos"], axis=1)
        p.both("rollout_step", 1e-3, 1, 0.0)
        if t % 50 == 49 or t == ticks - 1:
            assert rel_err(p.gpu.get(abi.F_Q), p.ref.get(abi.F_Q))[0] <= 1e-8, t
            assert rel_err(p.gpu.get(abi.F_DQ), p.ref.get(abi.F_DQ), 1e-3)[0] <= 1e-6, t
    p.both("tick", True)
    e1 = np.linalg.norm(p.ref.get(abi.F_CURRENT_POSITION, t0) - w["goal_pos"], axis=1)
    assert np.median(e1) < 0.2 * np.median(e0), (np.median(e0), np.median(e1))
    q_loop = p.gpu.get(abi.F_Q).copy()
    # the same roll-out as ONE call
    p.gpu.restore(snap)
    p.gpu.set(abi.F_Q, w["q"]); p.gpu.set(abi.F_DQ, np.zeros_like(w["dq"]))
    p.gpu.rollout(ticks, 1e-3, 1, 0.0, True)
    assert np.array_equal(p.gpu.get(abi.F_Q), q_loop)


def test_step_large_batch_block_shape(oracle_built):
    """batches of 224 * 148 robots and more run k_rollout_step2 with 224-thread blocks (two per SM): same parity bar"""
    chain = load_chain("panda")
    n, B = chain["n"], 224 * 148 + 37
    rng = np.random.default_rng(21)
    p = Pair(chain, B)
    q = random_q(chain, B, rng)
    dq = rng.uniform(-1, 1, (B, n))
    tau = rng.uniform(-5, 5, (B, n))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.set(abi.F_TAU, tau)
    p.both("rollout_step", 1e-3, 3, 0.1)
    assert rel_err(p.gpu.get(abi.F_NONLINEAR_EFFECTS), p.ref.get(abi.F_NONLINEAR_EFFECTS))[0] <= FP64_TOL
    assert rel_err(p.gpu.get(abi.F_DQ), p.ref.get(abi.F_DQ))[0] <= FP64_TOL
    assert rel_err(p.gpu.get(abi.F_Q), p.ref.get(abi.F_Q))[0] <= FP64_TOL


@pytest.mark.parametrize("name", ["panda", "panda_sliding_base"])
def test_step_fp32(oracle_built, name):
    """FP32 batch (both kernels: body-frame for the Panda, general for the sliding base) against the FP64 oracle, 1e-4"""
    chain = load_chain(name)
    n, B = chain["n"], 512
    rng = np.random.default_rng(33)
    p = Pair(chain, B, abi.PRECISION_F32)
    q = random_q(chain, B, rng)
    dq = rng.uniform(-1, 1, (B, n))
    tau = rng.uniform(-5, 5, (B, n))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.set(abi.F_TAU, tau)
    p.both("rollout_step", 1e-3, 1, 0.2)
    assert rel_err(p.gpu.get(abi.F_NONLINEAR_EFFECTS), p.ref.get(abi.F_NONLINEAR_EFFECTS))[0] <= FP32_TOL
    assert rel_err(p.gpu.get(abi.F_DQ), p.ref.get(abi.F_DQ))[0] <= FP32_TOL
    assert rel_err(p.gpu.get(abi.F_Q), p.ref.get(abi.F_Q))[0] <= FP32_TOL
