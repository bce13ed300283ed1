"""GPU parity of the device-side rollout step against the C++ oracle: b(q,dq) + g(q), one and many
symplectic-Euler steps, and the closed loop controller + dynamics that an MPC rollout runs without leaving the device."""
import copy

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


@pytest.mark.parametrize("name", ["panda", "panda_sliding_base"])
def test_step_tilted_base_and_full_inertias(oracle_built, name):
    """a base that is rotated and offset in the world, gravity off the world axes and inertia tensors with products of inertia:
    nothing the fixtures exercise, all of it enters b + g (the base acceleration that stands for gravity is expressed in base
    coordinates; the body-frame inertias about the joint origins pick up the off-diagonal terms)"""
    chain = copy.deepcopy(load_chain(name))
    cz, sz, cy, sy, cx, sx = np.cos(0.4), np.sin(0.4), np.cos(0.7), np.sin(0.7), np.cos(-0.3), np.sin(-0.3)
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]]); Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    chain["world_rot"] = (Rz @ Ry @ Rx).tolist()
    chain["world_xyz"] = [0.1, -0.2, 0.3]
    chain["gravity"] = [1.0, -2.0, -9.0]
    rng = np.random.default_rng(77)
    for b in chain["bodies"]:
        I = np.array(b["inertia"], float)
        A = rng.uniform(-0.3, 0.3, (3, 3))
        Q = np.linalg.qr(np.eye(3) + A)[0]
        b["inertia"] = (Q @ I @ Q.T).tolist()  # same principal moments, rotated axes: symmetric positive definite, full
    n, B = chain["n"], 256
    p = Pair(chain, B)
    q = random_q(chain, B, rng)
    dq = rng.uniform(-1, 1, (B, n))
    tau = rng.uniform(-5, 5, (B, n))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.set(abi.F_TAU, tau)
    p.both("rollout_step", 1e-3, 2, 0.1)
    h_ref = p.ref.get(abi.F_NONLINEAR_EFFECTS)
    assert np.abs(h_ref).max() > 1.0
    assert rel_err(p.gpu.get(abi.F_NONLINEAR_EFFECTS), h_ref)[0] <= FP64_TOL
    assert rel_err(p.gpu.get(abi.F_DQ), p.ref.get(abi.F_DQ))[0] <= FP64_TOL
    assert rel_err(p.gpu.get(abi.F_Q), p.ref.get(abi.F_Q))[0] <= FP64_TOL
