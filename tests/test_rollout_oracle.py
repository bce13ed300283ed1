"""CPU checks of the rollout step of the oracle (restatement only, parity unpinned: sai2-model /
Sai2Simulation are un-vendored): the recursive Newton-Euler b(q,dq) + g(q) against two independent routes, and the
integrator against physics identities."""
import copy

import numpy as np
import pytest

import osc_abi as abi
from common import load_chain, q0_of, random_q


def _batch(orc, chain, q, dq):
    o = orc.OracleBatch(chain, len(q), q0=q0_of(chain))
    o.set(abi.F_Q, q)
    o.set(abi.F_DQ, dq)
    o.update_model()
    return o


def _mass(orc, chain, q):
    o = orc.OracleBatch(chain, len(q), q0=q0_of(chain))
    o.set(abi.F_Q, q)
    o.update_model()
    n = chain["n"]
    return o.get(abi.F_M).reshape(len(q), n, n)


@pytest.mark.parametrize("name", ["panda", "panda_sliding_base", "puma560_synthetic", "rrrrbot"])
def test_coriolis_from_mass_matrix_derivatives(oracle_built, name):
    """b_i = sum_k (dM/dq_k dq)_i dq_k - 1/2 dq^T dM/dq_i dq  (Lagrange), dM/dq by central differences of the CRBA matrix"""
    orc = oracle_built
    chain = load_chain(name)
    n, B = chain["n"], 12
    rng = np.random.default_rng(11)
    q = random_q(chain, B, rng)
    dq = rng.uniform(-1, 1, (B, n))
    o = _batch(orc, chain, q, dq)
    h = o.nonlinear_effects()
    o0 = _batch(orc, chain, q, np.zeros((B, n)))
    g = o0.nonlinear_effects()
    eps = 1e-5
    dM = np.zeros((n, B, n, n))
    for k in range(n):
        e = np.zeros(n); e[k] = eps
        dM[k] = (_mass(orc, chain, q + e) - _mass(orc, chain, q - e)) / (2 * eps)
    b = np.zeros((B, n))
    for r in range(B):
        Mdot = sum(dM[k][r] * dq[r, k] for k in range(n))
        b[r] = Mdot @ dq[r] - 0.5 * np.array([dq[r] @ dM[i][r] @ dq[r] for i in range(n)])
    scale = np.abs(h - g).max()
    assert scale > 1e-2
    assert np.abs((h - g) - b).max() < 2e-6 * max(scale, 1.0)


@pytest.mark.parametrize("name", ["panda", "panda_sliding_base", "puma560_synthetic"])
def test_gravity_part_equals_joint_gravity_vector(oracle_built, name):
    """dq = 0: the Newton-Euler pass must return jointGravityVector (Jacobian-transpose sum, RobotController.cpp:71)"""
    orc = oracle_built
    chain = load_chain(name)
    n, B = chain["n"], 16
    rng = np.random.default_rng(5)
    q = random_q(chain, B, rng)
    o = _batch(orc, chain, q, np.zeros((B, n)))
    o.add_joint_task(abi.joint_task_desc(n))
    o.enable_gravity_compensation(True)
    o.update_task_models()
    o.compute_torques(True)
    g_ref = o.get(abi.F_GRAVITY)
    assert np.abs(g_ref).max() > 1.0
    assert np.abs(o.nonlinear_effects() - g_ref).max() < 1e-11 * np.abs(g_ref).max()


def test_holding_torque_and_energy(oracle_built):
    orc = oracle_built
    chain = load_chain("panda")
    n, B = chain["n"], 8
    rng = np.random.default_rng(3)
    q = random_q(chain, B, rng)
    dq = rng.uniform(-0.5, 0.5, (B, n))
    # tau = b + g  =>  qdd = 0: the velocity does not change over a step
    o = _batch(orc, chain, q, dq)
    o.set(abi.F_TAU, o.nonlinear_effects())
    o.rollout_step(1e-3, 1, 0.0)
    assert np.abs(o.get(abi.F_DQ) - dq).max() < 1e-12
    assert np.abs(o.get(abi.F_Q) - (q + dq * 1e-3)).max() < 1e-14
    # no gravity, no torque: kinetic energy is conserved; the symplectic Euler drift is first order in dt
    free = copy.deepcopy(chain)
    free["gravity"] = [0.0, 0.0, 0.0]

    def drift(dt, steps):
        b = _batch(orc, free, q, dq)
        b.set(abi.F_TAU, np.zeros((B, n)))
        M0 = b.get(abi.F_M).reshape(B, n, n)
        T0 = 0.5 * np.einsum("bi,bij,bj->b", dq, M0, dq)
        b.rollout_step(dt, steps, 0.0)
        b.update_model()
        v = b.get(abi.F_DQ)
        T1 = 0.5 * np.einsum("bi,bij,bj->b", v, b.get(abi.F_M).reshape(B, n, n), v)
        return np.abs(T1 - T0) / T0

    d1, d2 = drift(2e-4, 250), drift(1e-4, 500)
    assert d1.max() < 2e-2, d1
    ratio = d1 / np.maximum(d2, 1e-300)
    assert (ratio > 1.5).all() and (ratio < 2.6).all(), ratio
    # damping removes energy
    b = _batch(orc, free, q, dq)
    b.set(abi.F_TAU, np.zeros((B, n)))
    b.rollout_step(1e-3, 200, 2.0)
    assert (np.abs(b.get(abi.F_DQ)).max(axis=1) < np.abs(dq).max(axis=1)).all()


def test_tilted_base_full_inertias(oracle_built):
    """b from the derivatives of the mass matrix and g from the gravity vector, on a chain with a rotated base, off-axis gravity
    and products of inertia (the chain of tests/test_rollout_gpu.py::test_step_tilted_base_and_full_inertias)"""
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
    n, B = chain["n"], 8
    q = random_q(chain, B, rng)
    dq = rng.uniform(-1, 1, (B, n))
    h = _batch(orc, chain, q, dq).nonlinear_effects()
    o0 = _batch(orc, chain, q, np.zeros((B, n)))
    g = o0.nonlinear_effects()
    o0.add_joint_task(abi.joint_task_desc(n))
    o0.enable_gravity_compensation(True)
    o0.update_task_models()
    o0.compute_torques(True)
    g_ref = o0.get(abi.F_GRAVITY)
    assert np.abs(g_ref).max() > 1.0
    assert np.abs(g - g_ref).max() < 1e-11 * np.abs(g_ref).max()
    eps = 1e-5
    dM = np.zeros((n, B, n, n))
    for k in range(n):
        e = np.zeros(n); e[k] = eps
        dM[k] = (_mass(orc, chain, q + e) - _mass(orc, chain, q - e)) / (2 * eps)
    b = np.zeros((B, n))
    for r in range(B):
        Mdot = sum(dM[k][r] * dq[r, k] for k in range(n))
        b[r] = Mdot @ dq[r] - 0.5 * np.array([dq[r] @ dM[i][r] @ dq[r] for i in range(n)])
    scale = np.abs(h - g).max()
    assert scale > 1e-2
    assert np.abs((h - g) - b).max() < 2e-6 * max(scale, 1.0)
