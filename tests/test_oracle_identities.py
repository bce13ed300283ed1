"""CPU: physics identities that pin the C++ oracle's model functions (the un-vendored sai2-model part of the path,
SURVEY.md Appendix B) without any reference output: Jacobian vs finite differences of FK, mass matrix vs kinetic energy,
operational-space identities, closed-form planar 4R."""
import numpy as np
import pytest

import np_oracle as npo
import osc_abi as abi
from common import load_chain, q0_of, random_q


def _batch(orc, chain, q, dq=None, link="end-effector", p_c=(0, 0, 0.07), **kw):
    B = len(q)
    o = orc.OracleBatch(chain, B, q0=q0_of(chain))
    o.set(abi.F_Q, q)
    o.set(abi.F_DQ, np.zeros_like(q) if dq is None else dq)
    o.update_model()
    t0 = o.add_motion_force_task(abi.mft_desc(chain, link, p_c, **kw))
    return o, t0


@pytest.mark.parametrize("name,link,p_c", [("panda", "end-effector", (0, 0, 0.07)), ("puma560_synthetic", "end-effector", (0.07, 0, 0)),
                                           ("panda_sliding_base", "end-effector", (0, 0, 0.07)), ("rrrrbot", "link4", (0.5, 0, 0))])
def test_jacobian_is_the_derivative_of_fk(oracle_built, name, link, p_c):
    chain = load_chain(name)
    n = chain["n"]
    rng = np.random.default_rng(1)
    q = random_q(chain, 6, rng)
    o, t0 = _batch(oracle_built, chain, q, link=link, p_c=p_c)
    o.tick(False)
    J = o.get(abi.F_JACOBIAN, t0).reshape(-1, 6, n)
    h = 1e-6
    for j in range(n):
        e = np.zeros(n); e[j] = h
        xs, Rs = [], []
        for s in (+1, -1):
            o.set(abi.F_Q, q + s * e); o.tick(False)
            xs.append(o.get(abi.F_CURRENT_POSITION, t0)); Rs.append(o.get(abi.F_CURRENT_ORIENTATION, t0).reshape(-1, 3, 3))
        dv = (xs[0] - xs[1]) / (2 * h)
        dR = (Rs[0] - Rs[1]) / (2 * h)
        o.set(abi.F_Q, q); o.tick(False)
        R = o.get(abi.F_CURRENT_ORIENTATION, t0).reshape(-1, 3, 3)
        W = dR @ R.transpose(0, 2, 1)  # skew(omega)
        dw = np.stack([W[:, 2, 1], W[:, 0, 2], W[:, 1, 0]], axis=1)
        assert np.abs(J[:, :3, j] - dv).max() < 1e-8
        assert np.abs(J[:, 3:, j] - dw).max() < 1e-8


@pytest.mark.parametrize("name", ["panda", "panda_heavy_ee", "puma560_synthetic", "panda_sliding_base", "rrrrbot"])
def test_mass_matrix_is_the_kinetic_energy_hessian(oracle_built, name):
    chain = load_chain(name)
    n = chain["n"]
    rng = np.random.default_rng(2)
    q = random_q(chain, 8, rng)
    o = oracle_built.OracleBatch(chain, 8, q0=q0_of(chain))
    o.set(abi.F_Q, q); o.update_model()
    M = o.get(abi.F_M).reshape(-1, n, n)
    Minv = o.get(abi.F_MINV).reshape(-1, n, n)
    for r in range(8):
        m = npo.Model(chain); m.setQ(q[r]); m.updateModel()  # sum_b Jv^T m Jv + Jw^T I Jw, independent route
        assert np.abs(M[r] - m.M()).max() < 1e-12 * max(1.0, np.abs(M[r]).max())
        assert np.allclose(M[r], M[r].T, atol=0) or np.abs(M[r] - M[r].T).max() < 1e-14
        assert np.linalg.eigvalsh(M[r]).min() > 0
        assert np.abs(Minv[r] @ M[r] - np.eye(n)).max() < 1e-9


def test_planar_4r_closed_form(oracle_built):
    """rrrrbot (examples/11): planar chain, links 0.5 m, m = 1, com at 0.25: closed-form M and tip Jacobian"""
    chain = load_chain("rrrrbot")
    rng = np.random.default_rng(3)
    q = rng.uniform(-2, 2, (5, 4))
    o, t0 = _batch(oracle_built, chain, q, link="link4", p_c=(0.5, 0, 0))
    o.tick(False)
    J = o.get(abi.F_JACOBIAN, t0).reshape(-1, 6, 4)
    M = o.get(abi.F_M).reshape(-1, 4, 4)
    Lk, lc, mass = 0.5, 0.25, 1.0
    Izz = chain["bodies"][0]["inertia"][2][2]
    for r in range(5):
        th = np.cumsum(q[r])
        Jt = np.zeros((2, 4))
        for j in range(4):
            for k in range(j, 4):
                Jt[0, j] += -Lk * np.sin(th[k]); Jt[1, j] += Lk * np.cos(th[k])
        assert np.abs(J[r, :2] - Jt).max() < 1e-12 and np.abs(J[r, 5] - 1).max() < 1e-14 and np.abs(J[r, 2:5]).max() < 1e-14
        Mc = np.zeros((4, 4))
        for b in range(4):
            Jc = np.zeros((2, 4))
            for j in range(b + 1):
                for k in range(j, b + 1):
                    L_ = lc if k == b else Lk
                    Jc[0, j] += -L_ * np.sin(th[k]); Jc[1, j] += L_ * np.cos(th[k])
            w = np.zeros(4); w[:b + 1] = 1
            Mc += mass * Jc.T @ Jc + Izz * np.outer(w, w)
        assert np.abs(M[r] - Mc).max() < 1e-12


def test_operational_space_identities(oracle_built):
    """Lambda (J M^-1 J^T) = I, N^2 = N, J N = 0, N_total = N_2 N_1, tau_2 in the nullspace of the motion task"""
    chain = load_chain("panda")
    rng = np.random.default_rng(4)
    import workloads
    w = workloads.config2(chain, 12, 4)
    o, t0 = _batch(oracle_built, chain, w["q"], w["dq"])
    t1 = o.add_joint_task(abi.joint_task_desc(7))
    o.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); o.set(abi.F_JT_GOAL_POSITION, w["jt_goal"], t1)
    o.tick(True)
    J = o.get(abi.F_JACOBIAN, t0).reshape(-1, 6, 7)
    Minv = o.get(abi.F_MINV).reshape(-1, 7, 7)
    Lam = o.get(abi.F_LAMBDA_WORLD, t0).reshape(-1, 6, 6)
    N1 = o.get(abi.F_N_TASK, t0).reshape(-1, 7, 7)
    N2 = o.get(abi.F_N_TASK, t1).reshape(-1, 7, 7)
    Nt = o.get(abi.F_N_TOTAL).reshape(-1, 7, 7)
    tau2 = o.get(abi.F_TASK_TORQUES, t1)
    assert (o.get(abi.F_ALPHA, t0) == 1).all()
    for r in range(12):
        A = J[r] @ Minv[r] @ J[r].T
        assert np.abs(Lam[r] @ A - np.eye(6)).max() < 1e-8
        assert np.abs(N1[r] @ N1[r] - N1[r]).max() < 1e-9
        assert np.abs(J[r] @ N1[r]).max() < 1e-9
        assert np.abs(Nt[r] - N2[r] @ N1[r]).max() < 1e-12
        assert np.abs(J[r] @ Minv[r] @ tau2[r]).max() < 1e-8 * max(1.0, np.abs(tau2[r]).max())  # dynamically consistent


def test_svd_sign_flip_invariance():
    """tau and N do not depend on the sign convention of the SVD (hard part 2): numpy oracle with flipped U, V"""
    chain = load_chain("panda")
    rng = np.random.default_rng(5)
    q = random_q(chain, 1, rng)[0]
    taus = []
    orig = npo.svd_canonical
    try:
        for flip in (1.0, -1.0):
            def svd(A, flip=flip):
                U, s, V = orig(A)
                return U * flip, s, V * flip
            npo.svd_canonical = svd
            m = npo.Model(chain); m.setQ(q); m.setDq(rng.uniform(-0.3, 0.3, 7) * 0 + 0.1); m.updateModel()
            t0 = npo.MotionForceTask(m, "end-effector", np.array([0, 0, 0.07])); t1 = npo.JointTask(m)
            t0.goal_pos = t0.x + 0.05; t1.goal_pos = q + 0.1
            c = npo.RobotController(m, [t0, t1]); c.updateControllerTaskModels()
            taus.append(c.computeControlTorques())
    finally:
        npo.svd_canonical = orig
    assert np.abs(taus[0] - taus[1]).max() < 1e-9 * np.abs(taus[0]).max()
