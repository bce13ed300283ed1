"""Replay of the frozen golden vectors (tests/golden/*.npz, made by tools/make_golden.py from the numpy/LAPACK oracle):
on the C++ oracle (CPU, not gpu) and on the CUDA path through the C ABI (gpu).  FP64, 1e-9 relative."""
import os

import numpy as np
import pytest

import osc_abi as abi
from common import FP64_TOL, ROOT, load_chain, q0_of, rel_err, rel_err_each, task_kappa


def G(name):
    return dict(np.load(os.path.join(ROOT, "tests", "golden", name + ".npz")))


def make_batch(kind, chain, B):
    if kind == "oracle":
        import orc
        orc.build()
        return orc.OracleBatch(chain, B, q0=q0_of(chain))
    import osc_b200
    return osc_b200.Batch(chain, B, q0=q0_of(chain))


KINDS = ["oracle", pytest.param("cuda", marks=pytest.mark.gpu)]


def close(a, b, tol=FP64_TOL, floor=None):
    e = rel_err(np.asarray(a, float).reshape(len(b), -1), np.asarray(b, float).reshape(len(b), -1), floor)[0]
    assert e <= tol, e


@pytest.mark.parametrize("kind", KINDS)
def test_golden_panda_two_task(kind):
    g = G("panda_two_task")
    chain = load_chain("panda")
    B = len(g["q"])
    b = make_batch(kind, chain, B)
    b.set(abi.F_Q, g["q"]); b.set(abi.F_DQ, g["dq"]); b.update_model()
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", tuple(g["p_c"])))
    t1 = b.add_joint_task(abi.joint_task_desc(7))
    x0 = b.get(abi.F_CURRENT_POSITION, t0)
    b.set(abi.F_GOAL_POSITION, x0 + g["dpos"], t0); b.set(abi.F_GOAL_LINEAR_VELOCITY, g["gv"], t0)
    b.set(abi.F_GOAL_ANGULAR_VELOCITY, g["gw"], t0); b.set(abi.F_JT_GOAL_POSITION, g["gq"], t1)
    for t in range(len(g["dqt"])):
        b.set(abi.F_Q, g["q"] + g["dqt"][t])
        tau = b.tick(True)
    close(tau, g["tau"])
    close(b.get(abi.F_M), g["M"])
    close(b.get(abi.F_JACOBIAN, t0), g["J"])
    close(b.get(abi.F_LAMBDA_WORLD, t0), g["lam"])
    close(b.get(abi.F_N_TOTAL), g["Ntot"], floor=1.0)
    close(b.get(abi.F_ALPHA, t0), g["alpha"], floor=1.0)
    close(b.get(abi.F_CURRENT_POSITION, t0), g["x"])


@pytest.mark.parametrize("kind", KINDS)
def test_golden_panda_force_popc(kind):
    g = G("panda_force_popc")
    chain = load_chain("panda")
    B = len(g["q"])
    b = make_batch(kind, chain, B)
    b.set(abi.F_Q, g["q"]); b.set(abi.F_DQ, g["dq"]); b.update_model()
    dirs = [(1, 0, 0), (0, 1, 0), (0, 0, 1)]
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0), dirs_translation=dirs, dirs_rotation=[]))
    b.add_joint_task(abi.joint_task_desc(7))
    p = b.mft_params(t0)
    p.force_space_dimension = 1
    p.force_or_motion_axis[2] = 1.0
    p.closed_loop_force_control = 1
    p.passivity_enabled = 1
    b.set_mft_params(t0, p)
    b.set(abi.F_GOAL_FORCE, np.array([0, 0, -5.0]), t0)
    keep = {}
    for t in range(len(g["f"])):
        b.set(abi.F_Q, g["q"] + 0.02 * np.sin(0.01 * t))
        b.update_model(); b.update_task_models()
        b.update_sensed_force_and_moment(t0, g["f"][t], np.zeros((B, 3)))
        tau = b.compute_torques(True)
        if t in g["tick_ids"]:
            keep[t] = np.array(tau, float)
    for i, t in enumerate(g["tick_ids"]):
        close(keep[int(t)], g["tau"][i])
    close(b.get(abi.F_POPC_RC, t0), g["rc"][:, None], floor=1.0)
    close(b.get(abi.F_INTEGRATED_FORCE_ERROR, t0), g["integ"])


@pytest.mark.parametrize("kind", KINDS)
def test_golden_puma_singular(kind):
    g = G("puma_singular")
    chain = load_chain("puma560_synthetic")
    B = len(g["q"])
    b = make_batch(kind, chain, B)
    b.set(abi.F_Q, g["q"]); b.set(abi.F_DQ, g["dq"]); b.update_model()
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0)))
    p = b.mft_params(t0)
    p.kp_type_1, p.kv_type_1, p.kv_type_2 = 50.0, 20.0, 20.0
    b.set_mft_params(t0, p)
    b.set(abi.F_GOAL_POSITION, b.get(abi.F_CURRENT_POSITION, t0) + g["off"], t0)
    for t in range(int(g["ticks"])):
        b.set(abi.F_Q, g["q"] + 1e-3 * t)
        tau = b.tick(False)
    # blend-branch robots invert J_s M^-1 J_s^T at condition numbers up to 1e7: per-robot bound max(1e-9, 64 eps kappa)
    each = rel_err_each(np.asarray(tau, float), g["tau"], 1e-3)
    bound = np.maximum(FP64_TOL, 64 * 2.2e-16 * task_kappa(b, t0))
    assert (each <= bound).all(), (each / bound).max()
    assert bound.max() <= 1e-8
    close(b.get(abi.F_ALPHA, t0), g["alpha"][:, None], floor=1.0)
    assert (g["alpha"] < 1).any()


@pytest.mark.parametrize("kind", KINDS)
def test_golden_planar_partial(kind):
    g = G("planar_partial")
    chain = load_chain("rrrrbot")
    B = len(g["q"])
    b = make_batch(kind, chain, B)
    b.set(abi.F_Q, g["q"]); b.set(abi.F_DQ, g["dq"]); b.update_model()
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "link4", (0.5, 0, 0), dirs_translation=[(1, 0, 0), (0, 1, 0)], dirs_rotation=[(0, 0, 1)]))
    b.add_joint_task(abi.joint_task_desc(4))
    b.set(abi.F_GOAL_POSITION, b.get(abi.F_CURRENT_POSITION, t0) + g["off"], t0)
    tau = b.tick(True)
    close(tau, g["tau"])
    close(b.get(abi.F_M), g["M"])
