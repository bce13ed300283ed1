"""BASELINE config 4 on the GPU: robots ON and NEAR kinematic singularities, 300 consecutive ticks so that the 200-deep
classification history of the SingularityHandler matters (SingularityHandler.cpp:76-143, 233-305).

(a) Panda, 6-DoF MotionForceTask + JointTask, velocity saturation on (examples/18:130-131); q on the elbow-extended set
    (q4 at its upper limit -0.0698) and the wrist-aligned set (q6 = 0), offset by eps in {1e-1, 1e-2, 1e-3, 1e-4};
(b) 6R PUMA-geometry chain at the wrist-lock posture of examples/19:76-80 +- eps, setSingularityHandlingGains(50, 20, 20).

FP64 is compared with the C++ oracle tick by tick; FP32 is run beside it and its error distribution is reported per
eps (the 1e-4 FP32 bound of north_star is asserted away from the singular set only: inside it the blend weights are
discontinuous functions of the singular values and the two precisions may legitimately classify differently)."""
import numpy as np
import pytest

import osc_abi as abi
from common import FP32_TOL, FP64_TOL, Pair, load_chain, q0_of, task_kappa

pytestmark = pytest.mark.gpu
EPS = (5e-1, 1e-1, 1e-2, 1e-3, 1e-4)  # BASELINE config 4 names the last four; 0.5 is the control group outside the singular set
PER = 48  # robots per eps


def _tau_err(tg, tr):
    tg, tr = np.asarray(tg, float), np.asarray(tr, float)
    return np.abs(tg - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-3 * max(np.abs(tr).max(), 1e-300))


def _panda_q(rng):
    q0 = q0_of(load_chain("panda"))
    q = np.tile(q0, (len(EPS) * PER, 1)) + rng.normal(0, 0.2, (len(EPS) * PER, 7)) * np.array([1, 1, 1, 0, 1, 0, 1.0])
    for k, eps in enumerate(EPS):
        s = slice(k * PER, (k + 1) * PER)
        q[s, 3] = -0.0698 - np.abs(rng.normal(0, eps, PER))   # elbow extended (joint 4 at its upper limit)
        q[s, 5] = rng.normal(0, eps, PER)                     # wrist aligned (joint 6 about zero)
    return q


def _puma_q(rng):
    base = np.array([0.0, -np.pi / 4, 3 * np.pi / 4, np.pi / 2, 0.0, np.pi])  # examples/19:76-80
    q = np.tile(base, (len(EPS) * PER, 1))
    for k, eps in enumerate(EPS):
        q[k * PER:(k + 1) * PER] += rng.normal(0, eps, (PER, 6))
    return q


def _run(chain_name, q, precision, ticks, vel_sat, gains, rng_seed):
    chain = load_chain(chain_name)
    n, B = chain["n"], len(q)
    rng = np.random.default_rng(rng_seed)
    p = Pair(chain, B, precision, q0=q0_of(chain))
    t0 = p.add_mft("end-effector", (0, 0, 0.0))
    t1 = p.add_jt() if n > 6 else None
    p.set(abi.F_Q, q); p.set(abi.F_DQ, np.zeros((B, n)))
    p.both("update_model"); p.both("reinitialize_tasks")
    kw = {}
    if vel_sat:
        kw.update(use_velocity_saturation=1)
    if gains:
        kw.update(kp_type_1=gains[0], kv_type_1=gains[1], kv_type_2=gains[2])
    if kw:
        p.set_mft_params(t0, **kw)
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    p.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    amp = rng.normal(0, 0.02, (B, n))
    errs, alphas, kappas = [], [], []
    for tick in range(ticks):
        qt = q + amp * np.sin(2 * np.pi * 0.5 * tick * 0.001)  # sweeps in and out of the singular region
        dqt = amp * 2 * np.pi * 0.5 * np.cos(2 * np.pi * 0.5 * tick * 0.001)
        p.set(abi.F_Q, qt); p.set(abi.F_DQ, dqt)
        tg, tr = p.both("tick", t1 is not None)
        errs.append(_tau_err(tg, tr))
        alphas.append(p.ref.get(abi.F_ALPHA, t0)[:, 0].copy())
        kappas.append(task_kappa(p.ref, t0))
    return p, t0, np.array(errs), np.array(alphas), np.array(kappas)


@pytest.mark.parametrize("case", ["panda", "puma"])
def test_singular_fp64_parity_300_ticks(oracle_built, case):
    rng = np.random.default_rng(4)
    if case == "panda":
        p, t0, errs, alphas, kappas = _run("panda", _panda_q(rng), abi.PRECISION_F64, 300, True, None, 41)
    else:
        p, t0, errs, alphas, kappas = _run("puma560_synthetic", _puma_q(rng), abi.PRECISION_F64, 300, False, (50.0, 20.0, 20.0), 42)
    assert ((alphas > 0) & (alphas < 1)).any(), "no robot-tick inside the blend band (0 < alpha < 1)"
    assert (alphas == 0).any(), "no robot-tick with alpha = 0 (singular directions fully handed to the joint strategy)"
    # discrete state must agree exactly: branch (ns rank), singularity types and counters after the 300-tick history
    for fld in (abi.F_NS_RANK, abi.F_SINGULARITY_TYPES, abi.F_SINGULARITY_COUNTERS):
        assert np.array_equal(np.asarray(p.gpu.get(fld, t0), np.int64), np.asarray(np.rint(p.ref.get(fld, t0)), np.int64)), fld
    # torques: 1e-9 wherever the conditioning allows it.  Inside the blend the handler inverts J_s M^-1 J_s^T at condition
    # numbers up to 1e7 (SingularityHandler.cpp:117-120); the oracle and the product are then both accurate to c eps kappa
    # and no better, so the bound is per robot-tick: max(1e-9, 64 eps kappa), kappa rebuilt with LAPACK from the oracle.
    clean = alphas == 1
    if clean.any():  # (every robot of the Panda set sits in the blend: its q4 is on the elbow-extended limit)
        assert errs[clean].max() <= FP64_TOL, errs[clean].max()
    bound = np.maximum(FP64_TOL, 64 * 2.2e-16 * kappas)
    worst = (errs / bound).max()
    assert worst <= 1.0, "torque error above max(1e-9, 64 eps kappa): ratio %.2f (err %.2e, kappa %.2e)" % (
        worst, errs.ravel()[(errs / bound).argmax()], kappas.ravel()[(errs / bound).argmax()])
    assert (errs[kappas < 1e5] <= FP64_TOL).all(), "1e-9 must hold up to kappa = 1e5"
    assert np.abs(p.gpu.get(abi.F_ALPHA, t0) - p.ref.get(abi.F_ALPHA, t0)).max() <= 1e-9
    assert np.abs(p.gpu.get(abi.F_Q_PRIOR, t0) - p.ref.get(abi.F_Q_PRIOR, t0)).max() <= 1e-12


@pytest.mark.parametrize("case", ["panda", "puma"])
def test_singular_fp32_error_study(oracle_built, case, capsys):
    rng = np.random.default_rng(4)
    if case == "panda":
        p, t0, errs, alphas, kappas = _run("panda", _panda_q(rng), abi.PRECISION_F32, 300, True, None, 41)
    else:
        p, t0, errs, alphas, kappas = _run("puma560_synthetic", _puma_q(rng), abi.PRECISION_F32, 300, False, (50.0, 20.0, 20.0), 42)
    rows = []
    for k, eps in enumerate(EPS):
        e = errs[:, k * PER:(k + 1) * PER].ravel()
        e = e[np.isfinite(e)]
        rows.append((eps, np.percentile(e, 50), np.percentile(e, 90), np.percentile(e, 99), e.max(),
                     float((alphas[:, k * PER:(k + 1) * PER] < 1).mean())))
    with capsys.disabled():
        print("\nFP32 torque error vs the FP64 oracle, %s, 300 ticks x %d robots per eps" % (case, PER))
        print("  eps      p50       p90       p99       max       share of robot-ticks in the blend branch")
        for r in rows:
            print("  %-7g  %.2e  %.2e  %.2e  %.2e  %.2f" % r)
    assert np.isfinite(errs).all()
    # north star: 1e-4 in FP32.  An FP32 batch stores q, dq, goals in FP32 (relative input rounding 6e-8) and computes its
    # factorisation chain / task models in FP64 (csrc/osc_fused2.cuh, csrc/osc_kernels.cuh Row<T, S>), so what is left is the
    # amplification of the INPUT rounding by the conditioning of the inverted projected inertias: asserted as
    # max(1e-4, 64 eps_f32 kappa) per robot-tick, i.e. a flat 1e-4 for every robot-tick with kappa <= 26.
    clean = alphas == 1
    if clean.any():
        assert errs[clean].max() <= FP32_TOL, errs[clean].max()
    bound = np.maximum(FP32_TOL, 64 * 6e-8 * kappas)
    flips = (errs > bound).mean()
    with capsys.disabled():
        print("  max over all robot-ticks %.2e (outside the blend: %s); share above max(1e-4, 64 eps32 kappa): %.4f; share with kappa <= 26: %.2f" % (
            errs.max(), "%.2e" % errs[clean].max() if clean.any() else "none", flips, (kappas <= 26).mean()))
    # a robot-tick may still differ when the FP32-rounded q puts it on the other side of a discrete decision (type-1 vs type-2
    # vote, joint-limit direction latch): rare by construction
    assert flips <= 0.01, flips
