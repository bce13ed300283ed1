"""Shared helpers for the parity tests: build the same batch on the GPU (product, through the C ABI) and on the CPU
(oracle), drive both with the same seeded inputs, compare field by field."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import osc_abi as abi  # noqa: E402

FP64_TOL = 1e-9   # north_star: torques, Lambda, N, J within 1e-9 relative in FP64
FP32_TOL = 1e-4   # and 1e-4 in FP32 (away from singular configurations)


def load_chain(name):
    with open(os.path.join(ROOT, "fixtures", name + ".chain.json")) as f:
        return json.load(f)


def q0_of(chain):
    return np.array([j["q0"] for j in chain["joints"]], dtype=np.float64)


def limits(chain):
    return (np.array([j["lower"] for j in chain["joints"]]), np.array([j["upper"] for j in chain["joints"]]))


def random_q(chain, B, rng, margin=0.1):
    lo, hi = limits(chain)
    return rng.uniform(lo + margin, hi - margin, (B, chain["n"]))


def rel_err_each(a, b, floor=None):
    """per robot: ||a_r - b_r||_inf / max(||b_r||_inf, floor); floor defaults to 1e-3 x the batch-wide scale."""
    a = np.asarray(a, np.float64).reshape(len(a), -1)
    b = np.asarray(b, np.float64).reshape(len(b), -1)
    if floor is None:
        floor = 1e-3 * max(np.abs(b).max(), 1e-300)
        if floor == 0:
            floor = 1.0
    e = np.abs(a - b).max(axis=1) / np.maximum(np.abs(b).max(axis=1), floor)
    e[~np.isfinite(a).all(axis=1) | ~np.isfinite(b).all(axis=1)] = np.inf
    return e


def rel_err(a, b, floor=None):
    """max over robots of ||a_r - b_r||_inf / max(||b_r||_inf, floor); floor defaults to 1e-3 x the batch-wide scale."""
    a = np.asarray(a, np.float64).reshape(len(a), -1)
    b = np.asarray(b, np.float64).reshape(len(b), -1)
    if floor is None:
        floor = 1e-3 * max(np.abs(b).max(), 1e-300)
        if floor == 0:
            floor = 1.0
    num = np.abs(a - b).max(axis=1)
    den = np.maximum(np.abs(b).max(axis=1), floor)
    bad = ~np.isfinite(a).all(axis=1) | ~np.isfinite(b).all(axis=1)
    e = num / den
    e[bad] = np.inf
    return float(e.max()), int(e.argmax())


# fields compared after a tick: (field, owner kind, floor)
ROBOT_FIELDS = [(abi.F_TAU, None), (abi.F_M, None), (abi.F_MINV, None), (abi.F_N_TOTAL, 1.0)]
TASK_FIELDS = [(abi.F_TASK_TORQUES, None), (abi.F_N_TASK, 1.0), (abi.F_N_PREC, 1.0), (abi.F_N_TASK_AND_PREC, 1.0)]
MFT_FIELDS = [(abi.F_JACOBIAN, None), (abi.F_PROJECTED_JACOBIAN, 1e-3), (abi.F_LAMBDA_WORLD, None),
              (abi.F_LAMBDA_MODIFIED_WORLD, None), (abi.F_ALPHA, 1.0), (abi.F_SINGULAR_VALUES, None),
              (abi.F_CURRENT_POSITION, None), (abi.F_CURRENT_ORIENTATION, 1.0), (abi.F_CURRENT_LINEAR_VELOCITY, None),
              (abi.F_CURRENT_ANGULAR_VELOCITY, None), (abi.F_ORIENTATION_ERROR, 1e-3), (abi.F_UNIT_MASS_FORCE, None),
              (abi.F_INTEGRATED_POSITION_ERROR, 1e-6), (abi.F_INTEGRATED_ORIENTATION_ERROR, 1e-6),
              (abi.F_INTEGRATED_FORCE_ERROR, 1e-6), (abi.F_INTEGRATED_MOMENT_ERROR, 1e-6), (abi.F_POPC_RC, 1.0),
              (abi.F_POPC_OBSERVER, 1e-6), (abi.F_Q_PRIOR, 1.0), (abi.F_SENSED_FORCE_WORLD, 1e-3),
              (abi.F_SENSED_MOMENT_WORLD, 1e-3), (abi.F_DESIRED_LINEAR_VELOCITY, 1e-3), (abi.F_DESIRED_ANGULAR_VELOCITY, 1e-3)]
MFT_INT_FIELDS = [abi.F_NS_RANK, abi.F_SINGULARITY_TYPES, abi.F_SINGULARITY_COUNTERS]
JT_FIELDS = [(abi.F_JT_CURRENT_POSITION, None), (abi.F_JT_CURRENT_VELOCITY, None), (abi.F_JT_INTEGRATED_ERROR, 1e-6),
             (abi.F_JT_LAMBDA_WORLD, None), (abi.F_JT_LAMBDA_MODIFIED_WORLD, None), (abi.F_JT_DESIRED_VELOCITY, 1e-3)]
JT_INT_FIELDS = [abi.F_JT_RANGE_RANK]


def compare(gpu, orc, task_types, tol, skip=(), report=None, loose=None):
    """Compare every exported field of both batches; returns {name: err}; asserts err <= tol."""
    names = {v: k for k, v in vars(abi).items() if k.startswith("F_")}
    errs = {}

    def one(field, task, floor):
        if field in skip:
            return
        a, b = gpu.get(field, task), orc.get(field, task)
        each = rel_err_each(a, b, floor)
        bound = np.broadcast_to(np.asarray((loose or {}).get(field, tol), float), each.shape)  # scalar or per-robot bounds
        where = int((each / bound).argmax())
        e = float(each.max())
        errs["%s[t%d]" % (names[field], task)] = e
        assert (each <= bound).all(), "%s task %d: rel err %.3e > %.1e at robot %d\n gpu=%s\n ref=%s" % (
            names[field], task, each[where], bound[where], where, a[where], b[where])

    def one_int(field, task):
        if field in skip:
            return
        a, b = gpu.get(field, task), orc.get(field, task)
        assert np.array_equal(np.asarray(a, np.int64), np.asarray(np.rint(b), np.int64)), "%s task %d differs: %s vs %s" % (
            names[field], task, a[:4], b[:4])

    for f, fl in ROBOT_FIELDS:
        one(f, -1, fl)
    for t, ty in enumerate(task_types):
        for f, fl in TASK_FIELDS:
            one(f, t, fl)
        if ty == abi.TASK_MOTION_FORCE:
            for f, fl in MFT_FIELDS:
                one(f, t, fl)
            for f in MFT_INT_FIELDS:
                one_int(f, t)
        else:
            for f, fl in JT_FIELDS:
                one(f, t, fl)
            for f in JT_INT_FIELDS:
                one_int(f, t)
    if report is not None:
        report.update(errs)
    return errs


class Pair:
    """The same batch twice: `gpu` (osc_b200.Batch) and `ref` (orc.OracleBatch).  Calls are mirrored to both."""

    def __init__(self, chain, B, precision=abi.PRECISION_F64, q0=None):
        import orc
        import osc_b200
        self.chain, self.B, self.n = chain, B, chain["n"]
        q0 = q0_of(chain) if q0 is None else q0
        self.gpu = osc_b200.Batch(chain, B, precision=precision, q0=q0)
        self.ref = orc.OracleBatch(chain, B, q0=q0)
        self.types = []

    def both(self, name, *a, **k):
        r1 = getattr(self.gpu, name)(*a, **k)
        r2 = getattr(self.ref, name)(*a, **k)
        return r1, r2

    def add_mft(self, *a, **k):
        d = abi.mft_desc(self.chain, *a, **k)
        self.types.append(abi.TASK_MOTION_FORCE)
        t1, t2 = self.both("add_motion_force_task", d)
        assert t1 == t2
        return t1

    def add_jt(self, selection=None, dt=0.001):
        d = abi.joint_task_desc(self.n, selection, dt)
        self.types.append(abi.TASK_JOINT)
        t1, t2 = self.both("add_joint_task", d)
        assert t1 == t2
        return t1

    def set(self, field, data, task=-1):
        self.both("set", field, data, task)

    def set_mft_params(self, task, **kw):
        out = []
        for b in (self.gpu, self.ref):
            p = b.mft_params(task)
            for k, v in kw.items():
                cur = getattr(p, k)
                if hasattr(cur, "__len__"):
                    for i, x in enumerate(np.broadcast_to(np.asarray(v, float), (len(cur),))):
                        cur[i] = float(x)
                else:
                    setattr(p, k, v)
            out.append(b.set_mft_params(task, p))
        assert out[0] == out[1], "reset masks differ: %s" % out
        return out[0]

    def set_jt_params(self, task, **kw):
        for b in (self.gpu, self.ref):
            p = b.joint_task_params(task)
            for k, v in kw.items():
                cur = getattr(p, k)
                if hasattr(cur, "__len__"):
                    for i, x in enumerate(np.broadcast_to(np.asarray(v, float), (len(cur),))):
                        cur[i] = float(x)
                else:
                    setattr(p, k, v)
            b.set_joint_task_params(task, p)

    def compare(self, tol, skip=(), report=None, loose=None):
        return compare(self.gpu, self.ref, self.types, tol, skip, report, loose)


def task_kappa(ref, t0):
    """Per robot: the largest condition number of the projected inertias the SingularityHandler inverts this tick --
    J_ns M^-1 J_ns^T and, inside the blend branch, J_s M^-1 J_s^T (SingularityHandler.cpp:110-131) -- rebuilt with LAPACK
    from the oracle's getters (projected Jacobian, M^-1, rank of the non-singular part).  Both the oracle and the product
    are accurate to c * eps * kappa there and no better, whatever the route; tests scale their tolerance with it."""
    Jp = ref.get(abi.F_PROJECTED_JACOBIAN, t0)
    Minv = ref.get(abi.F_MINV)
    n = int(round(np.sqrt(Minv.shape[1])))
    Jp, Minv = Jp.reshape(len(Jp), 6, n), Minv.reshape(len(Minv), n, n)
    ns = np.asarray(ref.get(abi.F_NS_RANK, t0)).reshape(-1).astype(int)
    m = min(6, n)
    U, sv, _ = np.linalg.svd(Jp, full_matrices=False)
    out = np.ones(len(Jp))
    for r in range(len(Jp)):
        rank = int((sv[r] > 1e-12 * sv[r, 0]).sum()) if sv[r, 0] > 0 else 0
        top = min(m, rank)
        for lo, hi in ((0, ns[r]), (ns[r], top)):
            if hi > lo:
                Js = U[r][:, lo:hi].T @ Jp[r]
                out[r] = max(out[r], np.linalg.cond(Js @ Minv[r] @ Js.T))
    return out


def nullspace_noise_scale(ref, t0):
    """Per robot: 1 + |Jbar|_F |Jp|_F with Jbar = M^-1 Jp^T Lambda_ns (world form).  N = I - Jbar Jp is a cancellation: every
    implementation, the oracle included, carries absolute rounding noise c eps |Jbar| |Jp| in it, whatever the route."""
    Jp = ref.get(abi.F_PROJECTED_JACOBIAN, t0)
    Minv = ref.get(abi.F_MINV)
    n = int(round(np.sqrt(Minv.shape[1])))
    Jp, Minv = Jp.reshape(len(Jp), 6, n), Minv.reshape(len(Minv), n, n)
    Lam = ref.get(abi.F_LAMBDA_WORLD, t0).reshape(len(Jp), 6, 6)
    Jbar = Minv @ np.transpose(Jp, (0, 2, 1)) @ Lam
    return 1.0 + np.linalg.norm(Jbar, axis=(1, 2)) * np.linalg.norm(Jp, axis=(1, 2))
