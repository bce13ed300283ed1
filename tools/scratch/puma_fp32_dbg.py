import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in ("sai2-primitives-perso_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import osc_abi as abi
from common import Pair, load_chain, random_q
np.set_printoptions(linewidth=200, precision=4)
for prec in (abi.PRECISION_F64, abi.PRECISION_F32):
  for with_jt in (False, True):
    chain = load_chain("puma560_synthetic"); n = 6; B = 64
    rng = np.random.default_rng(5)
    p = Pair(chain, B, prec)
    t0 = p.add_mft("end-effector", (0, 0, 0))
    t1 = p.add_jt() if with_jt else None
    q = random_q(chain, B, rng, 0.1); dq = rng.uniform(-0.5, 0.5, (B, n))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.both("update_model"); p.both("reinitialize_tasks")
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    p.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    if with_jt: p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, (B, n)), t1)
    tg, tr = p.both("tick", with_jt)
    tg = np.asarray(tg, float)
    err = np.abs(tg - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-3 * np.abs(tr).max())
    print("prec", prec, "jt", with_jt, "fused", p.gpu.tick_is_fused(), "err p50 %.2e max %.2e" % (np.median(err), err.max()), "fallback", p.gpu.tick_fallback_count())
    w = int(err.argmax())
    print("  worst robot", w, "gpu", tg[w], "ref", tr[w])
    for t in range(2 if with_jt else 1):
        a, b = np.asarray(p.gpu.get(abi.F_TASK_TORQUES, t), float)[w], p.ref.get(abi.F_TASK_TORQUES, t)[w]
        print("   task", t, "gpu", a, "ref", b)
    if with_jt:
        print("   jt rank gpu", p.gpu.get(abi.F_JT_RANGE_RANK, t1)[:8].ravel(), "ref", p.ref.get(abi.F_JT_RANGE_RANK, t1)[:8].ravel())
    print("   status", p.gpu.get(abi.F_STATUS)[:8].ravel())
