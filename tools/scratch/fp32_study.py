"""FP32 batch against the FP64 oracle: torque error by conditioning class (run on the GPU box)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in ("sai2-primitives-perso_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import osc_abi as abi
from common import Pair, load_chain, random_q

def run(name, link, p_c, B=4096, seed=5):
    chain = load_chain(name); n = chain["n"]
    rng = np.random.default_rng(seed)
    p = Pair(chain, B, abi.PRECISION_F32)
    t0 = p.add_mft(link, p_c); t1 = p.add_jt()
    q = random_q(chain, B, rng, 0.1); dq = rng.uniform(-0.5, 0.5, (B, n))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.both("update_model"); p.both("reinitialize_tasks")
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    p.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_GOAL_LINEAR_VELOCITY, rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, (B, n)), t1)
    tg, tr = p.both("tick", True)
    tg = np.asarray(tg, float)
    sv = p.ref.get(abi.F_SINGULAR_VALUES, 0); alpha = p.ref.get(abi.F_ALPHA, 0)[:, 0]
    ratio = sv[:, 5] / sv[:, 0]
    err = np.abs(tg - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-3 * np.abs(tr).max())
    print(name, link, "fallback", p.gpu.tick_fallback_count(), "of", B)
    for lo, hi in ((0.09, 9), (0.075, 0.09), (0.06, 0.075), (0.0, 0.06)):
        m = (ratio >= lo) & (ratio < hi)
        if m.any():
            e = err[m]
            print("  ratio [%.3f, %.3f): n=%5d  p50 %.2e  p99 %.2e  max %.2e   (alpha<1: %d)" % (lo, hi, m.sum(), np.median(e), np.quantile(e, 0.99), e.max(), (alpha[m] < 1).sum()))

run("panda", "end-effector", (0, 0, 0.07))
run("panda", "link6", (0.05, 0, 0.1))
run("puma560_synthetic", "end-effector", (0, 0, 0))
