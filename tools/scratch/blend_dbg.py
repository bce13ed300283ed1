import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import osc_abi as abi
import orc
orc.build()
import test_fused3_gpu as T3
import test_singular_gpu as TS
from common import load_chain, q0_of, Pair

def report(p, t0, t1, stacked, tag):
    tg, tr = p.both("tick", stacked)
    tg = np.asarray(tg); tr = np.asarray(tr)
    err = np.abs(tg - tr).max(1) / np.maximum(np.abs(tr).max(1), 1e-3 * np.abs(tr).max())
    alpha = p.ref.get(abi.F_ALPHA, t0)[:, 0]
    bad = np.argsort(-err)[:8]
    print(tag, "stacked", stacked, "max err %.3e" % err.max(), "n(err>1e-9) =", int((err > 1e-9).sum()), "blend robots", int((alpha < 1).sum()), "fallback", p.gpu.tick_fallback_count())
    t1g = np.asarray(p.gpu.get(abi.F_TASK_TORQUES, t0)); t1r = p.ref.get(abi.F_TASK_TORQUES, t0)
    t2g = np.asarray(p.gpu.get(abi.F_TASK_TORQUES, t1)); t2r = p.ref.get(abi.F_TASK_TORQUES, t1)
    ins = p.ref.get(abi.F_NS_RANK, t0)[:, 0]; types = p.ref.get(abi.F_SINGULARITY_TYPES, t0); cnt = p.ref.get(abi.F_SINGULARITY_COUNTERS, t0)
    tyg = np.asarray(p.gpu.get(abi.F_SINGULARITY_TYPES, t0)); cng = np.asarray(p.gpu.get(abi.F_SINGULARITY_COUNTERS, t0))
    blend = alpha < 1
    print("   blend robots by (ins, types, counters) and share wrong:")
    import collections
    agg = collections.defaultdict(lambda: [0, 0])
    for r in np.nonzero(blend)[0]:
        key = (int(ins[r]), tuple(int(v) for v in types[r]), "c1>c2" if cnt[r][0] > cnt[r][1] else "c1<=c2")
        agg[key][0] += 1; agg[key][1] += int(err[r] > 1e-9)
    for k, v in sorted(agg.items()): print("    ", k, v)
    print("   types equal gpu/ref:", bool((tyg == types).all()), " counters equal:", bool((cng == cnt).all()))
    for r in bad[:4]:
        print("  robot %4d err %.2e alpha %.3f  tau1 err %.2e (|tau1| %.2e)  tau2 err %.2e (|tau2| %.2e)" % (r, err[r], alpha[r], np.abs(t1g[r] - t1r[r]).max(), np.abs(t1r[r]).max(), np.abs(t2g[r] - t2r[r]).max(), np.abs(t2r[r]).max()))

which = sys.argv[1] if len(sys.argv) > 1 else "3"
chain = load_chain("panda")
if which == "3":
    for dec in (abi.BOUNDED_INERTIA_ESTIMATES, abi.FULL_DYNAMIC_DECOUPLING):
        dirs_t, dirs_r = T3.CASES["config3_translation_only"]
        p, t0, t1, q, rng = T3._setup(chain, 1024, 7 + len("config3_translation_only"), dirs_t, dirs_r, dynamic_decoupling_type=dec)
        p.set_jt_params(t1, dynamic_decoupling_type=dec)
        for stacked in (True, False):
            report(p, t0, t1, stacked, "partial M=3 dec=%d" % dec)
else:
    rng = np.random.default_rng(4)
    q = TS._panda_q(rng)
    B = len(q)
    p = Pair(chain, B, abi.PRECISION_F64, q0=q0_of(chain))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, rng.uniform(-0.3, 0.3, q.shape)); p.both("update_model")
    t0 = p.add_mft("end-effector", (0, 0, 0.07)); t1 = p.add_jt()
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    p.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    p.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, q.shape), t1)
    for k in range(3):
        for stacked in (True, False):
            report(p, t0, t1, stacked, "full M=6 tick %d" % k)
