import sys, os
import numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, "sai2-primitives-perso_b200"); sys.path.insert(0, "oracle")
import osc_abi as abi, workloads, osc_b200, orc
from common import load_chain, q0_of, random_q
import test_config3_force_popc_gpu as T
chain = load_chain("panda")
B = 262144
rng = np.random.default_rng(3)
q = random_q(chain, B, rng, 0.3); dq = rng.uniform(-0.5, 0.5, (B, 7)); gq = q + rng.normal(0, 0.1, (B, 7))
g = osc_b200.Batch(chain, B, q0=q0_of(chain))
t0, t1 = T._configure(g, chain, q, dq, 3, gq)
seen = {}
for tick in range(60):
    f, _ = workloads.config3_wrench(B, tick)
    tg = np.asarray(T._drive(g, t0, q, tick, f))
    st = g.get(abi.F_STATUS)[:, 0]
    bad = np.nonzero(st & 8)[0]
    for b in bad:
        seen.setdefault(int(b), (tick, int(st[b])))
print("robots with status:", len(seen), list(seen.items())[:10])
bad = np.array(sorted(seen))
if len(bad):
    r = orc.OracleBatch(chain, len(bad), q0=q0_of(chain))
    T._configure(r, chain, q[bad], dq[bad], 3, gq[bad])
    g2 = osc_b200.Batch(chain, B, q0=q0_of(chain))
    T._configure(g2, chain, q, dq, 3, gq)
    for tick in range(60):
        f, _ = workloads.config3_wrench(B, tick)
        tg = np.asarray(T._drive(g2, t0, q, tick, f)); tr = np.asarray(T._drive(r, t0, q[bad], tick, f[bad]))
        e = np.abs(tg[bad] - tr).max(axis=1) / np.maximum(np.abs(tr).max(axis=1), 1e-3)
        if e.max() > 1e-9: print("tick", tick, "err", e.max(), "robot", bad[e.argmax()])
    print("oracle status", r.get(abi.F_STATUS)[:, 0][:10])
    print("sv gpu", g2.get(abi.F_SINGULAR_VALUES, t0)[bad][:4]); print("sv ref", r.get(abi.F_SINGULAR_VALUES, t0)[:4])
