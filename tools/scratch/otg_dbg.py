import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in ("sai2-primitives-perso_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import osc_abi as abi, workloads
from test_otg_gpu import _golden_batch
g = np.load(os.path.join(ROOT, "tests/golden/otg_tasks.npz"))
b, w, t0, t1 = _golden_batch(g)
rec = workloads.otg_rollout(b, w, t0, t1, int(g["T"]), abi, per_call=True)
np.set_printoptions(linewidth=200, precision=6)
for k in ("des_pos", "des_ori", "jt_des_pos", "jt_des_vel", "tau"):
    d = np.abs(rec[k] - g[k]).max(axis=(1, 2))
    print(k, "first bad sample", np.argmax(d > 1e-9), "max", d.max())
    print("   per-sample err", d[:45])
print("flags0 gpu", rec["flags0"].sum(axis=(1, 2)))
print("flags0 ref", g["flags0"].sum(axis=(1, 2)))
print("flags1 gpu", rec["flags1"].sum(axis=(1, 2)))
print("flags1 ref", g["flags1"].sum(axis=(1, 2)))
print("res0", np.unique(b.get(abi.F_OTG_RESULT, t0), return_counts=True), "res1", np.unique(b.get(abi.F_OTG_RESULT, t1), return_counts=True))
print("status", np.unique(b.get(abi.F_STATUS), return_counts=True))
# ---- raw state after enable and after one tick
b, w, t0, t1 = _golden_batch(g)
st = b.get(abi.F_OTG_STATE, t0)
D = 6
names = dict(TGT_P=0, TGT_V=D, IN_P=2*D, IN_V=3*D, IN_A=4*D, OUT_P=5*D, OUT_V=6*D, OUT_A=7*D, TRAJ=8*D, TIME=19*D, DUR=19*D+1, REF=19*D+2, GORI=19*D+11, GANG=19*D+20)
def dump(st, r=0):
    for k, o in names.items():
        w_ = dict(TRAJ=66, TIME=1, DUR=1, REF=9, GORI=9, GANG=3).get(k, 6)
        print("  %-6s" % k, st[r, o:o + w_])
print("after enable: cur pos", b.get(abi.F_CURRENT_POSITION, t0)[0], "flags", b.get(abi.F_OTG_FLAGS, t0)[0])
dump(st)
b.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); b.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0)
print("goal pos", w["goal_pos"][0], "readback", b.get(abi.F_GOAL_POSITION, t0)[0])
b.update_model(); b.update_task_models(); b.compute_torques(True)
print("after tick 0: flags", b.get(abi.F_OTG_FLAGS, t0)[0], "res", b.get(abi.F_OTG_RESULT, t0)[0], "des", b.get(abi.F_DESIRED_POSITION, t0)[0], "golden", g["des_pos"][0][0])
dump(b.get(abi.F_OTG_STATE, t0))
