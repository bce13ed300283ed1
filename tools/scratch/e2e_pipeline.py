"""e2e step (all nine input fields in, torques out, pinned host buffers) through osc_step_host: one batch against the same 65,536
robots as S shards on ONE device (copies of shard k + 1 overlapping tick + read-back of shard k), AOS and SOA host layouts."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import osc_abi as abi
import osc_b200
import workloads

chain = json.load(open(os.path.join(ROOT, "fixtures", "panda.chain.json")))
B, n = 65536, 7
w = workloads.config2(chain, B, seed=2)
q0 = [j["q0"] for j in chain["joints"]]


def make(devices):
    b = osc_b200.Batch(chain, B, q0=q0, devices=devices) if devices else osc_b200.Batch(chain, B, q0=q0)
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = b.add_joint_task(abi.joint_task_desc(n))
    b.set(abi.F_Q, w["q"]); b.set(abi.F_DQ, w["dq"]); b.update_model()
    return b, t0, t1


def buffers(t0, t1, layout):
    z3 = np.zeros((B, 3))
    fields = [(abi.F_Q, -1, w["q"]), (abi.F_DQ, -1, w["dq"]), (abi.F_GOAL_POSITION, t0, w["goal_pos"]), (abi.F_GOAL_ORIENTATION, t0, w["goal_ori"]),
              (abi.F_GOAL_LINEAR_VELOCITY, t0, w["goal_linvel"]), (abi.F_GOAL_ANGULAR_VELOCITY, t0, w["goal_angvel"]),
              (abi.F_GOAL_LINEAR_ACCELERATION, t0, z3), (abi.F_GOAL_ANGULAR_ACCELERATION, t0, z3), (abi.F_JT_GOAL_POSITION, t1, w["jt_goal"])]
    keep, ins = [], []
    for f, task, a in fields:
        a = np.ascontiguousarray(a if layout == abi.LAYOUT_AOS else a.T)
        t = torch.from_numpy(a).pin_memory()
        keep.append(t)
        ins.append((task, f, t.data_ptr(), layout))
    tau = torch.zeros((B, n) if layout == abi.LAYOUT_AOS else (n, B), dtype=torch.float64).pin_memory()
    return keep, ins, tau


def run(name, devices, layout, steps=40, per_call=False):
    b, t0, t1 = make(devices)
    keep, ins, tau = buffers(t0, t1, layout)

    def step():
        if per_call:
            for task, f, ptr, lay in ins:
                b.L.osc_set_field(b.h, task, f, ptr, lay, abi.MEM_HOST)
            assert b.L.osc_tick(b.h, 1, tau.data_ptr(), layout, abi.MEM_HOST) == 0
        else:
            b.step_host(ins, tau.data_ptr(), True, layout)
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t) / steps * 1e3
    ref = tau.clone()
    print("%-44s %.4f ms/step  %.3e robot-ticks/s" % (name, ms, B / (ms * 1e-3)), flush=True)
    out = ref.numpy() if layout == abi.LAYOUT_AOS else ref.numpy().T
    return np.array(out)


base = run("one batch, 9 x set_field + tick, SOA", None, abi.LAYOUT_SOA, per_call=True)
for lay, nm in ((abi.LAYOUT_SOA, "SOA"), (abi.LAYOUT_AOS, "AOS")):
    r = run("one batch, osc_step_host, %s" % nm, None, lay)
    assert np.array_equal(r, base), "single %s differs" % nm
    for S in (2, 4, 8):
        r = run("%d shards on one device, osc_step_host, %s" % (S, nm), [0] * S, lay)
        assert np.array_equal(r, base), "%d shards %s differs" % (S, nm)
print("all variants bit-equal")
