"""osc_tick_noalloc (include/osc_b200.h): the fused tick with the library's static device scratch instead of the
cudaMalloc'ed fallback list -- no counterpart in the reference.

Two identical batches are driven with the same inputs; one ticks through osc_tick, the other through osc_tick_noalloc
(torques into a caller-owned DEVICE buffer), twice in a row.  The batch contains robots on the Panda's singular sets so
that the scratch list (the robots handed to k_tick_tail) is really used, and it must not be empty on either call.
Collected last on purpose (pytest -x): a failure here must not hide another test."""
import ctypes as C

import numpy as np
import pytest

import osc_abi as abi
from common import FP64_TOL, load_chain, q0_of, rel_err

pytestmark = pytest.mark.gpu


def _batch(chain, q, dq, goal, jgoal):
    import osc_b200
    B = len(q)
    b = osc_b200.Batch(chain, B, q0=q0_of(chain))
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = b.add_joint_task(abi.joint_task_desc(chain["n"]))
    b.set(abi.F_Q, q); b.set(abi.F_DQ, dq); b.update_model(); b.reinitialize_tasks()
    b.set(abi.F_GOAL_POSITION, b.get(abi.F_CURRENT_POSITION, t0) + goal, t0)
    b.set(abi.F_JT_GOAL_POSITION, q + jgoal, t1)
    return b


def test_noalloc_tick_matches_tick_twice_in_a_row():
    import torch
    chain = load_chain("panda")
    n, B = chain["n"], 4096 + 17
    rng = np.random.default_rng(77)
    q = np.tile(q0_of(chain), (B, 1)) + rng.normal(0, 0.3, (B, n))
    sing = rng.random(B) < 0.25                      # a quarter of the robots on / near the singular sets -> fallback list
    q[sing, 3] = -0.0698 - np.abs(rng.normal(0, 1e-3, sing.sum()))
    q[sing, 5] = rng.normal(0, 1e-3, sing.sum())
    dq = rng.uniform(-0.5, 0.5, (B, n))
    goal, jgoal = rng.normal(0, 0.05, (B, 3)), rng.normal(0, 0.1, (B, n))
    a = _batch(chain, q, dq, goal, jgoal)            # existing entry point
    b = _batch(chain, q, dq, goal, jgoal)            # new entry point
    assert a.tick_is_fused() and b.tick_is_fused()
    out = torch.full((B, n), float("nan"), dtype=torch.float64, device="cuda")
    out_soa = torch.full((n, B), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    for call in range(2):
        dq2 = dq * (1.0 - 0.5 * call)                # the second call sees different inputs (stale scratch would show)
        a.set(abi.F_DQ, dq2); b.set(abi.F_DQ, dq2)
        want = a.tick(True)
        assert a.tick_fallback_count() > 0, "the batch was meant to exercise the scratch list"
        b.tick_noalloc(out.data_ptr(), True, abi.LAYOUT_AOS)
        b.synchronize()
        got = out.cpu().numpy()
        e, where = rel_err(got, want)
        assert e <= FP64_TOL, "call %d: osc_tick_noalloc differs from osc_tick: %.3e at robot %d" % (call, e, where)
        assert np.array_equal(b.get(abi.F_TAU), got)
        for f, t in ((abi.F_TASK_TORQUES, 0), (abi.F_TASK_TORQUES, 1), (abi.F_JT_INTEGRATED_ERROR, 1), (abi.F_ALPHA, 0)):
            e, where = rel_err(b.get(f, t), a.get(f, t), 1.0 if f == abi.F_ALPHA else None)
            assert e <= FP64_TOL, (call, f, t, e, where)
    # SOA destination, no destination, and a regular tick of the same batch afterwards (its own list is untouched)
    a.tick(True, fetch=False)
    b.tick_noalloc(out_soa.data_ptr(), True, abi.LAYOUT_SOA)
    b.synchronize()
    e, _ = rel_err(out_soa.cpu().numpy().T, a.get(abi.F_TAU))
    assert e <= FP64_TOL
    want = a.tick(True)
    b.tick_noalloc(None, True)
    e, _ = rel_err(b.get(abi.F_TAU), want)
    assert e <= FP64_TOL
    e, _ = rel_err(b.tick(True), a.tick(True))
    assert e <= FP64_TOL
    a.close(); b.close()


def test_noalloc_rejects_what_it_cannot_serve():
    import osc_b200
    chain = load_chain("panda")
    L = osc_b200.lib()
    big = osc_b200.Batch(chain, abi.NOALLOC_MAX_ROBOTS + 1, q0=q0_of(chain))
    big.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    big.add_joint_task(abi.joint_task_desc(chain["n"]))
    assert L.osc_tick_noalloc(big.h, 1, None, abi.LAYOUT_AOS) == abi.ERR_INVALID_ARGUMENT
    assert b"OSC_NOALLOC_MAX_ROBOTS" in L.osc_last_error()
    big.close()
    part = osc_b200.Batch(chain, 64, q0=q0_of(chain))  # a partial JointTask in front (examples/06) runs the per-call route: not served
    part.add_joint_task(abi.joint_task_desc(chain["n"], np.eye(chain["n"])[:2]))
    part.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    assert L.osc_tick_noalloc(part.h, 1, None, abi.LAYOUT_AOS) == abi.ERR_UNSUPPORTED
    part.close()
    jt_only = osc_b200.Batch(chain, 64, q0=q0_of(chain))
    jt_only.add_joint_task(abi.joint_task_desc(chain["n"]))
    assert L.osc_tick_noalloc(jt_only.h, 1, None, abi.LAYOUT_AOS) == abi.OK   # the full JointTask alone (examples/01): k_tick_joint, no scratch at all
    jt_only.close()
    one = osc_b200.Batch(chain, 64, q0=q0_of(chain))
    one.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    assert L.osc_tick_noalloc(one.h, 1, None, abi.LAYOUT_AOS) == abi.OK   # the task alone (examples/07): fused since round 2
    one.add_joint_task(abi.joint_task_desc(chain["n"]))
    assert L.osc_tick_noalloc(one.h, 1, C.c_void_p(8), abi.LAYOUT_UNIFORM) == abi.ERR_INVALID_ARGUMENT
    assert L.osc_tick_noalloc(one.h, 1, None, abi.LAYOUT_AOS) == abi.OK
    one.close()
