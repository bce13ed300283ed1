"""The multi-device batch behind the ordinary handle (osc_batch_create_multi, include/osc_b200.h): every
entry point fans out over contiguous shards, one host thread + stream per device, host buffers sliced per shard.

On a one-GPU box the shards are placed on device 0 twice / three times (the sharding, slicing, fan-out and joins are the same
code); with more GPUs visible the same test uses distinct devices.  Results must be IDENTICAL to a single batch."""
import numpy as np
import pytest

import osc_abi as abi
from common import FP64_TOL, load_chain, q0_of, random_q, rel_err

pytestmark = pytest.mark.gpu


def _devices(n):
    import osc_b200
    have = osc_b200.lib().osc_device_count()
    return [k % have for k in range(n)]


def _build(chain, B, q, dq, devices, rng_seed=5):
    import osc_b200
    rng = np.random.default_rng(rng_seed)
    b = osc_b200.Batch(chain, B, q0=q0_of(chain), devices=devices)
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = b.add_joint_task(abi.joint_task_desc(chain["n"]))
    b.set(abi.F_Q, q); b.set(abi.F_DQ, dq); b.update_model(); b.reinitialize_tasks()
    x = b.get(abi.F_CURRENT_POSITION, t0)
    b.set(abi.F_GOAL_POSITION, x + rng.normal(0, 0.05, (B, 3)), t0)
    b.set(abi.F_GOAL_LINEAR_VELOCITY, np.array([0.01, -0.02, 0.03]), t0)        # UNIFORM
    b.set(abi.F_JT_GOAL_POSITION, q + rng.normal(0, 0.1, (B, chain["n"])), t1)
    return b, t0, t1


@pytest.mark.parametrize("ndev", [2, 3])
def test_multi_batch_equals_single_batch(ndev):
    chain = load_chain("panda")
    n, B = 7, 1000 + ndev                      # not a multiple of the shard count, nor of 32
    rng = np.random.default_rng(11)
    q = random_q(chain, B, rng, 0.1); dq = rng.uniform(-0.5, 0.5, (B, n))
    one, t0, t1 = _build(chain, B, q, dq, None)
    many, _, _ = _build(chain, B, q, dq, _devices(ndev))
    sh = many.shards()
    assert len(sh) == ndev and sh[0][1] == 0 and sum(s[2] for s in sh) == B
    assert all(sh[k][1] + sh[k][2] == sh[k + 1][1] for k in range(ndev - 1))
    assert many.tick_is_fused() and many.field_size(-1, abi.F_Q) == n
    for tick in range(3):
        qq = q + 0.01 * tick
        one.set(abi.F_Q, qq); many.set(abi.F_Q, qq)
        a, b = one.tick(True), many.tick(True)                       # AOS host output, joined from the shards
        assert np.array_equal(a, b)
    # SOA host buffers: each shard reads / writes its columns of the [k][B_total] array
    qs = np.ascontiguousarray((q + 0.05).T)
    for bt in (one, many):
        bt.set_raw(abi.F_Q, qs.ctypes.data, abi.LAYOUT_SOA, abi.MEM_HOST)
    out1, out2 = np.empty((n, B)), np.full((n, B), np.nan)
    one.L.osc_tick(one.h, 1, out1.ctypes.data, abi.LAYOUT_SOA, abi.MEM_HOST)
    many.L.osc_tick(many.h, 1, out2.ctypes.data, abi.LAYOUT_SOA, abi.MEM_HOST)
    assert np.array_equal(out1, out2)
    for f, t in ((abi.F_M, -1), (abi.F_JACOBIAN, t0), (abi.F_LAMBDA_WORLD, t0), (abi.F_N_TOTAL, -1), (abi.F_TASK_TORQUES, t1), (abi.F_STATUS, -1),
                 (abi.F_SINGULARITY_TYPES, t0)):
        assert np.array_equal(one.get(f, t), many.get(f, t)), f
    assert one.tick_fallback_count() == many.tick_fallback_count() > 0
    # per-call route, parameters, sensed wrench, rollout step, snapshot / restore
    for bt in (one, many):
        p = bt.mft_params(t0)
        p.force_space_dimension = 1; p.force_or_motion_axis[2] = 1.0; p.closed_loop_force_control = 1
        assert bt.set_mft_params(t0, p) == 1                         # the linear reset mask of parametrizeForceMotionSpaces
        bt.update_model(); bt.update_task_models()
        bt.update_sensed_force_and_moment(t0, rng.normal(0, 1, (B, 3)) * 0 + np.arange(B)[:, None] * 1e-3, np.zeros((B, 3)))
    assert np.array_equal(one.compute_torques(True), many.compute_torques(True))
    snaps = [bt.snapshot() for bt in (one, many)]
    for bt in (one, many):
        bt.rollout_step(1e-3, 2, 0.1)
    assert np.array_equal(one.get(abi.F_Q), many.get(abi.F_Q)) and not np.array_equal(one.get(abi.F_Q), qs.T)
    for bt, s in zip((one, many), snaps):
        bt.restore(s)
    assert np.array_equal(one.tick(True), many.tick(True))
    # device buffers are refused on a multi-device batch, with a message
    assert many.L.osc_set_field(many.h, -1, abi.F_Q, qs.ctypes.data, abi.LAYOUT_SOA, abi.MEM_DEVICE) == abi.ERR_UNSUPPORTED
    assert b"HOST buffers only" in many.L.osc_last_error()
    # errors of a shard reach the caller
    assert many.L.osc_task_update_model(many.h, 7, -1) == abi.ERR_INVALID_ARGUMENT and b"task index" in many.L.osc_last_error()
    one.close(); many.close()


def test_multi_batch_against_oracle(oracle_built):
    import orc
    chain = load_chain("panda")
    n, B = 7, 333
    rng = np.random.default_rng(12)
    q = random_q(chain, B, rng, 0.1); dq = rng.uniform(-0.5, 0.5, (B, n))
    many, t0, t1 = _build(chain, B, q, dq, _devices(2))
    ref = orc.OracleBatch(chain, B, q0=q0_of(chain))
    r0 = ref.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07))); r1 = ref.add_joint_task(abi.joint_task_desc(n))
    ref.set(abi.F_Q, q); ref.set(abi.F_DQ, dq); ref.update_model(); ref.reinitialize_tasks()
    for f, t in ((abi.F_GOAL_POSITION, t0), (abi.F_GOAL_LINEAR_VELOCITY, t0), (abi.F_JT_GOAL_POSITION, t1)):
        ref.set(f, many.get(f, t), t)
    assert rel_err(many.tick(True), ref.tick(True))[0] <= FP64_TOL


@pytest.mark.parametrize("layout", ["aos", "soa"])
def test_step_host_one_call_cycle_single_and_sharded(layout):
    """osc_step_host: the input fields, the tick and the torques back in one call.  On a single batch it must equal the
    osc_set_field / osc_tick sequence; on a sharded batch (the shards are issued back to back without waiting, their copies
    overlapping the other shards' kernels) it must give the same bits, also for a batch size that does not divide evenly."""
    import osc_b200
    chain = load_chain("panda")
    B, n = 4099, chain["n"]
    rng = np.random.default_rng(21)
    q = random_q(chain, B, rng, 0.3); dq = rng.uniform(-0.5, 0.5, (B, n))
    lay = abi.LAYOUT_AOS if layout == "aos" else abi.LAYOUT_SOA
    q2 = q + rng.normal(0, 0.01, q.shape); dq2 = dq * 0.5; gq = q + rng.normal(0, 0.1, q.shape)

    def arr(a):
        return np.ascontiguousarray(a if lay == abi.LAYOUT_AOS else a.T)

    taus = []
    for devices, one_call in ((None, False), (None, True), (_devices(3), True)):
        b, t0, t1 = _build(chain, B, q, dq, devices)
        bufs = [arr(q2), arr(dq2), arr(gq)]
        ins = [(-1, abi.F_Q, bufs[0].ctypes.data, lay), (-1, abi.F_DQ, bufs[1].ctypes.data, lay), (t1, abi.F_JT_GOAL_POSITION, bufs[2].ctypes.data, lay)]
        out = np.zeros((B, n) if lay == abi.LAYOUT_AOS else (n, B))
        for _ in range(2):
            if one_call:
                b.step_host(ins, out.ctypes.data, True, lay)
            else:
                for task, f, ptr, l_ in ins:
                    b.set_raw(f, ptr, l_, abi.MEM_HOST, task)
                b._chk(b.L.osc_tick(b.h, 1, out.ctypes.data, lay, abi.MEM_HOST))
        taus.append(np.array(out if lay == abi.LAYOUT_AOS else out.T))
    assert np.array_equal(taus[0], taus[1]), "osc_step_host on one batch differs from set_field + tick"
    assert np.array_equal(taus[0], taus[2]), "osc_step_host on a sharded batch differs from the single batch"
    assert np.isfinite(taus[0]).all() and np.abs(taus[0]).max() > 0
