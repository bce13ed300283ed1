"""CPU, world_size 2 over gloo: the multi-rank host logic (contiguous shards, no data-path collective, MAX-over-ranks timing,
host-side gather).  The per-shard compute stand-in here is the CPU oracle (test infrastructure); on GPUs each rank drives
its own osc_b200.Batch exactly the same way (bench.py --gpus N)."""
import os
import socket

import numpy as np
import pytest

import sharding
from common import ROOT, load_chain, q0_of


def test_shard_ranges_cover_exactly():
    for n in (1, 7, 64, 65536, 10**6 + 3):
        for world in (1, 2, 3, 4, 8):
            r = [sharding.shard_range(n, world, k) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)


def _worker(rank, world, port, B, out_path):
    import sys
    for p in ("sai2-primitives-perso_b200", "oracle", "tests"):
        sys.path.insert(0, os.path.join(ROOT, p))
    import torch.distributed as dist
    import orc
    import osc_abi as abi
    import workloads
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    chain = load_chain("panda")
    w = sharding.shard(workloads.config2(chain, B, seed=7), B, world, rank)
    lo, hi = sharding.shard_range(B, world, rank)
    o = orc.OracleBatch(chain, hi - lo, q0=q0_of(chain))
    o.set(abi.F_Q, w["q"]); o.set(abi.F_DQ, w["dq"]); o.update_model()
    t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = o.add_joint_task(abi.joint_task_desc(7))
    o.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); o.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0); o.set(abi.F_JT_GOAL_POSITION, w["jt_goal"], t1)
    dist.barrier()
    tau = o.tick(True)
    slowest = sharding.max_over_ranks(1.0 + rank, dist)
    full = sharding.gather_rows(tau, B, dist)
    if rank == 0:
        np.savez(out_path, tau=full, slowest=slowest)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shards_reproduce_the_single_batch(tmp_path, oracle_built):
    import torch.multiprocessing as mp
    import osc_abi as abi
    import workloads
    B, world = 37, 2  # odd on purpose: ragged shards
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "out.npz")
    mp.spawn(_worker, args=(world, port, B, out), nprocs=world, join=True)
    got = np.load(out)
    assert float(got["slowest"]) == 2.0  # MAX over ranks
    chain = load_chain("panda")
    w = workloads.config2(chain, B, seed=7)
    o = oracle_built.OracleBatch(chain, B, q0=q0_of(chain))
    o.set(abi.F_Q, w["q"]); o.set(abi.F_DQ, w["dq"]); o.update_model()
    t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = o.add_joint_task(abi.joint_task_desc(7))
    o.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); o.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0); o.set(abi.F_JT_GOAL_POSITION, w["jt_goal"], t1)
    ref = o.tick(True)
    assert got["tau"].shape == ref.shape
    assert np.array_equal(got["tau"], ref), "sharding must not change a single bit: robots are independent"
