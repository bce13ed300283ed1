"""The C++ facade (include/Sai2PrimitivesBatched.h): compiles against the C ABI with plain g++ (CPU check), refuses to run
without a CUDA device (no CPU fallback), and on the GPU reproduces the oracle through the reference's method names."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import osc_abi as abi
from common import FP64_TOL, ROOT, load_chain, q0_of, random_q, rel_err

PKG = os.path.join(ROOT, "sai2-primitives-perso_b200")
EXE = os.path.join(ROOT, "tests", "cpp", "_build", "facade_demo")


def _build(name="facade_demo"):
    EXE = os.path.join(ROOT, "tests", "cpp", "_build", name)
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    src = os.path.join(ROOT, "tests", "cpp", name + ".cpp")
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < max(os.path.getmtime(src), os.path.getmtime(os.path.join(ROOT, "include", "Sai2PrimitivesBatched.h"))):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), src, "-o", EXE,
                               "-L", PKG, "-losc_b200", "-Wl,-rpath," + PKG])
    return EXE


def _inputs(tmp_path, chain, B, seed):
    rng = np.random.default_rng(seed)
    n = chain["n"]
    import workloads
    w = workloads.config2(chain, B, seed)
    (tmp_path / "chain.bin").write_bytes(bytes(abi.chain_desc(chain)))
    with open(tmp_path / "in.bin", "wb") as f:
        for a in (w["q"], w["dq"], w["goal_pos"], w["goal_ori"], w["jt_goal"]):
            f.write(np.ascontiguousarray(a, np.float64).tobytes())
    return w


def test_facade_compiles_and_refuses_cpu(tmp_path):
    import osc_b200
    osc_b200.lib()
    exe = _build()
    if osc_b200.lib().osc_device_count() > 0:
        pytest.skip("a CUDA device is present: covered by the gpu test")
    chain = load_chain("panda")
    _inputs(tmp_path, chain, 8, 1)
    r = subprocess.run([exe, str(tmp_path / "chain.bin"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "8", "1", "0"],
                       capture_output=True, text=True)
    assert r.returncode == 10, (r.returncode, r.stderr)
    assert "no CUDA device" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4, 5, 6, 7])
def test_facade_matches_oracle(tmp_path, oracle_built, mode):
    """Modes of tests/cpp/facade_demo.cpp.  0: the reference's three calls per cycle, which the facade's control-cycle fusion turns
    into one fused launch (the demo checks the count); 4: the same with fusion off; 5: the same with a getter / a setQ between the
    calls of two of the cycles (those cycles run per call); 6: closed-loop force control with updateSensedForceAndMoment() inside every
    (fused) cycle and a new q per cycle.  All must give the oracle's torques."""
    orc = oracle_built
    exe = _build()
    chain = load_chain("panda")
    B, ticks = 256, (60 if mode == 3 else (5 if mode == 5 else 3))
    w = _inputs(tmp_path, chain, B, 40 + mode)
    r = subprocess.run([exe, str(tmp_path / "chain.bin"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), str(B), str(ticks), str(mode)],
                       capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stderr)
    out = np.fromfile(tmp_path / "out.bin", dtype=np.float64)
    tau, x = out[:B * 7].reshape(B, 7), out[B * 7:].reshape(B, 3)
    o = orc.OracleBatch(chain, B, q0=q0_of(chain))
    o.set(abi.F_Q, w["q"]); o.set(abi.F_DQ, w["dq"]); o.update_model()
    t0 = o.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = o.add_joint_task(abi.joint_task_desc(7))
    p = o.mft_params(t0)
    for a in range(3):
        p.kp_pos[a], p.kv_pos[a], p.kp_ori[a], p.kv_ori[a] = 120.0, 22.0, 210.0, 29.0
    o.set_mft_params(t0, p)
    jp = o.joint_task_params(t1)
    for a in range(7):
        jp.kp[a], jp.kv[a] = 55.0, 15.0
    o.set_joint_task_params(t1, jp)
    if mode == 3:  # the reference defaults: both internal OTGs on (MotionForceTask.h:67-72, JointTask.h:38-41)
        o.enable_internal_otg(t0, [0.3, np.pi / 3], [2.0, 2 * np.pi])
        o.enable_internal_otg(t1, np.pi / 3, 2 * np.pi)
    o.set(abi.F_GOAL_POSITION, w["goal_pos"], t0); o.set(abi.F_GOAL_ORIENTATION, w["goal_ori"], t0)
    o.set(abi.F_GOAL_LINEAR_VELOCITY, np.array([0.01, -0.02, 0.03]), t0); o.set(abi.F_JT_GOAL_POSITION, w["jt_goal"], t1)
    if mode in (6, 7):
        p = o.mft_params(t0)
        p.force_space_dimension = 1
        p.force_or_motion_axis[0], p.force_or_motion_axis[1], p.force_or_motion_axis[2] = 0.0, 0.0, 1.0
        p.closed_loop_force_control = 1
        o.set_mft_params(t0, p)
        o.set(abi.F_GOAL_FORCE, np.array([0.0, 0.0, 5.0]), t0)
    q = np.array(w["q"], float)
    for t in range(ticks):
        if mode in (6, 7):
            q = q + 1e-3 * (np.arange(7) - 3.0)
            o.set(abi.F_Q, q); o.update_model()
            o.update_sensed_force_and_moment(t0, np.array([0.3, -0.2, 4.0 + 0.1 * t]), np.array([0.01, 0.02, -0.03]))
            # parametrizeForceMotionSpaces re-initialised the linear goal to the current position (MotionForceTask.cpp:830-858); the
            # demo writes its position goal again every cycle
            o.set(abi.F_GOAL_POSITION, w["goal_pos"], t0)
        ref = o.tick(mode != 2)
    assert rel_err(tau, ref)[0] <= FP64_TOL
    assert rel_err(x, o.get(abi.F_CURRENT_POSITION, t0))[0] <= FP64_TOL


PUMA_URDF = os.path.join(ROOT, "fixtures", "puma560_synthetic.urdf")


def _world(tmp_path):
    w = tmp_path / "world.urdf"
    w.write_text('<?xml version="1.0" ?>\n<world name="w" gravity="0.0 0.0 -3.5">\n <robot name="PUMA"><origin xyz="0.5 0 0.2" rpy="0 0 0.3" />'
                 '<model dir="." path="puma560_synthetic.urdf" name="puma" /></robot>\n</world>\n')
    return str(w)


def test_facade_urdf_loader_on_cpu(tmp_path):
    """a C++ user loads the URDF + world file through the facade with plain g++ (no GPU needed for that part)"""
    import osc_b200
    osc_b200.lib()
    exe = _build("facade_urdf_multi")
    r = subprocess.run([exe, PUMA_URDF, _world(tmp_path), "64", "2", str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert "dof 6 gravity 0.000 0.000 -3.500" in r.stdout and "end-effector body 5" in r.stdout, (r.stdout, r.stderr)
    if osc_b200.lib().osc_device_count() == 0:
        assert r.returncode == 10 and "no CUDA device" in r.stderr
    else:
        assert r.returncode == 0, r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("shards", [2, 3])
def test_facade_urdf_multi_device(tmp_path, shards):
    """RobotBatch(UrdfModel, B, devices): the sharded batch gives the single batch's torques, bit for bit, through the facade"""
    exe = _build("facade_urdf_multi")
    B = 1000
    r = subprocess.run([exe, PUMA_URDF, _world(tmp_path), str(B), str(shards), str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    out = np.fromfile(tmp_path / "out.bin", dtype=np.float64).reshape(2, B, 6)
    assert np.isfinite(out).all() and np.abs(out[0]).max() > 0.1
    assert np.array_equal(out[0], out[1])
