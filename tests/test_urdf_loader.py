"""The C++ URDF / world-file reader of the library (csrc/osc_urdf.cpp, osc_urdf_load: SURVEY.md 8(f) row f-3) -- CPU tests.

* against the Python reader that produced fixtures/*.chain.json (sai2-primitives-perso_b200/urdf_chain.py), on the synthetic PUMA
  URDF shipped in fixtures/ and, in the build container, on the reference's in-tree URDFs (/root/reference is absent on the
  GPU box: those cases skip there);
* against HAND-DERIVED values read off the Panda URDF text (examples/15-haptic_control_impedance_type/panda_arm.urdf,
  SURVEY.md Appendix C.1): merged end-effector mass / centre of mass / inertia, the calibration posture, limits, efforts,
  joint origins, and the forward kinematics of the calibration posture;
* the world file: gravity attribute and robot <origin> (examples/04-task_and_redundancy/world.urdf:3,5-7), written to tmp_path."""
import json
import math
import os

import numpy as np
import pytest

import osc_abi as abi
from common import ROOT, load_chain

REF = "/root/reference/examples"
URDFS = {"panda": REF + "/15-haptic_control_impedance_type/panda_arm.urdf", "panda_heavy_ee": REF + "/16-haptic_control_admittance_type/panda_arm.urdf",
         "panda_sliding_base": REF + "/06-partial_joint_task/panda_arm_sliding_base.urdf", "rrrrbot": REF + "/11-planar_robot_controller/rrrrbot.urdf",
         "puma560_synthetic": os.path.join(ROOT, "fixtures", "puma560_synthetic.urdf")}


def _same(a, b, tol=1e-14):
    return np.allclose(np.asarray(a, float), np.asarray(b, float), rtol=0, atol=tol)


@pytest.mark.parametrize("name", sorted(URDFS))
def test_cpp_reader_matches_fixture(name):
    import osc_b200
    if not os.path.exists(URDFS[name]):
        pytest.skip("URDF not on this machine (the reference tree only exists in the build container)")
    c, f = osc_b200.load_urdf(URDFS[name]), load_chain(name)
    assert c["n"] == f["n"] and list(c["links"]) == list(f["links"])
    for a, b in zip(c["joints"], f["joints"]):
        assert a["name"] == b["name"] and a["type"] == b["type"]
        for k in ("axis", "xyz", "rot", "lower", "upper", "effort", "q0"):
            assert _same(a[k], b[k]), (name, a["name"], k)
    for a, b in zip(c["bodies"], f["bodies"]):
        for k in ("mass", "com", "inertia"):
            assert _same(a[k], b[k]), (name, k)
    for k in f["links"]:
        assert c["links"][k]["body"] == f["links"][k]["body"] and _same(c["links"][k]["xyz"], f["links"][k]["xyz"]) and _same(c["links"][k]["rot"], f["links"][k]["rot"])
    assert _same(c["world_rot"], np.eye(3)) and _same(c["world_xyz"], [0, 0, 0]) and _same(c["gravity"], [0, 0, -9.81])
    # the chain goes straight into the C ABI struct
    assert abi.chain_desc(c).n == f["n"]


PANDA_TEXT = """<?xml version="1.0" ?>
<!-- hand-written copy of the KINEMATIC / INERTIAL numbers of SURVEY.md Appendix C.1 (the reference's panda_arm.urdf), typed in from
     the table there; visuals, collisions and materials dropped.  A comment with a <tag> inside it must not confuse the reader. -->
<robot name="panda_test">
  <link name="link0"><inertial><origin xyz="0 0 0.05" rpy="0 0 0"/><mass value="4"/><inertia ixx="0.4" iyy="0.4" izz="0.4" ixy="0" ixz="0" iyz="0"/></inertial></link>
  <link name="link1"><inertial><origin xyz="0 0 -0.07" rpy="0 0 0"/><mass value="3"/><inertia ixx="0.3" iyy="0.3" izz="0.3" ixy="0" ixz="0" iyz="0"/></inertial></link>
  <link name="link2"><inertial><origin xyz="0 -0.1 0" rpy="0 0 0"/><mass value="3"/><inertia ixx="0.3" iyy="0.3" izz="0.3" ixy="0" ixz="0" iyz="0"/></inertial></link>
  <link name="link3"><inertial><origin xyz="0.04 0 -0.05" rpy="0 0 0"/><mass value="2"/><inertia ixx="0.2" iyy="0.2" izz="0.2" ixy="0" ixz="0" iyz="0"/></inertial></link>
  <link name="link4"><inertial><origin xyz="-0.04 0.05 0" rpy="0 0 0"/><mass value="2"/><inertia ixx="0.2" iyy="0.2" izz="0.2" ixy="0" ixz="0" iyz="0"/></inertial></link>
  <link name="link5"><inertial><origin xyz="0 0 -0.15" rpy="0 0 0"/><mass value="2"/><inertia ixx="0.2" iyy="0.2" izz="0.2" ixy="0" ixz="0" iyz="0"/></inertial></link>
  <link name="link6"><inertial><origin xyz="0.06 0 0" rpy="0 0 0"/><mass value="1.5"/><inertia ixx="0.1" iyy="0.1" izz="0.1" ixy="0" ixz="0" iyz="0"/></inertial></link>
  <link name="link7"><inertial><origin xyz="0 0 0.17" rpy="0 0 0"/><mass value="1.8"/><inertia ixx="0.09" iyy="0.05" izz="0.07" ixy="0" ixz="0" iyz="0"/></inertial></link>
  <link name='end-effector'><inertial><origin xyz="0 0 0" rpy="0 0 0"/><mass value="0.2"/><inertia ixx="0.01" iyy="0.01" izz="0.01" ixy="0" ixz="0" iyz="0"/></inertial></link>
  <joint name="joint1" type="revolute"><origin xyz="0 0 0.333" rpy="0 0 0"/><parent link="link0"/><child link="link1"/><axis xyz="0 0 1"/>
    <calibration falling="0"/><limit lower="-2.8973" upper="2.8973" effort="87" velocity="2.1750"/></joint>
  <joint name="joint2" type="revolute"><origin xyz="0 0 0" rpy="-1.57079632679 0 0"/><parent link="link1"/><child link="link2"/><axis xyz="0 0 1"/>
    <calibration falling="-25"/><limit lower="-1.7628" upper="1.7628" effort="87" velocity="2.1750"/></joint>
  <joint name="joint3" type="revolute"><origin xyz="0 -0.316 0" rpy="1.57079632679 0 0"/><parent link="link2"/><child link="link3"/><axis xyz="0 0 1"/>
    <calibration falling="0"/><limit lower="-2.8973" upper="2.8973" effort="87" velocity="2.1750"/></joint>
  <joint name="joint4" type="revolute"><origin xyz="0.0825 0 0" rpy="1.57079632679 0 0"/><parent link="link3"/><child link="link4"/><axis xyz="0 0 1"/>
    <calibration falling="-135"/><limit lower="-3.0718" upper="-0.0698" effort="87" velocity="2.1750"/></joint>
  <joint name="joint5" type="revolute"><origin xyz="-0.0825 0.384 0" rpy="-1.57079632679 0 0"/><parent link="link4"/><child link="link5"/><axis xyz="0 0 1"/>
    <calibration falling="0"/><limit lower="-2.8973" upper="2.8973" effort="12" velocity="2.6100"/></joint>
  <joint name="joint6" type="revolute"><origin xyz="0 0 0" rpy="1.57079632679 0 0"/><parent link="link5"/><child link="link6"/><axis xyz="0 0 1"/>
    <calibration falling="105"/><limit lower="-0.0175" upper="3.7525" effort="12" velocity="2.6100"/></joint>
  <joint name="joint7" type="revolute"><origin xyz="0.088 0 0" rpy="1.57079632679 0 0"/><parent link="link6"/><child link="link7"/><axis xyz="0 0 1"/>
    <calibration falling="0"/><limit lower="-2.8973" upper="2.8973" effort="12" velocity="2.6100"/></joint>
  <joint name="joint_ee" type="fixed"><origin xyz="0 0 0.15" rpy="0 0 0"/><parent link="link7"/><child link="end-effector"/></joint>
</robot>
"""


def _rx(a):
    return np.array([[1, 0, 0], [0, math.cos(a), -math.sin(a)], [0, math.sin(a), math.cos(a)]])


def _rz(a):
    return np.array([[math.cos(a), -math.sin(a), 0], [math.sin(a), math.cos(a), 0], [0, 0, 1]])


def test_hand_derived_panda(tmp_path):
    import osc_b200
    path = tmp_path / "panda_test.urdf"
    path.write_text(PANDA_TEXT)
    c = osc_b200.load_urdf(str(path))
    assert c["n"] == 7 and [j["name"] for j in c["joints"]] == ["joint%d" % k for k in range(1, 8)]
    assert [round(math.degrees(j["q0"]), 9) for j in c["joints"]] == [0, -25, 0, -135, 0, 105, 0]      # <calibration falling> in degrees
    assert [j["effort"] for j in c["joints"]] == [87, 87, 87, 87, 12, 12, 12]
    assert (c["joints"][3]["lower"], c["joints"][3]["upper"]) == (-3.0718, -0.0698)
    assert c["links"]["link0"]["body"] == -1 and c["links"]["link7"]["body"] == 6
    ee = c["links"]["end-effector"]                                                              # fixed link: a frame of body 6
    assert ee["body"] == 6 and _same(ee["xyz"], [0, 0, 0.15]) and _same(ee["rot"], np.eye(3))
    # link7 (1.8 kg at z = 0.17, diag(0.09, 0.05, 0.07)) + end-effector (0.2 kg at z = 0.15, 0.01 I) merged by hand:
    m = 2.0
    cz = (1.8 * 0.17 + 0.2 * 0.15) / 2.0                                                         # 0.168
    d7, de = 0.17 - cz, 0.15 - cz
    Ixx = 0.09 + 1.8 * d7 * d7 + 0.01 + 0.2 * de * de
    Iyy = 0.05 + 1.8 * d7 * d7 + 0.01 + 0.2 * de * de
    Izz = 0.07 + 0.01
    b = c["bodies"][6]
    assert abs(b["mass"] - m) < 1e-15 and _same(b["com"], [0, 0, cz], 1e-15) and _same(b["inertia"], np.diag([Ixx, Iyy, Izz]), 1e-15)
    assert abs(cz - 0.168) < 1e-15
    assert c["bodies"][0]["mass"] == 3 and _same(c["bodies"][0]["com"], [0, 0, -0.07])           # link0 (4 kg) is welded: not a body
    # forward kinematics of the calibration posture, composed by hand from the joint table (URDF: translate, rotate by rpy, then Rz(q))
    rolls = [0, -1.57079632679, 1.57079632679, 1.57079632679, -1.57079632679, 1.57079632679, 1.57079632679]
    xyz = [(0, 0, 0.333), (0, 0, 0), (0, -0.316, 0), (0.0825, 0, 0), (-0.0825, 0.384, 0), (0, 0, 0), (0.088, 0, 0)]
    q0 = np.radians([0, -25, 0, -135, 0, 105, 0])
    R, p = np.eye(3), np.zeros(3)
    for k in range(7):
        p = p + R @ np.array(xyz[k]); R = R @ _rx(rolls[k]) @ _rz(q0[k])
    x_hand = p + R @ np.array([0, 0, 0.15 + 0.07])                                               # end-effector + (0, 0, 0.07), 04...cpp:110
    import workloads
    x, Rl, _ = workloads.batched_fk(c, q0[None], "end-effector", (0, 0, 0.07))
    assert _same(x[0], x_hand, 1e-13) and _same(Rl[0], R, 1e-13)
    assert 0.2 < x_hand[0] < 0.7 and abs(x_hand[1]) < 1e-9 and 0.2 < x_hand[2] < 0.8            # a Panda in its ready pose: in front of the base
    if os.path.exists(URDFS["panda"]):   # and the hand-typed table is the reference's file
        ref = osc_b200.load_urdf(URDFS["panda"])
        for a, bb in zip(c["joints"], ref["joints"]):
            assert _same(a["xyz"], bb["xyz"]) and _same(a["rot"], bb["rot"]) and abs(a["q0"] - bb["q0"]) < 1e-15
        for a, bb in zip(c["bodies"], ref["bodies"]):
            assert _same(a["com"], bb["com"]) and _same(a["inertia"], bb["inertia"]) and abs(a["mass"] - bb["mass"]) < 1e-15


def test_world_file_pose_gravity_and_errors(tmp_path):
    import osc_b200
    (tmp_path / "r.urdf").write_text(PANDA_TEXT)
    (tmp_path / "world.urdf").write_text("""<?xml version="1.0" ?>
<world name="demo_world" gravity="0.0 1.5 -9.0">
  <robot name="OTHER"><origin xyz="9 9 9" rpy="0 0 0" /><model dir="x" path="y.urdf" name="o" /></robot>
  <robot name="PANDA">
    <model dir="${SAI2_MODEL_URDF_FOLDER}/panda" path="panda_arm_sphere.urdf" name="panda" />
    <origin xyz="0.1 -0.2 0.3" rpy="0 0 0.5" />
  </robot>
  <static_object name="Floor"><origin xyz="0.0 0.0 0.0" rpy="0 0 0" /></static_object>
</world>
""")
    c = osc_b200.load_urdf(str(tmp_path / "r.urdf"), str(tmp_path / "world.urdf"), "PANDA")
    assert _same(c["gravity"], [0.0, 1.5, -9.0]) and _same(c["world_xyz"], [0.1, -0.2, 0.3]) and _same(c["world_rot"], _rz(0.5), 1e-15)
    first = osc_b200.load_urdf(str(tmp_path / "r.urdf"), str(tmp_path / "world.urdf"))
    assert _same(first["world_xyz"], [9, 9, 9])
    zero_g = osc_b200.load_urdf(str(tmp_path / "r.urdf"))
    assert _same(zero_g["gravity"], [0, 0, -9.81]) and _same(zero_g["world_rot"], np.eye(3))
    with pytest.raises(osc_b200.OscInvalidArgument, match="no <robot name"):
        osc_b200.load_urdf(str(tmp_path / "r.urdf"), str(tmp_path / "world.urdf"), "NOPE")
    with pytest.raises(osc_b200.OscInvalidArgument, match="cannot read"):
        osc_b200.load_urdf(str(tmp_path / "missing.urdf"))
    (tmp_path / "branch.urdf").write_text(PANDA_TEXT.replace('<parent link="link6"/><child link="link7"/>', '<parent link="link5"/><child link="link7"/>'))
    with pytest.raises(osc_b200.OscInvalidArgument, match="serial chain"):
        osc_b200.load_urdf(str(tmp_path / "branch.urdf"))
    (tmp_path / "broken.urdf").write_text(PANDA_TEXT.replace("</robot>", ""))
    with pytest.raises(osc_b200.OscInvalidArgument, match="parse error"):
        osc_b200.load_urdf(str(tmp_path / "broken.urdf"))
