"""CPU: the C-ABI library loads and exports every symbol include/osc_b200.h declares; the ctypes mirror (osc_abi.py)
matches the header (constants, struct sizes); compute entry points fail loudly without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

import osc_abi as abi
from common import ROOT, load_chain, q0_of

HEADER = os.path.join(ROOT, "include", "osc_b200.h")
LIB = os.path.join(ROOT, "sai2-primitives-perso_b200", "libosc_b200.so")


def _header_text():
    return open(HEADER).read()


def _declared_functions():
    txt = re.sub(r"/\*.*?\*/", "", _header_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(osc_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    import osc_b200
    L = osc_b200.lib()
    declared = _declared_functions()
    assert len(declared) >= 50
    nm = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\bT (osc_[a-z0-9_]+)", nm))
    missing = [f for f in declared if f not in exported]
    assert not missing, "declared in osc_b200.h but not exported: %s" % missing
    extra = sorted(f for f in exported if f not in declared)
    assert not extra, "exported but not declared (visibility leak): %s" % extra
    for f in declared:
        getattr(L, f)
    assert sorted(osc_b200.EXPORTED) == declared, "osc_b200.EXPORTED out of date"


def test_constants_match_header():
    txt = _header_text()
    enums = dict((k, int(eval(v))) for k, v in re.findall(r"\b(OSC_[A-Z0-9_]+)\s*=\s*(-?\d+(?:\s*<<\s*\d+)?)", txt))
    defines = dict((k, int(v)) for k, v in re.findall(r"#define\s+(OSC_[A-Z0-9_]+)\s+(\d+)", txt))
    assert defines["OSC_NOALLOC_MAX_ROBOTS"] == abi.NOALLOC_MAX_ROBOTS
    assert defines["OSC_MAX_DOF"] == abi.MAX_DOF and defines["OSC_POPC_RING_CAPACITY"] == abi.POPC_RING_CAPACITY
    checked = 0
    for name, val in vars(abi).items():
        if not isinstance(val, int) or name.startswith("_") or name in ("MAX_DOF", "MAX_TASKS", "POPC_RING_CAPACITY", "MAX_SINGULARITY_BUFFER", "NOALLOC_MAX_ROBOTS"):
            continue
        key = "OSC_" + name
        if key in enums:
            assert enums[key] == val, (key, enums[key], val)
            checked += 1
    assert checked >= 90, checked
    fields = {k: v for k, v in enums.items() if k.startswith("OSC_F_")}
    assert len(fields) == len([n for n in vars(abi) if n.startswith("F_")]), "a field id is missing from osc_abi.py"


def test_struct_layout_matches_ctypes(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "osc_b200.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(osc_chain_desc),sizeof(osc_mft_desc),sizeof(osc_mft_params),sizeof(osc_joint_task_desc),sizeof(osc_joint_task_params),'
                   'offsetof(osc_mft_params,apply_disturbance_compensation),offsetof(osc_chain_desc,gravity));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(abi.ChainDesc), C.sizeof(abi.MftDesc), C.sizeof(abi.MftParams), C.sizeof(abi.JointTaskDesc), C.sizeof(abi.JointTaskParams),
            abi.MftParams.apply_disturbance_compensation.offset, abi.ChainDesc.gravity.offset]
    assert got == want


def test_no_cpu_fallback():
    import osc_b200
    L = osc_b200.lib()
    if L.osc_device_count() > 0:
        pytest.skip("CUDA device present")
    chain = load_chain("panda")
    with pytest.raises(osc_b200.OscError) as e:
        osc_b200.Batch(chain, 4, q0=q0_of(chain))
    assert e.value.code == abi.ERR_NO_DEVICE
    t = C.c_double()
    assert L.osc_measure_fp64_peak(0, C.byref(t), None) == abi.ERR_NO_DEVICE


def test_defaults_match_reference_headers():
    """defaults: src/tasks/MotionForceTask.h:40-75, src/tasks/JointTask.h:31-45, src/tasks/SingularityHandler.cpp:10-20,
    MotionForceTask.cpp:197 (values restated in SURVEY.md A.2 / A.6 / A.8)"""
    import osc_b200
    L = osc_b200.lib()
    p = abi.MftParams()
    L.osc_mft_default_params(C.byref(p))
    assert list(p.kp_pos) == [100.0] * 3 and list(p.kv_pos) == [20.0] * 3 and list(p.ki_pos) == [0.0] * 3
    assert list(p.kp_ori) == [200.0] * 3 and list(p.kv_ori) == [28.3] * 3
    assert (p.kp_force, p.kv_force, p.ki_force) == (0.7, 10.0, 1.3) and (p.kp_moment, p.kv_moment, p.ki_moment) == (0.7, 10.0, 1.3)
    assert p.kff_force == 0.95 and p.kff_moment == 0.95
    assert p.max_force_control_feedback_output == 20.0 and p.max_moment_control_feedback_output == 10.0
    assert p.force_space_dimension == 0 and p.moment_space_dimension == 0 and not p.closed_loop_force_control and not p.passivity_enabled
    assert not p.use_velocity_saturation and p.linear_saturation_velocity == 0.3
    assert p.dynamic_decoupling_type == abi.BOUNDED_INERTIA_ESTIMATES and p.bie_threshold == 0.1
    assert (p.s_min, p.s_max) == (6e-3, 6e-2) and (p.kp_type_1, p.kv_type_1, p.kv_type_2) == (50.0, 14.0, 5.0)
    assert p.s_abs_tol == 1e-3 and p.type_1_tol == 0.5 and p.type_2_torque_ratio == 1e-2 and p.perturb_step_size == 5.0 and p.buffer_size == 200
    assert p.apply_disturbance_compensation == 0  # quirk Q1
    j = abi.JointTaskParams()
    L.osc_joint_task_default_params(C.byref(j), 7)
    assert list(j.kp)[:7] == [50.0] * 7 and list(j.kv)[:7] == [14.0] * 7 and list(j.ki)[:7] == [0.0] * 7
    assert j.dynamic_decoupling_type == abi.BOUNDED_INERTIA_ESTIMATES and j.bie_threshold == 0.1 and not j.use_velocity_saturation
