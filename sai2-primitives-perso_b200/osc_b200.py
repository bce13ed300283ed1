"""ctypes binding of libosc_b200.so (the C ABI of include/osc_b200.h) for tests, bench.py and Python users.

The product is the shared library (hand-written sm_100a kernels behind a C ABI) and the C++ facade
include/Sai2PrimitivesBatched.h; this module is only the thinnest possible Python view of the same entry points.
There is NO CPU fallback: importing works without a GPU (so that the ABI can be inspected), but creating a batch
without a CUDA device raises, and a missing library raises at import of `lib()`.
"""
import ctypes as C
import os

import numpy as np

import osc_abi as abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("OSC_B200_LIB") or os.path.join(_HERE, "libosc_b200.so")  # override: A/B runs of two builds
_LIB = None


class OscError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("osc_b200 error %d: %s" % (code, msg))
        self.code = code


class OscInvalidArgument(OscError, ValueError):
    """the reference throws std::invalid_argument in this situation"""


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise ImportError("libosc_b200.so is missing: run `python sai2-primitives-perso_b200/build.py` "
                          "(there is no CPU fallback for the batched controller)")
    L = C.CDLL(LIB_PATH)
    vp, i, i64 = C.c_void_p, C.c_int, C.c_int64
    sig = {
        "osc_last_error": (C.c_char_p, []),
        "osc_version": (i, []),
        "osc_device_count": (i, []),
        "osc_urdf_load": (i, [C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(vp)]),
        "osc_urdf_free": (None, [vp]),
        "osc_urdf_chain": (i, [vp, C.POINTER(abi.ChainDesc), C.POINTER(C.c_double)]),
        "osc_urdf_link": (i, [vp, C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "osc_urdf_link_count": (i, [vp]),
        "osc_urdf_link_name": (C.c_char_p, [vp, i]),
        "osc_urdf_joint_name": (C.c_char_p, [vp, i]),
        "osc_batch_create": (i, [C.POINTER(abi.ChainDesc), i64, i, i, C.POINTER(C.c_double), C.POINTER(vp)]),
        "osc_batch_create_multi": (i, [C.POINTER(abi.ChainDesc), i64, i, C.POINTER(C.c_int), i, C.POINTER(C.c_double), C.POINTER(vp)]),
        "osc_batch_shard_count": (i, [vp]),
        "osc_batch_shard": (vp, [vp, i, C.POINTER(C.c_int), C.POINTER(i64), C.POINTER(i64)]),
        "osc_batch_destroy": (None, [vp]),
        "osc_batch_set_stream": (i, [vp, vp]),
        "osc_synchronize": (i, [vp]),
        "osc_batch_size": (i64, [vp]),
        "osc_batch_dof": (i, [vp]),
        "osc_batch_precision": (i, [vp]),
        "osc_batch_add_motion_force_task": (i, [vp, C.POINTER(abi.MftDesc)]),
        "osc_batch_add_joint_task": (i, [vp, C.POINTER(abi.JointTaskDesc)]),
        "osc_batch_task_count": (i, [vp]),
        "osc_batch_task_type": (i, [vp, i]),
        "osc_batch_task_dof": (i, [vp, i]),
        "osc_mft_default_params": (None, [C.POINTER(abi.MftParams)]),
        "osc_joint_task_default_params": (None, [C.POINTER(abi.JointTaskParams), i]),
        "osc_mft_get_params": (i, [vp, i, C.POINTER(abi.MftParams)]),
        "osc_mft_set_params": (i, [vp, i, C.POINTER(abi.MftParams)]),
        "osc_joint_task_get_params": (i, [vp, i, C.POINTER(abi.JointTaskParams)]),
        "osc_joint_task_set_params": (i, [vp, i, C.POINTER(abi.JointTaskParams)]),
        "osc_enable_gravity_compensation": (i, [vp, i]),
        "osc_joint_task_set_gains_unsafe": (i, [vp, i, vp, vp, vp, i]),
        "osc_mft_selection_projectors": (i, [vp, i, vp, vp]),
        "osc_task_goal_reached": (i, [vp, i, i, C.c_double, vp, i, C.POINTER(i64)]),
        "osc_field_size": (i, [vp, i, i]),
        "osc_set_field": (i, [vp, i, i, vp, i, i]),
        "osc_get_field": (i, [vp, i, i, vp, i, i]),
        "osc_mft_update_sensed_force_and_moment": (i, [vp, i, vp, vp, i, i]),
        "osc_update_model": (i, [vp]),
        "osc_update_kinematics": (i, [vp]),
        "osc_defer_model_update": (i, [vp]),
        "osc_task_update_model": (i, [vp, i, i]),
        "osc_update_task_models": (i, [vp]),
        "osc_task_compute_torques": (i, [vp, i, i]),
        "osc_compute_torques": (i, [vp, i, vp, i, i]),
        "osc_tick": (i, [vp, i, vp, i, i]),
        "osc_tick_noalloc": (i, [vp, i, vp, i]),
        "osc_step_host": (i, [vp, vp, i, i, vp, i]),
        "osc_tick_is_fused": (i, [vp]),
        "osc_batch_set_tick_outputs": (i, [vp, i]),
        "osc_tick_fallback_count": (i64, [vp]),
        "osc_launch_count": (i64, [vp]),
        "osc_reinitialize_tasks": (i, [vp, i]),
        "osc_snapshot_bytes": (i64, [vp]),
        "osc_snapshot": (i, [vp, vp, i]),
        "osc_restore": (i, [vp, vp, i]),
        "osc_measure_fp64_peak": (i, [i, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "osc_task_enable_internal_otg_acceleration_limited": (i, [vp, i, vp, i]),
        "osc_task_enable_internal_otg_jerk_limited": (i, [vp, i, vp, i]),
        "osc_task_disable_internal_otg": (i, [vp, i]),
        "osc_task_internal_otg_enabled": (i, [vp, i]),
        "osc_rollout_step": (i, [vp, C.c_double, i, C.c_double]),
        "osc_rollout": (i, [vp, i, i, C.c_double, i, C.c_double, vp, i, i]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype = res
        f.argtypes = args
    _LIB = L
    return L


EXPORTED = ["osc_last_error", "osc_version", "osc_device_count", "osc_urdf_load", "osc_urdf_free", "osc_urdf_chain", "osc_urdf_link", "osc_urdf_link_count",
            "osc_urdf_link_name", "osc_urdf_joint_name", "osc_batch_create", "osc_batch_create_multi", "osc_batch_shard_count", "osc_batch_shard", "osc_batch_destroy",
            "osc_batch_set_stream", "osc_synchronize", "osc_batch_size", "osc_batch_dof", "osc_batch_precision",
            "osc_batch_add_motion_force_task", "osc_batch_add_joint_task", "osc_batch_task_count", "osc_batch_task_type",
            "osc_batch_task_dof", "osc_mft_default_params", "osc_joint_task_default_params", "osc_mft_get_params",
            "osc_mft_set_params", "osc_joint_task_get_params", "osc_joint_task_set_params",
            "osc_enable_gravity_compensation", "osc_joint_task_set_gains_unsafe", "osc_mft_selection_projectors", "osc_task_goal_reached", "osc_field_size", "osc_set_field", "osc_get_field",
            "osc_mft_update_sensed_force_and_moment", "osc_update_model", "osc_update_kinematics", "osc_defer_model_update",
            "osc_task_update_model", "osc_update_task_models", "osc_task_compute_torques", "osc_compute_torques",
            "osc_tick", "osc_tick_noalloc", "osc_step_host", "osc_tick_is_fused", "osc_batch_set_tick_outputs", "osc_tick_fallback_count", "osc_launch_count", "osc_reinitialize_tasks", "osc_snapshot_bytes",
            "osc_snapshot", "osc_restore", "osc_measure_fp64_peak", "osc_task_enable_internal_otg_acceleration_limited",
            "osc_task_enable_internal_otg_jerk_limited", "osc_task_disable_internal_otg", "osc_task_internal_otg_enabled",
            "osc_rollout_step", "osc_rollout"]


def load_urdf(urdf_path, world_path=None, robot_name=None):
    """osc_urdf_load (the library's C++ URDF / world-file reader) -> the chain dict of fixtures/*.chain.json"""
    L = lib()
    h = C.c_void_p()
    rc = L.osc_urdf_load(urdf_path.encode(), world_path.encode() if world_path else None, robot_name.encode() if robot_name else None, C.byref(h))
    if rc < 0:
        raise (OscInvalidArgument if rc == abi.ERR_INVALID_ARGUMENT else OscError)(rc, L.osc_last_error().decode())
    try:
        cd = abi.ChainDesc()
        q0 = (C.c_double * abi.MAX_DOF)()
        L.osc_urdf_chain(h, C.byref(cd), q0)
        n = cd.n
        m3 = lambda a: [[a[3 * r + c] for c in range(3)] for r in range(3)]
        chain = dict(n=n, joints=[], bodies=[], links={}, world_xyz=list(cd.world_xyz), world_rot=m3(cd.world_rot), gravity=list(cd.gravity))
        for k in range(n):
            chain["joints"].append(dict(name=L.osc_urdf_joint_name(h, k).decode(), type="prismatic" if cd.joint_type[k] == abi.JOINT_PRISMATIC else "revolute",
                                        axis=list(cd.joint_axis[k]), xyz=list(cd.joint_xyz[k]), rot=m3(cd.joint_rot[k]), lower=cd.q_lower[k],
                                        upper=cd.q_upper[k], effort=cd.effort[k], q0=q0[k]))
            chain["bodies"].append(dict(mass=cd.mass[k], com=list(cd.com[k]), inertia=m3(cd.inertia[k])))
        for k in range(L.osc_urdf_link_count(h)):
            name = L.osc_urdf_link_name(h, k)
            body = C.c_int(); xyz = (C.c_double * 3)(); rot = (C.c_double * 9)()
            L.osc_urdf_link(h, name, C.byref(body), xyz, rot)
            chain["links"][name.decode()] = dict(body=body.value, xyz=list(xyz), rot=m3(rot))
        return chain
    finally:
        L.osc_urdf_free(h)


def measure_fp64_peak(device=0):
    """sustained DFMA throughput in TFLOP/s (2 flop per FMA) measured by the library's microbenchmark"""
    t, ms = C.c_double(), C.c_double()
    rc = lib().osc_measure_fp64_peak(device, C.byref(t), C.byref(ms))
    if rc < 0:
        raise OscError(rc, lib().osc_last_error().decode())
    return t.value


class FieldIn(C.Structure):  # osc_field_in
    _fields_ = [("task", C.c_int32), ("field", C.c_int32), ("data", C.c_void_p), ("layout", C.c_int32)]


class Batch:
    """B robots on one GPU.  Method names follow the C ABI; host numpy arrays are AOS [B, k]."""

    def __init__(self, chain: dict, n_robots: int, precision=abi.PRECISION_F64, device=0, q0=None, devices=None):
        """devices = a list of CUDA device indices: the batch is sharded over them (osc_batch_create_multi), one host thread and
        one stream per device inside the library; every method below works unchanged on such a batch (HOST buffers only)."""
        self.L = lib()
        self.chain = chain
        self.n = chain["n"]
        self.B = int(n_robots)
        self.precision = precision
        self.dtype = np.float64 if precision == abi.PRECISION_F64 else np.float32
        self._cd = abi.chain_desc(chain)
        h = C.c_void_p()
        q0a = np.ascontiguousarray(q0, dtype=np.float64) if q0 is not None else None
        q0p = q0a.ctypes.data_as(C.POINTER(C.c_double)) if q0a is not None else None
        if devices is not None:
            ids = (C.c_int * len(devices))(*[int(d) for d in devices])
            rc = self.L.osc_batch_create_multi(C.byref(self._cd), self.B, precision, ids, len(devices), q0p, C.byref(h))
        else:
            rc = self.L.osc_batch_create(C.byref(self._cd), self.B, precision, device, q0p, C.byref(h))
        self.h = None
        self._chk(rc)
        self.h = h

    def shards(self):
        """[(device, first_robot, n_robots)] of a (multi-device) batch"""
        out = []
        for k in range(self.L.osc_batch_shard_count(self.h)):
            d, f, n = C.c_int(), C.c_int64(), C.c_int64()
            self.L.osc_batch_shard(self.h, k, C.byref(d), C.byref(f), C.byref(n))
            out.append((d.value, f.value, n.value))
        return out

    def close(self):
        if getattr(self, "h", None):
            self.L.osc_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def _chk(self, rc):
        if rc < 0:
            msg = self.L.osc_last_error().decode()
            raise (OscInvalidArgument if rc == abi.ERR_INVALID_ARGUMENT else OscError)(rc, msg)
        return rc

    # ---- hierarchy ----
    def add_motion_force_task(self, desc):
        return self._chk(self.L.osc_batch_add_motion_force_task(self.h, C.byref(desc)))

    def add_joint_task(self, desc):
        return self._chk(self.L.osc_batch_add_joint_task(self.h, C.byref(desc)))

    def task_dof(self, task):
        return self._chk(self.L.osc_batch_task_dof(self.h, task))

    def mft_params(self, task):
        p = abi.MftParams()
        self._chk(self.L.osc_mft_get_params(self.h, task, C.byref(p)))
        return p

    def set_mft_params(self, task, p):
        return self._chk(self.L.osc_mft_set_params(self.h, task, C.byref(p)))

    def joint_task_params(self, task):
        p = abi.JointTaskParams()
        self._chk(self.L.osc_joint_task_get_params(self.h, task, C.byref(p)))
        return p

    def set_joint_task_params(self, task, p):
        return self._chk(self.L.osc_joint_task_set_params(self.h, task, C.byref(p)))

    def set_gains_unsafe(self, task, kp, kv, ki):
        """JointTask::setGainsUnsafe: no sign check; scalars or length-k vectors"""
        a = [np.ascontiguousarray(np.atleast_1d(x), dtype=np.float64) for x in (kp, kv, ki)]
        self._chk(self.L.osc_joint_task_set_gains_unsafe(self.h, task, a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, len(a[0])))

    def selection_projectors(self, task):
        """posSelectionProjector(), oriSelectionProjector()"""
        p, o = np.zeros(9), np.zeros(9)
        self._chk(self.L.osc_mft_selection_projectors(self.h, task, p.ctypes.data, o.ctypes.data))
        return p.reshape(3, 3), o.reshape(3, 3)

    def goal_reached(self, task, tolerance, what=0):
        """goalPositionReached (what = 0) / goalOrientationReached (what = 1): per-robot flags [B] and their count"""
        flags = np.zeros(self.B, dtype=np.int32)
        n = C.c_int64()
        self._chk(self.L.osc_task_goal_reached(self.h, task, what, float(tolerance), flags.ctypes.data, abi.MEM_HOST, C.byref(n)))
        return flags, n.value

    def enable_gravity_compensation(self, e=True):
        self._chk(self.L.osc_enable_gravity_compensation(self.h, int(e)))

    # ---- fields ----
    def field_size(self, task, field):
        return self._chk(self.L.osc_field_size(self.h, task, field))

    def _dt(self, field):
        return np.int32 if field in abi.INT_FIELDS else self.dtype

    def get(self, field, task=-1):
        w = self.field_size(task, field)
        out = np.empty((self.B, w), dtype=self._dt(field))
        self._chk(self.L.osc_get_field(self.h, task, field, out.ctypes.data, abi.LAYOUT_AOS, abi.MEM_HOST))
        return out

    def set(self, field, data, task=-1):
        w = self.field_size(task, field)
        a = np.ascontiguousarray(data, dtype=self._dt(field))
        if a.size == w and self.B != 1:
            self._chk(self.L.osc_set_field(self.h, task, field, a.ctypes.data, abi.LAYOUT_UNIFORM, abi.MEM_HOST))
        else:
            a = np.ascontiguousarray(a.reshape(self.B, w))
            self._chk(self.L.osc_set_field(self.h, task, field, a.ctypes.data, abi.LAYOUT_AOS, abi.MEM_HOST))

    def set_raw(self, field, ptr, layout, mem, task=-1):
        self._chk(self.L.osc_set_field(self.h, task, field, ptr, layout, mem))

    def get_raw(self, field, ptr, layout, mem, task=-1):
        self._chk(self.L.osc_get_field(self.h, task, field, ptr, layout, mem))

    def update_sensed_force_and_moment(self, task, f, m):
        f = np.ascontiguousarray(f, dtype=self.dtype)
        m = np.ascontiguousarray(m, dtype=self.dtype)
        layout = abi.LAYOUT_UNIFORM if (f.size == 3 and self.B != 1) else abi.LAYOUT_AOS
        self._chk(self.L.osc_mft_update_sensed_force_and_moment(self.h, task, f.ctypes.data, m.ctypes.data, layout, abi.MEM_HOST))

    # ---- tick ----
    def update_model(self):
        self._chk(self.L.osc_update_model(self.h))

    def update_kinematics(self):
        self._chk(self.L.osc_update_kinematics(self.h))

    def task_update_model(self, task, n_prec_task=-1):
        self._chk(self.L.osc_task_update_model(self.h, task, n_prec_task))

    def update_task_models(self):
        self._chk(self.L.osc_update_task_models(self.h))

    def task_compute_torques(self, task, with_tau_prec=False):
        self._chk(self.L.osc_task_compute_torques(self.h, task, int(with_tau_prec)))

    def compute_torques(self, stacked=True):
        out = np.empty((self.B, self.n), dtype=self.dtype)
        self._chk(self.L.osc_compute_torques(self.h, int(stacked), out.ctypes.data, abi.LAYOUT_AOS, abi.MEM_HOST))
        return out

    def tick(self, stacked=True, fetch=True):
        if not fetch:
            self._chk(self.L.osc_tick(self.h, int(stacked), None, 0, 0))
            return None
        out = np.empty((self.B, self.n), dtype=self.dtype)
        self._chk(self.L.osc_tick(self.h, int(stacked), out.ctypes.data, abi.LAYOUT_AOS, abi.MEM_HOST))
        return out

    def step_host(self, inputs, tau_ptr, stacked=True, tau_layout=abi.LAYOUT_AOS):
        """osc_step_host: inputs = [(task, field, host pointer, layout), ...]; tau_ptr = host pointer or 0.  One call per control
        cycle; on a batch sharded with devices=[...] (also the same device several times) the shards' copies and ticks overlap."""
        arr = (FieldIn * len(inputs))()
        for k, (task, field, ptr, layout) in enumerate(inputs):
            arr[k].task, arr[k].field, arr[k].data, arr[k].layout = int(task), int(field), C.c_void_p(ptr), int(layout)
        self._chk(self.L.osc_step_host(self.h, arr, len(inputs), int(stacked), C.c_void_p(tau_ptr) if tau_ptr else None, tau_layout))

    def tick_noalloc(self, tau_device_ptr, stacked=True, layout=abi.LAYOUT_AOS):
        """osc_tick_noalloc: fused tick with the library's static device scratch; tau lands in DEVICE memory (or nowhere)"""
        self._chk(self.L.osc_tick_noalloc(self.h, int(stacked), C.c_void_p(tau_device_ptr) if tau_device_ptr else None, layout))

    def tick_is_fused(self):
        return bool(self.L.osc_tick_is_fused(self.h))

    def set_tick_outputs(self, outputs):
        self._chk(self.L.osc_batch_set_tick_outputs(self.h, int(outputs)))

    def tick_fallback_count(self):
        return int(self.L.osc_tick_fallback_count(self.h))

    def launch_count(self):
        return int(self.L.osc_launch_count(self.h))

    def reinitialize_tasks(self, task=-1):
        self._chk(self.L.osc_reinitialize_tasks(self.h, task))

    def enable_internal_otg(self, task, max_velocity, max_acceleration):
        """enableInternalOtgAccelerationLimited.  JointTask: scalars or length-k vectors; MotionForceTask: (lin, ang) pairs."""
        mv = np.atleast_1d(np.asarray(max_velocity, dtype=np.float64))
        ma = np.atleast_1d(np.asarray(max_acceleration, dtype=np.float64))
        lim = np.ascontiguousarray(np.concatenate([mv, ma]))
        self._chk(self.L.osc_task_enable_internal_otg_acceleration_limited(self.h, task, lim.ctypes.data, len(mv)))

    def disable_internal_otg(self, task):
        self._chk(self.L.osc_task_disable_internal_otg(self.h, task))

    def internal_otg_enabled(self, task):
        return bool(self.L.osc_task_internal_otg_enabled(self.h, task))

    def rollout_step(self, dt, substeps=1, joint_damping=0.0):
        """sim->integrate() stand-in: forward dynamics under the current OSC_F_TAU + symplectic Euler, on the device"""
        self._chk(self.L.osc_rollout_step(self.h, dt, substeps, joint_damping))

    def rollout(self, ticks, sim_dt, substeps=1, joint_damping=0.0, stacked=True):
        tau = np.empty((self.B, self.n), dtype=self.dtype)
        self._chk(self.L.osc_rollout(self.h, int(stacked), ticks, sim_dt, substeps, joint_damping, tau.ctypes.data, abi.LAYOUT_AOS, abi.MEM_HOST))
        return tau

    def synchronize(self):
        self._chk(self.L.osc_synchronize(self.h))

    def set_stream(self, cuda_stream_ptr):
        self._chk(self.L.osc_batch_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def snapshot(self):
        nbytes = int(self.L.osc_snapshot_bytes(self.h))
        buf = np.empty(nbytes, dtype=np.uint8)
        self._chk(self.L.osc_snapshot(self.h, buf.ctypes.data, abi.MEM_HOST))
        return buf

    def restore(self, buf):
        self._chk(self.L.osc_restore(self.h, buf.ctypes.data, abi.MEM_HOST))
