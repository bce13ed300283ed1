"""ctypes binding of the C++ CPU oracle (oracle/_build/liboscoracle.so).  TEST INFRASTRUCTURE ONLY.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(_HERE), "sai2-primitives-perso_b200"))
import osc_abi as abi  # noqa: E402

_LIBS = {}
VARIANTS = ("", "O3", "heap", "count")  # g++ -O2 (default) | -O3 -march=native | -DORACLE_HEAP_TEMPORARIES | -DORACLE_COUNT_FLOPS


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib(variant=""):
    """variant: "" = g++ -O2 (the checker and the default CPU baseline), "O3" = -O3 -march=native, "heap" = malloc'ed matrix
    temporaries like Eigen's dynamic MatrixXd, "count" = operation-counting build (oracle/Makefile)"""
    if variant not in _LIBS:
        path = os.path.join(_HERE, "_build", "liboscoracle%s.so" % ("_" + variant if variant else ""))
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.POINTER(abi.ChainDesc), C.c_int64, C.POINTER(C.c_double)]
        L.orc_last_error.restype = C.c_char_p
        for name in ("orc_destroy", "orc_add_motion_force_task", "orc_add_joint_task", "orc_mft_get_params",
                     "orc_mft_set_params", "orc_joint_task_get_params", "orc_joint_task_set_params",
                     "orc_enable_gravity_compensation", "orc_task_dof", "orc_field_size", "orc_get_field",
                     "orc_set_field", "orc_mft_update_sensed_force_and_moment", "orc_update_model",
                     "orc_update_kinematics", "orc_task_update_model", "orc_update_task_models",
                     "orc_task_compute_torques", "orc_compute_torques", "orc_tick", "orc_reinitialize_tasks"):
            getattr(L, name).restype = C.c_int
        L.orc_destroy.restype = None
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_add_motion_force_task.argtypes = [C.c_void_p, C.POINTER(abi.MftDesc)]
        L.orc_add_joint_task.argtypes = [C.c_void_p, C.POINTER(abi.JointTaskDesc)]
        L.orc_mft_get_params.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.MftParams)]
        L.orc_mft_set_params.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.MftParams)]
        L.orc_joint_task_get_params.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.JointTaskParams)]
        L.orc_joint_task_set_params.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.JointTaskParams)]
        L.orc_enable_gravity_compensation.argtypes = [C.c_void_p, C.c_int]
        L.orc_task_dof.argtypes = [C.c_void_p, C.c_int]
        L.orc_field_size.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_get_field.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_set_field.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_mft_update_sensed_force_and_moment.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_update_model.argtypes = [C.c_void_p, C.c_int]
        L.orc_update_kinematics.argtypes = [C.c_void_p, C.c_int]
        L.orc_task_update_model.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_update_task_models.argtypes = [C.c_void_p, C.c_int]
        L.orc_task_compute_torques.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_compute_torques.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_tick.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_reinitialize_tasks.argtypes = [C.c_void_p, C.c_int]
        L.orc_hardware_threads.restype = C.c_int
        L.orc_rollout_step.restype = C.c_int
        L.orc_rollout_step.argtypes = [C.c_void_p, C.c_double, C.c_int, C.c_double, C.c_int]
        L.orc_nonlinear_effects.restype = C.c_int
        L.orc_nonlinear_effects.argtypes = [C.c_void_p, C.c_void_p]
        # internal OTG (otg_oracle.hpp)
        L.orc_otg_use_reference.restype = C.c_int
        L.orc_otg_use_reference.argtypes = [C.c_char_p]
        L.orc_task_enable_internal_otg_acceleration_limited.restype = C.c_int
        L.orc_task_enable_internal_otg_acceleration_limited.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.orc_task_disable_internal_otg.restype = C.c_int
        L.orc_task_disable_internal_otg.argtypes = [C.c_void_p, C.c_int]
        for name in ("orc_rk_create", "orc_otgj_create", "orc_otgc_create"):
            getattr(L, name).restype = C.c_void_p
        L.orc_rk_create.argtypes = [C.c_int, C.c_double]
        L.orc_otgj_create.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.c_void_p]
        L.orc_otgc_create.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double]
        for name in ("orc_rk_destroy", "orc_otgj_destroy", "orc_otgc_destroy"):
            getattr(L, name).restype = None
            getattr(L, name).argtypes = [C.c_void_p]
        L.orc_rk_update.restype = C.c_int
        L.orc_rk_update.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_otgj_step.restype = C.c_int
        L.orc_otgj_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_otgc_step.restype = C.c_int
        L.orc_otgc_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_count_tick_flops.restype = C.c_longlong
        L.orc_count_tick_flops.argtypes = [C.c_void_p, C.c_int]
        L.orc_config1_run.restype = C.c_int
        L.orc_config1_run.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _LIBS[variant] = L
    return _LIBS[variant]


REF_RUCKIG = os.path.join(_HERE, "_ref", "libruckig_ref.so")


def have_reference_ruckig():
    """oracle/_ref/libruckig_ref.so = the UNMODIFIED Ruckig of /root/reference/ruckig behind oracle/ruckig_ref_capi.cpp"""
    return os.path.exists(REF_RUCKIG)


def otg_use_reference(on=True):
    """Every OTG object created afterwards runs on the real Ruckig (on=True) or on the restatement (on=False)."""
    rc = lib().orc_otg_use_reference(REF_RUCKIG.encode() if on else None)
    if rc != 0:
        raise RuntimeError(lib().orc_last_error().decode())


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class Ruckig2:
    """One Ruckig instance (second order, phase synchronisation): restatement or reference backend."""

    def __init__(self, n, dt):
        self.L, self.n = lib(), n
        self.h = self.L.orc_rk_create(n, dt)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_rk_destroy(self.h)
            self.h = None

    def update(self, cur_p, cur_v, cur_a, tgt_p, tgt_v, max_v, max_a):
        io = _f64(np.concatenate([_f64(x).ravel() for x in (cur_p, cur_v, cur_a, tgt_p, tgt_v, max_v, max_a)]))
        out = np.full(3 * self.n, np.nan)
        info = np.zeros(3)
        r = self.L.orc_rk_update(self.h, io.ctypes.data, out.ctypes.data, info.ctypes.data)
        return r, out.reshape(3, self.n), info


class OtgJoints:
    """src/helper_modules/OTG_joints.cpp, acceleration limited"""

    def __init__(self, p0, dt, max_v, max_a):
        self.L, self.n = lib(), len(p0)
        p0, mv, ma = _f64(p0), _f64(np.broadcast_to(max_v, (self.n,))), _f64(np.broadcast_to(max_a, (self.n,)))
        self.h = self.L.orc_otgj_create(p0.ctypes.data, self.n, dt, mv.ctypes.data, ma.ctypes.data)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_otgj_destroy(self.h)
            self.h = None

    def step(self, goal_p, goal_v=None):
        gp = _f64(goal_p)
        gv = _f64(goal_v) if goal_v is not None else np.zeros(self.n)
        out = np.zeros(3 * self.n)
        fl = np.zeros(1, dtype=np.int32)
        r = self.L.orc_otgj_step(self.h, gp.ctypes.data, gv.ctypes.data, out.ctypes.data, fl.ctypes.data)
        return r, out.reshape(3, self.n), int(fl[0])


class OtgCartesian:
    """src/helper_modules/OTG_6dof_cartesian.cpp, acceleration limited"""

    def __init__(self, p0, R0, dt, lin_v, lin_a, ang_v, ang_a):
        self.L = lib()
        p0, R0 = _f64(p0), _f64(R0)
        self.h = self.L.orc_otgc_create(p0.ctypes.data, R0.ctypes.data, dt, lin_v, lin_a, ang_v, ang_a)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_otgc_destroy(self.h)
            self.h = None

    def step(self, gp, gR, gv=None, gw=None):
        gp, gR = _f64(gp), _f64(gR)
        gv = _f64(gv) if gv is not None else np.zeros(3)
        gw = _f64(gw) if gw is not None else np.zeros(3)
        out = np.zeros(24)
        fl = np.zeros(1, dtype=np.int32)
        r = self.L.orc_otgc_step(self.h, gp.ctypes.data, gR.ctypes.data, gv.ctypes.data, gw.ctypes.data, out.ctypes.data, fl.ctypes.data)
        return r, out, int(fl[0])


class OracleBatch:
    """B independent reference-style robots on the CPU.  Mirrors the method names of osc_b200.Batch."""

    def __init__(self, chain: dict, n_robots: int, q0=None, nthreads: int = 1, variant: str = ""):
        self.L = lib(variant)
        self.chain = chain
        self.n = chain["n"]
        self.B = int(n_robots)
        self.nthreads = nthreads
        self._cd = abi.chain_desc(chain)
        q0a = np.ascontiguousarray(q0, dtype=np.float64) if q0 is not None else None
        self.h = self.L.orc_create(C.byref(self._cd), self.B, q0a.ctypes.data_as(C.POINTER(C.c_double)) if q0a is not None else None)
        self.task_types = []

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_destroy(self.h)
            self.h = None

    def _chk(self, rc):
        if rc < 0:
            raise ValueError(self.L.orc_last_error().decode())
        return rc

    def add_motion_force_task(self, desc):
        self.task_types.append(abi.TASK_MOTION_FORCE)
        return self._chk(self.L.orc_add_motion_force_task(self.h, C.byref(desc)))

    def add_joint_task(self, desc):
        self.task_types.append(abi.TASK_JOINT)
        return self._chk(self.L.orc_add_joint_task(self.h, C.byref(desc)))

    def mft_params(self, task):
        p = abi.MftParams()
        self.L.orc_mft_get_params(self.h, task, C.byref(p))
        return p

    def set_mft_params(self, task, p):
        return self._chk(self.L.orc_mft_set_params(self.h, task, C.byref(p)))

    def joint_task_params(self, task):
        p = abi.JointTaskParams()
        self.L.orc_joint_task_get_params(self.h, task, C.byref(p))
        return p

    def set_joint_task_params(self, task, p):
        return self._chk(self.L.orc_joint_task_set_params(self.h, task, C.byref(p)))

    def enable_gravity_compensation(self, e=True):
        self.L.orc_enable_gravity_compensation(self.h, int(e))

    def task_dof(self, task):
        return self.L.orc_task_dof(self.h, task)

    def field_size(self, task, field):
        return self._chk(self.L.orc_field_size(self.h, task, field))

    def get(self, field, task=-1):
        w = self.field_size(task, field)
        out = np.empty((self.B, w), dtype=np.float64)
        self._chk(self.L.orc_get_field(self.h, task, field, out.ctypes.data))
        return out

    def set(self, field, data, task=-1):
        w = self.field_size(task, field)
        a = np.ascontiguousarray(data, dtype=np.float64)
        uniform = a.size == w and self.B != 1
        if not uniform:
            a = np.ascontiguousarray(a.reshape(self.B, w))
        self._chk(self.L.orc_set_field(self.h, task, field, a.ctypes.data, int(uniform)))

    def update_sensed_force_and_moment(self, task, f, m):
        f = np.ascontiguousarray(f, dtype=np.float64)
        m = np.ascontiguousarray(m, dtype=np.float64)
        uniform = f.size == 3 and self.B != 1
        self.L.orc_mft_update_sensed_force_and_moment(self.h, task, f.ctypes.data, m.ctypes.data, int(uniform))

    def update_model(self):
        self.L.orc_update_model(self.h, self.nthreads)

    def update_kinematics(self):
        self.L.orc_update_kinematics(self.h, self.nthreads)

    def task_update_model(self, task, n_prec_task=-1):
        self.L.orc_task_update_model(self.h, task, n_prec_task, self.nthreads)

    def update_task_models(self):
        self.L.orc_update_task_models(self.h, self.nthreads)

    def task_compute_torques(self, task, with_tau_prec=False):
        self.L.orc_task_compute_torques(self.h, task, int(with_tau_prec), self.nthreads)

    def compute_torques(self, stacked=True):
        self.L.orc_compute_torques(self.h, int(stacked), self.nthreads)
        return self.get(abi.F_TAU)

    def tick(self, stacked=True):
        self.L.orc_tick(self.h, int(stacked), self.nthreads)
        return self.get(abi.F_TAU)

    def reinitialize_tasks(self, task=-1):
        self.L.orc_reinitialize_tasks(self.h, task)

    def rollout_step(self, dt, substeps=1, joint_damping=0.0):
        self.L.orc_rollout_step(self.h, dt, substeps, joint_damping, self.nthreads)

    def nonlinear_effects(self):
        """b(q, dq) + g(q) at the current q, dq (call update_kinematics / update_model first)"""
        out = np.empty((self.B, self.n), dtype=np.float64)
        self.L.orc_nonlinear_effects(self.h, out.ctypes.data)
        return out

    def enable_internal_otg(self, task, max_velocity, max_acceleration):
        """JointTask: scalars or length-k vectors.  MotionForceTask: (lin, ang) pairs."""
        mv, ma = np.atleast_1d(_f64(max_velocity)), np.atleast_1d(_f64(max_acceleration))
        lim = _f64(np.concatenate([mv, ma]))
        self._chk(self.L.orc_task_enable_internal_otg_acceleration_limited(self.h, task, lim.ctypes.data, len(mv)))

    def disable_internal_otg(self, task):
        self.L.orc_task_disable_internal_otg(self.h, task)

    def count_tick_flops(self, stacked=True):
        """FLOP of one tick of robot 0, counted by the primitives of the oracle (variant "count" only; -1 otherwise)"""
        return int(self.L.orc_count_tick_flops(self.h, int(stacked)))

    def config1_run(self, ticks):
        """BASELINE config 1 (04-task_and_redundancy loop, robot 0): per-cycle wall time in microseconds"""
        out = np.zeros(ticks)
        self._chk(self.L.orc_config1_run(self.h, ticks, out.ctypes.data))
        return out
