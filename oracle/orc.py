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

_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "_build", "liboscoracle.so")
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
        _LIB = L
    return _LIB


class OracleBatch:
    """B independent reference-style robots on the CPU.  Mirrors the method names of osc_b200.Batch."""

    def __init__(self, chain: dict, n_robots: int, q0=None, nthreads: int = 1):
        self.L = lib()
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
