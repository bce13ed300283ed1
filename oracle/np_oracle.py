"""numpy cross-check oracle for the OSC control tick (TEST INFRASTRUCTURE ONLY).

PARITY UNPINNED: the reference (sai2-primitives) has no tests or golden vectors for this path and its
numerical core (sai2-model / RBDL / Eigen) is not available on this machine.  This file is an
independent restatement, on LAPACK (numpy.linalg), of the reference's algorithm; it exists to cross-check the
C++ oracle (oracle/osc_oracle.cpp) and to generate tests/golden/*.npz.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline leg may import it; the product path never does.

Each class cites the reference file:line it follows (paths relative to /root/reference/src).
Un-vendored sai2-model functions follow SURVEY.md Appendix B.
"""
from __future__ import annotations

import math
from collections import deque

import numpy as np

FULL_DYNAMIC_DECOUPLING, BOUNDED_INERTIA_ESTIMATES, IMPEDANCE = 0, 1, 2
NO_SINGULARITY, TYPE_1_SINGULARITY, TYPE_2_SINGULARITY = 0, 1, 2


def skew(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])


def rot_axis(axis, q):
    """Rodrigues rotation about a unit axis."""
    K = skew(np.asarray(axis, float))
    return np.eye(3) + math.sin(q) * K + (1 - math.cos(q)) * (K @ K)


# ----------------------------------------------------------------------------------------------------------
# sai2-model free functions
# ----------------------------------------------------------------------------------------------------------
def orientation_error(R_des, R_cur):
    """Sai2Model::orientationError: dphi = -1/2 sum_i R_cur[:,i] x R_des[:,i] (called MotionForceTask.cpp:292,443)."""
    return -0.5 * sum(np.cross(R_cur[:, i], R_des[:, i]) for i in range(3))


def svd_canonical(A):
    """Thin SVD with the sign convention both oracles and the kernels share:
    the largest-|component| entry of every V column is positive."""
    U, s, Vt = np.linalg.svd(A, full_matrices=False)
    V = Vt.T.copy()
    U = U.copy()
    for c in range(V.shape[1]):
        k = int(np.argmax(np.abs(V[:, c])))
        if V[k, c] < 0:
            V[:, c] *= -1
            U[:, c] *= -1
    return U, s, V


def matrix_range_basis(A, tol=1e-3):
    """Sai2Model::matrixRangeBasis (called JointTask.cpp:233, MotionForceTask.cpp:66,78,147,149)."""
    A = np.atleast_2d(np.asarray(A, float))
    rows = A.shape[0]
    max_range = min(A.shape)
    if np.linalg.norm(A) < tol:
        return np.zeros((rows, 1))
    U, s, _ = np.linalg.svd(A, full_matrices=False)
    if s[0] < tol:
        return np.zeros((rows, 1))
    rank = max_range
    for i in range(len(s) - 1, 0, -1):
        if s[i] / s[0] < tol:
            rank -= 1
        else:
            break
    if rank == rows:
        return np.eye(rows)
    return U[:, :rank]


def pseudo_inverse_diag(d, tol=1e-6):
    """Sai2Model::computePseudoInverse restricted to the diagonal gain matrices it is called on
    (MotionForceTask.cpp:417,450; JointTask.cpp:328): COD threshold relative to the largest pivot."""
    d = np.asarray(d, float)
    mx = np.max(np.abs(d)) if d.size else 0.0
    out = np.zeros_like(d)
    keep = np.abs(d) > tol * mx
    out[keep] = 1.0 / d[keep]
    return out


# ----------------------------------------------------------------------------------------------------------
# Robot model (stands in for Sai2Model::Sai2Model over RBDL)
# ----------------------------------------------------------------------------------------------------------
class Model:
    def __init__(self, chain):
        self.chain = chain
        self.n = chain["n"]
        J = chain["joints"]
        self.jtype = [j["type"] for j in J]
        self.axis = [np.array(j["axis"], float) for j in J]
        self.xyz = [np.array(j["xyz"], float) for j in J]
        self.rot = [np.array(j["rot"], float) for j in J]
        self.mass = [b["mass"] for b in chain["bodies"]]
        self.com = [np.array(b["com"], float) for b in chain["bodies"]]
        self.inertia = [np.array(b["inertia"], float) for b in chain["bodies"]]
        self.q_lower = np.array([j["lower"] for j in J])
        self.q_upper = np.array([j["upper"] for j in J])
        self.effort = np.array([j["effort"] for j in J])
        self.Rw = np.array(chain["world_rot"], float)
        self.pw = np.array(chain["world_xyz"], float)
        self.gravity = np.array(chain["gravity"], float)
        self._q = np.array([j["q0"] for j in J], float)
        self._dq = np.zeros(self.n)
        self.updateModel()

    def dof(self):
        return self.n

    def q(self):
        return self._q.copy()

    def dq(self):
        return self._dq.copy()

    def setQ(self, q):
        self._q = np.array(q, float)

    def setDq(self, dq):
        self._dq = np.array(dq, float)

    def link(self, name):
        L = self.chain["links"][name]
        return L["body"], np.array(L["xyz"], float), np.array(L["rot"], float)

    def updateKinematics(self):
        """Base-frame poses of every movable body (R_i, o_i) and world axis directions z_i."""
        R, p = np.eye(3), np.zeros(3)
        self.R, self.o, self.z = [], [], []
        for i in range(self.n):
            p = p + R @ self.xyz[i]
            R = R @ self.rot[i]
            if self.jtype[i] == "revolute":
                R = R @ rot_axis(self.axis[i], self._q[i])
            else:
                p = p + R @ self.axis[i] * self._q[i]
            self.R.append(R.copy())
            self.o.append(p.copy())
            self.z.append(R @ self.axis[i])

    def updateModel(self):
        """updateKinematics + M (here: sum of per-body kinetic-energy Hessians, an independent route from the
        composite-rigid-body recursions of the C++ oracle and the kernels) + dense inverse."""
        self.updateKinematics()
        n = self.n
        M = np.zeros((n, n))
        for b in range(n):
            if self.mass[b] == 0 and not np.any(self.inertia[b]):
                continue
            c = self.o[b] + self.R[b] @ self.com[b]
            Jb = self._jacobian_base(b, c)
            Iw = self.R[b] @ self.inertia[b] @ self.R[b].T
            M += self.mass[b] * Jb[:3].T @ Jb[:3] + Jb[3:].T @ Iw @ Jb[3:]
        self._M = M
        self._Minv = np.linalg.inv(M)

    def M(self):
        return self._M.copy()

    def MInv(self):
        return self._Minv.copy()

    def _jacobian_base(self, body, p_base):
        J = np.zeros((6, self.n))
        for i in range(body + 1):
            if self.jtype[i] == "revolute":
                J[:3, i] = np.cross(self.z[i], p_base - self.o[i])
                J[3:, i] = self.z[i]
            else:
                J[:3, i] = self.z[i]
        return J

    # --- frame queries: (link name, point/rotation in link) ---
    def position(self, link, p_in_link=np.zeros(3)):
        b, lp, lR = self.link(link)
        pb = lp + lR @ p_in_link
        return (self.o[b] + self.R[b] @ pb) if b >= 0 else pb

    def rotation(self, link, R_in_link=np.eye(3)):
        b, lp, lR = self.link(link)
        Rb = lR @ R_in_link
        return (self.R[b] @ Rb) if b >= 0 else Rb

    def positionInWorld(self, link, p_in_link=np.zeros(3)):
        return self.pw + self.Rw @ self.position(link, p_in_link)

    def rotationInWorld(self, link, R_in_link=np.eye(3)):
        return self.Rw @ self.rotation(link, R_in_link)

    def JWorldFrame(self, link, p_in_link=np.zeros(3)):
        """6 x n, linear rows first, world frame (MotionForceTask.cpp:262,293-298)."""
        b, _, _ = self.link(link)
        J = self._jacobian_base(b, self.position(link, p_in_link))
        return np.vstack([self.Rw @ J[:3], self.Rw @ J[3:]])

    def operationalSpaceMatrices(self, J):
        """Lambda via a Cholesky solve against identity, Jbar, N."""
        A = J @ self._Minv @ J.T
        L = np.linalg.cholesky(A)
        Li = np.linalg.inv(L)
        Lam = Li.T @ Li
        Jbar = self._Minv @ J.T @ Lam
        N = np.eye(self.n) - Jbar @ J
        return Lam, Jbar, N

    def jointGravityVector(self):
        g = np.zeros(self.n)
        gw = self.Rw.T @ self.gravity
        for b in range(self.n):
            c = self.o[b] + self.R[b] @ self.com[b]
            Jb = self._jacobian_base(b, c)
            g -= Jb[:3].T @ (self.mass[b] * gw)
        return g


# ----------------------------------------------------------------------------------------------------------
# POPCExplicitForceControl  (helper_modules/POPCExplicitForceControl.cpp:5-96)
# ----------------------------------------------------------------------------------------------------------
class POPC:
    WINDOW, MAX_COUNTER = 250, 50  # POPCExplicitForceControl.h:37-38

    def __init__(self, dt):
        self.dt = dt
        self.enabled = False  # .cpp:6
        self.reInitialize()

    def reInitialize(self):  # .cpp:10-21
        self.PO = 0.0
        self.E_corr = 0.0
        self.E_stored = 0.0
        self.window = deque()
        self.counter = self.MAX_COUNTER
        self.Rc = 1.0
        self.vcl_sq_sum = 0.0

    def enable(self):
        self.enabled = True

    def disable(self):
        self.enabled = False
        self.reInitialize()

    def compute(self, fd, fs, vcl, vr, kv_force, kff):  # .cpp:30-96, kv_force is the diagonal (3,)
        if not self.enabled:
            return vcl - kv_force * vr
        F_cmd = kff * fd + self.Rc * vcl - kv_force * vr
        vc_sq = vcl @ vcl
        p = ((fs - fd) @ vcl - F_cmd @ vr) * self.dt
        self.PO += p
        self.window.append(p)
        if self.PO + self.E_stored + self.E_corr > 0:
            while len(self.window) > self.WINDOW:
                if self.PO + self.E_corr + self.E_stored > self.window[0]:
                    if self.window[0] > 0:
                        self.PO -= self.window[0]
                    self.window.popleft()
                else:
                    break
        if self.counter <= 0:
            self.counter = self.MAX_COUNTER
            old = self.Rc
            if self.PO + self.E_stored + self.E_corr < 0:
                with np.errstate(divide="ignore", invalid="ignore"):
                    self.Rc = 1 + np.float64(self.PO + self.E_stored + self.E_corr) / np.float64(self.vcl_sq_sum * self.dt)
                if self.Rc > 1:
                    self.Rc = 1.0
                if self.Rc < 0:
                    self.Rc = 0.0
            else:
                self.Rc = (1 + (0.1 * self.MAX_COUNTER - 1) * self.Rc) / (0.1 * self.MAX_COUNTER)
            self.E_corr += (1 - old) * self.vcl_sq_sum * self.dt
            self.vcl_sq_sum = 0.0
        self.counter -= 1
        self.vcl_sq_sum += vc_sq
        return self.Rc * vcl - kv_force * vr


# ----------------------------------------------------------------------------------------------------------
# SingularityHandler  (tasks/SingularityHandler.cpp)
# ----------------------------------------------------------------------------------------------------------
class SingularityHandler:
    def __init__(self, robot, link, p_c, R_c, task_rank):  # .cpp:24-73 (+ Q5: intended 1e-2 * effort)
        self.robot, self.link, self.p_c, self.R_c, self.m = robot, link, p_c, R_c, task_rank
        n = robot.dof()
        self.n = n
        self.q_upper, self.q_lower = robot.q_upper.copy(), robot.q_lower.copy()
        self.tau_upper, self.tau_lower = robot.effort.copy(), -robot.effort
        self.s_abs_tol, self.type_1_tol, self.type_2_torque_ratio = 1e-3, 0.5, 1e-2
        self.type_2_angle_threshold, self.perturb_step, self.buffer_size = 5 * math.pi / 180, 5.0, 200
        self.type_2_torque_vector = self.type_2_torque_ratio * robot.effort
        self.types = []
        self.q_prior = 0.5 * (self.q_lower + self.q_upper)
        self.dq_prior = np.zeros(n)
        self.kp1, self.kv1, self.kv2 = 50.0, 14.0, 5.0
        self.decoupling, self.bie_threshold = BOUNDED_INERTIA_ESTIMATES, 0.1
        self.c1 = self.c2 = 0
        self.history = deque()
        self.type_2_direction = np.ones(n)
        self.enforce_type_1, self.enforce_handling = False, True
        self.s_min, self.s_max = 6e-3, 6e-2
        self.alpha = 1.0
        self.N = np.eye(n)
        self.N_ns = np.eye(n)
        self.J_post = np.zeros((1, n))
        self.U_ns = np.zeros((6, 1)); self.U_s = np.zeros((6, 1)); self.V_s = np.zeros((n, 1))
        self.J_ns = np.zeros((1, n)); self.J_s = np.zeros((1, n))
        self.L_ns = self.L_ns_mod = np.zeros((1, 1)); self.L_s = self.L_s_mod = np.zeros((1, 1))
        self.L_js = self.L_js_mod = np.zeros((1, 1))
        self.sv = np.zeros(6)

    def updateTaskModel(self, Jp, N_prec):  # .cpp:75-228
        r, n, m = self.robot, self.n, self.m
        U, s, V = svd_canonical(Jp)
        self.sv = s.copy()
        Minv = r.MInv()
        fully_singular = False
        if s[0] < self.s_abs_tol:
            fully_singular = True
            self.alpha = 0.0
            self.U_ns = np.zeros((m, 1)); self.J_ns = np.zeros((m, n)); self.L_ns = np.zeros((m, m))
            self.U_s = U[:, :m]; self.V_s = V[:, :m]; self.J_s = self.U_s.T @ Jp
            self.L_s = np.linalg.pinv(self.J_s @ Minv @ self.J_s.T)
        else:
            rng = range(1, m) if m > 1 else [1]  # Q4: m == 1 is treated as the non-singular branch
            for i in rng:
                rho = s[i] / s[0] if m > 1 else 1.0
                if m > 1 and rho < self.s_max:
                    self.alpha = min(max((rho - self.s_min) / (self.s_max - self.s_min), 0.0), 1.0)
                    self.U_ns = U[:, :i]; self.J_ns = self.U_ns.T @ Jp
                    self.L_ns, _, self.N_ns = r.operationalSpaceMatrices(self.J_ns)
                    self.U_s = U[:, i:m]; self.V_s = V[:, i:m]; self.J_s = self.U_s.T @ Jp
                    with np.errstate(all="ignore"):
                        self.L_s = np.linalg.inv(self.J_s @ Minv @ self.J_s.T)
                    break
                elif i == m - 1 or m == 1:
                    self.alpha = 1.0
                    self.U_ns = U[:, :m]; self.J_ns = self.U_ns.T @ Jp
                    self.L_ns, _, self.N_ns = r.operationalSpaceMatrices(self.J_ns)
                    self.U_s = np.zeros((m, 1)); self.V_s = np.zeros((n, 1)); self.J_s = np.zeros((m, n))
                    self.L_s = np.zeros((m, m))
        has_s = np.linalg.norm(self.U_s) != 0
        has_ns = np.linalg.norm(self.U_ns) != 0
        posture = False
        if fully_singular:  # Q11: N = N_prec whether or not handling is enabled
            self.N = N_prec.copy(); self.L_js = np.zeros((1, 1))
        elif (not has_s) or (not self.enforce_handling):
            self.N = self.N_ns; self.L_js = np.zeros((1, 1))
        else:
            self.J_post = self.V_s.T @ self.N_ns @ N_prec
            self.L_js, _, N_post = r.operationalSpaceMatrices(self.J_post)
            self.N = N_post @ self.N_ns
            posture = True
        if self.decoupling == FULL_DYNAMIC_DECOUPLING:
            self.L_ns_mod, self.L_s_mod, self.L_js_mod = self.L_ns, self.L_s, self.L_js
        elif self.decoupling == IMPEDANCE:
            self.L_ns_mod = np.eye(self.U_ns.shape[1]); self.L_s_mod = np.eye(self.U_s.shape[1])
            self.L_js_mod = np.eye(self.V_s.shape[1])
        else:
            M_BIE = r.M()
            for i in range(n):
                if M_BIE[i, i] < self.bie_threshold:
                    M_BIE[i, i] = self.bie_threshold
            Mi = np.linalg.inv(M_BIE)
            self.L_ns_mod = np.linalg.inv(self.J_ns @ Mi @ self.J_ns.T) if has_ns else self.L_ns
            with np.errstate(all="ignore"):
                try:
                    self.L_s_mod = np.linalg.inv(self.J_s @ Mi @ self.J_s.T) if has_s else self.L_s
                except np.linalg.LinAlgError:
                    self.L_s_mod = np.full_like(self.L_s, np.nan)
            # Q6: the posture Jacobian only exists when it was computed this tick
            self.L_js_mod = np.linalg.inv(self.J_post @ Mi @ self.J_post.T) if (has_s and posture) else self.L_js
        self.classify()

    def classify(self):  # .cpp:230-295
        r = self.robot
        if len(self.types) == 0 or self.c2 > self.c1:
            self.q_prior = r.q(); self.dq_prior = r.dq()
        if np.linalg.norm(self.U_s) == 0:
            self.types = []; self.history.clear(); self.c1 = self.c2 = 0
            return
        ncol = self.U_s.shape[1]
        self.types = [NO_SINGULARITY] * ncol
        q0 = r.q()
        pos0 = r.position(self.link, self.p_c); ori0 = r.rotation(self.link, self.R_c)
        for i in range(ncol):
            r.setQ(q0 + self.perturb_step * self.V_s[:, i]); r.updateKinematics()
            dpos = r.position(self.link, self.p_c) - pos0
            dori = orientation_error(r.rotation(self.link, self.R_c), ori0)
            motion = abs(np.concatenate([dpos, dori]) @ self.U_s[:, i])
            self.types[i] = TYPE_1_SINGULARITY if motion > self.type_1_tol else TYPE_2_SINGULARITY
            r.setQ(q0); r.updateKinematics()
        if TYPE_1_SINGULARITY in self.types:
            self.history.append(TYPE_1_SINGULARITY); self.c1 += 1
        else:
            self.history.append(TYPE_2_SINGULARITY); self.c2 += 1
        if len(self.history) > self.buffer_size:
            if self.history[0] == TYPE_1_SINGULARITY:
                self.c1 -= 1
            else:
                self.c2 -= 1
            self.history.popleft()

    def computeTorques(self, F_star, F_f):  # .cpp:297-368
        r, n = self.robot, self.n
        if len(self.types) == 0:
            return self.J_ns.T @ (self.L_ns_mod @ self.U_ns.T @ F_star + self.U_ns.T @ F_f)
        if self.decoupling == IMPEDANCE:
            return self.J_ns.T @ (self.U_ns.T @ F_star + self.U_ns.T @ F_f)
        if np.linalg.norm(self.U_ns) == 0:
            return np.zeros(n)
        tau_ns = self.J_ns.T @ (self.L_ns_mod @ self.U_ns.T @ F_star + self.U_ns.T @ F_f)
        if not self.enforce_handling:
            return tau_ns
        q, dq = r.q(), r.dq()
        if self.c1 > self.c2 or self.enforce_type_1:
            unit = -self.kp1 * (q - self.q_prior) - self.kv1 * dq
            tau_j = self.J_post.T @ self.L_js_mod @ self.V_s.T @ unit
        else:
            for i in range(n):
                if self.V_s[i, 0] != 0:
                    if abs(q[i] - self.q_upper[i]) < self.type_2_angle_threshold:
                        self.type_2_direction[i] = -1
                    elif abs(q[i] - self.q_lower[i]) < self.type_2_angle_threshold:
                        self.type_2_direction[i] = 1
            f = F_star + F_f
            nf = np.linalg.norm(f)
            fTd = (f / nf if nf > 0 else f) @ self.U_s[:, 0]  # Eigen normalized() leaves a zero vector unchanged
            unit = self.type_2_direction * (abs(fTd) * self.type_2_torque_vector)
            tau_j = self.J_post.T @ self.V_s.T @ unit + self.J_post.T @ self.L_js_mod @ self.V_s.T @ (-self.kv2 * dq)
        with np.errstate(all="ignore"):
            tau_s = self.J_s.T @ (self.L_s_mod @ self.U_s.T @ F_star + self.U_s.T @ F_f)
        for i in range(n):
            if np.isnan(tau_s[i]):
                tau_s[i] = 0
            elif tau_s[i] > self.tau_upper[i]:
                tau_s[i] = self.tau_upper[i]
            elif tau_s[i] < self.tau_lower[i]:
                tau_s[i] = self.tau_lower[i]
        return tau_ns + self.alpha * tau_s + (1 - self.alpha) * tau_j

    def lambda_world(self):
        """Basis-independent export: U_ns Lambda_ns U_ns^T (6x6)."""
        if np.linalg.norm(self.U_ns) == 0:
            return np.zeros((6, 6))
        return self.U_ns @ self.L_ns @ self.U_ns.T


# ----------------------------------------------------------------------------------------------------------
# MotionForceTask  (tasks/MotionForceTask.cpp)
# ----------------------------------------------------------------------------------------------------------
class MotionForceTask:
    def __init__(self, robot, link, p_c=np.zeros(3), R_c=np.eye(3), dirs_t=None, dirs_r=None,
                 in_compliant_frame=False, dt=0.001, name="motion_force_task"):
        self.robot, self.link, self.dt, self.name = robot, link, dt, name
        self.p_c, self.R_c = np.array(p_c, float), np.array(R_c, float)
        self.in_compliant_frame = in_compliant_frame
        if dirs_t is None and dirs_r is None:  # .cpp:28
            self.P = np.eye(6)
        else:  # .cpp:47-87
            if len(dirs_t) == 0 and len(dirs_r) == 0:
                raise ValueError("controlled_directions_translation and controlled_directions_rotation cannot both be empty")
            Bt = matrix_range_basis(np.array(dirs_t, float).T) if len(dirs_t) else np.zeros((3, 1))
            Br = matrix_range_basis(np.array(dirs_r, float).T) if len(dirs_r) else np.zeros((3, 1))
            self.P = np.zeros((6, 6))
            self.P[:3, :3] = Bt @ Bt.T
            self.P[3:, 3:] = Br @ Br.T
        self._initial_setup()

    def _initial_setup(self):  # .cpp:92-202, defaults MotionForceTask.h:40-75
        n = self.robot.dof()
        self.T_cs_R, self.T_cs_p = np.eye(3), np.zeros(3)
        self.popc = POPC(self.dt)
        self.kp_pos, self.kv_pos, self.ki_pos = np.full(3, 100.0), np.full(3, 20.0), np.zeros(3)
        self.kp_ori, self.kv_ori, self.ki_ori = np.full(3, 200.0), np.full(3, 28.3), np.zeros(3)
        self.kp_force, self.kv_force, self.ki_force = np.full(3, 0.7), np.full(3, 10.0), np.full(3, 1.3)
        self.kp_moment, self.kv_moment, self.ki_moment = np.full(3, 0.7), np.full(3, 10.0), np.full(3, 1.3)
        self.use_vel_sat, self.lin_sat, self.ang_sat = False, 0.3, math.pi / 3
        self.kff_force = self.kff_moment = 0.95
        self.max_force_fb, self.max_moment_fb = 20.0, 10.0
        self.force_dim = self.moment_dim = 0
        self.force_axis = np.zeros(3); self.moment_axis = np.zeros(3)
        self.closed_loop_force = self.closed_loop_moment = False
        self.N_prec = np.eye(n); self.N = np.zeros((n, n))
        self.J = np.zeros((6, n)); self.Jp = np.zeros((6, n))
        rp = matrix_range_basis(self.P[:3, :3]); ro = matrix_range_basis(self.P[3:, 3:])
        self.pos_range = 0 if np.linalg.norm(rp) == 0 else rp.shape[1]
        self.ori_range = 0 if np.linalg.norm(ro) == 0 else ro.shape[1]
        if self.pos_range + self.ori_range == 0:
            raise ValueError("empty task")
        self.use_otg = False  # OTG is a 'next' row; parity runs disable it like 04...cpp:115
        self.handler = SingularityHandler(self.robot, self.link, self.p_c, self.R_c, self.pos_range + self.ori_range)
        self.integ_force = np.zeros(3); self.integ_moment = np.zeros(3)
        self.reInitializeTask()

    def reInitializeTask(self):  # .cpp:204-245
        r = self.robot
        self.x = r.positionInWorld(self.link, self.p_c); self.R = r.rotationInWorld(self.link, self.R_c)
        self.goal_pos = self.x.copy(); self.goal_ori = self.R.copy()
        self.des_pos = self.x.copy(); self.des_ori = self.R.copy()
        z = np.zeros(3)
        self.v = z.copy(); self.w = z.copy()
        self.goal_linvel = z.copy(); self.goal_angvel = z.copy(); self.goal_linacc = z.copy(); self.goal_angacc = z.copy()
        self.des_linvel = z.copy(); self.des_angvel = z.copy(); self.des_linacc = z.copy(); self.des_angacc = z.copy()
        self.ori_err = z.copy()
        self.integ_pos = z.copy(); self.integ_ori = z.copy()
        self.goal_force = z.copy(); self.goal_moment = z.copy()
        self.sensed_force = z.copy(); self.sensed_moment = z.copy()
        self.integ_force = z.copy(); self.integ_moment = z.copy()
        self.unit_mass_force = np.zeros(6)

    # --- setters used by the tests (MotionForceTask.h) ---
    def setPosControlGains(self, kp, kv, ki=0.0):
        self.kp_pos, self.kv_pos, self.ki_pos = (np.broadcast_to(np.asarray(g, float), (3,)).copy() for g in (kp, kv, ki))

    def setOriControlGains(self, kp, kv, ki=0.0):
        self.kp_ori, self.kv_ori, self.ki_ori = (np.broadcast_to(np.asarray(g, float), (3,)).copy() for g in (kp, kv, ki))

    def setForceControlGains(self, kp, kv, ki):
        self.kp_force, self.kv_force, self.ki_force = np.full(3, kp), np.full(3, kv), np.full(3, ki)

    def setMomentControlGains(self, kp, kv, ki):
        self.kp_moment, self.kv_moment, self.ki_moment = np.full(3, kp), np.full(3, kv), np.full(3, ki)

    def enableVelocitySaturation(self, lin=0.3, ang=math.pi / 3):
        self.use_vel_sat, self.lin_sat, self.ang_sat = True, lin, ang

    def setForceSensorFrame(self, T_link_sensor_R, T_link_sensor_p):  # .cpp:794-803
        self.T_cs_R = self.R_c.T @ T_link_sensor_R
        self.T_cs_p = self.R_c.T @ (T_link_sensor_p - self.p_c)

    def updateSensedForceAndMoment(self, f_s, m_s):  # .cpp:805-828
        Rw = self.robot.rotationInWorld(self.link, self.R_c)
        f = self.T_cs_R @ f_s
        m = np.cross(self.T_cs_p, f) + self.T_cs_R @ m_s
        self.sensed_force = Rw @ f; self.sensed_moment = Rw @ m

    def parametrizeForceMotionSpaces(self, dim, axis=np.zeros(3)):  # .cpp:830-858
        reset = dim != self.force_dim
        self.force_dim = dim
        if dim in (1, 2):
            a = np.asarray(axis, float) / np.linalg.norm(axis)
            reset = reset or not np.allclose(a, self.force_axis, rtol=1e-12, atol=0)
            self.force_axis = a
        if reset:
            self.goal_pos = self.x.copy(); self.goal_linvel[:] = 0; self.goal_linacc[:] = 0
            self.integ_pos[:] = 0; self.integ_force[:] = 0
        return reset

    def parametrizeMomentRotMotionSpaces(self, dim, axis=np.zeros(3)):  # .cpp:860-890
        reset = dim != self.moment_dim
        self.moment_dim = dim
        if dim in (1, 2):
            a = np.asarray(axis, float) / np.linalg.norm(axis)
            reset = reset or not np.allclose(a, self.moment_axis, rtol=1e-12, atol=0)
            self.moment_axis = a
        if reset:
            self.goal_ori = self.R.copy(); self.goal_angvel[:] = 0; self.goal_angacc[:] = 0
            self.integ_ori[:] = 0; self.integ_moment[:] = 0
        return reset

    def setClosedLoopForceControl(self, on=True):  # .cpp:973-979
        if self.closed_loop_force != on:
            self.closed_loop_force = on
            self.integ_pos[:] = 0; self.integ_force[:] = 0

    def setClosedLoopMomentControl(self, on=True):  # .cpp:980-986
        if self.closed_loop_moment != on:
            self.closed_loop_moment = on
            self.integ_ori[:] = 0; self.integ_moment[:] = 0

    # --- sigma matrices (.cpp:892-971) ---
    def _rot_param(self):
        return self.robot.rotationInWorld(self.link, self.R_c) if self.in_compliant_frame else np.eye(3)

    @staticmethod
    def _sigma(dim, Psel, rot, axis):
        if dim == 0:
            return np.zeros((3, 3))
        if dim == 3:
            return Psel.copy()
        a = rot @ axis
        aa = np.outer(a, a)
        return Psel @ (aa if dim == 1 else np.eye(3) - aa) @ Psel.T

    def sigmaForce(self):
        return self._sigma(self.force_dim, self.P[:3, :3], self._rot_param(), self.force_axis)

    def sigmaMoment(self):
        return self._sigma(self.moment_dim, self.P[3:, 3:], self._rot_param(), self.moment_axis)

    def sigmaPosition(self):
        Pt = self.P[:3, :3]
        return Pt @ (np.eye(3) - self.sigmaForce()) @ Pt.T

    def sigmaOrientation(self):
        Pr = self.P[3:, 3:]
        return Pr @ (np.eye(3) - self.sigmaMoment()) @ Pr.T

    def updateTaskModel(self, N_prec):  # .cpp:247-268
        self.N_prec = N_prec.copy()
        self.J = self.P @ self.robot.JWorldFrame(self.link, self.p_c)
        self.Jp = self.J @ self.N_prec
        self.handler.updateTaskModel(self.Jp, self.N_prec)
        self.N = self.handler.N

    def getTaskAndPreviousNullspace(self):
        return self.N @ self.N_prec

    def computeTorques(self, tau_prec=None):  # .cpp:270-509; the tau_prec term is identically 0 (Q1)
        r, dt = self.robot, self.dt
        self.J = self.P @ r.JWorldFrame(self.link, self.p_c)
        self.Jp = self.J @ self.N_prec
        self.x = r.positionInWorld(self.link, self.p_c); self.R = r.rotationInWorld(self.link, self.R_c)
        self.ori_err = orientation_error(self.goal_ori, self.R)
        dq = r.dq()
        self.v = self.J[:3] @ dq; self.w = self.J[3:] @ dq
        Sf, Sm, Sp, So = self.sigmaForce(), self.sigmaMoment(), self.sigmaPosition(), self.sigmaOrientation()
        rot = self._rot_param()
        goal_force, goal_moment = rot @ self.goal_force, rot @ self.goal_moment
        if self.closed_loop_force:
            self.integ_force = self.integ_force + Sf @ (self.sensed_force - goal_force) * dt
            fb = Sf @ (-self.kp_force * (self.sensed_force - goal_force) - self.ki_force * self.integ_force)
            if np.linalg.norm(fb) > self.max_force_fb:
                fb = fb * (self.max_force_fb / np.linalg.norm(fb))
            F_fb = self.popc.compute(Sf @ goal_force, Sf @ self.sensed_force, Sf @ fb, Sf @ self.v, self.kv_force, self.kff_force)
        else:
            F_fb = Sf @ (-self.kv_force * self.v)
        if self.closed_loop_moment:
            self.integ_moment = self.integ_moment + Sm @ (self.sensed_moment - goal_moment) * dt
            mb = Sm @ (-self.kp_moment * (self.sensed_moment - goal_moment) - self.ki_moment * self.integ_moment)
            if np.linalg.norm(mb) > self.max_moment_fb:
                mb = mb * (self.max_moment_fb / np.linalg.norm(mb))
            M_fb = Sm @ (mb - self.kv_moment * self.w)
        else:
            M_fb = Sm @ (-self.kv_moment * self.w)
        self.des_pos, self.des_ori = self.goal_pos.copy(), self.goal_ori.copy()
        self.des_linvel, self.des_angvel = self.goal_linvel.copy(), self.goal_angvel.copy()
        self.des_linacc, self.des_angacc = self.goal_linacc.copy(), self.goal_angacc.copy()
        self.integ_pos = self.integ_pos + Sp @ (self.x - self.des_pos) * dt
        if self.use_vel_sat:
            kvi = pseudo_inverse_diag(self.kv_pos)
            self.des_linvel = -self.kp_pos * (kvi * (Sp @ (self.x - self.des_pos))) - self.ki_pos * (kvi * self.integ_pos)
            nv = np.linalg.norm(self.des_linvel)
            if nv > self.lin_sat:
                self.des_linvel = self.des_linvel * (self.lin_sat / nv)
            F_p = Sp @ (self.des_linacc - self.kv_pos * (self.v - self.des_linvel))
        else:
            F_p = Sp @ (self.des_linacc - self.kp_pos * (self.x - self.des_pos) - self.kv_pos * (self.v - self.des_linvel)
                        - self.ki_pos * self.integ_pos)
        e_o = So @ orientation_error(self.des_ori, self.R)
        self.integ_ori = self.integ_ori + e_o * dt
        if self.use_vel_sat:
            kvi = pseudo_inverse_diag(self.kv_ori)
            self.des_angvel = -self.kp_ori * (kvi * e_o) - self.ki_ori * (kvi * self.integ_ori)
            nw = np.linalg.norm(self.des_angvel)
            if nw > self.ang_sat:
                self.des_angvel = self.des_angvel * (self.ang_sat / nw)
            F_o = So @ (self.des_angacc - self.kv_ori * (self.w - self.des_angvel))
        else:
            F_o = So @ (self.des_angacc - self.kp_ori * e_o - self.kv_ori * (self.w - self.des_angvel)
                        - self.ki_ori * self.integ_ori)
        F_star = np.concatenate([F_p, F_o])
        self.unit_mass_force = F_star.copy()
        F_fm = np.concatenate([F_fb, M_fb])
        ff = np.concatenate([Sf @ goal_force, Sm @ goal_moment])
        if self.closed_loop_force:  # Q3: the moment half is scaled under the *force* flag
            ff[:3] *= self.kff_force; ff[3:] *= self.kff_moment
        return self.handler.computeTorques(F_star, F_fm + ff)


# ----------------------------------------------------------------------------------------------------------
# JointTask  (tasks/JointTask.cpp)
# ----------------------------------------------------------------------------------------------------------
class JointTask:
    def __init__(self, robot, S=None, dt=0.001, name="joint_task"):  # .cpp:14-89, defaults JointTask.h:31-45
        self.robot, self.dt, self.name = robot, dt, name
        n = robot.dof()
        self.S = np.eye(n) if S is None else np.atleast_2d(np.array(S, float))
        if self.S.shape[1] != n:
            raise ValueError("joint selection matrix size not consistent with robot dof")
        if np.linalg.matrix_rank(self.S) != self.S.shape[0]:
            raise ValueError("joint selection matrix is not full rank")
        self.k = self.S.shape[0]
        self.decoupling, self.bie_threshold = BOUNDED_INERTIA_ESTIMATES, 0.1
        self.kp, self.kv, self.ki = np.full(self.k, 50.0), np.full(self.k, 14.0), np.zeros(self.k)
        self.use_vel_sat = False; self.sat_vel = np.full(self.k, math.pi / 3)
        self.N_prec = np.eye(n); self.N = np.zeros((n, n))
        self.Lam = np.eye(self.k); self.Lam_mod = np.eye(self.k)
        self.Jp = self.S.copy(); self.range = np.eye(self.k)
        self.use_otg = False
        self.reInitializeTask()

    def isFullJointTask(self):
        return self.k == self.robot.dof()

    def reInitializeTask(self):  # .cpp:91-107
        self.cur_pos = self.S @ self.robot.q(); self.cur_vel = np.zeros(self.k)
        self.goal_pos = self.cur_pos.copy(); self.des_pos = self.cur_pos.copy()
        self.goal_vel = np.zeros(self.k); self.goal_acc = np.zeros(self.k)
        self.des_vel = np.zeros(self.k); self.des_acc = np.zeros(self.k)
        self.integ = np.zeros(self.k)

    def setGains(self, kp, kv, ki=0.0):
        self.kp, self.kv, self.ki = (np.broadcast_to(np.asarray(g, float), (self.k,)).copy() for g in (kp, kv, ki))

    def enableVelocitySaturation(self, v=math.pi / 3):
        self.use_vel_sat = True; self.sat_vel = np.broadcast_to(np.asarray(v, float), (self.k,)).copy()

    def updateTaskModel(self, N_prec):  # .cpp:218-283
        r, n = self.robot, self.robot.dof()
        self.N_prec = N_prec.copy()
        self.Jp = self.S @ self.N_prec
        self.range = matrix_range_basis(self.Jp)
        if np.linalg.norm(self.range) == 0:
            self.N = np.eye(n)
            return
        Jr = self.range.T @ self.Jp
        self.Lam, _, self.N = r.operationalSpaceMatrices(Jr)
        if self.decoupling == FULL_DYNAMIC_DECOUPLING:
            self.Lam_mod = self.Lam
        elif self.decoupling == BOUNDED_INERTIA_ESTIMATES:
            M_BIE = r.M()
            for i in range(n):
                if M_BIE[i, i] < self.bie_threshold:
                    M_BIE[i, i] = self.bie_threshold
            self.Lam_mod = np.linalg.inv(Jr @ np.linalg.inv(M_BIE) @ Jr.T)
        else:
            self.Lam_mod = np.eye(self.range.shape[1])

    def getTaskAndPreviousNullspace(self):
        return self.N @ self.N_prec

    def computeTorques(self, tau_prec=None):  # .cpp:285-356
        r = self.robot
        self.Jp = self.S @ self.N_prec
        self.cur_pos = self.S @ r.q(); self.cur_vel = self.S @ r.dq()
        if np.linalg.norm(self.range) == 0:
            return np.zeros(r.dof())
        self.des_pos, self.des_vel, self.des_acc = self.goal_pos.copy(), self.goal_vel.copy(), self.goal_acc.copy()
        self.integ = self.integ + (self.cur_pos - self.des_pos) * self.dt
        if self.use_vel_sat:
            kvi = pseudo_inverse_diag(self.kv)
            self.des_vel = -self.kp * kvi * (self.cur_pos - self.des_pos) - self.ki * kvi * self.integ
            self.des_vel = np.clip(self.des_vel, -self.sat_vel, self.sat_vel)  # Q7: loop to k
            u = -self.kv * (self.cur_vel - self.des_vel)
        else:
            u = -self.kp * (self.cur_pos - self.des_pos) - self.kv * (self.cur_vel - self.des_vel) - self.ki * self.integ
        f = self.Lam @ self.range.T @ self.des_acc + self.Lam_mod @ self.range.T @ u
        tau = self.Jp.T @ self.range @ f
        if tau_prec is not None:
            tau = tau - self.Jp.T @ self.range @ self.Lam @ self.range.T @ self.S @ r.MInv() @ tau_prec
        return tau


# ----------------------------------------------------------------------------------------------------------
# RobotController  (RobotController.cpp)
# ----------------------------------------------------------------------------------------------------------
class RobotController:
    def __init__(self, robot, tasks):  # .cpp:8-51
        if len(tasks) == 0:
            raise ValueError("RobotController must have at least one task")
        names, closed = [], False
        for t in tasks:
            if t.robot is not robot:
                raise ValueError("All tasks must have the same robot model in RobotController")
            if t.dt != tasks[0].dt:
                raise ValueError("All tasks must have the same loop timestep in RobotController")
            if t.name in names:
                raise ValueError("Tasks in RobotController must have unique names")
            names.append(t.name)
            if closed:
                raise ValueError("task [%s] cannot be added to the controller because it is in the nullspace of a full joint task" % t.name)
            if isinstance(t, JointTask) and t.isFullJointTask():
                closed = True
        self.robot, self.tasks, self.gravity_comp = robot, list(tasks), False

    def updateControllerTaskModels(self):  # .cpp:53-60
        N_prec = np.eye(self.robot.dof())
        for t in self.tasks:
            t.updateTaskModel(N_prec)
            N_prec = t.getTaskAndPreviousNullspace()
        self.N_total = N_prec

    def computeControlTorques(self):  # .cpp:62-74
        tau = np.zeros(self.robot.dof())
        for t in self.tasks:
            tau = tau + t.computeTorques(tau)
        if self.gravity_comp:
            tau = tau + self.robot.jointGravityVector()
        return tau
