"""Code-independent pins for the control-law CONVENTIONS: tests/golden/pins_*.npz.

The reference has no golden vectors and cannot be built here; the two oracles (oracle/osc_oracle.hpp,
oracle/np_oracle.py) were written from the same reading of it.  This generator shares NO code and no algorithm with
either of them (and none with the CUDA kernels):

* the robot model comes from ANALYTICAL MECHANICS, not from recursions: M(q) is the Hessian of the kinetic energy
  T = sum_b 1/2 m |d c_b/dt|^2 + 1/2 w_b^T (R_b I_b R_b^T) w_b in the joint velocities, J(q) is the derivative of the
  forward kinematics map (position: dx/dq; orientation: vee(dR/dq R^T)).  For the planar 4R `rrrrbot` both are derived in
  closed form with sympy (Lagrange); for any serial chain (the Panda) the derivatives are central differences evaluated in
  50-digit mpmath arithmetic (step 1e-20: truncation and rounding errors below 1e-28);
* every matrix inverse is a 50-digit mpmath LU solve; Lambda, Jbar, N, the bounded-inertia variants and the range basis of
  the joint task follow the DEFINITIONS (Lambda = (J M^-1 J^T)^-1, Jbar = M^-1 J^T Lambda,
  N = I - Jbar J), not an SVD route;
* the control law is typed in from the reference text (MotionForceTask.cpp:278-509, JointTask.cpp:285-356,
  RobotController.cpp:53-74, POPCExplicitForceControl.cpp:30-96), with hand-derived closed forms for the pieces that have
  one: orientationError(R_d, Rot(a, theta) R_d) = sin(theta) a (proved symbolically below), the selection projectors
  a a^T / I - a a^T, the velocity-saturation clamp.

Outputs are float64 roundings of 50-digit results.  tests/test_pins.py replays the inputs on the C++ oracle, the numpy
oracle (CPU) and the CUDA path (GPU) and compares at 1e-9.  Run in the build container:  python tools/make_sympy_golden.py
"""
import json
import os
import sys
from collections import deque

import mpmath as mp
import numpy as np
import sympy as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")
mp.mp.dps = 50
H = mp.mpf(10) ** -20


def chain(name):
    return json.load(open(os.path.join(ROOT, "fixtures", name + ".chain.json")))


def M3(a):
    return mp.matrix([[mp.mpf(float(x)) for x in row] for row in a])


def V3(a):
    return mp.matrix([mp.mpf(float(x)) for x in a])


def rot(axis, ang):
    """Rodrigues in mpmath"""
    a = V3(axis); a = a / mp.norm(a)
    K = mp.matrix([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return mp.eye(3) + mp.sin(ang) * K + (1 - mp.cos(ang)) * (K * K)


def vee(S):
    return mp.matrix([(S[2, 1] - S[1, 2]) / 2, (S[0, 2] - S[2, 0]) / 2, (S[1, 0] - S[0, 1]) / 2])


class MpChain:
    """forward kinematics of a serial chain in mpmath; everything else is derived from it by differentiation"""

    def __init__(self, c):
        self.c, self.n = c, c["n"]

    def fk(self, q):
        """world pose (R, p) of every body frame"""
        R, p = M3(self.c["world_rot"]), V3(self.c["world_xyz"])
        out = []
        for i, j in enumerate(self.c["joints"]):
            p = p + R * V3(j["xyz"])
            R = R * M3(j["rot"])
            if j["type"] == "prismatic":
                p = p + R * V3(j["axis"]) * q[i]
            else:
                R = R * rot(j["axis"], q[i])
            out.append((R, p))
        return out

    def dfk(self, q):
        """d/dq_i of every body pose: central differences, step 1e-20 at 50 digits"""
        d = []
        for i in range(self.n):
            qp, qm = list(q), list(q)
            qp[i] = q[i] + H; qm[i] = q[i] - H
            fp, fm = self.fk(qp), self.fk(qm)
            d.append([((a[0] - b[0]) / (2 * H), (a[1] - b[1]) / (2 * H)) for a, b in zip(fp, fm)])
        return d  # d[i][b] = (dR_b/dq_i, dp_b/dq_i)

    def mass_matrix(self, q):
        """Hessian of the kinetic energy in dq (Lagrange), body by body"""
        n, f, d = self.n, self.fk(q), self.dfk(q)
        M = mp.zeros(n, n)
        for b, body in enumerate(self.c["bodies"]):
            Rb = f[b][0]
            m, com, I = mp.mpf(body["mass"]), V3(body["com"]), M3(body["inertia"])
            Iw = Rb * I * Rb.T
            dc = [d[i][b][1] + d[i][b][0] * com for i in range(n)]          # d c_b / d q_i
            w = [vee(d[i][b][0] * Rb.T) for i in range(n)]                  # angular velocity per unit dq_i
            for i in range(n):
                for j in range(n):
                    M[i, j] += m * (dc[i].T * dc[j])[0] + (w[i].T * Iw * w[j])[0]
        return M

    def point(self, q, link, p_in_link, R_in_link=None):
        """world pose of the compliant frame and its 6 x n Jacobian (linear rows first) by differentiation"""
        L = self.c["links"][link]
        b = L["body"]
        pl = V3(L["xyz"]) + M3(L["rot"]) * V3(p_in_link)
        Rl = M3(L["rot"]) * (mp.eye(3) if R_in_link is None else M3(R_in_link))
        f, d = self.fk(q), self.dfk(q)
        x, R = f[b][1] + f[b][0] * pl, f[b][0] * Rl
        J = mp.zeros(6, self.n)
        for i in range(self.n):
            dx = d[i][b][1] + d[i][b][0] * pl
            w = vee(d[i][b][0] * f[b][0].T)
            for a in range(3):
                J[a, i], J[3 + a, i] = dx[a], w[a]
        return x, R, J


def inv(A):
    return mp.inverse(A)  # LU with partial pivoting at 50 digits


def osm(J, Minv):
    """operationalSpaceMatrices by definition"""
    Lam = inv(J * Minv * J.T)
    Jbar = Minv * J.T * Lam
    return Lam, Jbar, mp.eye(J.cols) - Jbar * J


def bie(M, thr=mp.mpf("0.1")):
    Mb = M.copy()
    for i in range(M.rows):
        if Mb[i, i] < thr:
            Mb[i, i] = thr
    return Mb


def orientation_error(Rd, R):
    """Sai2Model::orientationError: -1/2 sum_i R[:,i] x Rd[:,i]"""
    e = mp.matrix([0, 0, 0])
    for i in range(3):
        a, b = R[:, i], Rd[:, i]
        e += mp.matrix([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]])
    return -e / 2


def f64(A):
    if isinstance(A, mp.matrix):
        return np.array([[float(A[i, j]) for j in range(A.cols)] for i in range(A.rows)]).squeeze()
    return np.array(A, dtype=np.float64)


def diag3(v):
    return mp.diag([mp.mpf(float(x)) for x in np.broadcast_to(np.asarray(v, float), (3,))])


def sigma(dim, axis, Rot, Psel):
    """MotionForceTask::sigmaForce / sigmaMoment (.cpp:892-971) -- closed forms: 0, a a^T, I - a a^T, I (in the rotated frame)"""
    if dim == 0:
        return mp.zeros(3, 3)
    if dim == 3:
        return Psel * Psel.T
    a = Rot * V3(axis)
    aa = a * a.T
    return Psel * (aa if dim == 1 else mp.eye(3) - aa) * Psel.T


class MftSpec:
    """plain data: the MotionForceTask parameters the law below reads (defaults MotionForceTask.h:40-75)"""

    def __init__(self, **kw):
        self.kp_pos, self.kv_pos, self.ki_pos = 100.0, 20.0, 0.0
        self.kp_ori, self.kv_ori, self.ki_ori = 200.0, 28.3, 0.0
        self.kp_f, self.kv_f, self.ki_f = 0.7, 10.0, 1.3
        self.kp_m, self.kv_m, self.ki_m = 0.7, 10.0, 1.3
        self.kff_f, self.kff_m, self.max_f, self.max_m = 0.95, 0.95, 20.0, 10.0
        self.force_dim, self.force_axis, self.moment_dim, self.moment_axis = 0, (0, 0, 1), 0, (0, 0, 1)
        self.closed_force, self.closed_moment, self.passivity = False, False, False
        self.vel_sat, self.lin_sat, self.ang_sat = False, 0.3, float(mp.pi / 3)
        self.in_compliant_frame = False
        self.Pt, self.Pr = mp.eye(3), mp.eye(3)
        self.dt = 0.001
        self.__dict__.update(kw)


class Popc:
    """POPCExplicitForceControl.cpp:10-96 typed in from the reference text, mpmath arithmetic, collections.deque window"""

    def __init__(self, dt):
        self.dt, self.PO, self.Ecorr, self.Estored, self.win, self.counter, self.Rc, self.vsq = mp.mpf(dt), mp.mpf(0), mp.mpf(0), mp.mpf(0), deque(), 50, mp.mpf(1), mp.mpf(0)

    def __call__(self, fd, fs, vcl, vr, kv, kff, enabled):
        if not enabled:
            return vcl - kv * vr
        Fcmd = kff * fd + self.Rc * vcl - kv * vr
        p = (((fs - fd).T * vcl)[0] - (Fcmd.T * vr)[0]) * self.dt
        self.PO += p
        self.win.append(p)
        if self.PO + self.Estored + self.Ecorr > 0:
            while len(self.win) > 250:
                if self.PO + self.Ecorr + self.Estored > self.win[0]:
                    if self.win[0] > 0:
                        self.PO -= self.win[0]
                    self.win.popleft()
                else:
                    break
        if self.counter <= 0:
            self.counter = 50
            old = self.Rc
            if self.PO + self.Estored + self.Ecorr < 0:
                self.Rc = min(mp.mpf(1), max(mp.mpf(0), 1 + (self.PO + self.Estored + self.Ecorr) / (self.vsq * self.dt)))
            else:
                self.Rc = (1 + (mp.mpf("0.1") * 50 - 1) * self.Rc) / (mp.mpf("0.1") * 50)
            self.Ecorr += (1 - old) * self.vsq * self.dt
            self.vsq = mp.mpf(0)
        self.counter -= 1
        self.vsq += (vcl.T * vcl)[0]
        return self.Rc * vcl - kv * vr


def mft_law(s, st, x, R, v, w, goal, sensed_f=None, sensed_m=None):
    """MotionForceTask.cpp:306-500: returns F* (6) and F_f = force/moment feedback + feed-forward (6); st = integrators + popc"""
    Rot = R if s.in_compliant_frame else mp.eye(3)
    Sf, Sm = sigma(s.force_dim, s.force_axis, Rot, s.Pt), sigma(s.moment_dim, s.moment_axis, Rot, s.Pr)
    Sp, So = s.Pt * (mp.eye(3) - Sf) * s.Pt.T, s.Pr * (mp.eye(3) - Sm) * s.Pr.T
    Fg, Mg = Rot * V3(goal.get("force", (0, 0, 0))), Rot * V3(goal.get("moment", (0, 0, 0)))
    dt = mp.mpf(s.dt)
    if s.closed_force:
        st["int_f"] += Sf * (sensed_f - Fg) * dt
        fb = Sf * (-mp.mpf(s.kp_f) * (sensed_f - Fg) - mp.mpf(s.ki_f) * st["int_f"])
        if mp.norm(fb) > s.max_f:
            fb = fb * (mp.mpf(s.max_f) / mp.norm(fb))
        Ffb = st["popc"](Sf * Fg, Sf * sensed_f, Sf * fb, Sf * v, mp.mpf(s.kv_f), mp.mpf(s.kff_f), s.passivity)
    else:
        Ffb = Sf * (-mp.mpf(s.kv_f) * v)
    if s.closed_moment:
        st["int_m"] += Sm * (sensed_m - Mg) * dt
        mb = Sm * (-mp.mpf(s.kp_m) * (sensed_m - Mg) - mp.mpf(s.ki_m) * st["int_m"])
        if mp.norm(mb) > s.max_m:
            mb = mb * (mp.mpf(s.max_m) / mp.norm(mb))
        Mfb = Sm * (mb - mp.mpf(s.kv_m) * w)
    else:
        Mfb = Sm * (-mp.mpf(s.kv_m) * w)
    xd, Rd = goal["pos"], goal["ori"]
    vd, wd = V3(goal.get("linvel", (0, 0, 0))), V3(goal.get("angvel", (0, 0, 0)))
    ad, ald = V3(goal.get("linacc", (0, 0, 0))), V3(goal.get("angacc", (0, 0, 0)))
    kp, kv, ki = diag3(s.kp_pos), diag3(s.kv_pos), diag3(s.ki_pos)
    st["int_p"] += Sp * (x - xd) * dt
    if s.vel_sat:
        vd = -kp * inv(kv) * Sp * (x - xd) - ki * inv(kv) * st["int_p"]
        if mp.norm(vd) > s.lin_sat:
            vd = vd * (mp.mpf(s.lin_sat) / mp.norm(vd))
        Fp = Sp * (ad - kv * (v - vd))
    else:
        Fp = Sp * (ad - kp * (x - xd) - kv * (v - vd) - ki * st["int_p"])
    eo = So * orientation_error(Rd, R)
    st["int_o"] += eo * dt
    kp, kv, ki = diag3(s.kp_ori), diag3(s.kv_ori), diag3(s.ki_ori)
    if s.vel_sat:
        wd = -kp * inv(kv) * eo - ki * inv(kv) * st["int_o"]
        if mp.norm(wd) > s.ang_sat:
            wd = wd * (mp.mpf(s.ang_sat) / mp.norm(wd))
        Fo = So * (ald - kv * (w - wd))
    else:
        Fo = So * (ald - kp * eo - kv * (w - wd) - ki * st["int_o"])
    ff_f, ff_m = Sf * Fg, Sm * Mg
    if s.closed_force:  # quirk Q3: the moment half is scaled by the FORCE flag (.cpp:484-487)
        ff_f, ff_m = ff_f * mp.mpf(s.kff_f), ff_m * mp.mpf(s.kff_m)
    Fstar = mp.matrix(list(Fp) + list(Fo))
    Ff = mp.matrix(list(Ffb + ff_f) + list(Mfb + ff_m))
    return Fstar, Ff, vd, wd, eo


def new_state(dt=0.001):
    z = lambda: mp.matrix([0, 0, 0])
    return dict(int_p=z(), int_o=z(), int_f=z(), int_m=z(), popc=Popc(dt), int_q=None)


def two_task_tick(model, q, dq, spec, st, link, p_c, goal, jgoal, B6=None, stacked=True, sensed=None, jt=True,
                  jkp=50.0, jkv=14.0, jki=0.0):
    """[MotionForceTask (rank = cols of B6), full JointTask] through RobotController.cpp:53-74, non-singular branch
    (asserted), bounded inertia estimates (default).  B6: 6 x m basis of the controlled directions (identity when full)."""
    n = model.n
    qm, dqm = [mp.mpf(float(x)) for x in q], mp.matrix([mp.mpf(float(x)) for x in dq])
    M = model.mass_matrix(qm)
    Minv, Mbinv = inv(M), inv(bie(M))
    x, R, J6 = model.point(qm, link, p_c)
    B6 = mp.eye(6) if B6 is None else B6
    P = B6 * B6.T
    J = P * J6                                   # .cpp:262: _jacobian = _partial_task_projection * JWorldFrame
    tw = J * dqm                                 # .cpp:293-298 (projected velocities for a partial task)
    v, w = tw[0:3, 0], tw[3:6, 0]
    g = dict(goal)
    g["pos"] = x + V3(goal["dpos"]) if "dpos" in goal else V3(goal["pos"])
    g["ori"] = (goal["dR"](R) if "dR" in goal else (M3(goal["ori"]) if "ori" in goal else R))
    sf = sm = None
    if sensed is not None:  # updateSensedForceAndMoment (.cpp:805-828), identity sensor frame: rotated into the world by R
        sf, sm = R * V3(sensed[0]), R * V3(sensed[1])
    Fstar, Ff, vd, wd, eo = mft_law(spec, st, x, R, v, w, g, sf, sm)
    Jr = B6.T * J                                # m x n: the task in its own range (any orthonormal basis gives the same tau)
    S = mp.svd_r(Jr, compute_uv=False)
    assert S[len(S) - 1] / S[0] > mp.mpf("0.07"), "pin robots must sit in the non-singular branch (s_max = 6e-2)"
    Lam, Jbar, N1 = osm(Jr, Minv)
    Lam_b = inv(Jr * Mbinv * Jr.T)
    tau1 = Jr.T * (Lam_b * (B6.T * Fstar) + B6.T * Ff)      # SingularityHandler.cpp:307-309
    out = dict(x=x, R=R, J=J, M=M, Minv=Minv, Fstar=Fstar, ori_err=eo, tau1=tau1, v=v, w=w,
               lam_world=B6 * Lam * B6.T, lam_mod_world=B6 * Lam_b * B6.T, N1=N1)
    if not jt:
        out["tau"] = tau1
        return out
    # full JointTask in the nullspace (JointTask.cpp:218-356): Jp = N1, range basis = any unit vector of its column space
    Jp = N1
    k = max(range(n), key=lambda c: mp.norm(Jp[:, c]))
    rank_gap = mp.svd_r(Jp, compute_uv=False)
    r_dim = sum(1 for sv in rank_gap if sv / rank_gap[0] > mp.mpf("1e-3"))
    # orthonormal basis of range(Jp) by Gram-Schmidt on its largest columns (definition of matrixRangeBasis, App. B)
    cols = sorted(range(n), key=lambda c: -mp.norm(Jp[:, c]))
    Rb = []
    for c in cols:
        vcol = Jp[:, c]
        for b in Rb:
            vcol = vcol - b * (b.T * vcol)[0]
        if len(Rb) < r_dim and mp.norm(vcol) > mp.mpf("1e-6"):
            Rb.append(vcol / mp.norm(vcol))
    Rm = mp.matrix(n, r_dim)
    for c, b in enumerate(Rb):
        Rm[:, c] = b
    J2 = Rm.T * Jp
    Lam2, Jbar2, N2 = osm(J2, Minv)
    Lam2_b = inv(J2 * Mbinv * J2.T)
    qd, dqd, ddqd = (mp.matrix([mp.mpf(float(x)) for x in jgoal[key]]) for key in ("pos", "vel", "acc"))
    qv = mp.matrix(qm)
    if st["int_q"] is None:
        st["int_q"] = mp.zeros(n, 1)
    st["int_q"] += (qv - qd) * mp.mpf(spec.dt)
    u = -mp.mpf(jkp) * (qv - qd) - mp.mpf(jkv) * (dqm - dqd) - mp.mpf(jki) * st["int_q"]
    tau2 = Jp.T * Rm * (Lam2 * (Rm.T * ddqd) + Lam2_b * (Rm.T * u))
    if stacked:
        tau2 = tau2 - Jp.T * Rm * (Lam2 * (Rm.T * (Minv * tau1)))      # JointTask.cpp:285-292, tau_prec = running sum
    out.update(tau2=tau2, tau=tau1 + tau2, N_total=N2 * N1, jt_rank=r_dim, jt_lam_world=Rm * Lam2 * Rm.T)
    return out


# ---------------------------------------------------------------------------------------------------------------------
def prove_orientation_error():
    """orientationError(R_d, Rot(a, theta) R_d) = sin(theta) a for every R_d: symbolic proof with sympy"""
    th, a1, a2, a3 = sp.symbols("theta a1 a2 a3", real=True)
    a = sp.Matrix([a1, a2, a3])
    K = sp.Matrix([[0, -a3, a2], [a3, 0, -a1], [-a2, a1, 0]])
    E = sp.eye(3) + sp.sin(th) * K + (1 - sp.cos(th)) * K * K
    # R_d = I without loss of generality: the sum over an orthonormal frame is basis independent (vee(E - E^T))
    e = sp.zeros(3, 1)
    for i in range(3):
        e += E[:, i].cross(sp.eye(3)[:, i])
    e = sp.simplify((-e / 2 - sp.sin(th) * a).subs(a3 ** 2, 1 - a1 ** 2 - a2 ** 2))
    assert e == sp.zeros(3, 1), e


def rrrr_closed_form():
    """planar 4R by Lagrange in sympy: M(q) = Hessian of T, J = d(x, y, phi)/dq at the tip; returns mpmath callables"""
    c = chain("rrrrbot")
    q = sp.symbols("q1:5", real=True)
    dq = sp.symbols("dq1:5", real=True)
    l = [sp.Rational(float(j["xyz"][0])) for j in c["joints"]][1:] + [sp.Rational(1, 2)]   # link lengths 0.5 (tip at 0.5 in link4)
    th = [sum(q[:i + 1]) for i in range(4)]
    T = 0
    org = sp.Matrix([0, 0])
    for i, body in enumerate(c["bodies"]):
        m, cx, Izz = sp.Rational(float(body["mass"])), sp.Rational(float(body["com"][0])), sp.Rational(float(body["inertia"][2][2]))
        pc = org + cx * sp.Matrix([sp.cos(th[i]), sp.sin(th[i])])
        vc = sum((pc.diff(q[k]) * dq[k] for k in range(4)), sp.zeros(2, 1))
        om = sum(dq[:i + 1])
        T += m * (vc.T * vc)[0] / 2 + Izz * om ** 2 / 2
        org = org + l[i] * sp.Matrix([sp.cos(th[i]), sp.sin(th[i])])
    Msym = sp.hessian(T, dq)
    tip = sp.Matrix([org[0], org[1], th[3]])
    Jsym = tip.jacobian(q)
    Mf = sp.lambdify(q, Msym, "mpmath")
    Jf = sp.lambdify(q, Jsym, "mpmath")
    tipf = sp.lambdify(q, tip, "mpmath")
    return Mf, Jf, tipf


def case_rrrr(rng):
    """examples/11-planar_robot_controller: partial MotionForceTask (X, Y translation + Z rotation, .cpp:105-115) at
    link4 + (0.5, 0, 0), full JointTask, RobotController.  The general differentiation model must reproduce the sympy
    closed forms to 1e-25 before anything is written."""
    c = chain("rrrrbot")
    model = MpChain(c)
    Mf, Jf, tipf = rrrr_closed_form()
    B = 6
    q = []
    while len(q) < B:       # rejection: keep the task away from the blend band (sigma_min / sigma_max > 0.12; s_max = 0.06)
        qc = rng.uniform(-2.0, 2.0, 4)
        Jn = f64(Jf(*[mp.mpf(float(v)) for v in qc]))
        sv = np.linalg.svd(Jn, compute_uv=False)
        if sv[-1] / sv[0] > 0.12:
            q.append(qc)
    q = np.array(q)
    dq = rng.uniform(-0.5, 0.5, (B, 4))
    dpos = rng.normal(0, 0.05, (B, 3)); dpos[:, 2] = 0
    dth = rng.normal(0, 0.2, B)
    gv, gw = rng.normal(0, 0.05, (B, 3)), rng.normal(0, 0.05, (B, 3))
    gq, gdq, gddq = q + rng.normal(0, 0.1, (B, 4)), rng.normal(0, 0.1, (B, 4)), rng.normal(0, 0.1, (B, 4))
    B6 = mp.zeros(6, 3); B6[0, 0] = 1; B6[1, 1] = 1; B6[5, 2] = 1
    spec = MftSpec(Pt=mp.diag([1, 1, 0]), Pr=mp.diag([0, 0, 1]))
    rec = {k: [] for k in ("tau", "tau1", "tau2", "M", "J", "lam_world", "lam_mod_world", "N1", "N_total", "x", "R", "Fstar", "ori_err", "goal_ori", "tau_plain")}
    for r in range(B):
        qm = [mp.mpf(float(v)) for v in q[r]]
        Mc, Jc, tipc = Mf(*qm), Jf(*qm), tipf(*qm)
        Mg = model.mass_matrix(qm)
        x, R, J6 = model.point(qm, "link4", (0.5, 0, 0))
        assert mp.norm(Mc - Mg) < mp.mpf("1e-25"), "differentiation model vs sympy Lagrange (M)"
        J3 = mp.matrix(3, 4)
        for col in range(4):
            J3[0, col], J3[1, col], J3[2, col] = J6[0, col], J6[1, col], J6[5, col]
        assert mp.norm(J3 - Jc) < mp.mpf("1e-25"), "differentiation model vs sympy (J)"
        assert abs(x[0] - tipc[0]) + abs(x[1] - tipc[1]) < mp.mpf("1e-30")
        goal = dict(dpos=dpos[r], dR=lambda Rc, a=dth[r]: rot((0, 0, 1), a) * Rc, linvel=gv[r], angvel=gw[r])
        jg = dict(pos=gq[r], vel=gdq[r], acc=gddq[r])
        o = two_task_tick(model, q[r], dq[r], spec, new_state(), "link4", (0.5, 0, 0), goal, jg, B6=B6)
        o2 = two_task_tick(model, q[r], dq[r], spec, new_state(), "link4", (0.5, 0, 0), goal, jg, B6=B6, stacked=False)
        assert abs(o["ori_err"][2] + mp.sin(dth[r])) < mp.mpf("1e-30")            # closed form: R = Rot(z, -dth) R_goal
        for k in rec:
            if k == "goal_ori":
                rec[k].append(f64(rot((0, 0, 1), dth[r]) * o["R"]).ravel())
            elif k == "tau_plain":
                rec[k].append(f64(o2["tau"]))
            else:
                rec[k].append(f64(o[k]).ravel())
    return dict(q=q, dq=dq, dpos=dpos, gv=gv, gw=gw, gq=gq, gdq=gdq, gddq=gddq, **{k: np.array(v) for k, v in rec.items()})


def well_conditioned_panda_q(model, rng, B):
    c = model.c
    lo = np.array([j["lower"] for j in c["joints"]]) + 0.3
    hi = np.array([j["upper"] for j in c["joints"]]) - 0.3
    out = []
    while len(out) < B:
        q = rng.uniform(lo, hi)
        _, _, J = model.point([mp.mpf(float(v)) for v in q], "end-effector", (0, 0, 0.07))
        s = np.linalg.svd(f64(J), compute_uv=False)
        if s[5] / s[0] > 0.12:
            out.append(q)
    return np.array(out)


def case_panda_two_task(rng):
    """BASELINE config 2 shape: 6-DoF MotionForceTask at end-effector + (0, 0, 0.07), full JointTask, RobotController"""
    model = MpChain(chain("panda"))
    B = 4
    q = well_conditioned_panda_q(model, rng, B)
    dq = rng.uniform(-0.5, 0.5, (B, 7))
    dpos, ax, ang = rng.normal(0, 0.05, (B, 3)), rng.normal(0, 1, (B, 3)), rng.normal(0, 0.2, B)
    gv, gw, ga, gal = (rng.normal(0, 0.05, (B, 3)) for _ in range(4))
    gq, gdq, gddq = q + rng.normal(0, 0.1, (B, 7)), rng.normal(0, 0.1, (B, 7)), rng.normal(0, 0.1, (B, 7))
    keys = ("tau", "tau1", "tau2", "M", "Minv", "J", "lam_world", "lam_mod_world", "N1", "N_total", "x", "R", "Fstar", "ori_err", "v", "w", "jt_lam_world")
    rec = {k: [] for k in keys + ("goal_ori", "tau_plain")}
    for r in range(B):
        goal = dict(dpos=dpos[r], dR=lambda Rc, a=ax[r], t=ang[r]: rot(a, -t) * Rc, linvel=gv[r], angvel=gw[r], linacc=ga[r], angacc=gal[r])
        jg = dict(pos=gq[r], vel=gdq[r], acc=gddq[r])
        o = two_task_tick(model, q[r], dq[r], MftSpec(), new_state(), "end-effector", (0, 0, 0.07), goal, jg)
        o2 = two_task_tick(model, q[r], dq[r], MftSpec(), new_state(), "end-effector", (0, 0, 0.07), goal, jg, stacked=False)
        a = V3(ax[r]) / mp.norm(V3(ax[r]))
        # R = Rot(a, ang) R_goal -> sin(ang) a, up to the orthonormality of the fixture's float64 joint rotations (1e-16)
        assert mp.norm(o["ori_err"] - mp.sin(ang[r]) * a) < mp.mpf("1e-14")
        assert o["jt_rank"] == 1
        for k in keys:
            rec[k].append(f64(o[k]).ravel())
        rec["goal_ori"].append(f64(rot(ax[r], -ang[r]) * o["R"]).ravel())
        rec["tau_plain"].append(f64(o2["tau"]))
    return dict(q=q, dq=dq, dpos=dpos, gv=gv, gw=gw, ga=ga, gal=gal, gq=gq, gdq=gdq, gddq=gddq, **{k: np.array(v) for k, v in rec.items()})


LAW_CASES = [  # (name, MftSpec overrides, goal force, goal moment)
    ("force1_z_world", dict(force_dim=1, force_axis=(0, 0, 1)), (0, 0, -5.0), (0, 0, 0)),
    ("force2_x_compliant", dict(force_dim=2, force_axis=(1, 0, 0), in_compliant_frame=True), (1.0, -2.0, 3.0), (0, 0, 0)),
    ("force3_moment1", dict(force_dim=3, moment_dim=1, moment_axis=(0, 1, 0)), (1.0, 2.0, -3.0), (0.1, 0.2, -0.3)),
    ("moment2_compliant", dict(moment_dim=2, moment_axis=(0.6, 0, 0.8), in_compliant_frame=True), (0, 0, 0), (0.3, -0.1, 0.2)),
    ("moment3", dict(moment_dim=3), (0, 0, 0), (0.3, -0.1, 0.2)),
    ("vel_sat_saturated", dict(vel_sat=True, lin_sat=0.01, ang_sat=0.01), (0, 0, 0), (0, 0, 0)),
    ("vel_sat_unsaturated", dict(vel_sat=True, lin_sat=50.0, ang_sat=50.0), (0, 0, 0), (0, 0, 0)),
    ("anisotropic_gains_ki", dict(kp_pos=(100.0, 50.0, 150.0), kv_pos=(20.0, 10.0, 30.0), ki_pos=(3.0, 4.0, 5.0), kp_ori=(200.0, 100.0, 50.0),
                                  kv_ori=(28.3, 20.0, 10.0), ki_ori=(1.0, 2.0, 3.0)), (0, 0, 0), (0, 0, 0)),
    ("closed_loop_force_moment", dict(force_dim=1, force_axis=(0, 0, 1), moment_dim=1, moment_axis=(1, 0, 0), closed_force=True, closed_moment=True),
     (0, 0, -5.0), (0.2, 0, 0)),
    ("closed_loop_saturating_feedback", dict(force_dim=3, closed_force=True, max_f=2.0), (0, 0, -5.0), (0, 0, 0)),
]


def case_panda_law(rng):
    """known answers of the control law on one Panda posture per case: Sigma projectors (dims 0-3, world / compliant frame),
    velocity saturation (clamped and not), anisotropic gains with integrators (two ticks), closed-loop force/moment feedback
    with its saturation, feed-forward scaling (Q3).  Single MotionForceTask (no joint task): tau = J^T (Lambda_mod F* + F_f)."""
    model = MpChain(chain("panda"))
    q = well_conditioned_panda_q(model, rng, len(LAW_CASES))
    dq = rng.uniform(-0.5, 0.5, (len(LAW_CASES), 7))
    out = dict(q=q, dq=dq, names=np.array([c[0] for c in LAW_CASES]))
    dpos, ax, ang = rng.normal(0, 0.05, (len(LAW_CASES), 3)), rng.normal(0, 1, (len(LAW_CASES), 3)), rng.normal(0, 0.2, len(LAW_CASES))
    sf, sm = rng.normal(0, 4, (len(LAW_CASES), 2, 3)), rng.normal(0, 0.5, (len(LAW_CASES), 2, 3))
    rec = {k: [] for k in ("tau_t0", "tau_t1", "Fstar_t1", "goal_ori", "int_f", "int_p", "des_linvel", "des_angvel")}
    for r, (name, kw, gf, gm) in enumerate(LAW_CASES):
        spec, st = MftSpec(**kw), new_state()
        res = []
        Rg = None
        for tick in range(2):   # two ticks so that the integrators matter; q advances by dq * dt (kinematic playback)
            qt = q[r] + dq[r] * 0.001 * tick
            if Rg is None:
                _, R0, _ = model.point([mp.mpf(float(v)) for v in qt], "end-effector", (0, 0, 0.07))
                x0 = model.point([mp.mpf(float(v)) for v in qt], "end-effector", (0, 0, 0.07))[0]
                Rg, xg = rot(ax[r], -ang[r]) * R0, x0 + V3(dpos[r])
            goal = dict(pos=f64(xg), ori=f64(Rg), force=gf, moment=gm)
            # the goal pose is what the product receives: float64 roundings, so use the same roundings here
            o = two_task_tick(model, qt, dq[r], spec, st, "end-effector", (0, 0, 0.07), goal, None, jt=False, sensed=(sf[r, tick], sm[r, tick]))
            res.append(o)
        _, _, vd, wd, _ = mft_law(spec, new_state(), res[1]["x"], res[1]["R"], res[1]["v"], res[1]["w"], dict(pos=V3(f64(xg)), ori=M3(f64(Rg)), force=gf, moment=gm),
                                  res[1]["R"] * V3(sf[r, 1]), res[1]["R"] * V3(sm[r, 1])) if spec.vel_sat else (0, 0, V3((0, 0, 0)), V3((0, 0, 0)), 0)
        rec["tau_t0"].append(f64(res[0]["tau"])); rec["tau_t1"].append(f64(res[1]["tau"])); rec["Fstar_t1"].append(f64(res[1]["Fstar"]))
        rec["goal_ori"].append(f64(Rg).ravel()); rec["int_f"].append(f64(st["int_f"])); rec["int_p"].append(f64(st["int_p"]))
        rec["des_linvel"].append(f64(vd)); rec["des_angvel"].append(f64(wd))
        out.setdefault("goal_pos", []).append(f64(xg))
    out.update(sensed_f=sf, sensed_m=sm, goal_force=np.array([c[2] for c in LAW_CASES], float), goal_moment=np.array([c[3] for c in LAW_CASES], float),
               goal_pos=np.array(out["goal_pos"]), **{k: np.array(v) for k, v in rec.items()})
    return out


def case_popc_window(rng):
    """one passivity window and more: closed-loop force control on z with the passivity observer/controller on
    (examples/07 style), 520 ticks at a FIXED posture (the model is evaluated once, the law 520 times): the observer window
    (250) slides, the 50-tick controller period fires 10 times, R_c drops below 1 and recovers."""
    model = MpChain(chain("panda"))
    q = well_conditioned_panda_q(model, rng, 1)[0]
    T = 520
    spec = MftSpec(force_dim=1, force_axis=(0, 0, 1), closed_force=True, passivity=True)
    qm = [mp.mpf(float(x)) for x in q]
    x, R, J = model.point(qm, "end-effector", (0, 0, 0.07))
    # the control point moves AGAINST the commanded force (v_z = +0.08 m/s while F_cmd ~ -4.75 N along z): the robot absorbs
    # energy and the observer stays positive until the sensed force bursts; then R_c drops (0.62, 0) and recovers ((1 + 4 R_c) / 5)
    dq = np.linalg.pinv(f64(J)) @ np.array([0.01, -0.02, 0.08, 0, 0, 0])
    dqm = mp.matrix([mp.mpf(float(v_)) for v_ in dq])
    M = model.mass_matrix(qm)
    Mbinv = inv(bie(M))
    tw = J * dqm
    v, w = tw[0:3, 0], tw[3:6, 0]
    Lam_b = inv(J * Mbinv * J.T)
    t = np.arange(T)
    fz = -5.0 * (1 + 0.05 * np.sin(2 * np.pi * 3 * t * 0.001)) + rng.normal(0, 0.02, T)
    fz[120:135] *= 1.5          # a burst of active behaviour
    fw = np.zeros((T, 3)); fw[:, 2] = fz; fw[:, :2] = rng.normal(0, 1.0, (T, 2)) + 0.3   # off-axis components must be projected out
    f = np.array([f64(R.T * V3(fw[k])) for k in range(T)])   # what the sensor reads (sensor frame = compliant frame), float64
    st = new_state()
    goal = dict(pos=x, ori=R, force=(0, 0, -5.0))
    taus, rcs, pos_, ints = [], [], [], []
    for k in range(T):
        Fstar, Ff, _, _, _ = mft_law(spec, st, x, R, v, w, goal, R * V3(f[k]), V3((0, 0, 0)))
        taus.append(f64(J.T * (Lam_b * Fstar + Ff)))
        rcs.append(float(st["popc"].Rc)); pos_.append(float(st["popc"].PO)); ints.append(f64(st["int_f"]))
    assert min(rcs) == 0 and 0 < rcs[199] < 1 and rcs[-1] > 0.5 and len(st["popc"].win) <= T
    return dict(q=q, dq=dq, sensed_f=f, tau=np.array(taus), rc=np.array(rcs), po=np.array(pos_), int_f=np.array(ints),
                window_len=np.array([len(st["popc"].win)]))


# ---------------------------------------------------------------------------------------------------------------------
# The singular-blend branch (SingularityHandler.cpp:75-368) retyped from the reference text, 50 digits, definitions only (inverses
# by LU, the SVD by mpmath): for the closed-form blend routine of the CUDA path (csrc/osc_blend.cuh) and for the C++ oracle this is a
# third, independent implementation.  Constants: SingularityHandler.cpp:11-19, bounds 6e-3 / 6e-2 (MotionForceTask.h).
SH = dict(s_abs_tol=mp.mpf("1e-3"), type_1_tol=mp.mpf("0.5"), ratio=mp.mpf("1e-2"), angle_thr=5 * mp.pi / 180, step=mp.mpf(5), kp1=mp.mpf(50), kv1=mp.mpf(14),
          kv2=mp.mpf(5), s_min=mp.mpf("6e-3"), s_max=mp.mpf("6e-2"))


def blend_tick(model, q, dq, goal, jgoal, link="end-effector", p_c=(0, 0, 0.07), stacked=True, jkp=50.0, jkv=14.0, jt=True, B6=None):
    """First control cycle of [6-DoF MotionForceTask, full JointTask] for a robot INSIDE the blend band: classification history empty,
    q_prior latched to q, type-2 directions all +1 (SingularityHandler.cpp:40-62)."""
    n = model.n
    c = model.c
    assert M3(c["world_rot"]) == mp.eye(3)   # base frame = world frame for this fixture (the perturbation is measured in the base frame)
    qm, dqm = [mp.mpf(float(x)) for x in q], mp.matrix([mp.mpf(float(x)) for x in dq])
    qv = mp.matrix(qm)
    M = model.mass_matrix(qm)
    Minv, Mbinv = inv(M), inv(bie(M))
    x, R, J6 = model.point(qm, link, p_c)
    B6 = mp.eye(6) if B6 is None else B6           # 6 x m orthonormal basis of the controlled directions (partial task)
    m = B6.cols
    P = B6 * B6.T
    J = P * J6                                     # MotionForceTask.cpp:262
    tw = J * dqm
    spec, st = MftSpec(Pt=P[0:3, 0:3], Pr=P[3:6, 3:6]), new_state()
    g = dict(goal)
    g["pos"] = x + V3(goal["dpos"]); g["ori"] = R
    Fstar, Ff, vd, wd, eo = mft_law(spec, st, x, R, tw[0:3, 0], tw[3:6, 0], g, None, None)
    Ur, S, Vh = mp.svd_r(B6.T * J)                 # the task in its own range: J_r = U_r diag(S) Vh (m x m, m, m x n); U = B6 U_r
    U = B6 * Ur
    assert S[0] >= SH["s_abs_tol"]
    ins = next((i for i in range(1, m) if S[i] / S[0] < SH["s_max"]), None)
    assert ins is not None, "pin robots must sit in the blend band"
    alpha = min(max((S[ins] / S[0] - SH["s_min"]) / (SH["s_max"] - SH["s_min"]), mp.mpf(0)), mp.mpf(1))
    # sign convention of the implementation-defined JacobiSVD: largest-|component| of every V column positive
    V = Vh.T.copy()
    for k in range(m):
        col = [V[i, k] for i in range(n)]
        big = max(col, key=lambda t: abs(t))
        if big < 0:
            for i in range(n):
                V[i, k] = -V[i, k]
            for i in range(6):
                U[i, k] = -U[i, k]
    Uns, Us, Vs = U[:, 0:ins], U[:, ins:m], V[:, ins:m]
    Jns, Js = Uns.T * J, Us.T * J
    Lam_ns, _, N_ns = osm(Jns, Minv)
    Lam_ns_mod = inv(Jns * Mbinv * Jns.T)          # .cpp:175-190
    Lam_s_mod = inv(Js * Mbinv * Js.T)             # .cpp:196-202
    Jpost = Vs.T * N_ns                            # .cpp:150 (N_prec = I)
    _, _, N_post = osm(Jpost, Minv)
    Lam_js_mod = inv(Jpost * Mbinv * Jpost.T)      # .cpp:207-213
    N_t = N_post * N_ns
    # classifySingularity (.cpp:230-295), first call: one sample in the history
    types = []
    for k in range(m - ins):
        qp = [qm[i] + SH["step"] * Vs[i, k] for i in range(n)]
        x2, R2, _ = model.point(qp, link, p_c)
        d6 = mp.matrix(list(x2 - x) + list(orientation_error(R2, R)))
        motion = abs(sum(d6[a] * Us[a, k] for a in range(6)))
        types.append(1 if motion > SH["type_1_tol"] else 2)
    c1, c2 = (1, 0) if 1 in types else (0, 1)
    # computeTorques (.cpp:297-368)
    tau_ns = Jns.T * (Lam_ns_mod * (Uns.T * Fstar) + Uns.T * Ff)
    effort = mp.matrix([mp.mpf(j["effort"]) for j in c["joints"]])
    if c1 > c2:
        unit = -SH["kp1"] * (qv - qv) - SH["kv1"] * dqm              # q_prior latched to q in this cycle
        tj = Jpost.T * (Lam_js_mod * (Vs.T * unit))
    else:
        ft = Fstar + Ff
        fTd = sum((ft[a] / mp.norm(ft)) * Us[a, 0] for a in range(6))
        direction = [mp.mpf(1)] * n
        for i, j in enumerate(c["joints"]):
            if Vs[i, 0] != 0:
                if abs(qm[i] - mp.mpf(j["upper"])) < SH["angle_thr"]:
                    direction[i] = mp.mpf(-1)
                elif abs(qm[i] - mp.mpf(j["lower"])) < SH["angle_thr"]:
                    direction[i] = mp.mpf(1)
        unit = mp.matrix([direction[i] * abs(fTd) * SH["ratio"] * effort[i] for i in range(n)])
        tj = Jpost.T * (Vs.T * unit) + Jpost.T * (Lam_js_mod * (Vs.T * (-SH["kv2"] * dqm)))
    ts = Js.T * (Lam_s_mod * (Us.T * Fstar) + Us.T * Ff)
    ts = mp.matrix([min(max(ts[i], -effort[i]), effort[i]) for i in range(n)])
    tau1 = tau_ns + alpha * ts + (1 - alpha) * tj
    if not jt:  # a 6-DoF arm under the one task (examples/19): nothing below it
        return dict(tau=tau1, tau1=tau1, alpha=alpha, ins=ins, types=types + [0] * (6 - len(types)), S=S)
    # full JointTask below it: range of N_t (one direction for n = 7, m = 6)
    Jp = N_t
    cols = sorted(range(n), key=lambda cc: -mp.norm(Jp[:, cc]))   # orthonormal basis of range(N_t), dimension n - m (matrixRangeBasis)
    Rb = []
    for cc in cols:
        vcol = Jp[:, cc]
        for bb in Rb:
            vcol = vcol - bb * (bb.T * vcol)[0]
        if len(Rb) < n - m and mp.norm(vcol) > mp.mpf("1e-6"):
            Rb.append(vcol / mp.norm(vcol))
    Rm = mp.matrix(n, n - m)
    for cc, bb in enumerate(Rb):
        Rm[:, cc] = bb
    J2 = Rm.T * Jp
    Lam2, _, _ = osm(J2, Minv)
    Lam2_b = inv(J2 * Mbinv * J2.T)
    qd, dqd, ddqd = (mp.matrix([mp.mpf(float(t)) for t in jgoal[key]]) for key in ("pos", "vel", "acc"))
    u = -mp.mpf(jkp) * (qv - qd) - mp.mpf(jkv) * (dqm - dqd)
    tau2 = Jp.T * Rm * (Lam2 * (Rm.T * ddqd) + Lam2_b * (Rm.T * u))
    if stacked:
        tau2 = tau2 - Jp.T * Rm * (Lam2 * (Rm.T * (Minv * tau1)))
    return dict(tau=tau1 + tau2, tau1=tau1, tau2=tau2, alpha=alpha, ins=ins, types=types + [0] * (6 - len(types)), S=S, N_t=N_t)


def case_panda_blend(rng):
    """Panda postures on the elbow-extended / wrist-aligned sets (BASELINE config 4), inside the blend band, both joint-space
    strategies represented."""
    model = MpChain(chain("panda"))
    q0 = np.array([j["q0"] for j in chain("panda")["joints"]])
    want, recs, seen = 6, [], {1: 0, 2: 0}
    keys = ("tau", "tau1", "tau2", "alpha", "ins", "types", "S", "N_t", "tau_plain")
    out = {k: [] for k in keys}
    inp = {k: [] for k in ("q", "dq", "dpos", "gv", "gq", "gdq", "gddq")}
    tries = 0
    while len(out["tau"]) < want and tries < 200:
        tries += 1
        q = q0 + rng.normal(0, 0.2, 7) * np.array([1, 1, 1, 0, 1, 0, 1.0])
        eps = 10.0 ** rng.uniform(-3, -1.3)
        if tries % 2:
            q[3] = -0.0698 - abs(rng.normal(0, eps)); q[5] = q0[5] + rng.normal(0, 0.2)
        else:
            q[5] = rng.normal(0, eps); q[3] = q0[3] + rng.normal(0, 0.2)
        dq = rng.uniform(-0.3, 0.3, 7)
        goal = dict(dpos=rng.normal(0, 0.05, 3), linvel=rng.normal(0, 0.05, 3), angvel=np.zeros(3), linacc=np.zeros(3), angacc=np.zeros(3))
        jg = dict(pos=q + rng.normal(0, 0.1, 7), vel=rng.normal(0, 0.1, 7), acc=rng.normal(0, 0.1, 7))
        try:
            o = blend_tick(model, q, dq, goal, jg)
        except AssertionError:
            continue
        kind = 1 if 1 in o["types"] else 2
        if o["alpha"] in (0, 1) and len(out["tau"]) >= 2 or seen[kind] >= want - 2:
            continue
        # keep away from the classification threshold and the band edges: a pin must not depend on the last digits
        seen[kind] += 1
        o2 = blend_tick(model, q, dq, goal, jg, stacked=False)
        for k in ("tau", "tau1", "tau2", "S", "N_t"):
            out[k].append(f64(o[k]).ravel())
        out["alpha"].append(float(o["alpha"])); out["ins"].append(o["ins"]); out["types"].append(o["types"]); out["tau_plain"].append(f64(o2["tau"]).ravel())
        for k, v in (("q", q), ("dq", dq), ("dpos", goal["dpos"]), ("gv", goal["linvel"]), ("gq", jg["pos"]), ("gdq", jg["vel"]), ("gddq", jg["acc"])):
            inp[k].append(v)
    assert len(out["tau"]) == want and seen[1] > 0 and seen[2] > 0, (len(out["tau"]), seen)
    return dict(**{k: np.array(v) for k, v in inp.items()}, **{k: np.array(v) for k, v in out.items()})


def case_panda_blend_partial(rng):
    """BASELINE config 3 shape inside ITS singular band: translation-only MotionForceTask (rank 3) + full JointTask (nullspace of
    dimension 4), arm towards full extension."""
    model = MpChain(chain("panda"))
    q0 = np.array([j["q0"] for j in chain("panda")["joints"]])
    B6 = mp.zeros(6, 3)
    for a in range(3):
        B6[a, a] = 1
    want = 4
    keys = ("tau", "tau1", "tau2", "alpha", "ins", "types", "S", "tau_plain")
    out = {k: [] for k in keys}
    inp = {k: [] for k in ("q", "dq", "dpos", "gv", "gq", "gdq", "gddq")}
    # candidate postures: the position Jacobian is rank deficient only near full extension (about 5 % of the joint range); the C++ oracle
    # is used to FIND such postures among random ones -- the pinned outputs below come from the mpmath model alone
    sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200"))
    import orc
    import osc_abi as abi
    from common import random_q
    orc.build()
    cj = chain("panda")
    N = 4096
    cand = random_q(cj, N, rng, 0.3)
    ob = orc.OracleBatch(cj, N, q0=[j["q0"] for j in cj["joints"]])
    ob.set(abi.F_Q, cand); ob.set(abi.F_DQ, np.zeros_like(cand)); ob.update_model()
    t0 = ob.add_motion_force_task(abi.mft_desc(cj, "end-effector", (0, 0, 0.07), dirs_translation=[(1, 0, 0), (0, 1, 0), (0, 0, 1)], dirs_rotation=[]))
    ob.add_joint_task(abi.joint_task_desc(7))
    ob.tick(True)
    al = ob.get(abi.F_ALPHA, t0)[:, 0]
    pool = list(cand[(al > 0.05) & (al < 0.95)][:4]) + list(cand[al == 0][:1])
    want = len(pool)
    assert want >= 3
    tries = 0
    for q in pool:
        tries += 1
        dq = rng.uniform(-0.3, 0.3, 7)
        goal = dict(dpos=rng.normal(0, 0.05, 3), linvel=rng.normal(0, 0.05, 3), angvel=np.zeros(3), linacc=np.zeros(3), angacc=np.zeros(3))
        jg = dict(pos=q + rng.normal(0, 0.1, 7), vel=rng.normal(0, 0.1, 7), acc=rng.normal(0, 0.1, 7))
        try:
            o = blend_tick(model, q, dq, goal, jg, B6=B6)
        except AssertionError:
            continue
        o2 = blend_tick(model, q, dq, goal, jg, B6=B6, stacked=False)
        for k in ("tau", "tau1", "tau2", "S"):
            out[k].append(f64(o[k]).ravel())
        out["alpha"].append(float(o["alpha"])); out["ins"].append(o["ins"]); out["types"].append(o["types"]); out["tau_plain"].append(f64(o2["tau"]).ravel())
        for k, v in (("q", q), ("dq", dq), ("dpos", goal["dpos"]), ("gv", goal["linvel"]), ("gq", jg["pos"]), ("gdq", jg["vel"]), ("gddq", jg["acc"])):
            inp[k].append(v)
    assert len(out["tau"]) == want, (len(out["tau"]), tries)
    return dict(**{k: np.array(v) for k, v in inp.items()}, **{k: np.array(v) for k, v in out.items()})


def case_puma_blend(rng):
    """examples/19-puma_singularity: the 6R arm at its wrist-lock posture +- eps under ONE 6-DoF task (no joint task), bounded inertia
    estimates clamping the three wrist joints (the general clamp pattern)."""
    model = MpChain(chain("puma560_synthetic"))
    base = np.array([0.0, -np.pi / 4, 3 * np.pi / 4, np.pi / 2, 0.0, np.pi])
    want = 5
    out = {k: [] for k in ("tau", "alpha", "ins", "types", "S")}
    inp = {k: [] for k in ("q", "dq", "dpos", "gv")}
    tries = 0
    while len(out["tau"]) < want and tries < 400:
        tries += 1
        q = base + rng.normal(0, 0.05, 6)
        q[4] = rng.choice([-1.0, 1.0]) * 10.0 ** rng.uniform(-2.5, 0.0)
        dq = rng.uniform(-0.3, 0.3, 6)
        goal = dict(dpos=rng.normal(0, 0.05, 3), linvel=rng.normal(0, 0.05, 3), angvel=np.zeros(3), linacc=np.zeros(3), angacc=np.zeros(3))
        try:
            o = blend_tick(model, q, dq, goal, None, p_c=(0, 0, 0), jt=False)
        except AssertionError:
            continue
        inside = 0 < o["alpha"] < 1
        if (inside and sum(1 for a in out["alpha"] if 0 < a < 1) >= 3) or (not inside and sum(1 for a in out["alpha"] if not 0 < a < 1) >= 2):
            continue
        assert sum(1 for i in range(6) if model.mass_matrix([mp.mpf(float(t)) for t in q])[i, i] < mp.mpf("0.1")) >= 2   # a general clamp pattern
        out["tau"].append(f64(o["tau"]).ravel()); out["S"].append(f64(o["S"]).ravel())
        out["alpha"].append(float(o["alpha"])); out["ins"].append(o["ins"]); out["types"].append(o["types"])
        for k, v in (("q", q), ("dq", dq), ("dpos", goal["dpos"]), ("gv", goal["linvel"])):
            inp[k].append(v)
    assert len(out["tau"]) == want, len(out["tau"])
    return dict(**{k: np.array(v) for k, v in inp.items()}, **{k: np.array(v) for k, v in out.items()})


def lagrange_nonlinear(model, q, dq):
    """b(q, dq) + g(q) from Lagrange's equations: Mdot dq - 1/2 grad_q(dq' M dq) + grad_q V, V = -sum_b m_b g . c_b(q); M(q) is the
    kinetic-energy Hessian above, its q-derivative a second central difference (step 1e-12 over an M that is good to 1e-30)."""
    n = model.n
    H2 = mp.mpf(10) ** -12
    g = V3(model.c["gravity"])

    def V(qq):
        f = model.fk(qq)
        return -sum(mp.mpf(body["mass"]) * (g.T * (f[b][1] + f[b][0] * V3(body["com"])))[0] for b, body in enumerate(model.c["bodies"]))

    dM, dV = [], []
    for k in range(n):
        qp, qm = list(q), list(q)
        qp[k] = q[k] + H2; qm[k] = q[k] - H2
        dM.append((model.mass_matrix(qp) - model.mass_matrix(qm)) / (2 * H2))
        qp[k] = q[k] + H; qm[k] = q[k] - H
        dV.append((V(qp) - V(qm)) / (2 * H))
    v = mp.matrix(dq)
    Mdot = mp.zeros(n, n)
    for k in range(n):
        Mdot += dM[k] * dq[k]
    c = Mdot * v
    for k in range(n):
        c[k] -= (v.T * dM[k] * v)[0] / 2
    return c + mp.matrix(dV)


def case_panda_rollout(rng):
    """Row f-2: (A) b + g and M^-1 at six states for one step under the caller's own torque; (B) three symplectic-Euler substeps of the
    unactuated arm with joint damping, entirely in the model (examples/05...cpp:219-233 without contact)."""
    model = MpChain(chain("panda"))
    c = chain("panda")
    lo, hi = np.array([j["lower"] for j in c["joints"]]), np.array([j["upper"] for j in c["joints"]])
    A = 6
    q = lo + (0.15 + 0.7 * rng.random((A, 7))) * (hi - lo)
    dq = rng.uniform(-1.0, 1.0, (A, 7))
    h, Minv = [], []
    for r in range(A):
        qq, dd = [mp.mpf(float(t)) for t in q[r]], [mp.mpf(float(t)) for t in dq[r]]
        h.append(f64(lagrange_nonlinear(model, qq, dd)).ravel())
        Minv.append(f64(inv(model.mass_matrix(qq))))
    # (B) passive fall: tau = 0
    Bn, dt, damping, sub = 3, mp.mpf("0.001"), mp.mpf("0.2"), 3
    q0 = lo + (0.2 + 0.6 * rng.random((Bn, 7))) * (hi - lo)
    dq0 = rng.uniform(-0.5, 0.5, (Bn, 7))
    q3, dq3, h0 = [], [], []
    for r in range(Bn):
        qq, dd = [mp.mpf(float(t)) for t in q0[r]], [mp.mpf(float(t)) for t in dq0[r]]
        for s in range(sub):
            hh = lagrange_nonlinear(model, qq, dd)
            if s == 0:
                h0.append(f64(hh).ravel())
            qdd = inv(model.mass_matrix(qq)) * (-hh - damping * mp.matrix(dd))
            dd = [dd[i] + qdd[i] * dt for i in range(7)]
            qq = [qq[i] + dd[i] * dt for i in range(7)]
        q3.append([float(t) for t in qq]); dq3.append([float(t) for t in dd])
    return dict(q=q, dq=dq, h=np.array(h), Minv=np.array(Minv), fall_q0=q0, fall_dq0=dq0, fall_h0=np.array(h0), fall_q=np.array(q3), fall_dq=np.array(dq3),
                fall_dt=float(dt), fall_damping=float(damping), fall_substeps=sub)


def range_basis(A, tol=mp.mpf("1e-3")):
    """orthonormal basis of the column space of A (definition of matrixRangeBasis, App. B: relative tolerance 1e-3 on the singular
    values) by Gram-Schmidt on its largest columns; any basis of that space gives the same torques"""
    sv = mp.svd_r(A, compute_uv=False)
    if sv[0] < mp.mpf("1e-12"):
        return None
    r_dim = sum(1 for x in sv if x / sv[0] > tol)
    gap = [sv[i] / sv[0] for i in range(len(sv))]
    assert all(g > 10 * tol or g < tol / 10 for g in gap), "pin robots must have a clear rank gap"
    Rb = []
    for c in sorted(range(A.cols), key=lambda c: -mp.norm(A[:, c])):
        v = A[:, c]
        for b in Rb:
            v = v - b * (b.T * v)[0]
        if len(Rb) < r_dim and mp.norm(v) > mp.mpf("1e-6"):
            Rb.append(v / mp.norm(v))
    Rm = mp.matrix(A.rows, r_dim)
    for c, b in enumerate(Rb):
        Rm[:, c] = b
    return Rm


def hierarchy_tick(model, q, dq, tasks, stacked=True, link="end-effector", p_c=(0, 0, 0.07)):
    """RobotController.cpp:53-74 over any list of tasks, every MotionForceTask in its non-singular branch (asserted), bounded inertia
    estimates, no OTG.  tasks: ("jt", S k x n, goal dict pos/vel/acc of size k) | ("mft", B6 6 x m or None, goal dict).
    JointTask.cpp:218-356 with _joint_selection = S and N_prec != I; MotionForceTask.cpp:247-276 (disturbance compensation == 0: the
    _Lambda it multiplies by is only ever setZero, quirk Q1); SingularityHandler.cpp:75-160,300-309."""
    n = model.n
    qm, dqm = [mp.mpf(float(x)) for x in q], mp.matrix([mp.mpf(float(x)) for x in dq])
    qv = mp.matrix(qm)
    M = model.mass_matrix(qm)
    Minv, Mbinv = inv(M), inv(bie(M))
    Np = mp.eye(n)
    tau = mp.zeros(n, 1)
    per_task, ranks = [], []
    for t in tasks:
        if t[0] == "jt":
            S = mp.matrix([[mp.mpf(float(x)) for x in row] for row in t[1]])
            g = t[2]
            Jp = S * Np
            Rm = range_basis(Jp)
            if Rm is None:
                per_task.append(mp.zeros(n, 1)); ranks.append(0)
                continue
            J = Rm.T * Jp
            Lam, Jbar, N = osm(J, Minv)
            Lam_b = inv(J * Mbinv * J.T)
            qd, dqd, ddqd = (mp.matrix([mp.mpf(float(x)) for x in g[key]]) for key in ("pos", "vel", "acc"))
            u = -mp.mpf(50) * (S * qv - qd) - mp.mpf(14) * (S * dqm - dqd)
            tt = Jp.T * Rm * (Lam * (Rm.T * ddqd) + Lam_b * (Rm.T * u))
            if stacked:
                tt = tt - Jp.T * Rm * (Lam * (Rm.T * (S * (Minv * tau))))
            ranks.append(Rm.cols)
        else:
            B6 = mp.eye(6) if t[1] is None else t[1]
            goal = t[2]
            x, R, J6 = model.point(qm, link, p_c)
            P = B6 * B6.T
            J = P * J6
            tw = J * dqm
            g = dict(goal)
            g["pos"], g["ori"] = x + V3(goal["dpos"]), R
            Fstar, Ff, _, _, _ = mft_law(MftSpec(), new_state(), x, R, tw[0:3, 0], tw[3:6, 0], g)
            Jr = B6.T * (J * Np)
            sv = mp.svd_r(Jr, compute_uv=False)
            assert sv[len(sv) - 1] / sv[0] > mp.mpf("0.07"), "pin robots must sit in the non-singular branch"
            Lam, Jbar, N = osm(Jr, Minv)
            Lam_b = inv(Jr * Mbinv * Jr.T)
            tt = Jr.T * (Lam_b * (B6.T * Fstar) + B6.T * Ff)
            ranks.append(B6.cols)
        Np = N * Np
        tau = tau + tt
        per_task.append(tt)
    return dict(tau=tau, per_task=per_task, N_total=Np, ranks=ranks)


def case_hierarchies(rng):
    """Row f-4: three hierarchies RobotController accepts beyond [MotionForceTask, JointTask]:
    h1 examples/06: sliding-base Panda (n = 8), partial JointTask (slider + last joint) > 6-DoF MotionForceTask;
    h2 sliding base: partial JointTask (slider) > translation-only MotionForceTask > full JointTask;
    h3 Panda: translation-only MotionForceTask (examples/09) > partial JointTask (joints 1-3) > full JointTask;
    h0 PUMA: the full JointTask alone (examples/01-joint_control)."""
    out = {}
    e3 = mp.matrix(6, 3)
    for i in range(3):
        e3[i, i] = 1
    for name, cname, build in (("h1", "panda_sliding_base", lambda n: [("jt", np.eye(n)[[0, n - 1]]), ("mft", None)]),
                               ("h2", "panda_sliding_base", lambda n: [("jt", np.eye(n)[:1]), ("mft", e3), ("jt", np.eye(n))]),
                               ("h3", "panda", lambda n: [("mft", e3), ("jt", np.eye(n)[:3]), ("jt", np.eye(n))]),
                               ("h0", "puma560_synthetic", lambda n: [("jt", np.eye(n))])):   # examples/01: the joint task alone
        c = chain(cname)
        model = MpChain(c)
        n = c["n"]
        q0 = np.array([j["q0"] for j in c["joints"]], float)
        want, rec = 3, {k: [] for k in ("q", "dq", "dpos", "gv", "tau", "tau_plain", "N_total", "t0", "t1", "t2", "jg0", "jg1", "jgv0", "jgv1", "jga0", "jga1")}
        tries = 0
        while len(rec["q"]) < want and tries < 60:
            tries += 1
            q = q0 + rng.normal(0, 0.25, n)
            dq = rng.uniform(-0.4, 0.4, n)
            dpos, gv = rng.normal(0, 0.05, 3), rng.normal(0, 0.05, 3)
            tasks, jgs = [], []
            for kind, arg in build(n):
                if kind == "jt":
                    k = len(arg)
                    g = dict(pos=arg @ q + rng.normal(0, 0.1, k), vel=rng.normal(0, 0.1, k), acc=rng.normal(0, 0.1, k))
                    jgs.append(g)
                    tasks.append(("jt", arg, g))
                else:
                    tasks.append(("mft", arg, dict(dpos=dpos, linvel=gv)))
            try:
                o = hierarchy_tick(model, q, dq, tasks, True)
                o2 = hierarchy_tick(model, q, dq, tasks, False)
            except AssertionError:
                continue
            for k, v in (("q", q), ("dq", dq), ("dpos", dpos), ("gv", gv), ("tau", f64(o["tau"]).ravel()), ("tau_plain", f64(o2["tau"]).ravel()), ("N_total", f64(o["N_total"]))):
                rec[k].append(v)
            for i in range(3):
                rec["t%d" % i].append(f64(o["per_task"][i]).ravel() if i < len(tasks) else np.zeros(n))
            for i, g in enumerate(jgs):
                rec["jg%d" % i].append(g["pos"]); rec["jgv%d" % i].append(g["vel"]); rec["jga%d" % i].append(g["acc"])
            rec.setdefault("ranks", []).append(o["ranks"] + [0] * (3 - len(o["ranks"])))
        assert len(rec["q"]) == want, (name, len(rec["q"]))
        for k, v in rec.items():
            if len(v):
                out[name + "_" + k] = np.array(v)
    return out


def main():
    prove_orientation_error()
    print("orientationError(R_d, Rot(a, theta) R_d) = sin(theta) a: proved")
    os.makedirs(OUT, exist_ok=True)
    for name, fn, seed in (("pins_rrrr_two_task", case_rrrr, 1101), ("pins_panda_two_task", case_panda_two_task, 1102),
                           ("pins_panda_law", case_panda_law, 1103), ("pins_popc_window", case_popc_window, 1104),
                           ("pins_panda_blend", case_panda_blend, 1105), ("pins_puma_blend", case_puma_blend, 1106), ("pins_panda_blend_partial", case_panda_blend_partial, 1107),
                           ("pins_panda_rollout", case_panda_rollout, 1108), ("pins_hierarchies", case_hierarchies, 1109)):
        if len(sys.argv) > 1 and name not in sys.argv[1:]:
            continue
        d = fn(np.random.default_rng(seed))
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, {k: np.asarray(v).shape for k, v in d.items()})


if __name__ == "__main__":
    main()
