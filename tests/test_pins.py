"""Replay of the code-independent pins (tests/golden/pins_*.npz, made by tools/make_sympy_golden.py: sympy closed forms for the
planar 4R, 50-digit mpmath differentiation of the kinematic map / kinetic energy for the Panda, the control law typed in from
the reference text, hand-derived known answers) on
  * the C++ oracle            (CPU, -m "not gpu")
  * the numpy / LAPACK oracle (CPU, two-task cases)
  * the CUDA path through the C ABI (GPU, -m gpu), fused and API-faithful routes.
FP64, 1e-9 relative.  What these pin that the oracle-vs-oracle checks could not: the sign and frame of
orientationError, the Sigma projectors (dims 0-3, world / compliant frame), the velocity-saturation law, feed-forward
scaling (Q3), the bounded-inertia variants, Lambda / Jbar / N by their definitions, the joint task's range basis, the
stacking of RobotController, and one full passivity window of POPCExplicitForceControl."""
import os

import numpy as np
import pytest

import osc_abi as abi
from common import FP64_TOL, ROOT, load_chain, q0_of, rel_err


def G(name):
    return dict(np.load(os.path.join(ROOT, "tests", "golden", name + ".npz")))


def make_batch(kind, chain, B):
    if kind == "oracle":
        import orc
        orc.build()
        return orc.OracleBatch(chain, B, q0=q0_of(chain))
    import osc_b200
    return osc_b200.Batch(chain, B, q0=q0_of(chain))


KINDS = ["oracle", pytest.param("cuda", marks=pytest.mark.gpu)]


def close(a, b, what, tol=FP64_TOL, floor=None):
    b = np.asarray(b, float)
    e, where = rel_err(np.asarray(a, float).reshape(len(b), -1), b.reshape(len(b), -1), floor)
    assert e <= tol, "%s: rel err %.3e at robot %d" % (what, e, where)


def _two_task(kind, g, chain, link, p_c, stacked, per_call, **mft_kw):
    B = len(g["q"])
    b = make_batch(kind, chain, B)
    b.set(abi.F_Q, g["q"]); b.set(abi.F_DQ, g["dq"]); b.update_model()
    t0 = b.add_motion_force_task(abi.mft_desc(chain, link, p_c, **mft_kw))
    t1 = b.add_joint_task(abi.joint_task_desc(chain["n"]))
    x0 = b.get(abi.F_CURRENT_POSITION, t0)
    close(x0, g["x"], "FK position")
    close(b.get(abi.F_CURRENT_ORIENTATION, t0), g["R"], "FK orientation", floor=1.0)
    b.set(abi.F_GOAL_POSITION, g["x"] + g["dpos"], t0); b.set(abi.F_GOAL_ORIENTATION, g["goal_ori"], t0)
    b.set(abi.F_GOAL_LINEAR_VELOCITY, g["gv"], t0); b.set(abi.F_GOAL_ANGULAR_VELOCITY, g["gw"], t0)
    if "ga" in g:
        b.set(abi.F_GOAL_LINEAR_ACCELERATION, g["ga"], t0); b.set(abi.F_GOAL_ANGULAR_ACCELERATION, g["gal"], t0)
    b.set(abi.F_JT_GOAL_POSITION, g["gq"], t1); b.set(abi.F_JT_GOAL_VELOCITY, g["gdq"], t1); b.set(abi.F_JT_GOAL_ACCELERATION, g["gddq"], t1)
    if per_call:
        b.update_model(); b.update_task_models()
        tau = b.compute_torques(stacked)
    else:
        tau = b.tick(stacked)
    return b, t0, t1, tau


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("per_call", [False, True])
def test_pin_planar_4r_partial_task_sympy(kind, per_call):
    """examples/11: partial MotionForceTask (X, Y, rot Z) + JointTask on the planar 4R against the sympy Lagrange closed forms"""
    g = G("pins_rrrr_two_task")
    chain = load_chain("rrrrbot")
    kw = dict(dirs_translation=[(1, 0, 0), (0, 1, 0)], dirs_rotation=[(0, 0, 1)])
    b, t0, t1, tau = _two_task(kind, g, chain, "link4", (0.5, 0, 0), True, per_call, **kw)
    close(tau, g["tau"], "tau (stacked)")
    close(b.get(abi.F_M), g["M"], "M")
    close(b.get(abi.F_JACOBIAN, t0), g["J"], "J")
    close(b.get(abi.F_LAMBDA_WORLD, t0), g["lam_world"], "Lambda")
    close(b.get(abi.F_LAMBDA_MODIFIED_WORLD, t0), g["lam_mod_world"], "Lambda (bounded inertia)")
    close(b.get(abi.F_N_TASK, t0), g["N1"], "N_1", floor=1.0)
    close(b.get(abi.F_N_TOTAL), g["N_total"], "N_total", floor=1.0)
    close(b.get(abi.F_TASK_TORQUES, t0), g["tau1"], "tau_1")
    close(b.get(abi.F_TASK_TORQUES, t1), g["tau2"], "tau_2")
    close(b.get(abi.F_UNIT_MASS_FORCE, t0), g["Fstar"], "F*")
    close(b.get(abi.F_ORIENTATION_ERROR, t0), g["ori_err"], "orientation error", floor=1e-3)
    b2, _, _, tau_plain = _two_task(kind, g, chain, "link4", (0.5, 0, 0), False, per_call, **kw)
    close(tau_plain, g["tau_plain"], "tau (plain sum)")


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("per_call", [False, True])
def test_pin_panda_two_task_mpmath(kind, per_call):
    """BASELINE config 2 shape against the 50-digit differentiation model (per_call = False is the fused tick on the GPU)"""
    g = G("pins_panda_two_task")
    chain = load_chain("panda")
    b, t0, t1, tau = _two_task(kind, g, chain, "end-effector", (0, 0, 0.07), True, per_call)
    close(tau, g["tau"], "tau (stacked)")
    close(b.get(abi.F_TASK_TORQUES, t0), g["tau1"], "tau_1")
    close(b.get(abi.F_TASK_TORQUES, t1), g["tau2"], "tau_2")
    close(b.get(abi.F_UNIT_MASS_FORCE, t0), g["Fstar"], "F*")
    close(b.get(abi.F_ORIENTATION_ERROR, t0), g["ori_err"], "orientation error = sin(theta) a", floor=1e-3)
    close(b.get(abi.F_CURRENT_LINEAR_VELOCITY, t0), g["v"], "v = J dq")
    close(b.get(abi.F_CURRENT_ANGULAR_VELOCITY, t0), g["w"], "w = J dq")
    close(b.get(abi.F_M), g["M"], "M")
    close(b.get(abi.F_MINV), g["Minv"], "M^-1")
    close(b.get(abi.F_JACOBIAN, t0), g["J"], "J")
    close(b.get(abi.F_LAMBDA_WORLD, t0), g["lam_world"], "Lambda")
    close(b.get(abi.F_LAMBDA_MODIFIED_WORLD, t0), g["lam_mod_world"], "Lambda (bounded inertia)")
    close(b.get(abi.F_N_TASK, t0), g["N1"], "N_1", floor=1.0)
    close(b.get(abi.F_N_TOTAL), g["N_total"], "N_total", floor=1.0)
    close(b.get(abi.F_JT_LAMBDA_WORLD, t1), g["jt_lam_world"], "joint task R Lambda R^T")
    assert (np.asarray(b.get(abi.F_JT_RANGE_RANK, t1)).ravel() == 1).all() and (np.asarray(b.get(abi.F_ALPHA, t0)).ravel() == 1).all()
    if kind == "cuda":
        assert b.tick_is_fused() and (per_call or b.tick_fallback_count() == 0), "the pins must exercise the fused fast path"
    _, _, _, tau_plain = _two_task(kind, g, chain, "end-effector", (0, 0, 0.07), False, per_call)
    close(tau_plain, g["tau_plain"], "tau (plain sum)")


def test_pin_two_task_numpy_oracle():
    """the second (numpy / LAPACK) oracle against the same pins"""
    import np_oracle as npo
    for name, chain_name, link, p_c, kw in (("pins_rrrr_two_task", "rrrrbot", "link4", (0.5, 0, 0),
                                             dict(dirs_t=[np.array([1.0, 0, 0]), np.array([0, 1.0, 0])], dirs_r=[np.array([0, 0, 1.0])])),
                                            ("pins_panda_two_task", "panda", "end-effector", (0, 0, 0.07), {})):
        g, chain = G(name), load_chain(chain_name)
        for r in range(len(g["q"])):
            m = npo.Model(chain); m.setQ(g["q"][r]); m.setDq(g["dq"][r]); m.updateModel()
            t0 = npo.MotionForceTask(m, link, np.array(p_c, float), **kw); t1 = npo.JointTask(m)
            t0.goal_pos = g["x"][r] + g["dpos"][r]; t0.goal_ori = g["goal_ori"][r].reshape(3, 3)
            t0.goal_linvel = g["gv"][r]; t0.goal_angvel = g["gw"][r]
            if "ga" in g:
                t0.goal_linacc = g["ga"][r]; t0.goal_angacc = g["gal"][r]
            t1.goal_pos = g["gq"][r]; t1.goal_vel = g["gdq"][r]; t1.goal_acc = g["gddq"][r]
            ctl = npo.RobotController(m, [t0, t1])
            ctl.updateControllerTaskModels()
            tau = ctl.computeControlTorques()
            close(tau[None], g["tau"][r][None], "%s robot %d tau" % (name, r))
            close(m.M().ravel()[None], g["M"][r][None], "M")


LAW = {  # name -> osc_mft_params overrides (tools/make_sympy_golden.py LAW_CASES)
    "force1_z_world": dict(force_space_dimension=1, force_or_motion_axis=(0, 0, 1)),
    "force2_x_compliant": dict(force_space_dimension=2, force_or_motion_axis=(1, 0, 0)),
    "force3_moment1": dict(force_space_dimension=3, moment_space_dimension=1, moment_or_rotmotion_axis=(0, 1, 0)),
    "moment2_compliant": dict(moment_space_dimension=2, moment_or_rotmotion_axis=(0.6, 0, 0.8)),
    "moment3": dict(moment_space_dimension=3),
    "vel_sat_saturated": dict(use_velocity_saturation=1, linear_saturation_velocity=0.01, angular_saturation_velocity=0.01),
    "vel_sat_unsaturated": dict(use_velocity_saturation=1, linear_saturation_velocity=50.0, angular_saturation_velocity=50.0),
    "anisotropic_gains_ki": dict(kp_pos=(100.0, 50.0, 150.0), kv_pos=(20.0, 10.0, 30.0), ki_pos=(3.0, 4.0, 5.0), kp_ori=(200.0, 100.0, 50.0),
                                 kv_ori=(28.3, 20.0, 10.0), ki_ori=(1.0, 2.0, 3.0)),
    "closed_loop_force_moment": dict(force_space_dimension=1, force_or_motion_axis=(0, 0, 1), moment_space_dimension=1, moment_or_rotmotion_axis=(1, 0, 0),
                                     closed_loop_force_control=1, closed_loop_moment_control=1),
    "closed_loop_saturating_feedback": dict(force_space_dimension=3, closed_loop_force_control=1, max_force_control_feedback_output=2.0),
}
COMPLIANT = {"force2_x_compliant", "moment2_compliant"}


def _set_params(b, t0, kw):
    p = b.mft_params(t0)
    for k, v in kw.items():
        cur = getattr(p, k)
        if hasattr(cur, "__len__"):
            for i, x in enumerate(v):
                cur[i] = float(x)
        else:
            setattr(p, k, v)
    b.set_mft_params(t0, p)


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("route", ["per_call", "tick"])
def test_pin_control_law_known_answers(kind, route):
    """one Panda per case, a single MotionForceTask, two ticks (integrators): Sigma projectors, velocity saturation, gains,
    closed-loop force / moment feedback and its saturation, feed-forward scaling.  Route "tick" on the GPU: the fused tick of a 7-DoF
    arm under one task (k_tick_fused2<.., SINGLE>, general and SIMPLE instantiations)"""
    g = G("pins_panda_law")
    chain = load_chain("panda")
    for r, name in enumerate(g["names"]):
        name = str(name)
        b = make_batch(kind, chain, 1)
        b.set(abi.F_Q, g["q"][r][None]); b.set(abi.F_DQ, g["dq"][r][None]); b.update_model()
        t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07), in_compliant_frame=name in COMPLIANT))
        _set_params(b, t0, LAW[name])   # parametrize* resets the goal pose to the current one: set the goals afterwards
        b.set(abi.F_GOAL_POSITION, g["goal_pos"][r][None], t0); b.set(abi.F_GOAL_ORIENTATION, g["goal_ori"][r][None], t0)
        b.set(abi.F_GOAL_FORCE, g["goal_force"][r][None], t0); b.set(abi.F_GOAL_MOMENT, g["goal_moment"][r][None], t0)
        for tick in range(2):
            b.set(abi.F_Q, (g["q"][r] + g["dq"][r] * 0.001 * tick)[None])
            if route == "per_call":
                b.update_model(); b.update_task_models()
                b.update_sensed_force_and_moment(t0, g["sensed_f"][r, tick][None], g["sensed_m"][r, tick][None])
                tau = b.compute_torques(True)
            else:
                b.update_model()
                b.update_sensed_force_and_moment(t0, g["sensed_f"][r, tick][None], g["sensed_m"][r, tick][None])
                tau = b.tick(True)
            close(tau, g["tau_t%d" % tick][r][None], "%s tick %d tau" % (name, tick))
        close(b.get(abi.F_UNIT_MASS_FORCE, t0), g["Fstar_t1"][r][None], name + " F*")
        close(b.get(abi.F_INTEGRATED_FORCE_ERROR, t0), g["int_f"][r][None], name + " integrated force error", floor=1e-6)
        close(b.get(abi.F_INTEGRATED_POSITION_ERROR, t0), g["int_p"][r][None], name + " integrated position error", floor=1e-9)
        if "vel_sat" in name:
            close(b.get(abi.F_DESIRED_LINEAR_VELOCITY, t0), g["des_linvel"][r][None], name + " saturated desired velocity")
            close(b.get(abi.F_DESIRED_ANGULAR_VELOCITY, t0), g["des_angvel"][r][None], name + " saturated desired angular velocity")
            n = np.linalg.norm(g["des_linvel"][r])
            assert (abs(n - 0.01) < 1e-12) if name == "vel_sat_saturated" else (n < 50.0)


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("route", ["per_call", "tick", "tick_single"])
def test_pin_popc_window(kind, route):
    """520 ticks of closed-loop force control with the passivity observer / controller at a fixed posture: the 250-sample window
    slides, the 50-tick controller period fires 10 times, R_c leaves 1 and recovers (POPCExplicitForceControl.cpp:30-96)"""
    g = G("pins_popc_window")
    chain = load_chain("panda")
    b = make_batch(kind, chain, 1)
    b.set(abi.F_Q, g["q"][None]); b.set(abi.F_DQ, g["dq"][None]); b.update_model()
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
    t1 = b.add_joint_task(abi.joint_task_desc(7)) if route == "tick" else None   # "tick": two-task fused tick; "tick_single": the task alone
    _set_params(b, t0, dict(force_space_dimension=1, force_or_motion_axis=(0, 0, 1), closed_loop_force_control=1, passivity_enabled=1))
    b.set(abi.F_GOAL_FORCE, np.array([[0, 0, -5.0]]), t0)
    if t1 is not None:
        _set_jt_zero_gains(b, t1)
    worst, rc_err, po_err = 0.0, 0.0, 0.0
    for k in range(len(g["sensed_f"])):
        if route == "per_call":
            b.update_model(); b.update_task_models()
            b.update_sensed_force_and_moment(t0, g["sensed_f"][k][None], np.zeros((1, 3)))
            b.compute_torques(True)
        else:
            b.update_sensed_force_and_moment(t0, g["sensed_f"][k][None], np.zeros((1, 3)))
            b.tick(True)
        tau = np.asarray(b.get(abi.F_TASK_TORQUES, t0), float)
        worst = max(worst, rel_err(tau, g["tau"][k][None])[0])
        rc_err = max(rc_err, abs(float(b.get(abi.F_POPC_RC, t0)[0, 0]) - g["rc"][k]))
        po_err = max(po_err, abs(float(b.get(abi.F_POPC_OBSERVER, t0)[0, 0]) - g["po"][k]) / max(1e-6, abs(g["po"][k])))
    assert worst <= FP64_TOL, worst
    assert rc_err <= 1e-9 and po_err <= 1e-9, (rc_err, po_err)
    assert g["rc"].min() == 0 and 0 < g["rc"][199] < 1 and g["rc"][-1] > 0.5   # drops, bottoms out, recovers
    close(b.get(abi.F_INTEGRATED_FORCE_ERROR, t0), g["int_f"][-1][None], "integrated force error")


def _set_jt_zero_gains(b, t1):
    p = b.joint_task_params(t1)
    for i in range(7):
        p.kp[i], p.kv[i], p.ki[i] = 0.0, 0.0, 0.0
    b.set_joint_task_params(t1, p)


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("per_call", [False, True])
def test_pin_panda_singular_blend_mpmath(kind, per_call):
    """BASELINE config 4 shape: six Panda postures INSIDE the blend band of SingularityHandler (elbow extended / wrist aligned), both
    joint-space strategies, against a 50-digit model of SingularityHandler.cpp:75-368 typed in from the reference text
    (tools/make_sympy_golden.py::blend_tick: mpmath SVD, every inverse by LU).  On the GPU per_call = False is the fused tick, whose
    certificate sends these robots to the closed-form blend routine (csrc/osc_blend.cuh); per_call = True the API-faithful kernels.
    Tolerance 1e-9 on alpha / singular values / the joint-task part, kappa-scaled on the torques of the singular block, which all
    three implementations obtain by inverting a matrix of condition 1 / (s_ins / s_0)^2."""
    g = G("pins_panda_blend")
    chain = load_chain("panda")
    B = len(g["q"])
    outs = {}
    for stacked in (True, False):
        b = make_batch(kind, chain, B)
        b.set(abi.F_Q, g["q"]); b.set(abi.F_DQ, g["dq"]); b.update_model()
        t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
        t1 = b.add_joint_task(abi.joint_task_desc(7))
        x0 = b.get(abi.F_CURRENT_POSITION, t0)
        b.set(abi.F_GOAL_POSITION, x0 + g["dpos"], t0); b.set(abi.F_GOAL_LINEAR_VELOCITY, g["gv"], t0)
        b.set(abi.F_JT_GOAL_POSITION, g["gq"], t1); b.set(abi.F_JT_GOAL_VELOCITY, g["gdq"], t1); b.set(abi.F_JT_GOAL_ACCELERATION, g["gddq"], t1)
        if per_call:
            b.update_model(); b.update_task_models()
            tau = b.compute_torques(stacked)
        else:
            tau = b.tick(stacked)
        outs[stacked] = (b, t0, t1, np.asarray(tau, float))
    b, t0, t1, tau = outs[True]
    ratio = g["S"][np.arange(B), g["ins"]] / g["S"][:, 0]
    tol = np.maximum(FP64_TOL, 64 * 2.2e-16 / ratio ** 2)            # per robot
    assert tol.max() < 1e-7
    close(b.get(abi.F_SINGULAR_VALUES, t0), g["S"], "singular values of J")
    assert np.abs(np.asarray(b.get(abi.F_ALPHA, t0)).ravel() - g["alpha"]).max() <= 1e-9
    assert (np.asarray(b.get(abi.F_NS_RANK, t0)).ravel() == g["ins"]).all()
    assert (np.asarray(b.get(abi.F_SINGULARITY_TYPES, t0)) == g["types"]).all(), "type 1 / type 2 classification"

    def close_each(a, ref, what):
        a, ref = np.asarray(a, float).reshape(B, -1), np.asarray(ref, float).reshape(B, -1)
        e = np.abs(a - ref).max(axis=1) / np.maximum(np.abs(ref).max(axis=1), 1e-3 * np.abs(ref).max())
        assert (e <= tol).all(), "%s: rel err %s > %s" % (what, e, tol)
    close_each(tau, g["tau"], "tau (stacked)")
    close_each(b.get(abi.F_TASK_TORQUES, t0), g["tau1"], "tau_1 = tau_ns + alpha tau_s + (1 - alpha) tau_joint")
    close_each(b.get(abi.F_TASK_TORQUES, t1), g["tau2"], "tau_2")
    close_each(outs[False][3], g["tau_plain"], "tau (plain sum)")
    close_each(b.get(abi.F_N_TASK, t0), g["N_t"], "N = N_post N_ns")
    assert (np.asarray(b.get(abi.F_JT_RANGE_RANK, t1)).ravel() == 1).all()
    if kind == "cuda" and not per_call:
        assert b.tick_is_fused() and b.tick_fallback_count() == B, "the pins must go through k_tick_tail"


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("per_call", [False, True])
def test_pin_puma_single_task_blend_mpmath(kind, per_call):
    """examples/19 / BASELINE config 4 (PUMA): a 6R arm at its wrist-lock posture under ONE 6-DoF MotionForceTask, bounded inertia
    estimates clamping several joints, against the same 50-digit model of SingularityHandler (no joint task below).  On the GPU
    per_call = False is the single-task fused tick (k_tick_fused2<6> with L_BIE for any clamp pattern + the n = 6 blend routine)."""
    g = G("pins_puma_blend")
    chain = load_chain("puma560_synthetic")
    B = len(g["q"])
    b = make_batch(kind, chain, B)
    b.set(abi.F_Q, g["q"]); b.set(abi.F_DQ, g["dq"]); b.update_model()
    t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0)))
    x0 = b.get(abi.F_CURRENT_POSITION, t0)
    b.set(abi.F_GOAL_POSITION, x0 + g["dpos"], t0); b.set(abi.F_GOAL_LINEAR_VELOCITY, g["gv"], t0)
    if per_call:
        b.update_model(); b.update_task_models()
        tau = b.compute_torques(True)
    else:
        tau = b.tick(True)
    ratio = np.maximum(g["S"][np.arange(B), g["ins"]] / g["S"][:, 0], 1e-4)
    tol = np.maximum(FP64_TOL, 64 * 2.2e-16 / ratio ** 2)
    close(b.get(abi.F_SINGULAR_VALUES, t0), g["S"], "singular values of J", floor=1e-3)
    assert np.abs(np.asarray(b.get(abi.F_ALPHA, t0)).ravel() - g["alpha"]).max() <= 1e-9
    assert (np.asarray(b.get(abi.F_NS_RANK, t0)).ravel() == g["ins"]).all()
    assert (np.asarray(b.get(abi.F_SINGULARITY_TYPES, t0)) == g["types"]).all()
    a, ref = np.asarray(tau, float), g["tau"]
    e = np.abs(a - ref).max(axis=1) / np.maximum(np.abs(ref).max(axis=1), 1e-3 * np.abs(ref).max())
    assert (e <= tol).all(), "tau: rel err %s > %s" % (e, tol)
    if kind == "cuda" and not per_call:
        assert b.tick_is_fused() and b.tick_fallback_count() == B


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("per_call", [False, True])
def test_pin_panda_partial_task_blend_mpmath(kind, per_call):
    """BASELINE config 3 shape inside ITS OWN singular band: translation-only MotionForceTask (rank 3, nullspace of dimension 4) + full
    JointTask on Panda postures near full extension, against the 50-digit model of SingularityHandler for a partial task (the SVD of
    the task in its range, U = B6 U_r).  GPU, per_call = False: k_tick_fused3's certificate -> closed-form blend routine, M = 3."""
    g = G("pins_panda_blend_partial")
    chain = load_chain("panda")
    B = len(g["q"])
    outs = {}
    for stacked in (True, False):
        b = make_batch(kind, chain, B)
        b.set(abi.F_Q, g["q"]); b.set(abi.F_DQ, g["dq"]); b.update_model()
        t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07), dirs_translation=[(1, 0, 0), (0, 1, 0), (0, 0, 1)], dirs_rotation=[]))
        t1 = b.add_joint_task(abi.joint_task_desc(7))
        x0 = b.get(abi.F_CURRENT_POSITION, t0)
        b.set(abi.F_GOAL_POSITION, x0 + g["dpos"], t0); b.set(abi.F_GOAL_LINEAR_VELOCITY, g["gv"], t0)
        b.set(abi.F_JT_GOAL_POSITION, g["gq"], t1); b.set(abi.F_JT_GOAL_VELOCITY, g["gdq"], t1); b.set(abi.F_JT_GOAL_ACCELERATION, g["gddq"], t1)
        if per_call:
            b.update_model(); b.update_task_models()
            tau = b.compute_torques(stacked)
        else:
            tau = b.tick(stacked)
        outs[stacked] = (b, t0, t1, np.asarray(tau, float))
    b, t0, t1, tau = outs[True]
    ratio = g["S"][np.arange(B), g["ins"]] / g["S"][:, 0]
    tol = np.maximum(FP64_TOL, 64 * 2.2e-16 / ratio ** 2)
    assert np.abs(np.asarray(b.get(abi.F_SINGULAR_VALUES, t0))[:, :3] - g["S"]).max() <= 1e-9
    assert np.abs(np.asarray(b.get(abi.F_ALPHA, t0)).ravel() - g["alpha"]).max() <= 1e-9
    assert (np.asarray(b.get(abi.F_NS_RANK, t0)).ravel() == g["ins"]).all()
    assert (np.asarray(b.get(abi.F_SINGULARITY_TYPES, t0)) == g["types"]).all()

    def close_each(a, ref, what):
        a, ref = np.asarray(a, float).reshape(B, -1), np.asarray(ref, float).reshape(B, -1)
        e = np.abs(a - ref).max(axis=1) / np.maximum(np.abs(ref).max(axis=1), 1e-3 * np.abs(ref).max())
        assert (e <= tol).all(), "%s: rel err %s > %s" % (what, e, tol)
    close_each(tau, g["tau"], "tau (stacked)")
    close_each(b.get(abi.F_TASK_TORQUES, t0), g["tau1"], "tau_1")
    close_each(b.get(abi.F_TASK_TORQUES, t1), g["tau2"], "tau_2 (4-dimensional nullspace)")
    close_each(outs[False][3], g["tau_plain"], "tau (plain sum)")
    assert (np.asarray(b.get(abi.F_JT_RANGE_RANK, t1)).ravel() == 4).all()
    if kind == "cuda" and not per_call:
        assert b.tick_is_fused() and b.tick_fallback_count() == B


@pytest.mark.parametrize("kind", KINDS)
def test_pin_rollout_lagrange_mpmath(kind):
    """Row f-2 (forward dynamics + symplectic Euler, examples/05...cpp:219-233 without contact) against Lagrange's equations at 50
    digits: b + g = Mdot dq - 1/2 grad_q(dq' M dq) + grad_q V, with M the kinetic-energy Hessian -- no Newton-Euler recursion in the
    model.  (A) one step under the torque of a two-task tick: h, and q / dq advanced with the model's M^-1 and h under the torque the
    implementation itself reports; (B) three damped substeps of the unactuated arm, every number from the model."""
    g = G("pins_panda_rollout")
    chain = load_chain("panda")
    b = make_batch(kind, chain, len(g["q"]))
    b.set(abi.F_Q, g["q"]); b.set(abi.F_DQ, g["dq"]); b.update_model()
    b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07))); b.add_joint_task(abi.joint_task_desc(7))
    tau = np.asarray(b.tick(True), float)
    dt, damping = 1e-3, 0.3
    b.rollout_step(dt, 1, damping)
    close(b.get(abi.F_NONLINEAR_EFFECTS), g["h"], "b + g")
    dq1 = g["dq"] + dt * np.einsum("rij,rj->ri", g["Minv"], tau - g["h"] - damping * g["dq"])
    close(b.get(abi.F_DQ), dq1, "dq after one step")
    close(b.get(abi.F_Q), g["q"] + dt * dq1, "q after one step")

    f = make_batch(kind, chain, len(g["fall_q0"]))
    f.set(abi.F_Q, g["fall_q0"]); f.set(abi.F_DQ, g["fall_dq0"]); f.update_model()
    f.rollout_step(float(g["fall_dt"]), int(g["fall_substeps"]), float(g["fall_damping"]))   # no tick yet: OSC_F_TAU is zero
    close(f.get(abi.F_NONLINEAR_EFFECTS), g["fall_h0"], "b + g of the falling arm")
    close(f.get(abi.F_DQ), g["fall_dq"], "dq after three substeps")
    close(f.get(abi.F_Q), g["fall_q"], "q after three substeps")


_HIER = {  # name -> (chain, task list): what RobotController accepts beyond [MotionForceTask, JointTask]
    "h0": ("puma560_synthetic", lambda n: [("jt", None)]),                                                        # examples/01
    "h1": ("panda_sliding_base", lambda n: [("jt", np.eye(n)[[0, n - 1]]), ("mft", {})]),                      # examples/06
    "h2": ("panda_sliding_base", lambda n: [("jt", np.eye(n)[:1]), ("mft", dict(dirs_translation=[(1, 0, 0), (0, 1, 0), (0, 0, 1)], dirs_rotation=[])), ("jt", None)]),
    "h3": ("panda", lambda n: [("mft", dict(dirs_translation=[(1, 0, 0), (0, 1, 0), (0, 0, 1)], dirs_rotation=[])), ("jt", np.eye(n)[:3]), ("jt", None)]),
}


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("per_call", [False, True])
@pytest.mark.parametrize("name", sorted(_HIER))
def test_pin_hierarchies_mpmath(kind, per_call, name):
    """Partial JointTask first (examples/06, n = 8 with a prismatic base), three-task hierarchies with a partial MotionForceTask in
    the middle or in front: the N_prec chain, JointTask.cpp:218-356 with a selection matrix, the running-sum disturbance
    compensation of the joint tasks and the (identically zero, quirk Q1) one of the motion task, from the 50-digit model."""
    G_ = G("pins_hierarchies")
    g = {k[len(name) + 1:]: v for k, v in G_.items() if k.startswith(name + "_")}
    cname, build = _HIER[name]
    chain = load_chain(cname)
    n = chain["n"]
    b = make_batch(kind, chain, len(g["q"]))
    b.set(abi.F_Q, g["q"]); b.set(abi.F_DQ, g["dq"]); b.update_model()
    ids, j = [], 0
    for what, arg in build(n):
        if what == "jt":
            t = b.add_joint_task(abi.joint_task_desc(n, arg))
            b.set(abi.F_JT_GOAL_POSITION, g["jg%d" % j], t); b.set(abi.F_JT_GOAL_VELOCITY, g["jgv%d" % j], t); b.set(abi.F_JT_GOAL_ACCELERATION, g["jga%d" % j], t)
            j += 1
        else:
            t = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07), **arg))
            b.set(abi.F_GOAL_POSITION, b.get(abi.F_CURRENT_POSITION, t) + g["dpos"], t); b.set(abi.F_GOAL_LINEAR_VELOCITY, g["gv"], t)
        ids.append(t)
    for stacked, key in ((True, "tau"), (False, "tau_plain")):
        if per_call:
            b.update_model(); b.update_task_models()
            tau = b.compute_torques(stacked)
        else:
            tau = b.tick(stacked)
        close(tau, g[key], key)
        if stacked:
            for i, t in enumerate(ids):
                close(b.get(abi.F_TASK_TORQUES, t), g["t%d" % i], "torques of task %d" % i, floor=float(np.abs(g["tau"]).max()))
            close(b.get(abi.F_N_TOTAL), g["N_total"], "N_total", floor=1.0)
    assert [int(r) for r in g["ranks"][0][:len(ids)]] == {"h0": [6], "h1": [2, 6], "h2": [1, 3, 4], "h3": [3, 3, 1]}[name]
