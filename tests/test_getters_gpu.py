"""The getters of the reference that are functions of the task state (missing #5): sigmaForce / sigmaPosition /
sigmaMoment / sigmaOrientation (MotionForceTask.h:611-614, .cpp:892-971), getPositionError / getOrientationError
(.cpp:540-546), getGoalForce / getGoalMoment (.cpp:755-769), goalPositionReached / goalOrientationReached (.cpp:548-578;
JointTask.cpp:437-446), posSelectionProjector / oriSelectionProjector (.h:653-659), JointTask::setGainsUnsafe (.cpp:136-156).
Expected values are the reference's formulas evaluated with numpy on quantities read from the ORACLE (poses, goals, errors)."""
import numpy as np
import pytest

import osc_abi as abi
from common import FP64_TOL, Pair, load_chain, q0_of, random_q, rel_err

pytestmark = pytest.mark.gpu


def _sigma(dim, axis, Rot, P):
    if dim == 0:
        return np.zeros((len(Rot), 3, 3))
    if dim == 3:
        return np.broadcast_to(P @ P.T, (len(Rot), 3, 3))
    a = Rot @ np.asarray(axis, float)
    aa = a[:, :, None] * a[:, None, :]
    mid = aa if dim == 1 else np.eye(3)[None] - aa
    return P[None] @ mid @ P.T[None]


@pytest.mark.parametrize("compliant", [False, True])
@pytest.mark.parametrize("partial", [False, True])
def test_state_function_getters(oracle_built, compliant, partial):
    chain = load_chain("panda")
    n, B = 7, 300
    rng = np.random.default_rng(90 + compliant + 2 * partial)
    q = random_q(chain, B, rng, 0.3); dq = rng.uniform(-0.3, 0.3, (B, n))
    p = Pair(chain, B, q0=q0_of(chain))
    p.set(abi.F_Q, q); p.set(abi.F_DQ, dq); p.both("update_model")
    kw = dict(dirs_translation=[(1, 0, 0), (0, 0, 1)], dirs_rotation=[(0, 1, 0), (0, 0, 1)]) if partial else {}
    t0 = p.add_mft("end-effector", (0, 0, 0.07), in_compliant_frame=compliant, **kw)
    t1 = p.add_jt()
    ax_f, ax_m = np.array([0.6, 0.0, 0.8]), np.array([0.0, 1.0, 0.0])
    p.set_mft_params(t0, force_space_dimension=1, force_or_motion_axis=ax_f, moment_space_dimension=2, moment_or_rotmotion_axis=ax_m)
    x = p.ref.get(abi.F_CURRENT_POSITION, t0)
    gpos = x + rng.normal(0, 0.05, (B, 3)); gpos[: B // 3] = x[: B // 3] + 1e-4      # a third of the robots at their goal
    gori = p.ref.get(abi.F_CURRENT_ORIENTATION, t0).copy()
    p.set(abi.F_GOAL_POSITION, gpos, t0)
    p.set(abi.F_GOAL_FORCE, np.array([1.0, -2.0, 3.0]), t0); p.set(abi.F_GOAL_MOMENT, np.array([0.1, 0.2, -0.3]), t0)
    gq = q + rng.normal(0, 0.1, (B, n)); gq[::2] = q[::2] + 1e-5
    p.set(abi.F_JT_GOAL_POSITION, gq, t1)
    p.both("tick", True)
    g = p.gpu
    R = p.ref.get(abi.F_CURRENT_ORIENTATION, t0).reshape(B, 3, 3)
    Rot = R if compliant else np.broadcast_to(np.eye(3), (B, 3, 3))
    Pt, Pr = g.selection_projectors(t0)
    if partial:
        assert np.allclose(Pt, np.diag([1.0, 0, 1.0])) and np.allclose(Pr, np.diag([0, 1.0, 1.0]))
    else:
        assert np.array_equal(Pt, np.eye(3)) and np.array_equal(Pr, np.eye(3))
    Sf, Sm = _sigma(1, ax_f, Rot, Pt), _sigma(2, ax_m, Rot, Pr)
    Sp, So = Pt[None] @ (np.eye(3)[None] - Sf) @ Pt.T[None], Pr[None] @ (np.eye(3)[None] - Sm) @ Pr.T[None]
    for f, want in ((abi.F_SIGMA_FORCE, Sf), (abi.F_SIGMA_POSITION, Sp), (abi.F_SIGMA_MOMENT, Sm), (abi.F_SIGMA_ORIENTATION, So)):
        assert rel_err(g.get(f, t0), want.reshape(B, 9), 1.0)[0] <= FP64_TOL, f
    e_pos = gpos - p.ref.get(abi.F_CURRENT_POSITION, t0)
    e_ori = p.ref.get(abi.F_ORIENTATION_ERROR, t0)
    pe = np.einsum("bij,bj->bi", Sp, e_pos); oe = np.einsum("bij,bj->bi", So, e_ori)
    assert rel_err(g.get(abi.F_POSITION_ERROR, t0), pe, 1e-3)[0] <= FP64_TOL
    assert rel_err(g.get(abi.F_SELECTED_ORIENTATION_ERROR, t0), oe, 1e-3)[0] <= FP64_TOL
    assert rel_err(g.get(abi.F_GOAL_FORCE_WORLD, t0), Rot @ np.array([1.0, -2.0, 3.0]), 1e-3)[0] <= FP64_TOL
    assert rel_err(g.get(abi.F_GOAL_MOMENT_WORLD, t0), Rot @ np.array([0.1, 0.2, -0.3]), 1e-3)[0] <= FP64_TOL
    npos = np.sqrt(np.maximum(np.einsum("bi,bi->b", e_pos, pe), 0)); nori = np.sqrt(np.maximum(np.einsum("bi,bi->b", e_ori, oe), 0))
    assert np.abs(g.get(abi.F_POSITION_ERROR_NORM, t0)[:, 0] - npos).max() <= 1e-12
    assert np.abs(g.get(abi.F_ORIENTATION_ERROR_NORM, t0)[:, 0] - nori).max() <= 1e-12
    flags, cnt = g.goal_reached(t0, 1e-3, 0)
    assert np.array_equal(flags, (npos < 1e-3).astype(np.int32)) and cnt == int((npos < 1e-3).sum()) and 0 < cnt < B
    flags, cnt = g.goal_reached(t0, 1e-6, 1)
    assert np.array_equal(flags, (nori < 1e-6).astype(np.int32)) and cnt == int((nori < 1e-6).sum())
    # JointTask::goalPositionReached: | R R^T (q - q_goal) | over the task's current range (a rank n - m direction set here)
    W = p.ref.get(abi.F_JT_LAMBDA_WORLD, t1).reshape(B, n, n)           # R Lambda R^T: same range as R
    ej = q - gq
    nj = np.empty(B)
    for r in range(B):
        u, s, _ = np.linalg.svd(W[r])
        Rr = u[:, s > 1e-9 * s[0]]
        nj[r] = np.linalg.norm(Rr.T @ ej[r])
    assert np.abs(g.get(abi.F_JT_POSITION_ERROR_NORM, t1)[:, 0] - nj).max() <= 1e-10
    flags, cnt = g.goal_reached(t1, 1e-2, 0)
    assert np.array_equal(flags, (nj < 1e-2).astype(np.int32)) and cnt == int((nj < 1e-2).sum()) and cnt >= B // 2


def test_set_gains_unsafe(oracle_built):
    import osc_b200
    chain = load_chain("panda")
    B = 64
    rng = np.random.default_rng(3)
    q = random_q(chain, B, rng, 0.3)
    outs = []
    for unsafe in (False, True):
        p = Pair(chain, B, q0=q0_of(chain))
        p.set(abi.F_Q, q); p.both("update_model")
        t0 = p.add_jt()
        p.set(abi.F_JT_GOAL_POSITION, q + 0.1, t0)
        if unsafe:
            p.gpu.set_gains_unsafe(t0, [30.0] * 7, [5.0] * 7, [0.0] * 7)
            p.set_jt_params(t0, kp=30.0, kv=5.0)            # oracle side (and a no-op on the product side)
        else:
            p.set_jt_params(t0, kp=30.0, kv=5.0)
        tg, tr = p.both("tick", True)
        assert rel_err(tg, tr)[0] <= FP64_TOL
        outs.append(tg)
    assert np.array_equal(outs[0], outs[1])
    g = p.gpu
    g.set_gains_unsafe(t0, -1.0, 2.0, 0.0)                  # isotropic, negative: accepted (that is what "unsafe" means)
    assert g.joint_task_params(t0).kp[3] == -1.0
    with pytest.raises(osc_b200.OscInvalidArgument, match="size of gain vectors"):
        g.set_gains_unsafe(t0, [1.0, 2.0], [1.0, 2.0], [0.0, 0.0])
    jp = g.joint_task_params(t0)
    jp.kp[0] = -5.0
    with pytest.raises(osc_b200.OscInvalidArgument, match="gains must be positive"):
        g.set_joint_task_params(t0, jp)


def test_defer_model_update_makes_the_sensed_wrench_follow_q(oracle_built):
    """osc_defer_model_update (what the facade's deferred updateModel() tells the batch): after a write of q the kinematic rows are
    stale, and a sensed wrench handed over before the tick must be rotated with the compliant frame at the NEW q -- exactly what
    robot->updateModel(); task->updateSensedForceAndMoment() does in the reference (MotionForceTask.cpp:805-828).  Closed-loop force
    control along z, four cycles with a new posture each; with an explicit osc_update_model, with the deferral, and -- to show the
    call is needed -- with neither, where the first cycle still uses the posture of the last model update."""
    import orc
    import osc_b200
    import workloads
    from common import load_chain, q0_of
    chain = load_chain("panda")
    B = 128
    w = workloads.config2(chain, B, 46)

    def setup(b):
        b.set(abi.F_Q, w["q"]); b.set(abi.F_DQ, w["dq"]); b.update_model()
        t0 = b.add_motion_force_task(abi.mft_desc(chain, "end-effector", (0, 0, 0.07)))
        b.add_joint_task(abi.joint_task_desc(7))
        p = b.mft_params(t0)
        p.force_space_dimension = 1
        p.force_or_motion_axis[0], p.force_or_motion_axis[1], p.force_or_motion_axis[2] = 0.0, 0.0, 1.0
        p.closed_loop_force_control = 1
        b.set_mft_params(t0, p)
        b.set(abi.F_GOAL_FORCE, np.array([0.0, 0.0, 5.0]), t0)
        return t0

    errs = {}
    for variant in ("update_model", "defer", "neither"):
        g = osc_b200.Batch(chain, B, q0=q0_of(chain)); o = orc.OracleBatch(chain, B, q0=q0_of(chain))
        t0 = setup(g); setup(o)
        q = np.array(w["q"], float)
        errs[variant] = []
        for t in range(4):
            q = q + 1e-3 * (np.arange(7) - 3.0)
            f, m = np.array([0.3, -0.2, 4.0 + 0.1 * t]), np.array([0.01, 0.02, -0.03])
            o.set(abi.F_Q, q); o.update_model(); o.update_sensed_force_and_moment(t0, f, m)
            ref = o.tick(True)
            g.set(abi.F_Q, q)
            if variant == "update_model":
                g.update_model()
            elif variant == "defer":
                g._chk(g.L.osc_defer_model_update(g.h))
            g.update_sensed_force_and_moment(t0, f, m)
            errs[variant].append(rel_err(g.tick(True), ref)[0])
            assert g.tick_is_fused()
    assert max(errs["update_model"]) <= FP64_TOL, errs
    assert max(errs["defer"]) <= FP64_TOL, errs
    assert errs["neither"][0] > 1e-6, "without the call the first cycle must show the stale compliant frame: %s" % errs["neither"]
