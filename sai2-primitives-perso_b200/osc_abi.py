"""ctypes mirror of the plain-data part of include/osc_b200.h (structs, enums, field ids).

Shared by the product binding (osc_b200.py) and by the test-side oracle binding (oracle/orc.py): it contains no
algorithm, only the ABI layout.  tests/test_abi.py checks every constant here against the header text.
"""
import ctypes as C

MAX_DOF = 8
MAX_TASKS = 4
POPC_RING_CAPACITY = 1024
MAX_SINGULARITY_BUFFER = 256
NOALLOC_MAX_ROBOTS = 262144

OK, ERR_INVALID_ARGUMENT, ERR_CUDA, ERR_NO_DEVICE, ERR_UNSUPPORTED = 0, -1, -2, -3, -4
PRECISION_F64, PRECISION_F32 = 64, 32
LAYOUT_AOS, LAYOUT_SOA, LAYOUT_UNIFORM = 0, 1, 2
MEM_HOST, MEM_DEVICE = 0, 1
JOINT_REVOLUTE, JOINT_PRISMATIC = 0, 1
TASK_UNDEFINED, TASK_JOINT, TASK_MOTION_FORCE = 0, 1, 2
FULL_DYNAMIC_DECOUPLING, BOUNDED_INERTIA_ESTIMATES, IMPEDANCE = 0, 1, 2
NO_SINGULARITY, TYPE_1_SINGULARITY, TYPE_2_SINGULARITY = 0, 1, 2

TICK_OUTPUTS_ALL, TICK_OUTPUTS_LAZY, TICK_OUTPUTS_TAU = 0, 1, 2

STATUS_NAN_TORQUE = 1 << 0
STATUS_POPC_RING_OVERFLOW = 1 << 1
STATUS_POPC_RC_NAN = 1 << 2
STATUS_SVD_NOT_CONVERGED = 1 << 3
STATUS_NOT_POSITIVE_DEFINITE = 1 << 4
STATUS_SINGULAR_BRANCH = 1 << 5
STATUS_BAD_GOAL_ORIENTATION = 1 << 6
STATUS_OTG_ERROR = 1 << 7

# field ids (enum in osc_b200.h)
F_Q, F_DQ, F_TAU, F_M, F_MINV, F_N_TOTAL, F_STATUS, F_GRAVITY = 0, 1, 2, 3, 4, 5, 6, 7
F_NONLINEAR_EFFECTS = 8
F_SIGMA_FORCE, F_SIGMA_POSITION, F_SIGMA_MOMENT, F_SIGMA_ORIENTATION = 110, 111, 112, 113
F_POSITION_ERROR, F_SELECTED_ORIENTATION_ERROR, F_GOAL_FORCE_WORLD, F_GOAL_MOMENT_WORLD = 114, 115, 116, 117
F_POSITION_ERROR_NORM, F_ORIENTATION_ERROR_NORM, F_JT_POSITION_ERROR_NORM = 118, 119, 120
F_TASK_TORQUES, F_N_TASK, F_N_PREC, F_N_TASK_AND_PREC = 20, 21, 22, 23
F_GOAL_POSITION, F_GOAL_ORIENTATION, F_GOAL_LINEAR_VELOCITY, F_GOAL_ANGULAR_VELOCITY = 40, 41, 42, 43
F_GOAL_LINEAR_ACCELERATION, F_GOAL_ANGULAR_ACCELERATION, F_GOAL_FORCE, F_GOAL_MOMENT = 44, 45, 46, 47
F_SENSED_FORCE_WORLD, F_SENSED_MOMENT_WORLD = 48, 49
F_CURRENT_POSITION, F_CURRENT_ORIENTATION, F_CURRENT_LINEAR_VELOCITY, F_CURRENT_ANGULAR_VELOCITY = 50, 51, 52, 53
F_ORIENTATION_ERROR, F_UNIT_MASS_FORCE, F_JACOBIAN, F_PROJECTED_JACOBIAN = 54, 55, 56, 57
F_LAMBDA_WORLD, F_LAMBDA_MODIFIED_WORLD, F_ALPHA, F_SINGULAR_VALUES, F_NS_RANK = 58, 59, 60, 61, 62
F_SINGULARITY_TYPES, F_SINGULARITY_COUNTERS = 63, 64
F_INTEGRATED_POSITION_ERROR, F_INTEGRATED_ORIENTATION_ERROR = 65, 66
F_INTEGRATED_FORCE_ERROR, F_INTEGRATED_MOMENT_ERROR = 67, 68
F_POPC_RC, F_POPC_OBSERVER, F_Q_PRIOR = 69, 70, 71
F_DESIRED_POSITION, F_DESIRED_ORIENTATION, F_DESIRED_LINEAR_VELOCITY, F_DESIRED_ANGULAR_VELOCITY = 72, 73, 74, 75
F_DESIRED_LINEAR_ACCELERATION, F_DESIRED_ANGULAR_ACCELERATION = 76, 77
F_JT_GOAL_POSITION, F_JT_GOAL_VELOCITY, F_JT_GOAL_ACCELERATION = 80, 81, 82
F_JT_CURRENT_POSITION, F_JT_CURRENT_VELOCITY, F_JT_INTEGRATED_ERROR = 83, 84, 85
F_JT_LAMBDA_WORLD, F_JT_LAMBDA_MODIFIED_WORLD, F_JT_RANGE_RANK, F_JT_DESIRED_VELOCITY = 86, 87, 88, 89
F_JT_DESIRED_POSITION, F_JT_DESIRED_ACCELERATION = 90, 91
F_OTG_FLAGS, F_OTG_RESULT, F_OTG_STATE = 100, 101, 102

INT_FIELDS = {F_STATUS, F_NS_RANK, F_SINGULARITY_TYPES, F_SINGULARITY_COUNTERS, F_JT_RANGE_RANK, F_OTG_FLAGS, F_OTG_RESULT}

d = C.c_double
i32 = C.c_int32


class ChainDesc(C.Structure):
    _fields_ = [("n", i32), ("joint_type", i32 * MAX_DOF), ("joint_axis", (d * 3) * MAX_DOF),
                ("joint_xyz", (d * 3) * MAX_DOF), ("joint_rot", (d * 9) * MAX_DOF), ("mass", d * MAX_DOF),
                ("com", (d * 3) * MAX_DOF), ("inertia", (d * 9) * MAX_DOF), ("q_lower", d * MAX_DOF),
                ("q_upper", d * MAX_DOF), ("effort", d * MAX_DOF), ("world_rot", d * 9), ("world_xyz", d * 3),
                ("gravity", d * 3)]


class MftDesc(C.Structure):
    _fields_ = [("body", i32), ("frame_xyz", d * 3), ("frame_rot", d * 9), ("partial", i32),
                ("n_dirs_translation", i32), ("dirs_translation", (d * 3) * 3), ("n_dirs_rotation", i32),
                ("dirs_rotation", (d * 3) * 3), ("parametrization_in_compliant_frame", i32), ("loop_timestep", d)]


class MftParams(C.Structure):
    _fields_ = [("kp_pos", d * 3), ("kv_pos", d * 3), ("ki_pos", d * 3), ("kp_ori", d * 3), ("kv_ori", d * 3),
                ("ki_ori", d * 3), ("kp_force", d), ("kv_force", d), ("ki_force", d), ("kp_moment", d),
                ("kv_moment", d), ("ki_moment", d), ("kff_force", d), ("kff_moment", d),
                ("max_force_control_feedback_output", d), ("max_moment_control_feedback_output", d),
                ("force_space_dimension", i32), ("force_or_motion_axis", d * 3), ("moment_space_dimension", i32),
                ("moment_or_rotmotion_axis", d * 3), ("closed_loop_force_control", i32),
                ("closed_loop_moment_control", i32), ("passivity_enabled", i32), ("use_velocity_saturation", i32),
                ("linear_saturation_velocity", d), ("angular_saturation_velocity", d),
                ("dynamic_decoupling_type", i32), ("bie_threshold", d), ("singularity_handling_enabled", i32),
                ("enforce_type_1_strategy", i32), ("s_min", d), ("s_max", d), ("kp_type_1", d), ("kv_type_1", d),
                ("kv_type_2", d), ("s_abs_tol", d), ("type_1_tol", d), ("type_2_torque_ratio", d),
                ("type_2_angle_threshold", d), ("perturb_step_size", d), ("buffer_size", i32), ("sensor_rot", d * 9),
                ("sensor_xyz", d * 3), ("apply_disturbance_compensation", i32)]


class JointTaskDesc(C.Structure):
    _fields_ = [("k", i32), ("selection", d * (MAX_DOF * MAX_DOF)), ("full", i32), ("loop_timestep", d)]


class JointTaskParams(C.Structure):
    _fields_ = [("kp", d * MAX_DOF), ("kv", d * MAX_DOF), ("ki", d * MAX_DOF), ("use_velocity_saturation", i32),
                ("saturation_velocity", d * MAX_DOF), ("dynamic_decoupling_type", i32), ("bie_threshold", d)]


def chain_desc(chain: dict) -> ChainDesc:
    """fixtures/*.chain.json dict -> osc_chain_desc"""
    n = chain["n"]
    if n > MAX_DOF:
        raise ValueError("chain has more than OSC_MAX_DOF joints")
    c = ChainDesc()
    c.n = n
    for i, j in enumerate(chain["joints"]):
        c.joint_type[i] = JOINT_PRISMATIC if j["type"] == "prismatic" else JOINT_REVOLUTE
        for k in range(3):
            c.joint_axis[i][k] = j["axis"][k]
            c.joint_xyz[i][k] = j["xyz"][k]
        for k in range(9):
            c.joint_rot[i][k] = j["rot"][k // 3][k % 3]
        c.q_lower[i], c.q_upper[i], c.effort[i] = j["lower"], j["upper"], j["effort"]
        b = chain["bodies"][i]
        c.mass[i] = b["mass"]
        for k in range(3):
            c.com[i][k] = b["com"][k]
        for k in range(9):
            c.inertia[i][k] = b["inertia"][k // 3][k % 3]
    for k in range(9):
        c.world_rot[k] = chain["world_rot"][k // 3][k % 3]
    for k in range(3):
        c.world_xyz[k] = chain["world_xyz"][k]
        c.gravity[k] = chain["gravity"][k]
    return c


def mft_desc(chain: dict, link: str, pos_in_link=(0, 0, 0), rot_in_link=None, dirs_translation=None,
             dirs_rotation=None, in_compliant_frame=False, dt=0.001) -> MftDesc:
    """MotionForceTask(robot, link_name, compliant_frame, ...) arguments -> osc_mft_desc.
    The link's own offset inside its movable body is folded into the compliant frame."""
    import numpy as np
    L = chain["links"][link]
    if L["body"] < 0:
        raise ValueError("control link is welded to the world")
    lR, lp = np.array(L["rot"], float), np.array(L["xyz"], float)
    R = np.eye(3) if rot_in_link is None else np.array(rot_in_link, float)
    p = lp + lR @ np.array(pos_in_link, float)
    Rf = lR @ R
    m = MftDesc()
    m.body = L["body"]
    for k in range(3):
        m.frame_xyz[k] = p[k]
    for k in range(9):
        m.frame_rot[k] = Rf[k // 3][k % 3]
    m.partial = 0 if (dirs_translation is None and dirs_rotation is None) else 1
    dt_ = [] if dirs_translation is None else list(dirs_translation)
    dr_ = [] if dirs_rotation is None else list(dirs_rotation)
    if len(dt_) > 3 or len(dr_) > 3:
        raise ValueError("at most 3 controlled directions per block")
    m.n_dirs_translation, m.n_dirs_rotation = len(dt_), len(dr_)
    for c_, v in enumerate(dt_):
        for k in range(3):
            m.dirs_translation[c_][k] = float(v[k])
    for c_, v in enumerate(dr_):
        for k in range(3):
            m.dirs_rotation[c_][k] = float(v[k])
    m.parametrization_in_compliant_frame = int(in_compliant_frame)
    m.loop_timestep = dt
    return m


def joint_task_desc(n: int, selection=None, dt=0.001) -> JointTaskDesc:
    import numpy as np
    j = JointTaskDesc()
    j.loop_timestep = dt
    if selection is None:
        j.full, j.k = 1, n
        return j
    S = np.atleast_2d(np.array(selection, float))
    if S.shape[1] != n:
        raise ValueError("joint selection matrix size not consistent with robot dof in JointTask constructor\n")
    j.full, j.k = 0, S.shape[0]
    flat = S.reshape(-1)
    for i in range(flat.size):
        j.selection[i] = flat[i]
    return j
