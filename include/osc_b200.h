/* osc_b200.h -- C ABI of the batched operational-space controller (B200 / sm_100a).
 *
 * This is the drop-in boundary for ONE hot path of sai2-primitives: the per-tick controller math
 *   robot->updateModel(); controller->updateControllerTaskModels(); controller->computeControlTorques()
 * (reference call sites: examples/05-using_robot_controller/05-using_robot_controller.cpp:143-148,195)
 * evaluated for B independent robot instances per call.  The reference has no FFI of its own: its boundary is
 * the C++ class API of a static library.  The C++ facade in include/Sai2PrimitivesBatched.h keeps
 * those class and method names and calls the entry points below; INTEGRATION.md shows the binding a maintainer adds.
 *
 * Conventions
 *   - plain C, opaque handles, plain pointers and sizes; no exceptions cross the boundary.
 *   - every function returns an int status: OSC_OK (0) or a negative OSC_ERR_* code; osc_last_error() returns the
 *     message (thread local), worded like the std::invalid_argument the reference throws at the cited line.
 *   - "real" buffers are double for an OSC_PRECISION_F64 batch and float for an OSC_PRECISION_F32 batch.
 *   - matrices in user buffers are ROW-major per robot (J is 6 x n with linear rows first, like
 *     Sai2Model::JWorldFrame as used at src/tasks/MotionForceTask.cpp:293-298).
 *   - batch buffers come in three layouts: AOS [B][k] (robot-major, what a host loop over robots produces),
 *     SOA [k][B] (the device-resident layout, zero-copy friendly) and UNIFORM [k] (one value broadcast to all robots).
 *   - buffers may live in host or device memory (OSC_MEM_*); device buffers must be on the batch's device.
 *   - calls are asynchronous on the batch's stream; a call that writes a HOST buffer returns after the data landed.
 *     A call that READS a host buffer (osc_set_field, osc_mft_update_sensed_force_and_moment, osc_step_host inputs) has consumed
 *     a PAGEABLE buffer when it returns (cudaMemcpyAsync stages pageable memory), but reads a PINNED (page-locked) buffer
 *     asynchronously: leave a pinned input buffer unchanged until the next call that returns data to the host, or
 *     osc_synchronize(), has returned.
 *   - one host thread drives one batch; a batch lives on one GPU; shard robots over GPUs with one batch per GPU
 *     (independent robots, no inter-GPU traffic, no collective: SURVEY.md 8e).
 */
#ifndef OSC_B200_H_
#define OSC_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* libosc_b200.so is built with -fvisibility=hidden: only this header is exported */
#endif

#define OSC_MAX_DOF 8
#define OSC_MAX_TASKS 4
#define OSC_POPC_RING_CAPACITY 1024 /* bounded stand-in for the unbounded std::queue, POPCExplicitForceControl.h:44 (quirk Q9) */
#define OSC_MAX_SINGULARITY_BUFFER 256 /* >= BUFFER_SIZE = 200, SingularityHandler.cpp:16 */
#define OSC_NOALLOC_MAX_ROBOTS 262144 /* largest batch osc_tick_noalloc accepts (its static device scratch is 4 bytes per robot) */

/* ---- status codes ---- */
enum {
	OSC_OK = 0,
	OSC_ERR_INVALID_ARGUMENT = -1, /* the reference throws std::invalid_argument */
	OSC_ERR_CUDA = -2,
	OSC_ERR_NO_DEVICE = -3, /* no CUDA device / extension cannot run: the product path never falls back to the CPU */
	OSC_ERR_UNSUPPORTED = -4
};

enum { OSC_PRECISION_F64 = 64, OSC_PRECISION_F32 = 32 };
enum { OSC_LAYOUT_AOS = 0, OSC_LAYOUT_SOA = 1, OSC_LAYOUT_UNIFORM = 2 };
enum { OSC_MEM_HOST = 0, OSC_MEM_DEVICE = 1 };
enum { OSC_JOINT_REVOLUTE = 0, OSC_JOINT_PRISMATIC = 1 };

/* TaskType, src/tasks/TemplateTask.h:19-23 */
enum { OSC_TASK_UNDEFINED = 0, OSC_TASK_JOINT = 1, OSC_TASK_MOTION_FORCE = 2 };
/* DynamicDecouplingType, src/helper_modules/Sai2PrimitivesCommonDefinitions.h:9-16 */
enum { OSC_FULL_DYNAMIC_DECOUPLING = 0, OSC_BOUNDED_INERTIA_ESTIMATES = 1, OSC_IMPEDANCE = 2 };
/* SingularityType, src/tasks/SingularityHandler.h:25-29 */
enum { OSC_NO_SINGULARITY = 0, OSC_TYPE_1_SINGULARITY = 1, OSC_TYPE_2_SINGULARITY = 2 };

/* per-robot status bits (field OSC_F_STATUS) */
enum {
	OSC_STATUS_NAN_TORQUE = 1 << 0,          /* a singular-task torque was NaN and zeroed, SingularityHandler.cpp:358 */
	OSC_STATUS_POPC_RING_OVERFLOW = 1 << 1,  /* Q9: the passivity window outgrew OSC_POPC_RING_CAPACITY, oldest dropped */
	OSC_STATUS_POPC_RC_NAN = 1 << 2,         /* Q8 */
	OSC_STATUS_SVD_NOT_CONVERGED = 1 << 3,
	OSC_STATUS_NOT_POSITIVE_DEFINITE = 1 << 4, /* a Cholesky pivot was <= 0 (the reference's llt() would return garbage) */
	OSC_STATUS_SINGULAR_BRANCH = 1 << 5,     /* informational: robot is in the blend / fully singular branch */
	OSC_STATUS_BAD_GOAL_ORIENTATION = 1 << 6, /* orientationError would throw (Appendix B): goal not orthonormal */
	OSC_STATUS_OTG_ERROR = 1 << 7            /* the internal OTG returned an error this tick (the reference prints a warning,
	                                            keeps the previous output and zeroes the trajectory velocity:
	                                            OTG_joints.cpp:143-149, OTG_6dof_cartesian.cpp:229-235) */
};

/* ---- robot model: stands in for Sai2Model::Sai2Model(urdf) (un-vendored sai2-model) ----
 * A serial chain with fixed links already merged into their movable parent.  Joint i connects body i-1 (or the
 * welded root for i = 0) to body i; its origin (xyz, rot) is expressed in the parent body frame. */
typedef struct osc_chain_desc {
	int32_t n;
	int32_t joint_type[OSC_MAX_DOF];
	double joint_axis[OSC_MAX_DOF][3];
	double joint_xyz[OSC_MAX_DOF][3];
	double joint_rot[OSC_MAX_DOF][9];
	double mass[OSC_MAX_DOF];
	double com[OSC_MAX_DOF][3];
	double inertia[OSC_MAX_DOF][9]; /* about the com, body frame */
	double q_lower[OSC_MAX_DOF], q_upper[OSC_MAX_DOF], effort[OSC_MAX_DOF]; /* Sai2Model::jointLimits(), SingularityHandler.cpp:43-51 */
	double world_rot[9], world_xyz[3]; /* T_world_robot */
	double gravity[3];
} osc_chain_desc;

/* ---- URDF (+ world file) -> osc_chain_desc on the host: stands in for Sai2Model::Sai2Model(urdf_file) as the reference uses it
 * (examples/04-task_and_redundancy/04-task_and_redundancy.cpp:67) and for the world file (examples/04-.../world.urdf:3 gravity
 * attribute, :5-7 <robot> with its optional <origin> = T_world_robot).  Root link welded to the world, fixed joints merged into
 * their movable parent (every link stays addressable by name), initial posture from <calibration falling="deg"> (rising, in
 * metres, for a prismatic joint), limits / efforts from <limit>.  Serial chains of at most OSC_MAX_DOF joints.
 * world_path may be NULL (identity pose, gravity (0, 0, -9.81)); robot_name selects the <robot name=...> of the world file
 * (NULL = its first robot). ---- */
typedef struct osc_urdf osc_urdf_t;
int osc_urdf_load(const char* urdf_path, const char* world_path, const char* robot_name, osc_urdf_t** out);
void osc_urdf_free(osc_urdf_t* urdf);
/* the chain constants and (q0 may be NULL) the initial posture q0[n] */
int osc_urdf_chain(const osc_urdf_t* urdf, struct osc_chain_desc* chain, double* q0);
/* a named link (fixed links included): the movable body it is fixed in (-1 = welded to the world) and its pose in that body
 * -- what osc_mft_desc.body / frame_* are built from (compose with the compliant frame given in the link) */
int osc_urdf_link(const osc_urdf_t* urdf, const char* link_name, int* body, double xyz[3], double rot[9]);
int osc_urdf_link_count(const osc_urdf_t* urdf);
const char* osc_urdf_link_name(const osc_urdf_t* urdf, int i);
const char* osc_urdf_joint_name(const osc_urdf_t* urdf, int i);

/* MotionForceTask constructors, src/tasks/MotionForceTask.h:96-110, MotionForceTask.cpp:16-90 */
typedef struct osc_mft_desc {
	int32_t body;         /* movable body the control link is fixed in (the link's own offset is folded into frame_*) */
	double frame_xyz[3];  /* compliant frame in that body's frame */
	double frame_rot[9];
	int32_t partial;      /* 0: full 6-DoF constructor (.cpp:16-31); 1: partial constructor (.cpp:33-90) */
	int32_t n_dirs_translation;
	double dirs_translation[3][3];
	int32_t n_dirs_rotation;
	double dirs_rotation[3][3];
	int32_t parametrization_in_compliant_frame;
	double loop_timestep;
} osc_mft_desc;

/* Runtime parameters of a MotionForceTask (uniform over the batch).  Defaults: MotionForceTask.h:40-75,
 * SingularityHandler.cpp:10-20, MotionForceTask.cpp:197.  osc_mft_set_params applies the reference's side effects
 * when force_space_dim/axis, moment_space_dim/axis (.cpp:830-890), closed_loop_* (.cpp:973-986) or
 * passivity_enabled (POPCExplicitForceControl.cpp:25-28) change. */
typedef struct osc_mft_params {
	double kp_pos[3], kv_pos[3], ki_pos[3];
	double kp_ori[3], kv_ori[3], ki_ori[3];
	double kp_force, kv_force, ki_force;
	double kp_moment, kv_moment, ki_moment;
	double kff_force, kff_moment;
	double max_force_control_feedback_output, max_moment_control_feedback_output;
	int32_t force_space_dimension;
	double force_or_motion_axis[3];
	int32_t moment_space_dimension;
	double moment_or_rotmotion_axis[3];
	int32_t closed_loop_force_control, closed_loop_moment_control;
	int32_t passivity_enabled;
	int32_t use_velocity_saturation;
	double linear_saturation_velocity, angular_saturation_velocity;
	int32_t dynamic_decoupling_type;
	double bie_threshold;
	int32_t singularity_handling_enabled; /* enable/disableSingularityHandling, SingularityHandler.h:113-121 */
	int32_t enforce_type_1_strategy;      /* handleAllSingularitiesAsType1 */
	double s_min, s_max;
	double kp_type_1, kv_type_1, kv_type_2;
	double s_abs_tol, type_1_tol, type_2_torque_ratio, type_2_angle_threshold, perturb_step_size;
	int32_t buffer_size;
	double sensor_rot[9], sensor_xyz[3]; /* _T_control_to_sensor, MotionForceTask.cpp:794-803 */
	int32_t apply_disturbance_compensation; /* 0 (default) reproduces quirk Q1: the tau_prec term is identically zero */
} osc_mft_params;

/* JointTask constructors, src/tasks/JointTask.h:56-75, JointTask.cpp:14-43 */
typedef struct osc_joint_task_desc {
	int32_t k;                                 /* task dof = rows of the joint selection matrix */
	double selection[OSC_MAX_DOF * OSC_MAX_DOF]; /* k x n row-major; ignored when full != 0 */
	int32_t full;
	double loop_timestep;
} osc_joint_task_desc;

/* defaults JointTask.h:31-45 */
typedef struct osc_joint_task_params {
	double kp[OSC_MAX_DOF], kv[OSC_MAX_DOF], ki[OSC_MAX_DOF];
	int32_t use_velocity_saturation;
	double saturation_velocity[OSC_MAX_DOF];
	int32_t dynamic_decoupling_type;
	double bie_threshold;
} osc_joint_task_params;

/* ---- per-robot fields (set / get).  "task" is the index in the hierarchy, or -1 for robot-level fields. ---- */
enum {
	/* robot level (task = -1) */
	OSC_F_Q = 0,            /* n   in/out  robot->setQ / q() */
	OSC_F_DQ = 1,           /* n   in/out */
	OSC_F_TAU = 2,          /* n   out     computeControlTorques() result */
	OSC_F_M = 3,            /* n*n out     Sai2Model::M() */
	OSC_F_MINV = 4,         /* n*n out     Sai2Model::MInv() */
	OSC_F_N_TOTAL = 5,      /* n*n out     N_prec after the last task, RobotController.cpp:58 */
	OSC_F_STATUS = 6,       /* 1 (int32) out */
	OSC_F_GRAVITY = 7,      /* n   out     jointGravityVector() (valid when gravity compensation is on) */
	OSC_F_NONLINEAR_EFFECTS = 8, /* n out  b(q,dq) + g(q) = Sai2Model::coriolisPlusGravity at the state osc_rollout_step started from */
	/* any task */
	OSC_F_TASK_TORQUES = 20,   /* n   out  this task's computeTorques result */
	OSC_F_N_TASK = 21,         /* n*n out  getTaskNullspace() */
	OSC_F_N_PREC = 22,         /* n*n out  getPreviousTasksNullspace() */
	OSC_F_N_TASK_AND_PREC = 23,/* n*n out  getTaskAndPreviousNullspace() */
	/* MotionForceTask */
	OSC_F_GOAL_POSITION = 40,             /* 3 in/out */
	OSC_F_GOAL_ORIENTATION = 41,          /* 9 in/out row-major */
	OSC_F_GOAL_LINEAR_VELOCITY = 42,      /* 3 */
	OSC_F_GOAL_ANGULAR_VELOCITY = 43,     /* 3 */
	OSC_F_GOAL_LINEAR_ACCELERATION = 44,  /* 3 */
	OSC_F_GOAL_ANGULAR_ACCELERATION = 45, /* 3 */
	OSC_F_GOAL_FORCE = 46,                /* 3 */
	OSC_F_GOAL_MOMENT = 47,               /* 3 */
	OSC_F_SENSED_FORCE_WORLD = 48,        /* 3 out (in: use osc_mft_update_sensed_force_and_moment) */
	OSC_F_SENSED_MOMENT_WORLD = 49,       /* 3 */
	OSC_F_CURRENT_POSITION = 50,          /* 3 out */
	OSC_F_CURRENT_ORIENTATION = 51,       /* 9 out */
	OSC_F_CURRENT_LINEAR_VELOCITY = 52,   /* 3 out */
	OSC_F_CURRENT_ANGULAR_VELOCITY = 53,  /* 3 out */
	OSC_F_ORIENTATION_ERROR = 54,         /* 3 out */
	OSC_F_UNIT_MASS_FORCE = 55,           /* 6 out getUnitMassForce() */
	OSC_F_JACOBIAN = 56,                  /* 6*n out _jacobian */
	OSC_F_PROJECTED_JACOBIAN = 57,        /* 6*n out _projected_jacobian */
	OSC_F_LAMBDA_WORLD = 58,              /* 36 out U_ns Lambda_ns U_ns^T: the basis-independent form of _Lambda_ns */
	OSC_F_LAMBDA_MODIFIED_WORLD = 59,     /* 36 out same for _Lambda_ns_modified */
	OSC_F_ALPHA = 60,                     /* 1 out _alpha */
	OSC_F_SINGULAR_VALUES = 61,           /* 6 out _svd_s (zero padded) */
	OSC_F_NS_RANK = 62,                   /* 1 (int32) out cols(_task_range_ns), 0 when fully singular */
	OSC_F_SINGULARITY_TYPES = 63,         /* 6 (int32) out _singularity_types zero padded */
	OSC_F_SINGULARITY_COUNTERS = 64,      /* 2 (int32) out _type_1_counter, _type_2_counter */
	OSC_F_INTEGRATED_POSITION_ERROR = 65, /* 3 in/out */
	OSC_F_INTEGRATED_ORIENTATION_ERROR = 66,
	OSC_F_INTEGRATED_FORCE_ERROR = 67,
	OSC_F_INTEGRATED_MOMENT_ERROR = 68,
	OSC_F_POPC_RC = 69,                   /* 1 out _Rc */
	OSC_F_POPC_OBSERVER = 70,             /* 1 out _passivity_observer_value */
	OSC_F_Q_PRIOR = 71,                   /* n in/out setType1Posture */
	OSC_F_DESIRED_POSITION = 72,          /* 3 out */
	OSC_F_DESIRED_ORIENTATION = 73,       /* 9 out */
	OSC_F_DESIRED_LINEAR_VELOCITY = 74,   /* 3 out */
	OSC_F_DESIRED_ANGULAR_VELOCITY = 75,  /* 3 out */
	OSC_F_DESIRED_LINEAR_ACCELERATION = 76, /* 3 out */
	OSC_F_DESIRED_ANGULAR_ACCELERATION = 77, /* 3 out */
	/* JointTask (k = task dof) */
	OSC_F_JT_GOAL_POSITION = 80,     /* k in/out */
	OSC_F_JT_GOAL_VELOCITY = 81,     /* k */
	OSC_F_JT_GOAL_ACCELERATION = 82, /* k */
	OSC_F_JT_CURRENT_POSITION = 83,  /* k out */
	OSC_F_JT_CURRENT_VELOCITY = 84,  /* k out */
	OSC_F_JT_INTEGRATED_ERROR = 85,  /* k in/out */
	OSC_F_JT_LAMBDA_WORLD = 86,      /* k*k out R Lambda R^T : basis-independent form of _M_partial */
	OSC_F_JT_LAMBDA_MODIFIED_WORLD = 87, /* k*k out */
	OSC_F_JT_RANGE_RANK = 88,        /* 1 (int32) out cols(_current_task_range), 0 when the range is empty */
	OSC_F_JT_DESIRED_VELOCITY = 89,  /* k out */
	OSC_F_JT_DESIRED_POSITION = 90,  /* k out */
	OSC_F_JT_DESIRED_ACCELERATION = 91, /* k out */
	/* MotionForceTask getters that are functions of the task state (evaluated on demand at the state of the last
	 * computeTorques): MotionForceTask.h:611-614 / .cpp:892-971, .h:268-269 / .cpp:540-546, .cpp:755-769, .cpp:548-578 */
	OSC_F_SIGMA_FORCE = 110,              /* 9 out sigmaForce() */
	OSC_F_SIGMA_POSITION = 111,           /* 9 out sigmaPosition() */
	OSC_F_SIGMA_MOMENT = 112,             /* 9 out sigmaMoment() */
	OSC_F_SIGMA_ORIENTATION = 113,        /* 9 out sigmaOrientation() */
	OSC_F_POSITION_ERROR = 114,           /* 3 out getPositionError() = sigmaPosition (goal - current) */
	OSC_F_SELECTED_ORIENTATION_ERROR = 115, /* 3 out getOrientationError() = sigmaOrientation * _orientation_error */
	OSC_F_GOAL_FORCE_WORLD = 116,         /* 3 out getGoalForce() (rotated when parametrised in the compliant frame) */
	OSC_F_GOAL_MOMENT_WORLD = 117,        /* 3 out getGoalMoment() */
	OSC_F_POSITION_ERROR_NORM = 118,      /* 1 out the norm goalPositionReached(tol) compares with tol */
	OSC_F_ORIENTATION_ERROR_NORM = 119,   /* 1 out the norm goalOrientationReached(tol) compares with tol */
	/* JointTask */
	OSC_F_JT_POSITION_ERROR_NORM = 120,   /* 1 out the norm JointTask::goalPositionReached(tol) compares with tol (JointTask.cpp:437-446) */
	/* internal OTG of a task (either type), valid once it has been enabled */
	OSC_F_OTG_FLAGS = 100,  /* 1 (int32) out: bit 0 = isGoalReached() (OTG_joints.h, OTG_6dof_cartesian.h) */
	OSC_F_OTG_RESULT = 101, /* 1 (int32) out: ruckig::Result of the last update (0 working, 1 finished, < 0 error) */
	OSC_F_OTG_STATE = 102   /* diagnostic, FP64 batches only: the raw state rows of the OTG (layout OtgL<D> in csrc/osc_otg.cuh) */
};

typedef struct osc_batch osc_batch_t;

/* ---- life cycle ---- */
const char* osc_last_error(void);
int osc_version(void);
int osc_device_count(void);
/* Sai2Model(urdf) + B copies of the robot on `device`; q is initialised to q0 (may be NULL = zeros), dq to zero. */
int osc_batch_create(const osc_chain_desc* chain, int64_t n_robots, int precision, int device, const double* q0, osc_batch_t** out);
/* The same batch sharded over several GPUs of one box (independent robots, contiguous shards
 * [k B / G, (k + 1) B / G), one host thread + one stream per device, no inter-GPU traffic, no collective).  The handle is an
 * ordinary osc_batch_t: EVERY entry point of this header accepts it and fans out over the shards from one thread per device;
 * per-robot HOST buffers span all n_robots (AOS [B][k], SOA [k][B] or UNIFORM [k]) and each shard reads / writes its own
 * slice; a call that writes a host buffer returns when every shard's slice has landed.  Device buffers belong to one device:
 * a multi-device batch rejects OSC_MEM_DEVICE (OSC_ERR_UNSUPPORTED) -- use osc_batch_shard() for per-device calls and streams.
 * The reference drives one robot from one thread (src/RobotController.h:24-42); this is its batch-of-robots-on-8-GPUs form.
 * device_ids may repeat a device (two shards on one GPU: used by the single-GPU tests). */
int osc_batch_create_multi(const osc_chain_desc* chain, int64_t n_robots, int precision, const int* device_ids, int n_devices, const double* q0,
                           osc_batch_t** out);
/* number of shards (1 for a single-device batch) and shard k itself: an ordinary single-device batch handle owned by the parent
 * (never destroy it), with its device, first robot and robot count; NULL if k is out of range */
int osc_batch_shard_count(const osc_batch_t* batch);
osc_batch_t* osc_batch_shard(osc_batch_t* batch, int k, int* device, int64_t* first_robot, int64_t* n_robots);
void osc_batch_destroy(osc_batch_t* batch);
int osc_batch_set_stream(osc_batch_t* batch, void* cuda_stream);
int osc_synchronize(osc_batch_t* batch);
int64_t osc_batch_size(const osc_batch_t* batch);
int osc_batch_dof(const osc_batch_t* batch);
int osc_batch_precision(const osc_batch_t* batch);

/* ---- hierarchy construction: RobotController::RobotController, src/RobotController.cpp:8-51 ----
 * Tasks are appended in priority order; the constructor rules of the reference are enforced (same dt, nothing after a
 * full joint task).  Both return the task index (>= 0) or an error code.  Adding a task reads the current model
 * (call osc_update_model first when q was changed), like initialSetup()/reInitializeTask() do. */
int osc_batch_add_motion_force_task(osc_batch_t* batch, const osc_mft_desc* desc);
int osc_batch_add_joint_task(osc_batch_t* batch, const osc_joint_task_desc* desc);
int osc_batch_task_count(const osc_batch_t* batch);
int osc_batch_task_type(const osc_batch_t* batch, int task);
int osc_batch_task_dof(const osc_batch_t* batch, int task); /* MotionForceTask: _pos_range + _ori_range; JointTask: _task_dof */

void osc_mft_default_params(osc_mft_params* p);
void osc_joint_task_default_params(osc_joint_task_params* p, int k);
int osc_mft_get_params(const osc_batch_t* batch, int task, osc_mft_params* out);
/* returns a bit mask >= 0 of the resets it performed: 1 linear (goal position <- current, linear integrators),
 * 2 angular, 4 passivity observer re-initialised */
int osc_mft_set_params(osc_batch_t* batch, int task, const osc_mft_params* p);
int osc_joint_task_get_params(const osc_batch_t* batch, int task, osc_joint_task_params* out);
int osc_joint_task_set_params(osc_batch_t* batch, int task, const osc_joint_task_params* p);
/* ---- internal online trajectory generation (Ruckig), acceleration-limited ----
 * JointTask::enableInternalOtgAccelerationLimited (src/tasks/JointTask.cpp:358-381): count = 1 (same limit for every task
 * dof, JointTask.h:262-271) or count = task dof; limits[0..count) = max velocity, limits[count..2*count) = max acceleration.
 * MotionForceTask::enableInternalOtgAccelerationLimited (src/tasks/MotionForceTask.cpp:511-523): count = 2 and
 * limits = {max_linear_velocity, max_angular_velocity, max_linear_acceleration, max_angular_acceleration}.
 * The reference constructs its tasks with the OTG ON (JointTask.h:38, MotionForceTask.h:67); a batch task starts with it
 * OFF (= after disableInternalOtg(), what every BASELINE config runs) and the C++ facade restores the reference default.
 * While it is on, the goal fields (OSC_F_GOAL_*, OSC_F_JT_GOAL_*) are the OTG's targets and the control law follows
 * OSC_F_DESIRED_* / OSC_F_JT_DESIRED_*.  The jerk-limited variant (enableInternalOtgJerkLimited) returns
 * OSC_ERR_UNSUPPORTED. */
int osc_task_enable_internal_otg_acceleration_limited(osc_batch_t* batch, int task, const double* limits, int count);
int osc_task_enable_internal_otg_jerk_limited(osc_batch_t* batch, int task, const double* limits, int count);
/* disableInternalOtg(), JointTask.h:320 / MotionForceTask.h:423 */
int osc_task_disable_internal_otg(osc_batch_t* batch, int task);
/* getInternalOtgEnabled() */
int osc_task_internal_otg_enabled(const osc_batch_t* batch, int task);

/* JointTask::setGainsUnsafe, src/tasks/JointTask.h:256 / JointTask.cpp:136-156: setGains without the sign check.  count = 1
 * (isotropic) or the task dof, else OSC_ERR_INVALID_ARGUMENT with the reference's message. */
int osc_joint_task_set_gains_unsafe(osc_batch_t* batch, int task, const double* kp, const double* kv, const double* ki, int count);
/* MotionForceTask::posSelectionProjector / oriSelectionProjector, src/tasks/MotionForceTask.h:653-659: the 3 x 3 diagonal blocks
 * of _partial_task_projection (row-major; uniform over the batch) */
int osc_mft_selection_projectors(const osc_batch_t* batch, int task, double pos_projector[9], double ori_projector[9]);
/* goalPositionReached(tolerance) / goalOrientationReached(tolerance) (src/tasks/MotionForceTask.h:508,521, .cpp:548-578;
 * src/tasks/JointTask.h:389, JointTask.cpp:437-446) for every robot: a device-side threshold of the error norm into one int32
 * flag per robot (flags_out: [B], host or device, may be NULL) and the number of robots that reached (n_reached, may be NULL).
 * what: 0 = position (MotionForceTask or JointTask), 1 = orientation (MotionForceTask). */
int osc_task_goal_reached(osc_batch_t* batch, int task, int what, double tolerance, int32_t* flags_out, int mem, int64_t* n_reached);

/* RobotController::enableGravityCompensation, src/RobotController.h:31-33 */
int osc_enable_gravity_compensation(osc_batch_t* batch, int enable);

/* ---- per-robot data ---- */
int osc_field_size(const osc_batch_t* batch, int task, int field); /* scalars per robot, < 0 on error */
int osc_set_field(osc_batch_t* batch, int task, int field, const void* data, int layout, int mem);
int osc_get_field(osc_batch_t* batch, int task, int field, void* out, int layout, int mem);
/* MotionForceTask::updateSensedForceAndMoment, src/tasks/MotionForceTask.cpp:805-828 (sensor-frame wrench in).  The compliant
 * frame is the one of the last osc_update_model / osc_update_kinematics, like the reference; in a loop built around osc_tick
 * (which runs updateModel itself) it is evaluated at the batch's current q, i.e. the reference's order of calls
 * robot->setQ(); robot->updateModel(); task->updateSensedForceAndMoment(); ... computeTorques(). */
int osc_mft_update_sensed_force_and_moment(osc_batch_t* batch, int task, const void* force, const void* moment, int layout, int mem);

/* ---- the tick ---- */
/* robot->updateModel(): kinematics, M (composite rigid body), M^-1.  examples/05...cpp:145 */
int osc_update_model(osc_batch_t* batch);
/* robot->updateKinematics() only */
int osc_update_kinematics(osc_batch_t* batch);
/* task->updateTaskModel(N_prec), src/tasks/TemplateTask.h:42.  n_prec_task = -1 uses identity, otherwise
 * getTaskAndPreviousNullspace() of that (already updated) task: the manual chaining of examples/04...cpp:141-150 */
int osc_task_update_model(osc_batch_t* batch, int task, int n_prec_task);
/* RobotController::updateControllerTaskModels, src/RobotController.cpp:53-60 */
int osc_update_task_models(osc_batch_t* batch);
/* task->computeTorques() (with_tau_prec = 0, TemplateTask.h:49) or task->computeTorques(tau_prec) (TemplateTask.h:58)
 * where tau_prec is the batch's running OSC_F_TAU; the result lands in OSC_F_TASK_TORQUES of the task. */
int osc_task_compute_torques(osc_batch_t* batch, int task, int with_tau_prec);
/* RobotController::computeControlTorques, src/RobotController.cpp:62-74 (stacked = 1), or the plain sum of the
 * no-argument computeTorques() of every task as in examples/04...cpp:188-206 (stacked = 0).
 * tau_out may be NULL (read OSC_F_TAU later). */
int osc_compute_torques(osc_batch_t* batch, int stacked, void* tau_out, int layout, int mem);
/* updateModel + updateControllerTaskModels + computeControlTorques in one call; uses the fused single-launch kernel
 * when the hierarchy is [MotionForceTask, full JointTask] (n = 6 or 7; the motion task full, or partial of rank 2-5 on n = 7) or
 * one full MotionForceTask alone (n = 6 or 7) on an all-revolute chain, or one full JointTask alone (any chain); otherwise the
 * sequence above (osc_tick_is_fused tells). */
int osc_tick(osc_batch_t* batch, int stacked, void* tau_out, int layout, int mem);
/* osc_tick for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same fused tick, same results as osc_tick, but the one scratch buffer the fused kernels use (the compacted list of the
 * robots that leave the fast path for the API-faithful route, 2 counters + one int per robot) lives in a file-scope
 * static __device__ array of the library instead of a cudaMalloc'ed one; the call performs no cudaMalloc / cudaFree and
 * calls no allocating library routine.  Limits and rules:
 *   - at most OSC_NOALLOC_MAX_ROBOTS (262,144) robots per batch, else OSC_ERR_INVALID_ARGUMENT; every GPU that loads the
 *     library holds the array (1 MiB + 256 B) until the process exits;
 *   - only hierarchies the fused tick serves (osc_tick_is_fused() == 1), else OSC_ERR_UNSUPPORTED; a task whose internal
 *     OTG is on must have gone through one osc_tick before (that is where the OTG's own list is allocated);
 *   - tau_out_device is DEVICE memory on the batch's device (AOS [B][n] or SOA [n][B]) or NULL (read OSC_F_TAU later);
 *   - the array is shared by every batch and every call on that GPU and keeps the last call's contents (the call clears
 *     the two counters it needs at zero): calls that overlap on the GPU -- two batches on different streams, or two host
 *     threads -- RACE on it.  Serialise them (same stream, or events).  osc_tick is unaffected and may overlap freely;
 *   - osc_tick_fallback_count() keeps reporting the last osc_tick, not this call. */
int osc_tick_noalloc(osc_batch_t* batch, int stacked, void* tau_out_device, int layout);
/* One control cycle with HOST buffers in one call (no counterpart in the reference, whose state lives in host memory): writes the
 * n_inputs fields (each as osc_set_field would: q, dq, goals, ...), runs osc_tick and returns when the torques are in tau_out
 * (host memory, AOS [B][n] or SOA [n][B]; NULL: read OSC_F_TAU later).  On a batch created over several shards
 * (osc_batch_create_multi -- also several shards on ONE device, e.g. device_ids = {0, 0, 0, 0}) the shards are issued one after
 * the other WITHOUT waiting in between: the host-to-device copies of shard k + 1 overlap the tick and the device-to-host copy of
 * shard k (PCIe is full duplex), and the call waits for all of them at the end.  Pinned host buffers are required for the overlap
 * (pageable memory makes every copy synchronous); AOS buffers give each shard one contiguous block (linear copies).
 * Measured (profiles/README.md r2m, 65,536 robots, 27 MB per step): on ONE device the overlap does not pay -- the step is bound by
 * the host-to-device copies, and cutting them into 2-8 x more, smaller (or pitched) copies costs more than hiding the 3.7 MB
 * read-back saves; use several shards for several devices, where every device has its own link. */
typedef struct osc_field_in {
	int32_t task;      /* -1: robot field (OSC_F_Q, OSC_F_DQ) */
	int32_t field;     /* OSC_F_* */
	const void* data;  /* host memory, whole batch */
	int32_t layout;    /* OSC_LAYOUT_AOS / _SOA / _UNIFORM */
} osc_field_in;
int osc_step_host(osc_batch_t* batch, const osc_field_in* inputs, int32_t n_inputs, int32_t stacked, void* tau_out, int32_t tau_layout);
/* Tells the batch that q / dq changed and that the caller DEFERS robot->updateModel() to the next osc_tick (control-cycle fusion of
 * the C++ facade): the kinematic and dynamic rows are treated as stale exactly as after a fused osc_tick -- a getter
 * evaluates them on demand at the current q, osc_mft_update_sensed_force_and_moment rotates through the compliant frame at the
 * current q, and an explicit osc_update_model still does the full update.  No kernel is launched.  No counterpart in the reference. */
int osc_defer_model_update(osc_batch_t* batch);
/* 1 if osc_tick runs the fused kernel for the current hierarchy / parameters */
int osc_tick_is_fused(const osc_batch_t* batch);
/* What osc_tick leaves behind besides tau and the persistent controller state (integrators, observers, history):
 *   OSC_TICK_OUTPUTS_LAZY (default)  vector-valued getters (task torques, poses, velocities, errors, unit-mass force, alpha)
 *        are written by the tick; matrix-valued getters (M, M^-1, J, Lambda, N, singular values ...) are evaluated on
 *        first use, at the batch's current q, by the per-call kernels -- every getter of the reference keeps working;
 *   OSC_TICK_OUTPUTS_ALL             the tick is the per-call kernel sequence (updateModel, updateTaskModel x tasks,
 *        computeTorques x tasks): every intermediate of the reference is materialised by the tick itself;
 *   OSC_TICK_OUTPUTS_TAU             tau and persistent state only (MPC rollouts / parallel environments). */
enum { OSC_TICK_OUTPUTS_ALL = 0, OSC_TICK_OUTPUTS_LAZY = 1, OSC_TICK_OUTPUTS_TAU = 2 };
int osc_batch_set_tick_outputs(osc_batch_t* batch, int outputs);
/* robots of the last fused tick that took the API-faithful route (singular / blend branch or not certifiable); synchronises */
int64_t osc_tick_fallback_count(osc_batch_t* batch);
/* number of kernels launched by this batch since creation (bench.py's gpu_launches) */
int64_t osc_launch_count(const osc_batch_t* batch);

/* RobotController::reinitializeTasks / TemplateTask::reInitializeTask; task = -1 for all */
int osc_reinitialize_tasks(osc_batch_t* batch, int task);

/* ---- device-side rollout (the caller of the tick in MPC rollouts / parallel environments) ----
 * Stands in for the simulation side of the reference's examples: sim->setJointTorques(tau); sim->integrate();
 * robot->setQ(sim->getJointPositions()); robot->setDq(sim->getJointVelocities())
 * (examples/05-using_robot_controller/05-using_robot_controller.cpp:219-233; Sai2Simulation is un-vendored).  Contact-free
 * forward dynamics qdd = M^-1 (tau - b(q,dq) - g(q) - joint_damping * dq) with tau = the batch's OSC_F_TAU held for
 * `substeps` symplectic-Euler steps of `dt` (dq += qdd dt; q += dq dt); q and dq are updated in place on the device.
 * Gravity is the chain's gravity vector (zero it to simulate a gravity-compensated robot). */
int osc_rollout_step(osc_batch_t* batch, double dt, int substeps, double joint_damping);
/* `ticks` x (osc_tick(stacked) ; osc_rollout_step(sim_dt, substeps, joint_damping)) queued without any host
 * synchronisation; tau_out (may be NULL) receives the torques of the last tick. */
int osc_rollout(osc_batch_t* batch, int stacked, int ticks, double sim_dt, int substeps, double joint_damping, void* tau_out, int layout, int mem);

/* ---- measurement helper: sustained FP64 FMA throughput of `device` in TFLOP/s (DFMA microbenchmark, 2 flop per FMA),
 * the denominator of the FP64 roofline bench.py reports (MEASURED_PEAKS.json has no FP64 entry). ---- */
int osc_measure_fp64_peak(int device, double* tflops, double* elapsed_ms);

/* ---- controller-state snapshot (MPC rollouts: fork, roll out, restore) ---- */
int64_t osc_snapshot_bytes(const osc_batch_t* batch);
int osc_snapshot(osc_batch_t* batch, void* dst, int mem);
int osc_restore(osc_batch_t* batch, const void* src, int mem);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* OSC_B200_H_ */
