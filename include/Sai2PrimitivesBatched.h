// Sai2PrimitivesBatched.h -- C++ host facade over the C ABI of osc_b200.h (header only, no Eigen, no CUDA headers).
//
// It keeps the class and method names of sai2-primitives so that a control loop written against
//   Sai2Primitives::RobotController / MotionForceTask / JointTask   (src/RobotController.h:24-42, src/tasks/MotionForceTask.h,
//   src/tasks/JointTask.h, src/tasks/TemplateTask.h:42-88)   and   Sai2Model::Sai2Model (setQ / setDq / updateModel)
// reads the same when it drives B robots at once.  Differences forced by the batch dimension:
//   * vectors / matrices are Sai2PrimitivesBatched::In / Out views (pointer + layout + memory space) instead of Eigen objects;
//     a std::array / std::vector / initializer list is a UNIFORM value broadcast to every robot;
//   * tasks attach to the robot batch when they are constructed, so construction order = priority order; RobotController checks it;
//   * updateTaskModel takes the PREVIOUS TASK (or nothing = identity) instead of an N_prec matrix: N_prec is that task's
//     getTaskAndPreviousNullspace(), exactly what examples/04-task_and_redundancy/04-task_and_redundancy.cpp:144-150 passes;
//   * errors: the std::invalid_argument of the reference, with the reference's message texts (osc_last_error()).
#ifndef SAI2_PRIMITIVES_BATCHED_H_
#define SAI2_PRIMITIVES_BATCHED_H_

#include <array>
#include <cstdint>
#include <initializer_list>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "osc_b200.h"

namespace Sai2PrimitivesBatched {

// src/helper_modules/Sai2PrimitivesCommonDefinitions.h:9-23
enum DynamicDecouplingType { FULL_DYNAMIC_DECOUPLING = OSC_FULL_DYNAMIC_DECOUPLING, BOUNDED_INERTIA_ESTIMATES = OSC_BOUNDED_INERTIA_ESTIMATES, IMPEDANCE = OSC_IMPEDANCE };
struct PIDGains {
	double kp, kv, ki;
	PIDGains(double kp_, double kv_, double ki_) : kp(kp_), kv(kv_), ki(ki_) {}
};
// src/tasks/TemplateTask.h:19-23
enum TaskType { UNDEFINED = OSC_TASK_UNDEFINED, JOINT_TASK = OSC_TASK_JOINT, MOTION_FORCE_TASK = OSC_TASK_MOTION_FORCE };
// src/tasks/SingularityHandler.h:25-29
enum SingularityType { NO_SINGULARITY = OSC_NO_SINGULARITY, TYPE_1_SINGULARITY = OSC_TYPE_1_SINGULARITY, TYPE_2_SINGULARITY = OSC_TYPE_2_SINGULARITY };

inline void check(int rc) {
	if (rc >= 0) return;
	const std::string msg = osc_last_error();
	if (rc == OSC_ERR_INVALID_ARGUMENT) throw std::invalid_argument(msg);
	throw std::runtime_error(msg);
}

// ---- batch views ----
struct In {  // read-only input: [B][k] AOS, [k][B] SOA or one [k] value for every robot
	const void* ptr;
	int layout, mem;
	std::vector<double> own;  // backing store of a uniform value
	In(const void* p, int layout_, int mem_ = OSC_MEM_HOST) : ptr(p), layout(layout_), mem(mem_) {}
	In(std::initializer_list<double> v) : layout(OSC_LAYOUT_UNIFORM), mem(OSC_MEM_HOST), own(v) { ptr = own.data(); }
	In(const std::vector<double>& v) : layout(OSC_LAYOUT_UNIFORM), mem(OSC_MEM_HOST), own(v) { ptr = own.data(); }
	template <size_t K> In(const std::array<double, K>& v) : layout(OSC_LAYOUT_UNIFORM), mem(OSC_MEM_HOST), own(v.begin(), v.end()) { ptr = own.data(); }
	In(const In& o) : ptr(o.ptr), layout(o.layout), mem(o.mem), own(o.own) { if (!own.empty()) ptr = own.data(); }
};
inline In hostAOS(const void* p) { return In(p, OSC_LAYOUT_AOS, OSC_MEM_HOST); }
inline In hostSOA(const void* p) { return In(p, OSC_LAYOUT_SOA, OSC_MEM_HOST); }
inline In deviceAOS(const void* p) { return In(p, OSC_LAYOUT_AOS, OSC_MEM_DEVICE); }
inline In deviceSOA(const void* p) { return In(p, OSC_LAYOUT_SOA, OSC_MEM_DEVICE); }
struct Out {
	void* ptr;
	int layout, mem;
	Out(void* p, int layout_ = OSC_LAYOUT_AOS, int mem_ = OSC_MEM_HOST) : ptr(p), layout(layout_), mem(mem_) {}
};

// ---- Sai2Model(urdf_file) + the world file: the chain constants, the initial posture and the named links, read by the library's
// own URDF reader (osc_urdf_load): examples/04-task_and_redundancy/04-task_and_redundancy.cpp:67 ----
class UrdfModel {
public:
	explicit UrdfModel(const std::string& urdf_path, const std::string& world_path = "", const std::string& robot_name = "") {
		check(osc_urdf_load(urdf_path.c_str(), world_path.empty() ? nullptr : world_path.c_str(), robot_name.empty() ? nullptr : robot_name.c_str(), &_u));
		check(osc_urdf_chain(_u, &_chain, _q0));
	}
	~UrdfModel() { osc_urdf_free(_u); }
	UrdfModel(const UrdfModel&) = delete;
	UrdfModel& operator=(const UrdfModel&) = delete;
	const osc_chain_desc& chain() const { return _chain; }
	const double* q0() const { return _q0; }
	int dof() const { return _chain.n; }
	osc_urdf_t* handle() const { return _u; }

private:
	osc_urdf_t* _u = nullptr;
	osc_chain_desc _chain;
	double _q0[OSC_MAX_DOF];
};

// ---- the robot batch: stands in for std::shared_ptr<Sai2Model::Sai2Model> (one model per robot instance) ----
class RobotBatch {
public:
	RobotBatch(const osc_chain_desc& chain, int64_t n_robots, int precision = OSC_PRECISION_F64, int device = 0, const double* q0 = nullptr) {
		check(osc_batch_create(&chain, n_robots, precision, device, q0, &_h));
	}
	// the same batch sharded over several GPUs of one box (osc_batch_create_multi): every call of this facade fans out over the
	// devices, one host thread and one stream per device inside the library; buffers are HOST buffers spanning all robots
	RobotBatch(const osc_chain_desc& chain, int64_t n_robots, const std::vector<int>& devices, int precision = OSC_PRECISION_F64, const double* q0 = nullptr) {
		check(osc_batch_create_multi(&chain, n_robots, precision, devices.data(), (int)devices.size(), q0, &_h));
	}
	// robots loaded from a URDF (+ world file): q is initialised to the URDF's calibration posture
	RobotBatch(const UrdfModel& model, int64_t n_robots, int precision = OSC_PRECISION_F64, int device = 0) {
		check(osc_batch_create(&model.chain(), n_robots, precision, device, model.q0(), &_h));
	}
	RobotBatch(const UrdfModel& model, int64_t n_robots, const std::vector<int>& devices, int precision = OSC_PRECISION_F64) {
		check(osc_batch_create_multi(&model.chain(), n_robots, precision, devices.data(), (int)devices.size(), model.q0(), &_h));
	}
	int shardCount() const { return osc_batch_shard_count(_h); }
	~RobotBatch() { osc_batch_destroy(_h); }
	RobotBatch(const RobotBatch&) = delete;
	RobotBatch& operator=(const RobotBatch&) = delete;

	// Every call made through handle() first applies a deferred control-cycle update (see setControlCycleFusion); rawHandle() does not.
	osc_batch_t* handle() const { flushControlCycle(); return _h; }
	osc_batch_t* rawHandle() const { return _h; }
	// Control-cycle fusion (on by default).  The reference's cycle is three calls -- robot->updateModel();
	// controller->updateControllerTaskModels(); controller->computeControlTorques() (examples/05...cpp:145-148,195).  With fusion on,
	// the first two are DEFERRED and computeControlTorques() issues the single fused launch (osc_tick) for all three: same results,
	// the rate of RobotController::tick() without touching the loop.  Goal setters and updateSensedForceAndMoment() keep the
	// deferral; any other call in between (setQ/setDq, a getter, a parameter change, a per-task update) first applies the deferred
	// updates with the API-faithful kernels, exactly where the reference would have run them, and that cycle then computes its
	// torques per call.  Turn it off to force the API-faithful kernels everywhere.
	void setControlCycleFusion(bool on) { flushControlCycle(); _fuse = on; }
	bool controlCycleFusion() const { return _fuse; }
	int64_t fusedControlCycles() const { return _fused_cycles; }  // how many computeControlTorques() calls took the fused launch
	void flushControlCycle() const {
		const int p = _pending;
		_pending = 0;
		if (p >= 1) check(osc_update_model(_h));
		if (p == 2) check(osc_update_task_models(_h));
	}
	int dof() const { return osc_batch_dof(_h); }
	int64_t size() const { return osc_batch_size(_h); }
	int precision() const { return osc_batch_precision(_h); }
	// Sai2Model::setQ / setDq / q / dq
	void setQ(const In& q) { set(-1, OSC_F_Q, q); }
	void setDq(const In& dq) { set(-1, OSC_F_DQ, dq); }
	void q(const Out& o) { get(-1, OSC_F_Q, o); }
	void dq(const Out& o) { get(-1, OSC_F_DQ, o); }
	// Sai2Model::updateModel / updateKinematics (examples/05-using_robot_controller/05-using_robot_controller.cpp:143-145)
	void updateModel() {
		flushControlCycle();
		if (_fuse) { _pending = 1; check(osc_defer_model_update(_h)); }  // rows stale: getters / the sensed wrench follow the current q
		else check(osc_update_model(_h));
	}
	void updateKinematics() { flushControlCycle(); check(osc_update_kinematics(_h)); }
	// the simulation side of examples/05...cpp:219-233 on the device (contact-free): sim->setJointTorques(tau); sim->integrate();
	// robot->setQ(...); robot->setDq(...) with tau = the controller's last torques
	void integrate(double dt, int substeps = 1, double joint_damping = 0.0) { flushControlCycle(); check(osc_rollout_step(_h, dt, substeps, joint_damping)); }
	void coriolisPlusGravity(const Out& o) { get(-1, OSC_F_NONLINEAR_EFFECTS, o); }  // of the state the last integrate() started from
	void M(const Out& o) { get(-1, OSC_F_M, o); }
	void MInv(const Out& o) { get(-1, OSC_F_MINV, o); }
	void jointGravityVector(const Out& o) { get(-1, OSC_F_GRAVITY, o); }
	void status(const Out& o) { get(-1, OSC_F_STATUS, o); }
	void synchronize() { check(osc_synchronize(_h)); }
	void setStream(void* cuda_stream) { check(osc_batch_set_stream(_h, cuda_stream)); }
	void setTickOutputs(int outputs) { flushControlCycle(); check(osc_batch_set_tick_outputs(_h, outputs)); }

	void set(int task, int field, const In& v) {
		if (task < 0) flushControlCycle();  // q, dq: the deferred updateModel() belongs to the state before this write
		check(osc_set_field(_h, task, field, v.ptr, v.layout, v.mem));
	}
	void get(int task, int field, const Out& o) { flushControlCycle(); check(osc_get_field(_h, task, field, o.ptr, o.layout, o.mem)); }
	int fieldSize(int task, int field) const { int n = osc_field_size(_h, task, field); check(n); return n; }

	// (used by RobotController)
	bool deferTaskModels() { if (_pending >= 1) { _pending = 2; return true; } return false; }
	bool takeFullControlCycle() { if (_pending == 2) { _pending = 0; ++_fused_cycles; return true; } return false; }

private:
	osc_batch_t* _h = nullptr;
	bool _fuse = true;
	mutable int _pending = 0;  // 1: updateModel() deferred, 2: + updateControllerTaskModels()
	int64_t _fused_cycles = 0;
};

// rigid transform (Eigen::Affine3d stand-in): p_parent = rot * p_child + xyz, rot row-major
struct Affine3 {
	double rot[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
	double xyz[3] = {0, 0, 0};
	static Affine3 Identity() { return Affine3(); }
	static Affine3 Translation(double x, double y, double z) { Affine3 a; a.xyz[0] = x; a.xyz[1] = y; a.xyz[2] = z; return a; }
	Affine3 operator*(const Affine3& o) const {
		Affine3 r;
		for (int i = 0; i < 3; ++i) {
			r.xyz[i] = xyz[i];
			for (int k = 0; k < 3; ++k) r.xyz[i] += rot[3 * i + k] * o.xyz[k];
			for (int j = 0; j < 3; ++j) {
				r.rot[3 * i + j] = 0;
				for (int k = 0; k < 3; ++k) r.rot[3 * i + j] += rot[3 * i + k] * o.rot[3 * k + j];
			}
		}
		return r;
	}
};
// A named link of the URDF: the movable body it is fixed in and its pose in that body (identity for a movable link).
// Stands in for the `link_name` string of the reference (fixed links are merged into their parent).
struct Link {
	int body = 0;
	Affine3 in_body;
	Link() {}
	Link(int body_, const Affine3& in_body_ = Affine3()) : body(body_), in_body(in_body_) {}
	// the `link_name` of the reference API, resolved by the URDF reader (fixed links are frames of their movable parent)
	Link(const UrdfModel& model, const std::string& link_name) { check(osc_urdf_link(model.handle(), link_name.c_str(), &body, in_body.xyz, in_body.rot)); }
};

// ---- src/tasks/TemplateTask.h:25-123 ----
class TemplateTask {
public:
	virtual ~TemplateTask() {}
	TaskType getTaskType() const { return _type; }
	const std::string& getTaskName() const { return _name; }
	double getLoopTimestep() const { return _dt; }
	int index() const { return _index; }
	const std::shared_ptr<RobotBatch>& robot() const { return _robot; }

	// updateTaskModel(N_prec): N_prec = identity ...
	void updateTaskModel() { check(osc_task_update_model(_robot->handle(), _index, -1)); }
	// ... or previous.getTaskAndPreviousNullspace()
	void updateTaskModel(const TemplateTask& previous) { check(osc_task_update_model(_robot->handle(), _index, previous._index)); }
	// VectorXd computeTorques(): result in getTaskTorques()
	void computeTorques() { check(osc_task_compute_torques(_robot->handle(), _index, 0)); }
	void computeTorques(const Out& tau) { computeTorques(); getTaskTorques(tau); }
	// VectorXd computeTorques(const VectorXd& tau_prec) with tau_prec = the controller's running sum (RobotController.cpp:66-68)
	void computeTorquesWithPreviousTorques() { check(osc_task_compute_torques(_robot->handle(), _index, 1)); }
	void getTaskTorques(const Out& o) { _robot->get(_index, OSC_F_TASK_TORQUES, o); }
	void reInitializeTask() { check(osc_reinitialize_tasks(_robot->handle(), _index)); }
	void getTaskNullspace(const Out& o) { _robot->get(_index, OSC_F_N_TASK, o); }
	void getPreviousTasksNullspace(const Out& o) { _robot->get(_index, OSC_F_N_PREC, o); }
	void getTaskAndPreviousNullspace(const Out& o) { _robot->get(_index, OSC_F_N_TASK_AND_PREC, o); }

protected:
	TemplateTask(const std::shared_ptr<RobotBatch>& robot, const std::string& name, TaskType type, double dt)
		: _robot(robot), _name(name), _type(type), _dt(dt) {}
	std::shared_ptr<RobotBatch> _robot;
	std::string _name;
	TaskType _type;
	double _dt;
	int _index = -1;
};

// ---- src/tasks/MotionForceTask.h ----
class MotionForceTask : public TemplateTask {
public:
	// MotionForceTask(robot, link_name, compliant_frame, task_name, is_force_motion_parametrization_in_compliant_frame,
	//                 loop_timestep)   src/tasks/MotionForceTask.h:96-101, MotionForceTask.cpp:16-31
	MotionForceTask(const std::shared_ptr<RobotBatch>& robot, const Link& link, const Affine3& compliant_frame = Affine3::Identity(),
					const std::string& task_name = "motion_force_task", bool is_force_motion_parametrization_in_compliant_frame = false,
					double loop_timestep = 0.001)
		: TemplateTask(robot, task_name, MOTION_FORCE_TASK, loop_timestep) {
		attach(link, compliant_frame, nullptr, nullptr, is_force_motion_parametrization_in_compliant_frame);
	}
	// partial task: controlled translation / rotation directions   MotionForceTask.h:103-110, MotionForceTask.cpp:33-90
	MotionForceTask(const std::shared_ptr<RobotBatch>& robot, const Link& link, const std::vector<std::array<double, 3>>& controlled_directions_translation,
					const std::vector<std::array<double, 3>>& controlled_directions_rotation, const Affine3& compliant_frame = Affine3::Identity(),
					const std::string& task_name = "partial_motion_force_task", bool is_force_motion_parametrization_in_compliant_frame = false,
					double loop_timestep = 0.001)
		: TemplateTask(robot, task_name, MOTION_FORCE_TASK, loop_timestep) {
		attach(link, compliant_frame, &controlled_directions_translation, &controlled_directions_rotation, is_force_motion_parametrization_in_compliant_frame);
	}

	// ---- goals (MotionForceTask.h:211-247, 362, 376) ----
	void setGoalPosition(const In& v) { _robot->set(_index, OSC_F_GOAL_POSITION, v); }
	void setGoalOrientation(const In& R) { _robot->set(_index, OSC_F_GOAL_ORIENTATION, R); }
	void setGoalLinearVelocity(const In& v) { _robot->set(_index, OSC_F_GOAL_LINEAR_VELOCITY, v); }
	void setGoalAngularVelocity(const In& v) { _robot->set(_index, OSC_F_GOAL_ANGULAR_VELOCITY, v); }
	void setGoalLinearAcceleration(const In& v) { _robot->set(_index, OSC_F_GOAL_LINEAR_ACCELERATION, v); }
	void setGoalAngularAcceleration(const In& v) { _robot->set(_index, OSC_F_GOAL_ANGULAR_ACCELERATION, v); }
	void setGoalForce(const In& v) { _robot->set(_index, OSC_F_GOAL_FORCE, v); }
	void setGoalMoment(const In& v) { _robot->set(_index, OSC_F_GOAL_MOMENT, v); }
	void getGoalPosition(const Out& o) { _robot->get(_index, OSC_F_GOAL_POSITION, o); }
	void getGoalOrientation(const Out& o) { _robot->get(_index, OSC_F_GOAL_ORIENTATION, o); }
	void getGoalLinearVelocity(const Out& o) { _robot->get(_index, OSC_F_GOAL_LINEAR_VELOCITY, o); }
	void getGoalAngularVelocity(const Out& o) { _robot->get(_index, OSC_F_GOAL_ANGULAR_VELOCITY, o); }
	// ---- state getters (MotionForceTask.h:121-205, 508-522) ----
	void getCurrentPosition(const Out& o) { _robot->get(_index, OSC_F_CURRENT_POSITION, o); }
	void getCurrentOrientation(const Out& o) { _robot->get(_index, OSC_F_CURRENT_ORIENTATION, o); }
	void getCurrentLinearVelocity(const Out& o) { _robot->get(_index, OSC_F_CURRENT_LINEAR_VELOCITY, o); }
	void getCurrentAngularVelocity(const Out& o) { _robot->get(_index, OSC_F_CURRENT_ANGULAR_VELOCITY, o); }
	void getDesiredPosition(const Out& o) { _robot->get(_index, OSC_F_DESIRED_POSITION, o); }
	void getDesiredOrientation(const Out& o) { _robot->get(_index, OSC_F_DESIRED_ORIENTATION, o); }
	void getDesiredLinearVelocity(const Out& o) { _robot->get(_index, OSC_F_DESIRED_LINEAR_VELOCITY, o); }
	void getDesiredAngularVelocity(const Out& o) { _robot->get(_index, OSC_F_DESIRED_ANGULAR_VELOCITY, o); }
	// Vector3d getPositionError() / getOrientationError() const (MotionForceTask.h:268-269, .cpp:540-546)
	void getPositionError(const Out& o) { _robot->get(_index, OSC_F_POSITION_ERROR, o); }
	void getOrientationError(const Out& o) { _robot->get(_index, OSC_F_SELECTED_ORIENTATION_ERROR, o); }
	// Matrix3d sigmaForce / sigmaPosition / sigmaMoment / sigmaOrientation() const (MotionForceTask.h:611-614): [B][9] row-major
	void sigmaForce(const Out& o) { _robot->get(_index, OSC_F_SIGMA_FORCE, o); }
	void sigmaPosition(const Out& o) { _robot->get(_index, OSC_F_SIGMA_POSITION, o); }
	void sigmaMoment(const Out& o) { _robot->get(_index, OSC_F_SIGMA_MOMENT, o); }
	void sigmaOrientation(const Out& o) { _robot->get(_index, OSC_F_SIGMA_ORIENTATION, o); }
	// Matrix3d posSelectionProjector() / oriSelectionProjector() const (MotionForceTask.h:653-659): uniform over the batch
	std::array<double, 9> posSelectionProjector() const { std::array<double, 9> a; check(osc_mft_selection_projectors(_robot->handle(), _index, a.data(), nullptr)); return a; }
	std::array<double, 9> oriSelectionProjector() const { std::array<double, 9> a; check(osc_mft_selection_projectors(_robot->handle(), _index, nullptr, a.data())); return a; }
	// getGoalForce() / getGoalMoment() (MotionForceTask.cpp:755-769: rotated into the world frame when parametrised in the compliant frame)
	void getGoalForce(const Out& o) { _robot->get(_index, OSC_F_GOAL_FORCE_WORLD, o); }
	void getGoalMoment(const Out& o) { _robot->get(_index, OSC_F_GOAL_MOMENT_WORLD, o); }
	// bool goalPositionReached(tolerance) / goalOrientationReached(tolerance) (MotionForceTask.h:508,521) for every robot:
	// returns how many robots reached; flags (int32 [B], host) may be null
	int64_t goalPositionReached(double tolerance, int32_t* flags = nullptr) {
		int64_t n = 0; check(osc_task_goal_reached(_robot->handle(), _index, 0, tolerance, flags, OSC_MEM_HOST, &n)); return n;
	}
	int64_t goalOrientationReached(double tolerance, int32_t* flags = nullptr) {
		int64_t n = 0; check(osc_task_goal_reached(_robot->handle(), _index, 1, tolerance, flags, OSC_MEM_HOST, &n)); return n;
	}
	void getUnitMassForce(const Out& o) { _robot->get(_index, OSC_F_UNIT_MASS_FORCE, o); }
	void getSensedForceControlWorldFrame(const Out& o) { _robot->get(_index, OSC_F_SENSED_FORCE_WORLD, o); }
	void getSensedMomentControlWorldFrame(const Out& o) { _robot->get(_index, OSC_F_SENSED_MOMENT_WORLD, o); }
	void getJacobian(const Out& o) { _robot->get(_index, OSC_F_JACOBIAN, o); }
	void getProjectedJacobian(const Out& o) { _robot->get(_index, OSC_F_PROJECTED_JACOBIAN, o); }
	void getLambdaWorldFrame(const Out& o) { _robot->get(_index, OSC_F_LAMBDA_WORLD, o); }
	void getLambdaModifiedWorldFrame(const Out& o) { _robot->get(_index, OSC_F_LAMBDA_MODIFIED_WORLD, o); }
	void getSingularityBlendingCoefficient(const Out& o) { _robot->get(_index, OSC_F_ALPHA, o); }
	void getSingularValues(const Out& o) { _robot->get(_index, OSC_F_SINGULAR_VALUES, o); }
	void getSingularityTypes(const Out& o) { _robot->get(_index, OSC_F_SINGULARITY_TYPES, o); }
	int getTaskDof() const { return osc_batch_task_dof(_robot->handle(), _index); }

	// ---- gains (MotionForceTask.h:272-355) ----
	void setPosControlGains(const PIDGains& g) { setPosControlGains(g.kp, g.kv, g.ki); }
	void setPosControlGains(double kp, double kv, double ki = 0) { setPosControlGains({kp, kp, kp}, {kv, kv, kv}, {ki, ki, ki}); }
	void setPosControlGains(const std::array<double, 3>& kp, const std::array<double, 3>& kv, const std::array<double, 3>& ki = {0, 0, 0}) {
		edit([&](osc_mft_params& p) { for (int a = 0; a < 3; ++a) { p.kp_pos[a] = kp[a]; p.kv_pos[a] = kv[a]; p.ki_pos[a] = ki[a]; } });
	}
	void setOriControlGains(const PIDGains& g) { setOriControlGains(g.kp, g.kv, g.ki); }
	void setOriControlGains(double kp, double kv, double ki = 0) { setOriControlGains({kp, kp, kp}, {kv, kv, kv}, {ki, ki, ki}); }
	void setOriControlGains(const std::array<double, 3>& kp, const std::array<double, 3>& kv, const std::array<double, 3>& ki = {0, 0, 0}) {
		edit([&](osc_mft_params& p) { for (int a = 0; a < 3; ++a) { p.kp_ori[a] = kp[a]; p.kv_ori[a] = kv[a]; p.ki_ori[a] = ki[a]; } });
	}
	void setForceControlGains(const PIDGains& g) { setForceControlGains(g.kp, g.kv, g.ki); }
	void setForceControlGains(double kp, double kv, double ki) { edit([&](osc_mft_params& p) { p.kp_force = kp; p.kv_force = kv; p.ki_force = ki; }); }
	void setMomentControlGains(const PIDGains& g) { setMomentControlGains(g.kp, g.kv, g.ki); }
	void setMomentControlGains(double kp, double kv, double ki) { edit([&](osc_mft_params& p) { p.kp_moment = kp; p.kv_moment = kv; p.ki_moment = ki; }); }
	void setFeedforwardForceGain(double k) { edit([&](osc_mft_params& p) { p.kff_force = k; }); }
	void setFeedforwardMomentGain(double k) { edit([&](osc_mft_params& p) { p.kff_moment = k; }); }
	void setMaxForceControlFeedbackOutput(double v) { edit([&](osc_mft_params& p) { p.max_force_control_feedback_output = v; }); }
	void setMaxMomentControlFeedbackOutput(double v) { edit([&](osc_mft_params& p) { p.max_moment_control_feedback_output = v; }); }

	// ---- force / motion space parametrisation (MotionForceTask.h:576-623, MotionForceTask.cpp:830-890, 973-986) ----
	// returns true when the parametrisation changed (goal <- current pose, integrators reset: .cpp:850-856)
	bool parametrizeForceMotionSpaces(int force_space_dimension, const std::array<double, 3>& force_or_motion_single_axis = {0, 0, 0}) {
		const int mask = edit([&](osc_mft_params& p) {
			p.force_space_dimension = force_space_dimension;
			for (int a = 0; a < 3; ++a) p.force_or_motion_axis[a] = force_or_motion_single_axis[a];
		});
		return (mask & 1) != 0;
	}
	bool parametrizeMomentRotMotionSpaces(int moment_space_dimension, const std::array<double, 3>& moment_or_rot_motion_single_axis = {0, 0, 0}) {
		const int mask = edit([&](osc_mft_params& p) {
			p.moment_space_dimension = moment_space_dimension;
			for (int a = 0; a < 3; ++a) p.moment_or_rotmotion_axis[a] = moment_or_rot_motion_single_axis[a];
		});
		return (mask & 2) != 0;
	}
	void setClosedLoopForceControl(bool closed_loop = true) { edit([&](osc_mft_params& p) { p.closed_loop_force_control = closed_loop; }); }
	void setClosedLoopMomentControl(bool closed_loop = true) { edit([&](osc_mft_params& p) { p.closed_loop_moment_control = closed_loop; }); }
	void enablePassivity() { edit([&](osc_mft_params& p) { p.passivity_enabled = 1; }); }
	void disablePassivity() { edit([&](osc_mft_params& p) { p.passivity_enabled = 0; }); }
	// MotionForceTask::setForceSensorFrame (.cpp:794-803): the sensor pose IN THE CONTROL LINK; stored relative to the compliant frame
	void setForceSensorFrame(const Affine3& transformation_in_link) {
		Affine3 inv;  // compliant_frame^-1
		for (int i = 0; i < 3; ++i) {
			for (int j = 0; j < 3; ++j) inv.rot[3 * i + j] = _compliant_frame.rot[3 * j + i];
		}
		for (int i = 0; i < 3; ++i) {
			inv.xyz[i] = 0;
			for (int k = 0; k < 3; ++k) inv.xyz[i] -= inv.rot[3 * i + k] * _compliant_frame.xyz[k];
		}
		const Affine3 T = inv * transformation_in_link;
		edit([&](osc_mft_params& p) {
			for (int k = 0; k < 9; ++k) p.sensor_rot[k] = T.rot[k];
			for (int k = 0; k < 3; ++k) p.sensor_xyz[k] = T.xyz[k];
		});
	}
	// MotionForceTask::updateSensedForceAndMoment (.cpp:805-828): sensor-frame wrench of every robot
	void updateSensedForceAndMoment(const In& sensed_force_sensor_frame, const In& sensed_moment_sensor_frame) {
		if (sensed_force_sensor_frame.layout != sensed_moment_sensor_frame.layout || sensed_force_sensor_frame.mem != sensed_moment_sensor_frame.mem)
			throw std::invalid_argument("force and moment must use the same layout and memory space in MotionForceTask::updateSensedForceAndMoment\n");
		check(osc_mft_update_sensed_force_and_moment(_robot->rawHandle(), _index, sensed_force_sensor_frame.ptr, sensed_moment_sensor_frame.ptr,
													 sensed_force_sensor_frame.layout, sensed_force_sensor_frame.mem));
	}

	// ---- velocity saturation, decoupling, singularity handling (MotionForceTask.h:430-432, 669-753) ----
	void enableVelocitySaturation(double linear_vel_sat = 0.3, double angular_vel_sat = 3.14159265358979323846 / 3) {
		if (linear_vel_sat <= 0 || angular_vel_sat <= 0)
			throw std::invalid_argument("Velocity saturation values should be strictly positive or zero in MotionForceTask::enableVelocitySaturation\n");
		edit([&](osc_mft_params& p) { p.use_velocity_saturation = 1; p.linear_saturation_velocity = linear_vel_sat; p.angular_saturation_velocity = angular_vel_sat; });
	}
	void disableVelocitySaturation() { edit([&](osc_mft_params& p) { p.use_velocity_saturation = 0; }); }
	void setDynamicDecouplingType(DynamicDecouplingType type) { edit([&](osc_mft_params& p) { p.dynamic_decoupling_type = type; }); }
	void setBoundedInertiaEstimateThreshold(double threshold) { edit([&](osc_mft_params& p) { p.bie_threshold = threshold; }); }
	void handleAllSingularitiesAsType1(bool flag) { edit([&](osc_mft_params& p) { p.enforce_type_1_strategy = flag; }); }
	void setType1Posture(const In& q_prior) { _robot->set(_index, OSC_F_Q_PRIOR, q_prior); }
	void enableSingularityHandling() { edit([&](osc_mft_params& p) { p.singularity_handling_enabled = 1; }); }
	void disableSingularityHandling() { edit([&](osc_mft_params& p) { p.singularity_handling_enabled = 0; }); }
	void setSingularityHandlingBounds(double s_min, double s_max) { edit([&](osc_mft_params& p) { p.s_min = s_min; p.s_max = s_max; }); }
	void setSingularityHandlingGains(double kp_type_1, double kv_type_1, double kv_type_2) {
		edit([&](osc_mft_params& p) { p.kp_type_1 = kp_type_1; p.kv_type_1 = kv_type_1; p.kv_type_2 = kv_type_2; });
	}
	// ---- internal OTG (MotionForceTask.h:387-427, MotionForceTask.cpp:511-538): on by default, acceleration limited ----
	void enableInternalOtgAccelerationLimited(double max_linear_velocity, double max_linear_acceleration, double max_angular_velocity,
											  double max_angular_acceleration) {
		const double lim[4] = {max_linear_velocity, max_angular_velocity, max_linear_acceleration, max_angular_acceleration};
		check(osc_task_enable_internal_otg_acceleration_limited(_robot->handle(), _index, lim, 2));
	}
	// third-order (jerk-limited) profiles are not built: OSC_ERR_UNSUPPORTED -> std::runtime_error
	void enableInternalOtgJerkLimited(double, double, double, double, double, double) {
		check(osc_task_enable_internal_otg_jerk_limited(_robot->handle(), _index, nullptr, 0));
	}
	void disableInternalOtg() { check(osc_task_disable_internal_otg(_robot->handle(), _index)); }
	bool getInternalOtgEnabled() const { return osc_task_internal_otg_enabled(_robot->handle(), _index) != 0; }
	void getDesiredLinearAcceleration(const Out& o) { _robot->get(_index, OSC_F_DESIRED_LINEAR_ACCELERATION, o); }
	void getDesiredAngularAcceleration(const Out& o) { _robot->get(_index, OSC_F_DESIRED_ANGULAR_ACCELERATION, o); }
	void resetIntegrators() {
		const std::array<double, 3> z = {0, 0, 0};
		_robot->set(_index, OSC_F_INTEGRATED_POSITION_ERROR, z);
		_robot->set(_index, OSC_F_INTEGRATED_ORIENTATION_ERROR, z);
		_robot->set(_index, OSC_F_INTEGRATED_FORCE_ERROR, z);
		_robot->set(_index, OSC_F_INTEGRATED_MOMENT_ERROR, z);
	}
	osc_mft_params getParams() const { osc_mft_params p; check(osc_mft_get_params(_robot->handle(), _index, &p)); return p; }

private:
	template <class F> int edit(F f) {
		osc_mft_params p = getParams();
		f(p);
		const int rc = osc_mft_set_params(_robot->handle(), _index, &p);
		check(rc);
		return rc;
	}
	void attach(const Link& link, const Affine3& compliant_frame, const std::vector<std::array<double, 3>>* dt, const std::vector<std::array<double, 3>>* dr,
				bool in_compliant_frame) {
		_compliant_frame = compliant_frame;
		const Affine3 f = link.in_body * compliant_frame;
		osc_mft_desc d{};
		d.body = link.body;
		for (int k = 0; k < 3; ++k) d.frame_xyz[k] = f.xyz[k];
		for (int k = 0; k < 9; ++k) d.frame_rot[k] = f.rot[k];
		d.partial = dt != nullptr;
		if (dt) {
			if (dt->size() > 3 || dr->size() > 3) throw std::invalid_argument("at most 3 controlled directions per block in MotionForceTask\n");
			d.n_dirs_translation = (int)dt->size();
			d.n_dirs_rotation = (int)dr->size();
			for (size_t c = 0; c < dt->size(); ++c) for (int k = 0; k < 3; ++k) d.dirs_translation[c][k] = (*dt)[c][k];
			for (size_t c = 0; c < dr->size(); ++c) for (int k = 0; k < 3; ++k) d.dirs_rotation[c][k] = (*dr)[c][k];
		}
		d.parametrization_in_compliant_frame = in_compliant_frame;
		d.loop_timestep = _dt;
		_index = osc_batch_add_motion_force_task(_robot->handle(), &d);
		check(_index);
		// DefaultParameters::use_internal_otg = true, acceleration limited (MotionForceTask.h:67-72, MotionForceTask.cpp:173-188)
		enableInternalOtgAccelerationLimited(0.3, 2.0, 3.14159265358979323846 / 3, 2 * 3.14159265358979323846);
	}
	Affine3 _compliant_frame;
};

// ---- src/tasks/JointTask.h ----
class JointTask : public TemplateTask {
public:
	// JointTask(robot, task_name, loop_timestep)   JointTask.h:56-60, JointTask.cpp:14-21
	JointTask(const std::shared_ptr<RobotBatch>& robot, const std::string& task_name = "joint_task", double loop_timestep = 0.001)
		: TemplateTask(robot, task_name, JOINT_TASK, loop_timestep) {
		osc_joint_task_desc d{};
		d.full = 1; d.k = robot->dof(); d.loop_timestep = loop_timestep;
		_k = d.k;
		_index = osc_batch_add_joint_task(_robot->handle(), &d);
		check(_index);
		enableInternalOtgAccelerationLimited(3.14159265358979323846 / 3, 2 * 3.14159265358979323846);  // JointTask.h:38-41, JointTask.cpp:73-83
	}
	// JointTask(robot, joint_selection_matrix (k x n, row-major, full row rank), task_name, loop_timestep)   JointTask.h:70-75, .cpp:23-43
	JointTask(const std::shared_ptr<RobotBatch>& robot, const std::vector<double>& joint_selection_matrix, int task_dof,
			  const std::string& task_name = "partial_joint_task", double loop_timestep = 0.001)
		: TemplateTask(robot, task_name, JOINT_TASK, loop_timestep) {
		const int n = robot->dof();
		if (task_dof < 1 || task_dof > n || (int)joint_selection_matrix.size() != task_dof * n)
			throw std::invalid_argument("joint selection matrix size not consistent with robot dof in JointTask constructor\n");
		osc_joint_task_desc d{};
		d.full = 0; d.k = task_dof; d.loop_timestep = loop_timestep;
		for (int i = 0; i < task_dof * n; ++i) d.selection[i] = joint_selection_matrix[i];
		_k = task_dof;
		_index = osc_batch_add_joint_task(_robot->handle(), &d);
		check(_index);
		enableInternalOtgAccelerationLimited(3.14159265358979323846 / 3, 2 * 3.14159265358979323846);  // JointTask.h:38-41, JointTask.cpp:73-83
	}
	bool isFullJointTask() const { return _k == _robot->dof(); }  // JointTask.h:123
	int getTaskDof() const { return _k; }

	// goals (JointTask.h:137-172)
	void setGoalPosition(const In& v) { _robot->set(_index, OSC_F_JT_GOAL_POSITION, v); }
	void setGoalVelocity(const In& v) { _robot->set(_index, OSC_F_JT_GOAL_VELOCITY, v); }
	void setGoalAcceleration(const In& v) { _robot->set(_index, OSC_F_JT_GOAL_ACCELERATION, v); }
	void getGoalPosition(const Out& o) { _robot->get(_index, OSC_F_JT_GOAL_POSITION, o); }
	void getCurrentPosition(const Out& o) { _robot->get(_index, OSC_F_JT_CURRENT_POSITION, o); }
	void getCurrentVelocity(const Out& o) { _robot->get(_index, OSC_F_JT_CURRENT_VELOCITY, o); }
	void getDesiredVelocity(const Out& o) { _robot->get(_index, OSC_F_JT_DESIRED_VELOCITY, o); }
	void getLambdaJointFrame(const Out& o) { _robot->get(_index, OSC_F_JT_LAMBDA_WORLD, o); }
	// gains (JointTask.h:234-254)
	void setGains(const PIDGains& g) { setGains(g.kp, g.kv, g.ki); }
	void setGains(double kp, double kv, double ki = 0) {
		edit([&](osc_joint_task_params& p) { for (int i = 0; i < _k; ++i) { p.kp[i] = kp; p.kv[i] = kv; p.ki[i] = ki; } });
	}
	void setGains(const std::vector<double>& kp, const std::vector<double>& kv, const std::vector<double>& ki = {}) {
		if ((int)kp.size() != _k || (int)kv.size() != _k || (!ki.empty() && (int)ki.size() != _k))
			throw std::invalid_argument("size of gain vectors inconsistent with number of task dofs in JointTask::setGains\n");
		edit([&](osc_joint_task_params& p) { for (int i = 0; i < _k; ++i) { p.kp[i] = kp[i]; p.kv[i] = kv[i]; p.ki[i] = ki.empty() ? 0 : ki[i]; } });
	}
	// setGainsUnsafe (JointTask.h:256, .cpp:136-156): no sign check; vectors of size 1 (isotropic) or task dof
	void setGainsUnsafe(const std::vector<double>& kp, const std::vector<double>& kv, const std::vector<double>& ki) {
		if (kp.size() != kv.size() || kp.size() != ki.size())
			throw std::invalid_argument("size of gain vectors inconsistent with number of task dofs in JointTask::setGains\n");
		check(osc_joint_task_set_gains_unsafe(_robot->handle(), _index, kp.data(), kv.data(), ki.data(), (int)kp.size()));
	}
	// bool goalPositionReached(tol = 1e-2) (JointTask.h:389) for every robot: how many reached; flags (int32 [B], host) may be null
	int64_t goalPositionReached(double tolerance = 1e-2, int32_t* flags = nullptr) {
		int64_t n = 0; check(osc_task_goal_reached(_robot->handle(), _index, 0, tolerance, flags, OSC_MEM_HOST, &n)); return n;
	}
	void enableVelocitySaturation(double saturation_velocity = 3.14159265358979323846 / 3) {
		edit([&](osc_joint_task_params& p) { p.use_velocity_saturation = 1; for (int i = 0; i < _k; ++i) p.saturation_velocity[i] = saturation_velocity; });
	}
	void enableVelocitySaturation(const std::vector<double>& saturation_velocity) {
		if ((int)saturation_velocity.size() != _k)
			throw std::invalid_argument("saturation velocity vector does not have the correct size in JointTask::enableVelocitySaturation\n");
		edit([&](osc_joint_task_params& p) { p.use_velocity_saturation = 1; for (int i = 0; i < _k; ++i) p.saturation_velocity[i] = saturation_velocity[i]; });
	}
	void disableVelocitySaturation() { edit([&](osc_joint_task_params& p) { p.use_velocity_saturation = 0; }); }
	void setDynamicDecouplingType(DynamicDecouplingType type) { edit([&](osc_joint_task_params& p) { p.dynamic_decoupling_type = type; }); }
	void setBoundedInertiaEstimateThreshold(double threshold) { edit([&](osc_joint_task_params& p) { p.bie_threshold = threshold; }); }
	// ---- internal OTG (JointTask.h:262-324, JointTask.cpp:358-406): on by default, acceleration limited ----
	void enableInternalOtgAccelerationLimited(double max_velocity, double max_acceleration) {
		const double lim[2] = {max_velocity, max_acceleration};
		check(osc_task_enable_internal_otg_acceleration_limited(_robot->handle(), _index, lim, 1));
	}
	void enableInternalOtgAccelerationLimited(const std::vector<double>& max_velocity, const std::vector<double>& max_acceleration) {
		if (max_velocity.size() == 1 && max_acceleration.size() == 1 && _k != 1) return enableInternalOtgAccelerationLimited(max_velocity[0], max_acceleration[0]);
		if ((int)max_velocity.size() != _k || (int)max_acceleration.size() != _k)
			throw std::invalid_argument("max velocity or max acceleration vector size not consistent with task dof in JointTask::enableInternalOtgAccelerationLimited\n");
		std::vector<double> lim(max_velocity);
		lim.insert(lim.end(), max_acceleration.begin(), max_acceleration.end());
		check(osc_task_enable_internal_otg_acceleration_limited(_robot->handle(), _index, lim.data(), _k));
	}
	void enableInternalOtgJerkLimited(double, double, double) { check(osc_task_enable_internal_otg_jerk_limited(_robot->handle(), _index, nullptr, 0)); }
	void disableInternalOtg() { check(osc_task_disable_internal_otg(_robot->handle(), _index)); }
	bool getInternalOtgEnabled() const { return osc_task_internal_otg_enabled(_robot->handle(), _index) != 0; }
	void getDesiredPosition(const Out& o) { _robot->get(_index, OSC_F_JT_DESIRED_POSITION, o); }
	void getDesiredAcceleration(const Out& o) { _robot->get(_index, OSC_F_JT_DESIRED_ACCELERATION, o); }
	void resetIntegrators() { _robot->set(_index, OSC_F_JT_INTEGRATED_ERROR, std::vector<double>((size_t)_k, 0.0)); }  // JointTask.h:394
	osc_joint_task_params getParams() const { osc_joint_task_params p; check(osc_joint_task_get_params(_robot->handle(), _index, &p)); return p; }

private:
	template <class F> void edit(F f) {
		osc_joint_task_params p = getParams();
		f(p);
		check(osc_joint_task_set_params(_robot->handle(), _index, &p));
	}
	int _k = 0;
};

// ---- src/RobotController.h:22-51 ----
class RobotController {
public:
	// RobotController(robot, tasks): the rules of RobotController.cpp:8-51.  Same-dt and nothing-after-a-full-joint-task are
	// enforced by the batch when the tasks are constructed; here: at least one task, same robot, unique names, priority
	// order = construction order.
	RobotController(const std::shared_ptr<RobotBatch>& robot, const std::vector<std::shared_ptr<TemplateTask>>& tasks) : _robot(robot), _tasks(tasks) {
		if (tasks.empty()) throw std::invalid_argument("RobotController must have at least one task");
		for (size_t i = 0; i < tasks.size(); ++i) {
			if (tasks[i]->robot() != robot) throw std::invalid_argument("All tasks must have the same robot model in RobotController");
			if (tasks[i]->getLoopTimestep() != tasks[0]->getLoopTimestep()) throw std::invalid_argument("All tasks must have the same loop timestep in RobotController");
			for (size_t j = 0; j < i; ++j)
				if (tasks[j]->getTaskName() == tasks[i]->getTaskName()) throw std::invalid_argument("Tasks in RobotController must have unique names");
			if (tasks[i]->index() != (int)i)
				throw std::invalid_argument("tasks must be given to RobotController in the order they were constructed on the robot batch (priority order)");
		}
		if ((int)tasks.size() != osc_batch_task_count(robot->handle()))
			throw std::invalid_argument("every task constructed on the robot batch must be given to RobotController");
	}
	void updateControllerTaskModels() {  // RobotController.cpp:53-60
		if (!_robot->deferTaskModels()) check(osc_update_task_models(_robot->handle()));
	}
	// VectorXd computeControlTorques()   RobotController.cpp:62-74.  After a deferred updateModel() + updateControllerTaskModels()
	// (RobotBatch::setControlCycleFusion) the three calls are one fused launch.
	void computeControlTorques(const Out& tau) {
		if (_robot->takeFullControlCycle()) check(osc_tick(_robot->rawHandle(), 1, tau.ptr, tau.layout, tau.mem));
		else check(osc_compute_torques(_robot->handle(), 1, tau.ptr, tau.layout, tau.mem));
	}
	void computeControlTorques() {
		if (_robot->takeFullControlCycle()) check(osc_tick(_robot->rawHandle(), 1, nullptr, 0, 0));
		else check(osc_compute_torques(_robot->handle(), 1, nullptr, 0, 0));
	}
	// robot->updateModel(); updateControllerTaskModels(); computeControlTorques() in one (fused) launch
	void tick(const Out& tau) { check(osc_tick(_robot->handle(), 1, tau.ptr, tau.layout, tau.mem)); }
	void tick() { check(osc_tick(_robot->handle(), 1, nullptr, 0, 0)); }
	// `ticks` control cycles with the device-side dynamics in between (MPC rollout): no host synchronisation inside
	void rollout(int ticks, double sim_dt, int substeps = 1, double joint_damping = 0.0) {
		check(osc_rollout(_robot->handle(), 1, ticks, sim_dt, substeps, joint_damping, nullptr, 0, 0));
	}
	void getControlTorques(const Out& tau) { _robot->get(-1, OSC_F_TAU, tau); }
	void enableGravityCompensation(bool enable) { check(osc_enable_gravity_compensation(_robot->handle(), enable)); }  // RobotController.h:31-33
	void reinitializeTasks() { check(osc_reinitialize_tasks(_robot->handle(), -1)); }
	std::vector<std::string> getTaskNames() const {
		std::vector<std::string> n;
		for (auto& t : _tasks) n.push_back(t->getTaskName());
		return n;
	}
	std::shared_ptr<JointTask> getJointTaskByName(const std::string& name) const {
		for (auto& t : _tasks)
			if (t->getTaskName() == name) {
				if (t->getTaskType() != JOINT_TASK) throw std::invalid_argument("Task [" + name + "] is not a JointTask");
				return std::static_pointer_cast<JointTask>(t);
			}
		throw std::invalid_argument("Task [" + name + "] not found");
	}
	std::shared_ptr<MotionForceTask> getMotionForceTaskByName(const std::string& name) const {
		for (auto& t : _tasks)
			if (t->getTaskName() == name) {
				if (t->getTaskType() != MOTION_FORCE_TASK) throw std::invalid_argument("Task [" + name + "] is not a MotionForceTask");
				return std::static_pointer_cast<MotionForceTask>(t);
			}
		throw std::invalid_argument("Task [" + name + "] not found");
	}

private:
	std::shared_ptr<RobotBatch> _robot;
	std::vector<std::shared_ptr<TemplateTask>> _tasks;
};

}  // namespace Sai2PrimitivesBatched
#endif  // SAI2_PRIMITIVES_BATCHED_H_
