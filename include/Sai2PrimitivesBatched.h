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

// ---- the robot batch: stands in for std::shared_ptr<Sai2Model::Sai2Model> (one model per robot instance) ----
class RobotBatch {
public:
	RobotBatch(const osc_chain_desc& chain, int64_t n_robots, int precision = OSC_PRECISION_F64, int device = 0, const double* q0 = nullptr) {
		check(osc_batch_create(&chain, n_robots, precision, device, q0, &_h));
	}
	~RobotBatch() { osc_batch_destroy(_h); }
	RobotBatch(const RobotBatch&) = delete;
	RobotBatch& operator=(const RobotBatch&) = delete;

	osc_batch_t* handle() const { return _h; }
	int dof() const { return osc_batch_dof(_h); }
	int64_t size() const { return osc_batch_size(_h); }
	int precision() const { return osc_batch_precision(_h); }
	// Sai2Model::setQ / setDq / q / dq
	void setQ(const In& q) { set(-1, OSC_F_Q, q); }
	void setDq(const In& dq) { set(-1, OSC_F_DQ, dq); }
	void q(const Out& o) { get(-1, OSC_F_Q, o); }
	void dq(const Out& o) { get(-1, OSC_F_DQ, o); }
	// Sai2Model::updateModel / updateKinematics (examples/05-using_robot_controller/05-using_robot_controller.cpp:143-145)
	void updateModel() { check(osc_update_model(_h)); }
	void updateKinematics() { check(osc_update_kinematics(_h)); }
	void M(const Out& o) { get(-1, OSC_F_M, o); }
	void MInv(const Out& o) { get(-1, OSC_F_MINV, o); }
	void jointGravityVector(const Out& o) { get(-1, OSC_F_GRAVITY, o); }
	void status(const Out& o) { get(-1, OSC_F_STATUS, o); }
	void synchronize() { check(osc_synchronize(_h)); }
	void setStream(void* cuda_stream) { check(osc_batch_set_stream(_h, cuda_stream)); }
	void setTickOutputs(int outputs) { check(osc_batch_set_tick_outputs(_h, outputs)); }

	void set(int task, int field, const In& v) { check(osc_set_field(_h, task, field, v.ptr, v.layout, v.mem)); }
	void get(int task, int field, const Out& o) { check(osc_get_field(_h, task, field, o.ptr, o.layout, o.mem)); }
	int fieldSize(int task, int field) const { int n = osc_field_size(_h, task, field); check(n); return n; }

private:
	osc_batch_t* _h = nullptr;
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
