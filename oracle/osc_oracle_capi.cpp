// C entry points of the CPU oracle (ctypes-friendly).  TEST INFRASTRUCTURE ONLY -- see the header of osc_oracle.hpp.
// Same field ids / parameter structs as include/osc_b200.h so that tests can compare field by field, but a different,
// much simpler calling convention: all per-robot buffers are host double arrays in AOS layout [B][k].
// "One thread per robot instance": the batch is statically partitioned over `nthreads` std::threads and every robot
// owns its model copy and task objects.
#include <chrono>
#include <string>
#include <thread>

#include "osc_oracle.hpp"

using namespace orc;

struct orc_batch {
	osc_chain_desc chain;
	int64_t B;
	int n;
	std::vector<Robot*> robots;
	std::vector<int> task_types;
	~orc_batch() { for (Robot* r : robots) delete r; }
};

static thread_local std::string g_err;

template <class F>
static void parallel_for(orc_batch* b, int nthreads, F f) {
	if (nthreads <= 1 || b->B < 2 * nthreads) {
		for (int64_t i = 0; i < b->B; ++i) f(*b->robots[i]);
		return;
	}
	std::vector<std::thread> th;
	for (int t = 0; t < nthreads; ++t) {
		int64_t lo = b->B * t / nthreads, hi = b->B * (t + 1) / nthreads;
		th.emplace_back([=, &f]() { for (int64_t i = lo; i < hi; ++i) f(*b->robots[i]); });
	}
	for (auto& t : th) t.join();
}

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

orc_batch* orc_create(const osc_chain_desc* chain, int64_t B, const double* q0) {
	orc_batch* b = new orc_batch;
	b->chain = *chain; b->B = B; b->n = chain->n;
	b->robots.reserve(B);
	for (int64_t i = 0; i < B; ++i) {
		Robot* r = new Robot(*chain);
		if (q0) for (int j = 0; j < chain->n; ++j) r->model.q[j] = q0[j];
		r->model.updateModel();
		b->robots.push_back(r);
	}
	return b;
}
void orc_destroy(orc_batch* b) { delete b; }

int orc_add_motion_force_task(orc_batch* b, const osc_mft_desc* d) {
	try {
		for (Robot* r : b->robots) r->tasks.push_back(new MotionForceTask(&r->model, *d));
	} catch (std::exception& e) { g_err = e.what(); return OSC_ERR_INVALID_ARGUMENT; }
	b->task_types.push_back(OSC_TASK_MOTION_FORCE);
	return (int)b->task_types.size() - 1;
}
int orc_add_joint_task(orc_batch* b, const osc_joint_task_desc* d) {
	try {
		for (Robot* r : b->robots) r->tasks.push_back(new JointTask(&r->model, *d));
	} catch (std::exception& e) { g_err = e.what(); return OSC_ERR_INVALID_ARGUMENT; }
	b->task_types.push_back(OSC_TASK_JOINT);
	return (int)b->task_types.size() - 1;
}
void orc_mft_default_params(osc_mft_params* p) { MotionForceTask::default_params(p); }
void orc_joint_task_default_params(osc_joint_task_params* p, int k) { JointTask::default_params(p, k); }
int orc_mft_get_params(orc_batch* b, int task, osc_mft_params* p) {
	*p = static_cast<MotionForceTask*>(b->robots[0]->tasks[task])->prm;
	return 0;
}
int orc_mft_set_params(orc_batch* b, int task, const osc_mft_params* p) {
	int resets = 0;
	for (Robot* r : b->robots) resets = static_cast<MotionForceTask*>(r->tasks[task])->setParams(*p);
	return resets;
}
int orc_joint_task_get_params(orc_batch* b, int task, osc_joint_task_params* p) {
	*p = static_cast<JointTask*>(b->robots[0]->tasks[task])->prm;
	return 0;
}
int orc_joint_task_set_params(orc_batch* b, int task, const osc_joint_task_params* p) {
	for (Robot* r : b->robots) static_cast<JointTask*>(r->tasks[task])->prm = *p;
	return 0;
}
int orc_enable_gravity_compensation(orc_batch* b, int e) {
	for (Robot* r : b->robots) r->gravity_comp = e != 0;
	return 0;
}
int orc_task_dof(orc_batch* b, int task) {
	Task* t = b->robots[0]->tasks[task];
	if (t->type == OSC_TASK_JOINT) return static_cast<JointTask*>(t)->k;
	auto* m = static_cast<MotionForceTask*>(t);
	return m->pos_range + m->ori_range;
}

static void put3(double* o, const V3& v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; }
static void putM3(double* o, const M3& m) { for (int i = 0; i < 9; ++i) o[i] = m.m[i / 3][i % 3]; }
static void putMat(double* o, const Mat& m, int rows, int cols) {
	for (int i = 0; i < rows; ++i)
		for (int j = 0; j < cols; ++j) o[i * cols + j] = (i < m.r && j < m.c) ? m(i, j) : 0.0;
}
static V3 get3(const double* p) { return V3(p); }

// returns the per-robot width, writes into o when o != nullptr (get) or reads from in when in != nullptr (set)
static int field_io(orc_batch* b, Robot& r, int task, int field, double* o, const double* in) {
	const int n = b->n;
	if (task < 0) {
		switch (field) {
			case OSC_F_Q: if (o) for (int i = 0; i < n; ++i) o[i] = r.model.q[i]; if (in) for (int i = 0; i < n; ++i) r.model.q[i] = in[i]; return n;
			case OSC_F_DQ: if (o) for (int i = 0; i < n; ++i) o[i] = r.model.dq[i]; if (in) for (int i = 0; i < n; ++i) r.model.dq[i] = in[i]; return n;
			case OSC_F_TAU: if (o) for (int i = 0; i < n; ++i) o[i] = r.tau.a[i]; if (in) for (int i = 0; i < n; ++i) r.tau.a[i] = in[i]; return n;
			case OSC_F_M: if (o) putMat(o, r.model.Mm, n, n); return n * n;
			case OSC_F_MINV: if (o) putMat(o, r.model.Minv, n, n); return n * n;
			case OSC_F_N_TOTAL: if (o) putMat(o, r.N_total, n, n); return n * n;
			case OSC_F_GRAVITY: if (o) for (int i = 0; i < n; ++i) o[i] = r.gravity[i]; return n;
			case OSC_F_NONLINEAR_EFFECTS: if (o) for (int i = 0; i < n; ++i) o[i] = r.nle[i]; return n;
			case OSC_F_STATUS: if (o) o[0] = 0; return 1;
		}
		return -1;
	}
	if (task >= (int)r.tasks.size()) return -1;
	Task* t = r.tasks[task];
	switch (field) {
		case OSC_F_TASK_TORQUES: if (o) for (int i = 0; i < n; ++i) o[i] = t->tau.a[i]; return n;
		case OSC_F_N_TASK: if (o) putMat(o, t->N, n, n); return n * n;
		case OSC_F_N_PREC: if (o) putMat(o, t->N_prec, n, n); return n * n;
		case OSC_F_N_TASK_AND_PREC: if (o) putMat(o, t->getTaskAndPreviousNullspace(), n, n); return n * n;
	}
	if (t->type == OSC_TASK_MOTION_FORCE) {
		auto* m = static_cast<MotionForceTask*>(t);
		auto* h = m->handler;
#define V3FIELD(id, member) case id: if (o) put3(o, m->member); if (in) m->member = get3(in); return 3;
#define M3FIELD(id, member) case id: if (o) putM3(o, m->member); if (in) m->member = M3(in); return 9;
		switch (field) {
			V3FIELD(OSC_F_GOAL_POSITION, goal_pos)
			M3FIELD(OSC_F_GOAL_ORIENTATION, goal_ori)
			V3FIELD(OSC_F_GOAL_LINEAR_VELOCITY, goal_linvel)
			V3FIELD(OSC_F_GOAL_ANGULAR_VELOCITY, goal_angvel)
			V3FIELD(OSC_F_GOAL_LINEAR_ACCELERATION, goal_linacc)
			V3FIELD(OSC_F_GOAL_ANGULAR_ACCELERATION, goal_angacc)
			V3FIELD(OSC_F_GOAL_FORCE, goal_force)
			V3FIELD(OSC_F_GOAL_MOMENT, goal_moment)
			V3FIELD(OSC_F_SENSED_FORCE_WORLD, sensed_force)
			V3FIELD(OSC_F_SENSED_MOMENT_WORLD, sensed_moment)
			V3FIELD(OSC_F_CURRENT_POSITION, x)
			M3FIELD(OSC_F_CURRENT_ORIENTATION, R)
			V3FIELD(OSC_F_CURRENT_LINEAR_VELOCITY, v)
			V3FIELD(OSC_F_CURRENT_ANGULAR_VELOCITY, w)
			V3FIELD(OSC_F_ORIENTATION_ERROR, ori_err)
			V3FIELD(OSC_F_INTEGRATED_POSITION_ERROR, integ_pos)
			V3FIELD(OSC_F_INTEGRATED_ORIENTATION_ERROR, integ_ori)
			V3FIELD(OSC_F_INTEGRATED_FORCE_ERROR, integ_force)
			V3FIELD(OSC_F_INTEGRATED_MOMENT_ERROR, integ_moment)
			V3FIELD(OSC_F_DESIRED_POSITION, des_pos)
			M3FIELD(OSC_F_DESIRED_ORIENTATION, des_ori)
			V3FIELD(OSC_F_DESIRED_LINEAR_VELOCITY, des_linvel)
			V3FIELD(OSC_F_DESIRED_ANGULAR_VELOCITY, des_angvel)
			V3FIELD(OSC_F_DESIRED_LINEAR_ACCELERATION, des_linacc)
			V3FIELD(OSC_F_DESIRED_ANGULAR_ACCELERATION, des_angacc)
			case OSC_F_OTG_FLAGS: if (!m->otg_) return -1; if (o) o[0] = m->otg_->goal_reached ? 1 : 0; return 1;
			case OSC_F_OTG_RESULT: if (!m->otg_) return -1; if (o) o[0] = m->otg_->result_value; return 1;
			case OSC_F_UNIT_MASS_FORCE: if (o) for (int i = 0; i < 6; ++i) o[i] = m->unit_mass_force[i]; return 6;
			case OSC_F_JACOBIAN: if (o) putMat(o, m->J, 6, n); return 6 * n;
			case OSC_F_PROJECTED_JACOBIAN: if (o) putMat(o, m->Jp, 6, n); return 6 * n;
			case OSC_F_LAMBDA_WORLD:
			case OSC_F_LAMBDA_MODIFIED_WORLD:
				if (o) {
					if (h->U_ns.norm() == 0 || h->U_ns.r != 6) putMat(o, Mat::Zero(6, 6), 6, 6);
					else putMat(o, h->U_ns * (field == OSC_F_LAMBDA_WORLD ? h->L_ns : h->L_ns_mod) * h->U_ns.T(), 6, 6);
				}
				return 36;
			case OSC_F_ALPHA: if (o) o[0] = h->alpha; return 1;
			case OSC_F_SINGULAR_VALUES: if (o) for (int i = 0; i < 6; ++i) o[i] = i < h->svd.k ? h->svd.s[i] : 0.0; return 6;
			case OSC_F_NS_RANK: if (o) o[0] = h->U_ns.norm() == 0 ? 0 : h->U_ns.c; return 1;
			case OSC_F_SINGULARITY_TYPES: if (o) for (int i = 0; i < 6; ++i) o[i] = i < (int)h->singularity_types.size() ? h->singularity_types[i] : 0; return 6;
			case OSC_F_SINGULARITY_COUNTERS: if (o) { o[0] = h->c1; o[1] = h->c2; } return 2;
			case OSC_F_POPC_RC: if (o) o[0] = m->popc.Rc; return 1;
			case OSC_F_POPC_OBSERVER: if (o) o[0] = m->popc.PO; return 1;
			case OSC_F_Q_PRIOR: if (o) for (int i = 0; i < n; ++i) o[i] = h->q_prior[i]; if (in) for (int i = 0; i < n; ++i) h->q_prior[i] = in[i]; return n;
		}
		return -1;
	}
	auto* j = static_cast<JointTask*>(t);
	const int k = j->k;
#define KFIELD(id, member) case id: if (o) for (int i = 0; i < k; ++i) o[i] = j->member[i]; if (in) for (int i = 0; i < k; ++i) j->member[i] = in[i]; return k;
	switch (field) {
		KFIELD(OSC_F_JT_GOAL_POSITION, goal_pos)
		KFIELD(OSC_F_JT_GOAL_VELOCITY, goal_vel)
		KFIELD(OSC_F_JT_GOAL_ACCELERATION, goal_acc)
		KFIELD(OSC_F_JT_CURRENT_POSITION, cur_pos)
		KFIELD(OSC_F_JT_CURRENT_VELOCITY, cur_vel)
		KFIELD(OSC_F_JT_INTEGRATED_ERROR, integ)
		KFIELD(OSC_F_JT_DESIRED_VELOCITY, des_vel)
		KFIELD(OSC_F_JT_DESIRED_POSITION, des_pos)
		KFIELD(OSC_F_JT_DESIRED_ACCELERATION, des_acc)
		case OSC_F_OTG_FLAGS: if (!j->otg_) return -1; if (o) o[0] = j->otg_->goal_reached ? 1 : 0; return 1;
		case OSC_F_OTG_RESULT: if (!j->otg_) return -1; if (o) o[0] = j->otg_->result_value; return 1;
		case OSC_F_JT_LAMBDA_WORLD:
		case OSC_F_JT_LAMBDA_MODIFIED_WORLD:
			if (o) {
				if (j->range.norm() == 0) putMat(o, Mat::Zero(k, k), k, k);
				else putMat(o, j->range * (field == OSC_F_JT_LAMBDA_WORLD ? j->Lam : j->Lam_mod) * j->range.T(), k, k);
			}
			return k * k;
		case OSC_F_JT_RANGE_RANK: if (o) o[0] = j->range.norm() == 0 ? 0 : j->range.c; return 1;
	}
	return -1;
}

int orc_field_size(orc_batch* b, int task, int field) { return field_io(b, *b->robots[0], task, field, nullptr, nullptr); }
int orc_get_field(orc_batch* b, int task, int field, double* out) {
	int w = orc_field_size(b, task, field);
	if (w < 0) return OSC_ERR_INVALID_ARGUMENT;
	for (int64_t i = 0; i < b->B; ++i) field_io(b, *b->robots[i], task, field, out + i * w, nullptr);
	return 0;
}
// uniform != 0: data is one [k] record applied to every robot
int orc_set_field(orc_batch* b, int task, int field, const double* data, int uniform) {
	int w = orc_field_size(b, task, field);
	if (w < 0) return OSC_ERR_INVALID_ARGUMENT;
	for (int64_t i = 0; i < b->B; ++i) field_io(b, *b->robots[i], task, field, nullptr, uniform ? data : data + i * w);
	return 0;
}
int orc_mft_update_sensed_force_and_moment(orc_batch* b, int task, const double* f, const double* m, int uniform) {
	for (int64_t i = 0; i < b->B; ++i)
		static_cast<MotionForceTask*>(b->robots[i]->tasks[task])->updateSensedForceAndMoment(V3(uniform ? f : f + 3 * i), V3(uniform ? m : m + 3 * i));
	return 0;
}

int orc_update_model(orc_batch* b, int nthreads) { parallel_for(b, nthreads, [](Robot& r) { r.model.updateModel(); }); return 0; }
int orc_update_kinematics(orc_batch* b, int nthreads) { parallel_for(b, nthreads, [](Robot& r) { r.model.updateKinematics(); }); return 0; }
int orc_task_update_model(orc_batch* b, int task, int n_prec_task, int nthreads) {
	parallel_for(b, nthreads, [=](Robot& r) {
		Mat N_prec = n_prec_task < 0 ? Mat::Identity(r.model.dof()) : r.tasks[n_prec_task]->getTaskAndPreviousNullspace();
		r.tasks[task]->updateTaskModel(N_prec);
	});
	return 0;
}
int orc_update_task_models(orc_batch* b, int nthreads) { parallel_for(b, nthreads, [](Robot& r) { r.updateControllerTaskModels(); }); return 0; }
int orc_task_compute_torques(orc_batch* b, int task, int with_tau_prec, int nthreads) {
	parallel_for(b, nthreads, [=](Robot& r) {
		Task* t = r.tasks[task];
		t->tau = with_tau_prec ? t->computeTorques(r.tau) : t->computeTorques();
	});
	return 0;
}
int orc_compute_torques(orc_batch* b, int stacked, int nthreads) { parallel_for(b, nthreads, [=](Robot& r) { r.computeControlTorques(stacked != 0); }); return 0; }
int orc_tick(orc_batch* b, int stacked, int nthreads) {
	parallel_for(b, nthreads, [=](Robot& r) {
		r.model.updateModel();
		r.updateControllerTaskModels();
		r.computeControlTorques(stacked != 0);
	});
	return 0;
}
int orc_rollout_step(orc_batch* b, double dt, int substeps, double damping, int nthreads) {
	parallel_for(b, nthreads, [=](Robot& r) { r.model.rolloutStep(&r.tau.a[0], dt, substeps, damping, r.nle); });
	return 0;
}
int orc_nonlinear_effects(orc_batch* b, double* out) {
	for (int64_t i = 0; i < b->B; ++i) b->robots[i]->model.nonlinearEffects(out + i * b->n);
	return 0;
}
int orc_reinitialize_tasks(orc_batch* b, int task) {
	for (Robot* r : b->robots)
		for (int t = 0; t < (int)r->tasks.size(); ++t)
			if (task < 0 || t == task) r->tasks[t]->reInitializeTask();
	return 0;
}
// ---- internal OTG (otg_oracle.hpp) ----
// 1 = run the wrappers on the REAL Ruckig (oracle/_ref/libruckig_ref.so) for every OTG created afterwards; 0 = restatement
int orc_otg_use_reference(const char* path) {
	if (!path) { otg::use_reference_backend() = false; return 0; }
	if (!otg::RefLib::get().load(path)) { g_err = std::string("cannot load ") + path; return OSC_ERR_INVALID_ARGUMENT; }
	otg::use_reference_backend() = true;
	return 0;
}
// same argument convention as osc_task_enable_internal_otg_acceleration_limited (include/osc_b200.h)
int orc_task_enable_internal_otg_acceleration_limited(orc_batch* b, int task, const double* limits, int count) {
	try {
		for (Robot* r : b->robots) {
			Task* t = r->tasks[task];
			if (t->type == OSC_TASK_MOTION_FORCE) {
				if (count != 2) throw std::invalid_argument("MotionForceTask OTG limits: 4 numbers");
				static_cast<MotionForceTask*>(t)->enableInternalOtgAccelerationLimited(limits[0], limits[2], limits[1], limits[3]);
			} else {
				auto* j = static_cast<JointTask*>(t);
				if (count != 1 && count != j->k) throw std::invalid_argument("max velocity or max acceleration vector size not consistent with task dof in JointTask::enableInternalOtgAccelerationLimited\n");
				double mv[8], ma[8];
				for (int i = 0; i < j->k; ++i) { mv[i] = limits[count == 1 ? 0 : i]; ma[i] = limits[count + (count == 1 ? 0 : i)]; }
				j->enableInternalOtgAccelerationLimited(mv, ma);
			}
		}
	} catch (std::exception& e) { g_err = e.what(); return OSC_ERR_INVALID_ARGUMENT; }
	return 0;
}
int orc_task_disable_internal_otg(orc_batch* b, int task) {
	for (Robot* r : b->robots) {
		Task* t = r->tasks[task];
		if (t->type == OSC_TASK_MOTION_FORCE) static_cast<MotionForceTask*>(t)->disableInternalOtg();
		else static_cast<JointTask*>(t)->disableInternalOtg();
	}
	return 0;
}

// stand-alone objects for the pinning tests (tests/test_otg_oracle.py)
void* orc_rk_create(int n, double dt) { return new otg::Engine(n, dt); }
void orc_rk_destroy(void* h) { delete (otg::Engine*)h; }
// one Ruckig::update; io = [cur_p cur_v cur_a tgt_p tgt_v max_v max_a] (7 x n), out = [new_p new_v new_a] (3 x n, left
// untouched on an error), info = {time, duration, new_calculation}
int orc_rk_update(void* h, const double* io, double* out, double* info) {
	otg::Engine* e = (otg::Engine*)h;
	const int n = e->n;
	otg::Input in;
	for (int i = 0; i < n; ++i) {
		in.cur_p[i] = io[i]; in.cur_v[i] = io[n + i]; in.cur_a[i] = io[2 * n + i]; in.tgt_p[i] = io[3 * n + i]; in.tgt_v[i] = io[4 * n + i];
		in.max_v[i] = io[5 * n + i]; in.max_a[i] = io[6 * n + i];
	}
	const int r = e->update(in, out, out + n, out + 2 * n);
	info[0] = e->last_time; info[1] = e->last_duration; info[2] = e->last_new_calculation ? 1 : 0;
	return r;
}
void* orc_otgj_create(const double* p0, int n, double dt, const double* max_v, const double* max_a) {
	auto* o = new otg::OtgJoints(p0, n, dt);
	o->setLimits(max_v, max_a);
	o->disableJerkLimits();
	return o;
}
void orc_otgj_destroy(void* h) { delete (otg::OtgJoints*)h; }
// setGoalPositionAndVelocity + update; out = [p v a] (3 x n); returns result_value, flags[0] = goal reached
int orc_otgj_step(void* h, const double* goal_p, const double* goal_v, double* out, int* flags) {
	auto* o = (otg::OtgJoints*)h;
	o->setGoalPositionAndVelocity(goal_p, goal_v);
	o->update();
	for (int i = 0; i < o->dim; ++i) { out[i] = o->out_p[i]; out[o->dim + i] = o->out_v[i]; out[2 * o->dim + i] = o->out_a[i]; }
	flags[0] = o->goal_reached ? 1 : 0;
	return o->result_value;
}
void* orc_otgc_create(const double* p0, const double* R0, double dt, double lin_v, double lin_a, double ang_v, double ang_a) {
	auto* o = new otg::OtgCartesian(p0, R0, dt);
	o->setLimits(lin_v, lin_a, ang_v, ang_a);
	return o;
}
void orc_otgc_destroy(void* h) { delete (otg::OtgCartesian*)h; }
// out = [pos 3, R 9, linvel 3, angvel 3, linacc 3, angacc 3] = 24; returns result_value, or -1000 for an invalid goal rotation
int orc_otgc_step(void* h, const double* gp, const double* gR, const double* gv, const double* gw, double* out, int* flags) {
	auto* o = (otg::OtgCartesian*)h;
	o->setGoalPositionAndLinearVelocity(gp, gv);
	if (!o->setGoalOrientationAndAngularVelocity(gR, gw)) return -1000;
	o->update();
	for (int i = 0; i < 3; ++i) { out[i] = o->out_p[i]; out[12 + i] = o->out_v[i]; out[18 + i] = o->out_a[i]; }
	o->getNextOrientation(out + 3);
	o->getNextAngularVelocity(out + 15);
	o->getNextAngularAcceleration(out + 21);
	flags[0] = o->goal_reached ? 1 : 0;
	return o->result_value;
}

// FLOP of one control tick of robot 0 (updateModel + updateControllerTaskModels + computeControlTorques), counted by the
// primitives of osc_oracle.hpp; -1 unless this library was built with -DORACLE_COUNT_FLOPS (make count).
long long orc_count_tick_flops(orc_batch* b, int stacked) {
#ifdef ORACLE_COUNT_FLOPS
	Robot& r = *b->robots[0];
	const long long before = flop_counter();
	r.model.updateModel();
	r.updateControllerTaskModels();
	r.computeControlTorques(stacked != 0);
	return flop_counter() - before;
#else
	(void)b; (void)stacked;
	return -1;
#endif
}

// BASELINE config 1 on the CPU: ONE robot (robot 0 of the batch, hierarchy [MotionForceTask, JointTask] already added), `ticks`
// cycles of the control loop of examples/04-task_and_redundancy/04-task_and_redundancy.cpp:133-206 -- setQ/setDq (kinematic
// playback q0 + amp sin(w_j t): no simulator here), updateModel, the two updateTaskModel calls, the goal generator of
// :155-185 (orientation +-pi/8 about Y at 0.2 Hz, circle r = 0.05 m at 0.33 Hz in y-z), the two computeTorques calls, the
// joint goal step at cycle 5000 (:196-201) -- with every cycle timed by steady_clock into out_us[ticks].
int orc_config1_run(orc_batch* b, int ticks, double* out_us) {
	Robot& r = *b->robots[0];
	if (r.tasks.size() != 2 || r.tasks[0]->type != OSC_TASK_MOTION_FORCE) { g_err = "config 1 needs [MotionForceTask, JointTask]"; return -1; }
	auto* m = static_cast<MotionForceTask*>(r.tasks[0]);
	auto* j = static_cast<JointTask*>(r.tasks[1]);
	const int n = b->n;
	double q0[8];
	for (int i = 0; i < n; ++i) q0[i] = r.model.q[i];
	r.model.updateModel();
	m->reInitializeTask(); j->reInitializeTask();
	const V3 x0 = m->x;
	const M3 R0 = m->R;
	const double wj[8] = {0.11, 0.13, 0.17, 0.19, 0.23, 0.29, 0.31, 0.37}, amp[8] = {0.3, 0.15, 0.3, 0.09, 0.3, 0.15, 0.3, 0.1};
	const double two_pi = 6.283185307179586;
	for (int tick = 0; tick < ticks; ++tick) {
		const double t = tick * 0.001;
		const auto t_start = std::chrono::steady_clock::now();
		for (int i = 0; i < n; ++i) {
			r.model.q[i] = q0[i] + amp[i] * std::sin(two_pi * wj[i] * t);
			r.model.dq[i] = amp[i] * two_pi * wj[i] * std::cos(two_pi * wj[i] * t);
		}
		r.model.updateModel();
		Mat N_prec = Mat::Identity(n);
		m->updateTaskModel(N_prec);
		N_prec = m->getTaskAndPreviousNullspace();
		j->updateTaskModel(N_prec);
		const double w_o = two_pi * 0.2, a_o = 3.141592653589793 / 8;
		const double ang = a_o * std::sin(w_o * t), angv = a_o * w_o * std::cos(w_o * t), anga = -a_o * w_o * w_o * std::sin(w_o * t);
		M3 Ry = M3::Identity();
		Ry.m[0][0] = std::cos(ang); Ry.m[0][2] = std::sin(ang); Ry.m[2][0] = -std::sin(ang); Ry.m[2][2] = std::cos(ang);
		m->goal_ori = Ry.T() * R0;
		m->goal_angvel = V3(0, angv, 0);
		m->goal_angacc = V3(0, anga, 0);
		const double rad = 0.05, w_c = two_pi * 0.33;
		m->goal_pos = x0 + rad * V3(0, std::sin(w_c * t), 1 - std::cos(w_c * t));
		m->goal_linvel = (rad * w_c) * V3(0, std::cos(w_c * t), std::sin(w_c * t));
		m->goal_linacc = (rad * w_c * w_c) * V3(0, -std::sin(w_c * t), std::cos(w_c * t));
		Mat tau = m->computeTorques();
		Mat tj = j->computeTorques();
		if (tick >= 5000) tau = tau + tj;
		if (tick == 5000) {
			j->reInitializeTask();
			for (int i = 0; i < n; ++i) j->goal_pos[i] = q0[i];
			j->goal_pos[0] += 1.5;
		}
		r.tau = tau;
		out_us[tick] = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_start).count();
	}
	return 0;
}

int orc_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
