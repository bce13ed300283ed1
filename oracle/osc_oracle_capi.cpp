// C entry points of the CPU oracle (ctypes-friendly).  TEST INFRASTRUCTURE ONLY -- see the header of osc_oracle.hpp.
// Same field ids / parameter structs as include/osc_b200.h so that tests can compare field by field, but a different,
// much simpler calling convention: all per-robot buffers are host double arrays in AOS layout [B][k].
// "One thread per robot instance": the batch is statically partitioned over `nthreads` std::threads and every robot
// owns its model copy and task objects.
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
int orc_reinitialize_tasks(orc_batch* b, int task) {
	for (Robot* r : b->robots)
		for (int t = 0; t < (int)r->tasks.size(); ++t)
			if (task < 0 || t == task) r->tasks[t]->reInitializeTask();
	return 0;
}
int orc_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
