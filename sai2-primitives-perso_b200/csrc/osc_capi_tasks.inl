// osc_capi_tasks.inl -- hierarchy construction and parameters (cold path, host side)

namespace osc {

// projector onto the range of the 3 x k direction matrix, with Sai2Model::matrixRangeBasis' rank rule (tolerance 1e-3):
// the B B^T blocks of MotionForceTask.cpp:82-87.  Returns the rank (0 when the list is empty / numerically zero).
static int range_projector_3(const double dirs[3][3], int k, double P[3][3], double Bout[3][3]) {
	memset(P, 0, sizeof(double) * 9);
	memset(Bout, 0, sizeof(double) * 9);
	if (k == 0) return 0;
	double C[3][3] = {{0}}, V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
	double fro = 0;
	for (int c = 0; c < k; ++c)
		for (int i = 0; i < 3; ++i) {
			fro += dirs[c][i] * dirs[c][i];
			for (int j = 0; j < 3; ++j) C[i][j] += dirs[c][i] * dirs[c][j];
		}
	if (sqrt(fro) < 1e-3) return 0;
	for (int sweep = 0; sweep < 50; ++sweep) {  // cyclic Jacobi on the symmetric 3x3
		double off = fabs(C[0][1]) + fabs(C[0][2]) + fabs(C[1][2]);
		if (off < 1e-300) break;
		for (int p = 0; p < 2; ++p)
			for (int q = p + 1; q < 3; ++q) {
				if (C[p][q] == 0) continue;
				double th = (C[q][q] - C[p][p]) / (2 * C[p][q]);
				double t = (th >= 0 ? 1 : -1) / (fabs(th) + sqrt(th * th + 1));
				double cs = 1 / sqrt(t * t + 1), sn = t * cs;
				for (int i = 0; i < 3; ++i) { double a = C[i][p], bq = C[i][q]; C[i][p] = cs * a - sn * bq; C[i][q] = sn * a + cs * bq; }
				for (int i = 0; i < 3; ++i) { double a = C[p][i], bq = C[q][i]; C[p][i] = cs * a - sn * bq; C[q][i] = sn * a + cs * bq; }
				for (int i = 0; i < 3; ++i) { double a = V[i][p], bq = V[i][q]; V[i][p] = cs * a - sn * bq; V[i][q] = sn * a + cs * bq; }
			}
	}
	int order[3] = {0, 1, 2};
	std::sort(order, order + 3, [&](int a, int bb) { return C[a][a] > C[bb][bb]; });
	double s[3];
	for (int i = 0; i < 3; ++i) s[i] = sqrt(std::max(C[order[i]][order[i]], 0.0));
	if (s[0] < 1e-3) return 0;
	int max_range = std::min(3, k), rank = max_range;
	for (int i = max_range - 1; i > 0; --i) {
		if (s[i] / s[0] < 1e-3) --rank;
		else break;
	}
	if (rank == 3) { P[0][0] = P[1][1] = P[2][2] = 1; Bout[0][0] = Bout[1][1] = Bout[2][2] = 1; return 3; }
	for (int c = 0; c < rank; ++c)
		for (int i = 0; i < 3; ++i) {
			Bout[i][c] = V[i][order[c]];
			for (int j = 0; j < 3; ++j) P[i][j] += V[i][order[c]] * V[j][order[c]];
		}
	return rank;
}

static int add_task_common(Batch& b, TaskH& t, double dt) {
	if (b.closed) {
		return invalid("task cannot be added to the controller because it is in the nullspace of a full joint task");  // RobotController.cpp:36-41
	}
	if ((int)b.tasks.size() >= OSC_MAX_TASKS) return invalid("too many tasks (OSC_MAX_TASKS)");
	if (!b.tasks.empty() && dt != b.dt) return invalid("All tasks must have the same loop timestep in RobotController");  // .cpp:23-27
	if (int rc = use_device(b)) return rc;
	if (int rc = b.ops->materialise(b)) return rc;
	b.ops->task_rows(t.type, t.p_rows, t.s_rows, t.i_rows);
	if (int rc = alloc_rows(b, &t.P, t.p_rows, b.esz)) return rc;
	if (int rc = alloc_rows(b, &t.S, t.s_rows, b.esz)) return rc;
	if (int rc = alloc_rows(b, (void**)&t.I, t.i_rows, 4)) return rc;
	b.dt = dt;
	b.tasks.push_back(t);
	const int idx = (int)b.tasks.size() - 1;
	if (int rc = b.ops->task_init(b, idx)) return rc;
	return idx;
}

}  // namespace osc

extern "C" {

void osc_mft_default_params(osc_mft_params* p) {  // MotionForceTask.h:40-75, SingularityHandler.cpp:10-20, MotionForceTask.cpp:197
	memset(p, 0, sizeof *p);
	for (int i = 0; i < 3; ++i) {
		p->kp_pos[i] = 100.0; p->kv_pos[i] = 20.0; p->ki_pos[i] = 0.0;
		p->kp_ori[i] = 200.0; p->kv_ori[i] = 28.3; p->ki_ori[i] = 0.0;
	}
	p->kp_force = 0.7; p->kv_force = 10.0; p->ki_force = 1.3;
	p->kp_moment = 0.7; p->kv_moment = 10.0; p->ki_moment = 1.3;
	p->kff_force = 0.95; p->kff_moment = 0.95;
	p->max_force_control_feedback_output = 20.0; p->max_moment_control_feedback_output = 10.0;
	p->linear_saturation_velocity = 0.3; p->angular_saturation_velocity = M_PI / 3;
	p->dynamic_decoupling_type = OSC_BOUNDED_INERTIA_ESTIMATES; p->bie_threshold = 0.1;
	p->singularity_handling_enabled = 1;
	p->s_min = 6e-3; p->s_max = 6e-2;
	p->kp_type_1 = 50; p->kv_type_1 = 14; p->kv_type_2 = 5;
	p->s_abs_tol = 1e-3; p->type_1_tol = 0.5; p->type_2_torque_ratio = 1e-2;
	p->type_2_angle_threshold = 5 * M_PI / 180; p->perturb_step_size = 5; p->buffer_size = 200;
	p->sensor_rot[0] = p->sensor_rot[4] = p->sensor_rot[8] = 1;
}
void osc_joint_task_default_params(osc_joint_task_params* p, int k) {  // JointTask.h:31-45
	memset(p, 0, sizeof *p);
	for (int i = 0; i < k && i < OSC_MAX_DOF; ++i) { p->kp[i] = 50.0; p->kv[i] = 14.0; p->ki[i] = 0.0; p->saturation_velocity[i] = M_PI / 3; }
	p->dynamic_decoupling_type = OSC_BOUNDED_INERTIA_ESTIMATES; p->bie_threshold = 0.1;
}

int osc_batch_add_motion_force_task(osc_batch_t* batch, const osc_mft_desc* d) {
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_batch_add_motion_force_task(s, d); });
	Batch* b = (Batch*)batch;
	if (!b || !d) return invalid("null argument");
	if (d->body < 0 || d->body >= b->n) return invalid("control link body index out of range");
	TaskH t;
	t.type = OSC_TASK_MOTION_FORCE;
	t.md = *d;
	osc_mft_default_params(&t.mp);
	memset(t.P6, 0, sizeof t.P6);
	if (!d->partial) {  // MotionForceTask.cpp:28
		for (int i = 0; i < 6; ++i) t.P6[i][i] = 1;
		for (int i = 0; i < 3; ++i) t.Bt[i][i] = t.Br[i][i] = 1;
		t.mt = t.mr = 3;
		t.dof = 6;
	} else {  // MotionForceTask.cpp:47-87, 146-160
		if (d->n_dirs_translation < 0 || d->n_dirs_translation > 3 || d->n_dirs_rotation < 0 || d->n_dirs_rotation > 3)
			return invalid("at most 3 controlled directions per block");
		if (d->n_dirs_translation == 0 && d->n_dirs_rotation == 0)
			return invalid("controlled_directions_translation and controlled_directions_rotation cannot both be empty in MotionForceTask::MotionForceTask\n");
		double Pt[3][3], Pr[3][3];
		const int rt = range_projector_3(d->dirs_translation, d->n_dirs_translation, Pt, t.Bt);
		const int rr = range_projector_3(d->dirs_rotation, d->n_dirs_rotation, Pr, t.Br);
		t.mt = rt; t.mr = rr;
		for (int i = 0; i < 3; ++i)
			for (int j = 0; j < 3; ++j) { t.P6[i][j] = Pt[i][j]; t.P6[3 + i][3 + j] = Pr[i][j]; }
		t.dof = rt + rr;
		if (t.dof == 0)
			return invalid("controlled_directions_translation and controlled_directions_rotation cannot both be empty in MotionForceTask::MotionForceTask\n");
	}
	if (t.dof > std::min(6, b->n)) return invalid("task rank exceeds the robot dof");
	// compliant frame into the canonical body frame
	double QT[3][3], R[3][3];
	h_tr3(b->Qc[d->body], QT);
	h_mv3(QT, d->frame_xyz, t.frame_p);
	for (int k = 0; k < 9; ++k) R[k / 3][k % 3] = d->frame_rot[k];
	h_mm3(QT, R, t.frame_R);
	return add_task_common(*b, t, d->loop_timestep);
}

int osc_batch_add_joint_task(osc_batch_t* batch, const osc_joint_task_desc* d) {
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_batch_add_joint_task(s, d); });
	Batch* b = (Batch*)batch;
	if (!b || !d) return invalid("null argument");
	TaskH t;
	t.type = OSC_TASK_JOINT;
	t.jd = *d;
	const int n = b->n;
	if (d->full) {
		t.dof = n;
	} else {
		if (d->k < 1 || d->k > n) return invalid("joint selection matrix size not consistent with robot dof in JointTask constructor\n");
		// full row rank check (FullPivLU::rank, JointTask.cpp:33-39) by Gaussian elimination with full pivoting
		std::vector<double> A(d->selection, d->selection + d->k * n);
		double maxpiv = 0;
		int rank = 0;
		std::vector<int> used_r(d->k, 0), used_c(n, 0);
		for (int step = 0; step < d->k; ++step) {
			int pr = -1, pc = -1;
			double best = 0;
			for (int i = 0; i < d->k; ++i)
				if (!used_r[i])
					for (int j = 0; j < n; ++j)
						if (!used_c[j] && fabs(A[i * n + j]) > best) { best = fabs(A[i * n + j]); pr = i; pc = j; }
			if (step == 0) maxpiv = best;
			if (pr < 0 || best <= maxpiv * 1e-12 * std::max(d->k, n)) break;
			++rank;
			used_r[pr] = used_c[pc] = 1;
			for (int i = 0; i < d->k; ++i)
				if (!used_r[i]) {
					const double f = A[i * n + pc] / A[pr * n + pc];
					for (int j = 0; j < n; ++j) A[i * n + j] -= f * A[pr * n + j];
				}
		}
		if (rank != d->k) return invalid("joint selection matrix is not full rank in JointTask constructor\n");
		t.dof = d->k;
	}
	osc_joint_task_default_params(&t.jp, t.dof);
	const int idx = add_task_common(*b, t, d->loop_timestep);
	if (idx >= 0 && t.dof == n) b->closed = true;  // isFullJointTask(), RobotController.cpp:44-49
	return idx;
}

int osc_mft_get_params(const osc_batch_t* batch, int task, osc_mft_params* out) {
	if (is_multi(batch)) batch = (const osc_batch_t*)shard0(batch);
	const Batch* b = (const Batch*)batch;
	if (!b || !out || task < 0 || task >= (int)b->tasks.size() || b->tasks[task].type != OSC_TASK_MOTION_FORCE) return invalid("not a MotionForceTask");
	*out = b->tasks[task].mp;
	return OSC_OK;
}

int osc_mft_set_params(osc_batch_t* batch, int task, const osc_mft_params* np) {
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_mft_set_params(s, task, np); });
	Batch* b = (Batch*)batch;
	if (!b || !np || task < 0 || task >= (int)b->tasks.size() || b->tasks[task].type != OSC_TASK_MOTION_FORCE) return invalid("not a MotionForceTask");
	TaskH& t = b->tasks[task];
	osc_mft_params p = *np;
	const osc_mft_params& o = t.mp;
	for (int i = 0; i < 3; ++i) {
		if (p.kp_pos[i] < 0 || p.kv_pos[i] < 0 || p.ki_pos[i] < 0) return invalid("all gains should be positive or zero in MotionForceTask::setPosControlGains\n");
		if (p.kp_ori[i] < 0 || p.kv_ori[i] < 0 || p.ki_ori[i] < 0) return invalid("all gains should be positive or zero in MotionForceTask::setOriControlGains\n");
	}
	if (p.use_velocity_saturation && (p.linear_saturation_velocity <= 0 || p.angular_saturation_velocity <= 0))
		return invalid("Velocity saturation values should be strictly positive or zero in MotionForceTask::enableVelocitySaturation\n");
	if (p.force_space_dimension < 0 || p.force_space_dimension > 3) return invalid("Force space dimension should be between 0 and 3 in MotionForceTask::parametrizeForceMotionSpaces\n");
	if (p.moment_space_dimension < 0 || p.moment_space_dimension > 3) return invalid("Moment space dimension should be between 0 and 3 in MotionForceTask::parametrizeMomentRotMotionSpaces\n");
	if (p.buffer_size < 1 || p.buffer_size > OSC_MAX_SINGULARITY_BUFFER) return invalid("singularity buffer size must be between 1 and OSC_MAX_SINGULARITY_BUFFER");
	auto norm3 = [](const double* a) { return sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); };
	auto changed = [](const double* a, const double* c) {  // !isApprox (Eigen, precision 1e-12)
		double d = 0, na = 0, nc = 0;
		for (int i = 0; i < 3; ++i) { d += (a[i] - c[i]) * (a[i] - c[i]); na += a[i] * a[i]; nc += c[i] * c[i]; }
		return !(d <= 1e-24 * std::min(na, nc));
	};
	int flags = 0, ret = 0;
	bool reset_lin = p.force_space_dimension != o.force_space_dimension;  // MotionForceTask.cpp:838
	if (p.force_space_dimension == 1 || p.force_space_dimension == 2) {
		const double nn = norm3(p.force_or_motion_axis);
		if (nn < 1e-2) return invalid("Force or motion axis should be a non singular vector in MotionForceTask::parametrizeForceMotionSpaces\n");
		for (double& x : p.force_or_motion_axis) x /= nn;
		reset_lin = reset_lin || changed(p.force_or_motion_axis, o.force_or_motion_axis);
	} else memcpy(p.force_or_motion_axis, o.force_or_motion_axis, sizeof p.force_or_motion_axis);
	bool reset_ang = p.moment_space_dimension != o.moment_space_dimension;  // .cpp:868
	if (p.moment_space_dimension == 1 || p.moment_space_dimension == 2) {
		const double nn = norm3(p.moment_or_rotmotion_axis);
		if (nn < 1e-2) return invalid("Moment or rot motion axis should be a non singular vector in MotionForceTask::parametrizeMomentRotMotionSpaces\n");
		for (double& x : p.moment_or_rotmotion_axis) x /= nn;
		reset_ang = reset_ang || changed(p.moment_or_rotmotion_axis, o.moment_or_rotmotion_axis);
	} else memcpy(p.moment_or_rotmotion_axis, o.moment_or_rotmotion_axis, sizeof p.moment_or_rotmotion_axis);
	if (reset_lin) { flags |= 1; ret |= 1; }
	if (reset_ang) { flags |= 2; ret |= 2; }
	if ((p.closed_loop_force_control != 0) != (o.closed_loop_force_control != 0)) flags |= 8;    // .cpp:973-979
	if ((p.closed_loop_moment_control != 0) != (o.closed_loop_moment_control != 0)) flags |= 16;  // .cpp:980-986
	if (!p.passivity_enabled && o.passivity_enabled) { flags |= 4; ret |= 4; }
	if (int rc = use_device(*b)) return rc;
	if (p.passivity_enabled && !t.popc_ring) {
		OSC_CUDA(cudaMalloc(&t.popc_ring, (size_t)OSC_POPC_RING_CAPACITY * b->stride * b->esz));
	}
	t.mp = p;
	if (flags) {
		if (int rc = b->ops->mft_reset(*b, task, flags)) return rc;
		if (t.otgO && (flags & 3))  // _otg->reInitializeLinear / reInitializeAngular, MotionForceTask.cpp:854,886
			if (int rc = b->ops->otg_reinit(*b, task, ((flags & 1) ? 2 : 0) | ((flags & 2) ? 4 : 0) | (t.otg_on ? 64 : 0))) return rc;
	}
	return ret;
}

int osc_task_enable_internal_otg_acceleration_limited(osc_batch_t* batch, int task, const double* limits, int count) {
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_task_enable_internal_otg_acceleration_limited(s, task, limits, count); });
	Batch* b = (Batch*)batch;
	if (!b || !limits || task < 0 || task >= (int)b->tasks.size()) return invalid("task index out of range");
	TaskH& t = b->tasks[task];
	double mv[8] = {0}, ma[8] = {0};
	if (t.type == OSC_TASK_MOTION_FORCE) {
		if (count != 2) return invalid("MotionForceTask OTG limits: {max_linear_velocity, max_angular_velocity, max_linear_acceleration, max_angular_acceleration}");
		for (int i = 0; i < 3; ++i) { mv[i] = limits[0]; mv[3 + i] = limits[1]; ma[i] = limits[2]; ma[3 + i] = limits[3]; }
		for (int i = 0; i < 6; ++i)  // OTG_6dof_cartesian.cpp:86-129
			if (!(mv[i] > 0) || !(ma[i] > 0)) return invalid("max velocity or acceleration set to 0 or negative value in some directions in OTG_6dof_cartesian\n");
	} else {
		if (count != 1 && count != t.dof)  // JointTask.cpp:366-372
			return invalid("max velocity or max acceleration vector size not consistent with task dof in JointTask::enableInternalOtgAccelerationLimited\n");
		for (int i = 0; i < t.dof; ++i) { mv[i] = limits[count == 1 ? 0 : i]; ma[i] = limits[count + (count == 1 ? 0 : i)]; }
		for (int i = 0; i < t.dof; ++i)  // OTG_joints.cpp:49-53,63-67
			if (!(mv[i] > 0) || !(ma[i] > 0)) return invalid("max velocity or acceleration cannot be 0 or negative in any directions in OTG_joints\n");
	}
	if (int rc = use_device(*b)) return rc;
	if (int rc = b->ops->materialise(*b)) return rc;
	bool fresh = false;
	if (!t.otgO) {
		b->ops->otg_rows(*b, task, t.otg_o_rows, t.otg_i_rows, t.otg_g_rows);
		if (int rc = alloc_rows(*b, (void**)&t.otgO, t.otg_o_rows, 8)) return rc;
		if (int rc = alloc_rows(*b, (void**)&t.otgI, t.otg_i_rows, 4)) return rc;
		if (int rc = alloc_rows(*b, &t.otgG, t.otg_g_rows, b->esz)) return rc;
		// ping-pong counters + compacted list of the robots on the general path: allocated here so that no tick ever allocates
		OSC_CUDA(cudaMalloc((void**)&t.otg_slow, (size_t)(b->stride + 2) * sizeof(int)));
		OSC_CUDA(cudaMemsetAsync(t.otg_slow, 0, 2 * sizeof(int), b->stream));
		t.otg_parity = 0;
		fresh = true;
	}
	const bool limits_changed = memcmp(t.otg_max_v, mv, sizeof mv) != 0 || memcmp(t.otg_max_a, ma, sizeof ma) != 0;
	memcpy(t.otg_max_v, mv, sizeof mv);
	memcpy(t.otg_max_a, ma, sizeof ma);
	if (t.otg_on && limits_changed)
		if (int rc = b->ops->otg_reinit(*b, task, 128)) return rc;
	if (!t.otg_on) {  // reInitialize(_current_position) when it was off (JointTask.cpp:373-375); the user's goals move to G
		if (int rc = b->ops->otg_reinit(*b, task, 1 | 8 | (fresh ? 32 : 0))) return rc;
	}
	t.otg_on = true;
	t.otg_goals_dirty = true;
	return OSC_OK;
}
int osc_task_enable_internal_otg_jerk_limited(osc_batch_t* batch, int task, const double* limits, int count) {
	(void)batch; (void)task; (void)limits; (void)count;
	set_error("jerk-limited internal OTG (third-order Ruckig profiles) is not built; use the acceleration-limited mode");
	return OSC_ERR_UNSUPPORTED;
}
int osc_task_disable_internal_otg(osc_batch_t* batch, int task) {
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_task_disable_internal_otg(s, task); });
	Batch* b = (Batch*)batch;
	if (!b || task < 0 || task >= (int)b->tasks.size()) return invalid("task index out of range");
	TaskH& t = b->tasks[task];
	if (!t.otg_on) return OSC_OK;
	if (int rc = use_device(*b)) return rc;
	if (int rc = b->ops->otg_reinit(*b, task, 16)) return rc;  // the goal rows are the user's goals again
	t.otg_on = false;
	return OSC_OK;
}
int osc_task_internal_otg_enabled(const osc_batch_t* batch, int task) {
	if (is_multi(batch)) batch = (const osc_batch_t*)shard0(batch);
	const Batch* b = (const Batch*)batch;
	if (!b || task < 0 || task >= (int)b->tasks.size()) return 0;
	return b->tasks[task].otg_on ? 1 : 0;
}

int osc_joint_task_get_params(const osc_batch_t* batch, int task, osc_joint_task_params* out) {
	if (is_multi(batch)) batch = (const osc_batch_t*)shard0(batch);
	const Batch* b = (const Batch*)batch;
	if (!b || !out || task < 0 || task >= (int)b->tasks.size() || b->tasks[task].type != OSC_TASK_JOINT) return invalid("not a JointTask");
	*out = b->tasks[task].jp;
	return OSC_OK;
}
int osc_joint_task_set_params(osc_batch_t* batch, int task, const osc_joint_task_params* p) {
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_joint_task_set_params(s, task, p); });
	Batch* b = (Batch*)batch;
	if (!b || !p || task < 0 || task >= (int)b->tasks.size() || b->tasks[task].type != OSC_TASK_JOINT) return invalid("not a JointTask");
	TaskH& t = b->tasks[task];
	for (int i = 0; i < t.dof; ++i) {
		if (p->kp[i] < 0 || p->kv[i] < 0 || p->ki[i] < 0) return invalid("gains must be positive or zero in JointTask::setGains\n");
		if (p->use_velocity_saturation && p->saturation_velocity[i] <= 0) return invalid("saturation velocity must be positive in JointTask::enableVelocitySaturation\n");
	}
	if (p->dynamic_decoupling_type < 0 || p->dynamic_decoupling_type > 2) return invalid("Dynamic decoupling type not recognized in JointTask::updateTaskModel\n");
	t.jp = *p;
	return OSC_OK;
}


int osc_joint_task_set_gains_unsafe(osc_batch_t* batch, int task, const double* kp, const double* kv, const double* ki, int count) {  // JointTask.cpp:136-156
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_joint_task_set_gains_unsafe(s, task, kp, kv, ki, count); });
	Batch* b = (Batch*)batch;
	if (!b || !kp || !kv || !ki || task < 0 || task >= (int)b->tasks.size() || b->tasks[task].type != OSC_TASK_JOINT) return invalid("not a JointTask");
	TaskH& t = b->tasks[task];
	if (count != 1 && count != t.dof) return invalid("size of gain vectors inconsistent with number of task dofs in JointTask::setGains\n");
	for (int i = 0; i < t.dof; ++i) {
		const int s = count == 1 ? 0 : i;
		t.jp.kp[i] = kp[s]; t.jp.kv[i] = kv[s]; t.jp.ki[i] = ki[s];
	}
	return OSC_OK;
}
int osc_mft_selection_projectors(const osc_batch_t* batch, int task, double pos_projector[9], double ori_projector[9]) {  // MotionForceTask.h:653-659
	if (is_multi(batch)) batch = (const osc_batch_t*)shard0(batch);
	const Batch* b = (const Batch*)batch;
	if (!b || task < 0 || task >= (int)b->tasks.size() || b->tasks[task].type != OSC_TASK_MOTION_FORCE) return invalid("not a MotionForceTask");
	const TaskH& t = b->tasks[task];
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j) {
			if (pos_projector) pos_projector[3 * i + j] = t.P6[i][j];
			if (ori_projector) ori_projector[3 * i + j] = t.P6[3 + i][3 + j];
		}
	return OSC_OK;
}

}  // extern "C"
