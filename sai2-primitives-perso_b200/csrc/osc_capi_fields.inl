// osc_capi_fields.inl -- per-robot field access

namespace osc {
// fields the fused tick does not write (include/osc_b200.h, OSC_TICK_OUTPUTS_LAZY): model and task-model matrices
static bool lazily_materialised(int task, int field) {
	if (task < 0) return field == OSC_F_M || field == OSC_F_MINV || field == OSC_F_N_TOTAL;
	switch (field) {
		case OSC_F_N_TASK: case OSC_F_N_PREC: case OSC_F_N_TASK_AND_PREC: case OSC_F_JACOBIAN: case OSC_F_PROJECTED_JACOBIAN:
		case OSC_F_LAMBDA_WORLD: case OSC_F_LAMBDA_MODIFIED_WORLD: case OSC_F_SINGULAR_VALUES:
		case OSC_F_JT_LAMBDA_WORLD: case OSC_F_JT_LAMBDA_MODIFIED_WORLD: case OSC_F_JT_RANGE_RANK: case OSC_F_NS_RANK: case OSC_F_ALPHA: case OSC_F_JT_POSITION_ERROR_NORM:
			return true;
	}
	return false;
}
}  // namespace osc

extern "C" {

int osc_field_size(const osc_batch_t* batch, int task, int field) {
	if (is_multi(batch)) batch = (const osc_batch_t*)shard0(batch);
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	Loc l;
	if (!b->ops->locate(*b, task, field, l)) return invalid("unknown field for this task");
	return l.width;
}

int osc_set_field(osc_batch_t* batch, int task, int field, const void* data, int layout, int mem) {
	OSC_RANGE("osc_set_field");
	if (is_multi(batch)) {
		Batch* p = (Batch*)batch;
		if (!data) return invalid("null argument");
		if (int rc = multi_host_only(mem)) return rc;
		Loc l;
		if (!shard0(p)->ops->locate(*shard0(p), task, field, l)) return invalid("unknown field for this task");
		const size_t esz = l.is_int ? 4 : p->esz;
		return multi_fan(p, [=](osc_batch_t* s, int k) {
			const int rc = osc_set_field(s, task, field, multi_slice(p, k, data, l.width, layout, esz), layout, mem);
			((Batch*)s)->io_pitch = 0;
			return rc;
		});
	}
	Batch* b = (Batch*)batch;
	if (!b || !data) return invalid("null argument");
	Loc l;
	if (!b->ops->locate(*b, task, field, l)) return invalid("unknown field for this task");
	if (!l.writable || l.special) return invalid("field is read-only");
	if (int rc = use_device(*b)) return rc;
	if (l.arena == 3) b->tasks[l.owner].otg_goals_dirty = true;  // the next k_otg_advance runs the goal filter
	return import_rows(*b, row_ptr(*b, l), l.width, l.is_int, data, layout, mem);
}

int osc_get_field(osc_batch_t* batch, int task, int field, void* out, int layout, int mem) {
	OSC_RANGE("osc_get_field");
	if (is_multi(batch)) {
		Batch* p = (Batch*)batch;
		if (!out) return invalid("null argument");
		if (int rc = multi_host_only(mem)) return rc;
		Loc l;
		if (!shard0(p)->ops->locate(*shard0(p), task, field, l)) return invalid("unknown field for this task");
		const size_t esz = l.is_int ? 4 : p->esz;
		return multi_fan(p, [=](osc_batch_t* s, int k) {
			const int rc = osc_get_field(s, task, field, multi_slice(p, k, out, l.width, layout, esz), layout, mem);
			((Batch*)s)->io_pitch = 0;
			return rc;
		});
	}
	Batch* b = (Batch*)batch;
	if (!b || !out) return invalid("null argument");
	Loc l;
	if (!b->ops->locate(*b, task, field, l)) return invalid("unknown field for this task");
	if (int rc = use_device(*b)) return rc;
	if (b->stale && lazily_materialised(task, field))
		if (int rc = b->ops->materialise(*b)) return rc;
	const void* rows = nullptr;
	if (l.special) {
		void* tmp = b->ensure_stage2((size_t)l.width * b->stride * 8);
		if (!tmp) return invalid("out of device memory (staging)");
		if (int rc = b->ops->pack_special(*b, l, tmp)) return rc;
		rows = tmp;
	} else {
		rows = row_ptr(*b, l);
	}
	return export_rows(*b, rows, l.width, l.is_int, out, layout, mem);
}

int osc_mft_update_sensed_force_and_moment(osc_batch_t* batch, int task, const void* force, const void* moment, int layout, int mem) {
	if (is_multi(batch)) {
		Batch* p = (Batch*)batch;
		if (!force || !moment) return invalid("null argument");
		if (int rc = multi_host_only(mem)) return rc;
		return multi_fan(p, [=](osc_batch_t* s, int k) {
			const void* f = multi_slice(p, k, force, 3, layout, p->esz);
			const void* m = multi_slice(p, k, moment, 3, layout, p->esz);
			const int rc = osc_mft_update_sensed_force_and_moment(s, task, f, m, layout, mem);
			((Batch*)s)->io_pitch = 0;
			return rc;
		});
	}
	Batch* b = (Batch*)batch;
	if (!b || !force || !moment) return invalid("null argument");
	if (task < 0 || task >= (int)b->tasks.size() || b->tasks[task].type != OSC_TASK_MOTION_FORCE) return invalid("not a MotionForceTask");
	if (int rc = use_device(*b)) return rc;
	char* tmp = (char*)b->ensure_stage2((size_t)6 * b->stride * 8);
	if (!tmp) return invalid("out of device memory (staging)");
	if (int rc = import_rows(*b, tmp, 3, false, force, layout, mem)) return rc;
	if (int rc = import_rows(*b, tmp + (size_t)3 * b->stride * b->esz, 3, false, moment, layout, mem)) return rc;
	// the sensor-frame transform reads the kinematics of the last updateModel; after a fused tick (stale rows) k_mft_sensed
	// evaluates them from the current q itself, so a force-control loop around osc_tick stays at one extra small launch
	return b->ops->update_sensed(*b, task, tmp);
}


int osc_task_goal_reached(osc_batch_t* batch, int task, int what, double tolerance, int32_t* flags_out, int mem, int64_t* n_reached) {
	if (is_multi(batch)) {
		Batch* p = (Batch*)batch;
		if (flags_out)
			if (int rc = multi_host_only(mem)) return rc;
		std::vector<int64_t> cnt(p->shards.size(), 0);
		const int rc = multi_fan(p, [=, &cnt](osc_batch_t* s, int k) {
			return osc_task_goal_reached(s, task, what, tolerance, flags_out ? flags_out + p->shard_first[k] : nullptr, mem, &cnt[k]);
		});
		if (rc < 0) return rc;
		if (n_reached) { *n_reached = 0; for (int64_t c : cnt) *n_reached += c; }
		return OSC_OK;
	}
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (task < 0 || task >= (int)b->tasks.size()) return invalid("task index out of range");
	const bool mft = b->tasks[task].type == OSC_TASK_MOTION_FORCE;
	if (what != 0 && !(what == 1 && mft)) return invalid("goal reached: what = 0 (position) or 1 (orientation, MotionForceTask only)");
	const int field = mft ? (what == 0 ? OSC_F_POSITION_ERROR_NORM : OSC_F_ORIENTATION_ERROR_NORM) : OSC_F_JT_POSITION_ERROR_NORM;
	Loc l;
	if (!b->ops->locate(*b, task, field, l)) return invalid("unknown field for this task");
	if (int rc = use_device(*b)) return rc;
	if (b->stale && lazily_materialised(task, field))
		if (int rc = b->ops->materialise(*b)) return rc;
	void* norms = b->ensure_stage2((size_t)b->stride * 8);
	char* st = (char*)b->ensure_stage((size_t)b->stride * 4 + 16);
	if (!norms || !st) return invalid("out of device memory (staging)");
	if (int rc = b->ops->pack_special(*b, l, norms)) return rc;
	int32_t* flags = (int32_t*)st;
	int* count = (int*)(st + (size_t)b->stride * 4);
	OSC_CUDA(cudaMemsetAsync(count, 0, sizeof(int), b->stream));
	const unsigned g = (unsigned)((b->B + 255) / 256);
	if (b->esz == 8) k_below_tolerance<double><<<g, 256, 0, b->stream>>>((const double*)norms, tolerance, flags, count, b->B);
	else k_below_tolerance<float><<<g, 256, 0, b->stream>>>((const float*)norms, tolerance, flags, count, b->B);
	++b->launches;
	OSC_CUDA(cudaGetLastError());
	if (flags_out)
		OSC_CUDA(cudaMemcpyAsync(flags_out, flags, (size_t)b->B * 4, mem == OSC_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, b->stream));
	int c = 0;
	OSC_CUDA(cudaMemcpyAsync(&c, count, sizeof(int), cudaMemcpyDeviceToHost, b->stream));
	OSC_CUDA(cudaStreamSynchronize(b->stream));
	if (n_reached) *n_reached = c;
	return OSC_OK;
}

}  // extern "C"
