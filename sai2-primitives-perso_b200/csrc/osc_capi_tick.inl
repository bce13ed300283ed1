// osc_capi_tick.inl -- the per-tick entry points and the controller-state snapshot

namespace osc {
static int check_task(Batch* b, int task) {
	if (!b) return invalid("null batch");
	if (task < 0 || task >= (int)b->tasks.size()) return invalid("task index out of range");
	return OSC_OK;
}
static int emit_tau(Batch& b, void* tau_out, int layout, int mem) {
	if (!tau_out) return OSC_OK;
	Loc l;
	b.ops->locate(b, -1, OSC_F_TAU, l);
	return export_rows(b, row_ptr(b, l), l.width, false, tau_out, layout, mem);
}
}  // namespace osc

extern "C" {

int osc_update_model(osc_batch_t* batch) {
	OSC_RANGE("osc_update_model");
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_update_model(s); });
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (int rc = use_device(*b)) return rc;
	if (b->stale) return b->ops->materialise(*b);  // model + task models at the current q (see osc_batch_set_tick_outputs)
	return b->ops->model(*b, 1);
}
int osc_defer_model_update(osc_batch_t* batch) {
	if (is_multi(batch)) {  // host-side flag only: no need for the shard threads
		for (Batch* s : ((Batch*)batch)->shards) s->stale = true;
		return OSC_OK;
	}
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	b->stale = true;
	return OSC_OK;
}
int osc_update_kinematics(osc_batch_t* batch) {
	OSC_RANGE("osc_update_kinematics");
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_update_kinematics(s); });
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (int rc = use_device(*b)) return rc;
	if (b->stale) return b->ops->materialise(*b);
	return b->ops->model(*b, 0);
}
int osc_task_update_model(osc_batch_t* batch, int task, int n_prec_task) {
	OSC_RANGE("osc_task_update_model");
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_task_update_model(s, task, n_prec_task); });
	Batch* b = (Batch*)batch;
	if (int rc = check_task(b, task)) return rc;
	if (n_prec_task >= (int)b->tasks.size() || n_prec_task == task)
		return invalid("N_prec matrix size not consistent with robot dof in updateTaskModel\n");
	if (int rc = use_device(*b)) return rc;
	if (int rc = b->ops->materialise(*b)) return rc;
	return b->ops->task_update(*b, task, n_prec_task < 0 ? -1 : n_prec_task);
}
int osc_update_task_models(osc_batch_t* batch) {
	OSC_RANGE("osc_update_task_models");
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_update_task_models(s); });  // RobotController.cpp:53-60
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (b->tasks.empty()) return invalid("RobotController must have at least one task");
	if (int rc = use_device(*b)) return rc;
	if (int rc = b->ops->materialise(*b)) return rc;
	for (int t = 0; t < (int)b->tasks.size(); ++t)
		if (int rc = b->ops->task_update(*b, t, t - 1)) return rc;
	return OSC_OK;
}
int osc_task_compute_torques(osc_batch_t* batch, int task, int with_tau_prec) {
	OSC_RANGE("osc_task_compute_torques");
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_task_compute_torques(s, task, with_tau_prec); });
	Batch* b = (Batch*)batch;
	if (int rc = check_task(b, task)) return rc;
	if (int rc = use_device(*b)) return rc;
	if (int rc = b->ops->materialise(*b)) return rc;
	if (b->tasks[task].otg_on)
		if (int rc = b->ops->otg_tick(*b, task)) return rc;
	return b->ops->task_compute(*b, task, with_tau_prec, 0);
}
int osc_compute_torques(osc_batch_t* batch, int stacked, void* tau_out, int layout, int mem) {
	OSC_RANGE("osc_compute_torques");
	if (is_multi(batch)) {
		Batch* p = (Batch*)batch;
		if (tau_out)
			if (int rc = multi_host_only(mem)) return rc;
		return multi_fan(p, [=](osc_batch_t* s, int k) {
			const int rc = osc_compute_torques(s, stacked, multi_slice(p, k, tau_out, p->n, layout, p->esz), layout, mem);
			((Batch*)s)->io_pitch = 0;
			return rc;
		});
	}  // RobotController.cpp:62-74
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (b->tasks.empty()) return invalid("RobotController must have at least one task");
	if (int rc = use_device(*b)) return rc;
	if (int rc = b->ops->materialise(*b)) return rc;
	for (int t = 0; t < (int)b->tasks.size(); ++t) {
		if (b->tasks[t].otg_on)
			if (int rc = b->ops->otg_tick(*b, t)) return rc;
		if (int rc = b->ops->task_compute(*b, t, (stacked && t > 0) ? 1 : 0, t == 0 ? 1 : 2)) return rc;
	}
	if (b->gravity)
		if (int rc = b->ops->add_gravity(*b)) return rc;
	return emit_tau(*b, tau_out, layout, mem);
}
int osc_tick_is_fused(const osc_batch_t* batch) {
	if (is_multi(batch)) batch = (const osc_batch_t*)shard0(batch);
	Batch* b = (Batch*)batch;
	return b && b->ops->fused_supported(*b) ? 1 : 0;
}
int osc_tick(osc_batch_t* batch, int stacked, void* tau_out, int layout, int mem) {
	OSC_RANGE("osc_tick");
	if (is_multi(batch)) {
		Batch* p = (Batch*)batch;
		if (tau_out)
			if (int rc = multi_host_only(mem)) return rc;
		return multi_fan(p, [=](osc_batch_t* s, int k) {
			const int rc = osc_tick(s, stacked, multi_slice(p, k, tau_out, p->n, layout, p->esz), layout, mem);
			((Batch*)s)->io_pitch = 0;
			return rc;
		});
	}
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (b->tasks.empty()) return invalid("RobotController must have at least one task");
	if (int rc = use_device(*b)) return rc;
	if (b->ops->fused_supported(*b)) {
		if (int rc = b->ops->otg_tick(*b, -1)) return rc;  // the desired state of this tick, before the control law reads it
		if (int rc = b->ops->tick_fused(*b, stacked)) return rc;
		return emit_tau(*b, tau_out, layout, mem);
	}
	b->stale = false;
	if (int rc = b->ops->model(*b, 1)) return rc;
	if (int rc = osc_update_task_models(batch)) return rc;
	return osc_compute_torques(batch, stacked, tau_out, layout, mem);
}
int osc_step_host(osc_batch_t* batch, const osc_field_in* inputs, int32_t n_inputs, int32_t stacked, void* tau_out, int32_t tau_layout) {
	OSC_RANGE("osc_step_host");
	Batch* p = (Batch*)batch;
	if (!p) return invalid("null batch");
	if (n_inputs < 0 || (n_inputs > 0 && !inputs)) return invalid("null argument in osc_step_host");
	for (int i = 0; i < n_inputs; ++i)
		if (!inputs[i].data) return invalid("null input buffer in osc_step_host");
	if (!is_multi(p)) {
		for (int i = 0; i < n_inputs; ++i)
			if (int rc = osc_set_field(batch, inputs[i].task, inputs[i].field, inputs[i].data, inputs[i].layout, OSC_MEM_HOST)) return rc;
		return osc_tick(batch, stacked, tau_out, tau_layout, OSC_MEM_HOST);
	}
	// shard-major issue from this thread, nothing waits until every shard is queued: copy engine order = shard 0's inputs, shard 1's
	// inputs, ...; shard k's kernels start when ITS inputs have landed, its torques travel back while shard k + 1 is still uploading
	int rc = OSC_OK;
	const int ns = (int)p->shards.size();
	for (int k = 0; k < ns && rc == OSC_OK; ++k) {
		Batch* s = p->shards[k];
		s->defer_host_sync = true;
		for (int i = 0; i < n_inputs && rc == OSC_OK; ++i) {
			Loc l;
			if (!s->ops->locate(*s, inputs[i].task, inputs[i].field, l)) { rc = invalid("unknown field for this task"); break; }
			const size_t esz = l.is_int ? 4 : p->esz;
			rc = osc_set_field((osc_batch_t*)s, inputs[i].task, inputs[i].field, multi_slice(p, k, inputs[i].data, l.width, inputs[i].layout, esz), inputs[i].layout, OSC_MEM_HOST);
			s->io_pitch = 0;
		}
		if (rc == OSC_OK) {
			rc = osc_tick((osc_batch_t*)s, stacked, multi_slice(p, k, tau_out, p->n, tau_layout, p->esz), tau_layout, OSC_MEM_HOST);
			s->io_pitch = 0;
		}
		s->defer_host_sync = false;
	}
	const std::string err = rc == OSC_OK ? std::string() : last_error();
	for (int k = 0; k < ns; ++k) {
		const int r2 = osc_synchronize((osc_batch_t*)p->shards[k]);
		if (rc == OSC_OK && r2 != OSC_OK) rc = r2;
	}
	if (!err.empty()) set_error(err);
	return rc;
}
int osc_rollout_step(osc_batch_t* batch, double dt, int substeps, double joint_damping) {
	OSC_RANGE("osc_rollout_step");
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_rollout_step(s, dt, substeps, joint_damping); });
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (!(dt > 0) || substeps < 1 || substeps > 1000 || joint_damping < 0) return invalid("rollout step: dt > 0, 1 <= substeps <= 1000, damping >= 0");
	if (int rc = use_device(*b)) return rc;
	return b->ops->rollout_step(*b, dt, substeps, joint_damping);
}
int osc_rollout(osc_batch_t* batch, int stacked, int ticks, double sim_dt, int substeps, double joint_damping, void* tau_out, int layout, int mem) {
	OSC_RANGE("osc_rollout");
	if (is_multi(batch)) {
		Batch* p = (Batch*)batch;
		if (tau_out)
			if (int rc = multi_host_only(mem)) return rc;
		return multi_fan(p, [=](osc_batch_t* s, int k) {
			const int rc = osc_rollout(s, stacked, ticks, sim_dt, substeps, joint_damping, multi_slice(p, k, tau_out, p->n, layout, p->esz), layout, mem);
			((Batch*)s)->io_pitch = 0;
			return rc;
		});
	}
	if (ticks < 1) return invalid("rollout: ticks >= 1");
	for (int t = 0; t < ticks; ++t) {
		if (int rc = osc_tick(batch, stacked, t == ticks - 1 ? tau_out : nullptr, layout, mem)) return rc;
		if (int rc = osc_rollout_step(batch, sim_dt, substeps, joint_damping)) return rc;
	}
	return OSC_OK;
}
int osc_batch_set_tick_outputs(osc_batch_t* batch, int outputs) {
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_batch_set_tick_outputs(s, outputs); });
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (outputs != OSC_TICK_OUTPUTS_ALL && outputs != OSC_TICK_OUTPUTS_LAZY && outputs != OSC_TICK_OUTPUTS_TAU)
		return invalid("unknown tick output mode");
	b->tick_outputs = outputs;
	return OSC_OK;
}
int64_t osc_tick_fallback_count(osc_batch_t* batch) {
	if (is_multi(batch)) {
		int64_t n = 0;
		for (Batch* s : ((Batch*)batch)->shards) {
			const int64_t c = osc_tick_fallback_count((osc_batch_t*)s);
			if (c < 0) return c;
			n += c;
		}
		return n;
	}
	Batch* b = (Batch*)batch;
	if (!b || !b->fb) return 0;
	if (use_device(*b)) return -1;
	int n = 0;
	if (cudaMemcpyAsync(&n, b->fb + b->fb_last, sizeof(int), cudaMemcpyDeviceToHost, b->stream) != cudaSuccess) return -1;
	if (cudaStreamSynchronize(b->stream) != cudaSuccess) return -1;
	return n;
}

int osc_reinitialize_tasks(osc_batch_t* batch, int task) {
	if (is_multi(batch)) return multi_fan((Batch*)batch, [=](osc_batch_t* s, int) { return osc_reinitialize_tasks(s, task); });
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (task >= (int)b->tasks.size()) return invalid("task index out of range");
	if (int rc = use_device(*b)) return rc;
	if (int rc = b->ops->materialise(*b)) return rc;
	for (int t = 0; t < (int)b->tasks.size(); ++t)
		if (task < 0 || t == task) {
			if (int rc = b->ops->task_reinit(*b, t)) return rc;
			if (b->tasks[t].otgO)  // _otg->reInitialize(...), JointTask.cpp:106 / MotionForceTask.cpp:244
				if (int rc = b->ops->otg_reinit(*b, t, 1 | (b->tasks[t].otg_on ? 64 : 0))) return rc;
		}
	return OSC_OK;
}

// snapshot = persistent reals + persistent ints of the robot and of every task (+ the passivity ring when allocated)
static int64_t snapshot_walk(Batch& b, char* buf, int mem, int dir) {
	int64_t off = 0;
	auto seg = [&](void* dev, size_t bytes) -> int {
		if (buf && bytes) {
			cudaMemcpyKind kind = mem == OSC_MEM_HOST ? (dir ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost) : cudaMemcpyDeviceToDevice;
			cudaError_t e = dir ? cudaMemcpyAsync(dev, buf + off, bytes, kind, b.stream) : cudaMemcpyAsync(buf + off, dev, bytes, kind, b.stream);
			if (e != cudaSuccess) return cuda_fail(e, "snapshot copy");
		}
		off += (int64_t)bytes;
		return OSC_OK;
	};
	if (seg(b.P, (size_t)b.p_rows * b.stride * b.esz)) return -1;
	if (seg(b.I, (size_t)b.i_rows * b.stride * 4)) return -1;
	for (TaskH& t : b.tasks) {
		if (seg(t.P, (size_t)t.p_rows * b.stride * b.esz)) return -1;
		if (seg(t.I, (size_t)t.i_rows * b.stride * 4)) return -1;
		if (t.popc_ring && seg(t.popc_ring, (size_t)OSC_POPC_RING_CAPACITY * b.stride * b.esz)) return -1;
		if (t.otgO) {
			if (seg(t.otgO, (size_t)t.otg_o_rows * b.stride * 8)) return -1;
			if (seg(t.otgI, (size_t)t.otg_i_rows * b.stride * 4)) return -1;
			if (seg(t.otgG, (size_t)t.otg_g_rows * b.stride * b.esz)) return -1;
		}
	}
	return off;
}
int64_t osc_snapshot_bytes(const osc_batch_t* batch) {
	if (is_multi(batch)) {
		int64_t n = 0;
		for (Batch* s : ((Batch*)batch)->shards) n += osc_snapshot_bytes((osc_batch_t*)s);
		return n;
	}
	Batch* b = (Batch*)batch;
	if (!b) return 0;
	return snapshot_walk(*b, nullptr, OSC_MEM_DEVICE, 0);
}
int osc_snapshot(osc_batch_t* batch, void* dst, int mem) {
	if (is_multi(batch)) {  // the shards' snapshots back to back, in shard order
		Batch* p = (Batch*)batch;
		if (!dst) return invalid("null argument");
		if (int rc = multi_host_only(mem)) return rc;
		std::vector<int64_t> off(p->shards.size(), 0);
		for (size_t k = 1; k < off.size(); ++k) off[k] = off[k - 1] + osc_snapshot_bytes((osc_batch_t*)p->shards[k - 1]);
		return multi_fan(p, [=, &off](osc_batch_t* s, int k) { return osc_snapshot(s, (char*)dst + off[k], mem); });
	}
	Batch* b = (Batch*)batch;
	if (!b || !dst) return invalid("null argument");
	if (int rc = use_device(*b)) return rc;
	if (snapshot_walk(*b, (char*)dst, mem, 0) < 0) return OSC_ERR_CUDA;
	if (mem == OSC_MEM_HOST) OSC_CUDA(cudaStreamSynchronize(b->stream));
	return OSC_OK;
}
int osc_restore(osc_batch_t* batch, const void* src, int mem) {
	if (is_multi(batch)) {  // the shards' snapshots back to back, in shard order
		Batch* p = (Batch*)batch;
		if (!src) return invalid("null argument");
		if (int rc = multi_host_only(mem)) return rc;
		std::vector<int64_t> off(p->shards.size(), 0);
		for (size_t k = 1; k < off.size(); ++k) off[k] = off[k - 1] + osc_snapshot_bytes((osc_batch_t*)p->shards[k - 1]);
		return multi_fan(p, [=, &off](osc_batch_t* s, int k) { return osc_restore(s, (const char*)src + off[k], mem); });
	}
	Batch* b = (Batch*)batch;
	if (!b || !src) return invalid("null argument");
	if (int rc = use_device(*b)) return rc;
	if (snapshot_walk(*b, (char*)src, mem, 1) < 0) return OSC_ERR_CUDA;
	for (TaskH& t : b->tasks) t.otg_goals_dirty = true;
	if (mem == OSC_MEM_HOST) OSC_CUDA(cudaStreamSynchronize(b->stream));
	return OSC_OK;
}

}  // extern "C"

// ---- FP64 peak microbenchmark ----
namespace osc {
__global__ void __launch_bounds__(256) k_dfma_peak(double* out, int iters, double seed) {
	double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
	const double m = 1.0000001, c = 1e-9;
	for (int i = 0; i < iters; ++i) {
		a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
		a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
	}
	out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
}  // namespace osc

extern "C" int osc_measure_fp64_peak(int device, double* tflops, double* elapsed_ms) {
	if (!tflops) return invalid("null argument");
	if (osc_device_count() == 0) { set_error("no CUDA device available"); return OSC_ERR_NO_DEVICE; }
	OSC_CUDA(cudaSetDevice(device));
	cudaDeviceProp prop;
	OSC_CUDA(cudaGetDeviceProperties(&prop, device));
	const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 16;
	double* out = nullptr;
	OSC_CUDA(cudaMalloc(&out, (size_t)blocks * threads * sizeof(double)));
	cudaEvent_t e0, e1;
	OSC_CUDA(cudaEventCreate(&e0));
	OSC_CUDA(cudaEventCreate(&e1));
	float best = 1e30f;
	for (int rep = 0; rep < 4; ++rep) {
		OSC_CUDA(cudaEventRecord(e0));
		k_dfma_peak<<<blocks, threads>>>(out, iters, 1.0 + rep);
		OSC_CUDA(cudaEventRecord(e1));
		OSC_CUDA(cudaEventSynchronize(e1));
		float ms = 0;
		OSC_CUDA(cudaEventElapsedTime(&ms, e0, e1));
		if (rep > 0 && ms < best) best = ms;
	}
	cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
	const double flop = 2.0 * 8.0 * (double)iters * (double)blocks * threads;
	*tflops = flop / (best * 1e-3) / 1e12;
	if (elapsed_ms) *elapsed_ms = best;
	return OSC_OK;
}
