// osc_capi.cu -- the C ABI of include/osc_b200.h (host side).  No torch, no CPU fallback: without a CUDA device every
// compute entry point fails with OSC_ERR_NO_DEVICE.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "osc_capi_util.h"

using namespace osc;

extern "C" {

const char* osc_last_error(void) { return last_error().c_str(); }
int osc_version(void) { return 100; }
int osc_device_count(void) {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
	return n;
}

int osc_batch_create(const osc_chain_desc* chain, int64_t n_robots, int precision, int device, const double* q0, osc_batch_t** out) {
	if (!chain || !out) return invalid("null argument in osc_batch_create");
	*out = nullptr;
	if (chain->n < 1 || chain->n > OSC_MAX_DOF) return invalid("robot dof must be between 1 and OSC_MAX_DOF");
	if (n_robots < 1) return invalid("n_robots must be positive");
	if (precision != OSC_PRECISION_F64 && precision != OSC_PRECISION_F32) return invalid("precision must be 64 or 32");
	int ndev = osc_device_count();
	if (ndev == 0) { set_error("no CUDA device available: the batched controller has no CPU fallback"); return OSC_ERR_NO_DEVICE; }
	if (device < 0 || device >= ndev) return invalid("device index out of range");
	Ops* ops = make_ops(precision, chain->n);
	if (!ops) { set_error("this (precision, dof) combination was not compiled into libosc_b200"); return OSC_ERR_UNSUPPORTED; }
	Batch* b = new (std::nothrow) Batch();
	if (!b) return invalid("out of host memory");
	b->ops = ops;
	b->chain_user = *chain;
	b->n = chain->n;
	b->B = n_robots;
	b->stride = (n_robots + 31) / 32 * 32;
	b->precision = precision;
	b->esz = precision == OSC_PRECISION_F64 ? 8 : 4;
	b->device = device;
	canonicalise_chain(*b);
	ops->set_chain(*b);
	int rc = OSC_OK;
	do {
		if ((rc = use_device(*b))) break;
		ops->robot_rows(b->p_rows, b->s_rows, b->i_rows);
		if ((rc = alloc_rows(*b, &b->P, b->p_rows, b->esz))) break;
		if ((rc = alloc_rows(*b, &b->S, b->s_rows, b->esz))) break;
		if ((rc = alloc_rows(*b, (void**)&b->I, b->i_rows, 4))) break;
		Loc l;
		ops->locate(*b, -1, OSC_F_Q, l);
		if (q0) {
			if ((rc = import_rows(*b, row_ptr(*b, l), l.width, false, q0_as_real(*b, q0), OSC_LAYOUT_UNIFORM, OSC_MEM_HOST))) break;
		}
		if ((rc = ops->model(*b, 1))) break;
	} while (0);
	if (rc != OSC_OK) { osc_batch_destroy((osc_batch_t*)b); return rc; }
	*out = (osc_batch_t*)b;
	return OSC_OK;
}

void osc_batch_destroy(osc_batch_t* batch) {
	Batch* b = (Batch*)batch;
	if (!b) return;
	cudaSetDevice(b->device);
	cudaStreamSynchronize(b->stream);
	for (TaskH& t : b->tasks) { cudaFree(t.P); cudaFree(t.S); cudaFree(t.I); cudaFree(t.popc_ring); cudaFree(t.otgO); cudaFree(t.otgI); cudaFree(t.otgG); cudaFree(t.otg_slow); }
	cudaFree(b->P); cudaFree(b->S); cudaFree(b->I); cudaFree(b->stage); cudaFree(b->stage2); cudaFree(b->fb);
	delete b->ops;
	delete b;
}

int osc_batch_set_stream(osc_batch_t* batch, void* s) {
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	b->stream = (cudaStream_t)s;
	return OSC_OK;
}
int osc_synchronize(osc_batch_t* batch) {
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (int rc = use_device(*b)) return rc;
	OSC_CUDA(cudaStreamSynchronize(b->stream));
	return OSC_OK;
}
int64_t osc_batch_size(const osc_batch_t* b) { return b ? ((const Batch*)b)->B : 0; }
int osc_batch_dof(const osc_batch_t* b) { return b ? ((const Batch*)b)->n : 0; }
int osc_batch_precision(const osc_batch_t* b) { return b ? ((const Batch*)b)->precision : 0; }
int64_t osc_launch_count(const osc_batch_t* b) { return b ? ((const Batch*)b)->launches : 0; }
int osc_batch_task_count(const osc_batch_t* b) { return b ? (int)((const Batch*)b)->tasks.size() : 0; }
int osc_batch_task_type(const osc_batch_t* batch, int task) {
	const Batch* b = (const Batch*)batch;
	if (!b || task < 0 || task >= (int)b->tasks.size()) return invalid("task index out of range");
	return b->tasks[task].type;
}
int osc_batch_task_dof(const osc_batch_t* batch, int task) {
	const Batch* b = (const Batch*)batch;
	if (!b || task < 0 || task >= (int)b->tasks.size()) return invalid("task index out of range");
	return b->tasks[task].dof;
}
int osc_enable_gravity_compensation(osc_batch_t* batch, int enable) {
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	b->gravity = enable != 0;
	return OSC_OK;
}

}  // extern "C"

#include "osc_capi_tasks.inl"
#include "osc_capi_fields.inl"
#include "osc_capi_tick.inl"
