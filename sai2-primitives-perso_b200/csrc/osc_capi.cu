// osc_capi.cu -- the C ABI of include/osc_b200.h (host side).  No torch, no CPU fallback: without a CUDA device every
// compute entry point fails with OSC_ERR_NO_DEVICE.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "osc_capi_util.h"
#include "osc_capi_multi.h"

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
	if (is_multi(b)) {
		delete b->pool;  // joins the shard threads
		for (Batch* s : b->shards) osc_batch_destroy((osc_batch_t*)s);
		delete b;
		return;
	}
	cudaSetDevice(b->device);
	cudaStreamSynchronize(b->stream);
	for (TaskH& t : b->tasks) { cudaFree(t.P); cudaFree(t.S); cudaFree(t.I); cudaFree(t.popc_ring); cudaFree(t.otgO); cudaFree(t.otgI); cudaFree(t.otgG); cudaFree(t.otg_slow); }
	cudaFree(b->P); cudaFree(b->S); cudaFree(b->I); cudaFree(b->stage); cudaFree(b->stage2); cudaFree(b->fb); cudaFree(b->tail_scratch);
	if (b->owned_stream) cudaStreamDestroy(b->owned_stream);
	delete b->ops;
	delete b;
}

int osc_batch_set_stream(osc_batch_t* batch, void* s) {
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (is_multi(b)) return invalid("a multi-device batch has one stream per device: set it on osc_batch_shard(batch, k)");
	b->stream = (cudaStream_t)s;
	return OSC_OK;
}
int osc_synchronize(osc_batch_t* batch) {
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (is_multi(b)) return multi_fan(b, [](osc_batch_t* s, int) { return osc_synchronize(s); });
	if (int rc = use_device(*b)) return rc;
	OSC_CUDA(cudaStreamSynchronize(b->stream));
	return OSC_OK;
}
int64_t osc_batch_size(const osc_batch_t* b) { return b ? ((const Batch*)b)->B : 0; }
int osc_batch_dof(const osc_batch_t* b) { return b ? ((const Batch*)b)->n : 0; }
int osc_batch_precision(const osc_batch_t* b) { return b ? ((const Batch*)b)->precision : 0; }
int64_t osc_launch_count(const osc_batch_t* b) {
	if (is_multi(b)) {
		int64_t n = 0;
		for (const Batch* s : ((const Batch*)b)->shards) n += s->launches;
		return n;
	}
	return b ? ((const Batch*)b)->launches : 0;
}
int osc_batch_task_count(const osc_batch_t* b) {
	if (is_multi(b)) b = (const osc_batch_t*)shard0(b);
	return b ? (int)((const Batch*)b)->tasks.size() : 0;
}
int osc_batch_task_type(const osc_batch_t* batch, int task) {
	if (is_multi(batch)) batch = (const osc_batch_t*)shard0(batch);
	const Batch* b = (const Batch*)batch;
	if (!b || task < 0 || task >= (int)b->tasks.size()) return invalid("task index out of range");
	return b->tasks[task].type;
}
int osc_batch_task_dof(const osc_batch_t* batch, int task) {
	if (is_multi(batch)) batch = (const osc_batch_t*)shard0(batch);
	const Batch* b = (const Batch*)batch;
	if (!b || task < 0 || task >= (int)b->tasks.size()) return invalid("task index out of range");
	return b->tasks[task].dof;
}
int osc_enable_gravity_compensation(osc_batch_t* batch, int enable) {
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (is_multi(b)) return multi_fan(b, [=](osc_batch_t* s, int) { return osc_enable_gravity_compensation(s, enable); });
	b->gravity = enable != 0;
	return OSC_OK;
}

int osc_batch_create_multi(const osc_chain_desc* chain, int64_t n_robots, int precision, const int* device_ids, int n_devices, const double* q0,
						   osc_batch_t** out) {
	if (!chain || !out || !device_ids) return invalid("null argument in osc_batch_create_multi");
	*out = nullptr;
	if (n_devices < 1 || n_devices > 64) return invalid("n_devices must be between 1 and 64");
	if (n_robots < n_devices) return invalid("fewer robots than devices");
	Batch* p = new (std::nothrow) Batch();
	if (!p) return invalid("out of host memory");
	p->n = chain->n; p->B = n_robots; p->precision = precision; p->esz = precision == OSC_PRECISION_F64 ? 8 : 4;
	p->chain_user = *chain; p->device = device_ids[0];
	for (int k = 0; k < n_devices; ++k) {  // contiguous shards [k B / G, (k + 1) B / G)  
		const int64_t lo = n_robots * k / n_devices, hi = n_robots * (k + 1) / n_devices;
		osc_batch_t* s = nullptr;
		const int rc = osc_batch_create(chain, hi - lo, precision, device_ids[k], q0, &s);
		if (rc != OSC_OK) {
			for (Batch* t : p->shards) osc_batch_destroy((osc_batch_t*)t);
			delete p;
			return rc;
		}
		// its own non-blocking stream, so that shards on one device (tests) and the default stream do not serialise each other
		cudaStream_t st = nullptr;
		if (cudaSetDevice(device_ids[k]) == cudaSuccess && cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) == cudaSuccess) {
			cudaStreamSynchronize(((Batch*)s)->stream);
			((Batch*)s)->stream = st;
			((Batch*)s)->owned_stream = st;
		}
		p->shards.push_back((Batch*)s);
		p->shard_first.push_back(lo);
	}
	p->pool = new MultiPool(n_devices);
	*out = (osc_batch_t*)p;
	return OSC_OK;
}
int osc_batch_shard_count(const osc_batch_t* batch) { return is_multi(batch) ? (int)((const Batch*)batch)->shards.size() : (batch ? 1 : 0); }
osc_batch_t* osc_batch_shard(osc_batch_t* batch, int k, int* device, int64_t* first_robot, int64_t* n_robots) {
	Batch* b = (Batch*)batch;
	if (!b) return nullptr;
	if (!is_multi(b)) {
		if (k != 0) return nullptr;
		if (device) *device = b->device;
		if (first_robot) *first_robot = 0;
		if (n_robots) *n_robots = b->B;
		return batch;
	}
	if (k < 0 || k >= (int)b->shards.size()) return nullptr;
	if (device) *device = b->shards[k]->device;
	if (first_robot) *first_robot = b->shard_first[k];
	if (n_robots) *n_robots = b->shards[k]->B;
	return (osc_batch_t*)b->shards[k];
}

}  // extern "C"

#include "osc_capi_tasks.inl"
#include "osc_capi_fields.inl"
#include "osc_capi_tick.inl"

// ---- osc_tick_noalloc: the fused tick for callers that may not allocate (no counterpart in the reference).
// The one piece of scratch memory the fused tick uses -- the compacted list of robots
// that take the API-faithful route, [2] ping-pong counters + one int per robot, normally cudaMalloc'ed by the first
// osc_tick of a batch -- lives in this file-scope device array instead.  Every GPU that loads the library holds its
// (OSC_NOALLOC_MAX_ROBOTS + 64) * 4 bytes until exit.  The array's name below is a HOST-side stand-in, not a device
// address: its device address is resolved once per device with cudaGetSymbolAddress, with that device current.
namespace osc {
static __device__ __align__(256) int g_noalloc_fallback[OSC_NOALLOC_MAX_ROBOTS + 64];

static int noalloc_scratch(int device, int** out) {
	static std::mutex mu;
	static int* ptr[64] = {};
	if (device < 0 || device >= 64) return invalid("osc_tick_noalloc: device index out of range");
	std::lock_guard<std::mutex> lock(mu);
	if (!ptr[device]) {
		int cur = -1;
		OSC_CUDA(cudaGetDevice(&cur));
		if (cur != device) OSC_CUDA(cudaSetDevice(device));
		void* p = nullptr;
		OSC_CUDA(cudaGetSymbolAddress(&p, g_noalloc_fallback));
		ptr[device] = (int*)p;
	}
	*out = ptr[device];
	return OSC_OK;
}
}  // namespace osc

extern "C" int osc_tick_noalloc(osc_batch_t* batch, int stacked, void* tau_out_device, int layout) {
	OSC_RANGE("osc_tick_noalloc");
	Batch* b = (Batch*)batch;
	if (!b) return invalid("null batch");
	if (is_multi(b)) {
		if (tau_out_device) return invalid("osc_tick_noalloc on a multi-device batch takes no output pointer (device memory belongs to one device): read OSC_F_TAU");
		return multi_fan(b, [=](osc_batch_t* s, int) { return osc_tick_noalloc(s, stacked, nullptr, layout); });
	}
	if (b->tasks.empty()) return invalid("RobotController must have at least one task");
	if (b->stride > OSC_NOALLOC_MAX_ROBOTS) return invalid("osc_tick_noalloc: the batch has more than OSC_NOALLOC_MAX_ROBOTS (262144) robots; use osc_tick");
	if (tau_out_device && layout != OSC_LAYOUT_AOS && layout != OSC_LAYOUT_SOA) return invalid("osc_tick_noalloc: tau_out layout must be AOS or SOA");
	if (!b->ops->fused_supported(*b)) {
		set_error("osc_tick_noalloc: only the hierarchies of the fused tick (osc_tick_is_fused() == 1, outputs LAZY or TAU) run without allocating; use osc_tick");
		return OSC_ERR_UNSUPPORTED;
	}
	for (const TaskH& t : b->tasks)
		if (t.otg_on && !t.otg_slow) {
			set_error("osc_tick_noalloc: the internal OTG of a task is on and has not ticked yet (its list is allocated by the first osc_tick)");
			return OSC_ERR_UNSUPPORTED;
		}
	if (int rc = use_device(*b)) return rc;
	int* scratch = nullptr;
	if (int rc = noalloc_scratch(b->device, &scratch)) return rc;
	// the array keeps the last call's contents: both counters must start at zero (the list itself is write-before-read)
	OSC_CUDA(cudaMemsetAsync(scratch, 0, 2 * sizeof(int), b->stream));
	int* const own_fb = b->fb;
	const int own_parity = b->fb_parity, own_last = b->fb_last;
	b->fb = scratch;
	b->fb_parity = 0;
	b->may_allocate = false;  // (k_tick_tail's compact scratch is used if an earlier osc_tick allocated it, else its robots work in place)
	int rc = b->ops->otg_tick(*b, -1);
	if (rc == OSC_OK) rc = b->ops->tick_fused(*b, stacked);
	b->may_allocate = true;
	b->fb = own_fb;
	b->fb_parity = own_parity;
	b->fb_last = own_last;
	if (rc != OSC_OK) return rc;
	if (!tau_out_device) return OSC_OK;
	Loc l;
	b->ops->locate(*b, -1, OSC_F_TAU, l);
	return export_rows(*b, row_ptr(*b, l), l.width, false, tau_out_device, layout, OSC_MEM_DEVICE);  // device destination: no staging
}
