// osc_host.h -- host-side state of a batch and the per-(precision, dof) operation table.
// The C ABI (osc_capi.cu) is precision/dof agnostic; everything that launches a templated kernel goes through Ops,
// implemented once per (T, N) in osc_ops_inst.cu (compiled once per instantiation so the builds run in parallel).
#pragma once
#include <mutex>
#include <unordered_map>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/osc_b200.h"

namespace osc {

struct TaskH {
	int type = OSC_TASK_UNDEFINED;
	osc_mft_desc md{};
	osc_mft_params mp{};
	osc_joint_task_desc jd{};
	osc_joint_task_params jp{};
	int dof = 0;           // MotionForceTask: _pos_range + _ori_range; JointTask: k
	double P6[6][6]{};     // _partial_task_projection
	double Bt[3][3]{}, Br[3][3]{};  // orthonormal bases of its two diagonal blocks (columns), ranks mt, mr
	int mt = 0, mr = 0;
	double frame_p[3]{}, frame_R[3][3]{};  // compliant frame in the CANONICAL body frame
	void *P = nullptr, *S = nullptr;
	int32_t* I = nullptr;
	int64_t p_rows = 0, s_rows = 0, i_rows = 0;
	void* popc_ring = nullptr;
	// internal OTG (osc_otg.cuh), allocated when first enabled: O = double state rows, OI = int rows, G = user goals (T)
	bool otg_on = false;
	double *otgO = nullptr;
	int32_t* otgI = nullptr;
	void* otgG = nullptr;
	int64_t otg_o_rows = 0, otg_i_rows = 0, otg_g_rows = 0;
	double otg_max_v[8]{}, otg_max_a[8]{};
	bool otg_goals_dirty = true;  // a goal field was written (or the state restored) since the last k_otg_advance
	int* otg_slow = nullptr;      // device: [2] ping-pong counters + [stride] list of the robots on the general path
	int otg_parity = 0;
};

struct Loc {
	int arena = 0;   // 0 P, 1 S, 2 I, 3 G (user goals while the internal OTG is on), 4 OTG ints
	int offset = 0;
	int width = 0;
	bool is_int = false;
	bool writable = false;
	int special = 0; // != 0: computed by Ops::pack_special into a temporary
	int owner = -1;  // task index or -1 (robot)
};

struct Batch;

struct Ops {
	virtual ~Ops() {}
	virtual void robot_rows(int64_t& p, int64_t& s, int64_t& i) = 0;
	virtual void task_rows(int type, int64_t& p, int64_t& s, int64_t& i) = 0;
	virtual int set_chain(Batch& b) = 0;  // convert the canonical chain to ChainC<T,N>
	virtual int model(Batch& b, int dynamics) = 0;
	virtual int task_init(Batch& b, int t) = 0;
	virtual int task_reinit(Batch& b, int t) = 0;
	virtual int mft_reset(Batch& b, int t, int flags) = 0;  // 1 linear, 2 angular, 4 popc, 8 lin integrators, 16 ang integrators
	virtual int task_update(Batch& b, int t, int src, int no_history = 0) = 0;
	virtual int task_compute(Batch& b, int t, int with_tau_prec, int accumulate) = 0;
	virtual int add_gravity(Batch& b) = 0;
	virtual int update_sensed(Batch& b, int t, const void* tmp6) = 0;
	virtual bool locate(Batch& b, int task, int field, Loc& loc) = 0;
	virtual int pack_special(Batch& b, const Loc& loc, void* tmp) = 0;
	virtual bool fused_supported(Batch& b) = 0;
	virtual int tick_fused(Batch& b, int stacked) = 0;
	virtual int materialise(Batch& b) = 0;
	virtual int rollout_step(Batch& b, double dt, int substeps, double damping) = 0;  // osc_rollout.cuh
	virtual void otg_rows(Batch& b, int t, int64_t& o, int64_t& i, int64_t& g) = 0;
	virtual int otg_reinit(Batch& b, int t, int mode) = 0;  // k_otg_reinit modes (osc_otg.cuh)
	virtual int otg_tick(Batch& b, int t) = 0;              // goal filter + update() + desired rows of task t (all tasks: t < 0)  // after a fused tick: run the per-call model / task-model kernels (no state advance)
};

Ops* make_ops(int precision, int n);  // nullptr when that (precision, dof) pair was not compiled in

struct Batch {
	osc_chain_desc chain_user{};   // as given
	osc_chain_desc chain{};        // canonical: every joint axis = local z
	double Qc[OSC_MAX_DOF][3][3]{}; // body-frame change: canonical body i = user body i * Qc[i]
	int n = 0;
	int64_t B = 0, stride = 0;
	int precision = 64, device = 0;
	size_t esz = 8;
	cudaStream_t stream = nullptr;
	void *P = nullptr, *S = nullptr;
	int32_t* I = nullptr;
	int64_t p_rows = 0, s_rows = 0, i_rows = 0;
	std::vector<TaskH> tasks;
	double dt = 0;
	bool closed = false;  // a full joint task was added: no further task (RobotController.cpp:36-49)
	bool gravity = false;
	int64_t launches = 0;
	Ops* ops = nullptr;
	int tick_outputs = OSC_TICK_OUTPUTS_LAZY;
	bool stale = false;        // a fused tick ran: the model / task-model scratch rows are older than tau
	int* fb = nullptr;         // device: [2] fallback counters (ping-pong) + [stride] robot list
	int fb_parity = 0;
	int fb_last = 0;           // which counter the last fused tick filled
	void* tail_scratch = nullptr;      // compact S rows of k_tick_tail, one slot per tail thread (osc_fused.cuh)
	int64_t tail_scratch_stride = 0;
	bool may_allocate = true;          // false inside osc_tick_noalloc
	void* chain_dev = nullptr;  // ChainC<T,N> host copy owned by ops
	// staging
	void* stage = nullptr; size_t stage_bytes = 0;
	void* stage2 = nullptr; size_t stage2_bytes = 0;
	void* ensure_stage(size_t bytes);
	void* ensure_stage2(size_t bytes);
	// ---- multi-device batch (osc_batch_create_multi): a parent owns one shard batch per device and a pool of one host
	// thread per shard; every entry point fans out over the shards (osc_capi_multi.h).  A parent holds no device memory.
	std::vector<Batch*> shards;
	std::vector<int64_t> shard_first;  // first robot of shard k in the parent's numbering
	struct MultiPool* pool = nullptr;
	// SOA user buffers of a shard call made by its parent: rows are `io_pitch` elements apart (0 = this batch's own size)
	int64_t io_pitch = 0;
	bool defer_host_sync = false;  // osc_step_host: a device-to-host export does not wait for the stream (the caller synchronises all shards at the end)
	cudaStream_t owned_stream = nullptr;  // created for a shard by osc_batch_create_multi, destroyed with it
};

void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what);
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is needed once per (kernel, device), not per launch.  Keyed by the kernel's
// address: kernels that differ only in a template argument have the same function type, so a static flag inside a generic
// lambda would be shared between them.
inline bool first_launch_on_device(const void* kernel, int device) {
	static std::mutex mu;
	static std::unordered_map<const void*, uint64_t> seen;
	std::lock_guard<std::mutex> lock(mu);
	uint64_t& mask = seen[kernel];
	if (device < 0 || device >= 64) return true;
	const uint64_t bit = 1ull << device;
	if (mask & bit) return false;
	mask |= bit;
	return true;
}

#define OSC_CUDA(x)                                      \
	do {                                                 \
		cudaError_t e_ = (x);                            \
		if (e_ != cudaSuccess) return cuda_fail(e_, #x); \
	} while (0)

}  // namespace osc
