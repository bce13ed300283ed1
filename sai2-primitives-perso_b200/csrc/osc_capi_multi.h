// osc_capi_multi.h -- the multi-device batch behind the same osc_batch_t handle (one host thread + one
// stream per device, contiguous shards [g B / G, (g+1) B / G), no inter-GPU traffic, no collective).
//
// osc_batch_create_multi() returns a PARENT batch: it owns one ordinary batch per device (its shards) and a pool of one host
// thread per shard.  Every C-ABI entry point called on a parent fans out: the same call runs on every shard from that shard's
// own thread (so that kernel launches and copies of the G devices are issued concurrently), per-robot buffers are sliced
// (AOS: a contiguous block; SOA: the shard's columns, copied with a row pitch of B_total; UNIFORM: as is), host-destination
// copies of the shards land in disjoint slices of the caller's array and the call returns when all of them have.
#pragma once
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>

#include "osc_capi_util.h"

namespace osc {

struct MultiPool {
	int n = 0;
	std::vector<std::thread> th;
	std::mutex mu;
	std::condition_variable cv_job, cv_done;
	std::atomic<uint64_t> gen{0};
	std::atomic<int> pending{0};
	bool stop = false;
	std::function<int(int)> job;
	std::vector<int> rc;
	std::vector<std::string> err;

	explicit MultiPool(int n_) : n(n_), rc(n_, 0), err(n_) {
		for (int k = 0; k < n; ++k) th.emplace_back([this, k] { run(k); });
	}
	~MultiPool() {
		{
			std::lock_guard<std::mutex> l(mu);
			stop = true;
			gen.fetch_add(1);
		}
		cv_job.notify_all();
		for (auto& t : th) t.join();
	}
	void run(int k) {
		uint64_t seen = 0;
		for (;;) {
			// a control loop posts a job every few tens of microseconds: spin briefly before sleeping on the condition variable
			for (int spin = 0; spin < 20000 && gen.load(std::memory_order_acquire) == seen; ++spin) {
#if defined(__x86_64__)
				__builtin_ia32_pause();
#endif
			}
			if (gen.load(std::memory_order_acquire) == seen) {
				std::unique_lock<std::mutex> l(mu);
				cv_job.wait(l, [&] { return gen.load() != seen; });
			}
			seen = gen.load(std::memory_order_acquire);
			if (stop) return;
			last_error().clear();
			rc[k] = job(k);
			err[k] = last_error();
			if (pending.fetch_sub(1, std::memory_order_acq_rel) == 1) {
				std::lock_guard<std::mutex> l(mu);
				cv_done.notify_all();
			}
		}
	}
	// runs f(k) for every shard k on its thread; returns the first error (its message becomes the caller's osc_last_error)
	int run_all(std::function<int(int)> f) {
		{
			std::lock_guard<std::mutex> l(mu);
			job = std::move(f);
			pending.store(n, std::memory_order_release);
			gen.fetch_add(1, std::memory_order_acq_rel);
		}
		cv_job.notify_all();
		for (int spin = 0; spin < 200000 && pending.load(std::memory_order_acquire) != 0; ++spin) {
#if defined(__x86_64__)
			__builtin_ia32_pause();
#endif
		}
		if (pending.load(std::memory_order_acquire) != 0) {
			std::unique_lock<std::mutex> l(mu);
			cv_done.wait(l, [&] { return pending.load() == 0; });
		}
		for (int k = 0; k < n; ++k)
			if (rc[k] < 0) { set_error(err[k]); return rc[k]; }
		return OSC_OK;
	}
};

inline bool is_multi(const void* batch) { return batch && !((const Batch*)batch)->shards.empty(); }
inline Batch* shard0(const void* batch) { return ((const Batch*)batch)->shards[0]; }

// fan a call out over the shards: f(shard handle, k) -> status.  Returns the first error, else the status of shard 0
// (task indices and reset masks are the same on every shard by construction).
template <class F> int multi_fan(Batch* p, F f) {
	if (int rc = p->pool->run_all([&](int k) { return f((osc_batch_t*)p->shards[k], k); })) return rc;
	return p->pool->rc[0];
}

// the part of a caller's per-robot buffer that belongs to shard k (sets / clears the shard's SOA row pitch)
inline void* multi_slice(Batch* p, int k, const void* data, int width, int layout, size_t esz) {
	Batch* s = p->shards[k];
	s->io_pitch = 0;
	if (!data || layout == OSC_LAYOUT_UNIFORM) return const_cast<void*>(data);
	const int64_t first = p->shard_first[k];
	if (layout == OSC_LAYOUT_AOS) return (char*)const_cast<void*>(data) + (size_t)first * width * esz;
	s->io_pitch = p->B;  // SOA [width][B_total]: the shard's columns
	return (char*)const_cast<void*>(data) + (size_t)first * esz;
}
inline int multi_host_only(int mem) {
	if (mem != OSC_MEM_HOST) {
		set_error("a multi-device batch takes HOST buffers only (device buffers belong to one device: use osc_batch_shard(batch, k) for per-device calls)");
		return OSC_ERR_UNSUPPORTED;
	}
	return OSC_OK;
}

}  // namespace osc
