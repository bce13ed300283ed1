// osc_capi_util.h -- helpers of the C ABI translation unit (errors, allocation, chain canonicalisation, layout moves).
#pragma once
#include <nvtx3/nvToolsExt.h>  // header-only NVTX v3: no link dependency, a no-op unless a profiler injects itself
#include <math.h>
#include <string.h>

#include <string>
#include <vector>

#include "osc_host.h"

namespace osc {
// NVTX range over one C-ABI call (nsys / ncu --nvtx show the control cycle by entry point: SURVEY.md 5 "tracing")
struct NvtxRange {
	explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
	~NvtxRange() { nvtxRangePop(); }
	NvtxRange(const NvtxRange&) = delete;
	NvtxRange& operator=(const NvtxRange&) = delete;
};
#define OSC_RANGE(name) ::osc::NvtxRange osc_nvtx_range_(name)


inline std::string& last_error() {
	static thread_local std::string e;
	return e;
}
void set_error(const std::string& msg) { last_error() = msg; }
int cuda_fail(cudaError_t e, const char* what) {
	set_error(std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what);
	return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? OSC_ERR_NO_DEVICE : OSC_ERR_CUDA;
}
inline int invalid(const std::string& msg) {
	set_error(msg);
	return OSC_ERR_INVALID_ARGUMENT;
}
inline int use_device(const Batch& b) {
	OSC_CUDA(cudaSetDevice(b.device));
	return OSC_OK;
}
inline int alloc_rows(Batch& b, void** p, int64_t rows, size_t esz) {
	const size_t bytes = (size_t)rows * b.stride * esz;
	OSC_CUDA(cudaMalloc(p, bytes ? bytes : 16));
	OSC_CUDA(cudaMemsetAsync(*p, 0, bytes, b.stream));
	return OSC_OK;
}
void* Batch::ensure_stage(size_t bytes) {
	if (bytes > stage_bytes) {
		cudaStreamSynchronize(stream);
		cudaFree(stage);
		stage = nullptr;
		if (cudaMalloc(&stage, bytes) != cudaSuccess) { stage_bytes = 0; return nullptr; }
		stage_bytes = bytes;
	}
	return stage;
}
void* Batch::ensure_stage2(size_t bytes) {
	if (bytes > stage2_bytes) {
		cudaStreamSynchronize(stream);
		cudaFree(stage2);
		stage2 = nullptr;
		if (cudaMalloc(&stage2, bytes) != cudaSuccess) { stage2_bytes = 0; return nullptr; }
		stage2_bytes = bytes;
	}
	return stage2;
}

// ---- small host 3x3 helpers ----
inline void h_mm3(const double a[3][3], const double b[3][3], double o[3][3]) {
	double t[3][3];
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j) t[i][j] = a[i][0] * b[0][j] + a[i][1] * b[1][j] + a[i][2] * b[2][j];
	memcpy(o, t, sizeof t);
}
inline void h_tr3(const double a[3][3], double o[3][3]) {
	double t[3][3];
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j) t[i][j] = a[j][i];
	memcpy(o, t, sizeof t);
}
inline void h_mv3(const double a[3][3], const double v[3], double o[3]) {
	double t[3];
	for (int i = 0; i < 3; ++i) t[i] = a[i][0] * v[0] + a[i][1] * v[1] + a[i][2] * v[2];
	memcpy(o, t, sizeof t);
}
// rotation Q with Q e_z = a (a unit): minimal rotation; a = -e_z -> rotation by pi about x
inline void h_align_z(const double a[3], double Q[3][3]) {
	const double c = a[2];
	if (c > 1.0 - 1e-15) { const double I[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}; memcpy(Q, I, sizeof I); return; }
	if (c < -1.0 + 1e-15) { const double F[3][3] = {{1, 0, 0}, {0, -1, 0}, {0, 0, -1}}; memcpy(Q, F, sizeof F); return; }
	// axis k = e_z x a = (-a_y, a_x, 0), sin = |k|
	const double kx = -a[1], ky = a[0], s = sqrt(kx * kx + ky * ky);
	const double ux = kx / s, uy = ky / s, v = 1 - c;
	const double R[3][3] = {{c + ux * ux * v, ux * uy * v, uy * s}, {uy * ux * v, c + uy * uy * v, -ux * s}, {-uy * s, ux * s, c}};
	memcpy(Q, R, sizeof R);
}

// Re-express the chain so that every joint axis is the local z axis of its body frame (canonical body i = user body i * Qc[i]).
// World-frame quantities (positions, axes, J, M) are unchanged; body-frame constants are rotated accordingly.
inline void canonicalise_chain(Batch& b) {
	const osc_chain_desc& u = b.chain_user;
	osc_chain_desc& c = b.chain;
	c = u;
	double Qprev[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
	for (int i = 0; i < u.n; ++i) {
		double a[3] = {u.joint_axis[i][0], u.joint_axis[i][1], u.joint_axis[i][2]};
		const double na = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
		for (double& x : a) x /= (na > 0 ? na : 1);
		double Q[3][3], QpT[3][3], QT[3][3];
		h_align_z(a, Q);
		h_tr3(Qprev, QpT);
		h_tr3(Q, QT);
		double rot[3][3], t[3][3];
		for (int k = 0; k < 9; ++k) rot[k / 3][k % 3] = u.joint_rot[i][k];
		h_mm3(QpT, rot, t);
		h_mm3(t, Q, t);
		for (int k = 0; k < 9; ++k) c.joint_rot[i][k] = t[k / 3][k % 3];
		h_mv3(QpT, u.joint_xyz[i], c.joint_xyz[i]);
		c.joint_axis[i][0] = 0; c.joint_axis[i][1] = 0; c.joint_axis[i][2] = 1;
		h_mv3(QT, u.com[i], c.com[i]);
		double I[3][3];
		for (int k = 0; k < 9; ++k) I[k / 3][k % 3] = u.inertia[i][k];
		h_mm3(QT, I, t);
		h_mm3(t, Q, t);
		for (int k = 0; k < 9; ++k) c.inertia[i][k] = t[k / 3][k % 3];
		memcpy(b.Qc[i], Q, sizeof Q);
		memcpy(Qprev, Q, sizeof Q);
	}
}

// ---- layout moves between user buffers and SoA rows ----
// AOS [B][w] <-> SoA rows, through a shared-memory tile of 32 robots x w per warp (two warps per block, <= 33 KB): the AOS side is read / written as ONE
// contiguous, fully coalesced range (32 w elements), the SoA side row by row (32 consecutive robots each).  The direct form
// (aos[r * w + e] per thread) is a stride-w gather: w x the sectors.  Row pitch w | 1 keeps the tile conflict free.
constexpr int OSC_TILE_MAX_W = 64;
template <class E> __global__ void __launch_bounds__(64) k_aos_to_soa(const E* __restrict__ aos, E* __restrict__ soa, int w, int64_t B, int64_t stride) {
	extern __shared__ __align__(16) unsigned char osc_tile_raw[];
	const int wp = w | 1, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	E* tile = reinterpret_cast<E*>(osc_tile_raw) + (size_t)warp * 32 * wp;
	const int64_t r0 = ((int64_t)blockIdx.x * 2 + warp) * 32;
	if (r0 >= B) return;
	const int64_t base = r0 * w, end = B * (int64_t)w;
	for (int i = lane; i < 32 * w; i += 32)
		if (base + i < end) tile[(i / w) * wp + (i % w)] = aos[base + i];
	__syncwarp();
	if (r0 + lane < B)
		for (int e = 0; e < w; ++e) soa[(int64_t)e * stride + r0 + lane] = tile[lane * wp + e];
}
template <class E> __global__ void __launch_bounds__(64) k_soa_to_aos(const E* __restrict__ soa, E* __restrict__ aos, int w, int64_t B, int64_t stride) {
	extern __shared__ __align__(16) unsigned char osc_tile_raw[];
	const int wp = w | 1, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	E* tile = reinterpret_cast<E*>(osc_tile_raw) + (size_t)warp * 32 * wp;
	const int64_t r0 = ((int64_t)blockIdx.x * 2 + warp) * 32;
	if (r0 >= B) return;
	if (r0 + lane < B)
		for (int e = 0; e < w; ++e) tile[lane * wp + e] = soa[(int64_t)e * stride + r0 + lane];
	__syncwarp();
	const int64_t base = r0 * w, end = B * (int64_t)w;
	for (int i = lane; i < 32 * w; i += 32)
		if (base + i < end) aos[base + i] = tile[(i / w) * wp + (i % w)];
}
// widths beyond the tile (none of the reference's fields: the widest getter is 7 x 7): one thread per robot
template <class E> __global__ void k_aos_to_soa_wide(const E* __restrict__ aos, E* __restrict__ soa, int w, int64_t B, int64_t stride) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	for (int e = 0; e < w; ++e) soa[(int64_t)e * stride + r] = aos[r * w + e];
}
template <class E> __global__ void k_soa_to_aos_wide(const E* __restrict__ soa, E* __restrict__ aos, int w, int64_t B, int64_t stride) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	for (int e = 0; e < w; ++e) aos[r * w + e] = soa[(int64_t)e * stride + r];
}
struct Uniform64 { unsigned char bytes[64 * 8]; };
template <class E> __global__ void k_broadcast(const __grid_constant__ Uniform64 u, E* __restrict__ soa, int w, int64_t B, int64_t stride) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const E* v = reinterpret_cast<const E*>(u.bytes);
	for (int e = 0; e < w; ++e) soa[(int64_t)e * stride + r] = v[e];
}

// per-robot flag value < tolerance (goalPositionReached / goalOrientationReached) + the number of robots that reached
template <class St> __global__ void k_below_tolerance(const St* v, double tol, int32_t* flags, int* count, int64_t B) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const bool ok = r < B && (double)v[r] < tol;
	if (r < B) flags[r] = ok ? 1 : 0;
	const unsigned m = __ballot_sync(0xffffffffu, ok);
	if ((threadIdx.x & 31) == 0 && m) atomicAdd(count, __popc(m));
}

inline void* arena_base(Batch& b, const Loc& l) {
	if (l.owner < 0) return l.arena == 0 ? b.P : (l.arena == 1 ? b.S : (void*)b.I);
	TaskH& t = b.tasks[l.owner];
	if (l.arena == 3) return t.otgG;
	if (l.arena == 4) return (void*)t.otgI;
	if (l.arena == 5) return (void*)t.otgO;
	return l.arena == 0 ? t.P : (l.arena == 1 ? t.S : (void*)t.I);
}
inline void* row_ptr(Batch& b, const Loc& l) {
	const size_t esz = l.is_int ? 4 : b.esz;
	return (char*)arena_base(b, l) + (size_t)l.offset * b.stride * esz;
}

template <class E> int import_rows_t(Batch& b, E* rows, int w, const void* data, int layout, int mem) {
	const dim3 g((unsigned)((b.B + 255) / 256));
	if (layout == OSC_LAYOUT_UNIFORM) {
		if (w > 64) return invalid("uniform layout supports at most 64 scalars per robot");
		Uniform64 u;
		if (mem == OSC_MEM_HOST) memcpy(u.bytes, data, (size_t)w * sizeof(E));
		else OSC_CUDA(cudaMemcpy(u.bytes, data, (size_t)w * sizeof(E), cudaMemcpyDeviceToHost));
		k_broadcast<E><<<g, 256, 0, b.stream>>>(u, rows, w, b.B, b.stride);
		++b.launches;
		OSC_CUDA(cudaGetLastError());
		return OSC_OK;
	}
	if (layout == OSC_LAYOUT_SOA) {
		const cudaMemcpyKind kind = mem == OSC_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
		if (b.io_pitch) {  // a shard's columns of its parent's [k][B_total] buffer
			OSC_CUDA(cudaMemcpy2DAsync(rows, (size_t)b.stride * sizeof(E), data, (size_t)b.io_pitch * sizeof(E), (size_t)b.B * sizeof(E), w, kind, b.stream));
			return OSC_OK;
		}
		if (b.stride == b.B) {  // rows are contiguous: one linear copy (the 2-D engine path is slower over PCIe)
			OSC_CUDA(cudaMemcpyAsync(rows, data, (size_t)b.B * sizeof(E) * w, kind, b.stream));
			return OSC_OK;
		}
		OSC_CUDA(cudaMemcpy2DAsync(rows, (size_t)b.stride * sizeof(E), data, (size_t)b.B * sizeof(E), (size_t)b.B * sizeof(E), w,
								   mem == OSC_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, b.stream));
		return OSC_OK;
	}
	if (layout != OSC_LAYOUT_AOS) return invalid("unknown layout");
	const E* src = (const E*)data;
	if (mem == OSC_MEM_HOST) {
		void* st = b.ensure_stage((size_t)b.B * w * sizeof(E));
		if (!st) return invalid("out of device memory (staging)");
		OSC_CUDA(cudaMemcpyAsync(st, data, (size_t)b.B * w * sizeof(E), cudaMemcpyHostToDevice, b.stream));
		src = (const E*)st;
	}
	if (w <= OSC_TILE_MAX_W) k_aos_to_soa<E><<<dim3((unsigned)((b.B + 63) / 64)), 64, (size_t)2 * 32 * (w | 1) * sizeof(E), b.stream>>>(src, rows, w, b.B, b.stride);
	else k_aos_to_soa_wide<E><<<g, 256, 0, b.stream>>>(src, rows, w, b.B, b.stride);
	++b.launches;
	OSC_CUDA(cudaGetLastError());
	return OSC_OK;
}
inline int import_rows(Batch& b, void* rows, int w, bool is_int, const void* data, int layout, int mem) {
	if (is_int || b.esz == 4) return import_rows_t<uint32_t>(b, (uint32_t*)rows, w, data, layout, mem);
	return import_rows_t<uint64_t>(b, (uint64_t*)rows, w, data, layout, mem);
}

template <class E> int export_rows_t(Batch& b, const E* rows, int w, void* out, int layout, int mem) {
	const dim3 g((unsigned)((b.B + 255) / 256));
	if (layout == OSC_LAYOUT_SOA && b.io_pitch) {
		OSC_CUDA(cudaMemcpy2DAsync(out, (size_t)b.io_pitch * sizeof(E), rows, (size_t)b.stride * sizeof(E), (size_t)b.B * sizeof(E), w,
								   mem == OSC_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, b.stream));
	} else if (layout == OSC_LAYOUT_SOA && b.stride == b.B) {
		OSC_CUDA(cudaMemcpyAsync(out, rows, (size_t)b.B * sizeof(E) * w, mem == OSC_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, b.stream));
	} else if (layout == OSC_LAYOUT_SOA) {
		OSC_CUDA(cudaMemcpy2DAsync(out, (size_t)b.B * sizeof(E), rows, (size_t)b.stride * sizeof(E), (size_t)b.B * sizeof(E), w,
								   mem == OSC_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, b.stream));
	} else if (layout == OSC_LAYOUT_AOS) {
		E* dst = (E*)out;
		if (mem == OSC_MEM_HOST) {
			dst = (E*)b.ensure_stage((size_t)b.B * w * sizeof(E));
			if (!dst) return invalid("out of device memory (staging)");
		}
		if (w <= OSC_TILE_MAX_W) k_soa_to_aos<E><<<dim3((unsigned)((b.B + 63) / 64)), 64, (size_t)2 * 32 * (w | 1) * sizeof(E), b.stream>>>(rows, dst, w, b.B, b.stride);
		else k_soa_to_aos_wide<E><<<g, 256, 0, b.stream>>>(rows, dst, w, b.B, b.stride);
		++b.launches;
		OSC_CUDA(cudaGetLastError());
		if (mem == OSC_MEM_HOST) OSC_CUDA(cudaMemcpyAsync(out, dst, (size_t)b.B * w * sizeof(E), cudaMemcpyDeviceToHost, b.stream));
	} else {
		return invalid("layout must be AOS or SOA for output buffers");
	}
	if (mem == OSC_MEM_HOST && !b.defer_host_sync) OSC_CUDA(cudaStreamSynchronize(b.stream));
	return OSC_OK;
}
inline int export_rows(Batch& b, const void* rows, int w, bool is_int, void* out, int layout, int mem) {
	if (is_int || b.esz == 4) return export_rows_t<uint32_t>(b, (const uint32_t*)rows, w, out, layout, mem);
	return export_rows_t<uint64_t>(b, (const uint64_t*)rows, w, out, layout, mem);
}

// q0 (double[n]) as a real[n] buffer of the batch precision
inline const void* q0_as_real(Batch& b, const double* q0) {
	static thread_local double d[OSC_MAX_DOF];
	static thread_local float f[OSC_MAX_DOF];
	if (b.esz == 8) { memcpy(d, q0, sizeof(double) * b.n); return d; }
	for (int i = 0; i < b.n; ++i) f[i] = (float)q0[i];
	return f;
}

}  // namespace osc
