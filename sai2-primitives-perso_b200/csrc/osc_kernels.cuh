// osc_kernels.cuh -- the API-faithful kernels: one launch per reference call
//   k_model        robot->updateModel() / updateKinematics()
//   k_mft_update   MotionForceTask::updateTaskModel   (src/tasks/MotionForceTask.cpp:247-268 + SingularityHandler.cpp:75-295)
//   k_jt_update    JointTask::updateTaskModel         (src/tasks/JointTask.cpp:218-283)
//   k_mft_compute  MotionForceTask::computeTorques    (src/tasks/MotionForceTask.cpp:270-509 + SingularityHandler.cpp:297-368
//                                                      + helper_modules/POPCExplicitForceControl.cpp:30-96)
//   k_jt_compute   JointTask::computeTorques          (src/tasks/JointTask.cpp:285-356)
// One robot per thread, SoA state in HBM (osc_layout.cuh).  These follow the reference step by step (full Jacobi SVD,
// every intermediate stored so that every getter works); the single-launch fast path lives in osc_fused.cuh.
#pragma once
#include "osc_layout.cuh"
#include "osc_model.cuh"

namespace osc {

// ---- uniform (per batch) task constants, passed by value as kernel parameters (constant bank) ----
template <class T> struct MftC {
	int body, partial, rank, in_compliant_frame;
	T p_c[3], R_c[3][3];
	T Pt[3][3], Pr[3][3];  // diagonal blocks of _partial_task_projection
	T Bt[3][3], Br[3][3];  // orthonormal bases of the controlled directions (columns 0..mt-1 / 0..mr-1; Pt = Bt Bt^T, Pr = Br Br^T)
	int mt, mr;            // _pos_range, _ori_range (MotionForceTask.cpp:146-152)
	T dt;
	T kp_pos[3], kv_pos[3], ki_pos[3], kp_ori[3], kv_ori[3], ki_ori[3];
	T kp_force, kv_force, ki_force, kp_moment, kv_moment, ki_moment, kff_force, kff_moment, max_force_fb, max_moment_fb;
	int force_dim, moment_dim;
	T force_axis[3], moment_axis[3];
	int closed_loop_force, closed_loop_moment, passivity, use_vel_sat;
	T lin_sat, ang_sat;
	int decoupling;
	T bie_threshold;
	int handling, enforce_type_1;
	T s_min, s_max, kp1, kv1, kv2, s_abs_tol, type_1_tol, type_2_torque_ratio, type_2_angle_threshold, perturb_step;
	int buffer_size;
	T sensor_R[3][3], sensor_p[3];
	int apply_compensation;
	void* popc_ring;  // [OSC_POPC_RING_CAPACITY][stride] in the batch STORAGE type, or nullptr until passivity is first enabled
};

template <class T, int N> struct JtC {
	int k, full;
	T S[N][N];  // rows >= k are zero
	T dt;
	T kp[N], kv[N], ki[N], sat[N];
	int use_vel_sat, decoupling;
	T bie_threshold;
};

// ---- SoA access ----
// T = the type the kernel computes in, S = the type the rows are stored in (S = T for an FP64 batch and for the fused kernels).
// The API-faithful kernels of an FP32 batch run with T = double over S = float rows: their task models (Lambda, Jbar and above
// all N = I - Jbar J, a cancellation whose FP32 rounding noise eps |Jbar| |J| reaches O(1) near singular postures and is then
// mistaken for range by the next task of the hierarchy) need FP64 arithmetic to hold the 1e-4 FP32 tolerance of the north star.
template <class T, class S> struct Ref {  // element proxy of a converting row
	S* p;
	OSC_DEV operator T() const { return (T)*p; }
	OSC_DEV const Ref& operator=(T v) const { *p = (S)v; return *this; }
	OSC_DEV const Ref& operator=(const Ref& o) const { *p = *o.p; return *this; }
	OSC_DEV const Ref& operator+=(T v) const { *p = (S)((T)*p + v); return *this; }
	OSC_DEV const Ref& operator-=(T v) const { *p = (S)((T)*p - v); return *this; }
};
template <class T, class S = T> struct Row {
	S* base;
	int64_t stride, r;
	OSC_DEV Ref<T, S> operator()(int e) const { return Ref<T, S>{base + (int64_t)e * stride + r}; }
	OSC_DEV T ro(int e) const { return (T)__ldg(base + (int64_t)e * stride + r); }
};
template <class T> struct Row<T, T> {
	T* base;
	int64_t stride, r;
	OSC_DEV T& operator()(int e) const { return base[(int64_t)e * stride + r]; }
	// read-only load (ld.global.nc): ONLY for rows the running kernel never writes; lets the compiler hoist it over stores
	OSC_DEV T ro(int e) const { return __ldg(base + (int64_t)e * stride + r); }
};
// the S (scratch) row of an owner: the robot's own, or the compact slot of a k_tick_tail thread (Arena::s_stride / s_index)
// (REDIR is a compile-time property of the caller: only k_tick_tail's instantiations of the per-robot functions pay for the selects)
template <class T, bool REDIR, class S> OSC_DEV Row<T, S> srow(const Arena<S>& a, int64_t stride, int64_t r) {
	if constexpr (REDIR) return Row<T, S>{a.S, a.s_stride ? a.s_stride : stride, a.s_index >= 0 ? a.s_index : r};
	else return Row<T, S>{a.S, stride, r};
}
template <class T, int R, int C, class S = T>
OSC_DEV void ld_mat(const Row<T, S>& a, int off, T (&A)[R][C]) {
	OSC_UNROLL for (int i = 0; i < R; ++i)
		OSC_UNROLL for (int j = 0; j < C; ++j) A[i][j] = a(off + i * C + j);
}
template <class T, int R, int C, class S = T>
OSC_DEV void st_mat(const Row<T, S>& a, int off, const T (&A)[R][C]) {
	OSC_UNROLL for (int i = 0; i < R; ++i)
		OSC_UNROLL for (int j = 0; j < C; ++j) a(off + i * C + j) = A[i][j];
}
template <class T, int R, class S = T>
OSC_DEV void ld_vec(const Row<T, S>& a, int off, T (&v)[R]) {
	OSC_UNROLL for (int i = 0; i < R; ++i) v[i] = a(off + i);
}
template <class T, int R, class S = T>
OSC_DEV void st_vec(const Row<T, S>& a, int off, const T (&v)[R]) {
	OSC_UNROLL for (int i = 0; i < R; ++i) a(off + i) = v[i];
}
template <class T, int N, class S = T>
OSC_DEV void ld_kin(const Row<T, S>& Sr, Kin<T, N>& k) {
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		OSC_UNROLL for (int a = 0; a < 3; ++a) {
			k.o[i][a] = Sr(RobotL<N>::KO + 3 * i + a);
			OSC_UNROLL for (int b = 0; b < 3; ++b) k.R[i][a][b] = Sr(RobotL<N>::KR + 9 * i + 3 * a + b);
		}
	}
}

#define OSC_THREADS 128
// resident blocks per SM of the API-faithful kernels: 2 (255 registers).  4 (128 registers) was measured slower (r2j: config 2 per-call
// tick 0.66 -> 0.79 ms, config 3 2.89 -> 3.74 ms): these kernels already spill 1-11 KB at 255 registers
#ifndef OSC_PERCALL_BLOCKS
#define OSC_PERCALL_BLOCKS 2
#endif

// ------------------------------------------------------------------------------------------------------------------
// robot->updateModel()
// ------------------------------------------------------------------------------------------------------------------
template <class T, int N, class S = T, bool REDIR = false>
__device__ void model_robot(const ChainC<T, N>& ch, const Arena<S>& rob, int64_t stride, int64_t r, int dynamics, int gravity) {
	const Row<T, S> P{rob.P, stride, r}, Sr = srow<T, REDIR>(rob, stride, r);
	T q[N];
	ld_vec<T, N>(P, RobotL<N>::Q, q);
	Kin<T, N> kin;
	forward_kinematics<T, N>(ch, q, kin);
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		OSC_UNROLL for (int a = 0; a < 3; ++a) {
			Sr(RobotL<N>::KO + 3 * i + a) = kin.o[i][a];
			OSC_UNROLL for (int b = 0; b < 3; ++b) Sr(RobotL<N>::KR + 9 * i + 3 * a + b) = kin.R[i][a][b];
		}
	}
	if (dynamics) {
		T M[N][N], Minv[N][N];
		mass_matrix<T, N>(ch, kin, M);
		const bool ok = spd_inverse<T, N>(M, Minv);
		st_mat<T, N, N>(Sr, RobotL<N>::M, M);
		st_mat<T, N, N>(Sr, RobotL<N>::MINV, Minv);
		if (!ok) rob.I[(int64_t)RobotL<N>::STATUS * stride + r] |= 16;  // OSC_STATUS_NOT_POSITIVE_DEFINITE
	}
	if (gravity) {
		T g[N];
		gravity_vector<T, N>(ch, kin, g);
		st_vec<T, N>(Sr, RobotL<N>::GRAV, g);
	}
}
template <class T, int N, class S = T>
__global__ void __launch_bounds__(OSC_THREADS, OSC_PERCALL_BLOCKS) k_model(const __grid_constant__ ChainC<T, N> ch, Arena<S> rob, int64_t B, int64_t stride, int dynamics, int gravity) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	model_robot<T, N>(ch, rob, stride, r, dynamics, gravity);
}

// bounded-inertia-estimate inverse: M with every diagonal entry raised to >= threshold, inverted
// (SingularityHandler.cpp:176-182, JointTask.cpp:254-260)
template <class T, int N, class S = T>
OSC_DEV void bie_inverse(const Row<T, S>& robS, T thr, T (&W)[N][N]) {
	T Mb[N][N];
	ld_mat<T, N, N>(robS, RobotL<N>::M, Mb);
	OSC_UNROLL for (int i = 0; i < N; ++i)
		if (Mb[i][i] < thr) Mb[i][i] = thr;
	spd_inverse<T, N>(Mb, W);
}

// Sai2Model::operationalSpaceMatrices on a masked K x N Jacobian: Lambda (masked), N = I - Minv J^T Lambda J
template <class T, int K, int N>
OSC_DEV bool op_space(const T (&Jk)[K][N], const T (&Minv)[N][N], unsigned mask, T (&Lam)[K][K], T (&Nn)[N][N]) {
	T A[K][K];
	congruence<T, K, N>(Jk, Minv, A);
	const bool ok = spd_inverse_masked<T, K>(A, mask, Lam);
	T MJt[N][K];
	mmt<T, N, N, K>(Minv, Jk, MJt);
	T Jbar[N][K];
	mm<T, N, K, K>(MJt, Lam, Jbar);
	OSC_UNROLL for (int i = 0; i < N; ++i)
		OSC_UNROLL for (int j = 0; j < N; ++j) {
			T s = (i == j) ? T(1) : T(0);
			OSC_UNROLL for (int k = 0; k < K; ++k) s = Num<T>::fma_(-Jbar[i][k], Jk[k][j], s);
			Nn[i][j] = s;
		}
	return ok;
}

}  // namespace osc
