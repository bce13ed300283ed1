// osc_fused.cuh -- single-launch control tick:  robot->updateModel(); controller->updateControllerTaskModels();
// controller->computeControlTorques()  (examples/05-using_robot_controller/05-using_robot_controller.cpp:143-148,195;
// src/RobotController.cpp:53-74) for the hierarchy [6-DoF MotionForceTask, full JointTask] (BASELINE configs 1, 2, 5).
//
// k_tick_fused: one robot per thread, everything between q and tau stays in registers.  It serves the robots that are
// in the reference's NON-SINGULAR branch this tick (SingularityHandler.cpp:123-141, alpha = 1); in that branch the SVD
// bases cancel out of every result the tick produces:
//     tau_1 = J^T ( (J W J^T)^-1 F* + F_f )                   W = M^-1 or M_BIE^-1   (.cpp:307-309 with U_ns square)
//     N_1   = I - M^-1 J^T (J M^-1 J^T)^-1 J                                          (.cpp:130-134)
//     JointTask range basis R = orth(range(N_1)) (JointTask.cpp:233): with M = L L^T, Y = L^-1 J^T and u the unit vector
//     orthogonal to range(Y):  N_1 = (L^-T u)(L u)^T,  R = L^-T u / |L^-T u|,  Lambda = 1 / |L^-T u|^2.
// so the two Jacobi SVDs of the reference's route are replaced by (a) a CERTIFICATE that the robot is in the non-singular
// branch -- sigma_0 >= s_abs_tol and lambda_min(J J^T) > s_max^2 * lambda_max, proven with an upper bound of lambda_max
// (tr((JJ^T)^4)^(1/4)) and a Sylvester inertia test -- and (b) Cholesky factorisations.  A robot the certificate cannot
// prove (singular, in the blend band, or within 25 % of it) is appended to a fallback list and gets the API-faithful
// route (full Jacobi SVDs, every branch of SingularityHandler) from k_tick_tail in the same stream.  Results agree with
// the reference route to O(eps * cond(J M^-1 J^T)): the parity tests hold both against the oracle at 1e-9.
#pragma once
#include "osc_aux_kernels.cuh"
#include "osc_host.h"

namespace osc {

#define OSC_FUSED_THREADS 64

enum { OSC_EXPORT_VECTORS = 1, OSC_EXPORT_NO_JT = 2 };  // NO_JT: single-task hierarchy (6-DoF arm, one 6-DoF MotionForceTask): tk1 aliases tk0, never touched

template <class T, int N, int MINB, class S = T>
__global__ void __launch_bounds__(OSC_FUSED_THREADS, MINB) k_tick_fused(const __grid_constant__ ChainC<T, N> ch, const __grid_constant__ MftC<T> tc,
																   const __grid_constant__ JtC<T, N> jc, Arena<S> rob, Arena<S> tk0, Arena<S> tk1,
																   int64_t B, int64_t stride, int stacked, int gravity, int exports,
																   int* __restrict__ fb_count, int* __restrict__ fb_list) {
	static_assert(N == 6 || N == 7, "the fused tick covers n - 6 <= 1 (one-dimensional joint-task range)");
	using ML = MftL<N>;
	using JL = JtL<N>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T, S> rP{rob.P, stride, r}, rS{rob.S, stride, r}, mP{tk0.P, stride, r}, mS{tk0.S, stride, r}, jP{tk1.P, stride, r}, jS{tk1.S, stride, r};
	const Row<int32_t> mI{tk0.I, stride, r};
	const bool ex = (exports & OSC_EXPORT_VECTORS) != 0;
	int status = 0;

	// ---- robot->updateModel(): kinematics, Jacobian at the compliant frame, mass matrix ----
	T q[N];
	ld_vec<T, N>(rP, RobotL<N>::Q, q);
	T J[6][N], x[3], Rc[3][3], M[N][N], grav[N];
	{
		Kin<T, N> kin;
		forward_kinematics<T, N>(ch, q, kin);
		mass_matrix<T, N>(ch, kin, M);
		if (gravity) gravity_vector<T, N>(ch, kin, grav);
		frame_pose<T, N>(kin, tc.body, tc.p_c, tc.R_c, x, Rc);
		point_jacobian<T, N>(ch, kin, tc.body, x, J);
	}

	// ---- the branch certificate on G = J J^T (singular values of Jp = J: N_prec = I) ----
	{
		T G[6][6];
		OSC_UNROLL for (int i = 0; i < 6; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) {
				T s = T(0);
				OSC_UNROLL for (int k = 0; k < N; ++k) s = Num<T>::fma_(J[i][k], J[j][k], s);
				G[i][j] = s; G[j][i] = s;
			}
		T f4 = T(0);  // tr(G^4) = |G^2|_F^2
		OSC_UNROLL for (int i = 0; i < 6; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) {
				T s = T(0);
				OSC_UNROLL for (int k = 0; k < 6; ++k) s = Num<T>::fma_(G[i][k], G[k][j], s);
				f4 = Num<T>::fma_(i == j ? s : T(2) * s, s, f4);
			}
		const T lmax_hi = Num<T>::sqrt_(Num<T>::sqrt_(f4)) * T(1.0000001);  // lambda_max <= lmax_hi <= 6^(1/4) lambda_max
		const T lmax_lo = lmax_hi * T(0.6389431);                            // 6^(-1/4)
		const bool s0_ok = lmax_lo > tc.s_abs_tol * tc.s_abs_tol * T(1.001);
		const bool sep_ok = spd_above<T, 6>(G, tc.s_max * tc.s_max * lmax_hi * T(1.001));
		if (!(s0_ok && sep_ok)) {  // not provably non-singular: API-faithful route
			fb_list[atomicAdd(fb_count, 1)] = (int)r;
			return;
		}
	}

	// ---- M = L L^T;  Y = L^-1 J^T (rows of J whitened);  A = J M^-1 J^T = Y^T Y ----
	Chol<T, N> cm;
	if (!chol_factor<T, N>(M, cm)) status |= 16;
	T Y[6][N];  // Y[c][:] = L^-1 J[c][:]^T
	OSC_UNROLL for (int c = 0; c < 6; ++c) {
		OSC_UNROLL for (int i = 0; i < N; ++i) Y[c][i] = J[c][i];
		chol_fwd<T, N>(cm, Y[c]);
	}
	Chol<T, 6> ca;
	{
		T A[6][6];
		OSC_UNROLL for (int i = 0; i < 6; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) {
				T s = T(0);
				OSC_UNROLL for (int k = 0; k < N; ++k) s = Num<T>::fma_(Y[i][k], Y[j][k], s);
				A[i][j] = s;
			}
		if (!chol_factor<T, 6>(A, ca)) status |= 16;
	}

	// ---- Lambda_ns_modified F* (SingularityHandler.cpp:160-225, 307-309) ----
	T dq[N];
	ld_vec<T, N>(rP, RobotL<N>::DQ, dq);
	T Fstar[6], Ff[6];
	mft_control_law<T, N>(tc, mP, mS, mI, stride, r, J, dq, x, Rc, Fstar, Ff, status, ex);
	T f[6];
	OSC_UNROLL for (int i = 0; i < 6; ++i) f[i] = Fstar[i];
	Chol<T, N> cb;  // factor of M_BIE (bounded inertia estimates), shared with the joint task below
	bool bie_any = false;
	if (tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES || jc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES) {
		// both tasks use the same threshold by default; the (rare) different-threshold case is refused by fused_tick_supported
		const T thr = tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES ? tc.bie_threshold : jc.bie_threshold;
		OSC_UNROLL for (int i = 0; i < N; ++i)
			if (M[i][i] < thr) { M[i][i] = thr; bie_any = true; }
		if (bie_any) chol_factor<T, N>(M, cb);
	}
	if (tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES && bie_any) {
		T Yb[6][N], Ab[6][6];
		OSC_UNROLL for (int c = 0; c < 6; ++c) {
			OSC_UNROLL for (int i = 0; i < N; ++i) Yb[c][i] = J[c][i];
			chol_fwd<T, N>(cb, Yb[c]);
		}
		OSC_UNROLL for (int i = 0; i < 6; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) {
				T s = T(0);
				OSC_UNROLL for (int k = 0; k < N; ++k) s = Num<T>::fma_(Yb[i][k], Yb[j][k], s);
				Ab[i][j] = s;
			}
		Chol<T, 6> cab;
		chol_factor<T, 6>(Ab, cab);
		chol_fwd<T, 6>(cab, f);
		chol_bwd<T, 6>(cab, f);
	} else if (tc.decoupling != OSC_IMPEDANCE) {
		chol_fwd<T, 6>(ca, f);
		chol_bwd<T, 6>(ca, f);
	}
	OSC_UNROLL for (int i = 0; i < 6; ++i) f[i] += Ff[i];
	T tau1[N];
	mtv<T, 6, N>(J, f, tau1);  // tau_1 = J^T (Lambda_mod F* + F_f)

	// ---- classification state of a non-singular robot (SingularityHandler.cpp:237-244): history cleared ----
	// and the q_prior latch (.cpp:233-236).  The latch of a non-singular robot is unobservable by the controller (it is
	// re-latched on entry into a singularity before its first use), so the tau-only mode skips the store; _dq_prior is
	// never read by the reference and is left alone.
	{
		const int ntypes = mI(ML::NTYPES);
		bool latch = ntypes == 0;
		if (ntypes != 0 || mI(ML::HIST_LEN) != 0) {
			latch = latch || mI(ML::C2) > mI(ML::C1);
			mI(ML::NTYPES) = 0; mI(ML::TYPES) = 0; mI(ML::C1) = 0; mI(ML::C2) = 0; mI(ML::HIST_LEN) = 0; mI(ML::HIST_HEAD) = 0;
		}
		if (ex && latch) st_vec<T, N>(mP, ML::Q_PRIOR, q);
	}

	// ---- JointTask in the nullspace (JointTask.cpp:218-356) ----
	T tau[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) tau[i] = tau1[i];
	T tau2[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) tau2[i] = T(0);
	if constexpr (N == 7) {
		// W = C^-1 Y (rows orthonormal); P = I - W^T W is the rank-1 orthogonal projector on range(Y)^perp
		T W[6][N];
		OSC_UNROLL for (int k = 0; k < N; ++k) {
			T col[6];
			OSC_UNROLL for (int c = 0; c < 6; ++c) col[c] = Y[c][k];
			chol_fwd<T, 6>(ca, col);
			OSC_UNROLL for (int c = 0; c < 6; ++c) W[c][k] = col[c];
		}
		// pick the unit vector e_k with the largest residual P_kk = 1 - |W[:,k]|^2
		T best = T(-1);
		int kb = 0;
		OSC_UNROLL for (int k = 0; k < N; ++k) {
			T s = T(1);
			OSC_UNROLL for (int c = 0; c < 6; ++c) s = Num<T>::fma_(-W[c][k], W[c][k], s);
			if (s > best) { best = s; kb = k; }
		}
		T u[N];
		OSC_UNROLL for (int i = 0; i < N; ++i) u[i] = (i == kb) ? T(1) : T(0);
		OSC_UNROLL for (int pass = 0; pass < 2; ++pass) {  // project twice ("twice is enough")
			T wu[6];
			OSC_UNROLL for (int c = 0; c < 6; ++c) {
				T s = T(0);
				OSC_UNROLL for (int i = 0; i < N; ++i) s = Num<T>::fma_(W[c][i], u[i], s);
				wu[c] = s;
			}
			OSC_UNROLL for (int i = 0; i < N; ++i)
				OSC_UNROLL for (int c = 0; c < 6; ++c) u[i] = Num<T>::fma_(-W[c][i], wu[c], u[i]);
		}
		{
			T nu = T(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) nu = Num<T>::fma_(u[i], u[i], nu);
			const T inv = Num<T>::rsqrt_(nu);
			OSC_UNROLL for (int i = 0; i < N; ++i) u[i] *= inv;
		}
		T w[N], v[N];  // N_1 = v w^T with v = L^-T u, w = L u
		chol_mul<T, N>(cm, u, w);
		OSC_UNROLL for (int i = 0; i < N; ++i) v[i] = u[i];
		chol_bwd<T, N>(cm, v);
		T nv2 = T(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) nv2 = Num<T>::fma_(v[i], v[i], nv2);
		const T lam = T(1) / nv2;              // Lambda = (R^T N_1 M^-1 N_1^T R)^-1 = 1 / |v|^2   (|L^-1 w| = |u| = 1)
		T lam_mod = lam;                        // JointTask.cpp:253-267
		if (jc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES && bie_any) {
			T t[N];
			OSC_UNROLL for (int i = 0; i < N; ++i) t[i] = w[i];
			chol_fwd<T, N>(cb, t);
			T s = T(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) s = Num<T>::fma_(t[i], t[i], s);
			lam_mod = T(1) / (nv2 * s);
		} else if (jc.decoupling == OSC_IMPEDANCE) {
			lam_mod = T(1);
		}
		T upd[N], acc[N];
		jt_pd_law<T, N>(jc, jP, jS, q, dq, upd, acc, ex);
		// R^T x = (v . x) / |v|;  tau_2 = N_1^T R s = w |v| s
		T va = T(0), vu = T(0), vc = T(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) { va = Num<T>::fma_(v[i], acc[i], va); vu = Num<T>::fma_(v[i], upd[i], vu); }
		if (stacked) {  // JointTask.cpp:285-292: - Jp^T R Lambda R^T S M^-1 tau_prec
			T t[N];
			OSC_UNROLL for (int i = 0; i < N; ++i) t[i] = tau1[i];
			chol_fwd<T, N>(cm, t);
			chol_bwd<T, N>(cm, t);
			OSC_UNROLL for (int i = 0; i < N; ++i) vc = Num<T>::fma_(v[i], t[i], vc);
		}
		const T sc = lam * (va - vc) + lam_mod * vu;  // the 1/|v| of R^T and the |v| of N_1^T R cancel
		OSC_UNROLL for (int i = 0; i < N; ++i) { tau2[i] = w[i] * sc; tau[i] += tau2[i]; }
	}
	if (ex) {
		if constexpr (N == 7) tk1.I[(int64_t)JL::RANK * stride + r] = 1;
		else tk1.I[(int64_t)JL::RANK * stride + r] = 0;
		st_vec<T, N>(jS, JL::CUR_POS, q);
		st_vec<T, N>(jS, JL::CUR_VEL, dq);
		st_vec<T, N>(jS, JL::TASK_TAU, tau2);
		st_vec<T, N>(mS, ML::TASK_TAU, tau1);
		mS(ML::ALPHA) = T(1);
		mI(ML::INS) = 6;
	}
	if (gravity) {
		OSC_UNROLL for (int i = 0; i < N; ++i) tau[i] += grav[i];
		if (ex) st_vec<T, N>(rS, RobotL<N>::GRAV, grav);
	}
	st_vec<T, N>(rS, RobotL<N>::TAU, tau);
	if (status) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status);
}

// osc_blend.cuh: the singular-blend branch in closed form (tried first by k_tick_tail<.., MB > 0>)
template <class T, int N, int M, class S>
__device__ __forceinline__ bool blend_robot(const ChainC<T, N>& ch, const MftC<T>& tc, const JtC<T, N>& jc, const Arena<S>& rob, const Arena<S>& tk0,
											const Arena<S>& tk1, const Row<T, S>& xm, int64_t stride, int64_t r, int stacked, int gravity, int reuse_law,
											int exports, T* __restrict__ sm);

// The API-faithful chain for one listed robot (every branch of SingularityHandler, full Jacobi SVDs): see k_tick_tail.
template <class T, int N, class S>
__device__ __forceinline__ void tail_chain_robot(const ChainC<T, N>& ch, const MftC<T>& tc, const JtC<T, N>& jc, const Arena<S>& rob, const Arena<S>& tk0,
												  const Arena<S>& tk1, const Arena<S>& xr, const Arena<S>& x0, const Arena<S>& x1, int64_t stride, int64_t r,
												  int stacked, int gravity, int reuse_law, int exports) {
	using ML = MftL<N>;
	using JL = JtL<N>;
	const bool redirected = xr.s_stride != 0;
	const Row<T, S> rS{rob.S, stride, r}, mS{tk0.S, stride, r}, jS{tk1.S, stride, r};                       // the robot's own rows
	const Row<T, S> xS = srow<T, true>(xr, stride, r), xm = srow<T, true>(x0, stride, r), xj = srow<T, true>(x1, stride, r);   // where the chain works
	if (redirected) {
		if (reuse_law & 1) {  // F*, F_f handed over by the fused kernel
			for (int i = 0; i < 6; ++i) { xm(ML::UMF + i) = (T)mS(ML::UMF + i); xm(ML::FFB + i) = (T)mS(ML::FFB + i); }
		}
		if (gravity) for (int i = 0; i < N; ++i) xS(RobotL<N>::GRAV + i) = (T)rS(RobotL<N>::GRAV + i);
	}
	const bool has_jt = !(exports & OSC_EXPORT_NO_JT);
	model_robot<T, N, S, true>(ch, xr, stride, r, 1, gravity);
	mft_update_robot<T, N, S, true>(ch, tc, xr, x0, (const S*)nullptr, stride, r, 0);
	if (has_jt) jt_update_robot<T, N, S, true>(jc, xr, x1, x0.S + (int64_t)ML::NTP * (x0.s_stride ? x0.s_stride : stride), stride, r, x0.s_stride, x0.s_index);
	mft_compute_robot<T, N, S, true>(ch, tc, xr, x0, stride, r, 0, 1, reuse_law & 1);
	if (has_jt) jt_compute_robot<T, N, S, true>(jc, xr, x1, stride, r, stacked, 2, reuse_law & 2);
	if (gravity) {
		for (int i = 0; i < N; ++i) xS(RobotL<N>::TAU + i) += (T)xS(RobotL<N>::GRAV + i);
	}
	if (redirected) {
		for (int i = 0; i < N; ++i) rS(RobotL<N>::TAU + i) = (T)xS(RobotL<N>::TAU + i);
		if (exports & OSC_EXPORT_VECTORS) {  // vector getters of the LAZY output mode (leading rows of the task arenas, osc_layout.cuh): the task
			// torques and alpha always; the rows of a control law only when this kernel ran that law (else the fused kernel
			// already wrote them at the robot's own rows)
			for (int e = 0; e < N; ++e) mS(ML::TASK_TAU + e) = (T)xm(ML::TASK_TAU + e);
			if (has_jt) for (int e = 0; e < N; ++e) jS(JL::TASK_TAU + e) = (T)xj(JL::TASK_TAU + e);
			mS(ML::ALPHA) = (T)xm(ML::ALPHA);
			if (!(reuse_law & 1)) for (int e = N; e < ML::J; ++e) mS(e) = (T)xm(e);
			if (has_jt && !(reuse_law & 2)) for (int e = N; e < JL::NT; ++e) jS(e) = (T)xj(e);
		}
	}
}
// out of line for the kernels that try the closed-form blend first (MB > 0): the chain is their rare second resort, and one copy
// of its 70 k instructions serves every MB
template <class T, int N, class S>
__device__ __noinline__ void tail_chain_robot_call(const ChainC<T, N>& ch, const MftC<T>& tc, const JtC<T, N>& jc, const Arena<S>& rob, const Arena<S>& tk0,
													const Arena<S>& tk1, const Arena<S>& xr, const Arena<S>& x0, const Arena<S>& x1, int64_t stride, int64_t r,
													int stacked, int gravity, int reuse_law, int exports) {
	tail_chain_robot<T, N, S>(ch, tc, jc, rob, tk0, tk1, xr, x0, x1, stride, r, stacked, gravity, reuse_law, exports);
}

// The robots the fused kernels could not certify as non-singular, one per thread over the compacted list.  MB > 0 (n = 7, task rank
// MB): the closed-form singular-blend routine of osc_blend.cuh first; what it declines -- and everything when MB == 0 -- takes the
// API-faithful chain.  Also re-arms the other fallback counter for the next tick.
template <class T, int N, class S = T, int MB = 0>
__global__ void __launch_bounds__(OSC_THREADS) k_tick_tail(const __grid_constant__ ChainC<T, N> ch, const __grid_constant__ MftC<T> tc,
															const __grid_constant__ JtC<T, N> jc, Arena<S> rob, Arena<S> tk0, Arena<S> tk1, int64_t stride,
															int stacked, int gravity, const int* fb_count, int* fb_count_next, const int* fb_list, int reuse_law,
															S* scratch, int64_t scratch_stride, int exports) {
	extern __shared__ __align__(16) unsigned char osc_stash_raw[];
	// launched with programmatic stream serialisation behind the fused kernel: its launch latency overlaps the fused kernel's
	// run; the fallback list is complete (and visible) once the wait returns.  A no-op under a plain launch.
	asm volatile("griddepcontrol.wait;" ::: "memory");
	const int n = *fb_count;
	if (blockIdx.x == 0 && threadIdx.x == 0) *fb_count_next = 0;
	using ML = MftL<N>;
	const int64_t slot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	// The listed robots are scattered over the batch.  With a scratch arena (one slot per tail thread) every S row of the
	// API-faithful chain below lives at [row][slot]: consecutive threads, consecutive addresses.  What the rest of the library reads
	// at the robot's own rows -- tau, and in the LAZY output mode the vector getters -- is copied out at the end.
	Arena<S> xr = rob, x0 = tk0, x1 = tk1;
	if (scratch && slot < scratch_stride) {
		xr.S = scratch; x0.S = scratch + (int64_t)RobotL<N>::S_SIZE * scratch_stride; x1.S = x0.S + (int64_t)ML::S_SIZE * scratch_stride;
		xr.s_stride = x0.s_stride = x1.s_stride = scratch_stride;
		xr.s_index = x0.s_index = x1.s_index = slot;
	}
	for (int idx = (int)slot; idx < n; idx += gridDim.x * blockDim.x) {
		const int64_t r = fb_list[idx];
		if constexpr (MB > 0) {
			const Row<T, S> xm = srow<T, true>(x0, stride, r);
			if (blend_robot<T, N, MB, S>(ch, tc, jc, rob, tk0, tk1, xm, stride, r, stacked, gravity, reuse_law, exports, reinterpret_cast<T*>(osc_stash_raw))) continue;
			tail_chain_robot_call<T, N, S>(ch, tc, jc, rob, tk0, tk1, xr, x0, x1, stride, r, stacked, gravity, reuse_law, exports);
		} else {
			tail_chain_robot<T, N, S>(ch, tc, jc, rob, tk0, tk1, xr, x0, x1, stride, r, stacked, gravity, reuse_law, exports);
		}
	}
}

// ---- [full JointTask] alone (examples/01-joint_control, 02-joint_control_internal_otg) -------------------------------------------
// N_prec = I, so the projected Jacobian and its range basis are the identity, Lambda = (M^-1)^-1 = M and Lambda_mod = M_BIE, M or I
// (JointTask.cpp:218-283); tau = M a_des + Lambda_mod u (:342-355): the mass matrix and two products, no factorisation at all.
// Any serial chain (prismatic joints included), one robot per thread.
template <class T, int N>
__global__ void __launch_bounds__(OSC_THREADS) k_tick_joint(const __grid_constant__ ChainC<T, N> ch, const __grid_constant__ JtC<T, N> jc, Arena<T> rob, Arena<T> tk,
															 int64_t B, int64_t stride, int gravity, int exports) {
	using JL = JtL<N>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T> rP{rob.P, stride, r}, rS{rob.S, stride, r}, jP{tk.P, stride, r}, jS{tk.S, stride, r};
	const bool ex = (exports & OSC_EXPORT_VECTORS) != 0;
	T q[N], dq[N], u[N], acc[N];
	ld_vec<T, N>(rP, RobotL<N>::Q, q);
	ld_vec<T, N>(rP, RobotL<N>::DQ, dq);
	jt_pd_law<T, N>(jc, jP, jS, q, dq, u, acc, ex);
	Kin<T, N> kin;
	forward_kinematics<T, N>(ch, q, kin);
	T M[N][N], tau[N];
	mass_matrix<T, N>(ch, kin, M);
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		T s = T(0);
		if (jc.decoupling == OSC_IMPEDANCE) {
			OSC_UNROLL for (int j = 0; j < N; ++j) s = Num<T>::fma_(M[i][j], acc[j], s);
			s += u[i];
		} else {
			OSC_UNROLL for (int j = 0; j < N; ++j) s = Num<T>::fma_(M[i][j], acc[j] + u[j], s);
			if (jc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES && M[i][i] < jc.bie_threshold) s = Num<T>::fma_(jc.bie_threshold - M[i][i], u[i], s);  // :253-267
		}
		tau[i] = s;
	}
	if (ex) {
		st_vec<T, N>(jS, JL::TASK_TAU, tau);
		st_vec<T, N>(jS, JL::CUR_POS, q);
		st_vec<T, N>(jS, JL::CUR_VEL, dq);
		tk.I[(int64_t)JL::RANK * stride + r] = N;
	}
	if (gravity) {  // RobotController.cpp:70-72
		T g[N];
		gravity_vector<T, N>(ch, kin, g);
		OSC_UNROLL for (int i = 0; i < N; ++i) tau[i] += g[i];
	}
	st_vec<T, N>(rS, RobotL<N>::TAU, tau);
}
inline bool joint_only_hierarchy(const Batch& b) { return b.tasks.size() == 1 && b.tasks[0].type == OSC_TASK_JOINT && b.tasks[0].jd.full; }

// hierarchy / parameter combinations the fused tick serves (everything else: the kernel sequence of osc_capi_tick.inl)
template <class T, int N> bool fused_tick_supported(Batch& b) {
	if (b.tick_outputs != OSC_TICK_OUTPUTS_ALL && joint_only_hierarchy(b)) return true;
	if constexpr (N != 6 && N != 7) return false;
	if (b.tick_outputs == OSC_TICK_OUTPUTS_ALL) return false;
	if (b.tasks.empty() || b.tasks[0].type != OSC_TASK_MOTION_FORCE) return false;
	const TaskH& m = b.tasks[0];
	if (b.tasks.size() == 1) {
		// single-task hierarchy, second-generation kernel only: a 6-DoF arm under one full MotionForceTask (examples/19 on the PUMA; BASELINE
		// config 4), or a 7-DoF arm whose RobotController holds that one task (examples/07-surface_surface_contact on the Panda)
		if (!((N == 6 || N == 7) && !m.md.partial && m.dof == 6)) return false;
		if (m.md.body != b.n - 1) return false;
		for (int i = 0; i < b.n; ++i)
			if (b.chain.joint_type[i] != OSC_JOINT_REVOLUTE) return false;
	} else {
		if (b.tasks.size() != 2) return false;
		const TaskH& j = b.tasks[1];
		if (j.type != OSC_TASK_JOINT) return false;
		if (!j.jd.full) return false;
		if (m.md.partial) {  // partial task: k_tick_fused3, instantiated for the (n, task rank) pairs of OSC_FUSED3_SUPPORTED (osc_fused3.cuh)
			if (!(N == 7 && (m.dof == 3 || m.dof == 2 || m.dof == 4 || m.dof == 5))) return false;
			if (b.tasks[0].md.body != b.n - 1) return false;
			for (int i = 0; i < b.n; ++i)
				if (b.chain.joint_type[i] != OSC_JOINT_REVOLUTE) return false;
		} else if (m.dof != 6) return false;
		if (m.mp.dynamic_decoupling_type == OSC_BOUNDED_INERTIA_ESTIMATES && j.jp.dynamic_decoupling_type == OSC_BOUNDED_INERTIA_ESTIMATES &&
			m.mp.bie_threshold != j.jp.bie_threshold)
			return false;
	}
	if (!(m.mp.s_max > 0) || !(m.mp.s_abs_tol > 0)) return false;
	// The branch certificate of the fused kernels is an FP32 bound on G = J J^T (absolute error ~ 4e-7 lambda_max): its shift
	// s_max^2 lambda_max must stay far above that noise, or a truly singular robot could be certified non-singular
	// (ADVICE r1).  Below s_max = 1e-2 (default 6e-2, setSingularityHandlingBounds) the tick takes the API-faithful kernels.
	if (m.mp.s_max < 1e-2) return false;
	return true;
}

}  // namespace osc
