// osc_aux_kernels.cuh -- cold-path kernels: task construction / reInitializeTask / parameter side effects / sensor
// wrench transform / packed getters.
#pragma once
#include "osc_jt.cuh"
#include "osc_mft_torque.cuh"

namespace osc {

// MotionForceTask::initialSetup (handler part, SingularityHandler.cpp:53-63) + POPC reInitialize
// (POPCExplicitForceControl.cpp:10-21)
template <class T, int N, class S = T>
__global__ void k_mft_init(const __grid_constant__ ChainC<T, N> ch, Arena<S> tk, int64_t B, int64_t stride) {
	using L = MftL<N>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T, S> tP{tk.P, stride, r};
	const Row<int32_t> tI{tk.I, stride, r};
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		tP(L::Q_PRIOR + i) = T(0.5) * (ch.q_lower[i] + ch.q_upper[i]);
		tP(L::DQ_PRIOR + i) = T(0);
		tP(L::T2DIR + i) = T(1);
	}
	for (int i = 0; i < L::I_SIZE; ++i) tI(i) = 0;
	tI(L::POPC_COUNTER) = 50;
	tP(L::POPC_PO) = T(0); tP(L::POPC_ECORR) = T(0); tP(L::POPC_RC) = T(1); tP(L::POPC_VSQ) = T(0);
}

// MotionForceTask::reInitializeTask (MotionForceTask.cpp:204-245)
template <class T, int N, class S = T>
__global__ void k_mft_reinit(const __grid_constant__ MftC<T> tc, Arena<S> rob, Arena<S> tk, int64_t B, int64_t stride) {
	using L = MftL<N>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T, S> rS{rob.S, stride, r}, tP{tk.P, stride, r}, tS{tk.S, stride, r};
	Kin<T, N> kin;
	ld_kin<T, N>(rS, kin);
	T x[3], R[3][3];
	frame_pose<T, N>(kin, tc.body, tc.p_c, tc.R_c, x, R);
	for (int e = L::GOAL_POS; e < L::POPC_PO; ++e) tP(e) = T(0);  // goals, sensed wrench, caches, integrators
	st_vec<T, 3>(tP, L::GOAL_POS, x);
	st_vec<T, 3>(tP, L::CUR_POS, x);
	st_mat<T, 3, 3>(tP, L::GOAL_ORI, R);
	st_mat<T, 3, 3>(tP, L::CUR_ORI, R);
	for (int e = L::TASK_TAU; e < L::J; ++e) tS(e) = T(0);
	st_vec<T, 3>(tS, L::DES_POS, x);
	st_mat<T, 3, 3>(tS, L::DES_ORI, R);
}

// side effects of parametrizeForceMotionSpaces / parametrizeMomentRotMotionSpaces (MotionForceTask.cpp:850-856, 882-888),
// setClosedLoopForce/MomentControl (.cpp:973-986) and disablePassivity (POPCExplicitForceControl.cpp:25-28)
template <class T, int N, class S = T>
__global__ void k_mft_reset(Arena<S> tk, int64_t B, int64_t stride, int flags) {
	using L = MftL<N>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T, S> tP{tk.P, stride, r};
	const Row<int32_t> tI{tk.I, stride, r};
	if (flags & 1) {
		for (int a = 0; a < 3; ++a) { tP(L::GOAL_POS + a) = tP(L::CUR_POS + a); tP(L::GOAL_LINVEL + a) = T(0); tP(L::GOAL_LINACC + a) = T(0); }
	}
	if (flags & (1 | 8)) {
		for (int a = 0; a < 3; ++a) { tP(L::INT_POS + a) = T(0); tP(L::INT_F + a) = T(0); }
	}
	if (flags & 2) {
		for (int a = 0; a < 9; ++a) tP(L::GOAL_ORI + a) = tP(L::CUR_ORI + a);
		for (int a = 0; a < 3; ++a) { tP(L::GOAL_ANGVEL + a) = T(0); tP(L::GOAL_ANGACC + a) = T(0); }
	}
	if (flags & (2 | 16)) {
		for (int a = 0; a < 3; ++a) { tP(L::INT_ORI + a) = T(0); tP(L::INT_M + a) = T(0); }
	}
	if (flags & 4) {
		tP(L::POPC_PO) = T(0); tP(L::POPC_ECORR) = T(0); tP(L::POPC_RC) = T(1); tP(L::POPC_VSQ) = T(0);
		tI(L::POPC_COUNTER) = 50; tI(L::POPC_HEAD) = 0; tI(L::POPC_LEN) = 0;
	}
}

// JointTask::reInitializeTask (JointTask.cpp:91-107)
template <class T, int N, class S = T>
__global__ void k_jt_reinit(const __grid_constant__ JtC<T, N> jc, Arena<S> rob, Arena<S> tk, int64_t B, int64_t stride, int init) {
	using L = JtL<N>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T, S> rP{rob.P, stride, r}, tP{tk.P, stride, r}, tS{tk.S, stride, r};
	T q[N], cp[N];
	ld_vec<T, N>(rP, RobotL<N>::Q, q);
	mv<T, N, N>(jc.S, q, cp);
	for (int e = 0; e < L::P_SIZE; ++e) tP(e) = T(0);
	st_vec<T, N>(tP, L::GOAL_POS, cp);
	st_vec<T, N>(tS, L::CUR_POS, cp);
	for (int i = 0; i < N; ++i) { tS(L::CUR_VEL + i) = T(0); tS(L::DES_VEL + i) = T(0); tS(L::TASK_TAU + i) = T(0); }
	if (init) {  // JointTask.cpp:63-68
		tk.I[(int64_t)L::RANK * stride + r] = jc.k;
		for (int i = 0; i < N; ++i)
			for (int j = 0; j < N; ++j) {
				const T id = (i == j && i < jc.k) ? T(1) : T(0);
				tS(L::RANGE + i * N + j) = id; tS(L::LAM + i * N + j) = id; tS(L::LAM_MOD + i * N + j) = id;
				tS(L::NPREC + i * N + j) = (i == j) ? T(1) : T(0);
				tS(L::NT + i * N + j) = T(0); tS(L::NTP + i * N + j) = T(0);
			}
	}
}

// MotionForceTask::updateSensedForceAndMoment (MotionForceTask.cpp:805-828); in6 = SoA [6][stride] sensor-frame wrench
template <class T, int N, class S = T>
__global__ void k_mft_sensed(const __grid_constant__ ChainC<T, N> ch, const __grid_constant__ MftC<T> tc, Arena<S> rob, Arena<S> tk, const S* in6,
							 int64_t B, int64_t stride, int kinematics_from_q) {
	using L = MftL<N>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T, S> rP{rob.P, stride, r}, rS{rob.S, stride, r}, tP{tk.P, stride, r};
	Kin<T, N> kin;
	if (kinematics_from_q) {
		// after a fused tick the kinematics rows are older than q (the fused kernels keep them on chip): the compliant frame is
		// evaluated here from the current q, which is what robot->updateModel() before updateSensedForceAndMoment gives the
		// reference (examples/07-surface_surface_contact order of calls) -- without re-running the model and task-model kernels
		T q[N];
		ld_vec<T, N>(rP, RobotL<N>::Q, q);
		forward_kinematics<T, N>(ch, q, kin);
	} else {
		ld_kin<T, N>(rS, kin);
	}
	T x[3], R[3][3];
	frame_pose<T, N>(kin, tc.body, tc.p_c, tc.R_c, x, R);
	T fs[3], ms[3], f[3], m[3], pxf[3], fw[3], mw[3];
	for (int a = 0; a < 3; ++a) { fs[a] = in6[(int64_t)a * stride + r]; ms[a] = in6[(int64_t)(3 + a) * stride + r]; }
	mv3(tc.sensor_R, fs, f);
	mv3(tc.sensor_R, ms, m);
	cross3(tc.sensor_p, f, pxf);
	for (int a = 0; a < 3; ++a) m[a] += pxf[a];
	mv3(R, f, fw);
	mv3(R, m, mw);
	st_vec<T, 3>(tP, L::SENSED_F, fw);
	st_vec<T, 3>(tP, L::SENSED_M, mw);
}

template <class T, int N, class S = T>
__global__ void k_add_gravity(Arena<S> rob, int64_t B, int64_t stride) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T, S> rS{rob.S, stride, r};
	for (int i = 0; i < N; ++i) rS(RobotL<N>::TAU + i) += rS(RobotL<N>::GRAV + i);
}

// out[(i*C + j)] = sum_ab U[i][a] L[a][b] U[j][b] : the basis-independent export of a Lambda stored in a range basis
template <class T, int R, int K, class St = T>
__global__ void k_pack_ulut(const St* S, int offU, int offL, St* out, int64_t B, int64_t stride, int out_dim) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T, St> s{const_cast<St*>(S), stride, r};
	T U[R][K], Lm[K][K], UL[R][K];
	ld_mat<T, R, K>(s, offU, U);
	ld_mat<T, K, K>(s, offL, Lm);
	mm<T, R, K, K>(U, Lm, UL);
	for (int i = 0; i < out_dim; ++i)
		for (int j = 0; j < out_dim; ++j) {
			T v = T(0);
			for (int k = 0; k < K; ++k) v += UL[i][k] * U[j][k];
			out[(int64_t)(i * out_dim + j) * stride + r] = (St)v;
		}
}
// singularity types unpack / counters
// Getters of MotionForceTask that are functions of the task state rather than stored members (MotionForceTask.h:268-269,
// 611-614; .cpp:540-578, 755-769, 892-971): the four Sigma selection matrices, getPositionError / getOrientationError, the goal
// force / moment in the world frame, and the two error norms goalPositionReached / goalOrientationReached compare with their
// tolerance.  The compliant-frame rotation is _current_orientation of the last computeTorques (or reInitializeTask).
//   what: 0-3 sigmaForce / sigmaPosition / sigmaMoment / sigmaOrientation (9)   4 position error (3)   5 orientation error (3)
//         6 goal force (3)   7 goal moment (3)   8 position error norm (1)   9 orientation error norm (1)
template <class T, int N, class St = T>
__global__ void k_pack_mft_getters(const __grid_constant__ MftC<T> tc, const St* P, const St* Sc, St* out, int64_t B, int64_t stride, int what) {
	using L = MftL<N>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T, St> tP{const_cast<St*>(P), stride, r}, tS{const_cast<St*>(Sc), stride, r}, o{out, stride, r};
	T R[3][3], rot[3][3];
	ld_mat<T, 3, 3>(tP, L::CUR_ORI, R);
	OSC_UNROLL for (int i = 0; i < 3; ++i)
		OSC_UNROLL for (int j = 0; j < 3; ++j) rot[i][j] = tc.in_compliant_frame ? R[i][j] : (i == j ? T(1) : T(0));
	T Sf[3][3], Sm[3][3], Sp[3][3], So[3][3];
	sigma_matrix<T>(tc.force_dim, tc.Pt, rot, tc.force_axis, Sf);
	sigma_matrix<T>(tc.moment_dim, tc.Pr, rot, tc.moment_axis, Sm);
	sigma_complement<T>(tc.Pt, Sf, Sp);
	sigma_complement<T>(tc.Pr, Sm, So);
	if (what < 4) {
		const T (&M)[3][3] = what == 0 ? Sf : (what == 1 ? Sp : (what == 2 ? Sm : So));
		OSC_UNROLL for (int i = 0; i < 3; ++i)
			OSC_UNROLL for (int j = 0; j < 3; ++j) o(3 * i + j) = M[i][j];
		return;
	}
	T v[3], res[3];
	if (what == 4 || what == 8) {  // sigmaPosition() * (_goal_position - _current_position), .cpp:540-542, 548-553
		OSC_UNROLL for (int a = 0; a < 3; ++a) v[a] = (T)tP(L::GOAL_POS + a) - (T)tP(L::CUR_POS + a);
		mv3(Sp, v, res);
	} else if (what == 5 || what == 9) {  // sigmaOrientation() * _orientation_error, .cpp:544-546, 565-569
		ld_vec<T, 3>(tS, L::ORI_ERR, v);
		mv3(So, v, res);
	} else {  // getGoalForce / getGoalMoment, .cpp:755-769
		ld_vec<T, 3>(tP, what == 6 ? L::GOAL_FORCE : L::GOAL_MOMENT, v);
		mv3(rot, v, res);
	}
	if (what >= 8) {
		const T q = dot3(v, res);
		o(0) = Num<T>::sqrt_(q > T(0) ? q : T(0));
	} else {
		OSC_UNROLL for (int a = 0; a < 3; ++a) o(a) = res[a];
	}
}
// JointTask::goalPositionReached (JointTask.cpp:437-446): | R R^T (q_task - q_goal) | with R = _current_task_range
template <class T, int N, class St = T>
__global__ void k_pack_jt_error_norm(const St* P, const St* Sc, const int32_t* I, St* out, int64_t B, int64_t stride, int k) {
	using L = JtL<N>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T, St> tP{const_cast<St*>(P), stride, r}, tS{const_cast<St*>(Sc), stride, r};
	const int rank = I[(int64_t)L::RANK * stride + r];
	T e[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) e[i] = i < k ? (T)tS(L::CUR_POS + i) - (T)tP(L::GOAL_POS + i) : T(0);
	T s2 = T(0);
	for (int c = 0; c < rank; ++c) {  // range basis stored k x rank, row-major N x N padded
		T d = T(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) d = Num<T>::fma_((T)tS(L::RANGE + i * N + c), e[i], d);
		s2 = Num<T>::fma_(d, d, s2);
	}
	out[r] = (St)Num<T>::sqrt_(s2);
}
template <int N>
__global__ void k_pack_types(const int32_t* I, int32_t* out, int64_t B, int64_t stride, int what) {
	using L = MftL<N>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	if (what == 0) {
		const int n = I[(int64_t)L::NTYPES * stride + r], t = I[(int64_t)L::TYPES * stride + r];
		for (int c = 0; c < 6; ++c) out[(int64_t)c * stride + r] = c < n ? ((t >> (2 * c)) & 3) : 0;
	} else {
		out[r] = I[(int64_t)L::C1 * stride + r];
		out[stride + r] = I[(int64_t)L::C2 * stride + r];
	}
}

}  // namespace osc
