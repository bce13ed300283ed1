// osc_mft_torque.cuh -- SingularityHandler::computeTorques (src/tasks/SingularityHandler.cpp:297-368) and the
// k_mft_compute kernel (MotionForceTask::computeTorques(), ::computeTorques(tau_prec), MotionForceTask.cpp:270-509).
#pragma once
#include "osc_mft_compute.cuh"

namespace osc {

// J_x^T (L_x U_x^T F* + U_x^T F_f) for a stored (U, J, Lambda_modified) triple
template <class T, int N, int K, class S = T>
OSC_DEV void range_torque(const Row<T, S>& tS, int offU, int offJ, int offL, const T (&Fstar)[6], const T (&Ff)[6], bool use_lambda, T (&tau)[N]) {
	T U[6][K], Jx[K][N];
	ld_mat<T, 6, K>(tS, offU, U);
	ld_mat<T, K, N>(tS, offJ, Jx);
	T a[K], b[K], f[K];
	mtv<T, 6, K>(U, Fstar, a);
	mtv<T, 6, K>(U, Ff, b);
	if (use_lambda) {
		T Lm[K][K];
		ld_mat<T, K, K>(tS, offL, Lm);
		mv<T, K, K>(Lm, a, f);
	} else {
		OSC_UNROLL for (int i = 0; i < K; ++i) f[i] = a[i];
	}
	OSC_UNROLL for (int i = 0; i < K; ++i) f[i] += b[i];
	mtv<T, K, N>(Jx, f, tau);
}

template <class T, int N, class S = T, bool REDIR = false>
__device__ void mft_compute_robot(const ChainC<T, N>& ch, const MftC<T>& tc, const Arena<S>& rob, const Arena<S>& tk, int64_t stride, int64_t r,
								  int with_tau_prec, int accumulate, int reuse_law = 0) {
	using L = MftL<N>;
	constexpr int K = L::K;
	const Row<T, S> rP{rob.P, stride, r}, rS = srow<T, REDIR>(rob, stride, r), tP{tk.P, stride, r}, tS = srow<T, REDIR>(tk, stride, r);
	const Row<int32_t> tI{tk.I, stride, r};
	int status = 0;

	// J, Jp, x, R from the CURRENT kinematics (MotionForceTask.cpp:280-298)
	Kin<T, N> kin;
	ld_kin<T, N>(rS, kin);
	T x[3], R[3][3];
	frame_pose<T, N>(kin, tc.body, tc.p_c, tc.R_c, x, R);
	T J[6][N];
	task_jacobian<T, N>(ch, tc, kin, x, J);
	{
		T Np[N][N], Jp[6][N];
		ld_mat<T, N, N>(tS, L::NPREC, Np);
		mm<T, 6, N, N>(J, Np, Jp);
		st_mat<T, 6, N>(tS, L::J, J);
		st_mat<T, 6, N>(tS, L::JP, Jp);
	}
	T q[N], dq[N];
	ld_vec<T, N>(rP, RobotL<N>::Q, q);
	ld_vec<T, N>(rP, RobotL<N>::DQ, dq);
	T Fstar[6], Ff[6];
	if (reuse_law) {  // k_tick_tail: the fused kernel already ran the control law (and advanced its state) for this robot
		ld_vec<T, 6>(tS, L::UMF, Fstar);
		ld_vec<T, 6>(tS, L::FFB, Ff);
	} else {
		mft_control_law<T, N>(tc, tP, tS, tI, stride, r, J, dq, x, R, Fstar, Ff, status);
	}

	// ---- SingularityHandler::computeTorques ----
	T tau[N];
	const int ntypes = tI(L::NTYPES);
	const int ins = tI(L::INS);
	if (ntypes == 0) {  // .cpp:307-309
		range_torque<T, N, K>(tS, L::U_NS, L::J_NS, L::L_NS_MOD, Fstar, Ff, true, tau);
	} else if (tc.decoupling == 2) {  // .cpp:310-312
		range_torque<T, N, K>(tS, L::U_NS, L::J_NS, L::L_NS_MOD, Fstar, Ff, false, tau);
	} else if (ins == 0) {  // fully singular: pass through (.cpp:317-318)
		OSC_UNROLL for (int i = 0; i < N; ++i) tau[i] = T(0);
	} else {
		range_torque<T, N, K>(tS, L::U_NS, L::J_NS, L::L_NS_MOD, Fstar, Ff, true, tau);
		if (tc.handling) {
			const T alpha = tS(L::ALPHA);
			T Vs[N][K], Jpost[K][N], Ljm[K][K];
			ld_mat<T, N, K>(tS, L::V_S, Vs);
			ld_mat<T, K, N>(tS, L::J_POST, Jpost);
			ld_mat<T, K, K>(tS, L::L_JS_MOD, Ljm);
			T tj[N];
			const int c1 = tI(L::C1), c2 = tI(L::C2);
			if (c1 > c2 || tc.enforce_type_1) {  // .cpp:328-331
				T unit[N], a[K], f[K];
				OSC_UNROLL for (int i = 0; i < N; ++i) unit[i] = -tc.kp1 * (q[i] - tP(L::Q_PRIOR + i)) - tc.kv1 * dq[i];
				mtv<T, N, K>(Vs, unit, a);
				mv<T, K, K>(Ljm, a, f);
				mtv<T, K, N>(Jpost, f, tj);
			} else {  // .cpp:332-351
				T Us[6][K];
				ld_mat<T, 6, K>(tS, L::U_S, Us);
				T nf = T(0);
				OSC_UNROLL for (int k = 0; k < 6; ++k) { const T f = Fstar[k] + Ff[k]; nf = Num<T>::fma_(f, f, nf); }
				nf = Num<T>::sqrt_(nf);
				// first singular column = column `ins` of the padded arrays
				T fTd = T(0);
				T v0[N];
				OSC_UNROLL for (int i = 0; i < N; ++i) v0[i] = T(0);
				OSC_UNROLL for (int c = 0; c < K; ++c) {
					if (c == ins) {
						OSC_UNROLL for (int k = 0; k < 6; ++k) { const T f = Fstar[k] + Ff[k]; fTd += (nf > T(0) ? f / nf : f) * Us[k][c]; }
						OSC_UNROLL for (int i = 0; i < N; ++i) v0[i] = Vs[i][c];
					}
				}
				T unit[N], damp[N];
				OSC_UNROLL for (int i = 0; i < N; ++i) {
					T dir = tP(L::T2DIR + i);
					if (v0[i] != T(0)) {
						if (Num<T>::abs_(q[i] - ch.q_upper[i]) < tc.type_2_angle_threshold) dir = T(-1);
						else if (Num<T>::abs_(q[i] - ch.q_lower[i]) < tc.type_2_angle_threshold) dir = T(1);
						tP(L::T2DIR + i) = dir;
					}
					unit[i] = dir * (Num<T>::abs_(fTd) * tc.type_2_torque_ratio * ch.effort[i]);  // quirk Q5: intended ratio
					damp[i] = -tc.kv2 * dq[i];
				}
				T a[K], b[K], f[K];
				mtv<T, N, K>(Vs, unit, a);
				mtv<T, N, K>(Vs, damp, b);
				mv<T, K, K>(Ljm, b, f);
				OSC_UNROLL for (int i = 0; i < K; ++i) f[i] += a[i];
				mtv<T, K, N>(Jpost, f, tj);
			}
			T ts[N];
			range_torque<T, N, K>(tS, L::U_S, L::J_S, L::L_S_MOD, Fstar, Ff, true, ts);
			OSC_UNROLL for (int i = 0; i < N; ++i) {  // .cpp:357-365
				if (ts[i] != ts[i]) { ts[i] = T(0); status |= 1; }
				else if (ts[i] > ch.effort[i]) ts[i] = ch.effort[i];
				else if (ts[i] < -ch.effort[i]) ts[i] = -ch.effort[i];
				tau[i] = tau[i] + alpha * ts[i] + (T(1) - alpha) * tj[i];
			}
		}
	}
	// computeTorques(tau_prec): the reference's compensation multiplies by a _Lambda that is identically zero (quirk Q1);
	// the intended term (world-frame Lambda) is available behind a non-default flag.
	if (with_tau_prec && tc.apply_compensation && ins > 0) {
		T tp[N], Minv[N][N], Mt[N], Jt[6], Uns[6][K], Lns[K][K], a[K], b[K], c6[6], Jp[6][N], comp[N];
		ld_vec<T, N>(rS, RobotL<N>::TAU, tp);
		ld_mat<T, N, N>(rS, RobotL<N>::MINV, Minv);
		mv<T, N, N>(Minv, tp, Mt);
		mv<T, 6, N>(J, Mt, Jt);
		ld_mat<T, 6, K>(tS, L::U_NS, Uns);
		ld_mat<T, K, K>(tS, L::L_NS, Lns);
		mtv<T, 6, K>(Uns, Jt, a);
		mv<T, K, K>(Lns, a, b);
		mv<T, 6, K>(Uns, b, c6);
		ld_mat<T, 6, N>(tS, L::JP, Jp);
		mtv<T, 6, N>(Jp, c6, comp);
		OSC_UNROLL for (int i = 0; i < N; ++i) tau[i] -= comp[i];
	}
	st_vec<T, N>(tS, L::TASK_TAU, tau);
	if (accumulate == 1) {  // first task of the hierarchy: tau = task torque
		st_vec<T, N>(rS, RobotL<N>::TAU, tau);
	} else if (accumulate == 2) {
		OSC_UNROLL for (int i = 0; i < N; ++i) rS(RobotL<N>::TAU + i) += tau[i];
	}
	if (status) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status);
}

template <class T, int N, class S = T>
__global__ void __launch_bounds__(OSC_THREADS, OSC_PERCALL_BLOCKS) k_mft_compute(const __grid_constant__ ChainC<T, N> ch, const __grid_constant__ MftC<T> tc,
															  Arena<S> rob, Arena<S> tk, int64_t B, int64_t stride, int with_tau_prec, int accumulate) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	mft_compute_robot<T, N>(ch, tc, rob, tk, stride, r, with_tau_prec, accumulate);
}

}  // namespace osc
