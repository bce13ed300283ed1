// osc_mft_update.cuh -- MotionForceTask::updateTaskModel + SingularityHandler::updateTaskModel / classifySingularity
// (src/tasks/MotionForceTask.cpp:247-268, src/tasks/SingularityHandler.cpp:75-295).
#pragma once
#include "osc_kernels.cuh"

namespace osc {

// J = P * JWorldFrame(link, p_c)   (MotionForceTask.cpp:261-263)
template <class T, int N>
OSC_DEV void task_jacobian(const ChainC<T, N>& ch, const MftC<T>& tc, const Kin<T, N>& kin, const T (&x)[3], T (&J)[6][N]) {
	point_jacobian<T, N>(ch, kin, tc.body, x, J);
	if (tc.partial) {
		OSC_UNROLL for (int j = 0; j < N; ++j) {
			const T l[3] = {J[0][j], J[1][j], J[2][j]}, a[3] = {J[3][j], J[4][j], J[5][j]};
			T pl[3], pa[3];
			mv3(tc.Pt, l, pl);
			mv3(tc.Pr, a, pa);
			OSC_UNROLL for (int k = 0; k < 3; ++k) { J[k][j] = pl[k]; J[3 + k][j] = pa[k]; }
		}
	}
}

// One robot.  `no_history` re-evaluates the task model without pushing a classification sample (lazy materialisation
// of the getters after a fused tick: the history was already advanced by the tick itself).
template <class T, int N, class S = T, bool REDIR = false>
__device__ void mft_update_robot(const ChainC<T, N>& ch, const MftC<T>& tc, const Arena<S>& rob, const Arena<S>& tk, const S* nprec_src,
								 int64_t stride, int64_t r, int no_history) {
	using L = MftL<N>;
	constexpr int K = L::K;
	const Row<T, S> rP{rob.P, stride, r}, rS = srow<T, REDIR>(rob, stride, r), tP{tk.P, stride, r}, tS = srow<T, REDIR>(tk, stride, r);
	const Row<int32_t> tI{tk.I, stride, r};
	int status = 0;

	Kin<T, N> kin;
	ld_kin<T, N>(rS, kin);
	T x[3], Rc[3][3];
	frame_pose<T, N>(kin, tc.body, tc.p_c, tc.R_c, x, Rc);
	T J[6][N];
	task_jacobian<T, N>(ch, tc, kin, x, J);
	T Np[N][N];
	if (nprec_src) {
		const Row<T, S> src{const_cast<S*>(nprec_src), stride, r};
		ld_mat<T, N, N>(src, 0, Np);
	} else {
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < N; ++j) Np[i][j] = (i == j) ? T(1) : T(0);
	}
	T Jp[6][N];
	mm<T, 6, N, N>(J, Np, Jp);  // MotionForceTask.cpp:264
	st_mat<T, 6, N>(tS, L::J, J);
	st_mat<T, 6, N>(tS, L::JP, Jp);
	st_mat<T, N, N>(tS, L::NPREC, Np);

	// ---- SingularityHandler::updateTaskModel ----
	T U[6][K], V[N][K], s[K];
	bool conv;
	if constexpr (N >= 6) {
		T A[N][6], W[6][6];
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < 6; ++j) A[i][j] = Jp[j][i];
		conv = svd_onesided<T, N, 6>(A, W, s);
		OSC_UNROLL for (int i = 0; i < 6; ++i)
			OSC_UNROLL for (int j = 0; j < K; ++j) U[i][j] = W[i][j];
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < K; ++j) V[i][j] = A[i][j];
	} else {
		T A[6][N], W[N][N];
		OSC_UNROLL for (int i = 0; i < 6; ++i)
			OSC_UNROLL for (int j = 0; j < N; ++j) A[i][j] = Jp[i][j];
		conv = svd_onesided<T, 6, N>(A, W, s);
		OSC_UNROLL for (int i = 0; i < 6; ++i)
			OSC_UNROLL for (int j = 0; j < K; ++j) U[i][j] = A[i][j];
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < K; ++j) V[i][j] = W[i][j];
	}
	if (!conv) status |= 8;
	// sign convention: largest-|component| of every V column positive
	OSC_UNROLL for (int c = 0; c < K; ++c) {
		T best = T(0), val = T(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			const T a = Num<T>::abs_(V[i][c]);
			if (a > best) { best = a; val = V[i][c]; }
		}
		if (val < T(0)) {
			OSC_UNROLL for (int i = 0; i < N; ++i) V[i][c] = -V[i][c];
			OSC_UNROLL for (int i = 0; i < 6; ++i) U[i][c] = -U[i][c];
		}
	}
	st_vec<T, K>(tS, L::SV, s);

	const int m = tc.rank;
	const unsigned mask_m = (1u << m) - 1u;
	unsigned mask_ns, mask_s;
	T alpha;
	bool fully = false;
	int ins = m;
	if (s[0] < tc.s_abs_tol) {  // SingularityHandler.cpp:83-98
		fully = true; alpha = T(0); mask_ns = 0u; mask_s = mask_m; ins = 0;
	} else {  // .cpp:100-142 (m == 1: non-singular, quirk Q4)
		T rho_ins = T(1);
		OSC_UNROLL for (int i = 1; i < K; ++i) {
			const T rho = s[i] / s[0];
			if (i < m && ins == m && rho < tc.s_max) { ins = i; rho_ins = rho; }
		}
		if (ins < m) {
			alpha = (rho_ins - tc.s_min) / (tc.s_max - tc.s_min);
			alpha = alpha < T(0) ? T(0) : (alpha > T(1) ? T(1) : alpha);
			mask_ns = (1u << ins) - 1u;
			mask_s = mask_m & ~mask_ns;
		} else {
			alpha = T(1); mask_ns = mask_m; mask_s = 0u;
		}
	}
	tS(L::ALPHA) = alpha;
	tI(L::INS) = fully ? 0 : ins;
	if (mask_s) status |= 32;

	T Minv[N][N];
	ld_mat<T, N, N>(rS, RobotL<N>::MINV, Minv);
	// modified inverse inertia for the *_modified Lambdas (.cpp:160-225)
	T Wm[N][N];
	if (tc.decoupling == 1) bie_inverse<T, N>(rS, tc.bie_threshold, Wm);
	else {
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < N; ++j) Wm[i][j] = Minv[i][j];
	}

	// non-singular part
	T Uns[6][K], Jns[K][N], Lns[K][K], Nns[N][N];
	OSC_UNROLL for (int i = 0; i < 6; ++i)
		OSC_UNROLL for (int c = 0; c < K; ++c) Uns[i][c] = ((mask_ns >> c) & 1u) ? U[i][c] : T(0);
	mtm<T, 6, K, N>(Uns, Jp, Jns);
	if (!op_space<T, K, N>(Jns, Minv, mask_ns, Lns, Nns)) status |= 16;
	{
		T Lmod[K][K], A[K][K];
		congruence<T, K, N>(Jns, Wm, A);
		spd_inverse_masked<T, K>(A, mask_ns, Lmod);
		if (tc.decoupling == 2) {
			OSC_UNROLL for (int i = 0; i < K; ++i)
				OSC_UNROLL for (int j = 0; j < K; ++j) Lmod[i][j] = (i == j && ((mask_ns >> i) & 1u)) ? T(1) : T(0);
		}
		st_mat<T, K, K>(tS, L::L_NS_MOD, Lmod);
	}
	st_mat<T, 6, K>(tS, L::U_NS, Uns);
	st_mat<T, K, N>(tS, L::J_NS, Jns);
	st_mat<T, K, K>(tS, L::L_NS, Lns);

	// singular part + posture + nullspace (.cpp:117-120, 145-158)
	T Nt[N][N];
	T Us[6][K], Vs[N][K];
	OSC_UNROLL for (int c = 0; c < K; ++c) {
		const bool on = (mask_s >> c) & 1u;
		OSC_UNROLL for (int i = 0; i < 6; ++i) Us[i][c] = on ? U[i][c] : T(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) Vs[i][c] = on ? V[i][c] : T(0);
	}
	if (mask_s) {
		T Js[K][N], Lsm[K][K], A[K][K];
		mtm<T, 6, K, N>(Us, Jp, Js);
		congruence<T, K, N>(Js, Wm, A);
		spd_inverse_masked<T, K>(A, mask_s, Lsm);
		if (tc.decoupling == 2) {
			OSC_UNROLL for (int i = 0; i < K; ++i)
				OSC_UNROLL for (int j = 0; j < K; ++j) Lsm[i][j] = (i == j && ((mask_s >> i) & 1u)) ? T(1) : T(0);
		}
		st_mat<T, K, N>(tS, L::J_S, Js);
		st_mat<T, K, K>(tS, L::L_S_MOD, Lsm);
	}
	st_mat<T, 6, K>(tS, L::U_S, Us);
	st_mat<T, N, K>(tS, L::V_S, Vs);
	if (fully) {  // quirk Q11: pass the task through
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < N; ++j) Nt[i][j] = Np[i][j];
	} else if (mask_s == 0u || !tc.handling) {
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < N; ++j) Nt[i][j] = Nns[i][j];
	} else {
		T NN[N][N], Jpost[K][N], Ljs[K][K], Npost[N][N];
		mm<T, N, N, N>(Nns, Np, NN);
		mtm<T, N, K, N>(Vs, NN, Jpost);  // V_s^T N_ns N_prec
		if (!op_space<T, K, N>(Jpost, Minv, mask_s, Ljs, Npost)) status |= 16;
		mm<T, N, N, N>(Npost, Nns, Nt);
		T Ljm[K][K], A[K][K];
		congruence<T, K, N>(Jpost, Wm, A);
		spd_inverse_masked<T, K>(A, mask_s, Ljm);
		if (tc.decoupling == 2) {
			OSC_UNROLL for (int i = 0; i < K; ++i)
				OSC_UNROLL for (int j = 0; j < K; ++j) Ljm[i][j] = (i == j && ((mask_s >> i) & 1u)) ? T(1) : T(0);
		}
		st_mat<T, K, N>(tS, L::J_POST, Jpost);
		st_mat<T, K, K>(tS, L::L_JS_MOD, Ljm);
	}
	st_mat<T, N, N>(tS, L::NT, Nt);
	{
		T NTP[N][N];
		mm<T, N, N, N>(Nt, Np, NTP);
		st_mat<T, N, N>(tS, L::NTP, NTP);
	}

	// ---- classifySingularity (.cpp:230-295) ----
	T q[N];
	ld_vec<T, N>(rP, RobotL<N>::Q, q);
	int c1 = tI(L::C1), c2 = tI(L::C2);
	if (!no_history && (tI(L::NTYPES) == 0 || c2 > c1)) {
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			tP(L::Q_PRIOR + i) = q[i];
			tP(L::DQ_PRIOR + i) = rP(RobotL<N>::DQ + i);
		}
	}
	if (mask_s == 0u) {
		tI(L::NTYPES) = 0; tI(L::TYPES) = 0; tI(L::C1) = 0; tI(L::C2) = 0; tI(L::HIST_LEN) = 0; tI(L::HIST_HEAD) = 0;
	} else if (!no_history) {
		int types = 0, ntypes = 0;
		bool any1 = false;
		const int first = fully ? 0 : ins;
		OSC_UNROLL for (int c = 0; c < K; ++c) {
			if ((mask_s >> c) & 1u) {
				T qp[N];
				OSC_UNROLL for (int i = 0; i < N; ++i) qp[i] = Num<T>::fma_(tc.perturb_step, Vs[i][c], q[i]);
				Kin<T, N> kp;
				forward_kinematics<T, N>(ch, qp, kp);
				T x2[3], R2[3][3], dw[3], ew[3], db[3], eb[3];
				frame_pose<T, N>(kp, tc.body, tc.p_c, tc.R_c, x2, R2);
				OSC_UNROLL for (int a = 0; a < 3; ++a) dw[a] = x2[a] - x[a];
				orientation_error<T>(R2, Rc, ew);
				// the reference measures the perturbation in the robot BASE frame (position()/rotation(), .cpp:250-260)
				mtv3(ch.world_rot, dw, db);
				mtv3(ch.world_rot, ew, eb);
				T motion = T(0);
				OSC_UNROLL for (int a = 0; a < 3; ++a) motion += db[a] * Us[a][c] + eb[a] * Us[3 + a][c];
				const int type = Num<T>::abs_(motion) > tc.type_1_tol ? 1 : 2;
				any1 = any1 || type == 1;
				types |= type << (2 * (c - first));
				++ntypes;
			}
		}
		tI(L::NTYPES) = ntypes; tI(L::TYPES) = types;
		// history ring of bits (1 = type 1), capacity OSC_MAX_SINGULARITY_BUFFER = 256
		int len = tI(L::HIST_LEN), head = tI(L::HIST_HEAD);
		// the deque of the reference pushes, then pops the front once it is longer than buffer_size (.cpp:276-294).  With
		// buffer_size == 256 the slot of the new bit IS the front's: pop first (the outgoing bit is read before it is overwritten)
		if (len >= tc.buffer_size) {
			const int32_t w = tk.I[(int64_t)(L::HIST + (head >> 5)) * stride + r];
			if ((w >> (head & 31)) & 1) --c1; else --c2;
			head = (head + 1) & 255;
			--len;
		}
		{
			const int pos = (head + len) & 255;
			int32_t& w = tk.I[(int64_t)(L::HIST + (pos >> 5)) * stride + r];
			w = any1 ? (w | (1 << (pos & 31))) : (w & ~(1 << (pos & 31)));
			++len;
			if (any1) ++c1; else ++c2;
		}
		tI(L::HIST_LEN) = len; tI(L::HIST_HEAD) = head; tI(L::C1) = c1; tI(L::C2) = c2;
	}
	if (status) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status);
}

template <class T, int N, class S = T>
__global__ void __launch_bounds__(OSC_THREADS, OSC_PERCALL_BLOCKS) k_mft_update(const __grid_constant__ ChainC<T, N> ch, const __grid_constant__ MftC<T> tc,
															 Arena<S> rob, Arena<S> tk, const S* nprec_src, int64_t B, int64_t stride, int no_history) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	mft_update_robot<T, N>(ch, tc, rob, tk, nprec_src, stride, r, no_history);
}

}  // namespace osc
