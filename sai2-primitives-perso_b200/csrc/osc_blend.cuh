// osc_blend.cuh -- the SINGULAR-BLEND branch of the tick in closed form, for the robots the fused kernels could not certify as
// non-singular (SingularityHandler.cpp:75-368 for the hierarchy [MotionForceTask of rank M, full JointTask], n = 7, revolute chain,
// control link on the last body).  Called by k_tick_tail ahead of the API-faithful chain; a robot this routine cannot serve
// (returns false BEFORE touching any state) takes that chain as before.
//
// Why: the API-faithful chain materialises every matrix the reference exposes (1,200 scratch doubles per robot, 71 k SASS
// instructions, 5 KB of local memory) and runs ~0.48 ms per thread whatever the number of listed robots (r2e).  The torques need
// far less.  With M = L L^T, J_r = B6^T J = U S V^T (one-sided Jacobi on the n x M matrix J_r^T), X = L^-1 V (n x M, WELL
// conditioned: X^T X = V^T M^-1 V whatever the singular values) and D = diag(1, .., 1, rho) the last-pivot bounded-inertia scaling
// (L_BIE^-1 L = D, beta = 1 - rho^2), every projected inertia of SingularityHandler::updateTaskModel is a block of ONE M x M matrix:
//     J_x W_m J_x^T = S_x (X^T D^2 X)[x,x] S_x            for x = ns (columns < ins) and x = s (columns >= ins)     (.cpp:160-204)
//     tau_x = J_x^T (Lambda_x,mod U_x^T F* + U_x^T F_f) = L X_x ( G_x^-1 (a_x / s_x) + s_x b_x ),  a = U^T F*, b = U^T F_f
// The posture Jacobian J_post = V_s^T N_ns (.cpp:117-120, 145-158) has L^-1 J_post^T = (I - P_ns) X_s = W_s R_ss, the s-block of the
// QR factorisation X = W R taken in column order, so J_post W_m J_post^T = R_ss^T (I - beta w w^T) R_ss (w = last row of W_s) and
//     tau_joint = J_post^T ( Lambda_js,mod V_s^T u [+ V_s^T u'] ) = L W_s R_ss ( ... ),   V_s^T u = X_s^T (L^T u).
// N_post N_ns = L^-T (I - W W^T) L^T = L^-T Q Q^T L^T: the nullspace handed to the joint task is the one of the FULL task, and the
// joint task is the R x R solve of k_tick_fused3 (osc_fused3.cuh P9, P11).  No matrix larger than M x M is inverted and nothing is
// written but the results; classification (perturbed forward kinematics per singular direction, history ring, q_prior latch,
// type-2 directions) follows the reference statement by statement.
#pragma once
#include "osc_fused3.cuh"

namespace osc {

template <int N, int M> struct BL {
	static constexpr int R = N - M;
	static constexpr int SC = 0;                                    // sin/cos (2N), later Q (N x R)
	static constexpr int QQ = 0;
	static constexpr int HEAD = 2 * N > N * R ? 2 * N : N * R;
	static constexpr int LL = HEAD;                                 // M lower triangle, then L strictly lower
	static constexpr int LI = LL + N * (N - 1) / 2;                 // 1 / L_ii
	static constexpr int XX = LL + N * (N + 1) / 2;                 // X = L^-1 V, column c at XX + c N + i
	static constexpr int LM = XX + N * M;                           // n = 6 only: L_BIE = chol(M_BIE) for ANY clamp pattern (strictly lower, then 1 / diagonal)
	static constexpr int LMI = LM + N * (N - 1) / 2;
	static constexpr int SLOTS = LM + (N == 6 ? N * (N + 1) / 2 : 0);
};

// pose of the control frame at joint angles q; rolled over the joints (small code: it runs once per singular direction)
template <class T, int N> OSC_DEV void blend_pose(const ChainC<T, N>& ch, const MftC<T>& tc, const T (&q)[N], T (&x)[3], T (&R)[3][3]) {
	T Rp[3][3], pp[3];
	OSC_UNROLL for (int a = 0; a < 3; ++a) {
		pp[a] = ch.world_xyz[a];
		OSC_UNROLL for (int b = 0; b < 3; ++b) Rp[a][b] = ch.world_rot[a][b];
	}
	_Pragma("unroll 1") for (int i = 0; i < N; ++i) {
		T pre[3][3], d[3], s, c;
		mm3(Rp, ch.rot[i], pre);
		mv3(Rp, ch.xyz[i], d);
		Num<T>::sincos_(q[i], &s, &c);
		OSC_UNROLL for (int a = 0; a < 3; ++a) {
			pp[a] += d[a];
			Rp[a][0] = c * pre[a][0] + s * pre[a][1];
			Rp[a][1] = c * pre[a][1] - s * pre[a][0];
			Rp[a][2] = pre[a][2];
		}
	}
	T d[3];
	mv3(Rp, tc.p_c, d);
	OSC_UNROLL for (int a = 0; a < 3; ++a) x[a] = pp[a] + d[a];
	mm3(Rp, tc.R_c, R);
}

// B6^T f for a world 6-vector f (translation rows first); M == 6: the identity
template <class T, int M> OSC_DEV void blend_range6(const MftC<T>& tc, const T (&f)[6], T (&o)[M]) {
	if constexpr (M == 6) {
		OSC_UNROLL for (int c = 0; c < 6; ++c) o[c] = f[c];
	} else {
		OSC_UNROLL for (int c = 0; c < M; ++c) {
			T s = T(0);
			if (c < tc.mt) { OSC_UNROLL for (int a = 0; a < 3; ++a) s = Num<T>::fma_(tc.Bt[a][c], f[a], s); }
			else {
				const int cr = (c - tc.mt) > 2 ? 2 : (c - tc.mt);
				OSC_UNROLL for (int a = 0; a < 3; ++a) s = Num<T>::fma_(tc.Br[a][cr], f[3 + a], s);
			}
			o[c] = s;
		}
	}
}

// One listed robot.  T = arithmetic type (double), S = storage type of the rows, sm = this block's stash (BL<N, M>::SLOTS x
// blockDim.x values of T).  xm: any scratch row set of the task (the U of the SVD is parked there between two phases).
template <class T, int N, int M, class S>
__device__ __forceinline__ bool blend_robot(const ChainC<T, N>& ch, const MftC<T>& tc, const JtC<T, N>& jc, const Arena<S>& rob, const Arena<S>& tk0,
											const Arena<S>& tk1, const Row<T, S>& xm, int64_t stride, int64_t r, int stacked, int gravity, int reuse_law,
											int exports, T* __restrict__ sm) {
	static_assert((N == 7 || N == 6) && M >= 2 && M <= 6 && M <= N, "blend_robot: n = 7 (task rank 2..6) or n = 6 (rank 6: no nullspace left)");
	using ML = MftL<N>;
	using JL = JtL<N>;
	using F = BL<N, M>;
	constexpr int R = F::R;
#define SM(slot) sm[(slot) * OSC_THREADS + threadIdx.x]
	// what this routine does not serve: a control law that was not run by the fused kernel, singularity handling off, IMPEDANCE
	// decoupling of the motion task (its torque has no singular terms, .cpp:310-312), a rank that is not this instantiation's
	const bool has_jt = !(exports & OSC_EXPORT_NO_JT);
	if (!(reuse_law & 1) || !tc.handling || tc.decoupling == OSC_IMPEDANCE || tc.rank != M || (has_jt && !jc.full) || (M < 6) != (tc.partial != 0)) return false;
	const Row<T, S> rP{rob.P, stride, r}, rS{rob.S, stride, r}, mP{tk0.P, stride, r}, mS{tk0.S, stride, r}, jP{tk1.P, stride, r}, jS{tk1.S, stride, r};
	const Row<int32_t> mI{tk0.I, stride, r};
	const bool ex = (exports & OSC_EXPORT_VECTORS) != 0;
	int status = 0;

	// ---- q, sin/cos ----
	T q[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		q[i] = rP(RobotL<N>::Q + i);
		T s, c;
		Num<T>::sincos_(q[i], &s, &c);
		SM(F::SC + 2 * i) = s;
		SM(F::SC + 2 * i + 1) = c;
	}
	// ---- mass matrix: body-frame composite-rigid-body recursion (k_tick_fused2 P2) ----
	{
		T cm = T(0), h0 = T(0), h1 = T(0), h2 = T(0), Ixx = T(0), Ixy = T(0), Ixz = T(0), Iyy = T(0), Iyz = T(0), Izz = T(0);
		OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
			cm += ch.mass[i];
			h0 += ch.hb[i][0]; h1 += ch.hb[i][1]; h2 += ch.hb[i][2];
			Ixx += ch.Ib[i][0]; Ixy += ch.Ib[i][1]; Ixz += ch.Ib[i][2]; Iyy += ch.Ib[i][3]; Iyz += ch.Ib[i][4]; Izz += ch.Ib[i][5];
			T p0 = -h1, p1 = h0, p2 = T(0), l0 = Ixz, l1 = Iyz, l2 = Izz;
			const int row = F::LL + i * (i + 1) / 2;
			SM(row + i) = l2;
			OSC_UNROLL for (int j = i; j >= 1; --j) {
				const T s = SM(F::SC + 2 * j), c = SM(F::SC + 2 * j + 1);
				const T a0 = c * p0 - s * p1, a1 = s * p0 + c * p1, b0 = c * l0 - s * l1, b1 = s * l0 + c * l1;
				p0 = ch.rot[j][0][0] * a0 + ch.rot[j][0][1] * a1 + ch.rot[j][0][2] * p2;
				p1 = ch.rot[j][1][0] * a0 + ch.rot[j][1][1] * a1 + ch.rot[j][1][2] * p2;
				p2 = ch.rot[j][2][0] * a0 + ch.rot[j][2][1] * a1 + ch.rot[j][2][2] * p2;
				const T m0 = ch.rot[j][0][0] * b0 + ch.rot[j][0][1] * b1 + ch.rot[j][0][2] * l2;
				const T m1 = ch.rot[j][1][0] * b0 + ch.rot[j][1][1] * b1 + ch.rot[j][1][2] * l2;
				const T m2 = ch.rot[j][2][0] * b0 + ch.rot[j][2][1] * b1 + ch.rot[j][2][2] * l2;
				l0 = m0 + ch.xyz[j][1] * p2 - ch.xyz[j][2] * p1;
				l1 = m1 + ch.xyz[j][2] * p0 - ch.xyz[j][0] * p2;
				l2 = m2 + ch.xyz[j][0] * p1 - ch.xyz[j][1] * p0;
				SM(row + j - 1) = l2;
			}
			if (i > 0) {
				const T s = SM(F::SC + 2 * i), c = SM(F::SC + 2 * i + 1);
				T E[3][3];
				OSC_UNROLL for (int a = 0; a < 3; ++a) {
					E[a][0] = ch.rot[i][a][0] * c + ch.rot[i][a][1] * s;
					E[a][1] = ch.rot[i][a][1] * c - ch.rot[i][a][0] * s;
					E[a][2] = ch.rot[i][a][2];
				}
				const T I3[3][3] = {{Ixx, Ixy, Ixz}, {Ixy, Iyy, Iyz}, {Ixz, Iyz, Izz}};
				T EI[3][3];
				mm3(E, I3, EI);
				const T g0 = E[0][0] * h0 + E[0][1] * h1 + E[0][2] * h2, g1 = E[1][0] * h0 + E[1][1] * h1 + E[1][2] * h2,
						g2 = E[2][0] * h0 + E[2][1] * h1 + E[2][2] * h2;
				const T r0 = ch.xyz[i][0], r1 = ch.xyz[i][1], r2 = ch.xyz[i][2];
				const T iso = cm * (r0 * r0 + r1 * r1 + r2 * r2) + T(2) * (r0 * g0 + r1 * g1 + r2 * g2);
				Ixx = EI[0][0] * E[0][0] + EI[0][1] * E[0][1] + EI[0][2] * E[0][2] + iso - cm * r0 * r0 - T(2) * r0 * g0;
				Iyy = EI[1][0] * E[1][0] + EI[1][1] * E[1][1] + EI[1][2] * E[1][2] + iso - cm * r1 * r1 - T(2) * r1 * g1;
				Izz = EI[2][0] * E[2][0] + EI[2][1] * E[2][1] + EI[2][2] * E[2][2] + iso - cm * r2 * r2 - T(2) * r2 * g2;
				Ixy = EI[0][0] * E[1][0] + EI[0][1] * E[1][1] + EI[0][2] * E[1][2] - cm * r0 * r1 - r0 * g1 - g0 * r1;
				Ixz = EI[0][0] * E[2][0] + EI[0][1] * E[2][1] + EI[0][2] * E[2][2] - cm * r0 * r2 - r0 * g2 - g0 * r2;
				Iyz = EI[1][0] * E[2][0] + EI[1][1] * E[2][1] + EI[1][2] * E[2][2] - cm * r1 * r2 - r1 * g2 - g1 * r2;
				h0 = g0 + cm * r0; h1 = g1 + cm * r1; h2 = g2 + cm * r2;
			}
		}
	}
	// ---- M = L L^T; bounded-inertia clamp pattern: last joint only, any other pattern -> the API-faithful chain ----
	T beta = T(0);
	{
		T Mx[N][N], li[N];
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) Mx[i][j] = SM(F::LL + i * (i + 1) / 2 + j);
		T delta = T(0);
		if constexpr (N == 6) {
			// no nullspace below a full task on a 6-DoF arm, so the last-pivot shortcut (which exists for the joint task's sake) is not
			// needed: keep L_BIE = chol(M_BIE) itself, whatever the clamp pattern (the PUMA clamps its three wrist joints)
			if (tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES) {
				T Mb[N][N], lb[N];
				OSC_UNROLL for (int i = 0; i < N; ++i)
					OSC_UNROLL for (int j = 0; j <= i; ++j) Mb[i][j] = (i == j && Mx[i][i] < tc.bie_threshold) ? tc.bie_threshold : Mx[i][j];
				if (!chol_inplace<T, N>(Mb, lb)) return false;
				OSC_UNROLL for (int i = 0; i < N; ++i) {
					OSC_UNROLL for (int j = 0; j < i; ++j) SM(F::LM + i * (i - 1) / 2 + j) = Mb[i][j];
					SM(F::LMI + i) = lb[i];
				}
			}
		} else
		if (tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES || (has_jt && jc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES)) {
			const T thr = tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES ? tc.bie_threshold : jc.bie_threshold;
			bool other = false;
			OSC_UNROLL for (int i = 0; i < N - 1; ++i) other = other || (Mx[i][i] < thr);
			if (other) return false;
			if (Mx[N - 1][N - 1] < thr) delta = thr - Mx[N - 1][N - 1];
		}
		if (!chol_inplace<T, N>(Mx, li)) return false;
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			OSC_UNROLL for (int j = 0; j < i; ++j) SM(F::LL + i * (i - 1) / 2 + j) = Mx[i][j];
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) SM(F::LI + i) = li[i];
		if (delta > T(0)) {
			const T d = Mx[N - 1][N - 1] * Mx[N - 1][N - 1];
			beta = T(1) - d / (d + delta);
		}
	}
	const T beta_m = tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES ? beta : T(0);  // the motion task's own decoupling type
	// ---- forward kinematics, A = J_r^T (n x M): column c = B6[:, c]^T J_world (k_tick_fused3 P4) ----
	T A[N][M];
	{
		T Rp[3][3], pp[3], Jz[N][3], Jo[N][3], x[3];
		OSC_UNROLL for (int a = 0; a < 3; ++a) {
			pp[a] = ch.world_xyz[a];
			OSC_UNROLL for (int b = 0; b < 3; ++b) Rp[a][b] = ch.world_rot[a][b];
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			T pre[3][3], d[3];
			mm3(Rp, ch.rot[i], pre);
			mv3(Rp, ch.xyz[i], d);
			const T s = SM(F::SC + 2 * i), c = SM(F::SC + 2 * i + 1);
			OSC_UNROLL for (int a = 0; a < 3; ++a) {
				pp[a] += d[a];
				Rp[a][0] = c * pre[a][0] + s * pre[a][1];
				Rp[a][1] = c * pre[a][1] - s * pre[a][0];
				Rp[a][2] = pre[a][2];
				Jz[i][a] = pre[a][2];
				Jo[i][a] = pp[a];
			}
		}
		T d[3];
		mv3(Rp, tc.p_c, d);
		OSC_UNROLL for (int a = 0; a < 3; ++a) x[a] = pp[a] + d[a];
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			const T e[3] = {x[0] - Jo[i][0], x[1] - Jo[i][1], x[2] - Jo[i][2]};
			T l[3];
			cross3(Jz[i], e, l);
			const T col[6] = {l[0], l[1], l[2], Jz[i][0], Jz[i][1], Jz[i][2]};
			T o[M];
			blend_range6<T, M>(tc, col, o);
			OSC_UNROLL for (int c = 0; c < M; ++c) A[i][c] = o[c];
		}
	}
	// ---- SVD: J_r^T = V S U_r^T (SingularityHandler.cpp:78); fully singular or not converged -> the API-faithful chain ----
	T W[M][M], s[M];
	if (!svd_onesided<T, N, M>(A, W, s)) return false;
	if (!(s[0] >= tc.s_abs_tol)) return false;
	// sign convention of the pair (v_c, u_c) (largest-|component| of every V column positive): the torques do not
	// depend on it, the classification does -- the pose change under q + step v_c is second order in a singular direction
	OSC_UNROLL for (int c = 0; c < M; ++c) {
		T best = T(0), val = T(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			const T av = Num<T>::abs_(A[i][c]);
			if (av > best) { best = av; val = A[i][c]; }
		}
		if (val < T(0)) {
			OSC_UNROLL for (int i = 0; i < N; ++i) A[i][c] = -A[i][c];
			OSC_UNROLL for (int k = 0; k < M; ++k) W[k][c] = -W[k][c];
		}
	}
	// a = U^T F*, b = U^T F_f (F*, F_f handed over by the fused kernel); |F* + F_f| for the type-2 direction (.cpp:336-339)
	T a[M], b[M], nf = T(0);
	{
		T f6[6], g6[6], fr[M], gr[M];
		OSC_UNROLL for (int k = 0; k < 6; ++k) {
			f6[k] = mS(ML::UMF + k);
			g6[k] = mS(ML::FFB + k);
			const T t = f6[k] + g6[k];
			nf = Num<T>::fma_(t, t, nf);
		}
		nf = Num<T>::sqrt_(nf);
		blend_range6<T, M>(tc, f6, fr);
		blend_range6<T, M>(tc, g6, gr);
		OSC_UNROLL for (int c = 0; c < M; ++c) {
			T sa = T(0), sb = T(0);
			OSC_UNROLL for (int k = 0; k < M; ++k) { sa = Num<T>::fma_(W[k][c], fr[k], sa); sb = Num<T>::fma_(W[k][c], gr[k], sb); }
			a[c] = sa; b[c] = sb;
		}
	}
	OSC_UNROLL for (int k = 0; k < M; ++k)
		OSC_UNROLL for (int c = 0; c < M; ++c) xm(ML::U_NS + k * M + c) = W[k][c];  // parked for the classification
	// ---- the split (.cpp:100-142) ----
	int ins = M;
	T alpha = T(1);
	{
		T rho_ins = T(1);
		OSC_UNROLL for (int i = 1; i < M; ++i) {
			const T rho = s[i] / s[0];
			if (ins == M && rho < tc.s_max) { ins = i; rho_ins = rho; }
		}
		if (ins < M) {
			alpha = (rho_ins - tc.s_min) / (tc.s_max - tc.s_min);
			alpha = alpha < T(0) ? T(0) : (alpha > T(1) ? T(1) : alpha);
		}
	}
	const unsigned mask_m = (1u << M) - 1u, mask_ns = (1u << ins) - 1u, mask_s = mask_m & ~mask_ns;
	if (mask_s) status |= 32;
	// ---- X = L^-1 V (forward substitution, in place), its Gram matrix and its last row ----
	OSC_UNROLL for (int k = 0; k < N; ++k) {
		T Lk[N];
		OSC_UNROLL for (int m = 0; m < k; ++m) Lk[m] = SM(F::LL + k * (k - 1) / 2 + m);
		const T li = SM(F::LI + k);
		OSC_UNROLL for (int c = 0; c < M; ++c) {
			T t = A[k][c];
			OSC_UNROLL for (int m = 0; m < k; ++m) t = Num<T>::fma_(-Lk[m], A[m][c], t);
			A[k][c] = t * li;
		}
	}
	OSC_UNROLL for (int c = 0; c < M; ++c)
		OSC_UNROLL for (int i = 0; i < N; ++i) SM(F::XX + c * N + i) = A[i][c];
	// ---- classifySingularity (.cpp:230-295): q_prior latch, type of every singular direction from a perturbed pose, history ----
	int c1 = mI(ML::C1), c2 = mI(ML::C2);
	if (mI(ML::NTYPES) == 0 || c2 > c1) {
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			mP(ML::Q_PRIOR + i) = q[i];
			mP(ML::DQ_PRIOR + i) = (T)rP(RobotL<N>::DQ + i);
		}
	}
	if (mask_s == 0u) {
		mI(ML::NTYPES) = 0; mI(ML::TYPES) = 0; mI(ML::C1) = 0; mI(ML::C2) = 0; mI(ML::HIST_LEN) = 0; mI(ML::HIST_HEAD) = 0;
		c1 = 0; c2 = 0;
	} else {
		int types = 0, ntypes = 0;
		bool any1 = false;
		T x0[3], R0[3][3];
		blend_pose<T, N>(ch, tc, q, x0, R0);
		_Pragma("unroll 1") for (int c = ins; c < M; ++c) {
			T qp[N];
			OSC_UNROLL for (int i = 0; i < N; ++i) {  // v_c = L x_c
				T t = SM(F::XX + c * N + i) / SM(F::LI + i);
				OSC_UNROLL for (int m = 0; m < i; ++m) t = Num<T>::fma_(SM(F::LL + i * (i - 1) / 2 + m), SM(F::XX + c * N + m), t);
				qp[i] = Num<T>::fma_(tc.perturb_step, t, q[i]);
			}
			T x2[3], R2[3][3], dw[3], ew[3], d6[6];
			blend_pose<T, N>(ch, tc, qp, x2, R2);
			OSC_UNROLL for (int k = 0; k < 3; ++k) dw[k] = x2[k] - x0[k];
			orientation_error<T>(R2, R0, ew);
			{  // the reference measures the perturbation in the robot BASE frame (position()/rotation(), .cpp:250-260)
				T db[3], eb[3];
				mtv3(ch.world_rot, dw, db);
				mtv3(ch.world_rot, ew, eb);
				OSC_UNROLL for (int k = 0; k < 3; ++k) { d6[k] = db[k]; d6[3 + k] = eb[k]; }
			}
			T dr[M];
			blend_range6<T, M>(tc, d6, dr);
			T motion = T(0);
			OSC_UNROLL for (int k = 0; k < M; ++k) motion = Num<T>::fma_(dr[k], (T)xm(ML::U_NS + k * M + c), motion);
			const int type = Num<T>::abs_(motion) > tc.type_1_tol ? 1 : 2;
			any1 = any1 || type == 1;
			types |= type << (2 * (c - ins));
			++ntypes;
		}
		mI(ML::NTYPES) = ntypes; mI(ML::TYPES) = types;
		int len = mI(ML::HIST_LEN), head = mI(ML::HIST_HEAD);
		if (len >= tc.buffer_size) {  // pop before push: with buffer_size == 256 the new bit takes the front's slot (osc_mft_update.cuh)
			const int32_t w = tk0.I[(int64_t)(ML::HIST + (head >> 5)) * stride + r];
			if ((w >> (head & 31)) & 1) --c1; else --c2;
			head = (head + 1) & 255;
			--len;
		}
		{
			const int pos = (head + len) & 255;
			int32_t& w = tk0.I[(int64_t)(ML::HIST + (pos >> 5)) * stride + r];
			w = any1 ? (w | (1 << (pos & 31))) : (w & ~(1 << (pos & 31)));
			++len;
			if (any1) ++c1; else ++c2;
		}
		mI(ML::HIST_LEN) = len; mI(ML::HIST_HEAD) = head; mI(ML::C1) = c1; mI(ML::C2) = c2;
	}
	mI(ML::INS) = ins;

	// ---- torques of the motion task (.cpp:297-368) ----
	T X[N][M];
	OSC_UNROLL for (int c = 0; c < M; ++c)
		OSC_UNROLL for (int i = 0; i < N; ++i) X[i][c] = SM(F::XX + c * N + i);
	T dq[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) dq[i] = rP(RobotL<N>::DQ + i);
	auto Lmul = [&](const T (&h)[N], T (&y)[N]) {  // y = L h
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			T t = h[i] / SM(F::LI + i);
			OSC_UNROLL for (int m = 0; m < i; ++m) t = Num<T>::fma_(SM(F::LL + i * (i - 1) / 2 + m), h[m], t);
			y[i] = t;
		}
	};
	const bool gen_bie = N == 6 && tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES;  // general clamp pattern: L_BIE in the stash
	auto bie_map = [&](const T (&y)[N], T (&z)[N]) {  // z = L_BIE^-1 L y  (= D y for the last-pivot pattern)
		T v[N];
		Lmul(y, v);
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			T t = v[i];
			OSC_UNROLL for (int m = 0; m < i; ++m) t = Num<T>::fma_(-SM(F::LM + i * (i - 1) / 2 + m), z[m], t);
			z[i] = t * SM(F::LMI + i);
		}
	};
	// G = X^T (L_BIE^-1 L)^T (L_BIE^-1 L) X (M x M): every projected inertia below is a block of it.  Last-pivot pattern:
	// L_BIE^-1 L = D = diag(1, .., 1, rho), G = X^T X - beta x_n x_n^T
	T G[M][M];
	if (gen_bie) {
		T Z[N][M];
		OSC_UNROLL for (int c = 0; c < M; ++c) {
			T y[N], z[N];
			OSC_UNROLL for (int i = 0; i < N; ++i) y[i] = X[i][c];
			bie_map(y, z);
			OSC_UNROLL for (int i = 0; i < N; ++i) Z[i][c] = z[i];
		}
		OSC_UNROLL for (int i = 0; i < M; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) {
				T t = T(0);
				OSC_UNROLL for (int k = 0; k < N; ++k) t = Num<T>::fma_(Z[k][i], Z[k][j], t);
				G[i][j] = t; G[j][i] = t;
			}
	} else {
		OSC_UNROLL for (int i = 0; i < M; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) {
				T t = T(0);
				OSC_UNROLL for (int k = 0; k < N; ++k) t = Num<T>::fma_(X[k][i], X[k][j], t);
				t = Num<T>::fma_(-beta_m * X[N - 1][i], X[N - 1][j], t);
				G[i][j] = t; G[j][i] = t;
			}
	}
	auto range_coef = [&](unsigned mask, T (&h)[N]) {  // h = X_x (G_x^-1 (a_x / s_x) + s_x b_x): tau_x = L h
		T Gc[M][M], Inv[M][M], rhs[M], coef[M];
		OSC_UNROLL for (int i = 0; i < M; ++i)
			OSC_UNROLL for (int j = 0; j < M; ++j) Gc[i][j] = G[i][j];
		if (!spd_inverse_masked<T, M>(Gc, mask, Inv)) status |= 16;
		OSC_UNROLL for (int c = 0; c < M; ++c) rhs[c] = ((mask >> c) & 1u) ? a[c] / s[c] : T(0);
		OSC_UNROLL for (int c = 0; c < M; ++c) {
			T t = T(0);
			OSC_UNROLL for (int k = 0; k < M; ++k) t = Num<T>::fma_(Inv[c][k], rhs[k], t);
			coef[c] = ((mask >> c) & 1u) ? Num<T>::fma_(s[c], b[c], t) : T(0);
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			T t = T(0);
			OSC_UNROLL for (int c = 0; c < M; ++c) t = Num<T>::fma_(X[i][c], coef[c], t);
			h[i] = t;
		}
	};
	T tau1[N];
	{
		T h[N];
		range_coef(mask_ns, h);
		Lmul(h, tau1);
	}
	if (mask_s) {
		T ts[N];
		{
			T h[N];
			range_coef(mask_s, h);
			Lmul(h, ts);
		}
		// joint-space strategy in the singular directions: u (and u' for type 2), projected on V_s = L X_s
		T pa[M], pb[M];
		bool type1 = c1 > c2 || tc.enforce_type_1;
		{
			T unit[N], damp[N];
			if (type1) {  // .cpp:328-331
				OSC_UNROLL for (int i = 0; i < N; ++i) { unit[i] = -tc.kp1 * (q[i] - (T)mP(ML::Q_PRIOR + i)) - tc.kv1 * dq[i]; damp[i] = T(0); }
			} else {  // .cpp:332-351
				T ab = T(0), xi[N];
				OSC_UNROLL for (int i = 0; i < N; ++i) xi[i] = T(0);
				OSC_UNROLL for (int c = 0; c < M; ++c)
					if (c == ins) {
						ab = a[c] + b[c];
						OSC_UNROLL for (int i = 0; i < N; ++i) xi[i] = X[i][c];
					}
				const T fTd = nf > T(0) ? ab / nf : ab;
				T v0[N];
				Lmul(xi, v0);  // first singular column of V
				OSC_UNROLL for (int i = 0; i < N; ++i) {
					T dir = mP(ML::T2DIR + i);
					if (v0[i] != T(0)) {
						if (Num<T>::abs_(q[i] - ch.q_upper[i]) < tc.type_2_angle_threshold) dir = T(-1);
						else if (Num<T>::abs_(q[i] - ch.q_lower[i]) < tc.type_2_angle_threshold) dir = T(1);
						mP(ML::T2DIR + i) = dir;
					}
					unit[i] = dir * (Num<T>::abs_(fTd) * tc.type_2_torque_ratio * ch.effort[i]);  // quirk Q5: intended ratio
					damp[i] = -tc.kv2 * dq[i];
				}
			}
			// V_s^T u = X_s^T (L^T u)
			T lu[N], ld[N];
			OSC_UNROLL for (int i = 0; i < N; ++i) {
				T t = unit[i] / SM(F::LI + i), d = damp[i] / SM(F::LI + i);
				OSC_UNROLL for (int k = i + 1; k < N; ++k) {
					const T l = SM(F::LL + k * (k - 1) / 2 + i);
					t = Num<T>::fma_(l, unit[k], t);
					d = Num<T>::fma_(l, damp[k], d);
				}
				lu[i] = t; ld[i] = d;
			}
			OSC_UNROLL for (int c = 0; c < M; ++c) {
				T t = T(0), d = T(0);
				OSC_UNROLL for (int i = 0; i < N; ++i) { t = Num<T>::fma_(X[i][c], lu[i], t); d = Num<T>::fma_(X[i][c], ld[i], d); }
				const bool on = (mask_s >> c) & 1u;
				pa[c] = on ? t : T(0);
				pb[c] = on ? d : T(0);
			}
		}
		// X = W Rm (modified Gram-Schmidt in column order, in place): L^-1 J_post^T = W_s R_ss
		T Rm[M][M];
		OSC_UNROLL for (int c = 0; c < M; ++c) {
			T nn = T(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) nn = Num<T>::fma_(X[i][c], X[i][c], nn);
			if (!(nn > T(0))) status |= 16;
			const T inv = Num<T>::rsqrt_(nn);
			Rm[c][c] = nn * inv;
			OSC_UNROLL for (int i = 0; i < N; ++i) X[i][c] *= inv;
			OSC_UNROLL for (int c2 = c + 1; c2 < M; ++c2) {
				T rr = T(0);
				OSC_UNROLL for (int i = 0; i < N; ++i) rr = Num<T>::fma_(X[i][c], X[i][c2], rr);
				Rm[c][c2] = rr;
				OSC_UNROLL for (int i = 0; i < N; ++i) X[i][c2] = Num<T>::fma_(-rr, X[i][c], X[i][c2]);
			}
		}
		T tj[N];
		{
			T yn[M], Gj[M][M], Ij[M][M], hh[M], e[M], hj[N];
			OSC_UNROLL for (int c = 0; c < M; ++c) {  // last row of W_s R_ss
				T t = T(0);
				OSC_UNROLL for (int p = 0; p <= c; ++p) t = (p >= ins) ? Num<T>::fma_(X[N - 1][p], Rm[p][c], t) : t;
				yn[c] = t;
			}
			if (gen_bie) {  // columns z_c = L_BIE^-1 L (W_s R_ss)[:, c]
				T Z[N][M];
				OSC_UNROLL for (int c = 0; c < M; ++c) {
					T y[N], z[N];
					OSC_UNROLL for (int i = 0; i < N; ++i) {
						T t = T(0);
						OSC_UNROLL for (int p = 0; p <= c; ++p) t = (p >= ins) ? Num<T>::fma_(X[i][p], Rm[p][c], t) : t;
						y[i] = t;
					}
					bie_map(y, z);
					OSC_UNROLL for (int i = 0; i < N; ++i) Z[i][c] = z[i];
				}
				OSC_UNROLL for (int i = 0; i < M; ++i)
					OSC_UNROLL for (int j = 0; j <= i; ++j) {
						T t = T(0);
						OSC_UNROLL for (int k = 0; k < N; ++k) t = Num<T>::fma_(Z[k][i], Z[k][j], t);
						Gj[i][j] = t; Gj[j][i] = t;
					}
			} else {
				OSC_UNROLL for (int i = 0; i < M; ++i)
					OSC_UNROLL for (int j = 0; j <= i; ++j) {
						T t = T(0);
						OSC_UNROLL for (int p = 0; p <= j; ++p) t = (p >= ins) ? Num<T>::fma_(Rm[p][i], Rm[p][j], t) : t;
						t = Num<T>::fma_(-beta_m * yn[i], yn[j], t);
						Gj[i][j] = t; Gj[j][i] = t;
					}
			}
			if (!spd_inverse_masked<T, M>(Gj, mask_s, Ij)) status |= 16;
			OSC_UNROLL for (int c = 0; c < M; ++c) {
				T t = T(0);
				OSC_UNROLL for (int k = 0; k < M; ++k) t = Num<T>::fma_(Ij[c][k], type1 ? pa[k] : pb[k], t);
				hh[c] = type1 ? t : t + pa[c];
			}
			OSC_UNROLL for (int p = 0; p < M; ++p) {  // e = R_ss h
				T t = T(0);
				OSC_UNROLL for (int c = p; c < M; ++c) t = Num<T>::fma_(Rm[p][c], hh[c], t);
				e[p] = (p >= ins) ? t : T(0);
			}
			OSC_UNROLL for (int i = 0; i < N; ++i) {
				T t = T(0);
				OSC_UNROLL for (int p = 0; p < M; ++p) t = Num<T>::fma_(X[i][p], e[p], t);
				hj[i] = t;
			}
			Lmul(hj, tj);
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) {  // .cpp:357-365
			if (ts[i] != ts[i]) { ts[i] = T(0); status |= 1; }
			else if (ts[i] > ch.effort[i]) ts[i] = ch.effort[i];
			else if (ts[i] < -ch.effort[i]) ts[i] = -ch.effort[i];
			tau1[i] = tau1[i] + alpha * ts[i] + (T(1) - alpha) * tj[i];
		}
	} else {
		// the orthonormal basis W of range(X) is needed below in any case
		OSC_UNROLL for (int c = 0; c < M; ++c) {
			T nn = T(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) nn = Num<T>::fma_(X[i][c], X[i][c], nn);
			if (!(nn > T(0))) status |= 16;
			const T inv = Num<T>::rsqrt_(nn);
			OSC_UNROLL for (int i = 0; i < N; ++i) X[i][c] *= inv;
			OSC_UNROLL for (int c2 = c + 1; c2 < M; ++c2) {
				T rr = T(0);
				OSC_UNROLL for (int i = 0; i < N; ++i) rr = Num<T>::fma_(X[i][c], X[i][c2], rr);
				OSC_UNROLL for (int i = 0; i < N; ++i) X[i][c2] = Num<T>::fma_(-rr, X[i][c], X[i][c2]);
			}
		}
	}

	// n = 6 under a full task: no nullspace left -- a joint task below it has range rank 0 and contributes nothing (JointTask.cpp:234-239);
	// n = 7 with the motion task as the whole hierarchy (examples/07): there is no joint task
	auto motion_task_only = [&]() {
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			T t = tau1[i];
			if (gravity) t += (T)rS(RobotL<N>::GRAV + i);
			rS(RobotL<N>::TAU + i) = t;
			if (ex) {
				mS(ML::TASK_TAU + i) = tau1[i];
				if (has_jt) { jS(JL::TASK_TAU + i) = T(0); jS(JL::CUR_POS + i) = q[i]; jS(JL::CUR_VEL + i) = dq[i]; }
			}
		}
	};
	if constexpr (R == 0) {
		motion_task_only();
		(void)jP; (void)stacked;
	} else if (!has_jt) {
		motion_task_only();
	} else {
	// ---- Q = orthonormal basis of the complement of range(W) (n x R): pivoted Gram-Schmidt on the unit vectors (k_tick_fused3 P9);
	// sin/cos are dead, Q takes their slots ----
	T k_last[R];
	{
		T res[N];
		OSC_UNROLL for (int k = 0; k < N; ++k) {
			T t = T(1);
			OSC_UNROLL for (int c = 0; c < M; ++c) t = Num<T>::fma_(-X[k][c], X[k][c], t);
			res[k] = t;
		}
		OSC_UNROLL for (int j = 0; j < R; ++j) {
			T best = T(-1);
			int kb = 0;
			OSC_UNROLL for (int k = 0; k < N; ++k)
				if (res[k] > best) { best = res[k]; kb = k; }
			T u[N];
			OSC_UNROLL for (int i = 0; i < N; ++i) u[i] = (i == kb) ? T(1) : T(0);
			OSC_UNROLL for (int pass = 0; pass < 2; ++pass) {
				T aa[M];
				OSC_UNROLL for (int c = 0; c < M; ++c) {
					T t = T(0);
					OSC_UNROLL for (int i = 0; i < N; ++i) t = Num<T>::fma_(X[i][c], u[i], t);
					aa[c] = t;
				}
				OSC_UNROLL for (int i = 0; i < N; ++i) {
					T t = u[i];
					OSC_UNROLL for (int c = 0; c < M; ++c) t = Num<T>::fma_(-X[i][c], aa[c], t);
					u[i] = t;
				}
				OSC_UNROLL for (int p = 0; p < R; ++p) {
					if (p < j) {
						T d = T(0);
						OSC_UNROLL for (int i = 0; i < N; ++i) d = Num<T>::fma_(SM(F::QQ + p * N + i), u[i], d);
						OSC_UNROLL for (int i = 0; i < N; ++i) u[i] = Num<T>::fma_(-d, SM(F::QQ + p * N + i), u[i]);
					}
				}
			}
			T nu = T(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) nu = Num<T>::fma_(u[i], u[i], nu);
			const T inv = Num<T>::rsqrt_(nu);
			OSC_UNROLL for (int i = 0; i < N; ++i) {
				u[i] *= inv;
				SM(F::QQ + j * N + i) = u[i];
				res[i] = Num<T>::fma_(-u[i], u[i], res[i]);
			}
			k_last[j] = u[N - 1];
		}
	}
	// ---- joint task in the R-dimensional nullspace (JointTask.cpp:218-356; k_tick_fused3 P11) ----
	T mt[N];  // M^-1 tau_1 = L^-T L^-1 tau_1
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		T t = tau1[i];
		OSC_UNROLL for (int m = 0; m < i; ++m) t = Num<T>::fma_(-SM(F::LL + i * (i - 1) / 2 + m), mt[m], t);
		mt[i] = t * SM(F::LI + i);
	}
	OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
		T t = mt[i];
		OSC_UNROLL for (int k = i + 1; k < N; ++k) t = Num<T>::fma_(-SM(F::LL + k * (k - 1) / 2 + i), mt[k], t);
		mt[i] = t * SM(F::LI + i);
	}
	T y1[N], upd[N];
	{
		T acc[N];
		jt_pd_law<T, N>(jc, jP, jS, q, dq, upd, acc, ex, !(reuse_law & 2));  // the SIMPLE fused kernel integrates before it can fall back
		OSC_UNROLL for (int i = 0; i < N; ++i) y1[i] = acc[i] - (stacked ? mt[i] : T(0));
		if (ex) {
			OSC_UNROLL for (int i = 0; i < N; ++i) { jS(JL::CUR_POS + i) = q[i]; jS(JL::CUR_VEL + i) = dq[i]; }
		}
	}
	const T beta_j = jc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES ? beta : T(0);
	T V[R][N], z1[R], z2[R], H[R][R];
	OSC_UNROLL for (int j = 0; j < R; ++j) {
		OSC_UNROLL for (int i = 0; i < N; ++i) V[j][i] = SM(F::QQ + j * N + i);
		OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
			T t = V[j][i];
			OSC_UNROLL for (int k = i + 1; k < N; ++k) t = Num<T>::fma_(-SM(F::LL + k * (k - 1) / 2 + i), V[j][k], t);
			V[j][i] = t * SM(F::LI + i);
		}
		T aa = T(0), bb = T(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) { aa = Num<T>::fma_(V[j][i], y1[i], aa); bb = Num<T>::fma_(V[j][i], upd[i], bb); }
		z1[j] = aa; z2[j] = bb;
		OSC_UNROLL for (int p = 0; p <= j; ++p) {
			T t = T(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) t = Num<T>::fma_(V[j][i], V[p][i], t);
			H[j][p] = t;
		}
	}
	T hi[R];
	if (!chol_inplace<T, R>(H, hi)) status |= 16;
	auto solve = [&](T (&z)[R]) {
		OSC_UNROLL for (int i = 0; i < R; ++i) {
			T t = z[i];
			OSC_UNROLL for (int k = 0; k < i; ++k) t = Num<T>::fma_(-H[i][k], z[k], t);
			z[i] = t * hi[i];
		}
		OSC_UNROLL for (int i = R - 1; i >= 0; --i) {
			T t = z[i];
			OSC_UNROLL for (int k = i + 1; k < R; ++k) t = Num<T>::fma_(-H[k][i], z[k], t);
			z[i] = t * hi[i];
		}
	};
	solve(z1);
	if (jc.decoupling != OSC_IMPEDANCE) {
		solve(z2);
		if (beta_j > T(0)) {  // K^-1 x = x + beta k (k . x) / (1 - beta |k|^2), k = Q^T e_n
			T kk = T(0), kx = T(0);
			OSC_UNROLL for (int j = 0; j < R; ++j) { kk = Num<T>::fma_(k_last[j], k_last[j], kk); kx = Num<T>::fma_(k_last[j], z2[j], kx); }
			const T coef = beta_j * kx / (T(1) - beta_j * kk);
			OSC_UNROLL for (int j = 0; j < R; ++j) z2[j] = Num<T>::fma_(coef, k_last[j], z2[j]);
		}
	}
	T sq[N], tau2[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		T t = T(0);
		OSC_UNROLL for (int j = 0; j < R; ++j) t = Num<T>::fma_(SM(F::QQ + j * N + i), z1[j] + z2[j], t);
		sq[i] = t;
	}
	Lmul(sq, tau2);
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		T t = tau1[i] + tau2[i];
		if (gravity) t += (T)rS(RobotL<N>::GRAV + i);
		rS(RobotL<N>::TAU + i) = t;
		if (ex) { mS(ML::TASK_TAU + i) = tau1[i]; jS(JL::TASK_TAU + i) = tau2[i]; }
	}
	}
	if (has_jt) tk1.I[(int64_t)JL::RANK * stride + r] = R;
	if (ex) mS(ML::ALPHA) = alpha;
	if (status) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status);
	return true;
#undef SM
}

}  // namespace osc
