// osc_fused3.cuh -- the single-launch tick for a PARTIAL MotionForceTask of rank M < 6 over a full JointTask (BASELINE config 3:
// translation-only task of examples/09-3d_position_force_controller/09-3d_position_force_controller.cpp:113-121 with force axes
// and passivity-based closed-loop force control, :176-182; MotionForceTask.cpp:33-90, 146-168, 892-971;
// POPCExplicitForceControl.cpp:30-96).  Same structure, certificate and fallback contract as k_tick_fused2 (osc_fused2.cuh); what
// changes is the algebra of the two places where the task rank enters:
//
//  * the task lives in its own range: with B6 = blkdiag(B_t, B_r) the 6 x M orthonormal basis of the controlled directions
//    (P = B6 B6^T is the reference's _partial_task_projection), J_r = B6^T J_world (M x n) is built directly, joint by joint, and
//    Y = L^-1 J_r^T, Y = C W (modified Gram-Schmidt on M rows), tau_1 = L W^T (C^-1 Lambda_mod-part + C^T B6^T F_f) exactly as in
//    the 6-DoF kernel.  Every result the reference exposes is invariant under the choice of basis inside the range
//    (U_ns of SingularityHandler.cpp:123-141 is another basis of the same range).
//  * the joint task below it sees a nullspace of dimension R = n - M > 1.  With Q (n x R) an orthonormal basis of the complement of
//    the rows of W, N_1 = L^-T Q Q^T L^T, V = L^-T Q, H = V^T V:
//        J_p^T Rg Lambda Rg^T      = L Q H^-1 V^T                       (Rg = matrixRangeBasis(N_1), JointTask.cpp:233-245)
//        J_p^T Rg Lambda_BIE Rg^T  = L Q K^-1 H^-1 V^T,  K = I - beta k k^T,  k = Q^T e_n   (last-pivot clamp, JointTask.cpp:253-267)
//        J_p^T Rg Rg^T             = L Q V^T                            (IMPEDANCE)
//    so tau_2 = L Q [ H^-1 V^T (qdd_des - M^-1 tau_1) + (K^-1) H^-1 V^T u_pd ]: one R x R Cholesky instead of the reference's 7 x 7
//    Jacobi SVD.  (R = 1 gives back the u-vector formulas of k_tick_fused2.)
#pragma once
#include "osc_fused2.cuh"

namespace osc {

template <int N, int M> struct F3 {
	static constexpr int R = N - M;
	// one contiguous scratch region that holds, in turn: sin/cos (2N) | F*, F_f (12) | C strictly lower + 1/C_cc -- and, once tau_1's
	// coefficient vector h is known, the complement basis Q (N x R) over all of it
	static constexpr int SC = 0;
	static constexpr int FS = 2 * N;
	static constexpr int CC = FS + 12;
	static constexpr int CI = CC + M * (M - 1) / 2;
	static constexpr int SCRATCH = CI + M;
	static constexpr int QQ = 0;
	static constexpr int SCR_END = SCRATCH > N * R ? SCRATCH : N * R;
	static constexpr int MM = SCR_END;                    // M lower triangle, later L strictly lower + 1/L_ii
	static constexpr int LL = MM;
	static constexpr int LI = MM + N * (N - 1) / 2;
	static constexpr int SLOTS = MM + N * (N + 1) / 2;
};

template <class T, int N, int M, int MAXREG, int BT = OSC_FUSED_THREADS, bool SYNC = false>
__global__ void __launch_bounds__(BT) __maxnreg__(MAXREG) k_tick_fused3(const __grid_constant__ ChainC<T, N> ch, const __grid_constant__ MftC<T> tc,
																			 const __grid_constant__ JtC<T, N> jc, Arena<T> rob, Arena<T> tk0, Arena<T> tk1,
																			 int64_t B, int64_t stride, int stacked, int gravity, int exports,
																			 int* __restrict__ fb_count, int* __restrict__ fb_list) {
	static_assert(M >= 1 && M < 6 && N > M, "partial task: 1 <= M < 6, nullspace dimension n - M >= 1");
	using ML = MftL<N>;
	using JL = JtL<N>;
	using F = F3<N, M>;
	constexpr int R = F::R;
	using A = double;  // accumulation type of the factorisation chain (see k_tick_fused2)
	extern __shared__ __align__(16) unsigned char osc_stash_raw[];
	A* const sm = reinterpret_cast<A*>(osc_stash_raw);
#define SM(slot) sm[(slot) * BT + threadIdx.x]
#define OSC_PHASE()                       \
	OSC_FENCE();                          \
	if constexpr (SYNC) __syncthreads()
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
#define OSC_FRESH_ROWS()                                                                                                            \
	int64_t r_ = r;                                                                                                                 \
	asm volatile("" : "+l"(r_));                                                                                                    \
	const Row<T> rP{rob.P, stride, r_}, rS{rob.S, stride, r_}, mP{tk0.P, stride, r_}, mS{tk0.S, stride, r_}, jP{tk1.P, stride, r_}, \
		jS{tk1.S, stride, r_};                                                                                                      \
	const Row<int32_t> mI{tk0.I, stride, r_};                                                                                       \
	(void)rP; (void)rS; (void)mP; (void)mS; (void)jP; (void)jS; (void)mI
	const bool ex = (exports & OSC_EXPORT_VECTORS) != 0;
	int status = 0;
	constexpr int CLS_RESET = 1 << 30, CLS_LATCH = 1 << 29;

	asm volatile("griddepcontrol.launch_dependents;");

	// ---- P1: q, dq -> stash; sin/cos ----
	{
		OSC_FRESH_ROWS();
		T q[N], dq[N];
		OSC_UNROLL for (int i = 0; i < N; ++i) { q[i] = rP.ro(RobotL<N>::Q + i); dq[i] = rP.ro(RobotL<N>::DQ + i); }
		{
			const int ntypes = mI(ML::NTYPES), hist_len = mI(ML::HIST_LEN);
			if (ntypes != 0 || hist_len != 0) status |= CLS_RESET;
			if (ntypes == 0) status |= CLS_LATCH;
		}
		{
			auto pf = [](const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p) : "memory"); };
			_Pragma("unroll 1") for (int e = ML::GOAL_POS; e < ML::CUR_POS; ++e) pf(&mP(e));
			_Pragma("unroll 1") for (int e = ML::INT_POS; e < ML::POPC_VSQ + 1; ++e) pf(&mP(e));
			_Pragma("unroll 1") for (int e = 0; e < JL::P_SIZE; ++e) pf(&jP(e));
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) { SM(F::SC + 2 * i) = q[i]; SM(F::MM + i) = dq[i]; }
		_Pragma("unroll 1") for (int i = 0; i < N; ++i) {
			T s, c;
			Num<T>::sincos_((T)SM(F::SC + 2 * i), &s, &c);
			SM(F::SC + 2 * i) = s;
			SM(F::SC + 2 * i + 1) = c;
		}
	}

	OSC_PHASE();
	// ---- P1b: rolled forward kinematics for the control law; v, w are the PROJECTED velocities P J dq of a partial task
	// (MotionForceTask.cpp:262, 293-298); then the control law itself (force axes, closed loop, POPC, saturation: .cpp:306-500) ----
	{
		T Rp[3][3], pp[3], w[3] = {T(0), T(0), T(0)}, mo[3] = {T(0), T(0), T(0)};
		OSC_UNROLL for (int a = 0; a < 3; ++a) {
			pp[a] = ch.world_xyz[a];
			OSC_UNROLL for (int b = 0; b < 3; ++b) Rp[a][b] = ch.world_rot[a][b];
		}
		_Pragma("unroll 1") for (int i = 0; i < N; ++i) {
			T pre[3][3], d[3];
			mm3(Rp, ch.rot[i], pre);
			mv3(Rp, ch.xyz[i], d);
			const T s = SM(F::SC + 2 * i), c = SM(F::SC + 2 * i + 1), dqi = SM(F::MM + i);
			OSC_UNROLL for (int a = 0; a < 3; ++a) {
				pp[a] += d[a];
				Rp[a][0] = c * pre[a][0] + s * pre[a][1];
				Rp[a][1] = c * pre[a][1] - s * pre[a][0];
				Rp[a][2] = pre[a][2];
			}
			const T z[3] = {pre[0][2], pre[1][2], pre[2][2]};
			T zxo[3];
			cross3(z, pp, zxo);
			OSC_UNROLL for (int a = 0; a < 3; ++a) { w[a] = Num<T>::fma_(dqi, z[a], w[a]); mo[a] = Num<T>::fma_(dqi, zxo[a], mo[a]); }
		}
		T x[3], Rc[3][3], d[3], vf[3], v[3], wp[3];
		mv3(Rp, tc.p_c, d);
		OSC_UNROLL for (int a = 0; a < 3; ++a) x[a] = pp[a] + d[a];
		mm3(Rp, tc.R_c, Rc);
		cross3(w, x, vf);
		OSC_UNROLL for (int a = 0; a < 3; ++a) vf[a] -= mo[a];
		mv3(tc.Pt, vf, v);
		mv3(tc.Pr, w, wp);
		T Fstar[6], Ff[6];
		OSC_FRESH_ROWS();
		mft_control_law_vw<T, N, false>(tc, mP, mS, mI, stride, r, v, wp, x, Rc, Fstar, Ff, status, ex);
		OSC_UNROLL for (int i = 0; i < 6; ++i) { SM(F::FS + i) = Fstar[i]; SM(F::FS + 6 + i) = Ff[i]; }
	}
#define OSC_FALL_BACK()                                                                      \
	{                                                                                        \
		OSC_UNROLL for (int i_ = 0; i_ < 6; ++i_) { mS(ML::UMF + i_) = SM(F::FS + i_); mS(ML::FFB + i_) = SM(F::FS + 6 + i_); } \
		if (status & ~(CLS_RESET | CLS_LATCH)) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status & ~(CLS_RESET | CLS_LATCH)); \
		fb_list[atomicAdd(fb_count, 1)] = (int)r;                                            \
		return;                                                                              \
	}
	const Row<T> mS{tk0.S, stride, r};

	OSC_PHASE();
	// ---- P2: mass matrix, body-frame composite-rigid-body recursion (identical to k_tick_fused2 P2) ----
	{
		T cm = T(0), h0 = T(0), h1 = T(0), h2 = T(0), Ixx = T(0), Ixy = T(0), Ixz = T(0), Iyy = T(0), Iyz = T(0), Izz = T(0);
		OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
			cm += ch.mass[i];
			h0 += ch.hb[i][0]; h1 += ch.hb[i][1]; h2 += ch.hb[i][2];
			Ixx += ch.Ib[i][0]; Ixy += ch.Ib[i][1]; Ixz += ch.Ib[i][2]; Iyy += ch.Ib[i][3]; Iyz += ch.Ib[i][4]; Izz += ch.Ib[i][5];
			T p0 = -h1, p1 = h0, p2 = T(0), l0 = Ixz, l1 = Iyz, l2 = Izz;
			const int row = F::MM + i * (i + 1) / 2;
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

	OSC_PHASE();
	// ---- P3: M = L L^T; bounded-inertia clamp pattern (last joint only, else the API-faithful route) ----
	A beta = A(0);
	{
		A Mx[N][N], li[N];
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) Mx[i][j] = SM(F::MM + i * (i + 1) / 2 + j);
		A delta = A(0);
		if (tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES || jc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES) {
			const A thr = tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES ? tc.bie_threshold : jc.bie_threshold;
			bool other = false;
			OSC_UNROLL for (int i = 0; i < N - 1; ++i) other = other || (Mx[i][i] < thr);
			if (other) OSC_FALL_BACK()
			if (Mx[N - 1][N - 1] < thr) delta = thr - Mx[N - 1][N - 1];
		}
		if (!chol_inplace<A, N>(Mx, li)) status |= 16;
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			OSC_UNROLL for (int j = 0; j < i; ++j) SM(F::LL + i * (i - 1) / 2 + j) = Mx[i][j];
			SM(F::LI + i) = li[i];
		}
		if (delta > A(0)) {
			const A d = Mx[N - 1][N - 1] * Mx[N - 1][N - 1];
			beta = A(1) - d / (d + delta);
		}
	}

	OSC_PHASE();
	// ---- P4: forward kinematics and the task Jacobian IN THE TASK'S RANGE: row c of J_r is B6[:, c]^T J_world, built joint by
	// joint (linear rows use B_t, angular rows B_r; rows are ordered translation first like _partial_task_projection) ----
	A J[M][N];
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
			OSC_UNROLL for (int c = 0; c < M; ++c) {
				T s = T(0);
				if (c < tc.mt) { OSC_UNROLL for (int a = 0; a < 3; ++a) s = Num<T>::fma_(tc.Bt[a][c], l[a], s); }
				else { OSC_UNROLL for (int a = 0; a < 3; ++a) s = Num<T>::fma_(tc.Br[a][(c - tc.mt) > 2 ? 2 : (c - tc.mt)], Jz[i][a], s); }
				J[c][i] = s;
			}
		}
	}

	OSC_PHASE();
	// ---- P5: branch certificate on the M x M Gram matrix G = J_r J_r^T, FP32 (see osc_fused.cuh): sigma_0 >= s_abs_tol and
	// sigma_{M-1} / sigma_0 > s_max, i.e. no blend for any i < M in the scan of SingularityHandler.cpp:100-142 ----
	{
		float G[M][M];
		OSC_UNROLL for (int i = 0; i < M; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) {
				float s = 0.f;
				OSC_UNROLL for (int k = 0; k < N; ++k) s = fmaf((float)J[i][k], (float)J[j][k], s);
				G[i][j] = s; G[j][i] = s;
			}
		float f4 = 0.f;
		OSC_UNROLL for (int i = 0; i < M; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) {
				float s = 0.f;
				OSC_UNROLL for (int k = 0; k < M; ++k) s = fmaf(G[i][k], G[k][j], s);
				f4 = fmaf(i == j ? s : 2.f * s, s, f4);
			}
		const float lmax_hi = sqrtf(sqrtf(f4)) * 1.001f;  // lambda_max <= lmax_hi <= M^(1/4) lambda_max
		constexpr float inv_root4[6] = {1.f, 1.f, 0.8408964f, 0.7598357f, 0.7071068f, 0.6687403f};  // M^(-1/4)
		const float tol = (float)tc.s_abs_tol, smax = (float)tc.s_max;
		const bool s0_ok = lmax_hi * inv_root4[M] > tol * tol * 1.02f;
		bool sep_ok = true;
		if constexpr (M > 1) sep_ok = spd_above<float, M>(G, smax * smax * lmax_hi * 1.02f);
		if (!(s0_ok && sep_ok)) OSC_FALL_BACK()
	}

	OSC_PHASE();
	// ---- P7: Y = L^-1 J_r^T in place ----
	OSC_UNROLL for (int k = 0; k < N; ++k) {
		A Lk[N];
		OSC_UNROLL for (int m = 0; m < k; ++m) Lk[m] = SM(F::LL + k * (k - 1) / 2 + m);
		const A li = SM(F::LI + k);
		OSC_UNROLL for (int c = 0; c < M; ++c) {
			A s = J[c][k];
			OSC_UNROLL for (int m = 0; m < k; ++m) s = Num<A>::fma_(-Lk[m], J[c][m], s);
			J[c][k] = s * li;
		}
	}
	A (&Y)[M][N] = J;

	OSC_PHASE();
	// ---- P8: Y = C W by modified Gram-Schmidt on the M rows; C (strictly lower + 1/C_cc) to the stash ----
	OSC_UNROLL for (int c = 0; c < M; ++c) {
		A nn = A(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) nn = Num<A>::fma_(Y[c][i], Y[c][i], nn);
		if (!(nn > A(0))) status |= 16;
		const A inv = Num<A>::rsqrt_(nn);
		SM(F::CI + c) = inv;
		OSC_UNROLL for (int i = 0; i < N; ++i) Y[c][i] *= inv;
		OSC_UNROLL for (int c2 = c + 1; c2 < M; ++c2) {
			A rr = A(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) rr = Num<A>::fma_(Y[c][i], Y[c2][i], rr);
			SM(F::CC + c2 * (c2 - 1) / 2 + c) = rr;
			OSC_UNROLL for (int i = 0; i < N; ++i) Y[c2][i] = Num<A>::fma_(-rr, Y[c][i], Y[c2][i]);
		}
	}
	A (&W)[M][N] = Y;

	OSC_PHASE();
	// ---- P10: tau_1 = J_r^T (Lambda_mod B6^T F* + B6^T F_f) = L h,  h = W^T g,  g = C^-1 f (+ Sherman-Morrison for the bounded
	// inertia estimate) + C^T ff   (SingularityHandler.cpp:160-225, 307-309 in the task's range) ----
	A h[N];
	{
		A f[M], ff[M], g[M];
		OSC_UNROLL for (int c = 0; c < M; ++c) {  // B6^T F*, B6^T F_f
			A s = A(0), t = A(0);
			if (c < tc.mt) {
				OSC_UNROLL for (int a = 0; a < 3; ++a) { s = Num<A>::fma_((A)tc.Bt[a][c], SM(F::FS + a), s); t = Num<A>::fma_((A)tc.Bt[a][c], SM(F::FS + 6 + a), t); }
			} else {
				const int cr = (c - tc.mt) > 2 ? 2 : (c - tc.mt);
				OSC_UNROLL for (int a = 0; a < 3; ++a) { s = Num<A>::fma_((A)tc.Br[a][cr], SM(F::FS + 3 + a), s); t = Num<A>::fma_((A)tc.Br[a][cr], SM(F::FS + 9 + a), t); }
			}
			f[c] = s; ff[c] = t;
		}
		if (tc.decoupling != OSC_IMPEDANCE) {
			OSC_UNROLL for (int i = 0; i < M; ++i) {  // g <- C^-1 f
				A s = f[i];
				OSC_UNROLL for (int m = 0; m < i; ++m) s = Num<A>::fma_(-SM(F::CC + i * (i - 1) / 2 + m), g[m], s);
				g[i] = s * SM(F::CI + i);
			}
			if (tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES && beta > A(0)) {
				A ww = A(0), wg = A(0);
				OSC_UNROLL for (int c = 0; c < M; ++c) { ww = Num<A>::fma_(W[c][N - 1], W[c][N - 1], ww); wg = Num<A>::fma_(W[c][N - 1], g[c], wg); }
				const A coef = beta * wg / (A(1) - beta * ww);
				OSC_UNROLL for (int c = 0; c < M; ++c) g[c] = Num<A>::fma_(coef, W[c][N - 1], g[c]);
			}
		} else {
			OSC_UNROLL for (int c = 0; c < M; ++c) { g[c] = A(0); ff[c] += f[c]; }  // Lambda_mod = I: C^T (f + ff)
		}
		OSC_UNROLL for (int c = 0; c < M; ++c) {  // g += C^T ff
			A s = ff[c] / SM(F::CI + c);
			OSC_UNROLL for (int c2 = c + 1; c2 < M; ++c2) s = Num<A>::fma_(SM(F::CC + c2 * (c2 - 1) / 2 + c), ff[c2], s);
			g[c] += s;
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			A s = A(0);
			OSC_UNROLL for (int c = 0; c < M; ++c) s = Num<A>::fma_(W[c][i], g[c], s);
			h[i] = s;
		}
	}
	// sin/cos, F*, F_f and C are dead: their slots receive Q.

	OSC_PHASE();
	// ---- P9: Q = orthonormal basis of the complement of the rows of W (n x R), by pivoted Gram-Schmidt on the unit vectors: each
	// step takes the e_k with the largest residual |(I - W^T W - Q Q^T) e_k|^2, projects it twice, normalises ----
	A k_last[R];  // last row of Q (k = Q^T e_n of the bounded-inertia correction)
	{
		A res[N];  // residual diagonal P_kk
		OSC_UNROLL for (int k = 0; k < N; ++k) {
			A s = A(1);
			OSC_UNROLL for (int c = 0; c < M; ++c) s = Num<A>::fma_(-W[c][k], W[c][k], s);
			res[k] = s;
		}
		OSC_UNROLL for (int j = 0; j < R; ++j) {
			A best = A(-1);
			int kb = 0;
			OSC_UNROLL for (int k = 0; k < N; ++k)
				if (res[k] > best) { best = res[k]; kb = k; }
			A u[N];
			OSC_UNROLL for (int i = 0; i < N; ++i) u[i] = (i == kb) ? A(1) : A(0);
			OSC_UNROLL for (int pass = 0; pass < 2; ++pass) {
				A a[M];
				OSC_UNROLL for (int c = 0; c < M; ++c) {
					A s = A(0);
					OSC_UNROLL for (int i = 0; i < N; ++i) s = Num<A>::fma_(W[c][i], u[i], s);
					a[c] = s;
				}
				OSC_UNROLL for (int i = 0; i < N; ++i) {
					A s = u[i];
					OSC_UNROLL for (int c = 0; c < M; ++c) s = Num<A>::fma_(-W[c][i], a[c], s);
					u[i] = s;
				}
				OSC_UNROLL for (int p = 0; p < R; ++p) {
					if (p < j) {
						A d = A(0);
						OSC_UNROLL for (int i = 0; i < N; ++i) d = Num<A>::fma_(SM(F::QQ + p * N + i), u[i], d);
						OSC_UNROLL for (int i = 0; i < N; ++i) u[i] = Num<A>::fma_(-d, SM(F::QQ + p * N + i), u[i]);
					}
				}
			}
			A nu = A(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) nu = Num<A>::fma_(u[i], u[i], nu);
			const A inv = Num<A>::rsqrt_(nu);
			OSC_UNROLL for (int i = 0; i < N; ++i) {
				u[i] *= inv;
				SM(F::QQ + j * N + i) = u[i];
				res[i] = Num<A>::fma_(-u[i], u[i], res[i]);
			}
			k_last[j] = u[N - 1];
		}
	}
	// W (= Y = J) is dead from here on.

	OSC_PHASE();
	{
	OSC_FRESH_ROWS();
	bool latch = (status & CLS_LATCH) != 0;
	if (status & CLS_RESET) {
		latch = latch || mI(ML::C2) > mI(ML::C1);
		mI(ML::NTYPES) = 0; mI(ML::TYPES) = 0; mI(ML::C1) = 0; mI(ML::C2) = 0; mI(ML::HIST_LEN) = 0; mI(ML::HIST_HEAD) = 0;
	}
	latch = latch && ex;
	// ---- P11: joint task in the R-dimensional nullspace ----
	// tau_1 = L h (exported per task), M^-1 tau_1 = L^-T h
	A mt[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) mt[i] = h[i];
	OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
		A st = mt[i];
		OSC_UNROLL for (int k = i + 1; k < N; ++k) st = Num<A>::fma_(-SM(F::LL + k * (k - 1) / 2 + i), mt[k], st);
		mt[i] = st * SM(F::LI + i);
	}
	// y1 = qdd_des - M^-1 tau_1 (stacked), u_pd: the joint PD law (JointTask.cpp:319-346)
	A y1[N], upd_[N];
	{
		T q[N], dq[N], upd[N], acc[N];
		ld_vec<T, N>(rP, RobotL<N>::Q, q);
		ld_vec<T, N>(rP, RobotL<N>::DQ, dq);
		jt_pd_law<T, N>(jc, jP, jS, q, dq, upd, acc, ex);
		OSC_UNROLL for (int i = 0; i < N; ++i) { y1[i] = (A)acc[i] - (stacked ? mt[i] : A(0)); upd_[i] = upd[i]; }
		if (ex) {
			st_vec<T, N>(jS, JL::CUR_POS, q);
			st_vec<T, N>(jS, JL::CUR_VEL, dq);
		}
		if (latch) {
			OSC_UNROLL for (int i = 0; i < N; ++i) mP(ML::Q_PRIOR + i) = q[i];
		}
	}
	// V = L^-T Q is never held: one column at a time, v_p = L^-T q_p gives z1[p] = v_p . y1 and z2[p] = v_p . u_pd, then
	// w_p = L^-1 v_p gives the column H[:, p] = V^T v_p = Q^T w_p with Q read back from the stash (R x 7 doubles of registers less:
	// the R = 4 instantiation of config 3 spilled 0.9 KB per thread and was bound by that traffic, r2e)
	A z1[R], z2[R], H[R][R];
	OSC_UNROLL for (int p = 0; p < R; ++p) {
		A v[N];
		OSC_UNROLL for (int i = 0; i < N; ++i) v[i] = SM(F::QQ + p * N + i);
		OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
			A s = v[i];
			OSC_UNROLL for (int k = i + 1; k < N; ++k) s = Num<A>::fma_(-SM(F::LL + k * (k - 1) / 2 + i), v[k], s);
			v[i] = s * SM(F::LI + i);
		}
		A a = A(0), b = A(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) { a = Num<A>::fma_(v[i], y1[i], a); b = Num<A>::fma_(v[i], upd_[i], b); }
		z1[p] = a; z2[p] = b;
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			A s = v[i];
			OSC_UNROLL for (int m = 0; m < i; ++m) s = Num<A>::fma_(-SM(F::LL + i * (i - 1) / 2 + m), v[m], s);
			v[i] = s * SM(F::LI + i);
		}
		OSC_UNROLL for (int j = p; j < R; ++j) {
			A s = A(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) s = Num<A>::fma_(SM(F::QQ + j * N + i), v[i], s);
			H[j][p] = s;
		}
	}
	// H = Lh Lh^T; a1 = H^-1 z1, a2 = H^-1 z2
	A hi[R];
	if (!chol_inplace<A, R>(H, hi)) status |= 16;
	auto solve = [&](A (&z)[R]) {
		OSC_UNROLL for (int i = 0; i < R; ++i) {
			A s = z[i];
			OSC_UNROLL for (int k = 0; k < i; ++k) s = Num<A>::fma_(-H[i][k], z[k], s);
			z[i] = s * hi[i];
		}
		OSC_UNROLL for (int i = R - 1; i >= 0; --i) {
			A s = z[i];
			OSC_UNROLL for (int k = i + 1; k < R; ++k) s = Num<A>::fma_(-H[k][i], z[k], s);
			z[i] = s * hi[i];
		}
	};
	solve(z1);
	if (jc.decoupling != OSC_IMPEDANCE) {
		solve(z2);
		if (jc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES && beta > A(0)) {  // K^-1 x = x + beta k (k . x) / (1 - beta |k|^2)
			A kk = A(0), kx = A(0);
			OSC_UNROLL for (int j = 0; j < R; ++j) { kk = Num<A>::fma_(k_last[j], k_last[j], kk); kx = Num<A>::fma_(k_last[j], z2[j], kx); }
			const A coef = beta * kx / (A(1) - beta * kk);
			OSC_UNROLL for (int j = 0; j < R; ++j) z2[j] = Num<A>::fma_(coef, k_last[j], z2[j]);
		}
	}
	// s = Q (a1 + a2); tau_2 = L s, tau_1 = L h, one sweep over L for both
	A sq[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		A s = A(0);
		OSC_UNROLL for (int j = 0; j < R; ++j) s = Num<A>::fma_(SM(F::QQ + j * N + i), z1[j] + z2[j], s);
		sq[i] = s;
	}
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		const A ld = A(1) / SM(F::LI + i);
		A s1 = ld * h[i], s2 = ld * sq[i];
		OSC_UNROLL for (int m = 0; m < i; ++m) {
			const A l = SM(F::LL + i * (i - 1) / 2 + m);
			s1 = Num<A>::fma_(l, h[m], s1);
			s2 = Num<A>::fma_(l, sq[m], s2);
		}
		A t = s1 + s2;
		if (gravity) t += rS(RobotL<N>::GRAV + i);
		rS(RobotL<N>::TAU + i) = t;
		if (ex) { mS(ML::TASK_TAU + i) = s1; jS(JL::TASK_TAU + i) = s2; }
	}
	if (ex) {
		tk1.I[(int64_t)JL::RANK * stride + r] = R;
		mS(ML::ALPHA) = T(1);
		mI(ML::INS) = M;
	}
	status &= ~(CLS_RESET | CLS_LATCH);
	if (status) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status);
	}
#undef OSC_FRESH_ROWS
#undef OSC_PHASE
#undef OSC_FALL_BACK
#undef SM
}

// hierarchy / parameters the partial-task kernel serves: [partial MotionForceTask of rank M, full JointTask], revolute chain, control
// link on the last body (the instantiated (N, M) pairs are listed in osc_ops_fields.inl)
inline bool fused3_supported(Batch& b, int& M) {
	if (b.tasks.size() != 2) return false;
	const TaskH &m = b.tasks[0], &j = b.tasks[1];
	if (m.type != OSC_TASK_MOTION_FORCE || j.type != OSC_TASK_JOINT || !m.md.partial || !j.jd.full) return false;
	if (!fused2_supported(b)) return false;
	M = m.dof;
	return true;
}

}  // namespace osc
