// osc_fused2.cuh -- second generation of the single-launch tick (see osc_fused.cuh for the algebra and the contract).
//
// Same mathematics as k_tick_fused, re-engineered around what ncu showed on B200 (profiles/README.md, r1b/r1c):
// the first version was a 124 KB straight-line kernel with 1.7-2.2 KB of local-memory spills per thread; at 16 warps/SM
// the spills miss L1 (36 % hit rate) and the warps sit on the long scoreboard, at 8 warps/SM on instruction fetch.
//   * no local memory: a 57-slot per-thread stash in SHARED memory (slot-major, conflict free) holds, in turn,
//     sin/cos -> F*, F_f | M -> L, 1/L_ii | chol(J M^-1 J^T); registers hold J (overwritten in place by Y = L^-1 J^T);
//   * the mass matrix comes from a ROLLED body-frame composite-rigid-body recursion (needs only sin/cos, 250 instructions
//     instead of 2,500); forward kinematics is unrolled but calls one out-of-line sincos;
//   * bounded inertia estimates without a second factorisation: when only the last diagonal entry of M is clamped (always
//     for the Panda: M_77 = 0.08 < 0.1) L_BIE differs from L in its last pivot only, so A_BIE = A - beta y y^T and
//     A_BIE^-1 F* follows from Sherman-Morrison; any other clamp pattern takes the API-faithful route;
//   * the branch certificate runs in FP32 on the FP32 pipe (it is a bound with a 2 % margin, not a result);
//   * M^-1 tau_1 = L^-T (Y f): the Jacobian is not needed after Y is formed (tau_1 = L (Y f)).
#pragma once
#include "osc_fused.cuh"

namespace osc {

template <int N> struct F2 {
	static constexpr int SC = 0;                          // sin/cos (2N), later F* (6) and F_f (6)
	static constexpr int MM = 2 * N;                      // M lower triangle (N(N+1)/2), later L strictly lower + 1/L_ii
	static constexpr int LL = MM;                         // L[i][j], j < i at LL + i(i-1)/2 + j
	static constexpr int LI = MM + N * (N - 1) / 2;       // 1 / L_ii
	static constexpr int CC = MM + N * (N + 1) / 2;       // chol(A) strictly lower (15)
	// 1 / C_ii: the spare slots behind F*, F_f first (2N - 12), the rest behind C
	static constexpr int SPARE = 2 * N - 12;
	__host__ __device__ static constexpr int cli(int i) { return i < SPARE ? SC + 12 + i : CC + 15 + (i - SPARE); }
	// SIMPLE kernel (no C stored): F* at CC..CC+5, the joint PID command u_pd (N) and the joint goal acceleration (N) behind it
	static constexpr int UPD = CC + 6;
	static constexpr int GAC = CC + 6 + N;
	static constexpr int SLOTS_GENERAL = CC + 15 + (6 - SPARE);
	static constexpr int SLOTS = SLOTS_GENERAL > GAC + N ? SLOTS_GENERAL : GAC + N;
};

template <class T> __device__ __noinline__ void sincos_outlined(T x, T* s, T* c) { Num<T>::sincos_(x, s, c); }

// SIMPLE = plain motion task over a plain joint task (no force / moment space, open loop, no velocity saturation): the
// host selects this instantiation from the uniform task parameters; F_f is identically zero there.
// BT threads per block; SYNC: block-wide barriers at the phase boundaries keep the warps of a block inside the same stretch of
// this 100 KB straight-line kernel, so that they share instruction-cache lines (r1k: instruction fetch was the top stall).
// SINGLE (n = 7): the MotionForceTask is the whole hierarchy (examples/07-surface_surface_contact: RobotController over one task):
// no joint task below it -- every joint-task argument aliases the motion task and is never touched, tau = tau_1.
template <class T, int N, int MAXREG, bool SIMPLE, int BT = OSC_FUSED_THREADS, bool SYNC = false, bool SINGLE = false>
__global__ void __launch_bounds__(BT) __maxnreg__(MAXREG) k_tick_fused2(const __grid_constant__ ChainC<T, N> ch, const __grid_constant__ MftC<T> tc,
																		 const __grid_constant__ JtC<T, N> jc, Arena<T> rob, Arena<T> tk0, Arena<T> tk1,
																		 int64_t B, int64_t stride, int stacked, int gravity, int exports,
																		 int* __restrict__ fb_count, int* __restrict__ fb_list) {
	static_assert(N == 6 || N == 7, "the fused tick covers n - 6 <= 1");
	static_assert(!SINGLE || N == 7, "n = 6 under one task needs no instantiation of its own: it has no nullspace");
	constexpr bool JT7 = N == 7 && !SINGLE;  // a full JointTask in the one-dimensional nullspace of the motion task
	using ML = MftL<N>;
	using JL = JtL<N>;
	using F = F2<N>;
	// Mixed precision of an FP32 batch: state in HBM, sin/cos, kinematics, the control law, the mass-matrix recursion and the
	// branch certificate are FP32; the factorisation chain (Cholesky of M, Y = L^-1 J^T, Gram-Schmidt, the triangular sweeps
	// of P10/P11) accumulates in FP64, because its rounding error is amplified by kappa(M) kappa(J)^2 (1e3-1e5 for the Panda)
	// and would otherwise eat the 1e-4 FP32 tolerance of the north star.  A = T for an FP64 batch: nothing changes there.
	using A = double;
	extern __shared__ __align__(16) unsigned char osc_stash_raw[];
	A* const sm = reinterpret_cast<A*>(osc_stash_raw);
#define SM(slot) sm[(slot) * BT + threadIdx.x]
#define OSC_PHASE()                       \
	OSC_FENCE();                          \
	if constexpr (SYNC) __syncthreads()
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T> rP{rob.P, stride, r}, rS{rob.S, stride, r}, mP{tk0.P, stride, r}, mS{tk0.S, stride, r}, jP{tk1.P, stride, r}, jS{tk1.S, stride, r};
	const Row<int32_t> mI{tk0.I, stride, r};
	// Row addresses are pure functions of r: left alone, the compiler computes them once at the top and SPILLS the 64-bit
	// addresses across the whole kernel (r1g: LDL of addresses ahead of every late load/store).  Each memory phase therefore
	// re-derives its rows from a laundered copy of r.
#define OSC_FRESH_ROWS()                                                                                                            \
	int64_t r_ = r;                                                                                                                 \
	asm volatile("" : "+l"(r_));                                                                                                    \
	const Row<T> rP{rob.P, stride, r_}, rS{rob.S, stride, r_}, mP{tk0.P, stride, r_}, mS{tk0.S, stride, r_}, jP{tk1.P, stride, r_}, \
		jS{tk1.S, stride, r_};                                                                                                      \
	const Row<int32_t> mI{tk0.I, stride, r_};                                                                                       \
	(void)rP; (void)rS; (void)mP; (void)mS; (void)jP; (void)jS; (void)mI
	const bool ex = (exports & OSC_EXPORT_VECTORS) != 0;
	int status = 0;
	constexpr int CLS_RESET = 1 << 30, CLS_LATCH = 1 << 29;  // private flags riding in `status`, masked out before it is published

	asm volatile("griddepcontrol.launch_dependents;");  // lets the launch of k_tick_tail proceed under this kernel

	// ---- P1: sin/cos of every joint.  q and dq are fetched as ONE batch of independent loads into the stash (a load per
	// rolled iteration would pay a DRAM round trip each: profiles/README.md r1f) ----
	if constexpr (SIMPLE && JT7) {
		// ... together with every input of the joint task: its PID command u_pd and the goal acceleration wait in the stash for
		// the range basis (P11), so that the kernel ends without a load phase.  The integrated error is stored here; a robot
		// that falls back tells k_tick_tail not to integrate again (reuse_law bit 1).
		// All 44 loads are issued back to back (a 22 MB burst over the whole grid: 3-4 us of pure HBM time since every warp
		// starts in lock step); q comes first and the sin/cos loop runs while the rest of the burst is still in flight.
		T q[N], dq[N], gp[N], gv[N], ga[N], ig[N];
		OSC_UNROLL for (int i = 0; i < N; ++i) q[i] = rP.ro(RobotL<N>::Q + i);
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			dq[i] = rP.ro(RobotL<N>::DQ + i);
			gp[i] = jP.ro(JL::GOAL_POS + i); gv[i] = jP.ro(JL::GOAL_VEL + i); ga[i] = jP.ro(JL::GOAL_ACC + i);
			ig[i] = jP(JL::INTEG + i);
		}
		{
			const int ntypes = mI(ML::NTYPES), hist_len = mI(ML::HIST_LEN);  // classification state: decided here, used after P10
			if (ntypes != 0 || hist_len != 0) status |= CLS_RESET;
			if (ntypes == 0) status |= CLS_LATCH;
		}
		{  // P0 (issued BEHIND the loads above so that q is not queued after them): pull the control-law rows towards L2
			auto pf = [](const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p) : "memory"); };
			_Pragma("unroll 1") for (int e = ML::GOAL_POS; e < (SIMPLE ? ML::GOAL_FORCE : ML::CUR_POS); ++e) pf(&mP(e));
			_Pragma("unroll 1") for (int e = ML::INT_POS; e < (SIMPLE ? ML::INT_F : ML::POPC_VSQ + 1); ++e) pf(&mP(e));
			if (!(SIMPLE && N == 7) && !SINGLE) _Pragma("unroll 1") for (int e = 0; e < JL::P_SIZE; ++e) pf(&jP(e));
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) SM(F::SC + 2 * i) = q[i];
		_Pragma("unroll 1") for (int i = 0; i < N; ++i) {
			T s, c;
			Num<T>::sincos_(SM(F::SC + 2 * i), &s, &c);
			SM(F::SC + 2 * i) = s;
			SM(F::SC + 2 * i + 1) = c;
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) {  // JointTask.cpp:319-346 without velocity saturation
			SM(F::MM + i) = dq[i];
			const T e = q[i] - gp[i];
			ig[i] = Num<T>::fma_(e, jc.dt, ig[i]);
			SM(F::UPD + i) = -jc.kp[i] * e - jc.kv[i] * (dq[i] - gv[i]) - jc.ki[i] * ig[i];
			SM(F::GAC + i) = ga[i];
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) jP(JL::INTEG + i) = ig[i];
		if (ex) {
			OSC_UNROLL for (int i = 0; i < N; ++i) { jS(JL::CUR_POS + i) = q[i]; jS(JL::CUR_VEL + i) = dq[i]; jS(JL::DES_VEL + i) = gv[i]; }
		}
	} else {
		T q[N], dq[N];
		OSC_UNROLL for (int i = 0; i < N; ++i) { q[i] = rP.ro(RobotL<N>::Q + i); dq[i] = rP.ro(RobotL<N>::DQ + i); }
		{
			const int ntypes = mI(ML::NTYPES), hist_len = mI(ML::HIST_LEN);
			if (ntypes != 0 || hist_len != 0) status |= CLS_RESET;
			if (ntypes == 0) status |= CLS_LATCH;
		}
		{  // P0 (issued BEHIND the loads above so that q is not queued after them): pull the control-law rows towards L2
			auto pf = [](const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p) : "memory"); };
			_Pragma("unroll 1") for (int e = ML::GOAL_POS; e < (SIMPLE ? ML::GOAL_FORCE : ML::CUR_POS); ++e) pf(&mP(e));
			_Pragma("unroll 1") for (int e = ML::INT_POS; e < (SIMPLE ? ML::INT_F : ML::POPC_VSQ + 1); ++e) pf(&mP(e));
			if (!(SIMPLE && N == 7) && !SINGLE) _Pragma("unroll 1") for (int e = 0; e < JL::P_SIZE; ++e) pf(&jP(e));
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) { SM(F::SC + 2 * i) = q[i]; SM(F::MM + i) = dq[i]; }
		_Pragma("unroll 1") for (int i = 0; i < N; ++i) {
			T s, c;
			Num<T>::sincos_(SM(F::SC + 2 * i), &s, &c);
			SM(F::SC + 2 * i) = s;
			SM(F::SC + 2 * i + 1) = c;
		}
	}

	OSC_PHASE();
	// ---- P1b: a ROLLED forward-kinematics pass for what the control law needs: x, R of the compliant frame and its twist
	// v = w x x - sum_i dq_i (z_i x o_i), w = sum_i dq_i z_i (= J dq, MotionForceTask.cpp:293-298) -- then the control law itself
	// (MotionForceTask.cpp:278-509) while nothing else is live.  F*, F_f are parked in the C slots until sin/cos are dead.
	// The control law runs BEFORE the branch certificate: a robot that later falls back hands F*, F_f to k_tick_tail
	// (ML::UMF, ML::FFB) instead of running the law a second time.
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
		T x[3], Rc[3][3], d[3], v[3];
		mv3(Rp, tc.p_c, d);
		OSC_UNROLL for (int a = 0; a < 3; ++a) x[a] = pp[a] + d[a];
		mm3(Rp, tc.R_c, Rc);
		cross3(w, x, v);
		OSC_UNROLL for (int a = 0; a < 3; ++a) v[a] -= mo[a];
		T Fstar[6], Ff[6];
		OSC_FRESH_ROWS();
		if constexpr (SIMPLE) {
			mft_control_law_motion_only<T, N>(tc, mP, mS, v, w, x, Rc, Fstar, ex);
			OSC_UNROLL for (int i = 0; i < 6; ++i) Ff[i] = T(0);
		} else {
			mft_control_law_vw<T, N, true>(tc, mP, mS, mI, stride, r, v, w, x, Rc, Fstar, Ff, status, ex);
		}
		OSC_UNROLL for (int i = 0; i < 6; ++i) {
			SM(F::CC + i) = Fstar[i];
			if constexpr (!SIMPLE) SM(F::CC + 6 + i) = Ff[i];
		}
	}
#define OSC_FALL_BACK()                                                                      \
	{                                                                                        \
		OSC_UNROLL for (int i_ = 0; i_ < 6; ++i_) { mS(ML::UMF + i_) = SM(F::CC + i_); mS(ML::FFB + i_) = SIMPLE ? T(0) : SM(F::CC + 6 + i_); } \
		if (status & ~(CLS_RESET | CLS_LATCH)) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status & ~(CLS_RESET | CLS_LATCH)); \
		fb_list[atomicAdd(fb_count, 1)] = (int)r;                                            \
		return;                                                                              \
	}

	OSC_PHASE();
	// ---- P2: mass matrix, composite rigid body recursion in BODY frames, tip to base (Sai2Model::updateModel) ----
	// composite of bodies >= i in frame i: mass cm, first moment h, inertia about the frame origin I (xx xy xz yy yz zz).
	// column i: momentum (p, l) of the composite under the unit joint twist (omega = z, v = 0), carried down the chain;
	// M[i][j] = z . l in frame j.  Frame change child -> parent: x_parent = xyz + rot * Rz(q) * x_child.
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
			if (i > 0) {  // express the composite in the parent frame
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
	// ---- P3: M = L L^T in place in registers, then parked in the stash; bounded-inertia clamp pattern ----
	A rho2 = A(1), beta = A(0);  // BIE: L_BIE = L except its last pivot; rho = L_nn / L_BIE,nn, beta = 1 - rho^2
	{
		A Mx[N][N], li[N];
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) Mx[i][j] = SM(F::MM + i * (i + 1) / 2 + j);
		A delta = A(0);
		if constexpr (N == 6) {
			// n = 6 under a full task: no nullspace, tau = J^T (Lambda_mod F* + F_f) is the only result, and with L_BIE = chol(M_BIE)
			// in place of L every step below yields exactly that (Y = L_BIE^-1 J^T = C W, tau = L_BIE W^T (C^-1 F* + C^T F_f)) for
			// ANY clamp pattern (SingularityHandler.cpp:175-182) -- the PUMA clamps its three wrist joints
			if (tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES) {
				OSC_UNROLL for (int i = 0; i < N; ++i)
					if (Mx[i][i] < (A)tc.bie_threshold) Mx[i][i] = (A)tc.bie_threshold;
			}
		} else
		if (tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES || (!SINGLE && jc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES)) {
			const A thr = tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES ? tc.bie_threshold : jc.bie_threshold;
			bool other = false;
			OSC_UNROLL for (int i = 0; i < N - 1; ++i) other = other || (Mx[i][i] < thr);
			if (other) OSC_FALL_BACK()  // a clamp that is not confined to the last joint: API-faithful route
			if (Mx[N - 1][N - 1] < thr) delta = thr - Mx[N - 1][N - 1];
		}
		if (!chol_inplace<A, N>(Mx, li)) status |= 16;
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			OSC_UNROLL for (int j = 0; j < i; ++j) SM(F::LL + i * (i - 1) / 2 + j) = Mx[i][j];
			SM(F::LI + i) = li[i];
		}
		if (delta > A(0)) {
			const A d = Mx[N - 1][N - 1] * Mx[N - 1][N - 1];
			rho2 = d / (d + delta);
			beta = A(1) - rho2;
		}
	}

	OSC_PHASE();
	// ---- P4: forward kinematics (revolute chain, control link on the last body) and the Jacobian at the compliant frame ----
	A J[6][N];  // built in the batch type, held in the accumulation type: Y = L^-1 J^T overwrites it in place
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
			OSC_UNROLL for (int a = 0; a < 3; ++a) { J[a][i] = l[a]; J[3 + a][i] = Jz[i][a]; }
		}
	}

	OSC_PHASE();
	// ---- P5: branch certificate in FP32 (a bound with a 2 % margin; see osc_fused.cuh) ----
	{
		float G[6][6];
		OSC_UNROLL for (int i = 0; i < 6; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) {
				float s = 0.f;
				OSC_UNROLL for (int k = 0; k < N; ++k) s = fmaf((float)J[i][k], (float)J[j][k], s);
				G[i][j] = s; G[j][i] = s;
			}
		float f4 = 0.f;  // tr(G^4) = |G^2|_F^2
		OSC_UNROLL for (int i = 0; i < 6; ++i)
			OSC_UNROLL for (int j = 0; j <= i; ++j) {
				float s = 0.f;
				OSC_UNROLL for (int k = 0; k < 6; ++k) s = fmaf(G[i][k], G[k][j], s);
				f4 = fmaf(i == j ? s : 2.f * s, s, f4);
			}
		const float lmax_hi = sqrtf(sqrtf(f4)) * 1.001f;  // lambda_max <= lmax_hi <= 6^(1/4) lambda_max
		const float tol = (float)tc.s_abs_tol, smax = (float)tc.s_max;
		const bool s0_ok = lmax_hi * 0.6389431f > tol * tol * 1.02f;
		const bool sep_ok = spd_above<float, 6>(G, smax * smax * lmax_hi * 1.02f);
		if (!(s0_ok && sep_ok)) OSC_FALL_BACK()  // not provably non-singular: API-faithful route
	}
	// sin/cos are dead: move F*, F_f into their slots (the C slots are needed by P8; the SIMPLE kernel stores no C)
	if constexpr (!SIMPLE) {
		OSC_UNROLL for (int i = 0; i < 12; ++i) SM(F::SC + i) = SM(F::CC + i);
	}

	OSC_PHASE();
	// ---- P7: Y = L^-1 J^T in place (J[c][:] <- L^-1 J[c][:]), L streamed row by row from the stash ----
	OSC_UNROLL for (int k = 0; k < N; ++k) {
		A Lk[N];
		OSC_UNROLL for (int m = 0; m < k; ++m) Lk[m] = SM(F::LL + k * (k - 1) / 2 + m);
		const A li = SM(F::LI + k);
		OSC_UNROLL for (int c = 0; c < 6; ++c) {
			A s = J[c][k];
			OSC_UNROLL for (int m = 0; m < k; ++m) s = Num<A>::fma_(-Lk[m], J[c][m], s);
			J[c][k] = s * li;
		}
	}
	A (&Y)[6][N] = J;

	OSC_PHASE();
	// ---- P8: orthonormalise the rows of Y in place (modified Gram-Schmidt): Y = C W with W W^T = I and C lower triangular,
	// i.e. C = chol(A) for A = Y Y^T = J M^-1 J^T without ever forming A (and without squaring its condition number).
	// C goes to the stash (strictly lower part + 1/C_cc); W stays in the registers of Y. ----
	// The SIMPLE kernel carries F* through the same row operations as an extra column: C^-1 [Y | F*] = [W | g] with
	// g = C^-1 F* = C^T Lambda F*, which is all P10 needs of C -- so C is not stored at all there.
	A g[6];
	if constexpr (SIMPLE) {
		OSC_UNROLL for (int c = 0; c < 6; ++c) g[c] = SM(F::CC + c);
	}
	OSC_UNROLL for (int c = 0; c < 6; ++c) {
		A nn = A(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) nn = Num<A>::fma_(Y[c][i], Y[c][i], nn);
		if (!(nn > A(0))) status |= 16;
		const A inv = Num<A>::rsqrt_(nn);
		if constexpr (SIMPLE) g[c] *= inv;
		else SM(F::cli(c)) = inv;
		OSC_UNROLL for (int i = 0; i < N; ++i) Y[c][i] *= inv;
		OSC_UNROLL for (int c2 = c + 1; c2 < 6; ++c2) {
			A rr = A(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) rr = Num<A>::fma_(Y[c][i], Y[c2][i], rr);
			if constexpr (SIMPLE) g[c2] = Num<A>::fma_(-rr, g[c], g[c2]);
			else SM(F::CC + c2 * (c2 - 1) / 2 + c) = rr;
			OSC_UNROLL for (int i = 0; i < N; ++i) Y[c2][i] = Num<A>::fma_(-rr, Y[c][i], Y[c2][i]);
		}
	}
	A (&W)[6][N] = Y;

	OSC_PHASE();
	// ---- P9: u = unit vector orthogonal to the rows of W (n = 7): P = I - W^T W = u u^T, N_1 = (L^-T u)(L u)^T ----
	A u[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) u[i] = A(0);
	if constexpr (JT7) {
		A best = A(-1);
		int kb = 0;
		OSC_UNROLL for (int k = 0; k < N; ++k) {  // residual of e_k: P_kk = 1 - |W[:,k]|^2
			A s = A(1);
			OSC_UNROLL for (int c = 0; c < 6; ++c) s = Num<A>::fma_(-W[c][k], W[c][k], s);
			if (s > best) { best = s; kb = k; }
		}
		A a[6];
		OSC_UNROLL for (int c = 0; c < 6; ++c) {
			A s = A(0);
			OSC_UNROLL for (int k = 0; k < N; ++k) s = (k == kb) ? W[c][k] : s;
			a[c] = s;
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			A s = (i == kb) ? A(1) : A(0);
			OSC_UNROLL for (int c = 0; c < 6; ++c) s = Num<A>::fma_(-W[c][i], a[c], s);
			u[i] = s;
		}
		// second projection ("twice is enough")
		OSC_UNROLL for (int c = 0; c < 6; ++c) {
			A s = A(0);
			OSC_UNROLL for (int i = 0; i < N; ++i) s = Num<A>::fma_(W[c][i], u[i], s);
			a[c] = s;
		}
		A nu = A(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			A s = u[i];
			OSC_UNROLL for (int c = 0; c < 6; ++c) s = Num<A>::fma_(-W[c][i], a[c], s);
			u[i] = s;
			nu = Num<A>::fma_(s, s, nu);
		}
		const A inv = Num<A>::rsqrt_(nu);
		OSC_UNROLL for (int i = 0; i < N; ++i) u[i] *= inv;
	}

	OSC_PHASE();
	// ---- P10: tau_1 = J^T (Lambda_mod F* + F_f) = L h with h = W^T g,  g = C^T (Lambda_mod F* + F_f)
	// (SingularityHandler.cpp:160-225, 307-309).  Lambda = A^-1 = C^-T C^-1, so C^T Lambda F* = C^-1 F*: one forward solve. ----
	A h[N];
	{
		if constexpr (!SIMPLE) {
			OSC_UNROLL for (int c = 0; c < 6; ++c) g[c] = SM(F::SC + c);
			if (tc.decoupling != OSC_IMPEDANCE) {
				OSC_UNROLL for (int i = 0; i < 6; ++i) {  // g <- C^-1 F*
					A s = g[i];
					OSC_UNROLL for (int m = 0; m < i; ++m) s = Num<A>::fma_(-SM(F::CC + i * (i - 1) / 2 + m), g[m], s);
					g[i] = s * SM(F::cli(i));
				}
			}
		}
		if (tc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES && beta > A(0)) {
			// A_BIE = A - beta y y^T with y = C w (w = last column of W):  C^T A_BIE^-1 F* = g + beta w (w . g) / (1 - beta |w|^2)
			A ww = A(0), wg = A(0);
			OSC_UNROLL for (int c = 0; c < 6; ++c) { ww = Num<A>::fma_(W[c][N - 1], W[c][N - 1], ww); wg = Num<A>::fma_(W[c][N - 1], g[c], wg); }
			const A coef = beta * wg / (A(1) - beta * ww);
			OSC_UNROLL for (int c = 0; c < 6; ++c) g[c] = Num<A>::fma_(coef, W[c][N - 1], g[c]);
		}
		if constexpr (!SIMPLE) {  // + C^T F_f (and C^T F* when Lambda_mod = I)
			A f[6];
			OSC_UNROLL for (int c = 0; c < 6; ++c) f[c] = SM(F::SC + 6 + c) + (tc.decoupling == OSC_IMPEDANCE ? g[c] : A(0));
			OSC_UNROLL for (int c = 0; c < 6; ++c) {
				A s = f[c] / SM(F::cli(c));
				OSC_UNROLL for (int c2 = c + 1; c2 < 6; ++c2) s = Num<A>::fma_(SM(F::CC + c2 * (c2 - 1) / 2 + c), f[c2], s);
				g[c] = (tc.decoupling == OSC_IMPEDANCE ? A(0) : g[c]) + s;
			}
		}
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			A s = A(0);
			OSC_UNROLL for (int c = 0; c < 6; ++c) s = Num<A>::fma_(W[c][i], g[c], s);
			h[i] = s;
		}
	}
	// Y (= J) is dead from here on.

	OSC_PHASE();
	{  // fresh rows for the closing memory phases
	OSC_FRESH_ROWS();
	// ---- classification state of a non-singular robot (SingularityHandler.cpp:233-244); see osc_fused.cuh ----
	bool latch = (status & CLS_LATCH) != 0;
	if (status & CLS_RESET) {
		latch = latch || mI(ML::C2) > mI(ML::C1);
		mI(ML::NTYPES) = 0; mI(ML::TYPES) = 0; mI(ML::C1) = 0; mI(ML::C2) = 0; mI(ML::HIST_LEN) = 0; mI(ML::HIST_HEAD) = 0;
	}
	latch = latch && ex;
	T qq[N];  // q_prior latch: loaded here, stored at the very end (the loads ride under P11)
	OSC_UNROLL for (int i = 0; i < N; ++i) qq[i] = latch ? rP.ro(RobotL<N>::Q + i) : T(0);

	OSC_PHASE();
	// ---- P11: tau_1 = L h and w = L u (one sweep over L), parked in the dead C slots ----
	constexpr int TAU1 = SIMPLE ? F::SC : F::CC, WW = TAU1 + N;  // dead slots: sin/cos (SIMPLE, F* lives in registers) or C
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		const A ld = A(1) / SM(F::LI + i);
		A s1 = ld * h[i], sw = ld * u[i];
		OSC_UNROLL for (int m = 0; m < i; ++m) {
			const A l = SM(F::LL + i * (i - 1) / 2 + m);
			s1 = Num<A>::fma_(l, h[m], s1);
			sw = Num<A>::fma_(l, u[m], sw);
		}
		if (ex) mS(ML::TASK_TAU + i) = s1;
		if constexpr (JT7) { SM(TAU1 + i) = s1; SM(WW + i) = sw; }
		else rS(RobotL<N>::TAU + i) = gravity ? s1 + rS(RobotL<N>::GRAV + i) : s1;
	}
	if constexpr (JT7) {
		OSC_FENCE();
		// v = L^-T u, M^-1 tau_1 = L^-T h: in place (u -> v, h -> mt), one backward sweep over L
		const A u_last = u[N - 1];
		OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
			A sv = u[i], st = h[i];
			OSC_UNROLL for (int k = i + 1; k < N; ++k) {
				const A l = SM(F::LL + k * (k - 1) / 2 + i);
				sv = Num<A>::fma_(-l, u[k], sv);
				st = Num<A>::fma_(-l, h[k], st);
			}
			const A li = SM(F::LI + i);
			u[i] = sv * li; h[i] = st * li;
		}
		A (&v)[N] = u;
		// JointTask in the nullspace (JointTask.cpp:218-356): R = v / |v|, Lambda = 1 / |v|^2,
		// tau_2 = w [Lambda (v.a - v.M^-1 tau_1) + Lambda_mod v.u_pd]
		A nv2 = A(0), vc = A(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) { nv2 = Num<A>::fma_(v[i], v[i], nv2); vc = Num<A>::fma_(v[i], h[i], vc); }
		const A lam = A(1) / nv2;
		A lam_mod = lam;  // JointTask.cpp:253-267
		if (jc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES) lam_mod = A(1) / (nv2 * (A(1) - beta * u_last * u_last));  // |L_BIE^-1 L u|^2
		else if (jc.decoupling == OSC_IMPEDANCE) lam_mod = A(1);
		A va = A(0), vu = A(0);
		if constexpr (SIMPLE) {  // u_pd and the goal acceleration were staged by P1
			OSC_UNROLL for (int i = 0; i < N; ++i) {
				va = Num<A>::fma_(v[i], SM(F::GAC + i), va);
				vu = Num<A>::fma_(v[i], SM(F::UPD + i), vu);
			}
		} else {
			T q[N], dq[N], upd[N], acc[N];
			ld_vec<T, N>(rP, RobotL<N>::Q, q);
			ld_vec<T, N>(rP, RobotL<N>::DQ, dq);
			jt_pd_law<T, N>(jc, jP, jS, q, dq, upd, acc, ex);
			OSC_UNROLL for (int i = 0; i < N; ++i) { va = Num<A>::fma_(v[i], acc[i], va); vu = Num<A>::fma_(v[i], upd[i], vu); }
			if (ex) {
				st_vec<T, N>(jS, JL::CUR_POS, q);
				st_vec<T, N>(jS, JL::CUR_VEL, dq);
			}
		}
		const A sc = lam * (va - (stacked ? vc : A(0))) + lam_mod * vu;
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			const A t2 = SM(WW + i) * sc;
			A t = SM(TAU1 + i) + t2;
			if (gravity) t += rS(RobotL<N>::GRAV + i);  // written by k_gravity ahead of this kernel
			rS(RobotL<N>::TAU + i) = t;
			if (ex) jS(JL::TASK_TAU + i) = t2;
		}
	} else if (ex && !(exports & OSC_EXPORT_NO_JT)) {
		OSC_UNROLL for (int i = 0; i < N; ++i) { jS(JL::CUR_POS + i) = rP(RobotL<N>::Q + i); jS(JL::CUR_VEL + i) = rP(RobotL<N>::DQ + i); jS(JL::TASK_TAU + i) = A(0); }
	}
	if (ex) {
		if (!(exports & OSC_EXPORT_NO_JT)) tk1.I[(int64_t)JL::RANK * stride + r] = JT7 ? 1 : 0;
		mS(ML::ALPHA) = T(1);
		mI(ML::INS) = 6;
	}
	if (latch) {
		OSC_UNROLL for (int i = 0; i < N; ++i) mP(ML::Q_PRIOR + i) = qq[i];
	}
	status &= ~(CLS_RESET | CLS_LATCH);
	if (status) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status);
	}
#undef OSC_FRESH_ROWS
#undef OSC_PHASE
#undef OSC_FALL_BACK
#undef SM
}

// jointGravityVector for the fused tick when gravity compensation is on (default off, RobotController.cpp:10): a small
// separate launch ahead of k_tick_fused2 keeps the gravity recursion out of the hot kernel's register budget.
template <class T, int N>
__global__ void __launch_bounds__(OSC_THREADS) k_gravity(const __grid_constant__ ChainC<T, N> ch, Arena<T> rob, int64_t B, int64_t stride) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T> P{rob.P, stride, r}, S{rob.S, stride, r};
	T q[N], g[N];
	ld_vec<T, N>(P, RobotL<N>::Q, q);
	Kin<T, N> kin;
	forward_kinematics<T, N>(ch, q, kin);
	gravity_vector<T, N>(ch, kin, g);
	st_vec<T, N>(S, RobotL<N>::GRAV, g);
}

// what the second-generation kernel adds to fused_tick_supported: a revolute chain with the control link on the last body
inline bool fused2_simple(const Batch& b) {
	const osc_mft_params& p = b.tasks[0].mp;
	return p.dynamic_decoupling_type != OSC_IMPEDANCE && p.force_space_dimension == 0 && p.moment_space_dimension == 0 && !p.closed_loop_force_control && !p.closed_loop_moment_control &&
		   !p.use_velocity_saturation && !(b.tasks.size() > 1 && b.tasks[1].jp.use_velocity_saturation);
}
inline bool fused2_supported(Batch& b) {
	if (b.tasks.empty() || b.tasks[0].md.body != b.n - 1) return false;
	for (int i = 0; i < b.n; ++i)
		if (b.chain.joint_type[i] != OSC_JOINT_REVOLUTE) return false;
	return true;
}

}  // namespace osc
