// osc_rollout.cuh -- device-side rollout step: the caller of the control tick in the
// MPC-rollout / parallel-environment use case, so that q and dq never leave the GPU between ticks.
//
// What it stands in for: the simulation side of the reference's examples,
//     sim->setJointTorques(robot_name, control_torques); sim->integrate(); robot->setQ(sim->getJointPositions(...));
//     robot->setDq(sim->getJointVelocities(...));        (examples/05-using_robot_controller/05-using_robot_controller.cpp:219-233)
// where sim is Sai2Simulation (un-vendored, articulated-body dynamics + contact).  Built here: contact-free rigid-body
// forward dynamics  qdd = M^-1 (tau - b(q, dq) - g(q) - d .* dq)  and semi-implicit (symplectic) Euler
//     dq += qdd dt ;  q += dq dt
// per robot, `substeps` times per call with the torque held (zero-order hold, like the example's control/sim threads).
// b + g come from one recursive Newton-Euler pass in WORLD coordinates about the world origin (the same convention as
// mass_matrix in osc_model.cuh): Sai2Model::coriolisPlusGravity.  One robot per thread, everything in registers.
#pragma once
#include "osc_kernels.cuh"

namespace osc {

// h = b(q, dq) + g(q): joint torques that hold qdd = 0.  World frame; gravity enters as a base acceleration of -g.
template <class T, int N>
OSC_DEV void nonlinear_effects(const ChainC<T, N>& ch, const Kin<T, N>& k, const T (&dq)[N], T (&h)[N]) {
	T w[3] = {T(0), T(0), T(0)}, al[3] = {T(0), T(0), T(0)};            // angular velocity / acceleration of the current body
	T a[3] = {-ch.gravity[0], -ch.gravity[1], -ch.gravity[2]};          // linear acceleration of the current body origin
	T op[3] = {k.o[0][0], k.o[0][1], k.o[0][2]};                         // previous origin (body -1 does not move: any point)
	T F[N][3], Mo[N][3];                                                 // per body: force, moment about the WORLD origin
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		const T z[3] = {k.R[i][0][2], k.R[i][1][2], k.R[i][2][2]};
		// origin i rides on body i-1
		const T d[3] = {k.o[i][0] - op[0], k.o[i][1] - op[1], k.o[i][2] - op[2]};
		T wd[3], wwd[3], ad[3];
		cross3(w, d, wd);
		cross3(w, wd, wwd);
		cross3(al, d, ad);
		OSC_UNROLL for (int c = 0; c < 3; ++c) a[c] += ad[c] + wwd[c];
		if (ch.prismatic[i]) {  // origin slides along z: Coriolis 2 w x z dq (no angular change)
			T wz[3];
			cross3(w, z, wz);
			OSC_UNROLL for (int c = 0; c < 3; ++c) a[c] += T(2) * dq[i] * wz[c];
		} else {
			T wz[3];
			cross3(w, z, wz);
			OSC_UNROLL for (int c = 0; c < 3; ++c) { al[c] += dq[i] * wz[c]; w[c] += dq[i] * z[c]; }
		}
		OSC_UNROLL for (int c = 0; c < 3; ++c) op[c] = k.o[i][c];
		// centre of mass of body i
		T rc[3];
		mv3(k.R[i], ch.com[i], rc);
		T wr[3], wwr[3], ar[3], ac[3];
		cross3(w, rc, wr);
		cross3(w, wr, wwr);
		cross3(al, rc, ar);
		OSC_UNROLL for (int c = 0; c < 3; ++c) ac[c] = a[c] + ar[c] + wwr[c];
		// inertia about the com in world axes applied to al and w:  R I R^T x
		const T Ib[3][3] = {{ch.inertia[i][0], ch.inertia[i][1], ch.inertia[i][2]},
							{ch.inertia[i][1], ch.inertia[i][3], ch.inertia[i][4]},
							{ch.inertia[i][2], ch.inertia[i][4], ch.inertia[i][5]}};
		T lw[3], la[3], t1[3], Iw[3], Ia[3], wIw[3];
		mtv3(k.R[i], w, lw);
		mtv3(k.R[i], al, la);
		mv3(Ib, lw, t1);
		mv3(k.R[i], t1, Iw);
		mv3(Ib, la, t1);
		mv3(k.R[i], t1, Ia);
		cross3(w, Iw, wIw);
		const T cw[3] = {k.o[i][0] + rc[0], k.o[i][1] + rc[1], k.o[i][2] + rc[2]};
		T f[3], cf[3];
		OSC_UNROLL for (int c = 0; c < 3; ++c) f[c] = ch.mass[i] * ac[c];
		cross3(cw, f, cf);
		OSC_UNROLL for (int c = 0; c < 3; ++c) { F[i][c] = f[c]; Mo[i][c] = Ia[c] + wIw[c] + cf[c]; }
	}
	T Fs[3] = {T(0), T(0), T(0)}, Ms[3] = {T(0), T(0), T(0)};  // wrench of bodies i..N-1 about the world origin
	OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
		OSC_UNROLL for (int c = 0; c < 3; ++c) { Fs[c] += F[i][c]; Ms[c] += Mo[i][c]; }
		const T z[3] = {k.R[i][0][2], k.R[i][1][2], k.R[i][2][2]};
		if (ch.prismatic[i]) {
			h[i] = dot3(z, Fs);
		} else {
			T oF[3];
			cross3(k.o[i], Fs, oF);
			const T m[3] = {Ms[0] - oF[0], Ms[1] - oF[1], Ms[2] - oF[2]};  // moment about joint origin i
			h[i] = dot3(z, m);
		}
	}
}

// qdd = M^-1 (tau - h - damping .* dq), then symplectic Euler; `substeps` times with tau held
template <class T, int N>
__global__ void __launch_bounds__(OSC_THREADS) k_rollout_step(const __grid_constant__ ChainC<T, N> ch, Arena<T> rob, int64_t B, int64_t stride, T dt,
															   int substeps, T damping, int export_h) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T> P{rob.P, stride, r}, S{rob.S, stride, r};
	T q[N], dq[N], tau[N];
	ld_vec<T, N>(P, RobotL<N>::Q, q);
	ld_vec<T, N>(P, RobotL<N>::DQ, dq);
	ld_vec<T, N>(S, RobotL<N>::TAU, tau);
	int status = 0;
	for (int s = 0; s < substeps; ++s) {
		Kin<T, N> kin;
		forward_kinematics<T, N>(ch, q, kin);
		T x[N];
		{  // the Newton-Euler pass first: its per-body wrenches are dead before the mass matrix is born
			T h[N];
			nonlinear_effects<T, N>(ch, kin, dq, h);
			if (export_h && s == 0) st_vec<T, N>(S, RobotL<N>::NLE, h);
			OSC_UNROLL for (int i = 0; i < N; ++i) x[i] = tau[i] - h[i] - damping * dq[i];
		}
		T M[N][N];
		mass_matrix<T, N>(ch, kin, M);
		Chol<T, N> L;
		if (!chol_factor<T, N>(M, L)) status |= 16;  // OSC_STATUS_NOT_POSITIVE_DEFINITE
		chol_fwd<T, N>(L, x);
		chol_bwd<T, N>(L, x);
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			dq[i] = Num<T>::fma_(x[i], dt, dq[i]);
			q[i] = Num<T>::fma_(dq[i], dt, q[i]);
		}
	}
	st_vec<T, N>(P, RobotL<N>::Q, q);
	st_vec<T, N>(P, RobotL<N>::DQ, dq);
	if (status) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status);
}

// ---- second generation (all-revolute chains): the structure of k_tick_fused2 --------------------------------------------
// k_rollout_step keeps world-frame kinematics of every body (84 values), the per-body wrenches (42) and M (49) alive in one
// 255-register straight line with 770 B of spills at 8 warps per SM.  Here everything is in BODY frames and needs sin/cos(q) only:
//   * recursive Newton-Euler in link coordinates, base to tip (omega, alpha, origin acceleration; gravity as a base acceleration)
//     then tip to base; the per-body force / moment about the body origin wait in a per-thread SHARED-memory stash;
//   * the mass matrix from the body-frame composite-rigid-body recursion of k_tick_fused2 (phase P2 there), into the same slots;
//   * Cholesky in registers, two triangular solves, symplectic Euler.
// 128 registers, 224-thread blocks x 2 per SM (63 stash slots per thread).
template <int N> struct R2 {
	static constexpr int SC = 0;            // sin/cos (2N)
	static constexpr int FN = 2 * N;        // per body f (3), n (3) in its own frame; later the lower triangle of M
	static constexpr int DQP = FN + N * (N + 1) / 2;  // dq parked behind M while the factorisation needs the registers
	static constexpr int QQ = 8 * N;                   // q, from the top of a substep to its integration (N slots of its own)
	static constexpr int SLOTS = DQP + N > 9 * N ? DQP + N : 9 * N;  // n = 7: 63 slots, 2 x (224 x 63 x 8 B + 1 KB) = 223 KB of the SM's 228 KB
};

template <class T, int N, int BT>
__global__ void __launch_bounds__(BT, 512 / BT) k_rollout_step2(const __grid_constant__ ChainC<T, N> ch, Arena<T> rob, int64_t B, int64_t stride,
																		 T dt, int substeps, T damping, int export_h) {
	using F = R2<N>;
	extern __shared__ __align__(16) unsigned char osc_stash_raw[];
	T* const sm = reinterpret_cast<T*>(osc_stash_raw);
#define ST(slot) sm[(slot) * BT + threadIdx.x]
// phase boundary: the compiler fence keeps nvcc, the block barrier keeps ptxas from interleaving the phases (register pressure);
// the barrier also keeps the warps of a block in the same stretch of code (instruction cache), as in k_tick_fused2
#define OSC_R2_PHASE() \
	OSC_FENCE();       \
	__syncthreads()
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T> P{rob.P, stride, r}, S{rob.S, stride, r};
	T q[N], dq[N];
	ld_vec<T, N>(P, RobotL<N>::Q, q);
	ld_vec<T, N>(P, RobotL<N>::DQ, dq);
	T ab[3];  // acceleration of the (fixed) base that stands for gravity, in base coordinates: world_rot^T (-g)
	OSC_UNROLL for (int a = 0; a < 3; ++a)
		ab[a] = -(ch.world_rot[0][a] * ch.gravity[0] + ch.world_rot[1][a] * ch.gravity[1] + ch.world_rot[2][a] * ch.gravity[2]);
	int status = 0;
	for (int step = 0; step < substeps; ++step) {
		OSC_UNROLL for (int i = 0; i < N; ++i) { ST(F::SC + 2 * i) = q[i]; ST(F::QQ + i) = q[i]; }
		_Pragma("unroll 1") for (int i = 0; i < N; ++i) {
			T s, c;
			Num<T>::sincos_(ST(F::SC + 2 * i), &s, &c);
			ST(F::SC + 2 * i) = s;
			ST(F::SC + 2 * i + 1) = c;
		}
		OSC_R2_PHASE();
		T h[N];
		{  // Newton-Euler, outward: body i in its own frame; x_parent = xyz_i + E_i x_i with E_i = rot_i Rz(q_i)
			T w[3] = {T(0), T(0), T(0)}, al[3] = {T(0), T(0), T(0)}, ac[3] = {ab[0], ab[1], ab[2]};
			OSC_UNROLL for (int i = 0; i < N; ++i) {
				T s = ST(F::SC + 2 * i), c = ST(F::SC + 2 * i + 1);
								T E[3][3];
				OSC_UNROLL for (int a = 0; a < 3; ++a) {
					E[a][0] = ch.rot[i][a][0] * c + ch.rot[i][a][1] * s;
					E[a][1] = ch.rot[i][a][1] * c - ch.rot[i][a][0] * s;
					E[a][2] = ch.rot[i][a][2];
				}
				T wr[3], wwr[3], ar[3], ap[3];
				cross3(w, ch.xyz[i], wr);
				cross3(w, wr, wwr);
				cross3(al, ch.xyz[i], ar);
				OSC_UNROLL for (int a = 0; a < 3; ++a) ap[a] = ac[a] + ar[a] + wwr[a];  // origin i, parent coordinates
				T wi[3], ali[3];
				mtv3(E, w, wi);
				mtv3(E, al, ali);
				mtv3(E, ap, ac);
				// joint rate about z_i:  alpha += omega_parent x (dq z),  omega += dq z
				al[0] = ali[0] + dq[i] * wi[1];
				al[1] = ali[1] - dq[i] * wi[0];
				al[2] = ali[2];
				w[0] = wi[0]; w[1] = wi[1]; w[2] = wi[2] + dq[i];
				// wrench that moves body i, about its origin:  f = m a + alpha x h + w x (w x h),  n = I alpha + w x (I w) + h x a
				const T Ib[3][3] = {{ch.Ib[i][0], ch.Ib[i][1], ch.Ib[i][2]}, {ch.Ib[i][1], ch.Ib[i][3], ch.Ib[i][4]}, {ch.Ib[i][2], ch.Ib[i][4], ch.Ib[i][5]}};
				T ah[3], wh[3], wwh[3], Ia[3], Iw[3], wIw[3], ha[3];
				cross3(al, ch.hb[i], ah);
				cross3(w, ch.hb[i], wh);
				cross3(w, wh, wwh);
				mv3(Ib, al, Ia);
				mv3(Ib, w, Iw);
				cross3(w, Iw, wIw);
				cross3(ch.hb[i], ac, ha);
				OSC_UNROLL for (int a = 0; a < 3; ++a) {
					ST(F::FN + 6 * i + a) = ch.mass[i] * ac[a] + ah[a] + wwh[a];
					ST(F::FN + 6 * i + 3 + a) = Ia[a] + wIw[a] + ha[a];
				}
				OSC_FENCE();  // one body at a time: left alone, ptxas builds every E_i up front and spills
			}
			OSC_R2_PHASE();
			// inward: wrench of bodies >= i about origin i; the joint torque is its moment about z_i
			T Fc[3] = {T(0), T(0), T(0)}, Nc[3] = {T(0), T(0), T(0)};
			OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
				OSC_UNROLL for (int a = 0; a < 3; ++a) { Fc[a] += ST(F::FN + 6 * i + a); Nc[a] += ST(F::FN + 6 * i + 3 + a); }
				h[i] = Nc[2];
				if (i > 0) {
					T s = ST(F::SC + 2 * i), c = ST(F::SC + 2 * i + 1);
										T E[3][3];
					OSC_UNROLL for (int a = 0; a < 3; ++a) {
						E[a][0] = ch.rot[i][a][0] * c + ch.rot[i][a][1] * s;
						E[a][1] = ch.rot[i][a][1] * c - ch.rot[i][a][0] * s;
						E[a][2] = ch.rot[i][a][2];
					}
					T Fp[3], Np[3], rF[3];
					mv3(E, Fc, Fp);
					mv3(E, Nc, Np);
					cross3(ch.xyz[i], Fp, rF);
					OSC_UNROLL for (int a = 0; a < 3; ++a) { Fc[a] = Fp[a]; Nc[a] = Np[a] + rF[a]; }
				}
				OSC_FENCE();
			}
		}
		if (export_h && step == 0) st_vec<T, N>(S, RobotL<N>::NLE, h);
		T x[N];
		{  // the torque is held over the substeps but not in registers: re-read (L1 / L2) where it is used
			T tau[N];
			ld_vec<T, N>(S, RobotL<N>::TAU, tau);
			OSC_UNROLL for (int i = 0; i < N; ++i) x[i] = tau[i] - h[i] - damping * dq[i];
		}
		// dq waits in slots the lower triangle of M leaves free while the factorisation needs the registers
		OSC_UNROLL for (int i = 0; i < N; ++i) ST(F::DQP + i) = dq[i];
		OSC_R2_PHASE();
		{  // mass matrix: the composite-rigid-body recursion in body frames of k_tick_fused2 (P2), lower triangle into the FN slots
			T cm = T(0), h0 = T(0), h1 = T(0), h2 = T(0), Ixx = T(0), Ixy = T(0), Ixz = T(0), Iyy = T(0), Iyz = T(0), Izz = T(0);
			OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
				cm += ch.mass[i];
				h0 += ch.hb[i][0]; h1 += ch.hb[i][1]; h2 += ch.hb[i][2];
				Ixx += ch.Ib[i][0]; Ixy += ch.Ib[i][1]; Ixz += ch.Ib[i][2]; Iyy += ch.Ib[i][3]; Iyz += ch.Ib[i][4]; Izz += ch.Ib[i][5];
				T p0 = -h1, p1 = h0, p2 = T(0), l0 = Ixz, l1 = Iyz, l2 = Izz;  // momentum of the composite under omega = z_i
				const int row = F::FN + i * (i + 1) / 2;
				ST(row + i) = l2;
				OSC_UNROLL for (int j = i; j >= 1; --j) {  // carried down the chain: M[i][j-1] = z . l in frame j-1
					T s = ST(F::SC + 2 * j), c = ST(F::SC + 2 * j + 1);
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
					ST(row + j - 1) = l2;
				}
				if (i > 0) {  // the composite in the parent frame
					T s = ST(F::SC + 2 * i), c = ST(F::SC + 2 * i + 1);
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
				OSC_FENCE();
			}
		}
		OSC_R2_PHASE();
		{  // M = L L^T in registers; x <- M^-1 x
			T Mx[N][N], li[N];
			OSC_UNROLL for (int i = 0; i < N; ++i)
				OSC_UNROLL for (int j = 0; j <= i; ++j) Mx[i][j] = ST(F::FN + i * (i + 1) / 2 + j);
			if (!chol_inplace<T, N>(Mx, li)) status |= 16;  // OSC_STATUS_NOT_POSITIVE_DEFINITE
			OSC_UNROLL for (int i = 0; i < N; ++i) {
				T a = x[i];
				OSC_UNROLL for (int k = 0; k < i; ++k) a = Num<T>::fma_(-Mx[i][k], x[k], a);
				x[i] = a * li[i];
			}
			OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
				T a = x[i];
				OSC_UNROLL for (int k = i + 1; k < N; ++k) a = Num<T>::fma_(-Mx[k][i], x[k], a);
				x[i] = a * li[i];
			}
		}
		OSC_FENCE();
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			dq[i] = Num<T>::fma_(x[i], dt, ST(F::DQP + i));
			q[i] = Num<T>::fma_(dq[i], dt, ST(F::QQ + i));
		}
	}
	st_vec<T, N>(P, RobotL<N>::Q, q);
	st_vec<T, N>(P, RobotL<N>::DQ, dq);
	if (status) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status);
#undef ST
#undef OSC_R2_PHASE
}

}  // namespace osc
