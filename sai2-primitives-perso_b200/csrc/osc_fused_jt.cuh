// osc_fused_jt.cuh -- second generation of the joint-task-alone tick (k_tick_joint in osc_fused.cuh is the first) for all-revolute
// chains, with the structure of k_tick_fused2 / k_rollout_step2: the mass matrix from the BODY-frame composite-rigid-body recursion
// (sin/cos(q) only, no world-frame kinematics), its lower triangle parked in a per-thread shared-memory stash, 128 registers,
// 224-thread blocks x 2 per SM with block barriers at the phase boundaries (64-thread blocks for small batches).
// tau = M a_des + Lambda_mod u with Lambda_mod = M_BIE | M | I (JointTask.cpp:218-283, 342-355; N_prec = I).
#pragma once
#include "osc_fused2.cuh"

namespace osc {

template <int N> struct FJ {
	static constexpr int SC = 0;       // sin/cos (2N)
	static constexpr int MM = 2 * N;   // lower triangle of M, row i at MM + i(i+1)/2
	static constexpr int SLOTS = 2 * N + N * (N + 1) / 2;
};

template <class T, int N, int BT>
__global__ void __launch_bounds__(BT, 512 / BT) k_tick_joint2(const __grid_constant__ ChainC<T, N> ch, const __grid_constant__ JtC<T, N> jc, Arena<T> rob,
															   Arena<T> tk, int64_t B, int64_t stride, int gravity, int exports) {
	using F = FJ<N>;
	using JL = JtL<N>;
	extern __shared__ __align__(16) unsigned char osc_stash_raw[];
	T* const sm = reinterpret_cast<T*>(osc_stash_raw);
#define ST(slot) sm[(slot) * BT + threadIdx.x]
#define OSC_FJ_PHASE() \
	OSC_FENCE();       \
	__syncthreads()
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<T> rP{rob.P, stride, r}, rS{rob.S, stride, r}, jP{tk.P, stride, r}, jS{tk.S, stride, r};
	const bool ex = (exports & OSC_EXPORT_VECTORS) != 0;
	OSC_UNROLL for (int i = 0; i < N; ++i) ST(F::SC + 2 * i) = rP.ro(RobotL<N>::Q + i);
	_Pragma("unroll 1") for (int i = 0; i < N; ++i) {
		T s, c;
		Num<T>::sincos_(ST(F::SC + 2 * i), &s, &c);
		ST(F::SC + 2 * i) = s;
		ST(F::SC + 2 * i + 1) = c;
	}
	OSC_FJ_PHASE();
	{  // composite rigid bodies tip -> base in body frames (k_tick_fused2 P2): column i = momentum of composite i under omega = z_i,
		// carried down the chain; M[i][j-1] = z . l in frame j-1
		T cm = T(0), h0 = T(0), h1 = T(0), h2 = T(0), Ixx = T(0), Ixy = T(0), Ixz = T(0), Iyy = T(0), Iyz = T(0), Izz = T(0);
		OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
			cm += ch.mass[i];
			h0 += ch.hb[i][0]; h1 += ch.hb[i][1]; h2 += ch.hb[i][2];
			Ixx += ch.Ib[i][0]; Ixy += ch.Ib[i][1]; Ixz += ch.Ib[i][2]; Iyy += ch.Ib[i][3]; Iyz += ch.Ib[i][4]; Izz += ch.Ib[i][5];
			T p0 = -h1, p1 = h0, p2 = T(0), l0 = Ixz, l1 = Iyz, l2 = Izz;
			const int row = F::MM + i * (i + 1) / 2;
			ST(row + i) = l2;
			OSC_UNROLL for (int j = i; j >= 1; --j) {
				const T s = ST(F::SC + 2 * j), c = ST(F::SC + 2 * j + 1);
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
			if (i > 0) {  // the composite in the parent frame: E = rot_i Rz(q_i), origin offset xyz_i
				const T s = ST(F::SC + 2 * i), c = ST(F::SC + 2 * i + 1);
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
	OSC_FJ_PHASE();
	T q[N], dq[N], u[N], acc[N], tau[N];
	ld_vec<T, N>(rP, RobotL<N>::Q, q);
	ld_vec<T, N>(rP, RobotL<N>::DQ, dq);
	jt_pd_law<T, N>(jc, jP, jS, q, dq, u, acc, ex);
	const bool impedance = jc.decoupling == OSC_IMPEDANCE;
	T x[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) x[i] = impedance ? acc[i] : acc[i] + u[i];
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		T s = T(0);
		OSC_UNROLL for (int j = 0; j < N; ++j) {
			const int hi = i > j ? i : j, lo = i > j ? j : i;
			s = Num<T>::fma_(ST(F::MM + hi * (hi + 1) / 2 + lo), x[j], s);
		}
		if (impedance) s += u[i];
		else if (jc.decoupling == OSC_BOUNDED_INERTIA_ESTIMATES) {
			const T mii = ST(F::MM + i * (i + 1) / 2 + i);
			if (mii < jc.bie_threshold) s = Num<T>::fma_(jc.bie_threshold - mii, u[i], s);  // JointTask.cpp:253-267
		}
		tau[i] = s;
	}
	if (ex) {
		st_vec<T, N>(jS, JL::TASK_TAU, tau);
		st_vec<T, N>(jS, JL::CUR_POS, q);
		st_vec<T, N>(jS, JL::CUR_VEL, dq);
		tk.I[(int64_t)JL::RANK * stride + r] = N;
	}
	if (gravity) {  // written by k_gravity ahead of this kernel (RobotController.cpp:70-72)
		OSC_UNROLL for (int i = 0; i < N; ++i) tau[i] += rS(RobotL<N>::GRAV + i);
	}
	st_vec<T, N>(rS, RobotL<N>::TAU, tau);
#undef ST
#undef OSC_FJ_PHASE
}

}  // namespace osc
