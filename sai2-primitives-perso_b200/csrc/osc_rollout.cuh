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

}  // namespace osc
