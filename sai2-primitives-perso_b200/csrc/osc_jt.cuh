// osc_jt.cuh -- JointTask::updateTaskModel (src/tasks/JointTask.cpp:218-283) and JointTask::computeTorques
// (src/tasks/JointTask.cpp:285-356).  The k x n selection matrix is padded to N x N with zero rows; padded rows give
// exactly-zero singular values, so the range logic of Sai2Model::matrixRangeBasis works on the padded shape.
#pragma once
#include "osc_kernels.cuh"

namespace osc {

// The joint-space control law of JointTask::computeTorques (JointTask.cpp:319-346): integrates the error, returns the
// PID (or velocity-saturated) command u and the feed-forward acceleration; rows >= k are zero.
template <class T, int N, bool NOSAT = false, class S = T>
OSC_DEV void jt_pd_law(const JtC<T, N>& jc, const Row<T, S>& tP, const Row<T, S>& tS, const T (&cp)[N], const T (&cv)[N], T (&u)[N], T (&acc)[N],
					   bool export_vectors, bool integrate = true) {
	using L = JtL<N>;
	T dv[N];
	const int k = jc.k;
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		const T e = cp[i] - tP(L::GOAL_POS + i);
		T ie = tP(L::INTEG + i);
		if (i < k && integrate) { ie += e * jc.dt; tP(L::INTEG + i) = ie; }  // JointTask.cpp:323-324
		dv[i] = tP(L::GOAL_VEL + i);
		acc[i] = (i < k) ? tP(L::GOAL_ACC + i) : T(0);
		if (!NOSAT && jc.use_vel_sat) {  // JointTask.cpp:327-340 (quirk Q7: loop to k)
			T mx = T(0);
			OSC_UNROLL for (int j = 0; j < N; ++j)
				if (j < k) mx = Num<T>::abs_(jc.kv[j]) > mx ? Num<T>::abs_(jc.kv[j]) : mx;
			const T kvi = Num<T>::abs_(jc.kv[i]) > T(1e-6) * mx ? T(1) / jc.kv[i] : T(0);
			T d = -jc.kp[i] * kvi * e - jc.ki[i] * kvi * ie;
			d = d > jc.sat[i] ? jc.sat[i] : (d < -jc.sat[i] ? -jc.sat[i] : d);
			dv[i] = d;
			u[i] = -jc.kv[i] * (cv[i] - d);
		} else {
			u[i] = -jc.kp[i] * e - jc.kv[i] * (cv[i] - dv[i]) - jc.ki[i] * ie;
		}
		if (i >= k) { u[i] = T(0); dv[i] = T(0); }
	}
	if (export_vectors) st_vec<T, N>(tS, L::DES_VEL, dv);
}

template <class T, int N, class S = T, bool REDIR = false>
__device__ void jt_update_robot(const JtC<T, N>& jc, const Arena<S>& rob, const Arena<S>& tk, const S* nprec_src, int64_t stride, int64_t r,
								int64_t np_stride = 0, int64_t np_index = -1) {
	using L = JtL<N>;
	const Row<T, S> rS = srow<T, REDIR>(rob, stride, r), tS = srow<T, REDIR>(tk, stride, r);
	const Row<int32_t> tI{tk.I, stride, r};
	int status = 0;
	T Np[N][N];
	if (nprec_src) {
		const Row<T, S> src{const_cast<S*>(nprec_src), (REDIR && np_stride) ? np_stride : stride, (REDIR && np_index >= 0) ? np_index : r};
		ld_mat<T, N, N>(src, 0, Np);
	} else {
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < N; ++j) Np[i][j] = (i == j) ? T(1) : T(0);
	}
	st_mat<T, N, N>(tS, L::NPREC, Np);
	T Jp[N][N];  // S * N_prec (JointTask.cpp:231)
	if (jc.full) {
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < N; ++j) Jp[i][j] = Np[i][j];
	} else {
		mm<T, N, N, N>(jc.S, Np, Jp);
	}
	const int k = jc.k;
	// Sai2Model::matrixRangeBasis(Jp, 1e-3)
	T fro = T(0);
	OSC_UNROLL for (int i = 0; i < N; ++i)
		OSC_UNROLL for (int j = 0; j < N; ++j) fro = Num<T>::fma_(Jp[i][j], Jp[i][j], fro);
	fro = Num<T>::sqrt_(fro);
	int rank = 0;
	T Rg[N][N];  // range basis, k x rank, zero padded
	OSC_UNROLL for (int i = 0; i < N; ++i)
		OSC_UNROLL for (int j = 0; j < N; ++j) Rg[i][j] = T(0);
	// Full joint task: Jp = N_prec is a PROJECTOR (oblique: its non-zero singular values are >= 1 and <= |N|_F, the others are
	// rounding noise), so the rank the reference reads off the singular values with its 1e-3 relative cut is trace(N) as long as
	// |N|_F < 500, and -- every result of the task being invariant under a rotation of the range basis -- any orthonormal basis
	// of range(N) serves: pivoted Gram-Schmidt on the columns instead of a 7x7 Jacobi SVD (15,100 of the 36,000 FLOP of a tick).
	bool fast = false;
	if (jc.full && fro >= T(1e-3) && fro < T(500)) {
		T tr = T(0);
		OSC_UNROLL for (int i = 0; i < N; ++i) tr += Jp[i][i];
		const int rk = (int)(tr + T(0.5));
		if (Num<T>::abs_(tr - T(rk)) <= T(4) * Num<T>::sqrt_(Num<T>::eps()) * fro && rk >= 1 && rk <= N) {  // integer trace: a projector
			fast = true;
			rank = rk;
			if (rk == N) {
				OSC_UNROLL for (int i = 0; i < N; ++i) Rg[i][i] = T(1);
			} else {
				T A[N][N];
				OSC_UNROLL for (int i = 0; i < N; ++i)
					OSC_UNROLL for (int j = 0; j < N; ++j) A[i][j] = Jp[i][j];
				OSC_UNROLL for (int step = 0; step < N - 1; ++step) {
					if (step < rk) {
						T best = T(-1);
						int kb = 0;
						OSC_UNROLL for (int j = 0; j < N; ++j) {
							T nj = T(0);
							OSC_UNROLL for (int i = 0; i < N; ++i) nj = Num<T>::fma_(A[i][j], A[i][j], nj);
							if (nj > best) { best = nj; kb = j; }
						}
						T v[N];
						OSC_UNROLL for (int i = 0; i < N; ++i) {
							T x = T(0);
							OSC_UNROLL for (int j = 0; j < N; ++j) x = (j == kb) ? A[i][j] : x;
							v[i] = x;
						}
						// re-orthogonalise against the basis so far ("twice is enough"), then normalise
						OSC_UNROLL for (int c = 0; c < N - 1; ++c) {
							if (c < step) {
								T d = T(0);
								OSC_UNROLL for (int i = 0; i < N; ++i) d = Num<T>::fma_(Rg[i][c], v[i], d);
								OSC_UNROLL for (int i = 0; i < N; ++i) v[i] = Num<T>::fma_(-d, Rg[i][c], v[i]);
							}
						}
						T nv = T(0);
						OSC_UNROLL for (int i = 0; i < N; ++i) nv = Num<T>::fma_(v[i], v[i], nv);
						const T inv = Num<T>::rsqrt_(nv);
						OSC_UNROLL for (int i = 0; i < N; ++i) { v[i] *= inv; Rg[i][step] = v[i]; }
						OSC_UNROLL for (int j = 0; j < N; ++j) {
							T d = T(0);
							OSC_UNROLL for (int i = 0; i < N; ++i) d = Num<T>::fma_(v[i], A[i][j], d);
							OSC_UNROLL for (int i = 0; i < N; ++i) A[i][j] = Num<T>::fma_(-d, v[i], A[i][j]);
						}
					}
				}
			}
		}
	}
	if (!fast && fro >= T(1e-3)) {
		T A[N][N], W[N][N], s[N];  // A = Jp^T (n x k): Jp = W S A^T  ->  U = W
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < N; ++j) A[i][j] = Jp[j][i];
		if (!svd_onesided<T, N, N>(A, W, s)) status |= 8;
		if (s[0] >= T(1e-3)) {
			rank = k;
			bool stop = false;
			OSC_UNROLL for (int i = N - 1; i > 0; --i) {
				if (i < k && !stop) {
					if (s[i] / s[0] < T(1e-3)) --rank;
					else stop = true;
				}
			}
			if (rank == k) {
				OSC_UNROLL for (int i = 0; i < N; ++i) Rg[i][i] = (i < k) ? T(1) : T(0);
			} else {
				OSC_UNROLL for (int i = 0; i < N; ++i)
					OSC_UNROLL for (int j = 0; j < N; ++j) Rg[i][j] = (j < rank) ? W[i][j] : T(0);
			}
		}
	}
	tI(L::RANK) = rank;
	st_mat<T, N, N>(tS, L::RANGE, Rg);
	T Nt[N][N];
	if (rank == 0) {  // JointTask.cpp:234-239
		OSC_UNROLL for (int i = 0; i < N; ++i)
			OSC_UNROLL for (int j = 0; j < N; ++j) Nt[i][j] = (i == j) ? T(1) : T(0);
	} else {
		const unsigned mask = (1u << rank) - 1u;
		T Jr[N][N], Lam[N][N], Minv[N][N];
		mtm<T, N, N, N>(Rg, Jp, Jr);  // R^T Jp, rows >= rank are zero
		ld_mat<T, N, N>(rS, RobotL<N>::MINV, Minv);
		if (!op_space<T, N, N>(Jr, Minv, mask, Lam, Nt)) status |= 16;
		st_mat<T, N, N>(tS, L::LAM, Lam);
		if (jc.decoupling == 1) {  // BOUNDED_INERTIA_ESTIMATES, JointTask.cpp:253-267
			T Wm[N][N], A[N][N], Lm[N][N];
			bie_inverse<T, N>(rS, jc.bie_threshold, Wm);
			congruence<T, N, N>(Jr, Wm, A);
			spd_inverse_masked<T, N>(A, mask, Lm);
			st_mat<T, N, N>(tS, L::LAM_MOD, Lm);
		} else if (jc.decoupling == 2) {
			OSC_UNROLL for (int i = 0; i < N; ++i)
				OSC_UNROLL for (int j = 0; j < N; ++j) tS(L::LAM_MOD + i * N + j) = (i == j && i < rank) ? T(1) : T(0);
		} else {
			st_mat<T, N, N>(tS, L::LAM_MOD, Lam);
		}
	}
	st_mat<T, N, N>(tS, L::NT, Nt);
	{
		T NTP[N][N];
		mm<T, N, N, N>(Nt, Np, NTP);
		st_mat<T, N, N>(tS, L::NTP, NTP);
	}
	if (status) atomicOr(&rob.I[(int64_t)RobotL<N>::STATUS * stride + r], status);
}

template <class T, int N, class S = T>
__global__ void __launch_bounds__(OSC_THREADS, OSC_PERCALL_BLOCKS) k_jt_update(const __grid_constant__ JtC<T, N> jc, Arena<S> rob, Arena<S> tk,
															const S* nprec_src, int64_t B, int64_t stride) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	jt_update_robot<T, N>(jc, rob, tk, nprec_src, stride, r);
}

template <class T, int N, class S = T, bool REDIR = false>
__device__ void jt_compute_robot(const JtC<T, N>& jc, const Arena<S>& rob, const Arena<S>& tk, int64_t stride, int64_t r, int with_tau_prec,
								 int accumulate, int integrated_already = 0) {
	using L = JtL<N>;
	const Row<T, S> rP{rob.P, stride, r}, rS = srow<T, REDIR>(rob, stride, r), tP{tk.P, stride, r}, tS = srow<T, REDIR>(tk, stride, r);
	const Row<int32_t> tI{tk.I, stride, r};
	T q[N], dq[N], cp[N], cv[N];
	ld_vec<T, N>(rP, RobotL<N>::Q, q);
	ld_vec<T, N>(rP, RobotL<N>::DQ, dq);
	if (jc.full) {
		OSC_UNROLL for (int i = 0; i < N; ++i) { cp[i] = q[i]; cv[i] = dq[i]; }
	} else {
		mv<T, N, N>(jc.S, q, cp);
		mv<T, N, N>(jc.S, dq, cv);
	}
	st_vec<T, N>(tS, L::CUR_POS, cp);
	st_vec<T, N>(tS, L::CUR_VEL, cv);
	const int rank = tI(L::RANK);
	T tau[N];
	OSC_UNROLL for (int i = 0; i < N; ++i) tau[i] = T(0);
	if (rank > 0) {
		T u[N], acc[N];
		jt_pd_law<T, N>(jc, tP, tS, cp, cv, u, acc, true, !integrated_already);  // k_tick_fused2 integrates before it can fall back
		T Rg[N][N], Lam[N][N], Lm[N][N], Np[N][N];
		ld_mat<T, N, N>(tS, L::RANGE, Rg);
		ld_mat<T, N, N>(tS, L::LAM, Lam);
		ld_mat<T, N, N>(tS, L::LAM_MOD, Lm);
		ld_mat<T, N, N>(tS, L::NPREC, Np);
		T ra[N], ru[N], f1[N], f2[N], f[N];
		mtv<T, N, N>(Rg, acc, ra);
		mtv<T, N, N>(Rg, u, ru);
		mv<T, N, N>(Lam, ra, f1);
		mv<T, N, N>(Lm, ru, f2);
		OSC_UNROLL for (int i = 0; i < N; ++i) f[i] = f1[i] + f2[i];
		if (with_tau_prec) {  // JointTask.cpp:285-292: - Jp^T R Lambda R^T S Minv tau_prec
			T tp[N], Minv[N][N], Mt[N], St[N], rt[N], lt[N];
			ld_vec<T, N>(rS, RobotL<N>::TAU, tp);
			ld_mat<T, N, N>(rS, RobotL<N>::MINV, Minv);
			mv<T, N, N>(Minv, tp, Mt);
			if (jc.full) {
				OSC_UNROLL for (int i = 0; i < N; ++i) St[i] = Mt[i];
			} else {
				mv<T, N, N>(jc.S, Mt, St);
			}
			mtv<T, N, N>(Rg, St, rt);
			mv<T, N, N>(Lam, rt, lt);
			OSC_UNROLL for (int i = 0; i < N; ++i) f[i] -= lt[i];
		}
		// Jp^T R f with Jp = S N_prec
		T Rf[N], SRf[N];
		mv<T, N, N>(Rg, f, Rf);
		if (jc.full) {
			OSC_UNROLL for (int i = 0; i < N; ++i) SRf[i] = Rf[i];
		} else {
			mtv<T, N, N>(jc.S, Rf, SRf);
		}
		mtv<T, N, N>(Np, SRf, tau);
	}
	st_vec<T, N>(tS, L::TASK_TAU, tau);
	if (accumulate == 1) {
		st_vec<T, N>(rS, RobotL<N>::TAU, tau);
	} else if (accumulate == 2) {
		OSC_UNROLL for (int i = 0; i < N; ++i) rS(RobotL<N>::TAU + i) += tau[i];
	}
}

template <class T, int N, class S = T>
__global__ void __launch_bounds__(OSC_THREADS, OSC_PERCALL_BLOCKS) k_jt_compute(const __grid_constant__ JtC<T, N> jc, Arena<S> rob, Arena<S> tk, int64_t B,
															 int64_t stride, int with_tau_prec, int accumulate) {
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	jt_compute_robot<T, N>(jc, rob, tk, stride, r, with_tau_prec, accumulate);
}

}  // namespace osc
