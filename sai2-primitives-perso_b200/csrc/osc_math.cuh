// osc_math.cuh -- per-thread small dense linear algebra for the batched OSC kernels (sm_100a).
//
// Design: ONE robot per thread.  Every matrix is a fixed-shape register/local array indexed with compile-time
// constants inside fully unrolled loops, so ptxas keeps the working set in registers and the FP64 pipe sees no idle
// lanes (a warp-per-robot mapping would leave 25/32 lanes idle on 7x7 data).  Variable ranks (task
// rank, singular split, joint-task range) are handled by MASKING fixed-shape operands, never by run-time shapes:
// no divergence inside the linear algebra and no dynamically indexed local memory.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#define OSC_DEV __device__ __forceinline__
#define OSC_UNROLL _Pragma("unroll")

namespace osc {

template <class T> struct Num;
template <> struct Num<double> {
	static OSC_DEV double rsqrt_(double x) { return ::rsqrt(x); }
	static OSC_DEV double sqrt_(double x) { return ::sqrt(x); }
	static OSC_DEV double abs_(double x) { return ::fabs(x); }
	static OSC_DEV void sincos_(double x, double* s, double* c) { ::sincos(x, s, c); }
	static OSC_DEV double fma_(double a, double b, double c) { return ::fma(a, b, c); }
	static OSC_DEV double eps() { return 2.220446049250313e-16; }
	static OSC_DEV double jacobi_tol() { return 1e-15; }  // ~ sqrt(rows) eps (LAPACK dgesvj); below eps the test can never pass
	static OSC_DEV double tiny_sv() { return 1e-14; }
};
template <> struct Num<float> {
	static OSC_DEV float rsqrt_(float x) { return ::rsqrtf(x); }
	static OSC_DEV float sqrt_(float x) { return ::sqrtf(x); }
	static OSC_DEV float abs_(float x) { return ::fabsf(x); }
	static OSC_DEV void sincos_(float x, float* s, float* c) { ::sincosf(x, s, c); }
	static OSC_DEV float fma_(float a, float b, float c) { return ::fmaf(a, b, c); }
	static OSC_DEV float eps() { return 1.1920929e-7f; }
	static OSC_DEV float jacobi_tol() { return 5e-7f; }
	static OSC_DEV float tiny_sv() { return 1e-6f; }
};

// ---- 3-vectors / 3x3 (row-major m[r][c]) ----
template <class T> OSC_DEV void cross3(const T a[3], const T b[3], T o[3]) {
	o[0] = a[1] * b[2] - a[2] * b[1];
	o[1] = a[2] * b[0] - a[0] * b[2];
	o[2] = a[0] * b[1] - a[1] * b[0];
}
template <class T> OSC_DEV T dot3(const T a[3], const T b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <class T> OSC_DEV T norm3(const T a[3]) { return Num<T>::sqrt_(dot3(a, a)); }
template <class T> OSC_DEV void mv3(const T m[3][3], const T v[3], T o[3]) {
	OSC_UNROLL for (int i = 0; i < 3; ++i) o[i] = m[i][0] * v[0] + m[i][1] * v[1] + m[i][2] * v[2];
}
template <class T> OSC_DEV void mtv3(const T m[3][3], const T v[3], T o[3]) {  // m^T v
	OSC_UNROLL for (int i = 0; i < 3; ++i) o[i] = m[0][i] * v[0] + m[1][i] * v[1] + m[2][i] * v[2];
}
template <class T> OSC_DEV void mm3(const T a[3][3], const T b[3][3], T o[3][3]) {
	OSC_UNROLL for (int i = 0; i < 3; ++i)
		OSC_UNROLL for (int j = 0; j < 3; ++j) o[i][j] = a[i][0] * b[0][j] + a[i][1] * b[1][j] + a[i][2] * b[2][j];
}
// Sai2Model::orientationError: -1/2 sum_i R[:,i] x Rd[:,i]
template <class T> OSC_DEV void orientation_error(const T Rd[3][3], const T R[3][3], T e[3]) {
	e[0] = e[1] = e[2] = T(0);
	OSC_UNROLL for (int i = 0; i < 3; ++i) {
		const T a[3] = {R[0][i], R[1][i], R[2][i]}, b[3] = {Rd[0][i], Rd[1][i], Rd[2][i]};
		T c[3];
		cross3(a, b, c);
		e[0] += c[0]; e[1] += c[1]; e[2] += c[2];
	}
	e[0] *= T(-0.5); e[1] *= T(-0.5); e[2] *= T(-0.5);
}

// ---- generic fixed-shape products ----
// C[R][C] = A[R][K] * B[K][C]
template <class T, int R, int K, int C> OSC_DEV void mm(const T (&A)[R][K], const T (&B)[K][C], T (&O)[R][C]) {
	OSC_UNROLL for (int i = 0; i < R; ++i)
		OSC_UNROLL for (int j = 0; j < C; ++j) {
			T s = T(0);
			OSC_UNROLL for (int k = 0; k < K; ++k) s = Num<T>::fma_(A[i][k], B[k][j], s);
			O[i][j] = s;
		}
}
// O[R][C] = A[K][R]^T * B[K][C]
template <class T, int K, int R, int C> OSC_DEV void mtm(const T (&A)[K][R], const T (&B)[K][C], T (&O)[R][C]) {
	OSC_UNROLL for (int i = 0; i < R; ++i)
		OSC_UNROLL for (int j = 0; j < C; ++j) {
			T s = T(0);
			OSC_UNROLL for (int k = 0; k < K; ++k) s = Num<T>::fma_(A[k][i], B[k][j], s);
			O[i][j] = s;
		}
}
// O[R][C] = A[R][K] * B[C][K]^T
template <class T, int R, int K, int C> OSC_DEV void mmt(const T (&A)[R][K], const T (&B)[C][K], T (&O)[R][C]) {
	OSC_UNROLL for (int i = 0; i < R; ++i)
		OSC_UNROLL for (int j = 0; j < C; ++j) {
			T s = T(0);
			OSC_UNROLL for (int k = 0; k < K; ++k) s = Num<T>::fma_(A[i][k], B[j][k], s);
			O[i][j] = s;
		}
}
template <class T, int R, int C> OSC_DEV void mv(const T (&A)[R][C], const T (&x)[C], T (&y)[R]) {
	OSC_UNROLL for (int i = 0; i < R; ++i) {
		T s = T(0);
		OSC_UNROLL for (int k = 0; k < C; ++k) s = Num<T>::fma_(A[i][k], x[k], s);
		y[i] = s;
	}
}
template <class T, int R, int C> OSC_DEV void mtv(const T (&A)[R][C], const T (&x)[R], T (&y)[C]) {  // A^T x
	OSC_UNROLL for (int j = 0; j < C; ++j) {
		T s = T(0);
		OSC_UNROLL for (int k = 0; k < R; ++k) s = Num<T>::fma_(A[k][j], x[k], s);
		y[j] = s;
	}
}
// S = A W A^T, A[R][K], W[K][K] symmetric -> S[R][R] symmetric (both triangles written)
template <class T, int R, int K> OSC_DEV void congruence(const T (&A)[R][K], const T (&W)[K][K], T (&S)[R][R]) {
	T AW[R][K];
	mm<T, R, K, K>(A, W, AW);
	OSC_UNROLL for (int i = 0; i < R; ++i)
		OSC_UNROLL for (int j = 0; j <= i; ++j) {
			T s = T(0);
			OSC_UNROLL for (int k = 0; k < K; ++k) s = Num<T>::fma_(AW[i][k], A[j][k], s);
			S[i][j] = s; S[j][i] = s;
		}
}

// ---- symmetric positive definite inverse through Cholesky (stands in for llt().solve(I) and, on the SPD matrices the
// reference feeds it, for PartialPivLU inverse()).  Returns false when a pivot is not positive (status bit). ----
template <class T, int N> OSC_DEV bool spd_inverse(const T (&A)[N][N], T (&Ainv)[N][N]) {
	T L[N][N];   // lower factor; diagonal holds 1/L_jj
	bool ok = true;
	OSC_UNROLL for (int j = 0; j < N; ++j) {
		T d = A[j][j];
		OSC_UNROLL for (int k = 0; k < j; ++k) d = Num<T>::fma_(-L[j][k], L[j][k], d);
		ok = ok && (d > T(0));
		const T inv = Num<T>::rsqrt_(d);
		L[j][j] = inv;
		OSC_UNROLL for (int i = j + 1; i < N; ++i) {
			T s = A[i][j];
			OSC_UNROLL for (int k = 0; k < j; ++k) s = Num<T>::fma_(-L[i][k], L[j][k], s);
			L[i][j] = s * inv;
		}
	}
	// Linv (lower): column by column
	T Li[N][N];
	OSC_UNROLL for (int c = 0; c < N; ++c) {
		OSC_UNROLL for (int i = 0; i < N; ++i) {
			if (i < c) { Li[i][c] = T(0); continue; }
			T s = (i == c) ? T(1) : T(0);
			OSC_UNROLL for (int k = c; k < i; ++k) s = Num<T>::fma_(-L[i][k], Li[k][c], s);
			Li[i][c] = s * L[i][i];
		}
	}
	// Ainv = Li^T Li
	OSC_UNROLL for (int i = 0; i < N; ++i)
		OSC_UNROLL for (int j = 0; j <= i; ++j) {
			T s = T(0);
			OSC_UNROLL for (int k = i; k < N; ++k) s = Num<T>::fma_(Li[k][i], Li[k][j], s);
			Ainv[i][j] = s; Ainv[j][i] = s;
		}
	return ok;
}

// Masked SPD inverse: rows/cols with mask bit 0 are treated as absent (unit diagonal inside, zero in the result).
template <class T, int N> OSC_DEV bool spd_inverse_masked(T (&A)[N][N], unsigned mask, T (&Ainv)[N][N]) {
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		const bool on = (mask >> i) & 1u;
		OSC_UNROLL for (int j = 0; j < N; ++j) {
			const bool onj = (mask >> j) & 1u;
			if (!(on && onj)) A[i][j] = (i == j) ? T(1) : T(0);
		}
	}
	const bool ok = spd_inverse<T, N>(A, Ainv);
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		const bool on = (mask >> i) & 1u;
		if (!on) Ainv[i][i] = T(0);
	}
	return ok;
}

// ---- triangular kernels of the fused tick (osc_fused.cuh) ----
// Packed lower factor: A = L L^T with L[i][k] (k < i) as is, 1/L[i][i] in Li[i] and L[i][i] in Ld[i].
template <class T, int N> struct Chol {
	T L[N][N];  // strictly lower part used
	T Li[N];    // 1 / L_ii
	T Ld[N];    // L_ii
};
// A symmetric (lower triangle read).  Returns false when a pivot is not positive.
template <class T, int N> OSC_DEV bool chol_factor(const T (&A)[N][N], Chol<T, N>& c) {
	bool ok = true;
	OSC_UNROLL for (int j = 0; j < N; ++j) {
		T d = A[j][j];
		OSC_UNROLL for (int k = 0; k < j; ++k) d = Num<T>::fma_(-c.L[j][k], c.L[j][k], d);
		ok = ok && (d > T(0));
		const T inv = Num<T>::rsqrt_(d);
		c.Li[j] = inv;
		c.Ld[j] = d * inv;
		OSC_UNROLL for (int i = j + 1; i < N; ++i) {
			T s = A[i][j];
			OSC_UNROLL for (int k = 0; k < j; ++k) s = Num<T>::fma_(-c.L[i][k], c.L[j][k], s);
			c.L[i][j] = s * inv;
		}
	}
	return ok;
}
// In-place Cholesky of the lower triangle of A: on exit A[i][j] (j < i) = L[i][j], li[i] = 1 / L_ii, A[i][i] = L_ii.
template <class T, int N> OSC_DEV bool chol_inplace(T (&A)[N][N], T (&li)[N]) {
	bool ok = true;
	OSC_UNROLL for (int j = 0; j < N; ++j) {
		T d = A[j][j];
		OSC_UNROLL for (int k = 0; k < j; ++k) d = Num<T>::fma_(-A[j][k], A[j][k], d);
		ok = ok && (d > T(0));
		const T inv = Num<T>::rsqrt_(d);
		li[j] = inv;
		A[j][j] = d * inv;
		OSC_UNROLL for (int i = j + 1; i < N; ++i) {
			T s = A[i][j];
			OSC_UNROLL for (int k = 0; k < j; ++k) s = Num<T>::fma_(-A[i][k], A[j][k], s);
			A[i][j] = s * inv;
		}
	}
	return ok;
}
// x <- L^-1 x
template <class T, int N> OSC_DEV void chol_fwd(const Chol<T, N>& c, T (&x)[N]) {
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		T s = x[i];
		OSC_UNROLL for (int k = 0; k < i; ++k) s = Num<T>::fma_(-c.L[i][k], x[k], s);
		x[i] = s * c.Li[i];
	}
}
// x <- L^-T x
template <class T, int N> OSC_DEV void chol_bwd(const Chol<T, N>& c, T (&x)[N]) {
	OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
		T s = x[i];
		OSC_UNROLL for (int k = i + 1; k < N; ++k) s = Num<T>::fma_(-c.L[k][i], x[k], s);
		x[i] = s * c.Li[i];
	}
}
// y = L x
template <class T, int N> OSC_DEV void chol_mul(const Chol<T, N>& c, const T (&x)[N], T (&y)[N]) {
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		T s = c.Ld[i] * x[i];
		OSC_UNROLL for (int k = 0; k < i; ++k) s = Num<T>::fma_(c.L[i][k], x[k], s);
		y[i] = s;
	}
}
// true when A - shift * I is positive definite (pivots of its LDL^T all positive): a Sylvester-inertia test
template <class T, int N> OSC_DEV bool spd_above(const T (&A)[N][N], T shift) {
	T L[N][N], dinv[N];
	bool ok = true;
	OSC_UNROLL for (int j = 0; j < N; ++j) {
		T d = A[j][j] - shift;
		OSC_UNROLL for (int k = 0; k < j; ++k) d = Num<T>::fma_(-L[j][k] * L[j][k], dinv[k], d);  // L holds l_jk * d_k
		ok = ok && (d > T(0));
		dinv[j] = T(1) / d;
		OSC_UNROLL for (int i = j + 1; i < N; ++i) {
			T s = A[i][j];
			OSC_UNROLL for (int k = 0; k < j; ++k) s = Num<T>::fma_(-L[i][k] * L[j][k], dinv[k], s);
			L[i][j] = s;
		}
	}
	return ok;
}

// ---- one-sided (Hestenes) Jacobi SVD of A[R][C], R >= C:  A = Q diag(s) W^T ----
// On exit: A's columns hold Q (normalised, numerically-null columns completed to an orthonormal set), W[C][C] is
// orthogonal, s[C] descending.  Stands in for Eigen::JacobiSVD (SingularityHandler.cpp:78, matrixRangeBasis).
// Returns false if the sweep limit was hit.
template <class T, int R, int C> OSC_DEV bool svd_onesided(T (&A)[R][C], T (&W)[C][C], T (&s)[C]) {
	OSC_UNROLL for (int i = 0; i < C; ++i)
		OSC_UNROLL for (int j = 0; j < C; ++j) W[i][j] = (i == j) ? T(1) : T(0);
	// columns below tiny_sv x |A|_F are numerically null: they are replaced by the Gram-Schmidt completion below, and
	// rotating rounding noise against rounding noise never meets the relative test (a rank-deficient N_prec would spin
	// through all 40 sweeps and raise OSC_STATUS_SVD_NOT_CONVERGED for nothing)
	T negl = T(0);
	OSC_UNROLL for (int i = 0; i < R; ++i)
		OSC_UNROLL for (int j = 0; j < C; ++j) negl = Num<T>::fma_(A[i][j], A[i][j], negl);
	negl *= Num<T>::tiny_sv() * Num<T>::tiny_sv();
	bool converged = false;
	for (int sweep = 0; sweep < 40 && !converged; ++sweep) {
		bool rotated = false;
		OSC_UNROLL for (int p = 0; p < C - 1; ++p) {
			OSC_UNROLL for (int q = p + 1; q < C; ++q) {
				T al = T(0), be = T(0), ga = T(0);
				OSC_UNROLL for (int i = 0; i < R; ++i) {
					al = Num<T>::fma_(A[i][p], A[i][p], al);
					be = Num<T>::fma_(A[i][q], A[i][q], be);
					ga = Num<T>::fma_(A[i][p], A[i][q], ga);
				}
				const bool skip = (ga == T(0)) || al <= negl || be <= negl || (Num<T>::abs_(ga) <= Num<T>::jacobi_tol() * Num<T>::sqrt_(al * be));
				if (!skip) {
					rotated = true;
					const T zeta = (be - al) / (T(2) * ga);
					const T t = (zeta >= T(0) ? T(1) : T(-1)) / (Num<T>::abs_(zeta) + Num<T>::sqrt_(T(1) + zeta * zeta));
					const T cs = Num<T>::rsqrt_(T(1) + t * t), sn = cs * t;
					OSC_UNROLL for (int i = 0; i < R; ++i) {
						const T wp = A[i][p], wq = A[i][q];
						A[i][p] = cs * wp - sn * wq;
						A[i][q] = sn * wp + cs * wq;
					}
					OSC_UNROLL for (int i = 0; i < C; ++i) {
						const T rp = W[i][p], rq = W[i][q];
						W[i][p] = cs * rp - sn * rq;
						W[i][q] = sn * rp + cs * rq;
					}
				}
			}
		}
		converged = !rotated;
	}
	OSC_UNROLL for (int j = 0; j < C; ++j) {
		T s2 = T(0);
		OSC_UNROLL for (int i = 0; i < R; ++i) s2 = Num<T>::fma_(A[i][j], A[i][j], s2);
		s[j] = Num<T>::sqrt_(s2);
	}
	// sort descending with a fixed compare-exchange network (odd-even transposition): static indexing only
	OSC_UNROLL for (int pass = 0; pass < C; ++pass) {
		OSC_UNROLL for (int j = (pass & 1); j + 1 < C; j += 2) {
			if (s[j] < s[j + 1]) {
				T ts = s[j]; s[j] = s[j + 1]; s[j + 1] = ts;
				OSC_UNROLL for (int i = 0; i < R; ++i) { T x = A[i][j]; A[i][j] = A[i][j + 1]; A[i][j + 1] = x; }
				OSC_UNROLL for (int i = 0; i < C; ++i) { T x = W[i][j]; W[i][j] = W[i][j + 1]; W[i][j + 1] = x; }
			}
		}
	}
	// normalise columns; complete numerically-null ones by Gram-Schmidt on unit vectors
	const T tiny = Num<T>::tiny_sv() * s[0];
	OSC_UNROLL for (int j = 0; j < C; ++j) {
		if (s[j] > tiny && s[j] > T(0)) {
			const T inv = T(1) / s[j];
			OSC_UNROLL for (int i = 0; i < R; ++i) A[i][j] *= inv;
		} else {
			bool done = false;
			_Pragma("unroll 1") for (int e = 0; e < R; ++e) {
				if (!done) {
					T v[R];
					OSC_UNROLL for (int i = 0; i < R; ++i) v[i] = (i == e) ? T(1) : T(0);
					_Pragma("unroll 1") for (int pass = 0; pass < 2; ++pass)
						OSC_UNROLL for (int c = 0; c < C; ++c) {
							if (c < j) {
								T d = T(0);
								OSC_UNROLL for (int i = 0; i < R; ++i) d = Num<T>::fma_(A[i][c], v[i], d);
								OSC_UNROLL for (int i = 0; i < R; ++i) v[i] = Num<T>::fma_(-d, A[i][c], v[i]);
							}
						}
					T nv = T(0);
					OSC_UNROLL for (int i = 0; i < R; ++i) nv = Num<T>::fma_(v[i], v[i], nv);
					if (nv > T(0.25)) {
						const T inv = Num<T>::rsqrt_(nv);
						OSC_UNROLL for (int i = 0; i < R; ++i) A[i][j] = v[i] * inv;
						done = true;
					}
				}
			}
		}
	}
	return converged;
}

}  // namespace osc
