// osc_oracle.hpp -- CPU restatement of the sai2-primitives control tick.  TEST INFRASTRUCTURE ONLY.
//
// PARITY UNPINNED: the reference has no tests / golden vectors for this path and cannot be compiled here (it needs
// Eigen3 and the un-vendored sai2-model -> RBDL), so this oracle is a restatement pinned only by
// (a) physics identities, (b) the independent numpy/LAPACK oracle oracle/np_oracle.py, (c) the golden vectors under
// tests/golden/ that those two agree on, and (d) the code-independent pins tests/golden/pins_*.npz (sympy closed forms, 50-digit
// mpmath models of M, J, Lambda, N, the control law, POPC, the SingularityHandler blend and the rollout step typed in from the
// reference text: tools/make_sympy_golden.py, replayed by tests/test_pins.py at 1e-9).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may build, load or call anything in this directory.  The product (libosc_b200.so) never does.
//
// Every function cites the reference file:line it follows (paths relative to /root/reference/src unless noted).
// sai2-model functions (not in the tree) follow SURVEY.md Appendix B; quirks Q1-Q13 follow SURVEY.md A.9.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <limits>
#include <stdexcept>
#include <vector>

#include "../include/osc_b200.h"

#include "otg_oracle.hpp"

namespace orc {

// ---- operation counter ("freeze the algorithmic FLOP count by instrumenting the CPU oracle") ----
// -DORACLE_COUNT_FLOPS: every matrix / vector primitive below adds its floating-point adds and multiplies (FMA = 2,
// divisions and square roots = 1, transcendental calls not counted) to a thread-local counter; orc_count_tick_flops()
// reads it around one tick.  Counting is at the granularity of the primitives (a product counts 2 r k c whatever the zeros).
#ifdef ORACLE_COUNT_FLOPS
inline long long& flop_counter() { static thread_local long long c = 0; return c; }
#define ORC_OPS(n) (orc::flop_counter() += (long long)(n))
#else
#define ORC_OPS(n) ((void)0)
#endif

// -DORACLE_HEAP_TEMPORARIES: every matrix temporary owns a malloc'ed buffer, like the dynamic-size Eigen::MatrixXd
// temporaries of the reference (one allocation per expression result): a closer proxy of the reference's CPU cost than the
// default fixed-capacity storage (SURVEY.md 8d "Timing the reference CPU path beside it").
#ifdef ORACLE_HEAP_TEMPORARIES
struct MatBuf {
	double* p;
	MatBuf() : p((double*)std::malloc(64 * sizeof(double))) {}
	MatBuf(const MatBuf& o) : p((double*)std::malloc(64 * sizeof(double))) { std::memcpy(p, o.p, 64 * sizeof(double)); }
	MatBuf& operator=(const MatBuf& o) { if (this != &o) std::memcpy(p, o.p, 64 * sizeof(double)); return *this; }
	~MatBuf() { std::free(p); }
	double& operator[](int i) { return p[i]; }
	double operator[](int i) const { return p[i]; }
};
#endif

// ------------------------------------------------------------------------------------------------------------------
// Tiny dense matrix with run-time shape (stands in for Eigen::MatrixXd on <= 8x8 data; no heap unless
// ORACLE_HEAP_TEMPORARIES).
// ------------------------------------------------------------------------------------------------------------------
struct Mat {
	int r = 0, c = 0;
#ifdef ORACLE_HEAP_TEMPORARIES
	MatBuf a;
#else
	double a[64];
#endif
	Mat() {}
	Mat(int r_, int c_) : r(r_), c(c_) { for (int i = 0; i < r * c; ++i) a[i] = 0.0; }
	double& operator()(int i, int j) { return a[i * c + j]; }
	double operator()(int i, int j) const { return a[i * c + j]; }
	static Mat Zero(int r, int c) { return Mat(r, c); }
	static Mat Identity(int n) {
		Mat m(n, n);
		for (int i = 0; i < n; ++i) m(i, i) = 1;
		return m;
	}
	Mat T() const {
		Mat t(c, r);
		for (int i = 0; i < r; ++i)
			for (int j = 0; j < c; ++j) t(j, i) = (*this)(i, j);
		return t;
	}
	double norm() const {
		ORC_OPS(2 * r * c + 1);
		double s = 0;
		for (int i = 0; i < r * c; ++i) s += a[i] * a[i];
		return std::sqrt(s);
	}
	Mat leftCols(int k) const { return block(0, 0, r, k); }
	Mat block(int i0, int j0, int nr, int nc) const {
		Mat b(nr, nc);
		for (int i = 0; i < nr; ++i)
			for (int j = 0; j < nc; ++j) b(i, j) = (*this)(i0 + i, j0 + j);
		return b;
	}
	Mat col(int j) const { return block(0, j, r, 1); }
	void setBlock(int i0, int j0, const Mat& b) {
		for (int i = 0; i < b.r; ++i)
			for (int j = 0; j < b.c; ++j) (*this)(i0 + i, j0 + j) = b(i, j);
	}
};
inline Mat operator*(const Mat& A, const Mat& B) {
	ORC_OPS(2LL * A.r * A.c * B.c);
	Mat C(A.r, B.c);
	for (int i = 0; i < A.r; ++i)
		for (int k = 0; k < A.c; ++k) {
			double aik = A(i, k);
			for (int j = 0; j < B.c; ++j) C(i, j) += aik * B(k, j);
		}
	return C;
}
inline Mat operator*(double s, const Mat& A) {
	ORC_OPS(A.r * A.c);
	Mat C = A;
	for (int i = 0; i < A.r * A.c; ++i) C.a[i] *= s;
	return C;
}
inline Mat operator+(const Mat& A, const Mat& B) {
	ORC_OPS(A.r * A.c);
	Mat C = A;
	for (int i = 0; i < A.r * A.c; ++i) C.a[i] += B.a[i];
	return C;
}
inline Mat operator-(const Mat& A, const Mat& B) {
	ORC_OPS(A.r * A.c);
	Mat C = A;
	for (int i = 0; i < A.r * A.c; ++i) C.a[i] -= B.a[i];
	return C;
}
inline Mat vec(const double* v, int n) {
	Mat m(n, 1);
	for (int i = 0; i < n; ++i) m.a[i] = v[i];
	return m;
}

struct V3 {
	double x = 0, y = 0, z = 0;
	V3() {}
	V3(double a, double b, double c) : x(a), y(b), z(c) {}
	explicit V3(const double* p) : x(p[0]), y(p[1]), z(p[2]) {}
	double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
	double& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
	double norm() const { ORC_OPS(6); return std::sqrt(x * x + y * y + z * z); }
	double dot(const V3& o) const { ORC_OPS(5); return x * o.x + y * o.y + z * o.z; }
	V3 cross(const V3& o) const { ORC_OPS(9); return V3(y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x); }
};
inline V3 operator+(const V3& a, const V3& b) { ORC_OPS(3); return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(const V3& a, const V3& b) { ORC_OPS(3); return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(const V3& a) { return V3(-a.x, -a.y, -a.z); }
inline V3 operator*(double s, const V3& a) { ORC_OPS(3); return V3(s * a.x, s * a.y, s * a.z); }
inline V3 operator*(const V3& a, double s) { ORC_OPS(3); return V3(s * a.x, s * a.y, s * a.z); }
inline V3 cwise(const V3& a, const V3& b) { ORC_OPS(3); return V3(a.x * b.x, a.y * b.y, a.z * b.z); }

struct M3 {
	double m[3][3];
	M3() { std::memset(m, 0, sizeof m); }
	explicit M3(const double* p) {
		for (int i = 0; i < 9; ++i) m[i / 3][i % 3] = p[i];
	}
	static M3 Identity() {
		M3 r;
		r.m[0][0] = r.m[1][1] = r.m[2][2] = 1;
		return r;
	}
	static M3 Diag(const V3& d) {
		M3 r;
		r.m[0][0] = d.x; r.m[1][1] = d.y; r.m[2][2] = d.z;
		return r;
	}
	static M3 Outer(const V3& a, const V3& b) {
		ORC_OPS(9);
		M3 r;
		for (int i = 0; i < 3; ++i)
			for (int j = 0; j < 3; ++j) r.m[i][j] = a[i] * b[j];
		return r;
	}
	M3 T() const {
		M3 r;
		for (int i = 0; i < 3; ++i)
			for (int j = 0; j < 3; ++j) r.m[i][j] = m[j][i];
		return r;
	}
	V3 col(int j) const { return V3(m[0][j], m[1][j], m[2][j]); }
};
inline M3 operator*(const M3& A, const M3& B) {
	ORC_OPS(45);
	M3 C;
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j)
			for (int k = 0; k < 3; ++k) C.m[i][j] += A.m[i][k] * B.m[k][j];
	return C;
}
inline V3 operator*(const M3& A, const V3& v) {
	ORC_OPS(15);
	return V3(A.m[0][0] * v.x + A.m[0][1] * v.y + A.m[0][2] * v.z, A.m[1][0] * v.x + A.m[1][1] * v.y + A.m[1][2] * v.z,
			  A.m[2][0] * v.x + A.m[2][1] * v.y + A.m[2][2] * v.z);
}
inline M3 operator+(const M3& A, const M3& B) {
	ORC_OPS(9);
	M3 C;
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j) C.m[i][j] = A.m[i][j] + B.m[i][j];
	return C;
}
inline M3 operator-(const M3& A, const M3& B) {
	ORC_OPS(9);
	M3 C;
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j) C.m[i][j] = A.m[i][j] - B.m[i][j];
	return C;
}
inline M3 operator*(double s, const M3& A) {
	ORC_OPS(9);
	M3 C;
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j) C.m[i][j] = s * A.m[i][j];
	return C;
}

// ------------------------------------------------------------------------------------------------------------------
// Eigen stand-ins
// ------------------------------------------------------------------------------------------------------------------
// A.llt().solve(Identity): Lambda in Sai2Model::operationalSpaceMatrices
inline Mat llt_solve_identity(const Mat& A) {
	const int n = A.r;
	ORC_OPS(n * n * n / 3 + 2LL * n * n * n);
	Mat L(n, n);
	for (int j = 0; j < n; ++j) {
		double d = A(j, j);
		for (int k = 0; k < j; ++k) d -= L(j, k) * L(j, k);
		d = std::sqrt(d);
		L(j, j) = d;
		for (int i = j + 1; i < n; ++i) {
			double s = A(i, j);
			for (int k = 0; k < j; ++k) s -= L(i, k) * L(j, k);
			L(i, j) = s / d;
		}
	}
	Mat X(n, n);
	for (int c = 0; c < n; ++c) {
		double y[8];
		for (int i = 0; i < n; ++i) {
			double s = (i == c) ? 1.0 : 0.0;
			for (int k = 0; k < i; ++k) s -= L(i, k) * y[k];
			y[i] = s / L(i, i);
		}
		for (int i = n - 1; i >= 0; --i) {
			double s = y[i];
			for (int k = i + 1; k < n; ++k) s -= L(k, i) * X(k, c);
			X(i, c) = s / L(i, i);
		}
	}
	return X;
}

// MatrixXd::inverse() on a dynamic matrix = PartialPivLU; a singular input yields inf/NaN like Eigen, never throws
// (used at SingularityHandler.cpp:120,182,190,201,212; JointTask.cpp:260-265)
inline Mat lu_inverse(const Mat& A) {
	const int n = A.r;
	ORC_OPS(2LL * n * n * n / 3 + 2LL * n * n * n);
	Mat LU = A;
	int piv[8];
	for (int i = 0; i < n; ++i) piv[i] = i;
	for (int k = 0; k < n; ++k) {
		int p = k;
		double best = std::fabs(LU(k, k));
		for (int i = k + 1; i < n; ++i)
			if (std::fabs(LU(i, k)) > best) { best = std::fabs(LU(i, k)); p = i; }
		if (p != k) {
			for (int j = 0; j < n; ++j) std::swap(LU(k, j), LU(p, j));
			std::swap(piv[k], piv[p]);
		}
		for (int i = k + 1; i < n; ++i) {
			LU(i, k) /= LU(k, k);
			for (int j = k + 1; j < n; ++j) LU(i, j) -= LU(i, k) * LU(k, j);
		}
	}
	Mat X(n, n);
	for (int c = 0; c < n; ++c) {
		double y[8];
		for (int i = 0; i < n; ++i) {
			double s = (piv[i] == c) ? 1.0 : 0.0;
			for (int k = 0; k < i; ++k) s -= LU(i, k) * y[k];
			y[i] = s;
		}
		for (int i = n - 1; i >= 0; --i) {
			double s = y[i];
			for (int k = i + 1; k < n; ++k) s -= LU(i, k) * X(k, c);
			X(i, c) = s / LU(i, i);
		}
	}
	return X;
}

// JacobiSVD<MatrixXd>(A, ComputeThinU | ComputeThinV): SingularityHandler.cpp:78; inside matrixRangeBasis.
// One-sided (Hestenes) Jacobi; sigma sorted descending, sigma >= 0.  U/V signs are implementation-defined in Eigen; here
// (and in the kernels) the largest-|component| entry of each V column is made positive.
struct Svd {
	Mat U, V;
	double s[8];
	int k = 0;
};
inline Svd jacobi_svd(const Mat& A) {
	const bool tr = A.r < A.c;
	Mat W = tr ? A.T() : A;  // rows >= cols
	const int m = W.r, n = W.c;
	Mat R = Mat::Identity(n);
	// Eigen's JacobiSVD stops on an ABSOLUTE threshold tied to the largest diagonal entry (precision x maxDiagEntry): columns at
	// rounding-noise level are not rotated against each other.  Same here: columns below 1e-14 |A|_F are numerically null.
	double negl = 0;
	for (int i = 0; i < m; ++i)
		for (int j = 0; j < n; ++j) negl += W(i, j) * W(i, j);
	negl *= 1e-28;
	for (int sweep = 0; sweep < 60; ++sweep) {
		bool rotated = false;
		for (int p = 0; p < n - 1; ++p)
			for (int q = p + 1; q < n; ++q) {
				ORC_OPS(6 * m);
				double al = 0, be = 0, ga = 0;
				for (int i = 0; i < m; ++i) {
					al += W(i, p) * W(i, p);
					be += W(i, q) * W(i, q);
					ga += W(i, p) * W(i, q);
				}
				if (ga == 0.0 || al <= negl || be <= negl || std::fabs(ga) <= 1e-15 * std::sqrt(al * be)) continue;
				rotated = true;
				ORC_OPS(14 + 6 * m + 6 * n);
				double zeta = (be - al) / (2.0 * ga);
				double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
				double cs = 1.0 / std::sqrt(1.0 + t * t), sn = cs * t;
				for (int i = 0; i < m; ++i) {
					double wp = W(i, p), wq = W(i, q);
					W(i, p) = cs * wp - sn * wq;
					W(i, q) = sn * wp + cs * wq;
				}
				for (int i = 0; i < n; ++i) {
					double rp = R(i, p), rq = R(i, q);
					R(i, p) = cs * rp - sn * rq;
					R(i, q) = sn * rp + cs * rq;
				}
			}
		if (!rotated) break;
	}
	double sv[8];
	int order[8];
	for (int j = 0; j < n; ++j) {
		double s2 = 0;
		for (int i = 0; i < m; ++i) s2 += W(i, j) * W(i, j);
		sv[j] = std::sqrt(s2);
		order[j] = j;
	}
	std::stable_sort(order, order + n, [&](int a, int b) { return sv[a] > sv[b]; });
	Mat Q(m, n), Rs(n, n);
	Svd out;
	out.k = n;
	const double tiny = 1e-14 * sv[order[0]];
	for (int jj = 0; jj < n; ++jj) {
		int j = order[jj];
		out.s[jj] = sv[j];
		for (int i = 0; i < n; ++i) Rs(i, jj) = R(i, j);
		if (sv[j] > tiny && sv[j] > 0) {
			for (int i = 0; i < m; ++i) Q(i, jj) = W(i, j) / sv[j];
		} else {
			// numerically zero singular value: complete the basis (Eigen returns an orthonormal completion)
			for (int e = 0; e < m; ++e) {
				double v[8];
				for (int i = 0; i < m; ++i) v[i] = (i == e) ? 1.0 : 0.0;
				for (int pass = 0; pass < 2; ++pass)
					for (int c = 0; c < jj; ++c) {
						double d = 0;
						for (int i = 0; i < m; ++i) d += Q(i, c) * v[i];
						for (int i = 0; i < m; ++i) v[i] -= d * Q(i, c);
					}
				double nv = 0;
				for (int i = 0; i < m; ++i) nv += v[i] * v[i];
				nv = std::sqrt(nv);
				if (nv > 0.5) {
					for (int i = 0; i < m; ++i) Q(i, jj) = v[i] / nv;
					break;
				}
			}
		}
	}
	// W = Q S Rs^T.  Not transposed: A = W -> U = Q, V = Rs.  Transposed: A = W^T = Rs S Q^T -> U = Rs, V = Q.
	out.U = tr ? Rs : Q;
	out.V = tr ? Q : Rs;
	for (int j = 0; j < n; ++j) {
		int kmax = 0;
		for (int i = 1; i < out.V.r; ++i)
			if (std::fabs(out.V(i, j)) > std::fabs(out.V(kmax, j))) kmax = i;
		if (out.V(kmax, j) < 0) {
			for (int i = 0; i < out.V.r; ++i) out.V(i, j) = -out.V(i, j);
			for (int i = 0; i < out.U.r; ++i) out.U(i, j) = -out.U(i, j);
		}
	}
	return out;
}

// CompleteOrthogonalDecomposition::pseudoInverse stand-in (SingularityHandler.cpp:96-98; result is never consumed: the
// fully singular branch returns zero torques, .cpp:317-318).  Via the SVD with Eigen's default rank threshold.
inline Mat pseudo_inverse(const Mat& A) {
	Svd s = jacobi_svd(A);
	Mat out(A.c, A.r);
	const double thr = std::numeric_limits<double>::epsilon() * std::max(A.r, A.c) * s.s[0];
	for (int k = 0; k < s.k; ++k) {
		if (!(s.s[k] > thr)) continue;
		for (int i = 0; i < A.c; ++i)
			for (int j = 0; j < A.r; ++j) out(i, j) += s.V(i, k) * s.U(j, k) / s.s[k];
	}
	return out;
}

// Sai2Model::matrixRangeBasis (free function of sai2-model; called JointTask.cpp:233,
// MotionForceTask.cpp:66,78,147,149)
inline Mat matrixRangeBasis(const Mat& A, double tol = 1e-3) {
	const int max_range = std::min(A.r, A.c);
	if (A.norm() < tol) return Mat::Zero(A.r, 1);
	Svd svd = jacobi_svd(A);
	const double s0 = svd.s[0];
	if (s0 < tol) return Mat::Zero(A.r, 1);
	int task_dof = max_range;
	for (int i = svd.k - 1; i > 0; --i) {
		if (svd.s[i] / s0 < tol) task_dof -= 1;
		else break;
	}
	if (task_dof == A.r) return Mat::Identity(A.r);
	return svd.U.leftCols(task_dof);
}

// Sai2Model::computePseudoInverse on the diagonal gain matrices it is called with (MotionForceTask.cpp:417,450;
// JointTask.cpp:328): COD with setThreshold(1e-6) relative to the largest pivot.
inline void pseudoInverseDiag(const double* d, int n, double* out, double tol = 1e-6) {
	double mx = 0;
	for (int i = 0; i < n; ++i) mx = std::max(mx, std::fabs(d[i]));
	for (int i = 0; i < n; ++i) out[i] = (std::fabs(d[i]) > tol * mx) ? 1.0 / d[i] : 0.0;
}

// Sai2Model::orientationError (free function; called MotionForceTask.cpp:292,443, SingularityHandler.cpp:260)
inline V3 orientationError(const M3& Rd, const M3& R) {
	V3 e;
	for (int i = 0; i < 3; ++i) e = e + R.col(i).cross(Rd.col(i));
	return -0.5 * e;
}

// ------------------------------------------------------------------------------------------------------------------
// Robot model: stands in for Sai2Model::Sai2Model over RBDL
// ------------------------------------------------------------------------------------------------------------------
struct Model {
	osc_chain_desc ch;
	int n;
	double q[8], dq[8];
	M3 R[8];   // base-frame orientation of body i
	V3 o[8];   // base-frame origin of body i
	Mat Mm, Minv;

	explicit Model(const osc_chain_desc& c) : ch(c), n(c.n) {
		for (int i = 0; i < 8; ++i) q[i] = dq[i] = 0;
	}
	int dof() const { return n; }

	static M3 axisRot(const V3& a, double ang) {
		const double c = std::cos(ang), s = std::sin(ang), v = 1 - c;
		M3 r;
		r.m[0][0] = c + a.x * a.x * v;       r.m[0][1] = a.x * a.y * v - a.z * s; r.m[0][2] = a.x * a.z * v + a.y * s;
		r.m[1][0] = a.y * a.x * v + a.z * s; r.m[1][1] = c + a.y * a.y * v;       r.m[1][2] = a.y * a.z * v - a.x * s;
		r.m[2][0] = a.z * a.x * v - a.y * s; r.m[2][1] = a.z * a.y * v + a.x * s; r.m[2][2] = c + a.z * a.z * v;
		return r;
	}
	// relative transform parent body -> body i for the current q: (E = child axes in parent coords, r = origin in parent)
	void jointTransform(int i, M3& E, V3& r) const {
		const V3 ax(ch.joint_axis[i]);
		E = M3(ch.joint_rot[i]);
		r = V3(ch.joint_xyz[i]);
		if (ch.joint_type[i] == OSC_JOINT_REVOLUTE) E = E * axisRot(ax, q[i]);
		else r = r + E * (q[i] * ax);
	}
	// Sai2Model::updateKinematics (RBDL UpdateKinematicsCustom): base-frame poses
	void updateKinematics() {
		M3 Rp = M3::Identity();
		V3 pp;
		for (int i = 0; i < n; ++i) {
			M3 E; V3 r;
			jointTransform(i, E, r);
			o[i] = pp + Rp * r;
			R[i] = Rp * E;
			Rp = R[i];
			pp = o[i];
		}
	}
	// Sai2Model::updateModel: kinematics + M by the composite-rigid-body algorithm in link coordinates (the structure of
	// RBDL's CompositeRigidBodyAlgorithm: composite inertias accumulated tip->base, one spatial force walked up per column)
	// + dense inverse (_M.inverse()).  examples/05...cpp:145
	void updateModel() {
		updateKinematics();
		struct RB { double m; V3 c; M3 I; };  // mass, com, inertia about the com, in the body's own frame
		RB Ic[8];
		M3 E[8]; V3 r[8];
		for (int i = 0; i < n; ++i) {
			Ic[i].m = ch.mass[i]; Ic[i].c = V3(ch.com[i]); Ic[i].I = M3(ch.inertia[i]);
			jointTransform(i, E[i], r[i]);
		}
		Mm = Mat::Zero(n, n);
		for (int i = n - 1; i >= 0; --i) {
			// spatial force of the composite body i under unit joint velocity, at origin i, frame i
			const V3 ax(ch.joint_axis[i]);
			V3 w, v;
			if (ch.joint_type[i] == OSC_JOINT_REVOLUTE) w = ax; else v = ax;
			const RB& B = Ic[i];
			V3 vc = v + w.cross(B.c);                        // com velocity
			V3 f = B.m * vc;                                 // linear momentum
			V3 nn = B.I * w + B.c.cross(f);                  // angular momentum about origin i
			Mm(i, i) = (ch.joint_type[i] == OSC_JOINT_REVOLUTE) ? ax.dot(nn) : ax.dot(f);
			int j = i;
			while (j > 0) {
				// express the force in the parent frame, about the parent origin
				V3 fp = E[j] * f;
				V3 np = E[j] * nn + r[j].cross(fp);
				f = fp; nn = np;
				j = j - 1;
				const V3 aj(ch.joint_axis[j]);
				double h = (ch.joint_type[j] == OSC_JOINT_REVOLUTE) ? aj.dot(nn) : aj.dot(f);
				Mm(i, j) = h; Mm(j, i) = h;
			}
			// add composite i to its parent
			if (i > 0) {
				RB& P = Ic[i - 1];
				RB C;  // child expressed in the parent frame
				C.m = B.m; C.c = r[i] + E[i] * B.c; C.I = E[i] * B.I * E[i].T();
				double mt = P.m + C.m;
				if (mt > 0) {
					V3 ct = (1.0 / mt) * (P.m * P.c + C.m * C.c);
					auto shift = [](const RB& b, const V3& to) {
						V3 d = b.c - to;
						return b.I + b.m * (d.dot(d) * M3::Identity() - M3::Outer(d, d));
					};
					M3 It = shift(P, ct) + shift(C, ct);
					P.m = mt; P.c = ct; P.I = It;
				}
			}
		}
		Minv = lu_inverse(Mm);
	}
	const Mat& M() const { return Mm; }
	const Mat& MInv() const { return Minv; }

	// position / rotation of a frame fixed in `body`, robot-base frame (SingularityHandler.cpp:250-251)
	V3 position(int body, const V3& p) const { return o[body] + R[body] * p; }
	M3 rotation(int body, const M3& Rl) const { return R[body] * Rl; }
	V3 positionInWorld(int body, const V3& p) const { return V3(ch.world_xyz) + M3(ch.world_rot) * position(body, p); }
	M3 rotationInWorld(int body, const M3& Rl) const { return M3(ch.world_rot) * rotation(body, Rl); }
	// Sai2Model::JWorldFrame: 6 x n, linear rows first (MotionForceTask.cpp:262,293-298)
	Mat JWorldFrame(int body, const V3& p) const {
		Mat J(6, n);
		const V3 pb = position(body, p);
		const M3 Rw(ch.world_rot);
		for (int i = 0; i <= body; ++i) {
			V3 z = R[i] * V3(ch.joint_axis[i]);
			V3 lin, ang;
			if (ch.joint_type[i] == OSC_JOINT_REVOLUTE) { lin = z.cross(pb - o[i]); ang = z; }
			else lin = z;
			lin = Rw * lin; ang = Rw * ang;
			for (int k = 0; k < 3; ++k) { J(k, i) = lin[k]; J(3 + k, i) = ang[k]; }
		}
		return J;
	}
	struct OpSpaceMatrices { Mat Lambda, Jbar, N; };
	// Sai2Model::operationalSpaceMatrices (called JointTask.cpp:241-243,
	// SingularityHandler.cpp:110-111,130-131,154-155)
	OpSpaceMatrices operationalSpaceMatrices(const Mat& J) const {
		OpSpaceMatrices m;
		Mat inv_inertia = J * Minv * J.T();
		m.Lambda = llt_solve_identity(inv_inertia);
		m.Jbar = Minv * J.T() * m.Lambda;
		m.N = Mat::Identity(n) - m.Jbar * J;
		return m;
	}
	// Sai2Model::coriolisPlusGravity: h = b(q, dq) + g(q) by the recursive Newton-Euler algorithm in LINK coordinates (the
	// structure of RBDL's InverseDynamics with qdd = 0: velocities / accelerations base -> tip in each body's own frame,
	// forces tip -> base).  Row f-2 of SURVEY.md 8(f); restatement (sai2-model / RBDL are un-vendored), pinned by Lagrange's
	// equations at 50 digits (tests/test_pins.py::test_pin_rollout_lagrange_mpmath).
	void nonlinearEffects(double* h) const {
		M3 E[8]; V3 r[8];
		V3 w[8], wd[8], a[8], F[8], Nn[8];
		const V3 gb = M3(ch.world_rot).T() * V3(ch.gravity);
		V3 wp, wdp, ap = -1.0 * gb;  // base: at rest, accelerating against gravity
		for (int i = 0; i < n; ++i) {
			jointTransform(i, E[i], r[i]);
			const V3 ax(ch.joint_axis[i]);
			const M3 Et = E[i].T();
			// acceleration of origin i as a point of the parent, in parent coordinates, then in frame i
			V3 ao = ap + wdp.cross(r[i]) + wp.cross(wp.cross(r[i]));
			const V3 wpi = Et * wp;
			if (ch.joint_type[i] == OSC_JOINT_REVOLUTE) {
				w[i] = wpi + dq[i] * ax;
				wd[i] = Et * wdp + wpi.cross(dq[i] * ax);
				a[i] = Et * ao;
			} else {
				w[i] = wpi;
				wd[i] = Et * wdp;
				a[i] = Et * ao + 2.0 * wpi.cross(dq[i] * ax);
			}
			const V3 c(ch.com[i]);
			const M3 I(ch.inertia[i]);
			const V3 ac = a[i] + wd[i].cross(c) + w[i].cross(w[i].cross(c));
			F[i] = ch.mass[i] * ac;
			Nn[i] = I * wd[i] + w[i].cross(I * w[i]) + c.cross(F[i]);  // moment about origin i
			wp = w[i]; wdp = wd[i]; ap = a[i];
		}
		V3 f, nn;  // wrench of the subtree, frame i, about origin i
		for (int i = n - 1; i >= 0; --i) {
			if (i < n - 1) {
				const V3 fc = E[i + 1] * f;
				nn = E[i + 1] * nn + r[i + 1].cross(fc);
				f = fc;
			}
			f = f + F[i];
			nn = nn + Nn[i];
			const V3 ax(ch.joint_axis[i]);
			h[i] = (ch.joint_type[i] == OSC_JOINT_REVOLUTE) ? ax.dot(nn) : ax.dot(f);
		}
	}
	// the simulation side of examples/05...cpp:219-233 without contact: qdd = M^-1 (tau - h - d dq), symplectic Euler
	void rolloutStep(const double* tau, double dt, int substeps, double damping, double* h_out) {
		for (int s = 0; s < substeps; ++s) {
			updateModel();
			double h[8];
			nonlinearEffects(h);
			if (s == 0 && h_out) for (int i = 0; i < n; ++i) h_out[i] = h[i];
			double qdd[8];
			for (int i = 0; i < n; ++i) {
				qdd[i] = 0;
				for (int j = 0; j < n; ++j) qdd[i] += Minv(i, j) * (tau[j] - h[j] - damping * dq[j]);
			}
			for (int i = 0; i < n; ++i) {
				dq[i] += qdd[i] * dt;
				q[i] += dq[i] * dt;
			}
		}
	}
	// Sai2Model::jointGravityVector (RobotController.cpp:71)
	void jointGravityVector(double* g) const {
		const V3 gw = M3(ch.world_rot).T() * V3(ch.gravity);
		for (int i = 0; i < n; ++i) g[i] = 0;
		for (int b = 0; b < n; ++b) {
			V3 c = position(b, V3(ch.com[b]));
			for (int i = 0; i <= b; ++i) {
				V3 z = R[i] * V3(ch.joint_axis[i]);
				V3 lin = (ch.joint_type[i] == OSC_JOINT_REVOLUTE) ? z.cross(c - o[i]) : z;
				g[i] -= ch.mass[b] * lin.dot(gw);
			}
		}
	}
};

// ------------------------------------------------------------------------------------------------------------------
// POPCExplicitForceControl  (helper_modules/POPCExplicitForceControl.cpp)
// ------------------------------------------------------------------------------------------------------------------
struct POPC {
	double dt;
	bool enabled = false;                      // .cpp:6
	static constexpr int WINDOW = 250, MAX_COUNTER = 50;  // POPCExplicitForceControl.h:37-38
	double PO, E_corr, E_stored, Rc, vcl_sq_sum;
	int counter;
	std::deque<double> window;
	explicit POPC(double dt_) : dt(dt_) { reInitialize(); }
	void reInitialize() {  // .cpp:10-21
		PO = 0; E_corr = 0; E_stored = 0;
		window.clear();
		counter = MAX_COUNTER;
		Rc = 1.0; vcl_sq_sum = 0;
	}
	void enable() { enabled = true; }
	void disable() { enabled = false; reInitialize(); }  // .cpp:25-28
	// .cpp:30-96; kv_force is the (isotropic) diagonal
	V3 computePassivitySaturatedForce(const V3& fd, const V3& fs, const V3& vcl, const V3& vr, const V3& kv_force, double kff) {
		if (!enabled) return vcl - cwise(kv_force, vr);
		V3 F_cmd = kff * fd + Rc * vcl - cwise(kv_force, vr);
		double vc_squared = vcl.dot(vcl);
		V3 f_diff = fs - fd;
		double p = (f_diff.dot(vcl) - F_cmd.dot(vr)) * dt;
		PO += p;
		window.push_back(p);
		if (PO + E_stored + E_corr > 0) {
			while ((int)window.size() > WINDOW) {
				if (PO + E_corr + E_stored > window.front()) {
					if (window.front() > 0) PO -= window.front();
					window.pop_front();
				} else break;
			}
		}
		if (counter <= 0) {
			counter = MAX_COUNTER;
			double old_Rc = Rc;
			if (PO + E_stored + E_corr < 0) {
				Rc = 1 + (PO + E_stored + E_corr) / (vcl_sq_sum * dt);
				if (Rc > 1) Rc = 1;
				if (Rc < 0) Rc = 0;
			} else {
				Rc = (1 + (0.1 * MAX_COUNTER - 1) * Rc) / (double)(0.1 * MAX_COUNTER);
			}
			E_corr += (1 - old_Rc) * vcl_sq_sum * dt;
			vcl_sq_sum = 0;
		}
		counter--;
		vcl_sq_sum += vc_squared;
		return Rc * vcl - cwise(kv_force, vr);
	}
};

// ------------------------------------------------------------------------------------------------------------------
// SingularityHandler  (tasks/SingularityHandler.cpp)
// ------------------------------------------------------------------------------------------------------------------
struct SingularityHandler {
	Model* robot;
	int body; V3 p_c; M3 R_c;
	int task_rank, dof;
	const osc_mft_params* prm;  // gains / thresholds live in the task's parameter block
	double q_upper[8], q_lower[8], tau_upper[8], tau_lower[8];
	std::vector<int> singularity_types;
	std::deque<int> history;
	int c1 = 0, c2 = 0;
	double q_prior[8], dq_prior[8], type_2_direction[8];
	Svd svd;
	double alpha = 1;
	Mat N, U_ns, U_s, V_s, J_ns, J_s, L_ns, N_ns, L_s, L_ns_mod, L_s_mod, L_js, L_js_mod, J_post;
	bool nan_seen = false;

	SingularityHandler(Model* r, int body_, const V3& p, const M3& Rc, int rank, const osc_mft_params* prm_)
		: robot(r), body(body_), p_c(p), R_c(Rc), task_rank(rank), dof(r->dof()), prm(prm_) {  // .cpp:24-73
		for (int i = 0; i < dof; ++i) {
			q_upper[i] = r->ch.q_upper[i]; q_lower[i] = r->ch.q_lower[i];
			tau_upper[i] = r->ch.effort[i]; tau_lower[i] = -r->ch.effort[i];
			q_prior[i] = 0.5 * (q_lower[i] + q_upper[i]);  // _joint_midrange, .cpp:47,55
			dq_prior[i] = 0;
			type_2_direction[i] = 1;
		}
		N = Mat::Identity(dof); N_ns = Mat::Identity(dof);
		U_ns = Mat::Zero(6, 1); U_s = Mat::Zero(6, 1); V_s = Mat::Zero(dof, 1);
		J_ns = Mat::Zero(1, dof); J_s = Mat::Zero(1, dof); J_post = Mat::Zero(1, dof);
		L_ns = L_ns_mod = L_s = L_s_mod = L_js = L_js_mod = Mat::Zero(1, 1);
		svd.k = 0;
	}

	void updateTaskModel(const Mat& Jp, const Mat& N_prec) {  // .cpp:75-228
		const int m = task_rank;
		svd = jacobi_svd(Jp);
		bool fully_singular = false;
		if (svd.s[0] < prm->s_abs_tol) {  // .cpp:83-98
			fully_singular = true;
			alpha = 0;
			U_ns = Mat::Zero(m, 1); J_ns = Mat::Zero(m, dof); L_ns = Mat::Zero(m, m);
			U_s = svd.U.leftCols(m); V_s = svd.V.leftCols(m);
			J_s = U_s.T() * Jp;
			L_s = pseudo_inverse(J_s * robot->MInv() * J_s.T());
		} else {
			auto non_singular = [&]() {  // .cpp:123-141
				alpha = 1;
				U_ns = svd.U.leftCols(m);
				J_ns = U_ns.T() * Jp;
				auto ns = robot->operationalSpaceMatrices(J_ns);
				L_ns = ns.Lambda; N_ns = ns.N;
				U_s = Mat::Zero(m, 1); V_s = Mat::Zero(dof, 1); J_s = Mat::Zero(m, dof); L_s = Mat::Zero(m, m);
			};
			if (m == 1) non_singular();  // Q4 (defined behaviour; the reference's loop body never runs for m = 1)
			for (int i = 1; i < m; ++i) {
				double inv_condition_number = svd.s[i] / svd.s[0];
				if (inv_condition_number < prm->s_max) {  // .cpp:103-121
					alpha = std::min(std::max((inv_condition_number - prm->s_min) / (prm->s_max - prm->s_min), 0.), 1.);
					U_ns = svd.U.leftCols(i);
					J_ns = U_ns.T() * Jp;
					auto ns = robot->operationalSpaceMatrices(J_ns);
					L_ns = ns.Lambda; N_ns = ns.N;
					U_s = svd.U.block(0, i, svd.U.r, m - i);
					V_s = svd.V.block(0, i, svd.V.r, m - i);
					J_s = U_s.T() * Jp;
					L_s = lu_inverse(J_s * robot->MInv() * J_s.T());
					break;
				} else if (i == m - 1) {
					non_singular();
				}
			}
		}
		// .cpp:145-158 (Q11: the fully singular task is passed through whether or not handling is enabled)
		const bool has_s = U_s.norm() != 0, has_ns = U_ns.norm() != 0;
		bool posture = false;
		if (fully_singular) {
			N = N_prec; L_js = Mat::Zero(1, 1);
		} else if (!has_s || !prm->singularity_handling_enabled) {
			N = N_ns; L_js = Mat::Zero(1, 1);
		} else {
			J_post = V_s.T() * N_ns * N_prec;
			auto op = robot->operationalSpaceMatrices(J_post);
			L_js = op.Lambda;
			N = op.N * N_ns;
			posture = true;
		}
		switch (prm->dynamic_decoupling_type) {  // .cpp:160-225
			case OSC_FULL_DYNAMIC_DECOUPLING:
				L_ns_mod = L_ns; L_s_mod = L_s; L_js_mod = L_js;
				break;
			case OSC_IMPEDANCE:
				L_ns_mod = Mat::Identity(U_ns.c); L_s_mod = Mat::Identity(U_s.c); L_js_mod = Mat::Identity(V_s.c);
				break;
			default: {  // BOUNDED_INERTIA_ESTIMATES
				Mat M_BIE = robot->M();
				for (int i = 0; i < dof; ++i)
					if (M_BIE(i, i) < prm->bie_threshold) M_BIE(i, i) = prm->bie_threshold;
				Mat M_inv_BIE = lu_inverse(M_BIE);
				L_ns_mod = has_ns ? lu_inverse(J_ns * M_inv_BIE * J_ns.T()) : L_ns;
				L_s_mod = has_s ? lu_inverse(J_s * M_inv_BIE * J_s.T()) : L_s;
				// Q6: only when the posture Jacobian was computed this tick
				L_js_mod = (has_s && posture) ? lu_inverse(J_post * M_inv_BIE * J_post.T()) : L_js;
			}
		}
		classifySingularity();
	}

	void classifySingularity() {  // .cpp:230-295
		if (singularity_types.size() == 0 || c2 > c1) {
			for (int i = 0; i < dof; ++i) { q_prior[i] = robot->q[i]; dq_prior[i] = robot->dq[i]; }
		}
		if (U_s.norm() == 0) {
			singularity_types.clear(); history.clear(); c1 = c2 = 0;
			return;
		}
		singularity_types.assign(U_s.c, OSC_NO_SINGULARITY);
		double curr_q[8];
		for (int i = 0; i < dof; ++i) curr_q[i] = robot->q[i];
		V3 curr_pos = robot->position(body, p_c);
		M3 curr_ori = robot->rotation(body, R_c);
		bool any_type_1 = false;
		for (int c = 0; c < U_s.c; ++c) {
			for (int i = 0; i < dof; ++i) robot->q[i] = curr_q[i] + prm->perturb_step_size * V_s(i, c);
			robot->updateKinematics();
			V3 pos_delta = robot->position(body, p_c) - curr_pos;
			V3 ori_delta = orientationError(robot->rotation(body, R_c), curr_ori);
			double motion = 0;
			for (int k = 0; k < 3; ++k) motion += pos_delta[k] * U_s(k, c) + ori_delta[k] * U_s(3 + k, c);
			motion = std::fabs(motion);
			singularity_types[c] = motion > prm->type_1_tol ? OSC_TYPE_1_SINGULARITY : OSC_TYPE_2_SINGULARITY;
			any_type_1 = any_type_1 || singularity_types[c] == OSC_TYPE_1_SINGULARITY;
			for (int i = 0; i < dof; ++i) robot->q[i] = curr_q[i];
			robot->updateKinematics();
		}
		if (any_type_1) { history.push_back(OSC_TYPE_1_SINGULARITY); c1++; }
		else { history.push_back(OSC_TYPE_2_SINGULARITY); c2++; }
		if ((int)history.size() > prm->buffer_size) {
			if (history.front() == OSC_TYPE_1_SINGULARITY) c1--;
			else if (history.front() == OSC_TYPE_2_SINGULARITY) c2--;
			history.pop_front();
		}
	}

	Mat computeTorques(const Mat& F_star, const Mat& F_f) {  // .cpp:297-368
		if (singularity_types.size() == 0) {
			return J_ns.T() * (L_ns_mod * U_ns.T() * F_star + U_ns.T() * F_f);
		} else if (prm->dynamic_decoupling_type == OSC_IMPEDANCE) {
			return J_ns.T() * (U_ns.T() * F_star + U_ns.T() * F_f);
		}
		Mat tau_ns = Mat::Zero(dof, 1);
		if (U_ns.norm() == 0) return tau_ns;
		tau_ns = J_ns.T() * (L_ns_mod * U_ns.T() * F_star + U_ns.T() * F_f);
		if (!prm->singularity_handling_enabled) return tau_ns;
		Mat joint_strategy;
		if (c1 > c2 || prm->enforce_type_1_strategy) {
			Mat unit(dof, 1);
			for (int i = 0; i < dof; ++i) unit.a[i] = -prm->kp_type_1 * (robot->q[i] - q_prior[i]) - prm->kv_type_1 * robot->dq[i];
			joint_strategy = J_post.T() * L_js_mod * V_s.T() * unit;
		} else {
			for (int i = 0; i < V_s.r; ++i) {
				if (V_s(i, 0) != 0) {
					if (std::fabs(robot->q[i] - q_upper[i]) < prm->type_2_angle_threshold) type_2_direction[i] = -1;
					else if (std::fabs(robot->q[i] - q_lower[i]) < prm->type_2_angle_threshold) type_2_direction[i] = 1;
				}
			}
			Mat f = F_star + F_f;
			double nf = f.norm();
			double fTd = 0;
			for (int k = 0; k < 6; ++k) fTd += (nf > 0 ? f.a[k] / nf : f.a[k]) * U_s(k, 0);
			Mat unit(dof, 1), damp(dof, 1);
			for (int i = 0; i < dof; ++i) {
				// Q5: the intended _type_2_torque_ratio * effort (the reference reads the ratio before initialising it)
				unit.a[i] = type_2_direction[i] * (std::fabs(fTd) * prm->type_2_torque_ratio * robot->ch.effort[i]);
				damp.a[i] = -prm->kv_type_2 * robot->dq[i];
			}
			joint_strategy = J_post.T() * V_s.T() * unit + J_post.T() * L_js_mod * V_s.T() * damp;
		}
		Mat tau_s = J_s.T() * (L_s_mod * U_s.T() * F_star + U_s.T() * F_f);
		for (int i = 0; i < dof; ++i) {
			if (std::isnan(tau_s.a[i])) { tau_s.a[i] = 0; nan_seen = true; }
			else if (tau_s.a[i] > tau_upper[i]) tau_s.a[i] = tau_upper[i];
			else if (tau_s.a[i] < tau_lower[i]) tau_s.a[i] = tau_lower[i];
		}
		return tau_ns + alpha * tau_s + (1 - alpha) * joint_strategy;
	}
};

// ------------------------------------------------------------------------------------------------------------------
// Tasks
// ------------------------------------------------------------------------------------------------------------------
struct Task {
	Model* robot;
	int type;
	double dt;
	Mat N_prec, N, tau;  // tau = last computeTorques result
	virtual ~Task() {}
	virtual void updateTaskModel(const Mat& N_prec) = 0;
	virtual Mat computeTorques() = 0;
	virtual Mat computeTorques(const Mat& tau_prec) = 0;
	virtual void reInitializeTask() = 0;
	Mat getTaskAndPreviousNullspace() const { return N * N_prec; }
};

struct MotionForceTask : Task {
	osc_mft_desc desc;
	osc_mft_params prm;
	int body; V3 p_c; M3 R_c;
	Mat P;  // _partial_task_projection
	int pos_range, ori_range;
	POPC popc;
	SingularityHandler* handler;
	Mat J, Jp;
	V3 x, v, w, ori_err;
	M3 R;
	V3 goal_pos, goal_linvel, goal_angvel, goal_linacc, goal_angacc, goal_force, goal_moment;
	M3 goal_ori;
	V3 des_pos, des_linvel, des_angvel, des_linacc, des_angacc;
	M3 des_ori;
	V3 sensed_force, sensed_moment;
	V3 integ_pos, integ_ori, integ_force, integ_moment;
	double unit_mass_force[6];
	// internal OTG (MotionForceTask.h:812-815).  The reference enables it by default (MotionForceTask.h:67, .cpp:173-188);
	// here it starts disabled (every BASELINE config runs with OTG off) and enableInternalOtgAccelerationLimited turns it on.
	bool use_internal_otg = false;
	otg::OtgCartesian* otg_ = nullptr;

	MotionForceTask(Model* r, const osc_mft_desc& d) : desc(d), body(d.body), p_c(d.frame_xyz), R_c(d.frame_rot), popc(d.loop_timestep) {
		robot = r; type = OSC_TASK_MOTION_FORCE; dt = d.loop_timestep;
		P = Mat::Identity(6);  // MotionForceTask.cpp:28
		if (d.partial) {       // MotionForceTask.cpp:47-87
			if (d.n_dirs_translation == 0 && d.n_dirs_rotation == 0)
				throw std::invalid_argument("controlled_directions_translation and controlled_directions_rotation cannot both be empty in MotionForceTask::MotionForceTask\n");
			Mat Bt = Mat::Zero(3, 1), Br = Mat::Zero(3, 1);
			if (d.n_dirs_translation) {
				Mat Vt(3, d.n_dirs_translation);
				for (int c = 0; c < d.n_dirs_translation; ++c)
					for (int k = 0; k < 3; ++k) Vt(k, c) = d.dirs_translation[c][k];
				Bt = matrixRangeBasis(Vt);
			}
			if (d.n_dirs_rotation) {
				Mat Vr(3, d.n_dirs_rotation);
				for (int c = 0; c < d.n_dirs_rotation; ++c)
					for (int k = 0; k < 3; ++k) Vr(k, c) = d.dirs_rotation[c][k];
				Br = matrixRangeBasis(Vr);
			}
			P = Mat::Zero(6, 6);
			P.setBlock(0, 0, Bt * Bt.T());
			P.setBlock(3, 3, Br * Br.T());
		}
		initialSetup();
	}
	~MotionForceTask() override { delete handler; delete otg_; }

	static void default_params(osc_mft_params* p) {  // MotionForceTask.h:40-75; SingularityHandler.cpp:10-20; MotionForceTask.cpp:197
		std::memset(p, 0, sizeof *p);
		for (int i = 0; i < 3; ++i) {
			p->kp_pos[i] = 100; p->kv_pos[i] = 20; p->ki_pos[i] = 0;
			p->kp_ori[i] = 200; p->kv_ori[i] = 28.3; p->ki_ori[i] = 0;
		}
		p->kp_force = 0.7; p->kv_force = 10; p->ki_force = 1.3;
		p->kp_moment = 0.7; p->kv_moment = 10; p->ki_moment = 1.3;
		p->kff_force = p->kff_moment = 0.95;
		p->max_force_control_feedback_output = 20; p->max_moment_control_feedback_output = 10;
		p->linear_saturation_velocity = 0.3; p->angular_saturation_velocity = M_PI / 3;
		p->dynamic_decoupling_type = OSC_BOUNDED_INERTIA_ESTIMATES; p->bie_threshold = 0.1;
		p->singularity_handling_enabled = 1;
		p->s_min = 6e-3; p->s_max = 6e-2;
		p->kp_type_1 = 50; p->kv_type_1 = 14; p->kv_type_2 = 5;
		p->s_abs_tol = 1e-3; p->type_1_tol = 0.5; p->type_2_torque_ratio = 1e-2;
		p->type_2_angle_threshold = 5 * M_PI / 180; p->perturb_step_size = 5; p->buffer_size = 200;
		p->sensor_rot[0] = p->sensor_rot[4] = p->sensor_rot[8] = 1;
	}

	void initialSetup() {  // MotionForceTask.cpp:92-202
		const int dof = robot->dof();
		default_params(&prm);
		J = Mat::Zero(6, dof); Jp = Mat::Zero(6, dof);
		N = Mat::Zero(dof, dof); N_prec = Mat::Identity(dof);
		tau = Mat::Zero(dof, 1);
		Mat range_pos = matrixRangeBasis(P.block(0, 0, 3, 3));
		Mat range_ori = matrixRangeBasis(P.block(3, 3, 3, 3));
		pos_range = range_pos.norm() == 0 ? 0 : range_pos.c;
		ori_range = range_ori.norm() == 0 ? 0 : range_ori.c;
		if (pos_range + ori_range == 0)
			throw std::invalid_argument("controlled_directions_translation and controlled_directions_rotation cannot both be empty in MotionForceTask::MotionForceTask\n");
		handler = new SingularityHandler(robot, body, p_c, R_c, pos_range + ori_range, &prm);
		reInitializeTask();
	}

	void reInitializeTask() override {  // MotionForceTask.cpp:204-245
		x = robot->positionInWorld(body, p_c);
		R = robot->rotationInWorld(body, R_c);
		goal_pos = des_pos = x;
		goal_ori = des_ori = R;
		v = w = goal_linvel = goal_angvel = des_linvel = des_angvel = V3();
		goal_linacc = goal_angacc = des_linacc = des_angacc = V3();
		ori_err = integ_pos = integ_ori = V3();
		goal_force = goal_moment = sensed_force = sensed_moment = V3();
		integ_force = integ_moment = V3();
		for (double& u : unit_mass_force) u = 0;
		if (otg_) otgReInitialize();  // .cpp:244
	}
	void otgReInitialize() {
		double p0[3] = {x.x, x.y, x.z}, R0[9];
		for (int i = 0; i < 9; ++i) R0[i] = R.m[i / 3][i % 3];
		otg_->reInitialize(p0, R0);
	}
	// MotionForceTask.cpp:511-523
	void enableInternalOtgAccelerationLimited(double max_lin_v, double max_lin_a, double max_ang_v, double max_ang_a) {
		if (max_lin_v <= 0 || max_lin_a <= 0 || max_ang_v <= 0 || max_ang_a <= 0)  // OTG_6dof_cartesian.cpp:88-129
			throw std::invalid_argument("max velocity / acceleration set to 0 or negative value in OTG_6dof_cartesian\n");
		if (!otg_) {  // constructor, .cpp:171-172
			double p0[3] = {x.x, x.y, x.z}, R0[9];
			for (int i = 0; i < 9; ++i) R0[i] = R.m[i / 3][i % 3];
			otg_ = new otg::OtgCartesian(p0, R0, dt);
		}
		if (!use_internal_otg) otgReInitialize();
		otg_->setLimits(max_lin_v, max_lin_a, max_ang_v, max_ang_a);
		use_internal_otg = true;
	}
	void disableInternalOtg() { use_internal_otg = false; }  // MotionForceTask.h:423

	// applies the side effects of parametrizeForceMotionSpaces (.cpp:830-858), parametrizeMomentRotMotionSpaces
	// (.cpp:860-890), setClosedLoopForce/MomentControl (.cpp:973-986), enable/disablePassivity
	int setParams(const osc_mft_params& np) {
		int resets = 0;
		auto axis_changed = [](const double* a, const double* b) {
			// !normalized().isApprox(old): Eigen isApprox, precision 1e-12 (relative, squared norms)
			double d = 0, na = 0, nb = 0;
			for (int i = 0; i < 3; ++i) { d += (a[i] - b[i]) * (a[i] - b[i]); na += a[i] * a[i]; nb += b[i] * b[i]; }
			return !(d <= 1e-24 * std::min(na, nb));
		};
		osc_mft_params p = np;
		auto normalize = [](double* a) {
			double nrm = std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
			if (nrm > 0) for (int i = 0; i < 3; ++i) a[i] /= nrm;
		};
		bool reset_lin = p.force_space_dimension != prm.force_space_dimension;
		if (p.force_space_dimension == 1 || p.force_space_dimension == 2) {
			normalize(p.force_or_motion_axis);
			reset_lin = reset_lin || axis_changed(p.force_or_motion_axis, prm.force_or_motion_axis);
		} else std::memcpy(p.force_or_motion_axis, prm.force_or_motion_axis, sizeof p.force_or_motion_axis);
		bool reset_ang = p.moment_space_dimension != prm.moment_space_dimension;
		if (p.moment_space_dimension == 1 || p.moment_space_dimension == 2) {
			normalize(p.moment_or_rotmotion_axis);
			reset_ang = reset_ang || axis_changed(p.moment_or_rotmotion_axis, prm.moment_or_rotmotion_axis);
		} else std::memcpy(p.moment_or_rotmotion_axis, prm.moment_or_rotmotion_axis, sizeof p.moment_or_rotmotion_axis);
		if (reset_lin) {
			goal_pos = x; goal_linvel = V3(); goal_linacc = V3();
			if (otg_) { double p0[3] = {x.x, x.y, x.z}; otg_->reInitializeLinear(p0); }  // .cpp:854
			integ_pos = V3(); integ_force = V3();
			resets |= 1;
		}
		if (reset_ang) {
			goal_ori = R; goal_angvel = V3(); goal_angacc = V3();
			if (otg_) { double R0[9]; for (int i = 0; i < 9; ++i) R0[i] = R.m[i / 3][i % 3]; otg_->reInitializeAngular(R0); }  // .cpp:886
			integ_ori = V3(); integ_moment = V3();
			resets |= 2;
		}
		if ((p.closed_loop_force_control != 0) != (prm.closed_loop_force_control != 0)) { integ_pos = V3(); integ_force = V3(); }
		if ((p.closed_loop_moment_control != 0) != (prm.closed_loop_moment_control != 0)) { integ_ori = V3(); integ_moment = V3(); }
		if (p.passivity_enabled && !prm.passivity_enabled) popc.enable();
		if (!p.passivity_enabled && prm.passivity_enabled) { popc.disable(); resets |= 4; }
		prm = p;
		return resets;
	}

	// MotionForceTask.cpp:805-828
	void updateSensedForceAndMoment(const V3& f_s, const V3& m_s) {
		M3 Rcs(prm.sensor_rot);
		V3 pcs(prm.sensor_xyz);
		M3 Rw = robot->rotationInWorld(body, R_c);
		V3 f = Rcs * f_s;
		V3 m = pcs.cross(f) + Rcs * m_s;
		sensed_force = Rw * f;
		sensed_moment = Rw * m;
	}

	M3 Pblock(int o) const {
		M3 r;
		for (int i = 0; i < 3; ++i)
			for (int j = 0; j < 3; ++j) r.m[i][j] = P(o + i, o + j);
		return r;
	}
	M3 rotParam() const { return desc.parametrization_in_compliant_frame ? robot->rotationInWorld(body, R_c) : M3::Identity(); }
	M3 sigma(int dim, const M3& Psel, const double* axis) const {  // .cpp:892-925 / 932-966
		if (dim == 0) return M3();
		if (dim == 3) return Psel;
		V3 a = rotParam() * V3(axis);
		M3 aa = M3::Outer(a, a);
		if (dim == 1) return Psel * aa * Psel.T();
		return Psel * (M3::Identity() - aa) * Psel.T();
	}
	M3 sigmaForce() const { return sigma(prm.force_space_dimension, Pblock(0), prm.force_or_motion_axis); }
	M3 sigmaMoment() const { return sigma(prm.moment_space_dimension, Pblock(3), prm.moment_or_rotmotion_axis); }
	M3 sigmaPosition() const { return Pblock(0) * (M3::Identity() - sigmaForce()) * Pblock(0).T(); }        // .cpp:927-930
	M3 sigmaOrientation() const { return Pblock(3) * (M3::Identity() - sigmaMoment()) * Pblock(3).T(); }    // .cpp:968-971

	void updateTaskModel(const Mat& N_prec_) override {  // MotionForceTask.cpp:247-268
		N_prec = N_prec_;
		J = P * robot->JWorldFrame(body, p_c);
		Jp = J * N_prec;
		handler->updateTaskModel(Jp, N_prec);
		N = handler->N;
	}

	Mat computeTorques(const Mat& tau_prec) override {  // MotionForceTask.cpp:270-276
		Mat t = computeTorques();
		if (prm.apply_disturbance_compensation) {
			// non-default: the *intended* compensation with the world-frame Lambda
			Mat Lw = handler->U_ns * handler->L_ns * handler->U_ns.T();
			if (handler->U_ns.norm() != 0) t = t - Jp.T() * Lw * J * robot->MInv() * tau_prec;
		}
		// Q1: _Lambda is only ever setZero(6,6) (.cpp:140), so the reference's compensation term is identically zero
		tau = t;
		return t;
	}

	Mat computeTorques() override {  // MotionForceTask.cpp:278-509
		const int dof = robot->dof();
		J = P * robot->JWorldFrame(body, p_c);
		Jp = J * N_prec;
		x = robot->positionInWorld(body, p_c);
		R = robot->rotationInWorld(body, R_c);
		ori_err = orientationError(goal_ori, R);
		v = w = V3();
		for (int k = 0; k < 3; ++k)
			for (int j = 0; j < dof; ++j) { v[k] += J(k, j) * robot->dq[j]; w[k] += J(3 + k, j) * robot->dq[j]; }
		const M3 Sf = sigmaForce(), Sm = sigmaMoment(), Sp = sigmaPosition(), So = sigmaOrientation();
		const M3 rot = rotParam();
		const V3 gf = rot * goal_force, gm = rot * goal_moment;  // getGoalForce/Moment, .cpp:755-769
		const V3 kp_f(prm.kp_force, prm.kp_force, prm.kp_force), kv_f(prm.kv_force, prm.kv_force, prm.kv_force), ki_f(prm.ki_force, prm.ki_force, prm.ki_force);
		const V3 kp_m(prm.kp_moment, prm.kp_moment, prm.kp_moment), kv_m(prm.kv_moment, prm.kv_moment, prm.kv_moment), ki_m(prm.ki_moment, prm.ki_moment, prm.ki_moment);
		const V3 kp_p(prm.kp_pos), kv_p(prm.kv_pos), ki_p(prm.ki_pos), kp_o(prm.kp_ori), kv_o(prm.kv_ori), ki_o(prm.ki_ori);
		V3 F_fb, M_fb, F_p, F_o;
		// Q2: the rotated gain matrices of .cpp:319-324 are dead code
		if (prm.closed_loop_force_control) {  // .cpp:327-349
			integ_force = integ_force + (Sf * (sensed_force - gf)) * dt;
			V3 fb = Sf * (-1.0 * cwise(kp_f, sensed_force - gf) - cwise(ki_f, integ_force));
			if (fb.norm() > prm.max_force_control_feedback_output) fb = fb * (prm.max_force_control_feedback_output / fb.norm());
			F_fb = popc.computePassivitySaturatedForce(Sf * gf, Sf * sensed_force, Sf * fb, Sf * v, kv_f, prm.kff_force);
		} else {
			F_fb = Sf * (-1.0 * cwise(kv_f, v));
		}
		if (prm.closed_loop_moment_control) {  // .cpp:357-378
			integ_moment = integ_moment + (Sm * (sensed_moment - gm)) * dt;
			V3 mb = Sm * (-1.0 * cwise(kp_m, sensed_moment - gm) - cwise(ki_m, integ_moment));
			if (mb.norm() > prm.max_moment_control_feedback_output) mb = mb * (prm.max_moment_control_feedback_output / mb.norm());
			M_fb = Sm * (mb - cwise(kv_m, w));
		} else {
			M_fb = Sm * (-1.0 * cwise(kv_m, w));
		}
		// .cpp:387-392
		des_pos = goal_pos; des_ori = goal_ori;
		des_linvel = goal_linvel; des_angvel = goal_angvel;
		des_linacc = goal_linacc; des_angacc = goal_angacc;
		if (use_internal_otg) {  // .cpp:394-407
			const double gp[3] = {goal_pos.x, goal_pos.y, goal_pos.z}, gv[3] = {goal_linvel.x, goal_linvel.y, goal_linvel.z},
						 gw[3] = {goal_angvel.x, goal_angvel.y, goal_angvel.z};
			double gR[9], Rn[9], wn[3], an[3];
			for (int i = 0; i < 9; ++i) gR[i] = goal_ori.m[i / 3][i % 3];
			otg_->setGoalPositionAndLinearVelocity(gp, gv);
			if (!otg_->setGoalOrientationAndAngularVelocity(gR, gw))
				throw std::invalid_argument("goal orientation is not a valid rotation matrix OTG_6dof_cartesian::setGoalOrientationAndAngularVelocity\n");
			otg_->update();
			des_pos = V3(otg_->out_p); des_linvel = V3(otg_->out_v); des_linacc = V3(otg_->out_a);
			otg_->getNextOrientation(Rn);
			des_ori = M3(Rn);
			otg_->getNextAngularVelocity(wn); otg_->getNextAngularAcceleration(an);
			des_angvel = V3(wn); des_angacc = V3(an);
		}
		integ_pos = integ_pos + (Sp * (x - des_pos)) * dt;  // .cpp:411-413
		if (prm.use_velocity_saturation) {                  // .cpp:416-429
			double kvi[3];
			pseudoInverseDiag(prm.kv_pos, 3, kvi);
			V3 kv_inv(kvi);
			des_linvel = -1.0 * cwise(kp_p, cwise(kv_inv, Sp * (x - des_pos))) - cwise(ki_p, cwise(kv_inv, integ_pos));
			if (des_linvel.norm() > prm.linear_saturation_velocity) des_linvel = des_linvel * (prm.linear_saturation_velocity / des_linvel.norm());
			F_p = Sp * (des_linacc - cwise(kv_p, v - des_linvel));
		} else {  // .cpp:431-436
			F_p = Sp * (des_linacc - cwise(kp_p, x - des_pos) - cwise(kv_p, v - des_linvel) - cwise(ki_p, integ_pos));
		}
		V3 step_ori_err = So * orientationError(des_ori, R);  // .cpp:441-443
		integ_ori = integ_ori + step_ori_err * dt;
		if (prm.use_velocity_saturation) {  // .cpp:449-461
			double kvi[3];
			pseudoInverseDiag(prm.kv_ori, 3, kvi);
			V3 kv_inv(kvi);
			des_angvel = -1.0 * cwise(kp_o, cwise(kv_inv, step_ori_err)) - cwise(ki_o, cwise(kv_inv, integ_ori));
			if (des_angvel.norm() > prm.angular_saturation_velocity) des_angvel = des_angvel * (prm.angular_saturation_velocity / des_angvel.norm());
			F_o = So * (des_angacc - cwise(kv_o, w - des_angvel));
		} else {  // .cpp:463-467
			F_o = So * (des_angacc - cwise(kp_o, step_ori_err) - cwise(kv_o, w - des_angvel) - cwise(ki_o, integ_ori));
		}
		Mat F_star(6, 1), F_fm(6, 1), ff(6, 1);
		V3 ffh = Sf * gf, fft = Sm * gm;
		if (prm.closed_loop_force_control) { ffh = prm.kff_force * ffh; fft = prm.kff_moment * fft; }  // Q3, .cpp:484-487
		for (int k = 0; k < 3; ++k) {
			F_star.a[k] = F_p[k]; F_star.a[3 + k] = F_o[k];
			F_fm.a[k] = F_fb[k]; F_fm.a[3 + k] = M_fb[k];
			ff.a[k] = ffh[k]; ff.a[3 + k] = fft[k];
			unit_mass_force[k] = F_p[k]; unit_mass_force[3 + k] = F_o[k];
		}
		if (pos_range + ori_range == 0) { tau = Mat::Zero(dof, 1); return tau; }  // .cpp:300-304 (unreachable: ctor throws)
		tau = handler->computeTorques(F_star, F_fm + ff);  // .cpp:505-506
		return tau;
	}
};

struct JointTask : Task {
	osc_joint_task_desc desc;
	osc_joint_task_params prm;
	Mat S;  // _joint_selection
	int k;
	Mat Jp, range, Lam, Lam_mod;
	double cur_pos[8], cur_vel[8], goal_pos[8], goal_vel[8], goal_acc[8], des_pos[8], des_vel[8], des_acc[8], integ[8];
	bool use_internal_otg = false;  // JointTask.h:440 (reference default: on; here off until enabled, see MotionForceTask)
	otg::OtgJoints* otg_ = nullptr;
	~JointTask() override { delete otg_; }
	// JointTask.cpp:358-381
	void enableInternalOtgAccelerationLimited(const double* max_v, const double* max_a) {
		for (int i = 0; i < k; ++i)
			if (max_v[i] <= 0 || max_a[i] <= 0) throw std::invalid_argument("max velocity / acceleration cannot be 0 or negative in OTG_joints\n");
		if (!otg_) otg_ = new otg::OtgJoints(cur_pos, k, dt);  // constructor, .cpp:71-72
		if (!use_internal_otg) otg_->reInitialize(cur_pos);
		otg_->setLimits(max_v, max_a);
		otg_->disableJerkLimits();
		use_internal_otg = true;
	}
	void disableInternalOtg() { use_internal_otg = false; }  // JointTask.h:320

	static void default_params(osc_joint_task_params* p, int k) {  // JointTask.h:31-45
		std::memset(p, 0, sizeof *p);
		for (int i = 0; i < k; ++i) { p->kp[i] = 50; p->kv[i] = 14; p->ki[i] = 0; p->saturation_velocity[i] = M_PI / 3; }
		p->dynamic_decoupling_type = OSC_BOUNDED_INERTIA_ESTIMATES; p->bie_threshold = 0.1;
	}

	JointTask(Model* r, const osc_joint_task_desc& d) : desc(d) {  // JointTask.cpp:14-89
		robot = r; type = OSC_TASK_JOINT; dt = d.loop_timestep;
		const int dof = r->dof();
		if (d.full) S = Mat::Identity(dof);
		else {
			S = Mat(d.k, dof);
			for (int i = 0; i < d.k * dof; ++i) S.a[i] = d.selection[i];
			// FullPivLU rank check, JointTask.cpp:33-39 (here: smallest singular value)
			Svd s = jacobi_svd(S);
			if (d.k > dof || s.s[d.k - 1] <= 1e-12 * s.s[0])
				throw std::invalid_argument("joint selection matrix is not full rank in JointTask constructor\n");
		}
		k = S.r;
		default_params(&prm, k);
		N_prec = Mat::Identity(dof); N = Mat::Zero(dof, dof);
		Lam = Mat::Identity(k); Lam_mod = Mat::Identity(k);
		Jp = S; range = Mat::Identity(k);
		tau = Mat::Zero(dof, 1);
		reInitializeTask();
	}
	bool isFullJointTask() const { return k == robot->dof(); }  // JointTask.h:123

	void reInitializeTask() override {  // JointTask.cpp:91-107
		for (int i = 0; i < k; ++i) {
			double s = 0;
			for (int j = 0; j < robot->dof(); ++j) s += S(i, j) * robot->q[j];
			cur_pos[i] = goal_pos[i] = des_pos[i] = s;
			cur_vel[i] = goal_vel[i] = goal_acc[i] = des_vel[i] = des_acc[i] = integ[i] = 0;
		}
		if (otg_) otg_->reInitialize(cur_pos);  // .cpp:106
	}

	void updateTaskModel(const Mat& N_prec_) override {  // JointTask.cpp:218-283
		const int dof = robot->dof();
		N_prec = N_prec_;
		Jp = S * N_prec;
		range = matrixRangeBasis(Jp);
		if (range.norm() == 0) { N = Mat::Identity(dof); return; }
		Mat Jr = range.T() * Jp;
		auto op = robot->operationalSpaceMatrices(Jr);
		Lam = op.Lambda; N = op.N;
		switch (prm.dynamic_decoupling_type) {
			case OSC_FULL_DYNAMIC_DECOUPLING: Lam_mod = Lam; break;
			case OSC_IMPEDANCE: Lam_mod = Mat::Identity(range.c); break;
			default: {
				Mat M_BIE = robot->M();
				for (int i = 0; i < dof; ++i)
					if (M_BIE(i, i) < prm.bie_threshold) M_BIE(i, i) = prm.bie_threshold;
				Mat M_inv_BIE = lu_inverse(M_BIE);
				Lam_mod = lu_inverse(range.T() * Jp * M_inv_BIE * Jp.T() * range);
			}
		}
	}

	Mat computeTorques(const Mat& tau_prec) override {  // JointTask.cpp:285-292
		Mat t = computeTorques();
		if (range.norm() != 0) {
			Mat comp = Jp.T() * range * Lam * range.T() * S * robot->MInv() * tau_prec;
			t = t - comp;
		}
		tau = t;
		return t;
	}

	Mat computeTorques() override {  // JointTask.cpp:294-356
		const int dof = robot->dof();
		Jp = S * N_prec;
		for (int i = 0; i < k; ++i) {
			double sp = 0, sv = 0;
			for (int j = 0; j < dof; ++j) { sp += S(i, j) * robot->q[j]; sv += S(i, j) * robot->dq[j]; }
			cur_pos[i] = sp; cur_vel[i] = sv;
		}
		if (range.norm() == 0) { tau = Mat::Zero(dof, 1); return tau; }
		Mat u(k, 1), acc(k, 1);
		for (int i = 0; i < k; ++i) { des_pos[i] = goal_pos[i]; des_vel[i] = goal_vel[i]; des_acc[i] = goal_acc[i]; }
		if (use_internal_otg) {  // .cpp:313-320
			otg_->setGoalPositionAndVelocity(goal_pos, goal_vel);
			otg_->update();
			for (int i = 0; i < k; ++i) { des_pos[i] = otg_->out_p[i]; des_vel[i] = otg_->out_v[i]; des_acc[i] = otg_->out_a[i]; }
		}
		for (int i = 0; i < k; ++i) integ[i] += (cur_pos[i] - des_pos[i]) * dt;
		if (prm.use_velocity_saturation) {
			double kvi[8];
			pseudoInverseDiag(prm.kv, k, kvi);
			for (int i = 0; i < k; ++i) {  // Q7: loop to k (the reference loops to dof)
				des_vel[i] = -prm.kp[i] * kvi[i] * (cur_pos[i] - des_pos[i]) - prm.ki[i] * kvi[i] * integ[i];
				if (des_vel[i] > prm.saturation_velocity[i]) des_vel[i] = prm.saturation_velocity[i];
				else if (des_vel[i] < -prm.saturation_velocity[i]) des_vel[i] = -prm.saturation_velocity[i];
				u.a[i] = -prm.kv[i] * (cur_vel[i] - des_vel[i]);
			}
		} else {
			for (int i = 0; i < k; ++i)
				u.a[i] = -prm.kp[i] * (cur_pos[i] - des_pos[i]) - prm.kv[i] * (cur_vel[i] - des_vel[i]) - prm.ki[i] * integ[i];
		}
		for (int i = 0; i < k; ++i) acc.a[i] = des_acc[i];
		Mat f = Lam * range.T() * acc + Lam_mod * range.T() * u;
		tau = Jp.T() * range * f;
		return tau;
	}
};

// ------------------------------------------------------------------------------------------------------------------
// One robot instance with its own model copy and task hierarchy (RobotController.cpp)
// ------------------------------------------------------------------------------------------------------------------
struct Robot {
	Model model;
	std::vector<Task*> tasks;
	bool gravity_comp = false;
	Mat N_total, tau;
	double gravity[8];
	double nle[8] = {0};  // b + g of the last rollout step
	explicit Robot(const osc_chain_desc& c) : model(c) {
		tau = Mat::Zero(c.n, 1); N_total = Mat::Identity(c.n);
		for (double& g : gravity) g = 0;
	}
	~Robot() { for (Task* t : tasks) delete t; }
	Robot(const Robot&) = delete;
	Robot& operator=(const Robot&) = delete;

	void updateControllerTaskModels() {  // RobotController.cpp:53-60
		Mat N_prec = Mat::Identity(model.dof());
		for (Task* t : tasks) {
			t->updateTaskModel(N_prec);
			N_prec = t->getTaskAndPreviousNullspace();
		}
		N_total = N_prec;
	}
	void computeControlTorques(bool stacked) {  // RobotController.cpp:62-74; stacked=false: examples/04...cpp:188-206
		Mat control_torques = Mat::Zero(model.dof(), 1);
		for (Task* t : tasks) {
			Mat tt = stacked ? t->computeTorques(control_torques) : t->computeTorques();
			t->tau = tt;
			control_torques = control_torques + tt;
		}
		if (gravity_comp) {
			model.jointGravityVector(gravity);
			for (int i = 0; i < model.dof(); ++i) control_torques.a[i] += gravity[i];
		}
		tau = control_torques;
	}
};

}  // namespace orc
