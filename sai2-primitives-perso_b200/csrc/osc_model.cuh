// osc_model.cuh -- per-thread robot model: forward kinematics, point Jacobian, mass matrix (world-frame composite
// rigid body), gravity vector.  Stands in for Sai2Model::updateKinematics / updateModel / JWorldFrame /
// jointGravityVector (un-vendored sai2-model; reference call sites
// examples/05-using_robot_controller/05-using_robot_controller.cpp:143-145, src/tasks/MotionForceTask.cpp:262,286-289).
//
// The host canonicalises the chain (osc_capi.cu: canonicalise_chain) so that EVERY joint axis is the local z axis and
// the root transform is T_world_robot; all quantities below are therefore world-frame directly.
#pragma once
#include "osc_math.cuh"

namespace osc {

template <class T, int N> struct ChainC {
	T rot[N][3][3];   // joint origin rotation in the parent body frame (axis-alignment folded in)
	T xyz[N][3];      // joint origin translation in the parent body frame
	int prismatic[N];
	T mass[N];
	T com[N][3];
	T inertia[N][6];  // xx xy xz yy yz zz about the com, body frame
	T hb[N][3];       // first moment m * com about the body (= joint) origin, body frame
	T Ib[N][6];       // inertia about the body origin, body frame: inertia + m (|c|^2 1 - c c^T)   (body-frame CRBA of osc_fused2.cuh)
	T q_lower[N], q_upper[N], effort[N];
	T world_rot[3][3], world_xyz[3], gravity[3];
};

template <class T, int N> struct Kin {
	T R[N][3][3];  // world orientation of body i (its z column is the joint axis)
	T o[N][3];     // world position of joint/body origin i
};

// Sai2Model::updateKinematics
template <class T, int N> OSC_DEV void forward_kinematics(const ChainC<T, N>& ch, const T (&q)[N], Kin<T, N>& k) {
	T Rp[3][3], pp[3];
	OSC_UNROLL for (int a = 0; a < 3; ++a) {
		pp[a] = ch.world_xyz[a];
		OSC_UNROLL for (int b = 0; b < 3; ++b) Rp[a][b] = ch.world_rot[a][b];
	}
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		T pre[3][3];
		mm3(Rp, ch.rot[i], pre);
		T d[3];
		mv3(Rp, ch.xyz[i], d);
		OSC_UNROLL for (int a = 0; a < 3; ++a) pp[a] += d[a];
		if (ch.prismatic[i]) {
			OSC_UNROLL for (int a = 0; a < 3; ++a) {
				pp[a] = Num<T>::fma_(q[i], pre[a][2], pp[a]);
				OSC_UNROLL for (int b = 0; b < 3; ++b) Rp[a][b] = pre[a][b];
			}
		} else {
			T s, c;
			Num<T>::sincos_(q[i], &s, &c);
			OSC_UNROLL for (int a = 0; a < 3; ++a) {
				Rp[a][0] = c * pre[a][0] + s * pre[a][1];
				Rp[a][1] = c * pre[a][1] - s * pre[a][0];
				Rp[a][2] = pre[a][2];
			}
		}
		OSC_UNROLL for (int a = 0; a < 3; ++a) {
			k.o[i][a] = pp[a];
			OSC_UNROLL for (int b = 0; b < 3; ++b) k.R[i][a][b] = Rp[a][b];
		}
	}
}

// pose of a frame (p_c, R_c) fixed in `body`: positionInWorld / rotationInWorld
template <class T, int N> OSC_DEV void frame_pose(const Kin<T, N>& k, int body, const T (&p_c)[3], const T (&R_c)[3][3], T (&x)[3], T (&R)[3][3]) {
	OSC_UNROLL for (int b = 0; b < N; ++b) {
		if (b == body) {
			T d[3];
			mv3(k.R[b], p_c, d);
			OSC_UNROLL for (int a = 0; a < 3; ++a) x[a] = k.o[b][a] + d[a];
			mm3(k.R[b], R_c, R);
		}
	}
}

// Sai2Model::JWorldFrame: 6 x N, linear rows first; columns of joints beyond `body` are zero
template <class T, int N> OSC_DEV void point_jacobian(const ChainC<T, N>& ch, const Kin<T, N>& k, int body, const T (&x)[3], T (&J)[6][N]) {
	OSC_UNROLL for (int i = 0; i < N; ++i) {
		const T z[3] = {k.R[i][0][2], k.R[i][1][2], k.R[i][2][2]};
		const bool on = i <= body;
		if (ch.prismatic[i]) {
			OSC_UNROLL for (int a = 0; a < 3; ++a) { J[a][i] = on ? z[a] : T(0); J[3 + a][i] = T(0); }
		} else {
			const T d[3] = {x[0] - k.o[i][0], x[1] - k.o[i][1], x[2] - k.o[i][2]};
			T l[3];
			cross3(z, d, l);
			OSC_UNROLL for (int a = 0; a < 3; ++a) { J[a][i] = on ? l[a] : T(0); J[3 + a][i] = on ? z[a] : T(0); }
		}
	}
}

// Mass matrix by the composite-rigid-body recursion carried out in WORLD coordinates about the world origin:
// every body contributes a 10-parameter spatial inertia (m, h = m c, I_O); composites are plain sums tip -> base;
// column i is the momentum (k, l) of composite i under the unit joint twist s_i, projected on every ancestor twist.
template <class T, int N> OSC_DEV void mass_matrix(const ChainC<T, N>& ch, const Kin<T, N>& k, T (&M)[N][N]) {
	T cm = T(0), ch_[3] = {T(0), T(0), T(0)}, cI[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};  // composite
	OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
		// body i in world coordinates
		T c[3];
		mv3(k.R[i], ch.com[i], c);
		OSC_UNROLL for (int a = 0; a < 3; ++a) c[a] += k.o[i][a];
		const T Ib[3][3] = {{ch.inertia[i][0], ch.inertia[i][1], ch.inertia[i][2]},
							{ch.inertia[i][1], ch.inertia[i][3], ch.inertia[i][4]},
							{ch.inertia[i][2], ch.inertia[i][4], ch.inertia[i][5]}};
		T RI[3][3];
		mm3(k.R[i], Ib, RI);
		const T m = ch.mass[i];
		const T cc = dot3(c, c);
		// I_O = R Ib R^T + m (|c|^2 I - c c^T), 6 unique entries xx xy xz yy yz zz
		int e = 0;
		OSC_UNROLL for (int a = 0; a < 3; ++a)
			OSC_UNROLL for (int b = a; b < 3; ++b) {
				T v = RI[a][0] * k.R[i][b][0] + RI[a][1] * k.R[i][b][1] + RI[a][2] * k.R[i][b][2];
				v += m * ((a == b ? cc : T(0)) - c[a] * c[b]);
				cI[e] += v;
				++e;
			}
		cm += m;
		OSC_UNROLL for (int a = 0; a < 3; ++a) ch_[a] = Num<T>::fma_(m, c[a], ch_[a]);
		// unit twist of joint i about the world origin
		const T z[3] = {k.R[i][0][2], k.R[i][1][2], k.R[i][2][2]};
		T w[3], v[3];
		if (ch.prismatic[i]) {
			OSC_UNROLL for (int a = 0; a < 3; ++a) { w[a] = T(0); v[a] = z[a]; }
		} else {
			OSC_UNROLL for (int a = 0; a < 3; ++a) w[a] = z[a];
			cross3(k.o[i], z, v);
		}
		// momentum of the composite: l = m v + w x h ; kk = I_O w + h x v
		T wxh[3], hxv[3], l[3], kk[3];
		cross3(w, ch_, wxh);
		cross3(ch_, v, hxv);
		OSC_UNROLL for (int a = 0; a < 3; ++a) l[a] = Num<T>::fma_(cm, v[a], wxh[a]);
		kk[0] = cI[0] * w[0] + cI[1] * w[1] + cI[2] * w[2] + hxv[0];
		kk[1] = cI[1] * w[0] + cI[3] * w[1] + cI[4] * w[2] + hxv[1];
		kk[2] = cI[2] * w[0] + cI[4] * w[1] + cI[5] * w[2] + hxv[2];
		OSC_UNROLL for (int j = 0; j < N; ++j) {
			if (j <= i) {
				const T zj[3] = {k.R[j][0][2], k.R[j][1][2], k.R[j][2][2]};
				T h;
				if (ch.prismatic[j]) {
					h = dot3(zj, l);
				} else {
					T vj[3];
					cross3(k.o[j], zj, vj);
					h = dot3(zj, kk) + dot3(vj, l);
				}
				M[i][j] = h; M[j][i] = h;
			}
		}
	}
}

// Sai2Model::jointGravityVector: g_i = -sum_b m_b (dv_cb/dq_i) . g_world, accumulated tip -> base
template <class T, int N> OSC_DEV void gravity_vector(const ChainC<T, N>& ch, const Kin<T, N>& k, T (&g)[N]) {
	T cm = T(0), mc[3] = {T(0), T(0), T(0)};
	OSC_UNROLL for (int i = N - 1; i >= 0; --i) {
		T c[3];
		mv3(k.R[i], ch.com[i], c);
		cm += ch.mass[i];
		OSC_UNROLL for (int a = 0; a < 3; ++a) mc[a] = Num<T>::fma_(ch.mass[i], c[a] + k.o[i][a], mc[a]);
		const T z[3] = {k.R[i][0][2], k.R[i][1][2], k.R[i][2][2]};
		if (ch.prismatic[i]) {
			g[i] = -cm * dot3(z, ch.gravity);
		} else {
			// sum_b m_b z x (c_b - o_i) = z x (mc - cm o_i)
			const T d[3] = {mc[0] - cm * k.o[i][0], mc[1] - cm * k.o[i][1], mc[2] - cm * k.o[i][2]};
			T l[3];
			cross3(z, d, l);
			g[i] = -dot3(l, ch.gravity);
		}
	}
}

}  // namespace osc
