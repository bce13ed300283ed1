// osc_mft_compute.cuh -- MotionForceTask::computeTorques (src/tasks/MotionForceTask.cpp:270-509), the sigma selection
// matrices (.cpp:892-971), POPCExplicitForceControl::computePassivitySaturatedForce
// (src/helper_modules/POPCExplicitForceControl.cpp:30-96) and SingularityHandler::computeTorques
// (src/tasks/SingularityHandler.cpp:297-368).
#pragma once
#include "osc_mft_update.cuh"

namespace osc {

// sigmaForce / sigmaMoment (.cpp:892-925 / 932-966): Psel (a a^T | I - a a^T) Psel^T with a = rot * axis
template <class T> OSC_DEV void sigma_matrix(int dim, const T (&Psel)[3][3], const T (&rot)[3][3], const T (&axis)[3], T (&S)[3][3]) {
	if (dim == 0) {
		OSC_UNROLL for (int i = 0; i < 3; ++i)
			OSC_UNROLL for (int j = 0; j < 3; ++j) S[i][j] = T(0);
		return;
	}
	if (dim == 3) {
		OSC_UNROLL for (int i = 0; i < 3; ++i)
			OSC_UNROLL for (int j = 0; j < 3; ++j) S[i][j] = Psel[i][j];
		return;
	}
	T a[3];
	mv3(rot, axis, a);
	T Mid[3][3], Tmp[3][3];
	OSC_UNROLL for (int i = 0; i < 3; ++i)
		OSC_UNROLL for (int j = 0; j < 3; ++j) {
			const T aa = a[i] * a[j];
			Mid[i][j] = (dim == 1) ? aa : ((i == j ? T(1) : T(0)) - aa);
		}
	mm3(Psel, Mid, Tmp);
	OSC_UNROLL for (int i = 0; i < 3; ++i)
		OSC_UNROLL for (int j = 0; j < 3; ++j) S[i][j] = Tmp[i][0] * Psel[j][0] + Tmp[i][1] * Psel[j][1] + Tmp[i][2] * Psel[j][2];
}
// P (I - S) P^T  (sigmaPosition / sigmaOrientation, .cpp:927-930, 968-971)
template <class T> OSC_DEV void sigma_complement(const T (&Psel)[3][3], const T (&S)[3][3], T (&O)[3][3]) {
	T Mid[3][3], Tmp[3][3];
	OSC_UNROLL for (int i = 0; i < 3; ++i)
		OSC_UNROLL for (int j = 0; j < 3; ++j) Mid[i][j] = (i == j ? T(1) : T(0)) - S[i][j];
	mm3(Psel, Mid, Tmp);
	OSC_UNROLL for (int i = 0; i < 3; ++i)
		OSC_UNROLL for (int j = 0; j < 3; ++j) O[i][j] = Tmp[i][0] * Psel[j][0] + Tmp[i][1] * Psel[j][1] + Tmp[i][2] * Psel[j][2];
}

// POPCExplicitForceControl::computePassivitySaturatedForce; the std::queue is a ring of OSC_POPC_RING_CAPACITY slots
// in HBM (quirk Q9).  Returns R_c * vcl - kv_f * vr.
template <class T, int N, class S = T>
OSC_DEV void popc_force(const MftC<T>& tc, const Row<T, S>& tP, const Row<int32_t>& tI, int64_t stride, int64_t r, const T (&fd)[3], const T (&fs)[3],
						const T (&vcl)[3], const T (&vr)[3], T (&out)[3], int& status) {
	using L = MftL<N>;
	const T kv = tc.kv_force;
	if (!tc.passivity || tc.popc_ring == nullptr) {  // .cpp:33-35
		OSC_UNROLL for (int a = 0; a < 3; ++a) out[a] = vcl[a] - kv * vr[a];
		return;
	}
	T PO = tP(L::POPC_PO), Ec = tP(L::POPC_ECORR), Rc = tP(L::POPC_RC), vsq = tP(L::POPC_VSQ);
	int counter = tI(L::POPC_COUNTER), head = tI(L::POPC_HEAD), len = tI(L::POPC_LEN);
	T Fcmd[3], fdiff[3];
	OSC_UNROLL for (int a = 0; a < 3; ++a) { Fcmd[a] = tc.kff_force * fd[a] + Rc * vcl[a] - kv * vr[a]; fdiff[a] = fs[a] - fd[a]; }
	const T vc2 = dot3(vcl, vcl);
	const T p = (dot3(fdiff, vcl) - dot3(Fcmd, vr)) * tc.dt;
	PO += p;
	S* const ring = static_cast<S*>(tc.popc_ring);  // the window lives in the batch storage type
	if (len == 1024) {  // overflow: drop the oldest sample (status bit), Q9
		head = (head + 1) & 1023; --len; status |= 2;
	}
	ring[(int64_t)((head + len) & 1023) * stride + r] = (S)p;
	++len;
	if (PO + Ec > T(0)) {  // _stored_energy_PO is identically 0 (.cpp:13)
		while (len > 250) {
			const T front = (T)ring[(int64_t)head * stride + r];
			if (PO + Ec > front) {
				if (front > T(0)) PO -= front;
				head = (head + 1) & 1023; --len;
			} else break;
		}
	}
	if (counter <= 0) {  // .cpp:64-90
		counter = 50;
		const T old = Rc;
		if (PO + Ec < T(0)) {
			Rc = T(1) + (PO + Ec) / (vsq * tc.dt);
			if (Rc > T(1)) Rc = T(1);
			if (Rc < T(0)) Rc = T(0);
			if (Rc != Rc) status |= 4;  // quirk Q8
		} else {
			Rc = (T(1) + (T(0.1) * T(50) - T(1)) * Rc) / (T(0.1) * T(50));
		}
		Ec += (T(1) - old) * vsq * tc.dt;
		vsq = T(0);
	}
	--counter;
	vsq += vc2;
	OSC_UNROLL for (int a = 0; a < 3; ++a) out[a] = Rc * vcl[a] - kv * vr[a];
	tP(L::POPC_PO) = PO; tP(L::POPC_ECORR) = Ec; tP(L::POPC_RC) = Rc; tP(L::POPC_VSQ) = vsq;
	tI(L::POPC_COUNTER) = counter; tI(L::POPC_HEAD) = head; tI(L::POPC_LEN) = len;
}

#define OSC_FENCE() asm volatile("" ::: "memory")  // keeps ptxas from interleaving independent phases (register pressure)

// One 3-dof half of the motion law (MotionForceTask.cpp:411-437 linear / 441-468 angular): integrates Sigma e, returns
// Sigma (a_des - kp e - kv (v - v_des) - ki int) or its velocity-saturated variant; vd is the desired velocity (in/out).
template <class T, class S = T>
OSC_DEV void motion_half(const T (&Sg)[3][3], const T (&err)[3], bool err_is_projected, const T (&vel)[3], T (&vd)[3], const T (&acc)[3],
						 const T (&kp)[3], const T (&kv)[3], const T (&ki)[3], bool use_sat, T sat, T dt, const Row<T, S>& tP, int off_int, T (&out)[3]) {
	T se[3], ip[3], raw[3];
	mv3(Sg, err, se);
	OSC_UNROLL for (int a = 0; a < 3; ++a) { ip[a] = tP(off_int + a) + se[a] * dt; tP(off_int + a) = ip[a]; }
	if (use_sat) {
		T mx = T(0);
		OSC_UNROLL for (int a = 0; a < 3; ++a) mx = Num<T>::abs_(kv[a]) > mx ? Num<T>::abs_(kv[a]) : mx;
		OSC_UNROLL for (int a = 0; a < 3; ++a) {
			const T kvi = Num<T>::abs_(kv[a]) > T(1e-6) * mx ? T(1) / kv[a] : T(0);  // computePseudoInverse of the diagonal gain
			vd[a] = -kp[a] * kvi * se[a] - ki[a] * kvi * ip[a];
		}
		const T nv = norm3(vd);
		if (nv > sat) {
			const T sc = sat / nv;
			OSC_UNROLL for (int a = 0; a < 3; ++a) vd[a] *= sc;
		}
		OSC_UNROLL for (int a = 0; a < 3; ++a) raw[a] = acc[a] - kv[a] * (vel[a] - vd[a]);
	} else {
		// the linear law uses the UNPROJECTED error in the kp term (.cpp:431-434), the angular one the projected error (.cpp:443)
		OSC_UNROLL for (int a = 0; a < 3; ++a) raw[a] = acc[a] - kp[a] * (err_is_projected ? se[a] : err[a]) - kv[a] * (vel[a] - vd[a]) - ki[a] * ip[a];
	}
	mv3(Sg, raw, out);
}

// The control law: returns F* (unit mass force) and F_f (force feedback + feed-forward); updates integrators and caches.
// The linear half (force space + position) is finished before the angular half starts, so that only one pair of
// selection matrices is live at a time.  FULL = the task controls all six directions (P = I).
template <class T, int N, bool FULL = false, class S = T>
OSC_DEV void mft_control_law_vw(const MftC<T>& tc, const Row<T, S>& tP, const Row<T, S>& tS, const Row<int32_t>& tI, int64_t stride, int64_t r,
								const T (&v)[3], const T (&w)[3], const T (&x)[3], const T (&R)[3][3], T (&Fstar)[6], T (&Ff)[6], int& status,
								bool export_vectors = true) {
	using L = MftL<N>;
	if (export_vectors) {
		st_vec<T, 3>(tS, L::CUR_V, v);
		st_vec<T, 3>(tS, L::CUR_W, w);
	}
	st_vec<T, 3>(tP, L::CUR_POS, x);
	st_mat<T, 3, 3>(tP, L::CUR_ORI, R);
	T rot[3][3];
	OSC_UNROLL for (int i = 0; i < 3; ++i)
		OSC_UNROLL for (int j = 0; j < 3; ++j) rot[i][j] = tc.in_compliant_frame ? R[i][j] : (i == j ? T(1) : T(0));
	const T Id[3][3] = {{T(1), T(0), T(0)}, {T(0), T(1), T(0)}, {T(0), T(0), T(1)}};

	{  // ---- linear half: force space (.cpp:327-354) and position (.cpp:411-437) ----
		T Sf[3][3], gf_[3], gf[3], Ffb[3];
		if constexpr (FULL) sigma_matrix<T>(tc.force_dim, Id, rot, tc.force_axis, Sf);
		else sigma_matrix<T>(tc.force_dim, tc.Pt, rot, tc.force_axis, Sf);
		ld_vec<T, 3>(tP, L::GOAL_FORCE, gf_);
		mv3(rot, gf_, gf);  // getGoalForce, .cpp:755-761
		if (tc.closed_loop_force) {
			T fs[3], e[3], se[3], ie[3], raw[3], fb[3];
			ld_vec<T, 3>(tP, L::SENSED_F, fs);
			OSC_UNROLL for (int a = 0; a < 3; ++a) e[a] = fs[a] - gf[a];
			mv3(Sf, e, se);
			OSC_UNROLL for (int a = 0; a < 3; ++a) { ie[a] = tP(L::INT_F + a) + se[a] * tc.dt; tP(L::INT_F + a) = ie[a]; }
			OSC_UNROLL for (int a = 0; a < 3; ++a) raw[a] = -tc.kp_force * e[a] - tc.ki_force * ie[a];
			mv3(Sf, raw, fb);
			const T nf = norm3(fb);
			if (nf > tc.max_force_fb) {
				const T sc = tc.max_force_fb / nf;
				OSC_UNROLL for (int a = 0; a < 3; ++a) fb[a] *= sc;
			}
			T a1[3], a2[3], a3[3], a4[3];
			mv3(Sf, gf, a1); mv3(Sf, fs, a2); mv3(Sf, fb, a3); mv3(Sf, v, a4);
			popc_force<T, N>(tc, tP, tI, stride, r, a1, a2, a3, a4, Ffb, status);
		} else {  // .cpp:352-353
			T raw[3];
			OSC_UNROLL for (int a = 0; a < 3; ++a) raw[a] = -tc.kv_force * v[a];
			mv3(Sf, raw, Ffb);
		}
		T ffh[3];
		mv3(Sf, gf, ffh);
		OSC_UNROLL for (int a = 0; a < 3; ++a) Ff[a] = Ffb[a] + (tc.closed_loop_force ? tc.kff_force * ffh[a] : ffh[a]);
		T Sp[3][3];
		if constexpr (FULL) sigma_complement<T>(Id, Sf, Sp);
		else sigma_complement<T>(tc.Pt, Sf, Sp);
		T xg[3], vd[3], ad[3], ex[3], Fp[3];
		ld_vec<T, 3>(tP, L::GOAL_POS, xg);
		ld_vec<T, 3>(tP, L::GOAL_LINVEL, vd);
		ld_vec<T, 3>(tP, L::GOAL_LINACC, ad);
		OSC_UNROLL for (int a = 0; a < 3; ++a) ex[a] = x[a] - xg[a];
		motion_half<T>(Sp, ex, false, v, vd, ad, tc.kp_pos, tc.kv_pos, tc.ki_pos, tc.use_vel_sat != 0, tc.lin_sat, tc.dt, tP, L::INT_POS, Fp);
		OSC_UNROLL for (int a = 0; a < 3; ++a) Fstar[a] = Fp[a];
		if (export_vectors) {
			st_vec<T, 3>(tS, L::DES_POS, xg);
			st_vec<T, 3>(tS, L::DES_LINVEL, vd);
		}
	}
	OSC_FENCE();
	{  // ---- angular half: moment space (.cpp:357-383) and orientation (.cpp:441-468) ----
		T Sm[3][3], gm_[3], gm[3], Mfb[3];
		if constexpr (FULL) sigma_matrix<T>(tc.moment_dim, Id, rot, tc.moment_axis, Sm);
		else sigma_matrix<T>(tc.moment_dim, tc.Pr, rot, tc.moment_axis, Sm);
		ld_vec<T, 3>(tP, L::GOAL_MOMENT, gm_);
		mv3(rot, gm_, gm);
		if (tc.closed_loop_moment) {
			T ms[3], e[3], se[3], ie[3], raw[3], mb[3];
			ld_vec<T, 3>(tP, L::SENSED_M, ms);
			OSC_UNROLL for (int a = 0; a < 3; ++a) e[a] = ms[a] - gm[a];
			mv3(Sm, e, se);
			OSC_UNROLL for (int a = 0; a < 3; ++a) { ie[a] = tP(L::INT_M + a) + se[a] * tc.dt; tP(L::INT_M + a) = ie[a]; }
			OSC_UNROLL for (int a = 0; a < 3; ++a) raw[a] = -tc.kp_moment * e[a] - tc.ki_moment * ie[a];
			mv3(Sm, raw, mb);
			const T nm = norm3(mb);
			if (nm > tc.max_moment_fb) {
				const T sc = tc.max_moment_fb / nm;
				OSC_UNROLL for (int a = 0; a < 3; ++a) mb[a] *= sc;
			}
			OSC_UNROLL for (int a = 0; a < 3; ++a) raw[a] = mb[a] - tc.kv_moment * w[a];
			mv3(Sm, raw, Mfb);
		} else {  // .cpp:381-382
			T raw[3];
			OSC_UNROLL for (int a = 0; a < 3; ++a) raw[a] = -tc.kv_moment * w[a];
			mv3(Sm, raw, Mfb);
		}
		T fft[3];
		mv3(Sm, gm, fft);
		// quirk Q3 (.cpp:484-487): the moment feed-forward is scaled under the *force* flag
		OSC_UNROLL for (int a = 0; a < 3; ++a) Ff[3 + a] = Mfb[a] + (tc.closed_loop_force ? tc.kff_moment * fft[a] : fft[a]);
		T So[3][3];
		if constexpr (FULL) sigma_complement<T>(Id, Sm, So);
		else sigma_complement<T>(tc.Pr, Sm, So);
		T Rg[3][3], oe[3], wd[3], bd[3], Fo[3];
		ld_mat<T, 3, 3>(tP, L::GOAL_ORI, Rg);
		orientation_error<T>(Rg, R, oe);  // .cpp:291-292 (_orientation_error, for getters); desired orientation == goal
		ld_vec<T, 3>(tP, L::GOAL_ANGVEL, wd);
		ld_vec<T, 3>(tP, L::GOAL_ANGACC, bd);
		motion_half<T>(So, oe, true, w, wd, bd, tc.kp_ori, tc.kv_ori, tc.ki_ori, tc.use_vel_sat != 0, tc.ang_sat, tc.dt, tP, L::INT_ORI, Fo);
		OSC_UNROLL for (int a = 0; a < 3; ++a) Fstar[3 + a] = Fo[a];
		if (export_vectors) {
			st_vec<T, 3>(tS, L::ORI_ERR, oe);
			st_mat<T, 3, 3>(tS, L::DES_ORI, Rg);
			st_vec<T, 3>(tS, L::DES_ANGVEL, wd);
		}
	}
	if (export_vectors) st_vec<T, 6>(tS, L::UMF, Fstar);
}

// The law for the plain motion task: no force / moment space (dimension 0), open loop, no velocity saturation, all six
// directions controlled.  Then Sigma_force = Sigma_moment = 0, Sigma_position = Sigma_orientation = I and F_f = 0
// (MotionForceTask.cpp:306-309, 352-353, 381-382, 427-467): F* = a_des - kp e - kv (v - v_des) - ki int e.
template <class T, int N, class S = T>
OSC_DEV void mft_control_law_motion_only(const MftC<T>& tc, const Row<T, S>& tP, const Row<T, S>& tS, const T (&v)[3], const T (&w)[3], const T (&x)[3],
										 const T (&R)[3][3], T (&Fstar)[6], bool export_vectors) {
	using L = MftL<N>;
	// every load first (goals are read-only rows of this kernel, the integrators are read-modify-write), then the arithmetic,
	// then the stores: a load placed after a store through the same arena cannot be hoisted by the compiler and costs a
	// serialised memory round trip (profiles/README.md r1f)
	T xg[3], vd[3], ad[3], Rg[3][3], wd[3], bd[3], ip[3], io[3], oe[3];
	OSC_UNROLL for (int a = 0; a < 3; ++a) {
		xg[a] = tP.ro(L::GOAL_POS + a); vd[a] = tP.ro(L::GOAL_LINVEL + a); ad[a] = tP.ro(L::GOAL_LINACC + a);
		wd[a] = tP.ro(L::GOAL_ANGVEL + a); bd[a] = tP.ro(L::GOAL_ANGACC + a);
		OSC_UNROLL for (int b = 0; b < 3; ++b) Rg[a][b] = tP.ro(L::GOAL_ORI + 3 * a + b);
		ip[a] = tP(L::INT_POS + a); io[a] = tP(L::INT_ORI + a);
	}
	orientation_error<T>(Rg, R, oe);
	OSC_UNROLL for (int a = 0; a < 3; ++a) {
		const T e = x[a] - xg[a];
		ip[a] = ip[a] + e * tc.dt;
		Fstar[a] = ad[a] - tc.kp_pos[a] * e - tc.kv_pos[a] * (v[a] - vd[a]) - tc.ki_pos[a] * ip[a];
		io[a] = io[a] + oe[a] * tc.dt;
		Fstar[3 + a] = bd[a] - tc.kp_ori[a] * oe[a] - tc.kv_ori[a] * (w[a] - wd[a]) - tc.ki_ori[a] * io[a];
	}
	st_vec<T, 3>(tP, L::INT_POS, ip);
	st_vec<T, 3>(tP, L::INT_ORI, io);
	st_vec<T, 3>(tP, L::CUR_POS, x);
	st_mat<T, 3, 3>(tP, L::CUR_ORI, R);
	if (export_vectors) {
		st_vec<T, 3>(tS, L::CUR_V, v);
		st_vec<T, 3>(tS, L::CUR_W, w);
		st_vec<T, 3>(tS, L::DES_POS, xg);
		st_vec<T, 3>(tS, L::DES_LINVEL, vd);
		st_vec<T, 3>(tS, L::ORI_ERR, oe);
		st_mat<T, 3, 3>(tS, L::DES_ORI, Rg);
		st_vec<T, 3>(tS, L::DES_ANGVEL, wd);
		st_vec<T, 6>(tS, L::UMF, Fstar);
	}
}

// same, from the task Jacobian: v = J[0:3] dq, w = J[3:6] dq (MotionForceTask.cpp:293-298)
template <class T, int N, class S = T>
OSC_DEV void mft_control_law(const MftC<T>& tc, const Row<T, S>& tP, const Row<T, S>& tS, const Row<int32_t>& tI, int64_t stride, int64_t r,
							 const T (&J)[6][N], const T (&dq)[N], const T (&x)[3], const T (&R)[3][3], T (&Fstar)[6], T (&Ff)[6], int& status,
							 bool export_vectors = true) {
	T v[3], w[3];
	OSC_UNROLL for (int a = 0; a < 3; ++a) {
		T sv = T(0), sw = T(0);
		OSC_UNROLL for (int j = 0; j < N; ++j) { sv = Num<T>::fma_(J[a][j], dq[j], sv); sw = Num<T>::fma_(J[3 + a][j], dq[j], sw); }
		v[a] = sv; w[a] = sw;
	}
	mft_control_law_vw<T, N>(tc, tP, tS, tI, stride, r, v, w, x, R, Fstar, Ff, status, export_vectors);
}

}  // namespace osc
