// otg_oracle.hpp -- TEST INFRASTRUCTURE (checker only; never linked into or called by the product path).
//
// CPU restatement of the reference's internal online trajectory generation, acceleration-limited
// mode (the default of both tasks: src/tasks/JointTask.h:38-41, src/tasks/MotionForceTask.h:67-72):
//   * Ruckig (vendored by the reference under ruckig/, community version) restricted to what that mode executes:
//     position interface, max_jerk = +inf (second order), Synchronization::Phase, continuous durations, no minimum
//     duration, symmetric limits.  Every function cites the Ruckig file:line it follows.
//   * the two wrappers src/helper_modules/OTG_joints.cpp and OTG_6dof_cartesian.cpp (goal filtering with isApprox, the
//     finished / working / error state machine, the angle-axis re-referencing of the orientation).
//
// PARITY PINNED for this row: the restated Ruckig is checked against the real library compiled from
// /root/reference/ruckig (oracle/_ref/libruckig_ref.so, built by `make -C oracle ref`) on randomised inputs and on the
// second-order known-answer cases of ruckig/test/test-target-known.cpp:262-304 (tests/test_otg_oracle.py), and the
// committed fixtures tests/golden/otg_*.npz were generated with the REAL library underneath the wrappers
// (tools/make_otg_golden.py).  The wrappers themselves need Eigen + the task classes and stay restatements.
//
// Backend switch: OtgBackend::Restated (default) or OtgBackend::Reference (dlopen of libruckig_ref.so) so that the
// whole task-level oracle can run on the real Ruckig.
//
// Defined deviation (SURVEY A.9 numbering continued):
//   Q14  OTG_joints::update, Finished with |v| >= 1e-3 calls setGoalPosition(_goal_position_eigen) with a member that is
//        never assigned (OTG_joints.cpp:131, OTG_joints.h:172): size mismatch -> std::invalid_argument.  Defined here
//        as what OTG_6dof_cartesian.cpp:216 does: keep the target position, zero the target velocity.
#pragma once
#include <dlfcn.h>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <string>

namespace otg {

constexpr int MAXD = 8;
constexpr double kInf = std::numeric_limits<double>::infinity();
constexpr double kEps = std::numeric_limits<double>::epsilon();

enum Result {  // ruckig/include/ruckig/result.hpp:7-18
	Working = 0, Finished = 1, Error = -1, ErrorInvalidInput = -100, ErrorTrajectoryDuration = -101, ErrorZeroLimits = -104,
	ErrorExecutionTimeCalculation = -110, ErrorSynchronizationCalculation = -111
};

// ruckig/include/ruckig/utils.hpp:43-49
inline void integrate(double t, double p0, double v0, double a0, double j, double& p, double& v, double& a) {
	p = p0 + t * (v0 + t * (a0 / 2 + t * j / 6));
	v = v0 + t * (a0 + t * j / 2);
	a = a0 + t * j;
}

// ruckig/include/ruckig/brake.hpp (second-order members only) + ruckig/src/ruckig/brake.cpp:82-102
struct Brake {
	double duration = 0.0, t0 = 0.0, a0 = 0.0, p0 = 0.0, v0 = 0.0;
	void second_order_position(double v0_, double vMax, double vMin, double aMax, double aMin) {
		constexpr double eps = 2.2e-14;  // brake.hpp:18
		t0 = 0.0;
		a0 = 0.0;
		if (aMax == 0.0 || aMin == 0.0) return;
		if (v0_ > vMax) {
			a0 = aMin;
			t0 = (vMax - v0_) / aMin + eps;
		} else if (v0_ < vMin) {
			a0 = aMax;
			t0 = (vMin - v0_) / aMax + eps;
		}
	}
	void finalize_second_order(double& ps, double& vs, double& as) {  // brake.hpp:65-75
		if (t0 <= 0.0) {
			duration = 0.0;
			return;
		}
		duration = t0;
		p0 = ps;
		v0 = vs;
		integrate(t0, ps, vs, a0, 0.0, ps, vs, as);
	}
};

enum Limits { ACC0 = 2, NONE = 7 };  // profile.hpp:55
enum Direction { UP = 0, DOWN = 1 };

// ruckig/include/ruckig/profile.hpp (second-order members)
struct Profile {
	static constexpr double v_eps = 1e-12, a_eps = 1e-12, p_precision = 1e-8, v_precision = 1e-8, t_max = 1e12;  // :34-43
	double t[7] = {0}, t_sum[7] = {0};
	double a[8] = {0}, v[8] = {0}, p[8] = {0};
	Brake brake;
	double pf = 0, vf = 0, af = 0;
	int limits = NONE, direction = UP;

	void set_boundary(const Profile& o) {  // :285-294
		a[0] = o.a[0]; v[0] = o.v[0]; p[0] = o.p[0];
		af = o.af; vf = o.vf; pf = o.pf;
		brake = o.brake;
	}
	void set_boundary(double p0, double v0, double a0, double pf_, double vf_, double af_) {  // :296-303
		a[0] = a0; v[0] = v0; p[0] = p0;
		af = af_; vf = vf_; pf = pf_;
	}
	// check_for_second_order<UDDU, limits>, :307-350
	bool check(int lim, double aUp, double aDown, double vMax, double vMin) {
		if (t[0] < 0) return false;
		t_sum[0] = t[0];
		for (int i = 0; i < 6; ++i) {
			if (t[i + 1] < 0) return false;
			t_sum[i + 1] = t_sum[i] + t[i + 1];
		}
		if (t_sum[6] > t_max) return false;
		a[0] = (t[0] > 0 ? aUp : 0); a[1] = 0; a[2] = (t[2] > 0 ? aDown : 0); a[3] = 0;
		a[4] = (t[4] > 0 ? aDown : 0); a[5] = 0; a[6] = (t[6] > 0 ? aUp : 0); a[7] = af;
		direction = (vMax > 0) ? UP : DOWN;
		const double vUppLim = (direction == UP ? vMax : vMin) + v_eps;
		const double vLowLim = (direction == UP ? vMin : vMax) - v_eps;
		for (int i = 0; i < 7; ++i) {
			v[i + 1] = v[i] + t[i] * a[i];
			p[i + 1] = p[i] + t[i] * (v[i] + t[i] * a[i] / 2);
		}
		limits = lim;
		return std::abs(p[7] - pf) < p_precision && std::abs(v[7] - vf) < v_precision && v[2] <= vUppLim && v[3] <= vUppLim &&
			   v[4] <= vUppLim && v[5] <= vUppLim && v[6] <= vUppLim && v[2] >= vLowLim && v[3] >= vLowLim && v[4] >= vLowLim &&
			   v[5] >= vLowLim && v[6] >= vLowLim;
	}
	// check_for_second_order_with_timing (7-argument form), :358-361
	bool check_timing_limits(int lim, double aUp, double aDown, double vMax, double vMin, double aMax, double aMin) {
		return (aMin - a_eps < aUp) && (aUp < aMax + a_eps) && (aMin - a_eps < aDown) && (aDown < aMax + a_eps) &&
			   check(lim, aUp, aDown, vMax, vMin);
	}
	double total() const { return t_sum[6] + brake.duration; }
};

// ruckig/include/ruckig/block.hpp
struct Interval {
	bool valid = false;
	double left = 0, right = 0;
	Profile profile;
	void set(const Profile& pl, const Profile& pr) {  // :30-43
		const double ld = pl.total(), rd = pr.total();
		valid = true;
		if (ld < rd) { left = ld; right = rd; profile = pr; }
		else { left = rd; right = ld; profile = pl; }
	}
};
struct Block {
	Profile p_min;
	double t_min = 0;
	Interval a, b;
	void set_min_profile(const Profile& p) {  // :46-51
		p_min = p;
		t_min = p.total();
		a.valid = false;
		b.valid = false;
	}
	bool is_blocked(double t) const {  // :133-135
		return (t < t_min) || (a.valid && a.left < t && t < a.right) || (b.valid && b.left < t && t < b.right);
	}
	// calculate_block<N, numerical_robust = true>, :58-131
	static bool calculate(Block& block, Profile* vp, int count) {
		auto remove = [&](int index) {
			for (int i = index; i < count - 1; ++i) vp[i] = vp[i + 1];
			count -= 1;
		};
		if (count == 1) {
			block.set_min_profile(vp[0]);
			return true;
		} else if (count == 2) {
			if (std::abs(vp[0].t_sum[6] - vp[1].t_sum[6]) < 8 * kEps) {
				block.set_min_profile(vp[0]);
				return true;
			}
			const int idx_min = (vp[0].t_sum[6] < vp[1].t_sum[6]) ? 0 : 1;
			const int idx_else = (idx_min + 1) % 2;
			block.set_min_profile(vp[idx_min]);
			block.a.set(vp[idx_min], vp[idx_else]);
			return true;
		} else if (count == 4) {
			if (std::abs(vp[0].t_sum[6] - vp[1].t_sum[6]) < 32 * kEps && vp[0].direction != vp[1].direction) remove(1);
			else if (std::abs(vp[2].t_sum[6] - vp[3].t_sum[6]) < 256 * kEps && vp[2].direction != vp[3].direction) remove(3);
			else if (std::abs(vp[0].t_sum[6] - vp[3].t_sum[6]) < 256 * kEps && vp[0].direction != vp[3].direction) remove(3);
			else return false;
		} else if (count % 2 == 0) {
			return false;
		}
		int idx_min = 0;
		for (int i = 1; i < count; ++i)
			if (vp[i].t_sum[6] < vp[idx_min].t_sum[6]) idx_min = i;
		block.set_min_profile(vp[idx_min]);
		if (count == 3) {
			block.a.set(vp[(idx_min + 1) % 3], vp[(idx_min + 2) % 3]);
			return true;
		} else if (count == 5) {
			const int e1 = (idx_min + 1) % 5, e2 = (idx_min + 2) % 5, e3 = (idx_min + 3) % 5, e4 = (idx_min + 4) % 5;
			if (vp[e1].direction == vp[e2].direction) { block.a.set(vp[e1], vp[e2]); block.b.set(vp[e3], vp[e4]); }
			else { block.a.set(vp[e1], vp[e4]); block.b.set(vp[e2], vp[e3]); }
			return true;
		}
		return false;
	}
};

// ruckig/src/ruckig/position-second-step1.cpp  (extremal profiles of one DoF)
struct Step1 {
	double v0, vf, vMax_, vMin_, aMax_, aMin_, pd;
	// The library declares 3 slots (position.hpp:108) and writes the boundary of slot [count] after every hit; it zeroes
	// t[3..6] of the FIRST slot only (:33-36), later slots keep whatever the stack held (zero in practice, every
	// second-order profile has t[3..6] = 0 except the two UU solutions of step 2).  Here every slot is zeroed.
	Profile valid[6];
	int count = 0;
	Step1(double p0, double v0_, double pf, double vf_, double vMax, double vMin, double aMax, double aMin)
		: v0(v0_), vf(vf_), vMax_(vMax), vMin_(vMin), aMax_(aMax), aMin_(aMin), pd(pf - p0) {}  // :7-9
	void add() {  // position.hpp:117-121
		valid[count + 1].set_boundary(valid[count]);
		++count;
	}
	void time_acc0(double vMax, double vMin, double aMax, double aMin) {  // :11-23
		Profile& pr = valid[count];
		pr.t[0] = (-v0 + vMax) / aMax;
		pr.t[1] = (aMin * v0 * v0 - aMax * vf * vf) / (2 * aMax * aMin * vMax) + vMax * (aMax - aMin) / (2 * aMax * aMin) + pd / vMax;
		pr.t[2] = (vf - vMax) / aMin;
		pr.t[3] = pr.t[4] = pr.t[5] = pr.t[6] = 0;
		if (pr.check(ACC0, aMax, aMin, vMax, vMin)) add();
	}
	void time_none(double vMax, double vMin, double aMax, double aMin, bool return_after_found) {  // :25-58
		double h1 = (aMax * vf * vf - aMin * v0 * v0 - 2 * aMax * aMin * pd) / (aMax - aMin);
		if (h1 >= 0.0) {
			h1 = std::sqrt(h1);
			{
				Profile& pr = valid[count];
				pr.t[3] = pr.t[4] = pr.t[5] = pr.t[6] = 0;
				pr.t[0] = -(v0 + h1) / aMax;
				pr.t[1] = 0;
				pr.t[2] = (vf + h1) / aMin;
				if (pr.check(NONE, aMax, aMin, vMax, vMin)) {
					add();
					if (return_after_found) return;
				}
			}
			{
				Profile& pr = valid[count];
				pr.t[3] = pr.t[4] = pr.t[5] = pr.t[6] = 0;
				pr.t[0] = (-v0 + h1) / aMax;
				pr.t[1] = 0;
				pr.t[2] = (vf - h1) / aMin;
				if (pr.check(NONE, aMax, aMin, vMax, vMin)) add();
			}
		}
	}
	bool get_profile(const Profile& input, Block& block) {  // :92-136 (the zero-limit special case :94-107 cannot occur:
		// the wrappers reject limits <= 0, OTG_joints.cpp:49,63)
		count = 0;
		valid[0].set_boundary(input);
		if (std::abs(vf) < DBL_EPSILON) {
			const double vMax = (pd >= 0) ? vMax_ : vMin_, vMin = (pd >= 0) ? vMin_ : vMax_;
			const double aMax = (pd >= 0) ? aMax_ : aMin_, aMin = (pd >= 0) ? aMin_ : aMax_;
			time_none(vMax, vMin, aMax, aMin, true);
			if (count == 0) time_acc0(vMax, vMin, aMax, aMin);
			if (count == 0) time_none(vMin, vMax, aMin, aMax, true);
			if (count == 0) time_acc0(vMin, vMax, aMin, aMax);
		} else {
			time_none(vMax_, vMin_, aMax_, aMin_, false);
			time_none(vMin_, vMax_, aMin_, aMax_, false);
			time_acc0(vMax_, vMin_, aMax_, aMin_);
			time_acc0(vMin_, vMax_, aMin_, aMax_);
		}
		return Block::calculate(block, valid, count);
	}
};

// ruckig/src/ruckig/position-second-step2.cpp  (profile of one DoF for a given duration tf)
struct Step2 {
	double v0, tf, vf, vMax_, vMin_, aMax_, aMin_, pd, vd;
	Step2(double tf_, double p0, double v0_, double pf, double vf_, double vMax, double vMin, double aMax, double aMin)
		: v0(v0_), tf(tf_), vf(vf_), vMax_(vMax), vMin_(vMin), aMax_(aMax), aMin_(aMin), pd(pf - p0), vd(vf_ - v0_) {}  // :9-12
	bool time_acc0(Profile& pr, double vMax, double vMin, double aMax, double aMin) {  // :14-70
		{
			const double h1 = std::sqrt((2 * aMax * (pd - tf * vf) - 2 * aMin * (pd - tf * v0) + vd * vd) / (aMax * aMin) + tf * tf);
			pr.t[0] = (aMax * vd - aMax * aMin * (tf - h1)) / (aMax * (aMax - aMin));
			pr.t[1] = h1;
			pr.t[2] = tf - (pr.t[0] + h1);
			pr.t[3] = pr.t[4] = pr.t[5] = pr.t[6] = 0;
			if (pr.check(ACC0, aMax, aMin, vMax, vMin)) { pr.pf = pr.p[7]; return true; }
		}
		{
			const double h1 = (-vd + aMax * tf);
			pr.t[0] = -vd * vd / (2 * aMax * h1) + (pd - v0 * tf) / h1;
			pr.t[1] = -vd / aMax + tf;
			pr.t[2] = pr.t[3] = pr.t[4] = pr.t[5] = 0;
			pr.t[6] = tf - (pr.t[0] + pr.t[1]);
			if (pr.check(ACC0, aMax, aMin, vMax, vMin)) { pr.pf = pr.p[7]; return true; }
		}
		{
			pr.t[0] = 0;
			pr.t[1] = -vd / aMax + tf;
			pr.t[2] = pr.t[3] = pr.t[4] = pr.t[5] = 0;
			pr.t[6] = vd / aMax;
			if (pr.check(ACC0, aMax, aMin, vMax, vMin)) { pr.pf = pr.p[7]; return true; }
		}
		return false;
	}
	bool time_none(Profile& pr, double vMax, double vMin, double aMax, double aMin) {  // :72-107
		if (std::abs(v0) < DBL_EPSILON && std::abs(vf) < DBL_EPSILON && std::abs(pd) < DBL_EPSILON) {
			pr.t[0] = 0;
			pr.t[1] = tf;
			pr.t[2] = pr.t[3] = pr.t[4] = pr.t[5] = pr.t[6] = 0;
			if (pr.check(NONE, aMax, aMin, vMax, vMin)) { pr.pf = pr.p[7]; return true; }
		}
		{
			const double h1 = 2 * (vf * tf - pd);
			pr.t[0] = h1 / vd;
			pr.t[1] = tf - pr.t[0];
			pr.t[2] = pr.t[3] = pr.t[4] = pr.t[5] = pr.t[6] = 0;
			const double af = vd * vd / h1;
			if ((aMin - 1e-12 < af) && (af < aMax + 1e-12) && pr.check(NONE, af, -af, vMax, vMin)) { pr.pf = pr.p[7]; return true; }
		}
		return false;
	}
	bool check_all(Profile& pr, double vMax, double vMin, double aMax, double aMin) {  // position.hpp:143-145
		return time_acc0(pr, vMax, vMin, aMax, aMin) || time_none(pr, vMax, vMin, aMax, aMin);
	}
	bool get_profile(Profile& pr) {  // :109-117
		if (pd > 0) return check_all(pr, vMax_, vMin_, aMax_, aMin_) || check_all(pr, vMin_, vMax_, aMin_, aMax_);
		return check_all(pr, vMin_, vMax_, aMin_, aMax_) || check_all(pr, vMax_, vMin_, aMax_, aMin_);
	}
};

struct Input {
	double cur_p[MAXD] = {0}, cur_v[MAXD] = {0}, cur_a[MAXD] = {0}, tgt_p[MAXD] = {0}, tgt_v[MAXD] = {0};
	double max_v[MAXD] = {0}, max_a[MAXD] = {0};
	bool same(const Input& o, int n) const {  // input_parameter.hpp:362-394 (the fields the wrappers can change)
		for (int i = 0; i < n; ++i)
			if (!(cur_p[i] == o.cur_p[i] && cur_v[i] == o.cur_v[i] && cur_a[i] == o.cur_a[i] && tgt_p[i] == o.tgt_p[i] &&
				  tgt_v[i] == o.tgt_v[i] && max_v[i] == o.max_v[i] && max_a[i] == o.max_a[i]))
				return false;
		return true;
	}
};

// ruckig/include/ruckig/calculator_target.hpp (TargetCalculator) + trajectory.hpp + ruckig.hpp: the second-order slice
struct Ruckig2 {
	int n;
	double dt;
	bool initialized = false;  // ruckig.hpp current_input_initialized
	Input current_input;
	Profile prof[MAXD];        // output.trajectory.profiles[0]
	double duration = 0;       // output.trajectory.duration
	double time = 0;           // output.time
	bool new_calculation = false;
	Ruckig2(int n_, double dt_) : n(n_), dt(dt_) {}
	void reset() { initialized = false; }

	// is_input_collinear, calculator_target.hpp:46-119
	bool collinear(const Input& in, int limiting_direction, int limiting_dof, double* new_phase_control) const {
		double pd[MAXD];
		for (int d = 0; d < n; ++d) pd[d] = in.tgt_p[d] - in.cur_p[d];
		const double* scale_vector = nullptr;
		int scale_dof = -1;
		static const double zeros[MAXD] = {0};
		for (int d = 0; d < n; ++d) {
			if (std::abs(pd[d]) > kEps) { scale_vector = pd; scale_dof = d; break; }
			else if (std::abs(in.cur_v[d]) > kEps) { scale_vector = in.cur_v; scale_dof = d; break; }
			else if (std::abs(in.cur_a[d]) > kEps) { scale_vector = in.cur_a; scale_dof = d; break; }
			else if (std::abs(in.tgt_v[d]) > kEps) { scale_vector = in.tgt_v; scale_dof = d; break; }
		}
		if (scale_dof < 0) return false;
		(void)zeros;
		const double scale = scale_vector[scale_dof];
		const double pd_scale = pd[scale_dof] / scale, v0_scale = in.cur_v[scale_dof] / scale, vf_scale = in.tgt_v[scale_dof] / scale,
					 a0_scale = in.cur_a[scale_dof] / scale, af_scale = 0.0 / scale;
		const double scale_limiting = scale_vector[limiting_dof];
		const double control_limiting = (limiting_direction == UP) ? in.max_a[limiting_dof] : -in.max_a[limiting_dof];  // :94-96
		for (int d = 0; d < n; ++d) {
			const double cs = scale_vector[d];
			if (std::abs(pd[d] - pd_scale * cs) > kEps || std::abs(in.cur_v[d] - v0_scale * cs) > kEps ||
				std::abs(in.cur_a[d] - a0_scale * cs) > kEps || std::abs(in.tgt_v[d] - vf_scale * cs) > kEps ||
				std::abs(0.0 - af_scale * cs) > kEps)
				return false;
			new_phase_control[d] = control_limiting * cs / scale_limiting;
		}
		return true;
	}

	// synchronize, calculator_target.hpp:121-205 (t_min = nullopt, continuous durations)
	bool synchronize(const Block* blocks, double& t_sync, int& limiting_dof, Profile* profiles) const {
		double cand[3 * MAXD + 1];
		int idx[3 * MAXD + 1];
		bool any_interval = false;
		for (int d = 0; d < n; ++d) {
			cand[d] = blocks[d].t_min;
			cand[n + d] = blocks[d].a.valid ? blocks[d].a.right : kInf;
			cand[2 * n + d] = blocks[d].b.valid ? blocks[d].b.right : kInf;
			any_interval |= blocks[d].a.valid || blocks[d].b.valid;
		}
		cand[3 * n] = kInf;
		const int cnt = any_interval ? 3 * n + 1 : n;
		for (int i = 0; i < cnt; ++i) idx[i] = i;
		// std::sort in the library (ties in unspecified order); stable here so that both sides of a parity test agree
		std::stable_sort(idx, idx + cnt, [&](int i, int j) { return cand[i] < cand[j]; });
		for (int k = n - 1; k < cnt; ++k) {
			const double t = cand[idx[k]];
			bool blocked = false;
			for (int d = 0; d < n; ++d)
				if (blocks[d].is_blocked(t)) { blocked = true; break; }
			if (blocked || t < 0.0 || std::isinf(t)) continue;
			t_sync = t;
			if (idx[k] == 3 * n) { limiting_dof = -1; return true; }
			limiting_dof = idx[k] % n;
			switch (idx[k] / n) {
				case 0: profiles[limiting_dof] = blocks[limiting_dof].p_min; break;
				case 1: profiles[limiting_dof] = blocks[limiting_dof].a.profile; break;
				case 2: profiles[limiting_dof] = blocks[limiting_dof].b.profile; break;
			}
			return true;
		}
		return false;
	}

	// TargetCalculator::calculate, calculator_target.hpp:227-523.  Works on a scratch copy: the caller keeps the
	// previous trajectory when the result is an error (the wrappers restore the whole output object).
	int calculate(const Input& in, Profile* profiles, double& dur) const {
		Block blocks[MAXD];
		for (int d = 0; d < n; ++d) {
			Profile& p = profiles[d];
			p = Profile();
			const double vMin = -in.max_v[d], aMin = -in.max_a[d];
			p.brake.second_order_position(in.cur_v[d], in.max_v[d], vMin, in.max_a[d], aMin);  // :261-263
			p.set_boundary(in.cur_p[d], in.cur_v[d], in.cur_a[d], in.tgt_p[d], in.tgt_v[d], 0.0);  // :265
			p.brake.finalize_second_order(p.p[0], p.v[0], p.a[0]);  // :283-285
			Step1 s1(p.p[0], p.v[0], p.pf, p.vf, in.max_v[d], vMin, in.max_a[d], aMin);  // :295-297
			if (!s1.get_profile(p, blocks[d])) return ErrorExecutionTimeCalculation;  // :315-333
		}
		if (n == 1) {  // :340-345
			dur = blocks[0].t_min;
			profiles[0] = blocks[0].p_min;
			return Working;
		}
		int limiting = -1;
		if (!synchronize(blocks, dur, limiting, profiles)) return ErrorSynchronizationCalculation;  // :347-374
		if (dur > 7.6e3) return ErrorTrajectoryDuration;  // :388-392
		if (dur == 0.0) {  // :394-400
			for (int d = 0; d < n; ++d) profiles[d] = blocks[d].p_min;
			return Working;
		}
		if (limiting >= 0) {  // phase synchronisation, :406-473
			const Profile& pl = profiles[limiting];
			double npc[MAXD];
			if (collinear(in, pl.direction, limiting, npc)) {
				bool found = true;
				for (int d = 0; d < n; ++d) {
					if (d == limiting) continue;
					Profile& p = profiles[d];
					for (int i = 0; i < 7; ++i) p.t[i] = pl.t[i];
					found &= p.check_timing_limits(NONE, npc[d], -npc[d], in.max_v[d], -in.max_v[d], in.max_a[d], -in.max_a[d]);  // :428
					p.limits = pl.limits;
				}
				if (found) return Working;
			}
		}
		for (int d = 0; d < n; ++d) {  // time synchronisation, :476-520
			if (d == limiting) continue;
			Profile& p = profiles[d];
			const double t_profile = dur - p.brake.duration;
			if (std::abs(t_profile - blocks[d].t_min) < 2 * kEps) { p = blocks[d].p_min; continue; }
			else if (blocks[d].a.valid && std::abs(t_profile - blocks[d].a.right) < 2 * kEps) { p = blocks[d].a.profile; continue; }
			else if (blocks[d].b.valid && std::abs(t_profile - blocks[d].b.right) < 2 * kEps) { p = blocks[d].b.profile; continue; }
			Step2 s2(t_profile, p.p[0], p.v[0], p.pf, p.vf, in.max_v[d], -in.max_v[d], in.max_a[d], -in.max_a[d]);
			if (!s2.get_profile(p)) return ErrorSynchronizationCalculation;
		}
		return Working;
	}

	// Trajectory::state_to_integrate_from + at_time, trajectory.hpp:65-143,182-193 (one section)
	void at_time(double tm, double* np, double* nv, double* na) const {
		for (int d = 0; d < n; ++d) {
			const Profile& p = prof[d];
			if (tm >= duration) {  // :66-76
				const double t_diff = tm - (p.brake.duration + p.t_sum[6]);
				integrate(t_diff, p.p[7], p.v[7], p.a[7], 0.0, np[d], nv[d], na[d]);
				continue;
			}
			double t_diff = tm;
			if (p.brake.duration > 0) {  // :90-102
				if (t_diff < p.brake.duration) {
					integrate(t_diff, p.brake.p0, p.brake.v0, p.brake.a0, 0.0, np[d], nv[d], na[d]);
					continue;
				}
				t_diff -= p.brake.duration;
			}
			if (t_diff >= p.t_sum[6]) {  // :128-132
				integrate(t_diff - p.t_sum[6], p.p[7], p.v[7], p.a[7], 0.0, np[d], nv[d], na[d]);
				continue;
			}
			int index = 0;  // std::upper_bound(t_sum, t_diff), :134-135
			while (index < 7 && !(t_diff < p.t_sum[index])) ++index;
			if (index > 0) t_diff -= p.t_sum[index - 1];
			integrate(t_diff, p.p[index], p.v[index], p.a[index], 0.0, np[d], nv[d], na[d]);
		}
	}

	// Ruckig::update, ruckig.hpp:181-221 (+ validate_input: input_parameter.hpp:155-330, target velocity within limits)
	int update(const Input& in, double* np, double* nv, double* na) {
		new_calculation = false;
		int result = Working;
		if (!in.same(current_input, n) || !initialized) {
			for (int d = 0; d < n; ++d) {
				if (std::isnan(in.cur_p[d]) || std::isnan(in.cur_v[d]) || std::isnan(in.cur_a[d]) || std::isnan(in.tgt_p[d]) ||
					std::isnan(in.tgt_v[d]) || std::isnan(in.max_v[d]) || std::isnan(in.max_a[d]) || in.max_v[d] < 0 || in.max_a[d] < 0)
					return ErrorInvalidInput;
				if (in.tgt_v[d] > in.max_v[d] || in.tgt_v[d] < -in.max_v[d]) return ErrorInvalidInput;  // :289-300
			}
			Profile scratch[MAXD];
			double dur = 0;
			result = calculate(in, scratch, dur);
			if (result != Working) return result;
			for (int d = 0; d < n; ++d) prof[d] = scratch[d];
			duration = dur;
			current_input = in;
			initialized = true;
			time = 0.0;
			new_calculation = true;
		}
		time += dt;
		at_time(time, np, nv, na);
		for (int d = 0; d < n; ++d) {  // output.pass_to_input(current_input), :215
			current_input.cur_p[d] = np[d];
			current_input.cur_v[d] = nv[d];
			current_input.cur_a[d] = na[d];
		}
		if (time > duration) return Finished;
		return result;
	}
};

// ---- backend: restated Ruckig, or the real one from oracle/_ref/libruckig_ref.so ----
struct RefLib {
	void* h = nullptr;
	void* (*create)(int, double) = nullptr;
	void (*destroy)(void*) = nullptr;
	int (*update)(void*, const double*, const double*, const double*, const double*, const double*, const double*, const double*,
				  const double*, double*, double*, double*, double*) = nullptr;
	void (*reset)(void*) = nullptr;
	static RefLib& get() {
		static RefLib r;
		return r;
	}
	bool load(const char* path) {
		if (h) return true;
		h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
		if (!h) return false;
		create = (decltype(create))dlsym(h, "rkref_create");
		destroy = (decltype(destroy))dlsym(h, "rkref_destroy");
		update = (decltype(update))dlsym(h, "rkref_update");
		reset = (decltype(reset))dlsym(h, "rkref_reset");
		return create && destroy && update && reset;
	}
};
inline bool& use_reference_backend() {
	static bool f = false;
	return f;
}

struct Engine {
	int n;
	Ruckig2 restated;
	void* ref = nullptr;
	double last_time = 0, last_duration = 0;
	bool last_new_calculation = false;
	Engine(int n_, double dt) : n(n_), restated(n_, dt) {
		if (use_reference_backend()) ref = RefLib::get().create(n_, dt);
	}
	Engine(const Engine&) = delete;
	~Engine() {
		if (ref) RefLib::get().destroy(ref);
	}
	int update(const Input& in, double* np, double* nv, double* na) {
		if (ref) {
			double info[3], p[MAXD], v[MAXD], a[MAXD];
			const int r = RefLib::get().update(ref, in.cur_p, in.cur_v, in.cur_a, in.tgt_p, in.tgt_v, in.max_v, in.max_a, nullptr, p, v, a, info);
			if (r >= 0)
				for (int d = 0; d < n; ++d) { np[d] = p[d]; nv[d] = v[d]; na[d] = a[d]; }
			last_time = info[0]; last_duration = info[1]; last_new_calculation = info[2] != 0;
			return r;
		}
		double p[MAXD], v[MAXD], a[MAXD];
		const int r = restated.update(in, p, v, a);
		if (r >= 0)  // on an error the wrappers restore the previous output (OTG_joints.cpp:144)
			for (int d = 0; d < n; ++d) { np[d] = p[d]; nv[d] = v[d]; na[d] = a[d]; }
		last_time = restated.time; last_duration = restated.duration; last_new_calculation = restated.new_calculation;
		return r;
	}
};

// Eigen's DenseBase::isApprox for vectors / matrices: |a-b|^2 <= prec^2 min(|a|^2,|b|^2)
inline bool is_approx(const double* a, const double* b, int n, double prec = 1e-12) {
	double d2 = 0, a2 = 0, b2 = 0;
	for (int i = 0; i < n; ++i) { d2 += (a[i] - b[i]) * (a[i] - b[i]); a2 += a[i] * a[i]; b2 += b[i] * b[i]; }
	return d2 <= prec * prec * std::min(a2, b2);
}

// src/helper_modules/OTG_joints.cpp
struct OtgJoints {
	int dim;
	Engine eng;
	Input in;                                            // _input
	double out_p[MAXD] = {0}, out_v[MAXD] = {0}, out_a[MAXD] = {0};  // _output.new_*
	bool goal_reached = false;
	int result_value = Finished;  // OTG_joints.h:170
	int error_count = 0, q14_count = 0;
	OtgJoints(const double* initial_position, int n, double loop_time) : dim(n), eng(n, loop_time) {  // :16-26
		for (int i = 0; i < n; ++i) in.max_v[i] = in.max_a[i] = kInf;
		reInitialize(initial_position);
	}
	void reInitialize(const double* p0) {  // :28-41
		setGoalPositionAndVelocity(p0, nullptr);
		for (int i = 0; i < dim; ++i) { out_p[i] = p0[i]; out_v[i] = 0; out_a[i] = 0; }
		pass_to_input();
	}
	void pass_to_input() {
		for (int i = 0; i < dim; ++i) { in.cur_p[i] = out_p[i]; in.cur_v[i] = out_v[i]; in.cur_a[i] = out_a[i]; }
	}
	void setLimits(const double* max_v, const double* max_a) {  // :43-70
		for (int i = 0; i < dim; ++i) { in.max_v[i] = max_v[i]; in.max_a[i] = max_a[i]; }
	}
	void disableJerkLimits() {  // :86-89
		for (int i = 0; i < dim; ++i) in.cur_a[i] = 0;
	}
	void setGoalPositionAndVelocity(const double* gp, const double* gv) {  // :96-113
		double zero[MAXD] = {0};
		if (!gv) gv = zero;
		if (is_approx(gp, in.tgt_p, dim) && is_approx(gv, in.tgt_v, dim)) return;
		goal_reached = false;
		for (int i = 0; i < dim; ++i) { in.tgt_p[i] = gp[i]; in.tgt_v[i] = gv[i]; }
	}
	void update() {  // :115-150
		if (goal_reached) return;
		result_value = eng.update(in, out_p, out_v, out_a);
		if (result_value == Finished) {
			double nrm = 0;
			for (int i = 0; i < dim; ++i) nrm += out_v[i] * out_v[i];
			if (std::sqrt(nrm) < 1e-3) {
				goal_reached = true;
			} else {  // Q14 (header): keep the target position, zero the target velocity
				++q14_count;
				double tp[MAXD];
				for (int i = 0; i < dim; ++i) tp[i] = in.tgt_p[i];
				setGoalPositionAndVelocity(tp, nullptr);
			}
			return;
		}
		if (result_value == Working) {
			pass_to_input();
			return;
		}
		++error_count;  // the output was kept (Engine::update); :144-149
		for (int i = 0; i < dim; ++i) { in.cur_v[i] = 0; in.cur_a[i] = 0; }
	}
};

// 3x3 helpers (row-major double[9])
inline void m3_mul(const double* A, const double* B, double* C) {
	double t[9];
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j) t[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
	std::memcpy(C, t, sizeof(t));
}
inline void m3_tmul(const double* A, const double* B, double* C) {  // A^T B
	double t[9];
	for (int i = 0; i < 3; ++i)
		for (int j = 0; j < 3; ++j) t[3 * i + j] = A[i] * B[j] + A[3 + i] * B[3 + j] + A[6 + i] * B[6 + j];
	std::memcpy(C, t, sizeof(t));
}
inline void m3_vec(const double* A, const double* v, double* o) {
	double t[3];
	for (int i = 0; i < 3; ++i) t[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
	std::memcpy(o, t, sizeof(t));
}
inline void m3_tvec(const double* A, const double* v, double* o) {
	double t[3];
	for (int i = 0; i < 3; ++i) t[i] = A[i] * v[0] + A[3 + i] * v[1] + A[6 + i] * v[2];
	std::memcpy(o, t, sizeof(t));
}
// Eigen::AngleAxisd(Matrix3d) = AngleAxis(Quaternion(mat)): Eigen/src/Geometry/Quaternion.h (matrix -> quaternion by the
// trace branches) then AngleAxis.h operator=(QuaternionBase): angle = 2 atan2(|vec|, |w|), axis = vec / (+-|vec|)
inline void angle_axis_from_matrix(const double* R, double& angle, double* axis) {
	double q[4];  // x y z w
	double t = R[0] + R[4] + R[8];
	if (t > 0) {
		t = std::sqrt(t + 1.0);
		q[3] = 0.5 * t;
		t = 0.5 / t;
		q[0] = (R[7] - R[5]) * t;
		q[1] = (R[2] - R[6]) * t;
		q[2] = (R[3] - R[1]) * t;
	} else {
		int i = 0;
		if (R[4] > R[0]) i = 1;
		if (R[8] > R[4 * i]) i = 2;
		const int j = (i + 1) % 3, k = (j + 1) % 3;
		t = std::sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
		q[i] = 0.5 * t;
		t = 0.5 / t;
		q[3] = (R[3 * k + j] - R[3 * j + k]) * t;
		q[j] = (R[3 * j + i] + R[3 * i + j]) * t;
		q[k] = (R[3 * k + i] + R[3 * i + k]) * t;
	}
	double nn = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
	if (nn != 0.0) {
		angle = 2.0 * std::atan2(nn, std::abs(q[3]));
		if (q[3] < 0) nn = -nn;
		axis[0] = q[0] / nn; axis[1] = q[1] / nn; axis[2] = q[2] / nn;
	} else {
		angle = 0;
		axis[0] = 1; axis[1] = 0; axis[2] = 0;
	}
}
// Eigen AngleAxis::toRotationMatrix
inline void matrix_from_angle_axis(double angle, const double* ax, double* R) {
	const double s = std::sin(angle), c = std::cos(angle);
	const double sa[3] = {s * ax[0], s * ax[1], s * ax[2]}, ca[3] = {(1 - c) * ax[0], (1 - c) * ax[1], (1 - c) * ax[2]};
	double tmp;
	tmp = ca[0] * ax[1]; R[1] = tmp - sa[2]; R[3] = tmp + sa[2];
	tmp = ca[0] * ax[2]; R[2] = tmp + sa[1]; R[6] = tmp - sa[1];
	tmp = ca[1] * ax[2]; R[5] = tmp - sa[0]; R[7] = tmp + sa[0];
	R[0] = ca[0] * ax[0] + c; R[4] = ca[1] * ax[1] + c; R[8] = ca[2] * ax[2] + c;
}

// src/helper_modules/OTG_6dof_cartesian.cpp
struct OtgCartesian {
	Engine eng;
	Input in;
	double out_p[MAXD] = {0}, out_v[MAXD] = {0}, out_a[MAXD] = {0};
	bool goal_reached = false;
	int result_value = Finished;
	int error_count = 0;
	double ref_frame[9], goal_ori[9] = {0}, goal_angvel[3] = {0};  // _reference_frame, _goal_*_in_base_frame
	OtgCartesian(const double* p0, const double* R0, double loop_time) : eng(6, loop_time) {  // :30-44
		for (int i = 0; i < 6; ++i) in.max_v[i] = in.max_a[i] = kInf;
		std::memcpy(ref_frame, R0, sizeof(ref_frame));
		reInitialize(p0, R0);
	}
	static bool valid_rotation(const double* R) {  // :18-27
		double RtR[9];
		m3_tmul(R, R, RtR);
		double s = 0;
		for (int i = 0; i < 9; ++i) { const double d = RtR[i] - (i % 4 == 0 ? 1.0 : 0.0); s += d * d; }
		if (std::sqrt(s) > 1e-3) return false;
		const double det = R[0] * (R[4] * R[8] - R[5] * R[7]) - R[1] * (R[3] * R[8] - R[5] * R[6]) + R[2] * (R[3] * R[7] - R[4] * R[6]);
		return !(std::abs(det - 1) > 1e-3);
	}
	void pass_to_input() {
		for (int i = 0; i < 6; ++i) { in.cur_p[i] = out_p[i]; in.cur_v[i] = out_v[i]; in.cur_a[i] = out_a[i]; }
	}
	void reInitialize(const double* p0, const double* R0) {  // :46-59
		setGoalPositionAndLinearVelocity(p0, nullptr);
		setGoalOrientationAndAngularVelocity(R0, nullptr);
		for (int i = 0; i < 6; ++i) { in.cur_p[i] = in.tgt_p[i]; in.cur_v[i] = 0; in.cur_a[i] = 0; out_p[i] = in.tgt_p[i]; out_v[i] = 0; out_a[i] = 0; }
	}
	void reInitializeLinear(const double* p0) {  // :61-71
		setGoalPositionAndLinearVelocity(p0, nullptr);
		for (int i = 0; i < 3; ++i) { in.cur_p[i] = in.tgt_p[i]; in.cur_v[i] = 0; in.cur_a[i] = 0; out_p[i] = in.tgt_p[i]; out_v[i] = 0; out_a[i] = 0; }
	}
	void reInitializeAngular(const double* R0) {  // :73-84
		setGoalOrientationAndAngularVelocity(R0, nullptr);
		for (int i = 3; i < 6; ++i) { in.cur_p[i] = in.tgt_p[i]; in.cur_v[i] = 0; in.cur_a[i] = 0; out_p[i] = in.tgt_p[i]; out_v[i] = 0; out_a[i] = 0; }
	}
	void setLimits(double lin_v, double lin_a, double ang_v, double ang_a) {  // :86-129
		for (int i = 0; i < 3; ++i) { in.max_v[i] = lin_v; in.max_a[i] = lin_a; in.max_v[3 + i] = ang_v; in.max_a[3 + i] = ang_a; }
	}
	void setGoalPositionAndLinearVelocity(const double* gp, const double* gv) {  // :146-155
		const double zero[3] = {0, 0, 0};
		if (!gv) gv = zero;
		if (is_approx(gp, in.tgt_p, 3, 1e-3) && is_approx(gv, in.tgt_v, 3, 1e-3)) return;
		goal_reached = false;
		for (int i = 0; i < 3; ++i) { in.tgt_p[i] = gp[i]; in.tgt_v[i] = gv[i]; }
	}
	bool setGoalOrientationAndAngularVelocity(const double* gR, const double* gw) {  // :157-199; false = invalid rotation (throws there)
		const double zero[3] = {0, 0, 0};
		if (!gw) gw = zero;
		if (!valid_rotation(gR)) return false;
		if (is_approx(goal_ori, gR, 9, 1e-3) && is_approx(goal_angvel, gw, 3, 1e-3)) return true;
		goal_reached = false;
		double new_ref[9], R_new_to_prev[9];
		getNextOrientation(new_ref);
		m3_tmul(new_ref, ref_frame, R_new_to_prev);
		std::memcpy(ref_frame, new_ref, sizeof(new_ref));
		std::memcpy(goal_ori, gR, sizeof(goal_ori));
		std::memcpy(goal_angvel, gw, sizeof(goal_angvel));
		out_p[3] = out_p[4] = out_p[5] = 0;
		m3_vec(R_new_to_prev, out_v + 3, out_v + 3);
		m3_vec(R_new_to_prev, out_a + 3, out_a + 3);
		pass_to_input();
		double ref_to_goal[9], angle, axis[3];
		m3_tmul(ref_frame, goal_ori, ref_to_goal);
		angle_axis_from_matrix(ref_to_goal, angle, axis);
		for (int i = 0; i < 3; ++i) in.tgt_p[3 + i] = angle * axis[i];
		m3_tvec(ref_frame, goal_angvel, in.tgt_v + 3);
		return true;
	}
	void update() {  // :201-236
		if (goal_reached) return;
		result_value = eng.update(in, out_p, out_v, out_a);
		if (result_value == Finished) {
			double nrm = 0;
			for (int i = 0; i < 6; ++i) nrm += out_v[i] * out_v[i];
			if (std::sqrt(nrm) < 1e-3) {
				goal_reached = true;
			} else {
				double tp[3] = {in.tgt_p[0], in.tgt_p[1], in.tgt_p[2]}, gR[9];
				setGoalPositionAndLinearVelocity(tp, nullptr);
				std::memcpy(gR, goal_ori, sizeof(gR));
				setGoalOrientationAndAngularVelocity(gR, nullptr);
			}
			return;
		}
		if (result_value == Working) {
			pass_to_input();
			return;
		}
		++error_count;
		for (int i = 0; i < 6; ++i) { in.cur_v[i] = 0; in.cur_a[i] = 0; }
	}
	void getNextOrientation(double* R) const {  // :238-249
		const double nrm = std::sqrt(out_p[3] * out_p[3] + out_p[4] * out_p[4] + out_p[5] * out_p[5]);
		double nxt[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
		if (!(nrm < 1e-3)) {
			const double ax[3] = {out_p[3] / nrm, out_p[4] / nrm, out_p[5] / nrm};
			matrix_from_angle_axis(nrm, ax, nxt);
		}
		m3_mul(ref_frame, nxt, R);
	}
	void getNextAngularVelocity(double* w) const { m3_vec(ref_frame, out_v + 3, w); }      // OTG_6dof_cartesian.h
	void getNextAngularAcceleration(double* w) const { m3_vec(ref_frame, out_a + 3, w); }  // OTG_6dof_cartesian.h
};

}  // namespace otg
