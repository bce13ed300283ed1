// osc_otg.cuh -- batched internal online trajectory generation, acceleration-limited mode.
//
// What it replaces, per robot and per task, once per tick (src/tasks/JointTask.cpp:313-320, MotionForceTask.cpp:394-407):
//     _otg->setGoalPositionAndVelocity(goal, goal_velocity);  _otg->update();  _desired_* = _otg->getNext*();
// with _otg an OTG_joints (src/helper_modules/OTG_joints.cpp) or OTG_6dof_cartesian (OTG_6dof_cartesian.cpp) wrapper
// around ruckig::Ruckig with max_jerk = +inf and Synchronization::Phase (OTG_joints.cpp:23,86-89).
//
// Design (one robot per thread, like every other kernel here; all arithmetic FP64 whatever the batch precision,
// because Ruckig validates its profiles against absolute 1e-8 / 1e-12 tolerances):
//   * the steady state is a STATE ADVANCE: add dt, locate the segment of the stored piecewise-constant-acceleration
//     profile, evaluate p, v, a.  A second-order Ruckig profile has at most five segments per DoF
//     (brake, t0, t1 (coast), t2, t6) - it is stored as 11 doubles per DoF (layout below), not as Ruckig's 7/8-entry
//     arrays, and p[i], v[i] at the segment starts are re-derived in registers by the same recurrence Ruckig uses;
//   * a recalculation happens only for the robots whose target changed this tick (divergent, rare): extremal profiles
//     per DoF (position-second-step1.cpp), blocked intervals (block.hpp), synchronised duration
//     (calculator_target.hpp:121-205), phase synchronisation when the boundary conditions are collinear (:46-119,
//     :406-473), else per-DoF time synchronisation (position-second-step2.cpp);
//   * Ruckig's `input != current_input` test (ruckig.hpp:193) is carried as a DIRTY flag that every path that edits the
//     input without Ruckig seeing it sets (target change, error recovery, re-initialisation, orientation re-reference);
//   * the wrapper's _input.current_* equals _output.new_* except after a Finished / error tick: the IN rows are only
//     materialised then (IN_SYNCED flag), which saves 3D stores per robot-tick;
//   * outputs are written straight into the rows the control law reads (the task's P.GOAL_* rows = "_desired_*"); while
//     the OTG is enabled the user's goals live in the G arena (same offsets), see OpsImpl::locate.
// Jerk-limited mode (third order, position-third-step{1,2}.cpp, 1.7 kLoC of root finding) is not built: the enabling
// entry point returns OSC_ERR_UNSUPPORTED for it.
#pragma once
#include <math_constants.h>

#include "osc_kernels.cuh"

namespace osc {

// ---- O arena (double rows) and OI arena (int rows) of one OTG instance of dimension D ----
template <int D> struct OtgL {
	static constexpr int TGT_P = 0, TGT_V = D, IN_P = 2 * D, IN_V = 3 * D, IN_A = 4 * D, OUT_P = 5 * D, OUT_V = 6 * D, OUT_A = 7 * D,
						 TRAJ = 8 * D,  // 11 per DoF: brake t, brake a, p start, v start, t0, t1, t2, t6, a0, a2, a6
						 TIME = 19 * D, DUR = TIME + 1,
						 REF = DUR + 1, GORI = REF + 9, GANG = GORI + 9,  // cartesian only: _reference_frame, _goal_*_in_base_frame
						 // current segment of every DoF, 6 per DoF (row e of DoF d at SEG + e * D + d): time offset (brake duration
						 // once past it), segment start / end in that clock, p, v at the start, acceleration.  end < 0 = stale.
						 SEG = GANG + 3, O_SIZE = SEG + 6 * D;
	static constexpr int FLAGS = 0, RESULT = 1, ERRORS = 2, I_SIZE = 3;
};
// OUT_LIVE: the OUT rows are stale, the current output is the stored trajectory evaluated at TIME (the steady-state
// advance does not store what it can re-derive)
enum : int { OTG_REACHED = 1, OTG_INIT = 2, OTG_DIRTY = 4, OTG_IN_SYNCED = 8, OTG_OUT_LIVE = 16 };
enum : int { OTG_WORKING = 0, OTG_FINISHED = 1, OTG_ERR_INPUT = -100, OTG_ERR_DURATION = -101, OTG_ERR_STEP1 = -110, OTG_ERR_SYNC = -111 };

struct OtgJob {
	double* O;
	int32_t* OI;
	void* G;          // user goals (T rows, same offsets as the head of the task's P arena)
	void* P;          // task P arena: GOAL_* rows receive the desired state
	int32_t* status;  // robot status row
	const int32_t* rank;  // joint task: _current_task_range columns (the reference returns before the OTG when 0, JointTask.cpp:302-306)
	int k;            // joints: task dof
	double dt;
	double max_v[8], max_a[8];
	int check_goals;  // 0: no goal field was written since the last tick (host knowledge): the goal filter is skipped
	int* slow_count;  // robots that need the general path this tick (compacted list, filled by k_otg_advance)
	int* slow_next;   // the counter of the next tick (zeroed by k_otg_slow)
	int* slow_list;
};

// ---- one DoF profile, compact ----
struct OtgProf {
	double brT, brA, pS, vS;        // brake segment and the state it starts from
	double t0, t1, t2, t6, a0, a2, a6;
	double p0, v0, pf, vf;          // boundary after the brake segment; target
	int up;                         // Profile::direction == UP
	__device__ double tsum() const { return ((t0 + t1) + t2) + t6; }  // t3..t5 = 0 (profile.hpp:313-319)
	__device__ double total() const { return tsum() + brT; }
};

// Profile::check_for_second_order<UDDU, *> (ruckig/include/ruckig/profile.hpp:307-350) on the compact form.
// Comparisons keep the library's orientation so that NaN times behave the same (NaN < 0 is false, |NaN| < eps is false).
__device__ inline bool otg_check(OtgProf& p, double aUp, double aDown, double vMax, double vMin) {
	if (p.t0 < 0 || p.t1 < 0 || p.t2 < 0 || p.t6 < 0) return false;
	const double ts = p.tsum();
	if (ts > 1e12) return false;
	p.a0 = p.t0 > 0 ? aUp : 0.0;
	p.a2 = p.t2 > 0 ? aDown : 0.0;
	p.a6 = p.t6 > 0 ? aUp : 0.0;
	p.up = vMax > 0;
	const double vUpp = (p.up ? vMax : vMin) + 1e-12, vLow = (p.up ? vMin : vMax) - 1e-12;
	const double v1 = p.v0 + p.t0 * p.a0, p1 = p.p0 + p.t0 * (p.v0 + p.t0 * p.a0 / 2);
	const double v2 = v1 + p.t1 * 0.0, p2 = p1 + p.t1 * (v1 + p.t1 * 0.0 / 2);
	const double v3 = v2 + p.t2 * p.a2, p3 = p2 + p.t2 * (v2 + p.t2 * p.a2 / 2);
	const double v7 = v3 + p.t6 * p.a6, p7 = p3 + p.t6 * (v3 + p.t6 * p.a6 / 2);
	return fabs(p7 - p.pf) < 1e-8 && fabs(v7 - p.vf) < 1e-8 && v2 <= vUpp && v3 <= vUpp && v2 >= vLow && v3 >= vLow;
}

// what step 1 decides for a DoF: the boundary / brake part of a candidate is the caller's `base`, so a block keeps only this
// (64 B instead of a 128 B profile: the blocks of all DoF live in local memory, which is what bounds k_otg_slow)
struct OtgTimes {
	double t0, t1, t2, t6, a0, a2, a6;
	int up;
	__device__ double tsum() const { return ((t0 + t1) + t2) + t6; }
};
__device__ inline OtgTimes otg_times(const OtgProf& p) { return OtgTimes{p.t0, p.t1, p.t2, p.t6, p.a0, p.a2, p.a6, p.up}; }
__device__ inline void otg_apply(OtgProf& p, const OtgTimes& t) {
	p.t0 = t.t0; p.t1 = t.t1; p.t2 = t.t2; p.t6 = t.t6; p.a0 = t.a0; p.a2 = t.a2; p.a6 = t.a6; p.up = t.up;
}
struct OtgBlock {  // ruckig/include/ruckig/block.hpp: fastest profile + at most one blocked interval in second order
	OtgTimes pmin, pa;
	double tmin, aL, aR;
	bool hasA;
	__device__ bool blocked(double t) const { return (t < tmin) || (hasA && aL < t && t < aR); }
};

// PositionSecondOrderStep1::get_profile (ruckig/src/ruckig/position-second-step1.cpp:92-136) + Block::calculate_block
// (block.hpp:58-131).  `base` carries the boundary and the brake segment.
static __device__ __noinline__ bool otg_step1(const OtgProf& base, double vMax_, double aMax_, OtgBlock& blk) {
	const double vMin_ = -vMax_, aMin_ = -aMax_;
	const double v0 = base.v0, vf = base.vf, pd = base.pf - base.p0;
	OtgTimes cand[4];
	int n = 0;
	auto try_none = [&](double vMax, double vMin, double aMax, double aMin, bool first_only) {  // :25-58
		double h1 = (aMax * vf * vf - aMin * v0 * v0 - 2 * aMax * aMin * pd) / (aMax - aMin);
		if (!(h1 >= 0.0)) return;
		h1 = sqrt(h1);
		for (int s = 0; s < 2 && n < 4; ++s) {
			OtgProf c = base;
			c.t1 = 0; c.t6 = 0;
			c.t0 = s == 0 ? -(v0 + h1) / aMax : (-v0 + h1) / aMax;
			c.t2 = s == 0 ? (vf + h1) / aMin : (vf - h1) / aMin;
			if (otg_check(c, aMax, aMin, vMax, vMin)) {
				cand[n++] = otg_times(c);
				if (first_only) return;
			}
		}
	};
	auto try_acc0 = [&](double vMax, double vMin, double aMax, double aMin) {  // :11-23
		if (n >= 4) return;
		OtgProf c = base;
		c.t0 = (-v0 + vMax) / aMax;
		c.t1 = (aMin * v0 * v0 - aMax * vf * vf) / (2 * aMax * aMin * vMax) + vMax * (aMax - aMin) / (2 * aMax * aMin) + pd / vMax;
		c.t2 = (vf - vMax) / aMin;
		c.t6 = 0;
		if (otg_check(c, aMax, aMin, vMax, vMin)) cand[n++] = otg_times(c);
	};
	if (fabs(vf) < 2.220446049250313e-16) {  // no blocked interval when vf == 0: first hit wins (:111-124)
		const bool fw = pd >= 0;
		const double vMax = fw ? vMax_ : vMin_, vMin = fw ? vMin_ : vMax_, aMax = fw ? aMax_ : aMin_, aMin = fw ? aMin_ : aMax_;
		try_none(vMax, vMin, aMax, aMin, true);
		if (n == 0) try_acc0(vMax, vMin, aMax, aMin);
		if (n == 0) try_none(vMin, vMax, aMin, aMax, true);
		if (n == 0) try_acc0(vMin, vMax, aMin, aMax);
	} else {  // :126-131
		try_none(vMax_, vMin_, aMax_, aMin_, false);
		try_none(vMin_, vMax_, aMin_, aMax_, false);
		try_acc0(vMax_, vMin_, aMax_, aMin_);
		try_acc0(vMin_, vMax_, aMin_, aMax_);
	}
	// block.hpp:58-131
	const double eps = 2.220446049250313e-16;
	blk.hasA = false;
	auto set_min = [&](const OtgTimes& p) { blk.pmin = p; blk.tmin = p.tsum() + base.brT; };  // Profile total = times + brake
	auto set_a = [&](const OtgTimes& l, const OtgTimes& r) {  // Interval(profile_left, profile_right), :30-43
		const double ld = l.tsum() + base.brT, rd = r.tsum() + base.brT;
		blk.hasA = true;
		if (ld < rd) { blk.aL = ld; blk.aR = rd; blk.pa = r; }
		else { blk.aL = rd; blk.aR = ld; blk.pa = l; }
	};
	if (n == 1) { set_min(cand[0]); return true; }
	if (n == 2) {
		if (fabs(cand[0].tsum() - cand[1].tsum()) < 8 * eps) { set_min(cand[0]); return true; }
		const int im = cand[0].tsum() < cand[1].tsum() ? 0 : 1;
		set_min(cand[im]);
		set_a(cand[im], cand[im ^ 1]);
		return true;
	}
	if (n == 4) {  // "only happens due to numerical issues": drop one of two identical profiles of opposite direction
		int drop = -1;
		if (fabs(cand[0].tsum() - cand[1].tsum()) < 32 * eps && cand[0].up != cand[1].up) drop = 1;
		else if (fabs(cand[2].tsum() - cand[3].tsum()) < 256 * eps && cand[2].up != cand[3].up) drop = 3;
		else if (fabs(cand[0].tsum() - cand[3].tsum()) < 256 * eps && cand[0].up != cand[3].up) drop = 3;
		else return false;
		for (int i = drop; i < 3; ++i) cand[i] = cand[i + 1];
		n = 3;
	}
	if (n != 3) return false;
	int im = 0;
	for (int i = 1; i < 3; ++i)
		if (cand[i].tsum() < cand[im].tsum()) im = i;
	set_min(cand[im]);
	set_a(cand[(im + 1) % 3], cand[(im + 2) % 3]);
	return true;
}

// PositionSecondOrderStep2::get_profile (ruckig/src/ruckig/position-second-step2.cpp:109-117): profile of duration tf
static __device__ __noinline__ bool otg_step2(OtgProf& p, double tf, double vMax_, double aMax_) {
	const double v0 = p.v0, vf = p.vf, pd = p.pf - p.p0, vd = p.vf - p.v0;
	auto all = [&](double vMax, double vMin, double aMax, double aMin) -> bool {
		{  // time_acc0, UD, :16-31
			const double h1 = sqrt((2 * aMax * (pd - tf * vf) - 2 * aMin * (pd - tf * v0) + vd * vd) / (aMax * aMin) + tf * tf);
			p.t0 = (aMax * vd - aMax * aMin * (tf - h1)) / (aMax * (aMax - aMin));
			p.t1 = h1;
			p.t2 = tf - (p.t0 + h1);
			p.t6 = 0;
			if (otg_check(p, aMax, aMin, vMax, vMin)) return true;
		}
		{  // UU, :34-50
			const double h1 = (-vd + aMax * tf);
			p.t0 = -vd * vd / (2 * aMax * h1) + (pd - v0 * tf) / h1;
			p.t1 = -vd / aMax + tf;
			p.t2 = 0;
			p.t6 = tf - (p.t0 + p.t1);
			if (otg_check(p, aMax, aMin, vMax, vMin)) return true;
		}
		{  // UU, two steps, :53-67
			p.t0 = 0;
			p.t1 = -vd / aMax + tf;
			p.t2 = 0;
			p.t6 = vd / aMax;
			if (otg_check(p, aMax, aMin, vMax, vMin)) return true;
		}
		const double eps = 2.220446049250313e-16;
		if (fabs(v0) < eps && fabs(vf) < eps && fabs(pd) < eps) {  // time_none, :73-86
			p.t0 = 0; p.t1 = tf; p.t2 = 0; p.t6 = 0;
			if (otg_check(p, aMax, aMin, vMax, vMin)) return true;
		}
		{  // :89-104
			const double h1 = 2 * (vf * tf - pd);
			p.t0 = h1 / vd;
			p.t1 = tf - p.t0;
			p.t2 = 0; p.t6 = 0;
			const double af = vd * vd / h1;
			if ((aMin - 1e-12 < af) && (af < aMax + 1e-12) && otg_check(p, af, -af, vMax, vMin)) return true;
		}
		return false;
	};
	const double vMin_ = -vMax_, aMin_ = -aMax_;
	if (pd > 0) return all(vMax_, vMin_, aMax_, aMin_) || all(vMin_, vMax_, aMin_, aMax_);
	return all(vMin_, vMax_, aMin_, aMax_) || all(vMax_, vMin_, aMax_, aMin_);
}

// TargetCalculator::calculate for the second-order position interface (calculator_target.hpp:227-523).
// in: current p, v, a; target p, v; limits.  out: one profile per DoF + the synchronised duration.
// __noinline__ (also otg_step1 / otg_step2): this is the rare, divergent path; kept out of line it costs the steady-state
// advance no registers, and the fully inlined 255-register build of nvcc 12.9 overlapped the caller's cp[] with prof[] in
// the local frame for D = 6 (caught by tests/test_otg_gpu.py against the golden vectors; the out-of-line build is correct).
template <int D>
__device__ __noinline__ int otg_calculate(int n, const double* cp, const double* cv, const double* ca, const double* tp, const double* tv, const double* mv,
							 const double* ma, OtgProf* prof, double& duration) {
	const double eps = 2.220446049250313e-16;
	OtgBlock blk[D];
	for (int d = 0; d < n; ++d) {
		OtgProf b{};
		b.pS = cp[d]; b.vS = cv[d];
		b.p0 = cp[d]; b.v0 = cv[d]; b.pf = tp[d]; b.vf = tv[d];
		if (cv[d] > mv[d]) {  // brake.cpp:82-102 (eps = 2.2e-14, brake.hpp:18), brake.hpp:65-75
			b.brA = -ma[d];
			b.brT = (mv[d] - cv[d]) / -ma[d] + 2.2e-14;
		} else if (cv[d] < -mv[d]) {
			b.brA = ma[d];
			b.brT = (-mv[d] - cv[d]) / ma[d] + 2.2e-14;
		}
		if (b.brT > 0.0) {
			b.p0 = b.pS + b.brT * (b.vS + b.brT * (b.brA / 2 + b.brT * 0.0 / 6));
			b.v0 = b.vS + b.brT * (b.brA + b.brT * 0.0 / 2);
		} else {
			b.brT = 0.0;
		}
		prof[d] = b;
		if (!otg_step1(b, mv[d], ma[d], blk[d])) return OTG_ERR_STEP1;
	}
	if (n == 1) {  // :340-345
		duration = blk[0].tmin;
		otg_apply(prof[0], blk[0].pmin);
		return OTG_WORKING;
	}
	// synchronize, :121-205: candidates t_min[d], a.right[d] in ascending order (ties by index), from the (n-1)-th on
	int limiting = -1;
	{
		bool any = false;
		for (int d = 0; d < n; ++d) any |= blk[d].hasA;
		const int cnt = any ? 2 * n : n;  // the b intervals and the optional minimum duration are +inf here
		auto val = [&](int i) -> double { return i < n ? blk[i].tmin : (blk[i - n].hasA ? blk[i - n].aR : CUDART_INF); };
		double lastv = -CUDART_INF;
		int lasti = -1;
		bool found = false;
		for (int rank = 0; rank < cnt && !found; ++rank) {
			double bv = CUDART_INF;
			int bi = -1;
			for (int i = 0; i < cnt; ++i) {  // next (value, index) pair after (lastv, lasti)
				const double v = val(i);
				const bool after = (v > lastv) || (v == lastv && i > lasti);
				if (after && (bi < 0 || v < bv)) { bv = v; bi = i; }
			}
			if (bi < 0) break;
			lastv = bv; lasti = bi;
			if (rank < n - 1) continue;
			bool blocked = false;
			for (int d = 0; d < n; ++d) blocked |= blk[d].blocked(bv);
			if (blocked || bv < 0.0 || isinf(bv)) continue;
			duration = bv;
			limiting = bi % n;
			otg_apply(prof[limiting], bi < n ? blk[limiting].pmin : blk[limiting].pa);
			found = true;
		}
		if (!found) return OTG_ERR_SYNC;
	}
	if (duration > 7.6e3) return OTG_ERR_DURATION;  // :388-392
	if (duration == 0.0) {                          // :394-400
		for (int d = 0; d < n; ++d) otg_apply(prof[d], blk[d].pmin);
		return OTG_WORKING;
	}
	{  // phase synchronisation, :46-119 and :406-473
		const OtgProf pl = prof[limiting];
		int sd = -1, which = 0;  // scale vector: 0 pd, 1 v0, 2 a0, 3 vf
		for (int d = 0; d < n && sd < 0; ++d) {
			if (fabs(tp[d] - cp[d]) > eps) { sd = d; which = 0; }
			else if (fabs(cv[d]) > eps) { sd = d; which = 1; }
			else if (fabs(ca[d]) > eps) { sd = d; which = 2; }
			else if (fabs(tv[d]) > eps) { sd = d; which = 3; }
		}
		if (sd >= 0) {
			auto sv = [&](int d) -> double { return which == 0 ? tp[d] - cp[d] : (which == 1 ? cv[d] : (which == 2 ? ca[d] : tv[d])); };
			const double scale = sv(sd);
			const double pd_s = (tp[sd] - cp[sd]) / scale, v0_s = cv[sd] / scale, vf_s = tv[sd] / scale, a0_s = ca[sd] / scale;
			const double scale_lim = sv(limiting);
			const double ctrl_lim = pl.up ? ma[limiting] : -ma[limiting];
			bool collinear = true;
			for (int d = 0; d < n; ++d) {
				const double cs = sv(d);
				if (fabs((tp[d] - cp[d]) - pd_s * cs) > eps || fabs(cv[d] - v0_s * cs) > eps || fabs(ca[d] - a0_s * cs) > eps ||
					fabs(tv[d] - vf_s * cs) > eps)
					collinear = false;
			}
			if (collinear) {
				bool ok = true;
				for (int d = 0; d < n; ++d) {
					if (d == limiting) continue;
					const double c = ctrl_lim * sv(d) / scale_lim;  // new_phase_control
					OtgProf& p = prof[d];
					p.t0 = pl.t0; p.t1 = pl.t1; p.t2 = pl.t2; p.t6 = pl.t6;
					const bool lim_ok = (-ma[d] - 1e-12 < c) && (c < ma[d] + 1e-12) && (-ma[d] - 1e-12 < -c) && (-c < ma[d] + 1e-12);
					ok &= lim_ok && otg_check(p, c, -c, mv[d], -mv[d]);
				}
				if (ok) return OTG_WORKING;
			}
		}
	}
	for (int d = 0; d < n; ++d) {  // time synchronisation, :476-520
		if (d == limiting) continue;
		OtgProf& p = prof[d];
		const double tprof = duration - p.brT;
		if (fabs(tprof - blk[d].tmin) < 2 * eps) { otg_apply(p, blk[d].pmin); continue; }
		if (blk[d].hasA && fabs(tprof - blk[d].aR) < 2 * eps) { otg_apply(p, blk[d].pa); continue; }
		if (!otg_step2(p, tprof, mv[d], ma[d])) return OTG_ERR_SYNC;
	}
	return OTG_WORKING;
}

// Trajectory::at_time for one DoF (ruckig/include/ruckig/trajectory.hpp:65-143): state at time tm of a profile
__device__ inline void otg_eval(const OtgProf& p, double tm, double duration, double& np, double& nv, double& na) {
	const double v1 = p.v0 + p.t0 * p.a0, p1 = p.p0 + p.t0 * (p.v0 + p.t0 * p.a0 / 2);
	const double p2 = p1 + p.t1 * (v1 + p.t1 * 0.0 / 2);
	const double v3 = v1 + p.t2 * p.a2, p3 = p2 + p.t2 * (v1 + p.t2 * p.a2 / 2);
	const double v7 = v3 + p.t6 * p.a6, p7 = p3 + p.t6 * (v3 + p.t6 * p.a6 / 2);
	const double s0 = p.t0, s1 = s0 + p.t1, s2 = s1 + p.t2, s6 = s2 + p.t6;
	double t, bp, bv, ba;
	if (tm >= duration) {  // :66-76
		t = tm - (p.brT + s6); bp = p7; bv = v7; ba = 0.0;
	} else if (p.brT > 0 && tm < p.brT) {  // :90-99
		t = tm; bp = p.pS; bv = p.vS; ba = p.brA;
	} else {
		const double td = p.brT > 0 ? tm - p.brT : tm;
		if (td >= s6) { t = td - s6; bp = p7; bv = v7; ba = 0.0; }  // :128-132
		else if (td < s0) { t = td; bp = p.p0; bv = p.v0; ba = p.a0; }
		else if (td < s1) { t = td - s0; bp = p1; bv = v1; ba = 0.0; }
		else if (td < s2) { t = td - s1; bp = p2; bv = v1; ba = p.a2; }
		else { t = td - s2; bp = p3; bv = v3; ba = p.a6; }
	}
	np = bp + t * (bv + t * (ba / 2 + t * 0.0 / 6));  // utils.hpp:43-49
	nv = bv + t * (ba + t * 0.0 / 2);
	na = ba + t * 0.0;
}

// The segment of a profile that contains time tm (< duration), in the form the advance kernel caches: the same branches
// and the same subtraction order as otg_eval, so that evaluating the cached segment reproduces otg_eval bit for bit.
struct OtgSeg { double off, ts, te, p, v, a; };
__device__ inline OtgSeg otg_segment(const OtgProf& p, double tm) {
	const double v1 = p.v0 + p.t0 * p.a0, p1 = p.p0 + p.t0 * (p.v0 + p.t0 * p.a0 / 2);
	const double p2 = p1 + p.t1 * (v1 + p.t1 * 0.0 / 2);
	const double v3 = v1 + p.t2 * p.a2, p3 = p2 + p.t2 * (v1 + p.t2 * p.a2 / 2);
	const double v7 = v3 + p.t6 * p.a6, p7 = p3 + p.t6 * (v3 + p.t6 * p.a6 / 2);
	const double s0 = p.t0, s1 = s0 + p.t1, s2 = s1 + p.t2, s6 = s2 + p.t6;
	if (p.brT > 0 && tm < p.brT) return OtgSeg{0.0, 0.0, p.brT, p.pS, p.vS, p.brA};
	const double off = p.brT > 0 ? p.brT : 0.0;
	const double td = p.brT > 0 ? tm - p.brT : tm;
	if (td >= s6) return OtgSeg{off, s6, CUDART_INF, p7, v7, 0.0};
	if (td < s0) return OtgSeg{off, 0.0, s0, p.p0, p.v0, p.a0};
	if (td < s1) return OtgSeg{off, s0, s1, p1, v1, 0.0};
	if (td < s2) return OtgSeg{off, s1, s2, p2, v1, p.a2};
	return OtgSeg{off, s2, s6, p3, v3, p.a6};
}

// Eigen DenseBase::isApprox: |a-b|^2 <= prec^2 min(|a|^2, |b|^2)
template <int K> __device__ inline bool otg_approx(const double* a, const double* b, int k, double prec) {
	double d2 = 0, a2 = 0, b2 = 0;
	OSC_UNROLL for (int i = 0; i < K; ++i)
		if (i < k) { d2 += (a[i] - b[i]) * (a[i] - b[i]); a2 += a[i] * a[i]; b2 += b[i] * b[i]; }
	return d2 <= prec * prec * fmin(a2, b2);
}

// rotation helpers (row-major 3x3 in double[9])
__device__ inline void otg_m3mul(const double* A, const double* B, double* C) {
	double t[9];
	OSC_UNROLL for (int i = 0; i < 3; ++i)
		OSC_UNROLL for (int j = 0; j < 3; ++j) t[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
	OSC_UNROLL for (int i = 0; i < 9; ++i) C[i] = t[i];
}
__device__ inline void otg_m3tmul(const double* A, const double* B, double* C) {
	double t[9];
	OSC_UNROLL for (int i = 0; i < 3; ++i)
		OSC_UNROLL for (int j = 0; j < 3; ++j) t[3 * i + j] = A[i] * B[j] + A[3 + i] * B[3 + j] + A[6 + i] * B[6 + j];
	OSC_UNROLL for (int i = 0; i < 9; ++i) C[i] = t[i];
}
__device__ inline void otg_m3v(const double* A, const double* v, double* o) {
	const double x = v[0], y = v[1], z = v[2];
	OSC_UNROLL for (int i = 0; i < 3; ++i) o[i] = A[3 * i] * x + A[3 * i + 1] * y + A[3 * i + 2] * z;
}
__device__ inline void otg_m3tv(const double* A, const double* v, double* o) {
	const double x = v[0], y = v[1], z = v[2];
	OSC_UNROLL for (int i = 0; i < 3; ++i) o[i] = A[i] * x + A[3 + i] * y + A[6 + i] * z;
}
// getNextOrientation (OTG_6dof_cartesian.cpp:238-249): ref * AngleAxis(|w|, w/|w|), identity below 1e-3
__device__ inline void otg_next_orientation(const double* ref, const double* w, double* R) {
	const double nrm = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
	if (nrm < 1e-3) {
		OSC_UNROLL for (int i = 0; i < 9; ++i) R[i] = ref[i];
		return;
	}
	const double ax[3] = {w[0] / nrm, w[1] / nrm, w[2] / nrm};
	double s, c;
	sincos(nrm, &s, &c);
	const double k = 1.0 - c;
	double E[9];
	E[0] = k * ax[0] * ax[0] + c;         E[1] = k * ax[0] * ax[1] - s * ax[2]; E[2] = k * ax[0] * ax[2] + s * ax[1];
	E[3] = k * ax[0] * ax[1] + s * ax[2]; E[4] = k * ax[1] * ax[1] + c;         E[5] = k * ax[1] * ax[2] - s * ax[0];
	E[6] = k * ax[0] * ax[2] - s * ax[1]; E[7] = k * ax[1] * ax[2] + s * ax[0]; E[8] = k * ax[2] * ax[2] + c;
	otg_m3mul(ref, E, R);
}
// angle * axis of a rotation matrix the way Eigen::AngleAxisd(Matrix3d) computes it (via the quaternion)
__device__ inline void otg_rotvec(const double* R, double* w) {
	double q[4];
	double t = R[0] + R[4] + R[8];
	if (t > 0) {
		t = sqrt(t + 1.0);
		q[3] = 0.5 * t;
		t = 0.5 / t;
		q[0] = (R[7] - R[5]) * t; q[1] = (R[2] - R[6]) * t; q[2] = (R[3] - R[1]) * t;
	} else {
		int i = 0;
		if (R[4] > R[0]) i = 1;
		if (R[8] > R[4 * i]) i = 2;
		const int j = (i + 1) % 3, k = (j + 1) % 3;
		t = sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
		q[i] = 0.5 * t;
		t = 0.5 / t;
		q[3] = (R[3 * k + j] - R[3 * j + k]) * t;
		q[j] = (R[3 * j + i] + R[3 * i + j]) * t;
		q[k] = (R[3 * k + i] + R[3 * i + k]) * t;
	}
	double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
	if (n != 0.0) {
		const double angle = 2.0 * atan2(n, fabs(q[3]));
		if (q[3] < 0) n = -n;
		w[0] = angle * (q[0] / n); w[1] = angle * (q[1] / n); w[2] = angle * (q[2] / n);
	} else {
		w[0] = w[1] = w[2] = 0.0;
	}
}
__device__ inline bool otg_valid_rotation(const double* R) {  // OTG_6dof_cartesian.cpp:18-27
	double RtR[9];
	otg_m3tmul(R, R, RtR);
	double s = 0;
	OSC_UNROLL for (int i = 0; i < 9; ++i) { const double d = RtR[i] - (i % 4 == 0 ? 1.0 : 0.0); s += d * d; }
	if (sqrt(s) > 1e-3) return false;
	const double det = R[0] * (R[4] * R[8] - R[5] * R[7]) - R[1] * (R[3] * R[8] - R[5] * R[6]) + R[2] * (R[3] * R[7] - R[4] * R[6]);
	return !(fabs(det - 1) > 1e-3);
}

// stored profile of DoF d (11 doubles) -> registers; p0, v0 re-derived from the brake segment like brake.hpp:65-75
template <int D> __device__ inline void otg_load_prof(const Row<double>& O, int d, OtgProf& p) {
	const int o = OtgL<D>::TRAJ + 11 * d;
	p.brT = O(o); p.brA = O(o + 1); p.pS = O(o + 2); p.vS = O(o + 3); p.t0 = O(o + 4); p.t1 = O(o + 5); p.t2 = O(o + 6);
	p.t6 = O(o + 7); p.a0 = O(o + 8); p.a2 = O(o + 9); p.a6 = O(o + 10);
	if (p.brT > 0.0) {
		p.p0 = p.pS + p.brT * (p.vS + p.brT * (p.brA / 2 + p.brT * 0.0 / 6));
		p.v0 = p.vS + p.brT * (p.brA + p.brT * 0.0 / 2);
	} else { p.p0 = p.pS; p.v0 = p.vS; }
}
// _output.new_* of a robot: the OUT rows, or the stored trajectory at TIME when the advance kernel skipped the stores
template <int D> __device__ inline void otg_current_output(const Row<double>& O, int flags, int k, double* out_p, double* out_v, double* out_a) {
	using L = OtgL<D>;
	if (flags & OTG_OUT_LIVE) {
		const double tm = O(L::TIME), dur = O(L::DUR);
		for (int d = 0; d < k; ++d) {
			OtgProf p;
			otg_load_prof<D>(O, d, p);
			otg_eval(p, tm, dur, out_p[d], out_v[d], out_a[d]);
		}
	} else {
		for (int d = 0; d < k; ++d) { out_p[d] = O(L::OUT_P + d); out_v[d] = O(L::OUT_V + d); out_a[d] = O(L::OUT_A + d); }
	}
}

// ---- the per-tick kernels ------------------------------------------------------------------------------------------
// k_otg_advance: every robot.  Goal filter as a pure comparison (only when the host saw a goal write), then the state
// advance along the stored trajectory: registers only, no local memory.  Robots that need anything else this tick (goal
// change, pending recalculation, trajectory finishing, invalid goal) are appended to a compacted list.
// k_otg_slow: grid-stride over that list with the general per-robot routine below (otg_general).
// D = robot dof (joint task, k <= D task dofs) or 6 (cartesian).  GOFF_* = offsets of the goal rows in the G / P arenas.
template <class T, int D, bool CART>
__device__ void otg_general(const OtgJob& job, int64_t stride, int64_t r) {
	using L = OtgL<D>;
	const Row<double> O{job.O, stride, r};
	const Row<int32_t> OI{job.OI, stride, r};
	const Row<T> G{(T*)job.G, stride, r}, P{(T*)job.P, stride, r};
	const int k = CART ? 6 : job.k;
	if (!CART && job.rank && job.rank[r] == 0) return;
	int flags = OI(L::FLAGS);
	const int flags_in = flags;
	double tgt_p[D], tgt_v[D];
	OSC_UNROLL for (int i = 0; i < D; ++i) { tgt_p[i] = i < k ? O(L::TGT_P + i) : 0.0; tgt_v[i] = i < k ? O(L::TGT_V + i) : 0.0; }
	double out_p[D], out_v[D], out_a[D];
	bool have_out = false, out_dirty = false;
	auto load_out = [&]() {
		if (have_out) return;
		otg_current_output<D>(O, flags, k, out_p, out_v, out_a);
		if (flags & OTG_OUT_LIVE) { flags &= ~OTG_OUT_LIVE; out_dirty = true; }
		have_out = true;
	};
	double ref[9], gori[9], gang[3];
	bool ref_dirty = false;

	// ---- goal filter --------------------------------------------------------------------------------------------
	if constexpr (!CART) {  // OTG_joints::setGoalPositionAndVelocity, OTG_joints.cpp:96-113
		double gp[D], gv[D];
		OSC_UNROLL for (int i = 0; i < D; ++i) { gp[i] = i < k ? (double)G.ro(i) : 0.0; gv[i] = i < k ? (double)G.ro(D + i) : 0.0; }
		if (!(otg_approx<D>(gp, tgt_p, k, 1e-12) && otg_approx<D>(gv, tgt_v, k, 1e-12))) {
			flags = (flags & ~OTG_REACHED) | OTG_DIRTY;
			OSC_UNROLL for (int i = 0; i < D; ++i)
				if (i < k) { tgt_p[i] = gp[i]; tgt_v[i] = gv[i]; O(L::TGT_P + i) = gp[i]; O(L::TGT_V + i) = gv[i]; }
		}
	} else {
		using ML = MftL<7>;  // goal row offsets do not depend on N
		double gp[3], gv[3], gR[9], gw[3];
		OSC_UNROLL for (int i = 0; i < 3; ++i) {
			gp[i] = (double)G.ro(ML::GOAL_POS + i); gv[i] = (double)G.ro(ML::GOAL_LINVEL + i); gw[i] = (double)G.ro(ML::GOAL_ANGVEL + i);
		}
		OSC_UNROLL for (int i = 0; i < 9; ++i) { gR[i] = (double)G.ro(ML::GOAL_ORI + i); gori[i] = O(L::GORI + i); ref[i] = O(L::REF + i); }
		OSC_UNROLL for (int i = 0; i < 3; ++i) gang[i] = O(L::GANG + i);
		// setGoalPositionAndLinearVelocity, OTG_6dof_cartesian.cpp:146-155
		if (!(otg_approx<3>(gp, tgt_p, 3, 1e-3) && otg_approx<3>(gv, tgt_v, 3, 1e-3))) {
			flags = (flags & ~OTG_REACHED) | OTG_DIRTY;
			OSC_UNROLL for (int i = 0; i < 3; ++i) { tgt_p[i] = gp[i]; tgt_v[i] = gv[i]; O(L::TGT_P + i) = gp[i]; O(L::TGT_V + i) = gv[i]; }
		}
		// setGoalOrientationAndAngularVelocity, :157-199
		if (!otg_valid_rotation(gR)) {
			atomicOr(&job.status[r], 64);  // OSC_STATUS_BAD_GOAL_ORIENTATION: the reference throws here; the goal is ignored
		} else if (!(otg_approx<9>(gori, gR, 9, 1e-3) && otg_approx<3>(gang, gw, 3, 1e-3))) {
			flags = (flags & ~OTG_REACHED) | OTG_DIRTY | OTG_IN_SYNCED;  // pass_to_input, :186
			load_out();
			double nref[9], Rnp[9];
			otg_next_orientation(ref, out_p + 3, nref);
			otg_m3tmul(nref, ref, Rnp);
			OSC_UNROLL for (int i = 0; i < 9; ++i) { ref[i] = nref[i]; gori[i] = gR[i]; }
			OSC_UNROLL for (int i = 0; i < 3; ++i) gang[i] = gw[i];
			out_p[3] = out_p[4] = out_p[5] = 0.0;
			otg_m3v(Rnp, out_v + 3, out_v + 3);
			otg_m3v(Rnp, out_a + 3, out_a + 3);
			double r2g[9];
			otg_m3tmul(ref, gori, r2g);
			otg_rotvec(r2g, tgt_p + 3);
			otg_m3tv(ref, gang, tgt_v + 3);
			OSC_UNROLL for (int i = 3; i < 6; ++i) { O(L::TGT_P + i) = tgt_p[i]; O(L::TGT_V + i) = tgt_v[i]; }
			ref_dirty = true; out_dirty = true;
		}
	}

	auto store_cart_frames = [&]() {
		if constexpr (CART) {
			OSC_UNROLL for (int i = 0; i < 9; ++i) { O(L::REF + i) = ref[i]; O(L::GORI + i) = gori[i]; }
			OSC_UNROLL for (int i = 0; i < 3; ++i) O(L::GANG + i) = gang[i];
		}
	};
	auto store_out = [&]() {
		OSC_UNROLL for (int i = 0; i < D; ++i)
			if (i < k) { O(L::OUT_P + i) = out_p[i]; O(L::OUT_V + i) = out_v[i]; O(L::OUT_A + i) = out_a[i]; }
	};
	auto store_desired = [&]() {  // _desired_* = _otg->getNext*()
		if constexpr (!CART) {
			OSC_UNROLL for (int i = 0; i < D; ++i)
				if (i < k) { P(i) = (T)out_p[i]; P(D + i) = (T)out_v[i]; P(2 * D + i) = (T)out_a[i]; }
		} else {
			using ML = MftL<7>;
			double Rn[9], w[3], al[3];
			otg_next_orientation(ref, out_p + 3, Rn);
			otg_m3v(ref, out_v + 3, w);
			otg_m3v(ref, out_a + 3, al);
			OSC_UNROLL for (int i = 0; i < 3; ++i) {
				P(ML::GOAL_POS + i) = (T)out_p[i]; P(ML::GOAL_LINVEL + i) = (T)out_v[i]; P(ML::GOAL_LINACC + i) = (T)out_a[i];
				P(ML::GOAL_ANGVEL + i) = (T)w[i]; P(ML::GOAL_ANGACC + i) = (T)al[i];
			}
			OSC_UNROLL for (int i = 0; i < 9; ++i) P(ML::GOAL_ORI + i) = (T)Rn[i];
		}
	};

	// ---- update(), OTG_joints.cpp:115-150 / OTG_6dof_cartesian.cpp:201-236 ----------------------------------------
	if (flags & OTG_REACHED) {
		if (flags != flags_in) OI(L::FLAGS) = flags;
		return;
	}
	OtgProf prof[D];
	double duration, time;
	int result = OTG_WORKING;
	if ((flags & OTG_DIRTY) || !(flags & OTG_INIT)) {  // ruckig.hpp:193
		double cp[D], cv[D], ca[D];
		if (flags & OTG_IN_SYNCED) {
			load_out();
			OSC_UNROLL for (int i = 0; i < D; ++i) { cp[i] = out_p[i]; cv[i] = out_v[i]; ca[i] = out_a[i]; }
		} else {
			OSC_UNROLL for (int i = 0; i < D; ++i)
				if (i < k) { cp[i] = O(L::IN_P + i); cv[i] = O(L::IN_V + i); ca[i] = O(L::IN_A + i); }
		}
		// validate_input: the target velocity must respect the limits (input_parameter.hpp:289-300), no NaN
		OSC_UNROLL for (int i = 0; i < D; ++i)
			if (i < k) {
				if (isnan(cp[i]) || isnan(cv[i]) || isnan(ca[i]) || isnan(tgt_p[i]) || isnan(tgt_v[i])) result = OTG_ERR_INPUT;
				if (tgt_v[i] > job.max_v[i] || tgt_v[i] < -job.max_v[i]) result = OTG_ERR_INPUT;
			}
#ifdef OSC_OTG_DEBUG
		if (r == 0) printf("otg dbg CART=%d flags=%d cp=%g %g %g %g %g %g out_p=%g %g %g have_out=%d\n", (int)CART, flags, cp[0], cp[1], cp[2], cp[3], cp[4], cp[5], out_p[0], out_p[1], out_p[2], (int)have_out);
#endif
		if (result == OTG_WORKING) result = otg_calculate<D>(k, cp, cv, ca, tgt_p, tgt_v, job.max_v, job.max_a, prof, duration);
#ifdef OSC_OTG_DEBUG
		if (r == 0) printf("otg dbg after calc res=%d dur=%g pS=%g %g %g\n", result, duration, prof[0].pS, prof[1].pS, prof[2].pS);
#endif
		if (result != OTG_WORKING) {
			// the wrappers keep the previous output and zero the input velocity / acceleration (OTG_joints.cpp:144-149)
			if (flags & OTG_IN_SYNCED) {
				OSC_UNROLL for (int i = 0; i < D; ++i)
					if (i < k) O(L::IN_P + i) = cp[i];
			}
			OSC_UNROLL for (int i = 0; i < D; ++i)
				if (i < k) { O(L::IN_V + i) = 0.0; O(L::IN_A + i) = 0.0; }
			flags = (flags & ~OTG_IN_SYNCED) | OTG_DIRTY;
			OI(L::FLAGS) = flags;
			OI(L::RESULT) = result;
			OI(L::ERRORS) = OI(L::ERRORS) + 1;
			atomicOr(&job.status[r], 128);  // OSC_STATUS_OTG_ERROR
			if (out_dirty) { store_out(); store_cart_frames(); }
			return;
		}
		OSC_UNROLL for (int d = 0; d < D; ++d)
			if (d < k) {
				const OtgProf& p = prof[d];
				const int o = L::TRAJ + 11 * d;
				O(o) = p.brT; O(o + 1) = p.brA; O(o + 2) = p.pS; O(o + 3) = p.vS; O(o + 4) = p.t0; O(o + 5) = p.t1; O(o + 6) = p.t2;
				O(o + 7) = p.t6; O(o + 8) = p.a0; O(o + 9) = p.a2; O(o + 10) = p.a6;
				O(L::SEG + 2 * D + d) = -1.0;  // cached segment stale
			}
		O(L::DUR) = duration;
		time = 0.0;
		flags = (flags | OTG_INIT) & ~OTG_DIRTY;
	} else {
		if (flags & OTG_IN_SYNCED) load_out();  // before the clock moves: a Finished tick needs the previous output as _input
		for (int d = 0; d < k; ++d) otg_load_prof<D>(O, d, prof[d]);
		duration = O(L::DUR);
		time = O(L::TIME);
	}
	time += job.dt;  // ruckig.hpp:207
	double np[D], nv[D], na[D];
	OSC_UNROLL for (int d = 0; d < D; ++d)
		if (d < k) otg_eval(prof[d], time, duration, np[d], nv[d], na[d]);
	O(L::TIME) = time;
	if (time > duration) {  // Result::Finished, ruckig.hpp:217
		result = OTG_FINISHED;
		double vn = 0;
		OSC_UNROLL for (int d = 0; d < D; ++d)
			if (d < k) vn += nv[d] * nv[d];
		// the wrapper returns before pass_to_input: _input keeps the state of the previous tick
		if (flags & OTG_IN_SYNCED) {
			load_out();
			OSC_UNROLL for (int i = 0; i < D; ++i)
				if (i < k) { O(L::IN_P + i) = out_p[i]; O(L::IN_V + i) = out_v[i]; O(L::IN_A + i) = out_a[i]; }
			flags &= ~OTG_IN_SYNCED;
		}
		OSC_UNROLL for (int d = 0; d < D; ++d)
			if (d < k) { out_p[d] = np[d]; out_v[d] = nv[d]; out_a[d] = na[d]; }
		have_out = true;
		if (sqrt(vn) < 1e-3) {
			flags |= OTG_REACHED;
		} else if constexpr (!CART) {  // Q14 (oracle/otg_oracle.hpp): keep the target position, zero the target velocity
			double zero[D];
			OSC_UNROLL for (int i = 0; i < D; ++i) zero[i] = 0.0;
			if (!otg_approx<D>(zero, tgt_v, k, 1e-12)) {
				flags = (flags & ~OTG_REACHED) | OTG_DIRTY;
				OSC_UNROLL for (int i = 0; i < D; ++i)
					if (i < k) O(L::TGT_V + i) = 0.0;
			}
		} else {  // OTG_6dof_cartesian.cpp:214-218
			const double zero[3] = {0, 0, 0};
			if (!otg_approx<3>(zero, tgt_v, 3, 1e-3)) {
				flags = (flags & ~OTG_REACHED) | OTG_DIRTY;
				OSC_UNROLL for (int i = 0; i < 3; ++i) O(L::TGT_V + i) = 0.0;
			}
			if (!otg_approx<3>(gang, zero, 3, 1e-3)) {  // re-reference at the new output with a zero goal angular velocity
				flags = (flags & ~OTG_REACHED) | OTG_DIRTY | OTG_IN_SYNCED;
				double nref[9], Rnp[9];
				otg_next_orientation(ref, out_p + 3, nref);
				otg_m3tmul(nref, ref, Rnp);
				OSC_UNROLL for (int i = 0; i < 9; ++i) ref[i] = nref[i];
				gang[0] = gang[1] = gang[2] = 0.0;
				out_p[3] = out_p[4] = out_p[5] = 0.0;
				otg_m3v(Rnp, out_v + 3, out_v + 3);
				otg_m3v(Rnp, out_a + 3, out_a + 3);
				double r2g[9], tw[3];
				otg_m3tmul(ref, gori, r2g);
				otg_rotvec(r2g, tw);
				OSC_UNROLL for (int i = 0; i < 3; ++i) { O(L::TGT_P + 3 + i) = tw[i]; O(L::TGT_V + 3 + i) = 0.0; }
				ref_dirty = true;
			}
		}
	} else {  // Result::Working: pass_to_input
		OSC_UNROLL for (int d = 0; d < D; ++d)
			if (d < k) { out_p[d] = np[d]; out_v[d] = nv[d]; out_a[d] = na[d]; }
		have_out = true;
		flags |= OTG_IN_SYNCED;
	}
	store_out();
	if (ref_dirty) store_cart_frames();
	OI(L::FLAGS) = flags;
	OI(L::RESULT) = result;
	store_desired();
}

// every OTG of the hierarchy in ONE launch: blockIdx.y selects the task
struct OtgJobs {
	OtgJob j[OSC_MAX_TASKS];
	int cart[OSC_MAX_TASKS];
};

// N = robot dof.  Launched with programmatic stream serialisation behind k_otg_advance: its launch latency hides under the
// advance kernel; the list is complete and visible once griddepcontrol.wait returns (a no-op under a plain launch).
// resident blocks per SM of k_otg_slow (register cap = 65536 / (128 * blocks)).  Measured on the all-robots-recalculate tick, 65,536
// robots x 2 tasks (r2j): 2 blocks (255 registers) 0.412 ms, 4 blocks (128 registers, 2.2 KB of spills) 0.349 ms, 6 blocks (80
// registers) 0.362 ms; fewer resident blocks than 2 (shared-memory limited) 0.59 ms.  The kernel is bound by the latency of one
// thread walking 13 DoF profiles, not by its local-memory traffic: it wants resident warps.
#ifndef OSC_OTG_SLOW_BLOCKS
#define OSC_OTG_SLOW_BLOCKS 4
#endif
template <class T, int N>
__global__ void __launch_bounds__(128, OSC_OTG_SLOW_BLOCKS) k_otg_slow(const __grid_constant__ OtgJobs js, int64_t stride) {
	asm volatile("griddepcontrol.wait;" ::: "memory");
	const OtgJob& job = js.j[blockIdx.y];
	const int n = *job.slow_count;
	if (blockIdx.x == 0 && threadIdx.x == 0) *job.slow_next = 0;  // re-arm the other counter (ping-pong, like k_tick_tail)
	if (js.cart[blockIdx.y]) {
		for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) otg_general<T, 6, true>(job, stride, job.slow_list[i]);
	} else {
		for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) otg_general<T, N, false>(job, stride, job.slow_list[i]);
	}
}

template <class T, int D, bool CART>
__device__ __forceinline__ void otg_advance_robot(const OtgJob& job, int64_t B, int64_t stride) {
	using L = OtgL<D>;
	const int64_t r0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const bool live = r0 < B && !(!CART && job.rank && job.rank[r0 < B ? r0 : 0] == 0);
	const int64_t r = r0 < B ? r0 : B - 1;  // out-of-range lanes stay in the warp for the ballot below
	const Row<double> O{job.O, stride, r};
	const Row<int32_t> OI{job.OI, stride, r};
	const Row<T> G{(T*)job.G, stride, r}, P{(T*)job.P, stride, r};
	const int k = CART ? 6 : job.k;
	const int flags = OI(L::FLAGS);
	bool slow = false;
	if (job.check_goals) {  // the comparisons of setGoal*And*Velocity without their side effects
		if constexpr (!CART) {
			double d2p = 0, g2p = 0, t2p = 0, d2v = 0, g2v = 0, t2v = 0;
			OSC_UNROLL for (int i = 0; i < D; ++i)
				if (i < k) {
					const double gp = (double)G.ro(i), gv = (double)G.ro(D + i), tp = O(L::TGT_P + i), tv = O(L::TGT_V + i);
					d2p += (gp - tp) * (gp - tp); g2p += gp * gp; t2p += tp * tp;
					d2v += (gv - tv) * (gv - tv); g2v += gv * gv; t2v += tv * tv;
				}
			slow = !(d2p <= 1e-24 * fmin(g2p, t2p) && d2v <= 1e-24 * fmin(g2v, t2v));
		} else {
			using ML = MftL<7>;
			double gp[3], gv[3], gw[3], gR[9], tp[3], tv[3], gori[9], gang[3];
			OSC_UNROLL for (int i = 0; i < 3; ++i) {
				gp[i] = (double)G.ro(ML::GOAL_POS + i); gv[i] = (double)G.ro(ML::GOAL_LINVEL + i); gw[i] = (double)G.ro(ML::GOAL_ANGVEL + i);
				tp[i] = O(L::TGT_P + i); tv[i] = O(L::TGT_V + i); gang[i] = O(L::GANG + i);
			}
			OSC_UNROLL for (int i = 0; i < 9; ++i) { gR[i] = (double)G.ro(ML::GOAL_ORI + i); gori[i] = O(L::GORI + i); }
			slow = !(otg_approx<3>(gp, tp, 3, 1e-3) && otg_approx<3>(gv, tv, 3, 1e-3)) || !otg_valid_rotation(gR) ||
				   !(otg_approx<9>(gori, gR, 9, 1e-3) && otg_approx<3>(gang, gw, 3, 1e-3));
		}
	}
	bool idle = false;
	if (!slow) {
		idle = (flags & OTG_REACHED) != 0;  // OTG_joints.cpp:116-118
		slow = !idle && ((flags & OTG_DIRTY) || !(flags & OTG_INIT));
	}
	// one batch of independent loads (clock, duration, the cached segment of every DoF) BEFORE any store: a load per DoF
	// iteration behind the previous iteration's possibly-aliasing store paid a DRAM round trip each (ncu r1p: 24 us -> see
	// profiles/README.md)
	double time = 0, duration = 0;
	double s_off[D], s_ts[D], s_te[D], s_p[D], s_v[D], s_a[D];
	double ref[CART ? 9 : 1];
	if (!slow && !idle) {
		time = O(L::TIME) + job.dt;  // ruckig.hpp:207
		duration = O(L::DUR);
		if constexpr (CART) {
			OSC_UNROLL for (int i = 0; i < 9; ++i) ref[i] = O(L::REF + i);
		}
		OSC_UNROLL for (int d = 0; d < D; ++d)
			if (d < k) {
				s_off[d] = O(L::SEG + d); s_ts[d] = O(L::SEG + D + d); s_te[d] = O(L::SEG + 2 * D + d); s_p[d] = O(L::SEG + 3 * D + d);
				s_v[d] = O(L::SEG + 4 * D + d); s_a[d] = O(L::SEG + 5 * D + d);
			}
		slow = time > duration;      // Result::Finished this tick: wrapper logic in the general routine
	}
	slow = slow && live;
	{  // warp-aggregated append: one atomic per warp, consecutive robots stay consecutive in the list (coalesced general path)
		const unsigned m = __ballot_sync(0xffffffffu, slow);
		if (m) {
			const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
			int base = 0;
			if (lane == leader) base = atomicAdd(job.slow_count, __popc(m));
			base = __shfl_sync(0xffffffffu, base, leader);
			if (slow) job.slow_list[base + __popc(m & ((1u << lane) - 1u))] = (int)r;
		}
	}
	if (slow || idle || !live) return;
	double np[D], nv[D], na[D];
	OSC_UNROLL for (int d = 0; d < D; ++d)
		if (d < k) {
			OtgSeg sg{s_off[d], s_ts[d], s_te[d], s_p[d], s_v[d], s_a[d]};
			double td = sg.off > 0 ? time - sg.off : time;
			if (!(td < sg.te) || td < sg.ts) {  // crossed into the next segment (or stale): rebuild from the stored profile
				OtgProf p;
				otg_load_prof<D>(O, d, p);
				sg = otg_segment(p, time);
				O(L::SEG + d) = sg.off; O(L::SEG + D + d) = sg.ts; O(L::SEG + 2 * D + d) = sg.te; O(L::SEG + 3 * D + d) = sg.p;
				O(L::SEG + 4 * D + d) = sg.v; O(L::SEG + 5 * D + d) = sg.a;
				td = sg.off > 0 ? time - sg.off : time;
			}
			const double t = td - sg.ts;
			np[d] = sg.p + t * (sg.v + t * (sg.a / 2 + t * 0.0 / 6));  // utils.hpp:43-49
			nv[d] = sg.v + t * (sg.a + t * 0.0 / 2);
			na[d] = sg.a + t * 0.0;
		}
	O(L::TIME) = time;
	if ((flags & (OTG_IN_SYNCED | OTG_OUT_LIVE)) != (OTG_IN_SYNCED | OTG_OUT_LIVE)) OI(L::FLAGS) = flags | OTG_IN_SYNCED | OTG_OUT_LIVE;
	if constexpr (!CART) {
		OSC_UNROLL for (int i = 0; i < D; ++i)
			if (i < k) { P(i) = (T)np[i]; P(D + i) = (T)nv[i]; P(2 * D + i) = (T)na[i]; }
	} else {
		using ML = MftL<7>;
		double Rn[9];
		const double rv[3] = {np[3], np[4], np[5]};
		otg_next_orientation(ref, rv, Rn);
		OSC_UNROLL for (int i = 0; i < 3; ++i) {
			P(ML::GOAL_POS + i) = (T)np[i]; P(ML::GOAL_LINVEL + i) = (T)nv[i]; P(ML::GOAL_LINACC + i) = (T)na[i];
			P(ML::GOAL_ANGVEL + i) = (T)(ref[3 * i] * nv[3] + ref[3 * i + 1] * nv[4] + ref[3 * i + 2] * nv[5]);
			P(ML::GOAL_ANGACC + i) = (T)(ref[3 * i] * na[3] + ref[3 * i + 1] * na[4] + ref[3 * i + 2] * na[5]);
		}
		OSC_UNROLL for (int i = 0; i < 9; ++i) P(ML::GOAL_ORI + i) = (T)Rn[i];
	}
}
template <class T, int N>
__global__ void __launch_bounds__(128, 4) k_otg_advance(const __grid_constant__ OtgJobs js, int64_t B, int64_t stride) {
	asm volatile("griddepcontrol.launch_dependents;");  // lets the launch of k_otg_slow proceed under this kernel
	if (js.cart[blockIdx.y]) otg_advance_robot<T, 6, true>(js.j[blockIdx.y], B, stride);
	else otg_advance_robot<T, N, false>(js.j[blockIdx.y], B, stride);
}

// ---- k_otg_advance_tma (EXPERIMENT builds only): the steady-state advance with bulk-copy (TMA) staging ------------------------
// A persistent kernel: each block loops over tiles of 128 robots x one task; ONE thread issues the rows of the tile as
// cp.async.bulk copies (global -> shared, completion on an mbarrier: SASS UBLKCP / SYNCS.ARRIVE.TRANS64), the 128 threads
// evaluate the tile out of shared memory; four blocks per SM (54 KB each).  Measured on B200 (r2f): it ties the plain kernel
// above at 65,536 and at 1,048,576 robots, and a two-stage variant (half the resident threads) is 30 % slower.  The advance
// moves ~1.3 KB of DRAM traffic per robot and tick (6-row cached segments, REF, DUR, flags on top of the 792 algorithmic bytes)
// at ~0.7 of the copy bandwidth either way: how the loads are issued is not its bound.  Kept out of the product build.
#ifdef OSC_EXPERIMENT
template <int N> struct OtgStage {
	static constexpr int DMAX = N > 6 ? N : 6;
	static constexpr int ROWS = 2 + 6 * DMAX + 9;          // doubles rows of 128 robots
	static constexpr int BYTES = ROWS * 128 * 8 + 128 * 4;  // + the int flags row
};
__device__ __forceinline__ uint32_t otg_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void otg_mbar_init(uint64_t* bar, int count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(otg_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void otg_mbar_expect(uint64_t* bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(otg_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void otg_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(otg_smem_u32(dst)), "l"(src),
				 "r"(bytes), "r"(otg_smem_u32(bar))
				 : "memory");
}
__device__ __forceinline__ void otg_mbar_wait(uint64_t* bar, uint32_t parity) {
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"OTG_WAIT_%=:\n"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		"@p bra OTG_DONE_%=;\n"
		"bra OTG_WAIT_%=;\n"
		"OTG_DONE_%=:\n"
		"}\n" ::"r"(otg_smem_u32(bar)),
		"r"(parity)
		: "memory");
}

// issue the rows of one tile (one thread): D = dof of the task's OTG
template <int D, bool CART> __device__ __forceinline__ void otg_tile_issue(const OtgJob& job, int64_t stride, int64_t r0, double* buf, int32_t* fl, uint64_t* bar) {
	using L = OtgL<D>;
	const int64_t left = stride - r0;
	const uint32_t cnt = (uint32_t)(left < 128 ? left : 128);  // stride is a multiple of 32 robots: every piece is a multiple of 256 B
	constexpr int NR = 2 + 6 * D + (CART ? 9 : 0);
	otg_mbar_expect(bar, NR * cnt * 8 + cnt * 4);
	const double* O = job.O + r0;
	otg_bulk_g2s(buf, O + (int64_t)L::TIME * stride, cnt * 8, bar);
	otg_bulk_g2s(buf + 128, O + (int64_t)L::DUR * stride, cnt * 8, bar);
	_Pragma("unroll 1") for (int e = 0; e < 6 * D; ++e) otg_bulk_g2s(buf + (2 + e) * 128, O + (int64_t)(L::SEG + e) * stride, cnt * 8, bar);
	if constexpr (CART) {
		_Pragma("unroll 1") for (int e = 0; e < 9; ++e) otg_bulk_g2s(buf + (2 + 6 * D + e) * 128, O + (int64_t)(L::REF + e) * stride, cnt * 8, bar);
	}
	otg_bulk_g2s(fl, job.OI + (int64_t)L::FLAGS * stride + r0, cnt * 4, bar);
}

// evaluate one staged tile: same arithmetic, same stores and the same slow-list protocol as otg_advance_robot (check_goals == 0)
template <class T, int D, bool CART>
__device__ __forceinline__ void otg_tile_eval(const OtgJob& job, int64_t B, int64_t stride, int64_t r0, const double* buf, const int32_t* fl) {
	using L = OtgL<D>;
	const int tid = threadIdx.x;
	const int64_t rr = r0 + tid;
	const int64_t r = rr < B ? rr : B - 1;
	const bool live = rr < B && !(!CART && job.rank && job.rank[r] == 0);
	const Row<double> O{job.O, stride, r};
	const Row<int32_t> OI{job.OI, stride, r};
	const Row<T> P{(T*)job.P, stride, r};
	const int k = CART ? 6 : job.k;
	const int flags = rr < B ? fl[tid] : (int)OTG_REACHED;
	const bool idle = (flags & OTG_REACHED) != 0;
	bool slow = !idle && ((flags & OTG_DIRTY) || !(flags & OTG_INIT));
	double time = 0;
	if (!slow && !idle) {
		time = buf[tid] + job.dt;
		slow = time > buf[128 + tid];
	}
	slow = slow && live;
	{
		const unsigned m = __ballot_sync(0xffffffffu, slow);
		if (m) {
			const int lane = tid & 31, leader = __ffs(m) - 1;
			int base = 0;
			if (lane == leader) base = atomicAdd(job.slow_count, __popc(m));
			base = __shfl_sync(0xffffffffu, base, leader);
			if (slow) job.slow_list[base + __popc(m & ((1u << lane) - 1u))] = (int)r;
		}
	}
	if (slow || idle || !live) return;
	double np[D], nv[D], na[D];
	OSC_UNROLL for (int d = 0; d < D; ++d)
		if (d < k) {
			const double* sg_ = buf + (2 + d) * 128 + tid;
			OtgSeg sg{sg_[0], sg_[D * 128], sg_[2 * D * 128], sg_[3 * D * 128], sg_[4 * D * 128], sg_[5 * D * 128]};
			double td = sg.off > 0 ? time - sg.off : time;
			if (!(td < sg.te) || td < sg.ts) {
				OtgProf p;
				otg_load_prof<D>(O, d, p);
				sg = otg_segment(p, time);
				O(L::SEG + d) = sg.off; O(L::SEG + D + d) = sg.ts; O(L::SEG + 2 * D + d) = sg.te; O(L::SEG + 3 * D + d) = sg.p;
				O(L::SEG + 4 * D + d) = sg.v; O(L::SEG + 5 * D + d) = sg.a;
				td = sg.off > 0 ? time - sg.off : time;
			}
			const double t = td - sg.ts;
			np[d] = sg.p + t * (sg.v + t * (sg.a / 2 + t * 0.0 / 6));
			nv[d] = sg.v + t * (sg.a + t * 0.0 / 2);
			na[d] = sg.a + t * 0.0;
		}
	O(L::TIME) = time;
	if ((flags & (OTG_IN_SYNCED | OTG_OUT_LIVE)) != (OTG_IN_SYNCED | OTG_OUT_LIVE)) OI(L::FLAGS) = flags | OTG_IN_SYNCED | OTG_OUT_LIVE;
	if constexpr (!CART) {
		OSC_UNROLL for (int i = 0; i < D; ++i)
			if (i < k) { P(i) = (T)np[i]; P(D + i) = (T)nv[i]; P(2 * D + i) = (T)na[i]; }
	} else {
		using ML = MftL<7>;
		double ref[9], Rn[9];
		OSC_UNROLL for (int i = 0; i < 9; ++i) ref[i] = buf[(2 + 6 * D + i) * 128 + tid];
		const double rv[3] = {np[3], np[4], np[5]};
		otg_next_orientation(ref, rv, Rn);
		OSC_UNROLL for (int i = 0; i < 3; ++i) {
			P(ML::GOAL_POS + i) = (T)np[i]; P(ML::GOAL_LINVEL + i) = (T)nv[i]; P(ML::GOAL_LINACC + i) = (T)na[i];
			P(ML::GOAL_ANGVEL + i) = (T)(ref[3 * i] * nv[3] + ref[3 * i + 1] * nv[4] + ref[3 * i + 2] * nv[5]);
			P(ML::GOAL_ANGACC + i) = (T)(ref[3 * i] * na[3] + ref[3 * i + 1] * na[4] + ref[3 * i + 2] * na[5]);
		}
		OSC_UNROLL for (int i = 0; i < 9; ++i) P(ML::GOAL_ORI + i) = (T)Rn[i];
	}
}

template <class T, int N>
__global__ void __launch_bounds__(128, 4) k_otg_advance_tma(const __grid_constant__ OtgJobs js, int64_t B, int64_t stride, int tiles_x, int n_tiles) {
	extern __shared__ __align__(128) unsigned char otg_stage_raw[];
	__shared__ __align__(8) uint64_t bar;
	asm volatile("griddepcontrol.launch_dependents;");
	double* buf = reinterpret_cast<double*>(otg_stage_raw);
	int32_t* fl = reinterpret_cast<int32_t*>(otg_stage_raw + (size_t)OtgStage<N>::ROWS * 128 * 8);
	if (threadIdx.x == 0) {
		otg_mbar_init(&bar, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	int it = 0;
	for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
		const int y = tile / tiles_x;
		const int64_t r0 = (int64_t)(tile - y * tiles_x) * 128;
		if (threadIdx.x == 0) {  // one stage: the other three resident blocks of the SM cover this block's copy latency
			if (js.cart[y]) otg_tile_issue<6, true>(js.j[y], stride, r0, buf, fl, &bar);
			else otg_tile_issue<N, false>(js.j[y], stride, r0, buf, fl, &bar);
		}
		otg_mbar_wait(&bar, it & 1);
		if (js.cart[y]) otg_tile_eval<T, 6, true>(js.j[y], B, stride, r0, buf, fl);
		else otg_tile_eval<T, N, false>(js.j[y], B, stride, r0, buf, fl);
		__syncthreads();  // every read of the stage is done before it is refilled
	}
}
#endif  // OSC_EXPERIMENT

// ---- cold path: reInitialize / enable / disable ---------------------------------------------------------------------
// mode bit 1: reInitialize (all), 2: reInitializeLinear, 4: reInitializeAngular (OTG_6dof_cartesian.cpp:46-84, OTG_joints.cpp:28-41);
// 64: the caller also reset the task's goals to the current state (reInitializeTask, parametrize*Spaces): mirror that in G;
// 32: fresh instance (constructor); 8: copy the task's goal rows into the G arena (enable: the user's goals move there);
// 16: copy the G arena back into the goal rows (disable) and return; 128: mark the input changed (new limits).
// cur: joints = the task's S.CUR_POS rows (T), cartesian = the task's P arena (CUR_POS / CUR_ORI rows).
template <class T, int D, bool CART>
__global__ void __launch_bounds__(128) k_otg_reinit(const __grid_constant__ OtgJob job, const T* cur, int64_t B, int64_t stride, int mode, int goal_rows) {
	using L = OtgL<D>;
	const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= B) return;
	const Row<double> O{job.O, stride, r};
	const Row<int32_t> OI{job.OI, stride, r};
	const Row<T> G{(T*)job.G, stride, r}, P{(T*)job.P, stride, r}, C{const_cast<T*>(cur), stride, r};
	if (mode & 16) {
		for (int e = 0; e < goal_rows; ++e) P(e) = G(e);
		return;
	}
	if (mode & 8)
		for (int e = 0; e < goal_rows; ++e) G(e) = P(e);
	if (mode & 32) {
		for (int e = 0; e < L::O_SIZE; ++e) O(e) = 0.0;
		OI(L::FLAGS) = 0; OI(L::RESULT) = OTG_FINISHED; OI(L::ERRORS) = 0;
		if constexpr (CART) {  // constructor: _reference_frame = initial orientation (OTG_6dof_cartesian.cpp:43)
			using ML = MftL<7>;
			for (int i = 0; i < 9; ++i) O(L::REF + i) = (double)C(ML::CUR_ORI + i);
		}
	}
	if (mode & 128) OI(L::FLAGS) = OI(L::FLAGS) | OTG_DIRTY;  // the limits changed: input != current_input (input_parameter.hpp:370-371)
	if (!(mode & 7)) return;
	const int k = CART ? 6 : job.k;
	int flags = OI(L::FLAGS);
	if constexpr (!CART) {
		double p0[D], tp[D], tv[D], zero[D];
		for (int i = 0; i < D; ++i) { p0[i] = i < k ? (double)C(i) : 0.0; tp[i] = i < k ? O(L::TGT_P + i) : 0.0; tv[i] = i < k ? O(L::TGT_V + i) : 0.0; zero[i] = 0.0; }
		if (!(otg_approx<D>(p0, tp, k, 1e-12) && otg_approx<D>(zero, tv, k, 1e-12))) {
			flags &= ~OTG_REACHED;
			for (int i = 0; i < k; ++i) { O(L::TGT_P + i) = p0[i]; O(L::TGT_V + i) = 0.0; }
		}
		for (int i = 0; i < k; ++i) { O(L::OUT_P + i) = p0[i]; O(L::OUT_V + i) = 0.0; O(L::OUT_A + i) = 0.0; }
		if (mode & 64)
			for (int i = 0; i < k; ++i) { G(i) = (T)p0[i]; G(D + i) = T(0); G(2 * D + i) = T(0); }
		flags = (flags & ~OTG_OUT_LIVE) | OTG_IN_SYNCED | OTG_DIRTY;
	} else {
		using ML = MftL<7>;
		double tp[6], tv[6], ref[9], gori[9], gang[3], out_p[6], out_v[6], out_a[6];
		for (int i = 0; i < 6; ++i) { tp[i] = O(L::TGT_P + i); tv[i] = O(L::TGT_V + i); }
		otg_current_output<6>(O, flags, 6, out_p, out_v, out_a);
		flags &= ~OTG_OUT_LIVE;
		for (int i = 0; i < 9; ++i) { ref[i] = O(L::REF + i); gori[i] = O(L::GORI + i); }
		for (int i = 0; i < 3; ++i) gang[i] = O(L::GANG + i);
		double in_p[6], in_v[6], in_a[6];
		for (int i = 0; i < 6; ++i) {
			in_p[i] = (flags & OTG_IN_SYNCED) ? out_p[i] : O(L::IN_P + i);
			in_v[i] = (flags & OTG_IN_SYNCED) ? out_v[i] : O(L::IN_V + i);
			in_a[i] = (flags & OTG_IN_SYNCED) ? out_a[i] : O(L::IN_A + i);
		}
		const double zero[3] = {0, 0, 0};
		if (mode & (1 | 2)) {  // setGoalPosition(initial_position)
			double p0[3];
			for (int i = 0; i < 3; ++i) p0[i] = (double)C(ML::CUR_POS + i);
			if (mode & 64)
				for (int i = 0; i < 3; ++i) { G(ML::GOAL_POS + i) = (T)p0[i]; G(ML::GOAL_LINVEL + i) = T(0); G(ML::GOAL_LINACC + i) = T(0); }
			if (!(otg_approx<3>(p0, tp, 3, 1e-3) && otg_approx<3>(zero, tv, 3, 1e-3))) {
				flags &= ~OTG_REACHED;
				for (int i = 0; i < 3; ++i) { tp[i] = p0[i]; tv[i] = 0.0; }
			}
		}
		if (mode & (1 | 4)) {  // setGoalOrientation(initial_orientation)
			double R0[9];
			for (int i = 0; i < 9; ++i) R0[i] = (double)C(ML::CUR_ORI + i);
			if (mode & 64) {
				for (int i = 0; i < 9; ++i) G(ML::GOAL_ORI + i) = (T)R0[i];
				for (int i = 0; i < 3; ++i) { G(ML::GOAL_ANGVEL + i) = T(0); G(ML::GOAL_ANGACC + i) = T(0); }
			}
			if (otg_valid_rotation(R0) && !(otg_approx<9>(gori, R0, 9, 1e-3) && otg_approx<3>(gang, zero, 3, 1e-3))) {
				flags &= ~OTG_REACHED;
				double nref[9], Rnp[9];
				otg_next_orientation(ref, out_p + 3, nref);
				otg_m3tmul(nref, ref, Rnp);
				for (int i = 0; i < 9; ++i) { ref[i] = nref[i]; gori[i] = R0[i]; }
				gang[0] = gang[1] = gang[2] = 0.0;
				out_p[3] = out_p[4] = out_p[5] = 0.0;
				otg_m3v(Rnp, out_v + 3, out_v + 3);
				otg_m3v(Rnp, out_a + 3, out_a + 3);
				for (int i = 0; i < 6; ++i) { in_p[i] = out_p[i]; in_v[i] = out_v[i]; in_a[i] = out_a[i]; }  // pass_to_input
				double r2g[9];
				otg_m3tmul(ref, gori, r2g);
				otg_rotvec(r2g, tp + 3);
				tv[3] = tv[4] = tv[5] = 0.0;
			}
		}
		const int lo = (mode & (1 | 2)) ? 0 : 3, hi = (mode & (1 | 4)) ? 6 : 3;
		for (int i = lo; i < hi; ++i) { in_p[i] = tp[i]; in_v[i] = 0.0; in_a[i] = 0.0; out_p[i] = tp[i]; out_v[i] = 0.0; out_a[i] = 0.0; }
		for (int i = 0; i < 6; ++i) {
			O(L::TGT_P + i) = tp[i]; O(L::TGT_V + i) = tv[i];
			O(L::OUT_P + i) = out_p[i]; O(L::OUT_V + i) = out_v[i]; O(L::OUT_A + i) = out_a[i];
			O(L::IN_P + i) = in_p[i]; O(L::IN_V + i) = in_v[i]; O(L::IN_A + i) = in_a[i];
		}
		for (int i = 0; i < 9; ++i) { O(L::REF + i) = ref[i]; O(L::GORI + i) = gori[i]; }
		for (int i = 0; i < 3; ++i) O(L::GANG + i) = gang[i];
		flags = (flags & ~OTG_IN_SYNCED) | OTG_DIRTY;
	}
	OI(L::FLAGS) = flags;
}

}  // namespace osc
