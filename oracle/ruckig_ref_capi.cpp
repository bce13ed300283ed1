// ruckig_ref_capi.cpp -- TEST INFRASTRUCTURE.  A C API over the UNMODIFIED Ruckig library that the reference vendors
// (/root/reference/ruckig, community version; used by src/helper_modules/OTG_joints.cpp:20 and
// OTG_6dof_cartesian.cpp:34 as Ruckig<DynamicDOFs/6, EigenVector>).  Compiled by oracle/Makefile target `ref` straight
// from the sources where they lie under /root/reference/ruckig/{include,src}; output oracle/_ref/libruckig_ref.so.
// Nothing of the reference is copied into this repository: this file is our own glue and only calls Ruckig's public
// API (ruckig::Ruckig<0>::update, InputParameter<0>, OutputParameter<0>).  The reference instantiates the vectors with
// Eigen (absent here); the arithmetic of Ruckig does not depend on the vector container, so StandardVector
// (std::vector) is used.  Only tests/, tools/make_otg_golden.py and the oracle (as a checker backend) load this.
#include <ruckig/ruckig.hpp>

#include <cstring>
#include <limits>

using namespace ruckig;

namespace {
struct Handle {
	size_t n;
	Ruckig<0> otg;
	InputParameter<0> in;
	OutputParameter<0> out;
	Handle(size_t n_, double dt) : n(n_), otg(n_, dt), in(n_), out(n_) {
		in.synchronization = Synchronization::Phase;  // OTG_joints.cpp:23, OTG_6dof_cartesian.cpp:37
	}
};
}  // namespace

extern "C" {

void* rkref_create(int dofs, double dt) { return new Handle((size_t)dofs, dt); }
void rkref_destroy(void* h) { delete (Handle*)h; }

// One Ruckig::update(input, output) with the acceleration-limited setting of the reference wrappers
// (max_jerk = +inf: OTG_joints.cpp:86-89, OTG_6dof_cartesian.h disableJerkLimits).  `max_jerk` may be null (= +inf).
// On a negative result the output object is restored to its previous value, exactly what the reference wrappers do
// (OTG_joints.cpp:124,144: previous_output = _output; ... _output = previous_output).
// Returns ruckig::Result.  new_* receive output.new_position/velocity/acceleration; info = {time, duration, new_calculation}.
int rkref_update(void* hv, const double* cur_p, const double* cur_v, const double* cur_a, const double* tgt_p, const double* tgt_v,
				 const double* max_v, const double* max_a, const double* max_j, double* new_p, double* new_v, double* new_a, double* info) {
	Handle* h = (Handle*)hv;
	for (size_t i = 0; i < h->n; ++i) {
		h->in.current_position[i] = cur_p[i];
		h->in.current_velocity[i] = cur_v[i];
		h->in.current_acceleration[i] = cur_a[i];
		h->in.target_position[i] = tgt_p[i];
		h->in.target_velocity[i] = tgt_v[i];
		h->in.target_acceleration[i] = 0.0;
		h->in.max_velocity[i] = max_v[i];
		h->in.max_acceleration[i] = max_a[i];
		h->in.max_jerk[i] = max_j ? max_j[i] : std::numeric_limits<double>::infinity();
	}
	OutputParameter<0> previous = h->out;
	const Result res = h->otg.update(h->in, h->out);
	if (res < 0) h->out = previous;
	for (size_t i = 0; i < h->n; ++i) {
		new_p[i] = h->out.new_position[i];
		new_v[i] = h->out.new_velocity[i];
		new_a[i] = h->out.new_acceleration[i];
	}
	if (info) {
		info[0] = h->out.time;
		info[1] = h->out.trajectory.get_duration();
		info[2] = h->out.new_calculation ? 1.0 : 0.0;
	}
	return (int)res;
}

// the profile of one DoF of the current trajectory: t[7], a[7] (segment accelerations), then brake {duration, a}
void rkref_profile(void* hv, int dof, double* t7, double* a7, double* brake2) {
	Handle* h = (Handle*)hv;
	const auto profiles = h->out.trajectory.get_profiles();
	const Profile& p = profiles[0][dof];
	for (int i = 0; i < 7; ++i) { t7[i] = p.t[i]; a7[i] = p.a[i]; }
	brake2[0] = p.brake.duration;
	brake2[1] = p.brake.a[0];
}

// Ruckig::reset(): forces a new calculation at the next update
void rkref_reset(void* hv) { ((Handle*)hv)->otg.reset(); }

}  // extern "C"
