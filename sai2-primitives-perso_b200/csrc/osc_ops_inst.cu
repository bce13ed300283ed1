// osc_ops_inst.cu -- one instantiation of the kernel set.  Compiled once per (OSC_T, OSC_N) pair:
//   nvcc -DOSC_T=double -DOSC_TNAME=f64 -DOSC_N=7 -c osc_ops_inst.cu -o ops_f64_7.o
#include <stdlib.h>

#include <algorithm>

#include "osc_aux_kernels.cuh"
#include "osc_fused3.cuh"
#include "osc_fused_jt.cuh"
#include "osc_blend.cuh"
#include "osc_otg.cuh"
#include "osc_rollout.cuh"
#include "osc_host.h"

#ifndef OSC_T
#define OSC_T double
#define OSC_TNAME f64
#define OSC_N 7
#endif

namespace osc {

#define OSC_CAT_(a, b, c) make_ops_##a##_##b
#define OSC_CAT(a, b) OSC_CAT_(a, b, 0)

namespace {

// T = the batch precision = the STORAGE type of every row.  The fused kernels compute in T (mixed with FP64 accumulation in
// their factorisation chain); the API-faithful kernels compute in C = double whatever the batch precision (osc_kernels.cuh, Row).
template <class T, int N> struct OpsImpl : Ops {
	using C = double;
	ChainC<T, N> ch;
	ChainC<C, N> chc;

	static dim3 grid(const Batch& b) { return dim3((unsigned)((b.B + OSC_THREADS - 1) / OSC_THREADS)); }
	static Arena<T> rob(Batch& b) { return Arena<T>{(T*)b.P, (T*)b.S, b.I}; }
	static Arena<T> tk(TaskH& t) { return Arena<T>{(T*)t.P, (T*)t.S, t.I}; }
	int done(Batch& b, const char* what) {
		++b.launches;
		cudaError_t e = cudaGetLastError();
		if (e != cudaSuccess) return cuda_fail(e, what);
		return OSC_OK;
	}

	void robot_rows(int64_t& p, int64_t& s, int64_t& i) override { p = RobotL<N>::P_SIZE; s = RobotL<N>::S_SIZE; i = RobotL<N>::I_SIZE; }
	void task_rows(int type, int64_t& p, int64_t& s, int64_t& i) override {
		if (type == OSC_TASK_MOTION_FORCE) { p = MftL<N>::P_SIZE; s = MftL<N>::S_SIZE; i = MftL<N>::I_SIZE; }
		else { p = JtL<N>::P_SIZE; s = JtL<N>::S_SIZE; i = JtL<N>::I_SIZE; }
	}

	template <class X> static void fill_chain(const osc_chain_desc& c, ChainC<X, N>& ch) {
		for (int i = 0; i < N; ++i) {
			for (int a = 0; a < 3; ++a) {
				ch.xyz[i][a] = (X)c.joint_xyz[i][a];
				ch.com[i][a] = (X)c.com[i][a];
				for (int k = 0; k < 3; ++k) ch.rot[i][a][k] = (X)c.joint_rot[i][3 * a + k];
			}
			ch.prismatic[i] = c.joint_type[i] == OSC_JOINT_PRISMATIC;
			ch.mass[i] = (X)c.mass[i];
			const double* I = c.inertia[i];
			ch.inertia[i][0] = (X)I[0]; ch.inertia[i][1] = (X)I[1]; ch.inertia[i][2] = (X)I[2];
			ch.inertia[i][3] = (X)I[4]; ch.inertia[i][4] = (X)I[5]; ch.inertia[i][5] = (X)I[8];
			ch.q_lower[i] = (X)c.q_lower[i]; ch.q_upper[i] = (X)c.q_upper[i]; ch.effort[i] = (X)c.effort[i];
			const double m = c.mass[i], *cm = c.com[i];
			const double cc = cm[0] * cm[0] + cm[1] * cm[1] + cm[2] * cm[2];
			int e = 0;
			for (int a = 0; a < 3; ++a) {
				ch.hb[i][a] = (X)(m * cm[a]);
				for (int k = a; k < 3; ++k) ch.Ib[i][e++] = (X)(I[3 * a + k] + m * ((a == k ? cc : 0.0) - cm[a] * cm[k]));
			}
		}
		for (int a = 0; a < 3; ++a) {
			ch.world_xyz[a] = (X)c.world_xyz[a]; ch.gravity[a] = (X)c.gravity[a];
			for (int k = 0; k < 3; ++k) ch.world_rot[a][k] = (X)c.world_rot[3 * a + k];
		}
	}
	int set_chain(Batch& b) override {
		fill_chain<T>(b.chain, ch);
		fill_chain<C>(b.chain, chc);
		return OSC_OK;
	}

	MftC<T> mftc(const TaskH& t) { return mftc_<T>(t); }
	MftC<C> mftcc(const TaskH& t) { return mftc_<C>(t); }
	JtC<T, N> jtc(const Batch& b, const TaskH& t) { return jtc_<T>(b, t); }
	JtC<C, N> jtcc(const Batch& b, const TaskH& t) { return jtc_<C>(b, t); }
	template <class X> static MftC<X> mftc_(const TaskH& t) {
		MftC<X> c{};
		const osc_mft_params& p = t.mp;
		c.body = t.md.body; c.partial = t.md.partial; c.rank = t.dof; c.in_compliant_frame = t.md.parametrization_in_compliant_frame;
		for (int a = 0; a < 3; ++a) {
			c.p_c[a] = (X)t.frame_p[a];
			for (int k = 0; k < 3; ++k) { c.R_c[a][k] = (X)t.frame_R[a][k]; c.Pt[a][k] = (X)t.P6[a][k]; c.Pr[a][k] = (X)t.P6[3 + a][3 + k]; c.Bt[a][k] = (X)t.Bt[a][k]; c.Br[a][k] = (X)t.Br[a][k]; }
			c.kp_pos[a] = (X)p.kp_pos[a]; c.kv_pos[a] = (X)p.kv_pos[a]; c.ki_pos[a] = (X)p.ki_pos[a];
			c.kp_ori[a] = (X)p.kp_ori[a]; c.kv_ori[a] = (X)p.kv_ori[a]; c.ki_ori[a] = (X)p.ki_ori[a];
			c.force_axis[a] = (X)p.force_or_motion_axis[a]; c.moment_axis[a] = (X)p.moment_or_rotmotion_axis[a];
			c.sensor_p[a] = (X)p.sensor_xyz[a];
			for (int k = 0; k < 3; ++k) c.sensor_R[a][k] = (X)p.sensor_rot[3 * a + k];
		}
		c.mt = t.mt; c.mr = t.mr;
		c.dt = (X)t.md.loop_timestep;
		c.kp_force = (X)p.kp_force; c.kv_force = (X)p.kv_force; c.ki_force = (X)p.ki_force;
		c.kp_moment = (X)p.kp_moment; c.kv_moment = (X)p.kv_moment; c.ki_moment = (X)p.ki_moment;
		c.kff_force = (X)p.kff_force; c.kff_moment = (X)p.kff_moment;
		c.max_force_fb = (X)p.max_force_control_feedback_output; c.max_moment_fb = (X)p.max_moment_control_feedback_output;
		c.force_dim = p.force_space_dimension; c.moment_dim = p.moment_space_dimension;
		c.closed_loop_force = p.closed_loop_force_control; c.closed_loop_moment = p.closed_loop_moment_control;
		c.passivity = p.passivity_enabled; c.use_vel_sat = p.use_velocity_saturation;
		c.lin_sat = (X)p.linear_saturation_velocity; c.ang_sat = (X)p.angular_saturation_velocity;
		c.decoupling = p.dynamic_decoupling_type; c.bie_threshold = (X)p.bie_threshold;
		c.handling = p.singularity_handling_enabled; c.enforce_type_1 = p.enforce_type_1_strategy;
		c.s_min = (X)p.s_min; c.s_max = (X)p.s_max; c.kp1 = (X)p.kp_type_1; c.kv1 = (X)p.kv_type_1; c.kv2 = (X)p.kv_type_2;
		c.s_abs_tol = (X)p.s_abs_tol; c.type_1_tol = (X)p.type_1_tol; c.type_2_torque_ratio = (X)p.type_2_torque_ratio;
		c.type_2_angle_threshold = (X)p.type_2_angle_threshold; c.perturb_step = (X)p.perturb_step_size;
		c.buffer_size = p.buffer_size;
		c.apply_compensation = p.apply_disturbance_compensation;
		c.popc_ring = t.popc_ring;
		return c;
	}
	template <class X> static JtC<X, N> jtc_(const Batch& b, const TaskH& t) {
		JtC<X, N> c{};
		c.k = t.dof; c.full = t.jd.full; c.dt = (X)t.jd.loop_timestep;
		for (int i = 0; i < N; ++i)
			for (int j = 0; j < N; ++j)
				c.S[i][j] = t.jd.full ? (X)(i == j) : (i < t.dof ? (X)t.jd.selection[i * b.n + j] : (X)0);
		for (int i = 0; i < N; ++i) {
			c.kp[i] = (X)t.jp.kp[i]; c.kv[i] = (X)t.jp.kv[i]; c.ki[i] = (X)t.jp.ki[i]; c.sat[i] = (X)t.jp.saturation_velocity[i];
		}
		c.use_vel_sat = t.jp.use_velocity_saturation; c.decoupling = t.jp.dynamic_decoupling_type; c.bie_threshold = (X)t.jp.bie_threshold;
		return c;
	}

	int model(Batch& b, int dynamics) override {
		k_model<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(chc, rob(b), b.B, b.stride, dynamics, b.gravity ? 1 : 0);
		return done(b, "k_model");
	}
	int task_init(Batch& b, int t) override {
		TaskH& th = b.tasks[t];
		if (th.type == OSC_TASK_MOTION_FORCE) {
			k_mft_init<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(chc, tk(th), b.B, b.stride);
			if (int rc = done(b, "k_mft_init")) return rc;
			k_mft_reinit<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(mftcc(th), rob(b), tk(th), b.B, b.stride);
			return done(b, "k_mft_reinit");
		}
		k_jt_reinit<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(jtcc(b, th), rob(b), tk(th), b.B, b.stride, 1);
		return done(b, "k_jt_reinit");
	}
	int task_reinit(Batch& b, int t) override {
		TaskH& th = b.tasks[t];
		if (th.type == OSC_TASK_MOTION_FORCE) {
			k_mft_reinit<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(mftcc(th), rob(b), tk(th), b.B, b.stride);
			return done(b, "k_mft_reinit");
		}
		k_jt_reinit<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(jtcc(b, th), rob(b), tk(th), b.B, b.stride, 0);
		return done(b, "k_jt_reinit");
	}
	int mft_reset(Batch& b, int t, int flags) override {
		k_mft_reset<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(tk(b.tasks[t]), b.B, b.stride, flags);
		return done(b, "k_mft_reset");
	}
	const T* nprec_ptr(Batch& b, int src) {
		if (src < 0) return nullptr;
		TaskH& s = b.tasks[src];
		const int off = s.type == OSC_TASK_MOTION_FORCE ? MftL<N>::NTP : JtL<N>::NTP;
		return (const T*)s.S + (int64_t)off * b.stride;
	}
	int task_update(Batch& b, int t, int src, int no_history) override {
		TaskH& th = b.tasks[t];
		if (th.type == OSC_TASK_MOTION_FORCE) {
			k_mft_update<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(chc, mftcc(th), rob(b), tk(th), nprec_ptr(b, src), b.B, b.stride, no_history);
			return done(b, "k_mft_update");
		}
		k_jt_update<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(jtcc(b, th), rob(b), tk(th), nprec_ptr(b, src), b.B, b.stride);
		return done(b, "k_jt_update");
	}
	int task_compute(Batch& b, int t, int with_tau_prec, int accumulate) override {
		TaskH& th = b.tasks[t];
		if (th.type == OSC_TASK_MOTION_FORCE) {
			k_mft_compute<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(chc, mftcc(th), rob(b), tk(th), b.B, b.stride, with_tau_prec, accumulate);
			return done(b, "k_mft_compute");
		}
		k_jt_compute<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(jtcc(b, th), rob(b), tk(th), b.B, b.stride, with_tau_prec, accumulate);
		return done(b, "k_jt_compute");
	}
	int add_gravity(Batch& b) override {
		k_add_gravity<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(rob(b), b.B, b.stride);
		return done(b, "k_add_gravity");
	}
	int update_sensed(Batch& b, int t, const void* tmp6) override {
		TaskH& th = b.tasks[t];
		k_mft_sensed<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(chc, mftcc(th), rob(b), tk(th), (const T*)tmp6, b.B, b.stride, b.stale ? 1 : 0);
		return done(b, "k_mft_sensed");
	}

	int rollout_step(Batch& b, double dt, int substeps, double damping) override {
		bool revolute = true;
		for (int i = 0; i < N; ++i) revolute = revolute && !ch.prismatic[i];
#ifdef OSC_EXPERIMENT
		static const int ver = [] { const char* e = getenv("OSC_ROLLOUT_V"); return e ? atoi(e) : 2; }();
#else
		constexpr int ver = 2;
#endif
		if (revolute && ver == 2) {  // body-frame kernel with a shared-memory stash (osc_rollout.cuh); block shape as the fused tick
			auto go = [&](auto kern, int threads) -> int {
				const size_t smem = (size_t)R2<N>::SLOTS * threads * sizeof(T);
				if (first_launch_on_device((const void*)kern, b.device)) {
					cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
					if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(k_rollout_step2)");
				}
				kern<<<dim3((unsigned)((b.B + threads - 1) / threads)), threads, smem, b.stream>>>(ch, rob(b), b.B, b.stride, (T)dt, substeps, (T)damping, 1);
				return OSC_OK;
			};
			if (int rc = b.B >= 224 * 148 ? go(k_rollout_step2<T, N, 224>, 224) : go(k_rollout_step2<T, N, 64>, 64)) return rc;
			return done(b, "k_rollout_step2");
		}
		k_rollout_step<T, N><<<grid(b), OSC_THREADS, 0, b.stream>>>(ch, rob(b), b.B, b.stride, (T)dt, substeps, (T)damping, 1);
		return done(b, "k_rollout_step");
	}

	// ---- internal OTG (osc_otg.cuh) ----
	OtgJob otg_job(Batch& b, TaskH& th) {
		OtgJob j{};
		j.O = th.otgO; j.OI = th.otgI; j.G = th.otgG; j.P = th.P;
		j.status = b.I + (int64_t)RobotL<N>::STATUS * b.stride;
		j.k = th.type == OSC_TASK_MOTION_FORCE ? 6 : th.dof;
		j.rank = th.type == OSC_TASK_MOTION_FORCE ? nullptr : th.I + (int64_t)JtL<N>::RANK * b.stride;
		j.dt = b.dt;
		for (int i = 0; i < 8; ++i) { j.max_v[i] = th.otg_max_v[i]; j.max_a[i] = th.otg_max_a[i]; }
		return j;
	}
	void otg_rows(Batch& b, int t, int64_t& o, int64_t& i, int64_t& g) override {
		if (b.tasks[t].type == OSC_TASK_MOTION_FORCE) { o = OtgL<6>::O_SIZE; i = OtgL<6>::I_SIZE; g = MftL<N>::GOAL_FORCE; }
		else { o = OtgL<N>::O_SIZE; i = OtgL<N>::I_SIZE; g = 3 * N; }
	}
	int otg_reinit(Batch& b, int t, int mode) override {
		TaskH& th = b.tasks[t];
		th.otg_goals_dirty = true;
		const dim3 g((unsigned)((b.B + 127) / 128));
		if (th.type == OSC_TASK_MOTION_FORCE)
			k_otg_reinit<T, 6, true><<<g, 128, 0, b.stream>>>(otg_job(b, th), (const T*)th.P, b.B, b.stride, mode, (int)th.otg_g_rows);
		else
			k_otg_reinit<T, N, false><<<g, 128, 0, b.stream>>>(otg_job(b, th), (const T*)th.S + (int64_t)JtL<N>::CUR_POS * b.stride, b.B, b.stride, mode,
															   (int)th.otg_g_rows);
		return done(b, "k_otg_reinit");
	}
	// goal filter + update() + desired rows of task t (t >= 0) or of every task whose OTG is on (t < 0): two launches
	int otg_tick(Batch& b, int t) override {
		OtgJobs js{};
		int cnt = 0;
		for (int i = 0; i < (int)b.tasks.size(); ++i) {
			TaskH& th = b.tasks[i];
			if (!th.otg_on || (t >= 0 && i != t)) continue;
			if (!th.otg_slow) {
				cudaError_t e = cudaMalloc((void**)&th.otg_slow, (size_t)(b.stride + 2) * sizeof(int));
				if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc(otg slow list)");
				e = cudaMemsetAsync(th.otg_slow, 0, 2 * sizeof(int), b.stream);
				if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(otg slow counters)");
				th.otg_parity = 0;
			}
			OtgJob& j = js.j[cnt];
			j = otg_job(b, th);
			j.check_goals = th.otg_goals_dirty ? 1 : 0;
			j.slow_count = th.otg_slow + th.otg_parity;
			j.slow_next = th.otg_slow + (th.otg_parity ^ 1);
			j.slow_list = th.otg_slow + 2;
			js.cart[cnt] = th.type == OSC_TASK_MOTION_FORCE;
			th.otg_parity ^= 1;
			th.otg_goals_dirty = false;
			++cnt;
		}
		if (!cnt) return OSC_OK;
		bool goal_write = false;
		for (int i = 0; i < cnt; ++i) goal_write |= js.j[i].check_goals != 0;
#ifdef OSC_EXPERIMENT
		// measured A/B (profiles/README.md, r2f): the bulk-copy staged advance ties the plain kernel (27.3 vs 26.9 us at 65,536 robots,
		// 312 vs 301 us at 1,048,576), so the product build launches the plain kernel only
		static const int tma_mode = getenv("OSC_OTG_TMA") ? atoi(getenv("OSC_OTG_TMA")) : 0;
		if (!goal_write && b.B >= 4096 && tma_mode) {
			const int tiles_x = (int)((b.B + 127) / 128), n_tiles = tiles_x * cnt;
			const size_t smem = (size_t)OtgStage<N>::BYTES;
			auto kern = k_otg_advance_tma<T, N>;
			if (first_launch_on_device((const void*)kern, b.device)) {
				cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
				if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(k_otg_advance_tma)");
			}
			const int blocks = std::min(n_tiles, 148 * 4);
			kern<<<blocks, 128, smem, b.stream>>>(js, b.B, b.stride, tiles_x, n_tiles);
		} else
#endif
		{
			const dim3 g((unsigned)((b.B + 127) / 128), cnt);
			k_otg_advance<T, N><<<g, 128, 0, b.stream>>>(js, b.B, b.stride);
		}
		cudaLaunchConfig_t cfg = {};
		// a goal write may send every robot to the general path (MPC: new goals each tick): full grid then, else a small one
		bool any_goal_write = false;
		for (int i = 0; i < cnt; ++i) any_goal_write |= js.j[i].check_goals != 0;
		cfg.gridDim = dim3((unsigned)(any_goal_write ? (b.B + 127) / 128 : std::min<int64_t>((b.B + 127) / 128, 148 * 2)), cnt);
		cfg.blockDim = dim3(128);
		cfg.stream = b.stream;
		cudaLaunchAttribute at[1];
		at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
		at[0].val.programmaticStreamSerializationAllowed = 1;
#ifdef OSC_EXPERIMENT
		static const bool pdl = !getenv("OSC_NO_PDL");
#else
		constexpr bool pdl = true;
#endif
		cfg.attrs = at;
		cfg.numAttrs = pdl ? 1 : 0;
#ifdef OSC_EXPERIMENT
		{  // occupancy experiment: dynamic shared memory the kernel does not use limits the resident blocks per SM
			static const int smem = [] { const char* e = getenv("OSC_OTG_SLOW_SMEM"); return e ? atoi(e) : 0; }();
			static const int cap = [] { const char* e = getenv("OSC_OTG_SLOW_GRID"); return e ? atoi(e) : 0; }();
			if (smem > 0) {
				if (first_launch_on_device((const void*)k_otg_slow<T, N>, b.device)) cudaFuncSetAttribute(k_otg_slow<T, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
				cfg.dynamicSmemBytes = smem;
			}
			if (cap > 0) cfg.gridDim.x = std::min<unsigned>(cfg.gridDim.x, (unsigned)cap);
		}
#endif
		cudaError_t e = cudaLaunchKernelEx(&cfg, k_otg_slow<T, N>, js, b.stride);
		if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(k_otg_slow)");
		return done(b, "k_otg_slow");
	}

	bool locate(Batch& b, int task, int field, Loc& l) override;
	int pack_special(Batch& b, const Loc& l, void* tmp) override;
	bool fused_supported(Batch& b) override;
	int tick_fused(Batch& b, int stacked) override;
	int materialise(Batch& b) override;
};

#include "osc_ops_fields.inl"

}  // namespace

Ops* OSC_CAT(OSC_TNAME, OSC_N)() { return new OpsImpl<OSC_T, OSC_N>(); }

}  // namespace osc
