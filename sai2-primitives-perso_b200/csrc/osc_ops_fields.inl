// osc_ops_fields.inl -- field id -> storage location (included inside the anonymous namespace of osc_ops_inst.cu)

template <class T, int N> bool OpsImpl<T, N>::locate(Batch& b, int task, int field, Loc& l) {
	l = Loc{};
	l.owner = task;
	auto set = [&](int arena, int off, int w, bool wr, bool is_int = false) {
		l.arena = arena; l.offset = off; l.width = w; l.writable = wr; l.is_int = is_int;
		return true;
	};
	if (task < 0) {
		using R = RobotL<N>;
		switch (field) {
			case OSC_F_Q: return set(0, R::Q, N, true);
			case OSC_F_DQ: return set(0, R::DQ, N, true);
			case OSC_F_TAU: return set(1, R::TAU, N, true);
			case OSC_F_M: return set(1, R::M, N * N, false);
			case OSC_F_MINV: return set(1, R::MINV, N * N, false);
			case OSC_F_GRAVITY: return set(1, R::GRAV, N, false);
			case OSC_F_NONLINEAR_EFFECTS: return set(1, R::NLE, N, false);
			case OSC_F_STATUS: return set(2, R::STATUS, 1, true, true);
			case OSC_F_N_TOTAL: {
				if (b.tasks.empty()) return false;
				const int last = (int)b.tasks.size() - 1;
				l.owner = last;
				return set(1, b.tasks[last].type == OSC_TASK_MOTION_FORCE ? MftL<N>::NTP : JtL<N>::NTP, N * N, false);
			}
		}
		return false;
	}
	if (task >= (int)b.tasks.size()) return false;
	const TaskH& t = b.tasks[task];
	if (t.type == OSC_TASK_MOTION_FORCE) {
		using L = MftL<N>;
		constexpr int K = L::K;
		switch (field) {
			case OSC_F_TASK_TORQUES: return set(1, L::TASK_TAU, N, false);
			case OSC_F_N_TASK: return set(1, L::NT, N * N, false);
			case OSC_F_N_PREC: return set(1, L::NPREC, N * N, false);
			case OSC_F_N_TASK_AND_PREC: return set(1, L::NTP, N * N, false);
			case OSC_F_GOAL_POSITION: return set(t.otg_on ? 3 : 0, L::GOAL_POS, 3, true);
			case OSC_F_GOAL_ORIENTATION: return set(t.otg_on ? 3 : 0, L::GOAL_ORI, 9, true);
			case OSC_F_GOAL_LINEAR_VELOCITY: return set(t.otg_on ? 3 : 0, L::GOAL_LINVEL, 3, true);
			case OSC_F_GOAL_ANGULAR_VELOCITY: return set(t.otg_on ? 3 : 0, L::GOAL_ANGVEL, 3, true);
			case OSC_F_GOAL_LINEAR_ACCELERATION: return set(t.otg_on ? 3 : 0, L::GOAL_LINACC, 3, true);
			case OSC_F_GOAL_ANGULAR_ACCELERATION: return set(t.otg_on ? 3 : 0, L::GOAL_ANGACC, 3, true);
			case OSC_F_GOAL_FORCE: return set(0, L::GOAL_FORCE, 3, true);
			case OSC_F_GOAL_MOMENT: return set(0, L::GOAL_MOMENT, 3, true);
			case OSC_F_SENSED_FORCE_WORLD: return set(0, L::SENSED_F, 3, true);
			case OSC_F_SENSED_MOMENT_WORLD: return set(0, L::SENSED_M, 3, true);
			case OSC_F_CURRENT_POSITION: return set(0, L::CUR_POS, 3, false);
			case OSC_F_CURRENT_ORIENTATION: return set(0, L::CUR_ORI, 9, false);
			case OSC_F_CURRENT_LINEAR_VELOCITY: return set(1, L::CUR_V, 3, false);
			case OSC_F_CURRENT_ANGULAR_VELOCITY: return set(1, L::CUR_W, 3, false);
			case OSC_F_ORIENTATION_ERROR: return set(1, L::ORI_ERR, 3, false);
			case OSC_F_UNIT_MASS_FORCE: return set(1, L::UMF, 6, false);
			case OSC_F_JACOBIAN: return set(1, L::J, 6 * N, false);
			case OSC_F_PROJECTED_JACOBIAN: return set(1, L::JP, 6 * N, false);
			case OSC_F_LAMBDA_WORLD: l.special = 1; return set(1, 0, 36, false);
			case OSC_F_LAMBDA_MODIFIED_WORLD: l.special = 2; return set(1, 0, 36, false);
			case OSC_F_ALPHA: return set(1, L::ALPHA, 1, false);
			case OSC_F_SINGULAR_VALUES: l.special = 5; return set(1, L::SV, 6, false);
			case OSC_F_NS_RANK: return set(2, L::INS, 1, false, true);
			case OSC_F_SINGULARITY_TYPES: l.special = 3; return set(2, 0, 6, false, true);
			case OSC_F_SINGULARITY_COUNTERS: l.special = 4; return set(2, 0, 2, false, true);
			case OSC_F_INTEGRATED_POSITION_ERROR: return set(0, L::INT_POS, 3, true);
			case OSC_F_INTEGRATED_ORIENTATION_ERROR: return set(0, L::INT_ORI, 3, true);
			case OSC_F_INTEGRATED_FORCE_ERROR: return set(0, L::INT_F, 3, true);
			case OSC_F_INTEGRATED_MOMENT_ERROR: return set(0, L::INT_M, 3, true);
			case OSC_F_POPC_RC: return set(0, L::POPC_RC, 1, false);
			case OSC_F_POPC_OBSERVER: return set(0, L::POPC_PO, 1, false);
			case OSC_F_Q_PRIOR: return set(0, L::Q_PRIOR, N, true);
			case OSC_F_DESIRED_POSITION: return set(1, L::DES_POS, 3, false);
			case OSC_F_DESIRED_ORIENTATION: return set(1, L::DES_ORI, 9, false);
			case OSC_F_DESIRED_LINEAR_VELOCITY: return set(1, L::DES_LINVEL, 3, false);
			case OSC_F_DESIRED_ANGULAR_VELOCITY: return set(1, L::DES_ANGVEL, 3, false);
			case OSC_F_DESIRED_LINEAR_ACCELERATION: return set(0, L::GOAL_LINACC, 3, false);
			case OSC_F_DESIRED_ANGULAR_ACCELERATION: return set(0, L::GOAL_ANGACC, 3, false);
			case OSC_F_SIGMA_FORCE: case OSC_F_SIGMA_POSITION: case OSC_F_SIGMA_MOMENT: case OSC_F_SIGMA_ORIENTATION:
				l.special = 10 + (field - OSC_F_SIGMA_FORCE); return set(1, 0, 9, false);
			case OSC_F_POSITION_ERROR: case OSC_F_SELECTED_ORIENTATION_ERROR: case OSC_F_GOAL_FORCE_WORLD: case OSC_F_GOAL_MOMENT_WORLD:
				l.special = 10 + (field - OSC_F_SIGMA_FORCE); return set(1, 0, 3, false);
			case OSC_F_POSITION_ERROR_NORM: case OSC_F_ORIENTATION_ERROR_NORM:
				l.special = 10 + (field - OSC_F_SIGMA_FORCE); return set(1, 0, 1, false);
			case OSC_F_OTG_FLAGS: if (!t.otgI) return false; return set(4, OtgL<6>::FLAGS, 1, false, true);
			case OSC_F_OTG_RESULT: if (!t.otgI) return false; return set(4, OtgL<6>::RESULT, 1, false, true);
			case OSC_F_OTG_STATE: if (!t.otgO || sizeof(T) != 8) return false; return set(5, 0, OtgL<6>::O_SIZE, false);
		}
		(void)K;
		return false;
	}
	using L = JtL<N>;
	const int k = t.dof;
	switch (field) {
		case OSC_F_TASK_TORQUES: return set(1, L::TASK_TAU, N, false);
		case OSC_F_N_TASK: return set(1, L::NT, N * N, false);
		case OSC_F_N_PREC: return set(1, L::NPREC, N * N, false);
		case OSC_F_N_TASK_AND_PREC: return set(1, L::NTP, N * N, false);
		case OSC_F_JT_GOAL_POSITION: return set(t.otg_on ? 3 : 0, L::GOAL_POS, k, true);
		case OSC_F_JT_GOAL_VELOCITY: return set(t.otg_on ? 3 : 0, L::GOAL_VEL, k, true);
		case OSC_F_JT_GOAL_ACCELERATION: return set(t.otg_on ? 3 : 0, L::GOAL_ACC, k, true);
		case OSC_F_JT_CURRENT_POSITION: return set(1, L::CUR_POS, k, false);
		case OSC_F_JT_CURRENT_VELOCITY: return set(1, L::CUR_VEL, k, false);
		case OSC_F_JT_INTEGRATED_ERROR: return set(0, L::INTEG, k, true);
		case OSC_F_JT_DESIRED_VELOCITY: return set(1, L::DES_VEL, k, false);
		case OSC_F_JT_LAMBDA_WORLD: l.special = 6; return set(1, 0, k * k, false);
		case OSC_F_JT_LAMBDA_MODIFIED_WORLD: l.special = 7; return set(1, 0, k * k, false);
		case OSC_F_JT_RANGE_RANK: return set(2, L::RANK, 1, false, true);
		// _desired_position / _desired_acceleration: the rows the control law reads (= the goal rows while the OTG is off)
		case OSC_F_JT_DESIRED_POSITION: return set(0, L::GOAL_POS, k, false);
		case OSC_F_JT_DESIRED_ACCELERATION: return set(0, L::GOAL_ACC, k, false);
		case OSC_F_JT_POSITION_ERROR_NORM: l.special = 20; return set(1, 0, 1, false);
		case OSC_F_OTG_FLAGS: if (!t.otgI) return false; return set(4, OtgL<N>::FLAGS, 1, false, true);
		case OSC_F_OTG_RESULT: if (!t.otgI) return false; return set(4, OtgL<N>::RESULT, 1, false, true);
		case OSC_F_OTG_STATE: if (!t.otgO || sizeof(T) != 8) return false; return set(5, 0, OtgL<N>::O_SIZE, false);
	}
	return false;
}

// tmp: SoA [width][stride] of T (or int32)
template <class T, int N> int OpsImpl<T, N>::pack_special(Batch& b, const Loc& l, void* tmp) {
	TaskH& t = b.tasks[l.owner];
	constexpr int K = MftL<N>::K;
	switch (l.special) {
		case 1:
		case 2:
			k_pack_ulut<C, 6, K, T><<<grid(b), OSC_THREADS, 0, b.stream>>>((const T*)t.S, MftL<N>::U_NS, l.special == 1 ? MftL<N>::L_NS : MftL<N>::L_NS_MOD,
																		 (T*)tmp, b.B, b.stride, 6);
			return done(b, "k_pack_ulut");
		case 3:
		case 4:
			k_pack_types<N><<<grid(b), OSC_THREADS, 0, b.stream>>>(t.I, (int32_t*)tmp, b.B, b.stride, l.special == 3 ? 0 : 1);
			return done(b, "k_pack_types");
		case 5: {  // singular values zero padded to 6
			cudaError_t e = cudaMemsetAsync(tmp, 0, (size_t)6 * b.stride * sizeof(T), b.stream);
			if (e != cudaSuccess) return cuda_fail(e, "memset");
			e = cudaMemcpyAsync(tmp, (const T*)t.S + (int64_t)MftL<N>::SV * b.stride, (size_t)K * b.stride * sizeof(T), cudaMemcpyDeviceToDevice, b.stream);
			if (e != cudaSuccess) return cuda_fail(e, "memcpy");
			return OSC_OK;
		}
		case 6:
		case 7:
			k_pack_ulut<C, N, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>((const T*)t.S, JtL<N>::RANGE, l.special == 6 ? JtL<N>::LAM : JtL<N>::LAM_MOD,
																		 (T*)tmp, b.B, b.stride, t.dof);
			return done(b, "k_pack_ulut");
	}
	if (l.special >= 10 && l.special <= 19) {
		k_pack_mft_getters<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>(mftcc(t), (const T*)t.P, (const T*)t.S, (T*)tmp, b.B, b.stride, l.special - 10);
		return done(b, "k_pack_mft_getters");
	}
	if (l.special == 20) {
		k_pack_jt_error_norm<C, N, T><<<grid(b), OSC_THREADS, 0, b.stream>>>((const T*)t.P, (const T*)t.S, t.I, (T*)tmp, b.B, b.stride, t.dof);
		return done(b, "k_pack_jt_error_norm");
	}
	set_error("internal: unknown packed field");
	return OSC_ERR_INVALID_ARGUMENT;
}

template <class T, int N> bool OpsImpl<T, N>::fused_supported(Batch& b) { return fused_tick_supported<T, N>(b); }
template <class T, int N> int OpsImpl<T, N>::tick_fused(Batch& b, int stacked) {
	if (joint_only_hierarchy(b)) {  // examples/01, 02: no motion task, no nullspace, no fallback list
		TaskH& j = b.tasks[0];
		const int exports = b.tick_outputs == OSC_TICK_OUTPUTS_TAU ? 0 : OSC_EXPORT_VECTORS;
		bool revolute = true;
		for (int i = 0; i < N; ++i) revolute = revolute && !ch.prismatic[i];
		if (revolute) {  // body-frame kernel with a shared-memory stash (osc_fused_jt.cuh); block shape as the fused tick
			if (b.gravity) {
				k_gravity<T, N><<<grid(b), OSC_THREADS, 0, b.stream>>>(ch, rob(b), b.B, b.stride);
				if (int rc = done(b, "k_gravity")) return rc;
			}
			auto go = [&](auto kern, int threads) -> int {
				const size_t smem = (size_t)FJ<N>::SLOTS * threads * sizeof(T);
				if (smem > 48 * 1024 && first_launch_on_device((const void*)kern, b.device)) {
					cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
					if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(k_tick_joint2)");
				}
				kern<<<dim3((unsigned)((b.B + threads - 1) / threads)), threads, smem, b.stream>>>(ch, jtc(b, j), rob(b), tk(j), b.B, b.stride, b.gravity ? 1 : 0, exports);
				return OSC_OK;
			};
			if (int rc = b.B >= 224 * 148 ? go(k_tick_joint2<T, N, 224>, 224) : go(k_tick_joint2<T, N, 64>, 64)) return rc;
			if (int rc = done(b, "k_tick_joint2")) return rc;
		} else {
			k_tick_joint<T, N><<<grid(b), OSC_THREADS, 0, b.stream>>>(ch, jtc(b, j), rob(b), tk(j), b.B, b.stride, b.gravity ? 1 : 0, exports);
			if (int rc = done(b, "k_tick_joint")) return rc;
		}
		b.stale = true;
		return OSC_OK;
	}
	if constexpr (N == 6 || N == 7) {
		if (!b.fb) {
			cudaError_t e = cudaMalloc((void**)&b.fb, (size_t)(b.stride + 2) * sizeof(int));
			if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc(fallback list)");
			e = cudaMemsetAsync(b.fb, 0, 2 * sizeof(int), b.stream);
			if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(fallback counters)");
			b.fb_parity = 0;
		}
		const bool single = b.tasks.size() == 1;  // one MotionForceTask is the whole hierarchy: every joint-task argument below aliases the motion task and is never used
		TaskH &m = b.tasks[0], &j = single ? b.tasks[0] : b.tasks[1];
		int* cnt = b.fb + b.fb_parity;
		int* nxt = b.fb + (b.fb_parity ^ 1);
		int* list = b.fb + 2;
		const int exports = (b.tick_outputs == OSC_TICK_OUTPUTS_TAU ? 0 : OSC_EXPORT_VECTORS) | (single ? OSC_EXPORT_NO_JT : 0);
		const dim3 g((unsigned)((b.B + OSC_FUSED_THREADS - 1) / OSC_FUSED_THREADS));
		const MftC<T> tc = mftc(m);
		const JtC<T, N> jc = jtc(b, j);
		// Experiment knobs (kernel generation, occupancy / block-shape variants, no-PDL, no-tail diagnosis) exist only in
		// builds made with -DOSC_EXPERIMENT (profiles/ runs); the product build has none of them and reads no environment.
#ifdef OSC_EXPERIMENT
		static const int ver = [] { const char* e = getenv("OSC_FUSED_V"); return e ? atoi(e) : 2; }();
		static const int bt_env = [] { const char* e = getenv("OSC_FUSED_BT"); return e ? atoi(e) : 0; }();
		static const bool nosimple = getenv("OSC_FUSED_NOSIMPLE") != nullptr, debug_occ = getenv("OSC_DEBUG_OCC") != nullptr;
		static const bool no_tail = getenv("OSC_DIAG_NO_TAIL") != nullptr, pdl = !getenv("OSC_NO_PDL");
#else
		constexpr int ver = 2, bt_env = 0;
		constexpr bool nosimple = false, debug_occ = false, no_tail = false, pdl = true;
#endif
#define OSC_FUSED_ARGS ch, tc, jc, rob(b), tk(m), tk(j), b.B, b.stride, stacked, b.gravity ? 1 : 0, exports, cnt, list
		int reuse_law = 0;
		int M3 = 0;
		if (fused3_supported(b, M3)) {
			if constexpr (N == 7) {
				reuse_law = 1;
				if (b.gravity) {
					k_gravity<T, N><<<grid(b), OSC_THREADS, 0, b.stream>>>(ch, rob(b), b.B, b.stride);
					if (int rc = done(b, "k_gravity")) return rc;
				}
				auto go3 = [&](auto kern, int threads, int slots) -> int {
					const size_t smem = (size_t)slots * threads * sizeof(double);
					if (first_launch_on_device((const void*)kern, b.device)) {
						cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
						if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(k_tick_fused3)");
					}
					const dim3 gg((unsigned)((b.B + threads - 1) / threads));
					kern<<<gg, threads, smem, b.stream>>>(OSC_FUSED_ARGS);
					return OSC_OK;
				};
				const bool big = b.B >= 224 * 148;
				int rc = OSC_ERR_UNSUPPORTED;
				// block shape / register cap, measured on config 3 (262,144 robots, M = 3; r2i): 128 registers x 224 threads 0.53 ms per
				// tick (0.4-1.4 KB of spills per thread, the kernel was bound by that local-memory traffic), 168 x 192 0.47 ms,
				// 255 x 128 0.45 ms: unlike k_tick_fused2 this kernel wants the registers more than the resident warps
#define OSC_F3_CASE(MM)                                                                                                                    \
				if (M3 == MM) rc = big ? go3(k_tick_fused3<T, N, MM, 255, 128, true>, 128, F3<N, MM>::SLOTS) : go3(k_tick_fused3<T, N, MM, 128>, 64, F3<N, MM>::SLOTS);
				OSC_F3_CASE(2) OSC_F3_CASE(3) OSC_F3_CASE(4) OSC_F3_CASE(5)
#undef OSC_F3_CASE
				if (rc) return rc;
			}
		} else if (ver == 2 && fused2_supported(b)) {
			reuse_law = 1;
			if (b.gravity) {
				k_gravity<T, N><<<grid(b), OSC_THREADS, 0, b.stream>>>(ch, rob(b), b.B, b.stride);
				if (int rc = done(b, "k_gravity")) return rc;
			}
			const bool simple = fused2_simple(b) && !nosimple;
			if (simple && N == 7 && !single) reuse_law |= 2;  // the joint PID error is integrated at the top of the SIMPLE kernel
			// block shape: 224 threads with phase barriers once the batch fills two such blocks per SM (r1l: 44.7 us against
			// 45.7 us for 64-thread blocks at 65,536 robots), 64-thread blocks below that so that small batches still spread
			// over the SMs
			const int bt = bt_env ? bt_env : (b.B >= 224 * 148 ? 224 : 64);
			auto go = [&](auto kern, int threads) -> int {
				const size_t smem = (size_t)F2<N>::SLOTS * threads * sizeof(double);  // the stash is FP64 for both batch precisions
				if (first_launch_on_device((const void*)kern, b.device)) {  // once per kernel and device, not per tick
					cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
					if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(k_tick_fused2)");
				}
				if (debug_occ) {
					int nb = 0;
					cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, threads, smem);
					fprintf(stderr, "k_tick_fused2<%d>: %d threads/block, %d blocks/SM, %d stash slots\n", N, threads, nb, F2<N>::SLOTS);
				}
				const dim3 gg((unsigned)((b.B + threads - 1) / threads));
				kern<<<gg, threads, smem, b.stream>>>(OSC_FUSED_ARGS);
				return OSC_OK;
			};
			int rc = OSC_OK;
			bool launched = false;
			if constexpr (N == 7) {
				if (single) {  // the motion task is the whole hierarchy (examples/07): SINGLE instantiations, same block shapes
					launched = true;
					if (bt == 224) rc = simple ? go(k_tick_fused2<T, N, 128, true, 224, true, true>, 224) : go(k_tick_fused2<T, N, 128, false, 224, true, true>, 224);
					else rc = simple ? go(k_tick_fused2<T, N, 128, true, OSC_FUSED_THREADS, false, true>, OSC_FUSED_THREADS)
									 : go(k_tick_fused2<T, N, 128, false, OSC_FUSED_THREADS, false, true>, OSC_FUSED_THREADS);
				}
			}
			if (launched) {}
			else if (bt == 224) rc = simple ? go(k_tick_fused2<T, N, 128, true, 224, true>, 224) : go(k_tick_fused2<T, N, 128, false, 224, true>, 224);
#ifdef OSC_EXPERIMENT
			else if (bt == 65) rc = simple ? go(k_tick_fused2<T, N, 128, true, 64, true>, 64) : go(k_tick_fused2<T, N, 128, false, 64, true>, 64);
			else if (bt == 448) rc = simple ? go(k_tick_fused2<T, N, 128, true, 448, true>, 448) : go(k_tick_fused2<T, N, 128, false, 448, true>, 448);
#endif
			else rc = simple ? go(k_tick_fused2<T, N, 128, true>, 64) : go(k_tick_fused2<T, N, 128, false>, 64);
			if (rc) return rc;
		} else {
			// first-generation kernel (osc_fused.cuh): control link not on the last body, or a prismatic joint in the chain
			k_tick_fused<T, N, 8><<<g, OSC_FUSED_THREADS, 0, b.stream>>>(OSC_FUSED_ARGS);
		}
#undef OSC_FUSED_ARGS
		if (int rc = done(b, "k_tick_fused")) return rc;
		const unsigned tail_blocks = (unsigned)std::min<int64_t>((b.B + OSC_THREADS - 1) / OSC_THREADS, 148 * 2);
		if (!b.tail_scratch && b.may_allocate) {  // one compact slot of S rows per tail thread (see k_tick_tail)
			const int64_t slots = ((int64_t)tail_blocks * OSC_THREADS + 31) / 32 * 32;
			const size_t rows = (size_t)RobotL<N>::S_SIZE + MftL<N>::S_SIZE + JtL<N>::S_SIZE;
			if (cudaMalloc(&b.tail_scratch, rows * slots * sizeof(T)) == cudaSuccess) b.tail_scratch_stride = slots;
			else { cudaGetLastError(); b.tail_scratch = nullptr; }  // not fatal: the tail then works on the robots' own rows
		}
		if (!no_tail) {  // (no_tail: diagnosis builds only -- fallback robots get no torques in such a run)
			cudaLaunchConfig_t cfg = {};
			cfg.gridDim = dim3(tail_blocks);
			cfg.blockDim = dim3(OSC_THREADS);
			cfg.stream = b.stream;
			cudaLaunchAttribute at[1];
			at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
			at[0].val.programmaticStreamSerializationAllowed = 1;
			cfg.attrs = at;
			cfg.numAttrs = pdl ? 1 : 0;
			// closed-form blend routine ahead of the API-faithful chain (osc_blend.cuh): n = 7, task rank 2..6 (3 = translation-only,
			// config 3; 6 = config 4), the fused kernel ran the control law, singularity handling on, no IMPEDANCE decoupling
			int mb = 0;
			if constexpr (N == 7) {
				if ((reuse_law & 1) && m.mp.singularity_handling_enabled && m.mp.dynamic_decoupling_type != OSC_IMPEDANCE && m.dof >= 2 && m.dof <= 6 &&
					(m.dof < 6) == (m.md.partial != 0))
					mb = m.dof;
#ifdef OSC_EXPERIMENT
				static const bool no_blend = getenv("OSC_NO_BLEND") != nullptr;
				if (no_blend) mb = 0;
#endif
			}
			cudaError_t e = cudaSuccess;
			auto go_tail = [&](auto kern, size_t smem) {
				if (smem > 48 * 1024 && first_launch_on_device((const void*)kern, b.device)) {
					e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
					if (e != cudaSuccess) return;
				}
				cfg.dynamicSmemBytes = smem;
				e = cudaLaunchKernelEx(&cfg, kern, chc, mftcc(m), jtcc(b, j), rob(b), tk(m), tk(j), b.stride, stacked, b.gravity ? 1 : 0,
									   (const int*)cnt, nxt, (const int*)list, reuse_law, (T*)b.tail_scratch, b.tail_scratch_stride, exports);
			};
			if constexpr (N == 7) {
				if (mb == 6) go_tail(k_tick_tail<C, N, T, 6>, (size_t)BL<N, 6>::SLOTS * OSC_THREADS * sizeof(C));
				else if (mb == 5) go_tail(k_tick_tail<C, N, T, 5>, (size_t)BL<N, 5>::SLOTS * OSC_THREADS * sizeof(C));
				else if (mb == 4) go_tail(k_tick_tail<C, N, T, 4>, (size_t)BL<N, 4>::SLOTS * OSC_THREADS * sizeof(C));
				else if (mb == 3) go_tail(k_tick_tail<C, N, T, 3>, (size_t)BL<N, 3>::SLOTS * OSC_THREADS * sizeof(C));
				else if (mb == 2) go_tail(k_tick_tail<C, N, T, 2>, (size_t)BL<N, 2>::SLOTS * OSC_THREADS * sizeof(C));
				else go_tail(k_tick_tail<C, N, T, 0>, 0);
			} else if constexpr (N == 6) {
				if ((reuse_law & 1) && m.mp.singularity_handling_enabled && m.mp.dynamic_decoupling_type != OSC_IMPEDANCE && m.dof == 6 && !m.md.partial)
					mb = 6;
#ifdef OSC_EXPERIMENT
				static const bool no_blend6 = getenv("OSC_NO_BLEND") != nullptr;
				if (no_blend6) mb = 0;
#endif
				if (mb == 6) go_tail(k_tick_tail<C, N, T, 6>, (size_t)BL<N, 6>::SLOTS * OSC_THREADS * sizeof(C));
				else go_tail(k_tick_tail<C, N, T, 0>, 0);
			} else {
				go_tail(k_tick_tail<C, N, T, 0>, 0);
			}
			if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(k_tick_tail)");
		}
		if (int rc = done(b, "k_tick_tail")) return rc;
		b.fb_last = b.fb_parity;
		b.fb_parity ^= 1;
		b.stale = true;
		return OSC_OK;
	} else {
		(void)stacked;
		set_error("the fused tick is not built for this dof");
		return OSC_ERR_UNSUPPORTED;
	}
}
// The getters of the reference that the fused tick did not write: recompute the model and the task models with the
// per-call kernels at the current q.  No persistent state advances (k_mft_update runs with no_history while b.stale).
template <class T, int N> int OpsImpl<T, N>::materialise(Batch& b) {
	if (!b.stale) return OSC_OK;
	if (int rc = model(b, 1)) return rc;
	for (int t = 0; t < (int)b.tasks.size(); ++t)
		if (int rc = task_update(b, t, t - 1, 1)) return rc;
	b.stale = false;
	return OSC_OK;
}
