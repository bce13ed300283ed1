// osc_layout.cuh -- HBM data layout of a batch (host + device).
//
// Every per-robot quantity is a structure-of-arrays row: element `e` of robot `r` lives at base[e * stride + r] with
// stride = B rounded up to 32, so the 32 threads of a warp (32 consecutive robots) read/write 32 consecutive scalars:
// fully coalesced 256-byte (FP64) transactions whatever the field.  Three arenas per owner (the robot itself, or one
// task): P = persistent reals (inputs, goals, integrators, observers: what osc_snapshot copies), S = scratch reals
// (recomputed every tick: kinematics, M, task models, torques), I = persistent ints.
#pragma once
#include <stdint.h>

namespace osc {

template <int N> struct RobotL {
	// P
	static constexpr int Q = 0, DQ = N, P_SIZE = 2 * N;
	// S
	static constexpr int TAU = 0, KR = TAU + N, KO = KR + 9 * N, M = KO + 3 * N, MINV = M + N * N, NTOT = MINV + N * N,
						 GRAV = NTOT + N * N, NLE = GRAV + N /* b(q,dq) + g(q), written by k_rollout_step */, S_SIZE = NLE + N;
	// I
	static constexpr int STATUS = 0, I_SIZE = 1;
};

template <int N> struct MftL {
	static constexpr int K = N < 6 ? N : 6;
	// P
	static constexpr int GOAL_POS = 0, GOAL_ORI = 3, GOAL_LINVEL = 12, GOAL_ANGVEL = 15, GOAL_LINACC = 18, GOAL_ANGACC = 21,
						 GOAL_FORCE = 24, GOAL_MOMENT = 27, SENSED_F = 30, SENSED_M = 33, CUR_POS = 36, CUR_ORI = 39,
						 INT_POS = 48, INT_ORI = 51, INT_F = 54, INT_M = 57, POPC_PO = 60, POPC_ECORR = 61, POPC_RC = 62,
						 POPC_VSQ = 63, Q_PRIOR = 64, DQ_PRIOR = Q_PRIOR + N, T2DIR = DQ_PRIOR + N, P_SIZE = T2DIR + N;
	// I
	static constexpr int POPC_COUNTER = 0, POPC_HEAD = 1, POPC_LEN = 2, NTYPES = 3, TYPES = 4, C1 = 5, C2 = 6, HIST_LEN = 7,
						 HIST_HEAD = 8, HIST = 9, INS = HIST + 8, I_SIZE = INS + 1;
	// S
	static constexpr int TASK_TAU = 0, CUR_V = TASK_TAU + N, CUR_W = CUR_V + 3, ORI_ERR = CUR_W + 3, UMF = ORI_ERR + 3,
						 DES_POS = UMF + 6, DES_ORI = DES_POS + 3, DES_LINVEL = DES_ORI + 9, DES_ANGVEL = DES_LINVEL + 3,
						 J = DES_ANGVEL + 3, JP = J + 6 * N, NT = JP + 6 * N, NPREC = NT + N * N, NTP = NPREC + N * N,
						 U_NS = NTP + N * N, J_NS = U_NS + 6 * K, L_NS = J_NS + K * N, L_NS_MOD = L_NS + K * K,
						 U_S = L_NS_MOD + K * K, V_S = U_S + 6 * K, J_S = V_S + N * K, L_S_MOD = J_S + K * N,
						 J_POST = L_S_MOD + K * K, L_JS_MOD = J_POST + K * N, ALPHA = L_JS_MOD + K * K, SV = ALPHA + 1,
						 FFB = SV + K,  // F_f of the control law, handed from k_tick_fused2 to k_tick_tail for robots that fall back
						 S_SIZE = FFB + 6;
};

template <int N> struct JtL {
	// P (task dof k <= N, rows padded to N)
	static constexpr int GOAL_POS = 0, GOAL_VEL = N, GOAL_ACC = 2 * N, INTEG = 3 * N, P_SIZE = 4 * N;
	// I
	static constexpr int RANK = 0, I_SIZE = 1;
	// S
	static constexpr int TASK_TAU = 0, CUR_POS = N, CUR_VEL = 2 * N, DES_VEL = 3 * N, NT = 4 * N, NPREC = NT + N * N,
						 NTP = NPREC + N * N, RANGE = NTP + N * N, LAM = RANGE + N * N, LAM_MOD = LAM + N * N,
						 S_SIZE = LAM_MOD + N * N;
};

// arena views handed to kernels
template <class T> struct Arena {
	T* P;
	T* S;
	int32_t* I;
	// k_tick_tail only: its robots are scattered over the batch, so the S rows (the ~1,200 intermediates a robot stores and reloads
	// between the API-faithful functions) are redirected to a compact scratch arena indexed by the tail THREAD (coalesced) instead of
	// by the robot (one 32-byte sector per 8-byte access: r2d capture, 180 KB of DRAM traffic per robot).  0 / -1 = the robot's own rows.
	int64_t s_stride = 0;
	int64_t s_index = -1;
};

}  // namespace osc
