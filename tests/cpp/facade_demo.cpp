// facade_demo.cpp -- the control loop of examples/05-using_robot_controller/05-using_robot_controller.cpp:103-196 written
// against the batched facade (include/Sai2PrimitivesBatched.h).  Built and driven by tests/test_facade_cpp.py:
//   facade_demo <chain.bin> <in.bin> <out.bin> <B> <ticks> <mode>
// chain.bin = raw osc_chain_desc; in.bin = q[B][n], dq[B][n], goal_pos[B][3], goal_ori[B][9], joint_goal[B][n] (double, AOS);
// out.bin = tau of the last tick [B][n], then the motion task's current position [B][3].
// mode 0: RobotController (stack B of SURVEY.md 3), the reference's three calls per cycle -- with the facade's control-cycle fusion (on
// by default) they become one fused launch, which the demo checks;  1: RobotController::tick (fused);  2: manual chaining of
// examples/04-task_and_redundancy/04-task_and_redundancy.cpp:137-206 (stack C, plain sum of the task torques);  3: mode 1 with the
// reference's default internal OTGs;  4: mode 0 with fusion off (API-faithful kernels);  5: mode 0 with a getter and a setQ between
// the calls of a cycle (the deferred updates are applied where the reference runs them);  6 (7: with fusion off): mode 0 with closed-loop force control
// along z and updateSensedForceAndMoment() inside every cycle (examples/09:176-182): the sensed wrench keeps the deferral and is
// rotated with the kinematics of the q that was just set.
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

#include "Sai2PrimitivesBatched.h"

using namespace Sai2PrimitivesBatched;

static std::vector<double> read_doubles(FILE* f, size_t n) {
	std::vector<double> v(n);
	if (fread(v.data(), sizeof(double), n, f) != n) { fprintf(stderr, "short read\n"); exit(3); }
	return v;
}

int main(int argc, char** argv) {
	if (argc < 7) { fprintf(stderr, "usage\n"); return 2; }
	osc_chain_desc chain;
	FILE* fc = fopen(argv[1], "rb");
	if (!fc || fread(&chain, sizeof chain, 1, fc) != 1) { fprintf(stderr, "cannot read chain\n"); return 2; }
	fclose(fc);
	const int64_t B = atoll(argv[4]);
	const int ticks = atoi(argv[5]), mode = atoi(argv[6]);
	const int n = chain.n;
	try {
		auto robot = std::make_shared<RobotBatch>(chain, B);
		FILE* fi = fopen(argv[2], "rb");
		if (!fi) { fprintf(stderr, "cannot read inputs\n"); return 2; }
		auto q = read_doubles(fi, B * n), dq = read_doubles(fi, B * n), gp = read_doubles(fi, B * 3), go = read_doubles(fi, B * 9),
			 gq = read_doubles(fi, B * n);
		fclose(fi);
		robot->setQ(hostAOS(q.data()));
		robot->setDq(hostAOS(dq.data()));
		robot->updateModel();
		// end-effector link = fixed link on the last body at (0, 0, 0.15) (panda_arm.urdf:179-183); control point (0, 0, 0.07)
		const Link ee(n - 1, Affine3::Translation(0, 0, 0.15));
		auto motion_force_task = std::make_shared<MotionForceTask>(robot, ee, Affine3::Translation(0, 0, 0.07), "cartesian_task");
		// mode 3 keeps the reference defaults (both internal OTGs on, MotionForceTask.h:67 / JointTask.h:38); the other modes run
		// OTG-off like examples/04...cpp:115
		if (!motion_force_task->getInternalOtgEnabled()) return 7;
		if (mode != 3) motion_force_task->disableInternalOtg();
		motion_force_task->setPosControlGains(120.0, 22.0);
		motion_force_task->setOriControlGains(PIDGains(210.0, 29.0, 0.0));
		auto joint_task = std::make_shared<JointTask>(robot, "joint_task");
		joint_task->setGains(55.0, 15.0);
		if (mode != 3) joint_task->disableInternalOtg();
		bool unsupported = false;
		try { joint_task->enableInternalOtgJerkLimited(1.0, 1.0, 1.0); } catch (const std::exception&) { unsupported = true; }
		if (!unsupported) return 8;
		std::vector<std::shared_ptr<TemplateTask>> task_list = {motion_force_task, joint_task};
		auto robot_controller = std::make_unique<RobotController>(robot, task_list);
		if (robot_controller->getTaskNames().size() != 2 || robot_controller->getJointTaskByName("joint_task") != joint_task) return 4;
		// error behaviour of the reference: std::invalid_argument
		bool thrown = false;
		try { motion_force_task->setPosControlGains(-1.0, 1.0); } catch (const std::invalid_argument&) { thrown = true; }
		if (!thrown) return 5;
		thrown = false;
		try { JointTask after_full(robot, "x"); } catch (const std::invalid_argument&) { thrown = true; }  // RobotController.cpp:36-41
		if (!thrown) return 6;
		motion_force_task->setGoalPosition(hostAOS(gp.data()));
		motion_force_task->setGoalOrientation(hostAOS(go.data()));
		motion_force_task->setGoalLinearVelocity({0.01, -0.02, 0.03});  // uniform over the batch
		joint_task->setGoalPosition(hostAOS(gq.data()));
		if (mode == 6 || mode == 7) {
			motion_force_task->parametrizeForceMotionSpaces(1, {0.0, 0.0, 1.0});
			motion_force_task->setClosedLoopForceControl(true);
			motion_force_task->setGoalForce({0.0, 0.0, 5.0});
		}
		std::vector<double> tau(B * n), x(B * 3);
		for (int t = 0; t < ticks; ++t) {
			if (mode == 1 || mode == 3) {
				robot_controller->tick(Out(tau.data()));
			} else if (mode == 0 || mode == 4 || mode == 5 || mode == 6 || mode == 7) {
				if (mode == 4 || mode == 7) robot->setControlCycleFusion(false);
				if (mode == 6 || mode == 7) {  // the robot moved: new q every cycle, then the force sensor reading of this cycle
					for (size_t i = 0; i < q.size(); ++i) q[i] += 1e-3 * ((i % 7) - 3.0);
					robot->setQ(hostAOS(q.data()));
				}
				robot->updateModel();
				if (mode == 6 || mode == 7) motion_force_task->updateSensedForceAndMoment({0.3, -0.2, 4.0 + 0.1 * t}, {0.01, 0.02, -0.03});
				if (mode == 5 && t == 1) motion_force_task->getCurrentPosition(Out(x.data()));  // a getter inside the cycle
				robot_controller->updateControllerTaskModels();
				if (mode == 5 && t == 2) robot->setQ(hostAOS(q.data()));  // same values: only the ordering is exercised
				motion_force_task->setGoalPosition(hostAOS(gp.data()));     // goal setters keep the deferral
				robot_controller->computeControlTorques(Out(tau.data()));
			} else {
				robot->updateModel();
				motion_force_task->updateTaskModel();
				joint_task->updateTaskModel(*motion_force_task);
				std::vector<double> t0(B * n), t1(B * n);
				motion_force_task->computeTorques(Out(t0.data()));
				joint_task->computeTorques(Out(t1.data()));
				for (size_t i = 0; i < tau.size(); ++i) tau[i] = t0[i] + t1[i];
			}
		}
		if ((mode == 0 || mode == 6) && robot->fusedControlCycles() != ticks) return 12;
		if (mode == 4 && robot->fusedControlCycles() != 0) return 13;
		if (mode == 5 && robot->fusedControlCycles() != ticks - 2) return 14;
		motion_force_task->getCurrentPosition(Out(x.data()));
		FILE* fo = fopen(argv[3], "wb");
		fwrite(tau.data(), sizeof(double), tau.size(), fo);
		fwrite(x.data(), sizeof(double), x.size(), fo);
		fclose(fo);
	} catch (const std::runtime_error& e) {
		fprintf(stderr, "runtime_error: %s\n", e.what());
		return 10;  // e.g. no CUDA device: the batched controller has no CPU fallback
	} catch (const std::exception& e) {
		fprintf(stderr, "exception: %s\n", e.what());
		return 11;
	}
	return 0;
}
