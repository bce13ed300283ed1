// facade_urdf_multi.cpp -- a C++ user of the facade who loads the robot from a URDF (+ world file) and drives it on SEVERAL devices
// through one RobotBatch.  Built and run by tests/test_facade_cpp.py:
//   facade_urdf_multi <robot.urdf> <world.urdf|-> <B> <n_shards> <out.bin>
// Loads the URDF with the library's reader, prints dof / links, then (needs a CUDA device; exit code 10 without one) runs the same
// three control ticks on a single-device batch and on a batch sharded over <n_shards> devices (device k % device_count) and
// writes both torque arrays ([B][n] each) to out.bin.  The control link is looked up BY NAME, like the reference API.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <memory>
#include <vector>

#include "Sai2PrimitivesBatched.h"

using namespace Sai2PrimitivesBatched;

static std::vector<double> run(const std::shared_ptr<RobotBatch>& robot, const UrdfModel& model, int64_t B) {
	const int n = model.dof();
	std::vector<double> q(B * n), tau(B * n);
	for (int64_t r = 0; r < B; ++r)
		for (int i = 0; i < n; ++i) q[r * n + i] = model.q0()[i] + 0.3 * std::sin(0.37 * (double)(r + 1) * (i + 1));
	robot->setQ(hostAOS(q.data()));
	robot->updateModel();
	auto task = std::make_shared<MotionForceTask>(robot, Link(model, "end-effector"), Affine3::Translation(0.02, 0, 0.05), "cartesian_task");
	task->disableInternalOtg();
	auto joint_task = std::make_shared<JointTask>(robot, "joint_task");
	joint_task->disableInternalOtg();
	std::vector<std::shared_ptr<TemplateTask>> tasks = {task, joint_task};
	RobotController controller(robot, tasks);
	std::vector<double> x(B * 3);
	task->getCurrentPosition(Out(x.data()));
	for (int64_t r = 0; r < B; ++r) x[3 * r + 2] += 0.05;
	task->setGoalPosition(hostAOS(x.data()));
	task->setGoalLinearVelocity({0.01, 0.0, -0.02});  // uniform
	for (int tick = 0; tick < 3; ++tick) {
		for (double& v : q) v += 1e-3;
		robot->setQ(hostAOS(q.data()));
		controller.tick(Out(tau.data()));
	}
	if (task->goalPositionReached(1.0) != B || task->goalPositionReached(1e-6) != 0) { fprintf(stderr, "goalPositionReached\n"); exit(5); }
	return tau;
}

int main(int argc, char** argv) {
	if (argc < 6) { fprintf(stderr, "usage\n"); return 2; }
	try {
		UrdfModel model(argv[1], strcmp(argv[2], "-") ? argv[2] : "");
		printf("dof %d gravity %.3f %.3f %.3f links %d\n", model.dof(), model.chain().gravity[0], model.chain().gravity[1], model.chain().gravity[2],
			   osc_urdf_link_count(model.handle()));
		const Link ee(model, "end-effector");
		printf("end-effector body %d\n", ee.body);
		bool threw = false;
		try { Link nope(model, "no-such-link"); } catch (const std::invalid_argument&) { threw = true; }
		if (!threw) return 4;
		const int64_t B = atoll(argv[3]);
		const int shards = atoi(argv[4]);
		if (osc_device_count() == 0) { fprintf(stderr, "no CUDA device: the batched controller has no CPU fallback\n"); return 10; }
		std::vector<int> devices;
		for (int k = 0; k < shards; ++k) devices.push_back(k % osc_device_count());
		auto one = std::make_shared<RobotBatch>(model, B);
		auto many = std::make_shared<RobotBatch>(model, B, devices);
		if (many->shardCount() != shards || many->size() != B) return 6;
		const std::vector<double> a = run(one, model, B), b = run(many, model, B);
		FILE* fo = fopen(argv[5], "wb");
		fwrite(a.data(), sizeof(double), a.size(), fo);
		fwrite(b.data(), sizeof(double), b.size(), fo);
		fclose(fo);
	} catch (const std::exception& e) {
		fprintf(stderr, "exception: %s\n", e.what());
		return 3;
	}
	return 0;
}
