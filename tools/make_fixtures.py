"""Regenerate fixtures/*.chain.json from the reference's in-tree URDFs (run in the build container only;
/root/reference does not exist on the GPU box) and from the synthetic PUMA URDF shipped in fixtures/."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sai2-primitives-perso_b200"))
from urdf_chain import load_urdf, save_chain
REF = "/root/reference/examples"
jobs = {
    "panda": f"{REF}/15-haptic_control_impedance_type/panda_arm.urdf",
    "panda_heavy_ee": f"{REF}/16-haptic_control_admittance_type/panda_arm.urdf",
    "panda_sliding_base": f"{REF}/06-partial_joint_task/panda_arm_sliding_base.urdf",
    "rrrrbot": f"{REF}/11-planar_robot_controller/rrrrbot.urdf",
    "puma560_synthetic": f"{ROOT}/fixtures/puma560_synthetic.urdf",
}
for name, path in jobs.items():
    if not os.path.exists(path):
        print("skip", name); continue
    c = load_urdf(path)
    save_chain(c, f"{ROOT}/fixtures/{name}.chain.json")
    print(name, "n =", c["n"], "links:", list(c["links"]))
