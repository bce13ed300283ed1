"""URDF -> serial-chain constants (host side, cold path).

Plays the role of `Sai2Model::Sai2Model(urdf)` (un-vendored sai2-model, see
SURVEY.md Appendix B): parse a URDF, weld the root link to the world, merge
links attached through fixed joints into their movable parent (mass, centre of
mass and inertia combined), and keep every link addressable by name as a frame
fixed in a movable body.  The reference reads its robots this way at
examples/04-task_and_redundancy/04-task_and_redundancy.cpp:67.

The result is a plain dict ("chain") that is written as JSON under fixtures/
and consumed by the oracles, the C ABI (`osc_chain_desc`) and the tests:

  n                      number of movable joints
  joints[i]              type ("revolute"|"prismatic"), axis (unit, joint frame),
                         xyz / rot (3x3 row-major) of the joint origin in the
                         parent *body* frame, lower/upper/effort/velocity, q0
  bodies[i]              mass, com (3), inertia (3x3 about the com, body frame)
  links[name]            body index, xyz, rot (3x3) of the link frame in the body
  world_xyz / world_rot  T_world_robot (identity unless a world file says otherwise)
  gravity                world gravity vector

Only serial chains are accepted (every movable joint's parent body is the
previous movable body, or the welded root).
"""
from __future__ import annotations

import json
import math
import xml.etree.ElementTree as ET

import numpy as np


def rpy_to_rot(rpy):
    """URDF fixed-axis roll/pitch/yaw: R = Rz(yaw) Ry(pitch) Rx(roll)."""
    r, p, y = rpy
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def _floats(s, n=3, default=0.0):
    if s is None:
        return [default] * n
    out = []
    for tok in s.split():
        try:
            out.append(float(tok))
        except ValueError:
            # the sliding-base fixture has a malformed "0.-75" (SURVEY.md C.1): read as 0
            out.append(0.0)
    return out


def _origin(elem):
    o = elem.find("origin") if elem is not None else None
    if o is None:
        return np.zeros(3), np.eye(3)
    return np.array(_floats(o.get("xyz"))), rpy_to_rot(_floats(o.get("rpy")))


def load_urdf(path, world_xyz=(0, 0, 0), world_rpy=(0, 0, 0), gravity=(0, 0, -9.81)):
    root = ET.parse(path).getroot()
    links = {}
    for l in root.findall("link"):
        inert = l.find("inertial")
        if inert is None:
            links[l.get("name")] = dict(mass=0.0, com=np.zeros(3), I=np.zeros((3, 3)))
            continue
        c, Rc = _origin(inert)
        m = float(inert.find("mass").get("value"))
        ie = inert.find("inertia")
        g = lambda k: float(ie.get(k, "0"))
        I = np.array([[g("ixx"), g("ixy"), g("ixz")], [g("ixy"), g("iyy"), g("iyz")], [g("ixz"), g("iyz"), g("izz")]])
        links[l.get("name")] = dict(mass=m, com=c, I=Rc @ I @ Rc.T)
    joints = []
    children = set()
    for j in root.findall("joint"):
        p, R = _origin(j)
        ax = j.find("axis")
        axis = np.array(_floats(ax.get("xyz"))) if ax is not None else np.array([1.0, 0, 0])
        lim = j.find("limit")
        cal = j.find("calibration")
        q0 = 0.0
        if cal is not None:
            if j.get("type") == "prismatic":
                q0 = float(cal.get("rising", cal.get("falling", "0")))
            else:
                q0 = math.radians(float(cal.get("falling", cal.get("rising", "0"))))
        joints.append(dict(name=j.get("name"), type=j.get("type"), parent=j.find("parent").get("link"),
                           child=j.find("child").get("link"), xyz=p, rot=R, axis=axis,
                           lower=float(lim.get("lower", "0")) if lim is not None else 0.0,
                           upper=float(lim.get("upper", "0")) if lim is not None else 0.0,
                           effort=float(lim.get("effort", "0")) if lim is not None else 0.0,
                           velocity=float(lim.get("velocity", "0")) if lim is not None else 0.0, q0=q0))
        children.add(j.find("child").get("link"))
    roots = [n for n in links if n not in children]
    if len(roots) != 1:
        raise ValueError("URDF must have exactly one root link")
    by_parent = {}
    for j in joints:
        by_parent.setdefault(j["parent"], []).append(j)

    chain_joints, bodies, link_frames = [], [], {}

    # walk the tree; body index -1 = welded root (does not enter M)
    def visit(link, body, p_in_body, R_in_body):
        link_frames[link] = dict(body=body, xyz=p_in_body.copy(), rot=R_in_body.copy())
        if body >= 0:
            L = links[link]
            if L["mass"] > 0:
                bodies[body]["parts"].append((L["mass"], p_in_body + R_in_body @ L["com"], R_in_body @ L["I"] @ R_in_body.T))
        movable_seen = 0
        for j in by_parent.get(link, []):
            p = p_in_body + R_in_body @ j["xyz"]
            R = R_in_body @ j["rot"]
            if j["type"] == "fixed":
                visit(j["child"], body, p, R)
            elif j["type"] in ("revolute", "continuous", "prismatic"):
                movable_seen += 1
                if len(chain_joints) != body + 1:
                    raise ValueError("not a serial chain (branching at link %s)" % link)
                ax = j["axis"] / np.linalg.norm(j["axis"])
                chain_joints.append(dict(name=j["name"], type="prismatic" if j["type"] == "prismatic" else "revolute",
                                         axis=ax, xyz=p, rot=R, lower=j["lower"], upper=j["upper"],
                                         effort=j["effort"], velocity=j["velocity"], q0=j["q0"]))
                bodies.append(dict(parts=[]))
                visit(j["child"], len(chain_joints) - 1, np.zeros(3), np.eye(3))
            else:
                raise ValueError("unsupported joint type " + j["type"])

    visit(roots[0], -1, np.zeros(3), np.eye(3))

    out_bodies = []
    for b in bodies:
        m = sum(p[0] for p in b["parts"])
        if m <= 0:
            out_bodies.append(dict(mass=0.0, com=[0.0] * 3, inertia=np.zeros((3, 3)).tolist()))
            continue
        c = sum(p[0] * p[1] for p in b["parts"]) / m
        I = np.zeros((3, 3))
        for pm, pc, pI in b["parts"]:
            d = pc - c
            I += pI + pm * (d @ d * np.eye(3) - np.outer(d, d))
        out_bodies.append(dict(mass=m, com=c.tolist(), inertia=I.tolist()))
    chain = dict(
        name=root.get("name", "robot"), n=len(chain_joints),
        joints=[dict(name=j["name"], type=j["type"], axis=j["axis"].tolist(), xyz=j["xyz"].tolist(),
                     rot=j["rot"].tolist(), lower=j["lower"], upper=j["upper"], effort=j["effort"],
                     velocity=j["velocity"], q0=j["q0"]) for j in chain_joints],
        bodies=out_bodies,
        links={k: dict(body=v["body"], xyz=v["xyz"].tolist(), rot=v["rot"].tolist()) for k, v in link_frames.items()},
        world_xyz=list(map(float, world_xyz)), world_rot=rpy_to_rot(world_rpy).tolist(),
        gravity=list(map(float, gravity)))
    return chain


def save_chain(chain, path):
    with open(path, "w") as f:
        json.dump(chain, f, indent=1)


def load_chain(path):
    with open(path) as f:
        return json.load(f)


if __name__ == "__main__":
    import sys
    save_chain(load_urdf(sys.argv[1]), sys.argv[2])
