"""Kinematic tables the hot path needs from the robot assets.

The reference gets these from Isaac Gym's asset loader at construction
(legged_robot.py:328-345 names/counts, :442-470 limits, :384-407 body indices).
Without PhysX the same facts are read once from the URDF `<limit>` tags and
link order (depth-first, fixed joints collapsed unless dont_collapse="true",
legged_robot_config.py:129) and tabulated here; tests/test_robot_tables.py
re-derives them from /root/reference/resources/robots/* when that tree exists.
"""
from dataclasses import dataclass, field
from typing import List


@dataclass
class RobotTable:
    name: str
    dof_names: List[str]
    dof_lower: List[float]
    dof_upper: List[float]
    dof_velocity: List[float]
    dof_effort: List[float]
    body_names: List[str]
    urdf: str = ""
    extra: dict = field(default_factory=dict)

    @property
    def num_dof(self):
        return len(self.dof_names)

    @property
    def num_bodies(self):
        return len(self.body_names)


def _legs(prefixes, joints):
    names, lo, hi, vel, eff = [], [], [], [], []
    for p in prefixes:
        for (j, a, b, v, e) in joints(p):
            names.append(j); lo.append(a); hi.append(b); vel.append(v); eff.append(e)
    return names, lo, hi, vel, eff


def _go2():
    def joints(p):
        rear = p.startswith('R')
        return [(f"{p}_hip_joint", -1.0472, 1.0472, 30.1, 23.7),
                (f"{p}_thigh_joint", -0.5236 if rear else -1.5708, 4.5379 if rear else 3.4907, 30.1, 23.7),
                (f"{p}_calf_joint", -2.7227, -0.83776, 20.07, 35.55)]
    legs = ["FL", "FR", "RL", "RR"]
    n, lo, hi, v, e = _legs(legs, joints)
    bodies = ["base", "Head_upper", "Head_lower"]
    for p in legs:
        bodies += [f"{p}_hip", f"{p}_thigh", f"{p}_calf", f"{p}_foot"]
    return RobotTable("go2", n, lo, hi, v, e, bodies, "resources/robots/go2/urdf/go2.urdf")


def _h1():
    def joints(p):
        return [(f"{p}_hip_yaw_joint", -0.43, 0.43, 23., 200.),
                (f"{p}_hip_roll_joint", -0.43, 0.43, 23., 200.),
                (f"{p}_hip_pitch_joint", -1.57, 1.57, 23., 200.),
                (f"{p}_knee_joint", -0.26, 2.05, 14., 300.),
                (f"{p}_ankle_joint", -0.87, 0.52, 9., 40.)]
    sides = ["left", "right"]
    n, lo, hi, v, e = _legs(sides, joints)
    bodies = ["pelvis"]
    for p in sides:
        bodies += [f"{p}_{k}_link" for k in ("hip_yaw", "hip_roll", "hip_pitch", "knee", "ankle")]
    return RobotTable("h1", n, lo, hi, v, e, bodies, "resources/robots/h1/urdf/h1.urdf")


def _g1():
    def joints(p):
        left = p == "left"
        return [(f"{p}_hip_pitch_joint", -2.5307, 2.8798, 32., 88.),
                (f"{p}_hip_roll_joint", -0.5236 if left else -2.9671, 2.9671 if left else 0.5236, 20., 139.),
                (f"{p}_hip_yaw_joint", -2.7576, 2.7576, 32., 88.),
                (f"{p}_knee_joint", -0.087267, 2.8798, 20., 139.),
                (f"{p}_ankle_pitch_joint", -0.87267, 0.5236, 37., 50.),
                (f"{p}_ankle_roll_joint", -0.2618, 0.2618, 37., 50.)]
    sides = ["left", "right"]
    n, lo, hi, v, e = _legs(sides, joints)
    bodies = ["pelvis"]
    for p in sides:
        bodies += [f"{p}_{k}_link" for k in ("hip_pitch", "hip_roll", "hip_yaw", "knee", "ankle_pitch", "ankle_roll")]
    return RobotTable("g1", n, lo, hi, v, e, bodies, "resources/robots/g1_description/g1_12dof.urdf")


def _h1_2():
    def joints(p):
        left = p == "left"
        return [(f"{p}_hip_yaw_joint", -0.43, 0.43, 23., 200.),
                (f"{p}_hip_pitch_joint", -3.14, 2.5, 23., 200.),
                (f"{p}_hip_roll_joint", -0.43 if left else -3.14, 3.14 if left else 0.43, 23., 200.),
                (f"{p}_knee_joint", -0.26, 2.05, 14., 300.),
                (f"{p}_ankle_pitch_joint", -0.897334, 0.523598, 9., 60.),
                (f"{p}_ankle_roll_joint", -0.261799, 0.261799, 9., 40.)]
    sides = ["left", "right"]
    n, lo, hi, v, e = _legs(sides, joints)
    bodies = ["pelvis"]
    for p in sides:
        bodies += [f"{p}_{k}_link" for k in ("hip_yaw", "hip_pitch", "hip_roll", "knee", "ankle_pitch", "ankle_roll")]
    return RobotTable("h1_2", n, lo, hi, v, e, bodies, "resources/robots/h1_2/h1_2_12dof.urdf")


ROBOTS = {"go2": _go2(), "h1": _h1(), "g1": _g1(), "h1_2": _h1_2()}
ROBOTS["go2_handstand"] = ROBOTS["go2"]      # task name -> table: the handstand task runs on the Go2 asset


def robot_for_asset(asset_file: str) -> RobotTable:
    """Map cfg.asset.file (e.g. '{LEGGED_GYM_ROOT_DIR}/resources/robots/go2/urdf/go2.urdf') to a table."""
    base = asset_file.replace("\\", "/").split("/")[-1]
    for r in ROBOTS.values():
        if r.urdf.endswith("/" + base):
            return r
    raise ValueError(f"no kinematic table for asset '{asset_file}'")


def parse_urdf(path: str) -> RobotTable:
    """Re-derive a table from a URDF file (used by tests and for new robots)."""
    import xml.etree.ElementTree as ET
    root = ET.parse(path).getroot()
    joints = root.findall("joint")
    by_parent = {}
    for j in joints:
        by_parent.setdefault(j.find("parent").get("link"), []).append(j)
    children = {j.find("child").get("link") for j in joints}
    base = [l.get("name") for l in root.findall("link") if l.get("name") not in children][0]
    bodies, dofs = [base], []

    def walk(link):
        for j in by_parent.get(link, []):
            child = j.find("child").get("link")
            kind = j.get("type")
            collapsed = kind == "fixed" and j.get("dont_collapse") != "true"
            if kind != "fixed":
                lim = j.find("limit")
                dofs.append((j.get("name"), float(lim.get("lower")), float(lim.get("upper")),
                             float(lim.get("velocity")), float(lim.get("effort"))))
            if not collapsed:
                bodies.append(child)
            walk(child)

    walk(base)
    return RobotTable(root.get("name", ""), [d[0] for d in dofs], [d[1] for d in dofs], [d[2] for d in dofs],
                      [d[3] for d in dofs], [d[4] for d in dofs], bodies, path)
