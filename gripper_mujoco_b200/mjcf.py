"""MJCF export of the model this package simulates: the only route to pin the physics of the hot
path against real MuJoCo, which is not on this box. `export_mjcf` writes the kinematic tree, inertias, joint
springs / dampers / limits, collision primitives, the seven welds, one force actuator per controlled joint and
(optionally) one live object with MuJoCo's defaults left as the reference leaves them (description/mujoco/xacro/
mujoco.xacro:6-12: Euler, pyramidal cone, solref 0.02 1, solimp 0.9 0.95 0.001 0.5 2, balanceinertia).

Conventions match `oracle/physics.c::or_kinematics` and the CUDA kernels: a body frame is the parent frame moved by
(`pos`, `quat`); its single joint sits at the body origin (slide along / hinge about `axis`, body frame).

Contact filtering reproduces the pairs the batched path evaluates (gripper-object, gripper-ground, object-ground,
no gripper-gripper): gripper geoms contype 1 / conaffinity 0, object 2 / 1, ground 0 / 3.

`tools/mujoco_pin.py` (run where `import mujoco` works) steps this file under the same PD control law and writes
the trajectories `tests/test_mjcf.py` compares with the oracle when the fixture is present. NOT validated against
MuJoCo in this container (library absent): see DESIGN.md section 3.
"""
from __future__ import annotations

import xml.etree.ElementTree as ET

import numpy as np

from .model import GEOM_BOX, GEOM_CYLINDER, GEOM_ELLIPSOID, Model, mat_to_quat, quat_to_mat
from .objects import ObjectSpec

WELD_NAMES = ["pris1_weld", "pris2_weld", "pris3_weld", "rev1_weld", "rev2_weld", "rev3_weld", "palm_weld"]


def _v(a) -> str:
    return " ".join(f"{float(x):.10g}" for x in np.asarray(a).reshape(-1))


def controlled_joints(mdl: Model) -> list[int]:
    """Dof indices that carry a PD force (luke::control, src/myfunctions.cpp:1912-2057): base z, the three
    prismatic and three revolute motor joints, the palm slide."""
    N = mdl.N
    out = [0]
    for f in range(3):
        out += [1 + f * (N + 2), 2 + f * (N + 2)]
    return out + [mdl.B - 1]


def weld_pairs(mdl: Model) -> list[tuple[int, int]]:
    """(body1, body2) of pris1-3, rev1-3, palm welds (xml_script.py:332-371)."""
    N = mdl.N
    pris = [(0, 1 + f * (N + 2)) for f in range(3)]
    rev = [(1 + f * (N + 2), 2 + f * (N + 2)) for f in range(3)]
    return pris + rev + [(0, mdl.B - 1)]


def _geom_attrs(gtype: int, size) -> dict:
    s = np.asarray(size, dtype=float)
    if gtype == GEOM_BOX:
        return {"type": "box", "size": _v(s)}
    if gtype == GEOM_CYLINDER:                        # (r, r, half length) here; MuJoCo: (r, half length), axis z
        return {"type": "cylinder", "size": _v([s[0], s[2]])}
    if gtype == GEOM_ELLIPSOID:
        return {"type": "ellipsoid", "size": _v(s)}
    raise ValueError(gtype)


def export_mjcf(mdl: Model, timestep: float = 2e-3, obj: ObjectSpec | None = None,
                obj_pose=(0.0, 0.0, 0.0)) -> str:
    """The model as an MJCF string. `obj`: one live object (only fillet-free primitives have an exact MuJoCo
    primitive; a rounded one raises), spawned at (x, y, z-rotation) like MjClass::spawn_object."""
    root = ET.Element("mujoco", model=f"gripper_b200_N{mdl.N}")
    ET.SubElement(root, "compiler", angle="radian", balanceinertia="true", inertiafromgeom="false")
    ET.SubElement(root, "option", timestep=f"{timestep:.10g}", gravity="0 0 -9.81")
    ET.SubElement(root, "size", njmax="3400", nconmax="800")
    dflt = ET.SubElement(root, "default")
    ET.SubElement(dflt, "geom", condim="3", solref="0.02 1", solimp="0.9 0.95 0.001 0.5 2", margin="0")
    world = ET.SubElement(root, "worldbody")
    ET.SubElement(world, "geom", name="ground_geom", type="plane", size="0 0 1", friction="1 0.005 0.005",
                  contype="0", conaffinity="3")
    elems: list[ET.Element] = []
    for i, b in enumerate(mdl.bodies):
        parent = world if b.parent < 0 else elems[b.parent]
        e = ET.SubElement(parent, "body", name=b.name, pos=_v(b.pos), quat=_v(mat_to_quat(b.R)))
        I = np.asarray(b.inertia)
        ET.SubElement(e, "inertial", pos=_v(b.ipos), mass=f"{b.mass:.10g}",
                      fullinertia=_v([I[0, 0], I[1, 1], I[2, 2], I[0, 1], I[0, 2], I[1, 2]]))
        j = {"name": f"{b.name}_joint", "type": "slide" if b.jtype == 0 else "hinge", "axis": _v(b.axis), "pos": "0 0 0",
             "stiffness": f"{b.stiffness:.10g}", "damping": f"{b.damping:.10g}", "springref": "0", "ref": "0"}
        if b.lo < b.hi:
            j.update(limited="true", range=_v([b.lo, b.hi]))
        ET.SubElement(e, "joint", **j)
        elems.append(e)
    for k, g in enumerate(mdl.geoms):
        a = _geom_attrs(g.gtype, g.size)
        a.update(name=f"{mdl.bodies[g.body].name}_geom{k}", pos=_v(g.pos), quat=_v(mat_to_quat(g.R)),
                 friction=_v([g.friction, 0.005 * g.friction, 1e-4 * g.friction]), contype="1", conaffinity="0")
        ET.SubElement(elems[g.body], "geom", **a)
    if obj is not None:
        o = obj.obj
        if o.round > 0:
            raise ValueError(f"{obj.name}: a filleted object has no exact MuJoCo primitive (the reference uses its mesh)")
        x, y, rot = obj_pose
        # the reference writes qpos[3..6] = (sin(-rot/2), 0, 0, cos(-rot/2)) (src/mjclass.cpp:2398-2409; MuJoCo reads
        # it as w, x, y, z): kept as is, here and in gmj_spawn / or_spawn
        e = ET.SubElement(world, "body", name="object", pos=_v([x, y, o.z_rest + 1e-6]),
                          quat=_v([np.sin(-rot / 2.0), 0, 0, np.cos(-rot / 2.0)]))
        ET.SubElement(e, "freejoint", name="object_joint")
        ET.SubElement(e, "inertial", pos="0 0 0", mass=f"{o.mass:.10g}", diaginertia=_v(list(o.inertia)))
        a = _geom_attrs(int(o.type), list(o.size))
        a.update(name="object_geom", quat=_v(list(o.quat)), friction=_v([o.friction, 0.005, 0.005]), contype="2", conaffinity="1")
        ET.SubElement(e, "geom", **a)
    eq = ET.SubElement(root, "equality")
    for name, (b1, b2) in zip(WELD_NAMES, weld_pairs(mdl)):
        ET.SubElement(eq, "weld", name=name, body1=mdl.bodies[b1].name, body2=mdl.bodies[b2].name, active="false",
                      solref="0.02 1", solimp="0.9 0.95 0.001 0.5 2")
    act = ET.SubElement(root, "actuator")
    for d in controlled_joints(mdl):
        ET.SubElement(act, "motor", name=f"{mdl.bodies[d].name}_motor", joint=f"{mdl.bodies[d].name}_joint", gear="1")
    ET.indent(root, space=" ")
    return ET.tostring(root, encoding="unicode")


def control_law(mdl: Model, qpos, qvel, target_x: float, target_th: float, target_z: float, base_z: float):
    """ctrl of the motors above, in `controlled_joints` order (luke::control; oracle/physics.c:269-278)."""
    e, N = mdl.extras, mdl.N
    out = [-((qpos[0] - base_z) * e["base_kp"] + qvel[0] * e["base_kd"])]
    for f in range(3):
        ip, ir = 1 + f * (N + 2), 2 + f * (N + 2)
        out.append(-((qpos[ip] - target_x) * e["kp"][0] + qvel[ip] * e["kd"][0]))
        out.append(-((qpos[ir] - target_th) * e["kp"][1] + qvel[ir] * e["kd"][1]))
    out.append(-((qpos[mdl.B - 1] - target_z) * e["kp"][2] + qvel[mdl.B - 1] * e["kd"][2]))
    return np.array(out)


__all__ = ["export_mjcf", "control_law", "controlled_joints", "weld_pairs", "WELD_NAMES", "quat_to_mat"]
