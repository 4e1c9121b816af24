"""MJCF export (gripper_mujoco_b200/mjcf.py): structure against the model, and -- when a maintainer has produced
tests/golden/mujoco_pin.npz with tools/mujoco_pin.py on a machine that has MuJoCo -- the oracle against real MuJoCo."""
import os
import xml.etree.ElementTree as ET

import numpy as np
import pytest

from gripper_mujoco_b200.mjcf import WELD_NAMES, control_law, controlled_joints, export_mjcf, weld_pairs
from gripper_mujoco_b200.model import GripperParams, build_model, forward_kinematics, quat_to_mat
from gripper_mujoco_b200.objects import build_object_set

PIN = os.path.join(os.path.dirname(__file__), "golden", "mujoco_pin.npz")


@pytest.mark.parametrize("N", [5, 8, 24])
def test_mjcf_structure_matches_the_model(N):
    mdl = build_model(GripperParams(num_segments=N))
    cube = next(o for o in build_object_set() if o.name.startswith("cube_fillet_0_7"))
    root = ET.fromstring(export_mjcf(mdl, 3e-3, obj=cube, obj_pose=(0.004, -0.002, 0.3)))
    bodies = {b.get("name"): b for b in root.iter("body")}
    assert len(bodies) == mdl.B + 1
    joints = list(root.iter("joint"))
    assert len(joints) == mdl.B and len(list(root.iter("freejoint"))) == 1
    assert sum(j.get("type") == "hinge" for j in joints) == 3 * (N + 1) and sum(j.get("type") == "slide" for j in joints) == 5
    masses = [float(i.get("mass")) for i in root.iter("inertial")]
    assert abs(sum(masses) - (sum(b.mass for b in mdl.bodies) + cube.obj.mass)) < 1e-9
    assert len(list(root.find("worldbody").iter("geom"))) == len(mdl.geoms) + 2      # + ground + object
    assert [w.get("name") for w in root.iter("weld")] == WELD_NAMES
    for w, (b1, b2) in zip(root.iter("weld"), weld_pairs(mdl)):
        assert (w.get("body1"), w.get("body2")) == (mdl.bodies[b1].name, mdl.bodies[b2].name) and w.get("active") == "false"
    assert len(list(root.iter("motor"))) == len(controlled_joints(mdl)) == 8
    stiff = [float(j.get("stiffness")) for j in joints]
    assert sum(s > 0 for s in stiff) == 3 * N                                # the finger segment hinges only
    # nesting reproduces the parent array; composing pos / quat down the tree gives the model's own FK at q = 0
    xpos, xmat = forward_kinematics(mdl, np.zeros(mdl.B))

    def walk(elem, P, R):
        for child in elem.findall("body"):
            if child.get("name") == "object":
                continue
            p = np.array(child.get("pos").split(), dtype=float); q = np.array(child.get("quat").split(), dtype=float)
            Pc, Rc = P + R @ p, R @ quat_to_mat(q)
            i = next(k for k, b in enumerate(mdl.bodies) if b.name == child.get("name"))
            assert np.abs(Pc - xpos[i]).max() < 1e-9 and np.abs(Rc - xmat[i]).max() < 1e-6
            walk(child, Pc, Rc)
    walk(root.find("worldbody"), np.zeros(3), np.eye(3))
    filleted = next(o for o in build_object_set() if o.obj.round > 0)
    with pytest.raises(ValueError, match="filleted"):
        export_mjcf(mdl, obj=filleted)


def test_control_law_is_the_oracles():
    """ctrl = -(kp (q - q*) + kd qdot) on the eight controlled dofs (oracle/physics.c:269-278)."""
    mdl = build_model(GripperParams(num_segments=5))
    rng = np.random.default_rng(0)
    q, v = rng.normal(size=mdl.B) * 0.01, rng.normal(size=mdl.B) * 0.1
    c = control_law(mdl, q, v, 0.12, 0.05, 0.01, -0.003)
    e = mdl.extras
    assert len(c) == 8
    assert np.isclose(c[0], -((q[0] + 0.003) * e["base_kp"] + v[0] * e["base_kd"]))
    assert np.isclose(c[1], -((q[1] - 0.12) * e["kp"][0] + v[1] * e["kd"][0]))
    assert np.isclose(c[2], -((q[2] - 0.05) * e["kp"][1] + v[2] * e["kd"][1]))
    assert np.isclose(c[7], -((q[mdl.B - 1] - 0.01) * e["kp"][2] + v[mdl.B - 1] * e["kd"][2]))


@pytest.mark.skipif(not os.path.exists(PIN), reason="tests/golden/mujoco_pin.npz not produced yet (tools/mujoco_pin.py needs MuJoCo)")
def test_oracle_against_real_mujoco():
    """Physics pin: the oracle's trajectories against MuJoCo's on the exported model."""
    from common import make_pair
    from oracle.binding import OracleEnv
    G = np.load(PIN)
    N = int(G["segments"])
    mdl, st = make_pair(N, float(G["dt"]), noise=False)
    B = mdl.B
    for name, steps, force in (("settle", 400, 0.0), ("tip_load", 1500, 0.25)):
        s = st.copy()
        s.curve_validation, s.tip_force_applied = force > 0, force
        for k in range(7):
            setattr(s, f"reset_qpos_g{k}", [mdl.params.xy_home, 0.0][k % 2] if k < 6 else mdl.params.z_home)
        s.reset_qpos_base = 0.0
        e = OracleEnv(mdl.to_struct(), s.pack()); e.reset(1, 0, 1)
        ref = G[name]
        for r in range(len(ref)):
            e.physics_step(10)
            assert np.abs(e.qpos()[:B] - ref[r][:B]).max() < 2e-4, (name, r)
