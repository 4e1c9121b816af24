#!/usr/bin/env python
"""Pin the restated physics against real MuJoCo -- run on a machine where `import mujoco` works (2.1.5 is the
reference's version; 2.2+ loads the same file). Not runnable in the build container or on the GPU box.

    python tools/mujoco_pin.py [--segments 8] [--out tests/golden/mujoco_pin.npz]

For each scenario it loads the MJCF written by gripper_mujoco_b200.mjcf.export_mjcf, steps mj_step under the
reference's PD law (luke::control) and stores the qpos trajectory; tests/test_mjcf.py replays the same scenarios on
the CPU oracle and compares (the CUDA path is tied to the oracle by tests/test_gpu_*.py):
  settle       welds on, 400 steps of 2 ms from the keyframe pose (luke::calibrate_reset)
  tip_load     welds on, 0.25 N calibration load on the last segment of every finger, 1500 steps of 2 ms
  object_rest  a fillet-free cube on the ground under the open gripper, 300 steps
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjcf import WELD_NAMES, control_law, controlled_joints, export_mjcf  # noqa: E402
from gripper_mujoco_b200.model import GripperParams, build_model  # noqa: E402
from gripper_mujoco_b200.objects import build_object_set  # noqa: E402


def run(mujoco, mdl, xml, steps, welds_on, tip_force=0.0, record_every=10):
    m = mujoco.MjModel.from_xml_string(xml)
    d = mujoco.MjData(m)
    B = mdl.B
    q0 = np.array([b.qpos0 for b in mdl.bodies])
    d.qpos[:B] = q0
    mujoco.mj_forward(m, d)
    if welds_on:
        for name in WELD_NAMES:                      # activate at the current pose, as luke::set_constraint does
            i = mujoco.mj_name2id(m, mujoco.mjtObj.mjOBJ_EQUALITY, name)
            if hasattr(d, "eq_active"):
                d.eq_active[i] = 1
            else:
                m.eq_active[i] = 1
    p = mdl.params
    tx, tth, tz, bz = p.xy_home, 0.0, p.z_home, 0.0
    out = []
    for k in range(steps):
        mujoco.mj_step1(m, d)
        d.ctrl[:] = control_law(mdl, d.qpos, d.qvel, tx, tth, tz, bz)
        d.xfrc_applied[:] = 0
        if tip_force:
            for f, tb in enumerate(mdl.extras["tip_body"]):
                bid = mujoco.mj_name2id(m, mujoco.mjtObj.mjOBJ_BODY, mdl.bodies[tb].name)
                d.xfrc_applied[bid, :3] = mdl.extras["tip_mat"][f] @ np.array([0, 0, -tip_force])
        mujoco.mj_step2(m, d)
        if (k + 1) % record_every == 0:
            out.append(d.qpos.copy())
    return np.array(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--segments", type=int, default=8)
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "mujoco_pin.npz"))
    a = ap.parse_args()
    import mujoco
    mdl = build_model(GripperParams(num_segments=a.segments))
    cube = next(o for o in build_object_set() if o.name.startswith("cube_fillet_0_7"))
    res = {"segments": a.segments, "mujoco_version": mujoco.__version__, "dt": 2e-3,
           "settle": run(mujoco, mdl, export_mjcf(mdl, 2e-3), 400, True),
           "tip_load": run(mujoco, mdl, export_mjcf(mdl, 2e-3), 1500, True, tip_force=0.25),
           "object_rest": run(mujoco, mdl, export_mjcf(mdl, 2e-3, obj=cube), 300, True),
           "controlled": np.array(controlled_joints(mdl))}
    np.savez_compressed(a.out, **res)
    print("wrote", a.out)


if __name__ == "__main__":
    main()
