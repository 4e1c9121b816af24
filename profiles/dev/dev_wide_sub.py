"""Development aid: one wide-kernel substep against the oracle, per-dof differences."""
import sys
import numpy as np
import os, sys  # development script (moved out of tests/): make tests/common.py importable
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from common import OracleBatch, make_pair, objects
from gripper_mujoco_b200.batch import GripperBatch
np.set_printoptions(linewidth=220, precision=3, suppress=False)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 16
dt = {5: 2e-3, 8: 1e-3, 16: 8e-4, 24: 4e-4}[N]
for welds in (0, 0x7f):
    mdl, st = make_pair(N, dt, steps_per_action=40, noise=False)
    gb = GripperBatch(1, mdl, st, precision="fp64"); gb.set_objects(objects())
    ob = OracleBatch(1, mdl, st); gb.reset(); ob.reset()
    gb.set_welds(welds)
    for i in range(7):
        ob.envs[0].set_weld(i, (welds >> i) & 1)
    for k in range(3):
        gb.substeps(1, sensors=False); ob.envs[0].substep(1)
        qg, vg = gb.state()
        qo, vo = ob.qpos(), ob.qvel()
        p = gb.probe()[0]
        print(f"welds={welds:#x} substep {k}: max dq {np.abs(qg-qo).max():.3e} max dv {np.abs(vg-vo).max():.3e} nrow(dev)={p[11]} ncon(dev)={p[10]} peek={ob.envs[0].peek().nrow if hasattr(ob.envs[0].peek(),'nrow') else '?'}")
        if np.abs(vg - vo).max() > 1e-6:
            print(" dev qvel", vg[0]); print(" ora qvel", vo[0]); break
    gb.close()
