"""Development driver: per-step largest gap of the continuous event values (fp32 CUDA vs oracle)."""
import numpy as np
import os, sys  # development script (moved out of tests/): make tests/common.py importable
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from common import OracleBatch, find_object, make_pair, objects
from gripper_mujoco_b200.batch import GripperBatch
from gripper_mujoco_b200.settings import LINEAR_NAMES
import sys
prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
OBJECTS = ["cube_fillet_0_3", "cylinder_12_upright_fillet_0_5", "sphere_", "cuboid_221_fillet_10_0", "ellipse_1p51_52"]
n = 5
mdl, st = make_pair(5, 2e-3, noise=False)
gb = GripperBatch(n, mdl, st, precision=prec); gb.set_objects(objects())
ob = OracleBatch(n, mdl, st)
gb.reset(); ob.reset()
idx = np.array([find_object(nm)[0] for nm in OBJECTS], dtype=np.int32)
x = np.array([0.0, 0.004, -0.004, 0.0, 0.002], dtype=np.float32)
r = np.array([0.0, 0.3, 0.0, 0.5, -0.2], dtype=np.float32)
gb.spawn_object(idx, x, -x, r); ob.spawn(idx, x, -x, r)
for t in range(16):
    a = np.full(n, 1 if t < 10 else (5 if t < 14 else gb.n_actions - 1), dtype=np.int32)
    o_g, r_g, d_g, e_g = (v.cpu().numpy() for v in gb.step(a))
    o_o, r_o, d_o, e_o = ob.step(a)
    C = [16 + 2 * k for k in range(21)]
    d = np.abs(e_g[:, C] - e_o[:, C])
    env, k = np.unravel_index(d.argmax(), d.shape)
    qg, vg = gb.state()
    p = gb.probe()
    print(f"t={t} max ev diff {d.max():.4f} env {env} {LINEAR_NAMES[k]} gpu {e_g[env, C[k]]:.4f} oracle {e_o[env, C[k]]:.4f} "
          f"dq {np.abs(qg - ob.qpos()).max():.2e} dv {np.abs(vg - ob.qvel()).max():.2e} "
          f"si_g {p[env, 5:10].round(3)} si_o {np.array(ob.envs[env].peek().si[:5]).round(3)} niter {[e.peek().niter for e in ob.envs]}")
