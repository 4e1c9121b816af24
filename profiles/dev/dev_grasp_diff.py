import numpy as np, sys
import os, sys  # development script (moved out of tests/): make tests/common.py importable
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from common import OracleBatch, find_object, make_pair, objects
from gripper_mujoco_b200.batch import GripperBatch
n = 1
mdl, st = make_pair(5, 7.02e-3, noise=False)
gb = GripperBatch(n, mdl, st, precision="fp64"); gb.set_objects(objects())
ob = OracleBatch(n, mdl, st)
gb.reset(); ob.reset()
idx = np.full(n, find_object("cube_fillet_0_3")[0], dtype=np.int32)
gb.spawn_object(idx, 0.0, 0.0, 0.0); ob.spawn(idx, 0.0, 0.0, 0.0)
S = int(st.sim_steps_per_action)
seq = [0] * 60 + [2] * 10
for t, a in enumerate(seq):
    act = np.full(n, a, dtype=np.int32)
    if t < 56:
        gb.step(act); ob.step(act); continue
    # substep-level comparison
    gb.set_action(act); [e.set_action(int(a)) for e in ob.envs]
    for k in range(S):
        it0 = gb.stats()["solver_iters"]
        gb.substeps(1, True); ob.envs[0].substep(1)
        its = gb.stats()["solver_iters"] - it0
        qg, vg = gb.state(); qo = ob.qpos(); p = ob.envs[0].peek()
        d = np.abs(qg - qo)[0]
        if d.max() > 1e-9 or k == S - 1 or its > 1:
            st_ = gb.stats()
            print(f"t={t} k={k} dq={d.max():.2e} at {d.argmax()} ncon_o={p.ncon} nrow_o={p.nrow} niter_o={p.niter} nrow_g={gb.probe()[0, 11]:.0f} pivots_g={its}")
            if d.max() > 1e-6: sys.exit()
    gb.substeps(0, True)
    gb.observe(); ob.envs[0].observation(); ob.envs[0].reward()
