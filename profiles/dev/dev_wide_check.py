"""Development aid (not collected by pytest): wide-kernel (N = 16 / 24) parity against the oracle, printing
the per-step differences instead of asserting. Usage: python profiles/dev/dev_wide_check.py [fp64|fp32] [N ...]"""
import sys
import time

import numpy as np

import os, sys  # development script (moved out of tests/): make tests/common.py importable
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from common import OracleBatch, find_object, make_pair, objects

DISCRETE = list(range(15)) + [15 + 2 * k for k in range(21)]


def run(N, dt, precision, with_obj, steps=6, S=40, n=6):
    from gripper_mujoco_b200.batch import GripperBatch
    mdl, st = make_pair(N, dt, steps_per_action=S, noise=False)
    gb = GripperBatch(n, mdl, st, precision=precision)
    gb.set_objects(objects())
    ob = OracleBatch(n, mdl, st)
    gb.reset(); ob.reset()
    if with_obj:
        names = ["cube_fillet_0_3", "cylinder_12_upright_fillet_0_5", "sphere_", "cuboid_221_fillet_10_0", "ellipse_1p51_52", "cube_fillet_0_3"]
        idx = np.array([find_object(nm)[0] for nm in names[:n]], dtype=np.int32)
        x = np.linspace(-0.01, 0.01, n).astype(np.float32)
        gb.spawn_object(idx, x, 0.0, 0.1); ob.spawn(idx, x, 0.0, 0.1)
    rng = np.random.default_rng(N)
    for t in range(steps):
        a = rng.integers(0, gb.n_actions - 1, n).astype(np.int32)
        if with_obj and t < 4:
            a[:] = 0 if t % 2 == 0 else 2          # close the fingers on the object
        t0 = time.time()
        o_g, r_g, d_g, e_g = (v.cpu().numpy() for v in gb.step(a))
        tg = time.time() - t0
        o_o, r_o, d_o, e_o = ob.step(a)
        qg, vg = gb.state()
        print(f"N={N} {precision} obj={with_obj} t={t}: dq={np.abs(qg - ob.qpos()).max():.3e} dv={np.abs(vg - ob.qvel()).max():.3e} "
              f"dobs={np.abs(o_g - o_o).max():.3e} dr={np.abs(r_g - r_o).max():.3e} ev_equal={np.array_equal(e_g[:, DISCRETE], e_o[:, DISCRETE])} "
              f"done_equal={np.array_equal(d_g.astype(bool), d_o)} finite={np.isfinite(qg).all()} gpu_s={tg:.3f}", flush=True)
    print("stats", gb.stats(), flush=True)
    gb.close()


if __name__ == "__main__":
    precs = [a for a in sys.argv[1:] if a.startswith("fp")] or ["fp64", "fp32"]
    Ns = [int(a) for a in sys.argv[1:] if a.isdigit()] or [16, 24]
    for N in Ns:
        dt = {5: 2e-3, 8: 1e-3, 16: 8e-4, 24: 4e-4}[N]
        for p in precs:
            for with_obj in (False, True):
                try:
                    run(N, dt, p, with_obj)
                except Exception as ex:   # keep going: one GPU call should report everything
                    print(f"N={N} {p} obj={with_obj} FAILED: {type(ex).__name__}: {ex}", flush=True)
