"""Development driver (not a pytest): run CUDA path vs oracle on a few scenarios and print the gaps.
Usage: python tests/dev_gpu_check.py [fp64|fp32] [N]"""
import sys
import time

import numpy as np
import torch

import os, sys  # development script (moved out of tests/): make tests/common.py importable
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from common import OracleBatch, find_object, make_pair, objects

from gripper_mujoco_b200.batch import GripperBatch


def run(precision, N, n_envs, n_steps, with_object, noise, scripted=False, seed=7):
    mdl, st = make_pair(N, 2e-3, noise=noise)
    gb = GripperBatch(n_envs, mdl, st, precision=precision)
    gb.set_objects(objects())
    ob = OracleBatch(n_envs, mdl, st)
    gb.reset(); ob.reset()          # episode 2 on both sides
    rng = np.random.default_rng(seed)
    if with_object:
        names = ["cube_fillet_0_3", "cylinder_12_upright_fillet_0_5", "sphere_", "cuboid_311_thinner_fillet_20_3", "ellipse_1p51_52"]
        idx = []
        for k in range(n_envs):
            try:
                idx.append(find_object(names[k % len(names)])[0])
            except KeyError:
                idx.append(find_object("cube")[0])
        idx = np.array(idx, dtype=np.int32)
        x = rng.uniform(-0.01, 0.01, n_envs).astype(np.float32); y = rng.uniform(-0.01, 0.01, n_envs).astype(np.float32)
        r = rng.uniform(-0.5, 0.5, n_envs).astype(np.float32)
        gb.spawn_object(idx, x, y, r); ob.spawn(idx, x, y, r)
    nact = gb.n_actions
    print(f"--- {precision} N={N} envs={n_envs} steps={n_steps} object={with_object} noise={noise} n_obs={gb.n_obs} n_act={nact}")
    worst = dict(q=0, v=0, obs=0, rew=0)
    for t in range(n_steps):
        if scripted:
            # close fingers (X-), then palm, then lift / terminate
            a = np.full(n_envs, 1 if t < 12 else (5 if t < 18 else nact - 1), dtype=np.int32)
        else:
            a = rng.integers(0, nact - 1, n_envs).astype(np.int32)
        o_g, r_g, d_g, e_g = gb.step(a)
        o_o, r_o, d_o, e_o = ob.step(a)
        qg, vg = gb.state()
        qo, vo = ob.qpos(), ob.qvel()
        o_g = o_g.cpu().numpy(); r_g = r_g.cpu().numpy(); d_g = d_g.cpu().numpy(); e_g = e_g.cpu().numpy()
        dq = np.abs(qg - qo).max(); dv = np.abs(vg - vo).max(); dob = np.abs(o_g - o_o).max(); dr = np.abs(r_g - r_o).max()
        ev_bad = int((e_g != e_o).sum()); dn_bad = int((d_g.astype(bool) != d_o).sum())
        p = gb.probe()
        print(f"t={t:3d} dq={dq:.3e} dv={dv:.3e} dobs={dob:.3e} drew={dr:.3e} ev_mismatch={ev_bad} done_mismatch={dn_bad} "
              f"ncon={p[:, 10].max():.0f} nrow={p[:, 11].max():.0f} niter={p[:, 12].max():.0f} unstable={p[:, 13].sum():.0f} "
              f"rew={r_o[:3]} done={d_o[:3].astype(int)}")
        if ev_bad and t < 3:
            bad = np.argwhere(e_g != e_o)[:6]
            for b in bad:
                print("   ev", b, e_g[tuple(b)], e_o[tuple(b)])
        if dob > 1e-3 and t < 3:
            bad = np.argwhere(np.abs(o_g - o_o) > 1e-3)[:8]
            for b in bad:
                print("   obs", b, o_g[tuple(b)], o_o[tuple(b)])
        if not np.isfinite(dq):
            print("non-finite state; qg row0:", qg[0]); break
    print("stats", gb.stats())
    gb.close()


def timing(precision, N, n_envs, n_steps, with_object):
    mdl, st = make_pair(N, 2e-3, noise=True)
    gb = GripperBatch(n_envs, mdl, st, precision=precision)
    gb.set_objects(objects())
    gb.reset()
    if with_object:
        gb.spawn_into_scene(0.01, 0.5)
    a = torch.randint(0, gb.n_actions - 1, (n_steps + 3, n_envs), device=gb.device, dtype=torch.int32)
    for k in range(3):
        gb.step(a[k])
    torch.cuda.synchronize()
    t0 = time.time()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for k in range(n_steps):
        gb.step(a[3 + k])
    ev1.record(); torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / n_steps
    print(f"timing {precision} N={N} envs={n_envs} object={with_object}: {ms:.2f} ms/step -> {n_envs / ms * 1e3:.0f} action-steps/s "
          f"(wall {time.time() - t0:.2f}s) stats={gb.stats()}")
    gb.close()


if __name__ == "__main__":
    prec = sys.argv[1] if len(sys.argv) > 1 else "fp64"
    N = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    run(prec, N, 4, 6, False, False)
    run(prec, N, 5, 25, True, False, scripted=True)
    run(prec, N, 5, 10, True, True)
    timing(prec, N, 4096, 5, False)
    timing(prec, N, 4096, 5, True)
