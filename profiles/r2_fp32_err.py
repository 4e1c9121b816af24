"""Measured FP32-vs-oracle errors at contract size (sets the bars of tests/test_gpu_contract.py).
usage: python profiles/r2_fp32_err.py [n] [N] [steps]"""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from common import make_pair, objects, oracle_rollout
from gripper_mujoco_b200.batch import GripperBatch
def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    N = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    T = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    sample = int(sys.argv[4]) if len(sys.argv) > 4 else n
    dt, S = {5: (6.857e-3, 30), 8: (3.225e-3, 63)}[N]
    mdl, st = make_pair(N, dt, steps_per_action=S, noise=True)
    gb = GripperBatch(n, mdl, st, precision="fp32"); gb.set_objects(objects())
    rng = np.random.default_rng(7)
    idx = rng.integers(0, len(objects()), n).astype(np.int32)
    xyr = np.stack([rng.integers(-5, 6, n) * 2e-3, rng.integers(-5, 6, n) * 2e-3, rng.integers(-15, 16, n) * (np.pi / 30)], 1).astype(np.float32)
    acts = rng.integers(0, gb.n_actions, (T, n)).astype(np.int32)
    gb.reset(); gb.spawn_with_retry(0.01, float(np.pi / 2), None, idx, tries=3, fallback_xy_mm=10, fallback_rot_deg=5)
    spawned = gb.spawned.cpu().numpy()
    print("spawn codes (1 = clear pose found, 2 = fallback pose without collision test):", np.bincount(spawned, minlength=3))
    sel = np.sort(rng.choice(n, sample, replace=False))
    t0 = time.time()
    ep = int(gb.probe_field('episode')[0])
    ora = oracle_rollout(N, st, sel, ep, idx[sel], xyr[sel], acts[:, sel], search=(0.01, float(np.pi / 2), 3, 10, 5))
    print(f"oracle: {sample * T} action-steps in {time.time() - t0:.1f} s on {len(os.sched_getaffinity(0))} cores")
    DISC = list(range(15)) + [15 + 2 * k for k in range(21)]; CONT = [16 + 2 * k for k in range(21)]
    B = mdl.B
    for t in range(T):
        o, r, d, e = (v.cpu().numpy() for v in gb.step(acts[t]))
        q, v = gb.state()
        o, r, d, e, q, v = o[sel], r[sel], d[sel], e[sel], q[sel], v[sel]
        oq = ora["qpos"][t]
        edge = np.abs(np.hypot(oq[:, B], oq[:, B + 1]) - 10e-3) < 1e-6        # object resting on the within_XY_distance threshold
        neq = e[:, DISC] != ora["events"][t][:, DISC]
        neq[:, 14] &= ~edge
        same = ~neq.any(1)
        dq = np.abs(q - ora["qpos"][t]); dv = np.abs(v - ora["qvel"][t])
        rel = dq[:, :B].max(1) / np.maximum(np.abs(ora["qpos"][t][:, :B]).max(1), 1e-3)
        if t in (0, T - 1):
            mism = neq.sum(0)
            print("   mismatching discrete columns (index in the 57-vector: count):", {DISC[k]: int(c) for k, c in enumerate(mism) if c})
            w = np.where(neq[:, 14])[0][:6]
            for k in w:
                print("   col14 env", sel[k], "code", spawned[sel[k]], "oracle xy", oq[k, B:B + 2], "device xy", q[k, B:B + 2], "hyp", np.hypot(oq[k, B], oq[k, B + 1]), np.hypot(q[k, B], q[k, B + 1]), "ev", e[k, 14], ora["events"][t][k, 14])
            w = np.where(same & (np.abs(r - ora["reward"][t]) > 0.05))[0][:4]
            for k in w:
                print("   reward env", sel[k], "code", spawned[sel[k]], r[k], ora["reward"][t][k], "cont diff cols", np.where(np.abs(e[k] - ora["events"][t][k]) > 1e-3)[0], e[k][np.abs(e[k] - ora["events"][t][k]) > 1e-3], ora["events"][t][k][np.abs(e[k] - ora["events"][t][k]) > 1e-3])
            bad = np.argsort(-rel)[:5]
            print("   worst envs", sel[bad], "rel", rel[bad], "obj", idx[sel[bad]], "nrow", gb.probe_field("nrow")[sel[bad]], "same", same[bad])
        calm = same & (spawned[sel] == 1)
        print(f"t={t} calm {calm.mean():.4f} | calm envs: qpos rel max {rel[calm].max():.2e} obj {dq[calm][:, B:B+3].max():.2e} qvel {dv[calm].max():.2e} obs {np.abs(o - ora['obs'][t])[calm].max():.2e} rew {np.abs(r - ora['reward'][t])[calm].max():.2e} evval {np.abs(e[:, CONT] - ora['events'][t][:, CONT])[calm].max():.2e} | median rel {np.median(rel):.2e} p90 {np.quantile(rel, .9):.2e}")
        print(f"t={t} events identical {same.mean():.4f} done identical {(d.astype(bool) == ora['done'][t]).mean():.4f} | "
              f"qpos gripper rel max {rel.max():.2e} p99 {np.quantile(rel, .99):.2e} | obj pos max {dq[:, B:B+3].max():.2e} p99 {np.quantile(dq[:, B:B+3].max(1), .99):.2e} | "
              f"qvel max {dv.max():.2e} p99 {np.quantile(dv.max(1), .99):.2e} | obs max {np.abs(o - ora['obs'][t]).max():.2e} p99 {np.quantile(np.abs(o - ora['obs'][t]).max(1), .99):.2e} | "
              f"reward max {np.abs(r - ora['reward'][t]).max():.2e} | event values max {np.abs(e[:, CONT] - ora['events'][t][:, CONT]).max():.2e} p99 {np.quantile(np.abs(e[:, CONT] - ora['events'][t][:, CONT]).max(1), .99):.2e}")
    print("stats", gb.stats())


if __name__ == "__main__":
    main()
