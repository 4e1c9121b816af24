"""What is the M9 gap (no gripper <-> gripper contacts in the CUDA path and in the oracle's default mode) worth in the
quantities the statistical protocol compares? The oracle runs the protocol's episodes (tests/test_gpu_contract.py:
actions biased towards closing, every 4th env a scripted grasp-and-lift) twice on the same draws, without and with
its self_collision switch (oracle.h), and the paired differences are printed. CPU only.
Usage: python profiles/r2_m9_effect.py [episodes]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    from common import make_pair, objects, oracle_rollout
    from test_gpu_contract import SEARCH, STAT, TIMING, episode_stats, grasp_script
    from gripper_mujoco_b200 import selfcheck
    from gripper_mujoco_b200.settings import BINARY_NAMES
    N = 5
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    _, T, every = STAT[N]
    dt, S = TIMING[N]
    mdl, st = make_pair(N, dt, steps_per_action=S, noise=True)
    n_act = 15
    rng = np.random.default_rng(100 + N)
    idx = rng.integers(0, len(objects()), n).astype(np.int32)
    acts = rng.integers(0, n_act, (T, n)).astype(np.int32)
    b = (rng.random((T, n)) < 0.4) & (acts != n_act - 1)
    acts[b] = rng.choice([1, 3, 5, 13], size=int(b.sum()))
    for k in range(0, n, every):
        acts[:, k] = grasp_script(rng, T, n_act)
    res = {}
    for on in ("0", "1"):
        os.environ["ORACLE_SELF_COLLISION"] = on
        res[on] = oracle_rollout(N, st, np.arange(n), 1, idx, np.zeros((n, 3), np.float32), acts, search=SEARCH, stop_at_done=True)
    r0, l0, _, f0, fin0 = episode_stats(res["0"], None, T)
    r1, l1, _, f1, fin1 = episode_stats(res["1"], None, T)
    d = r1 - r0
    ts = list(range(3, T, 8))
    over = sum(int(((selfcheck.self_overlap(mdl, res["0"]["qpos"][t], depth=1e-4)[0] > 0) & (t < l0)).sum()) for t in ts)
    live = sum(int((t < l0).sum()) for t in ts)
    print(f"N={N} episodes={n} T={T}: M9-state env-steps without the contacts {over} of {live} sampled")
    print(f"return without {r0.mean():.4f} with {r1.mean():.4f}  paired diff {d.mean():+.2e} +- {d.std(ddof=1) / np.sqrt(n):.2e}; "
          f"episodes with a different return (1e-3) {np.mean(np.abs(d) > 1e-3):.4f}, different length {np.mean(l0 != l1):.4f}")
    for nm in ("lifted", "object_contact", "successful_grasp", "object_stable", "target_height"):
        if nm in BINARY_NAMES:
            k = BINARY_NAMES.index(nm)
            print(f"   {nm:20s} without {f0[:, k].mean():.4f} with {f1[:, k].mean():.4f} disagreeing episodes {(f0[:, k] != f1[:, k]).mean():.4f}")


if __name__ == "__main__":
    main()
