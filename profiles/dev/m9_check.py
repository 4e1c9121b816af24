"""GPU (FP64, -DGMJ_SELF_COLLISION build via GMJ_LIB) against the oracle with or_set_self_collision: empty gripper tilted in
until the hooks meet, and tilted onto the extended palm. Prints the per-step qpos error and the contact counts."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from common import OracleBatch, make_pair, objects  # noqa: E402
from gripper_mujoco_b200.batch import GripperBatch  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 5
prec = sys.argv[2] if len(sys.argv) > 2 else "fp64"
dt, S = {5: (6.857e-3, 30), 8: (3.225e-3, 63)}[N]
mdl, st = make_pair(N, dt, steps_per_action=S)
scripts = {"tilt": [2] * 40, "palm": [4] * 30 + [2] * 45, "close_tilt": [0] * 25 + [2] * 40 + [3] * 10}
names = list(scripts)
T = max(len(v) for v in scripts.values())
n = len(names)
acts = np.array([[scripts[k][t] if t < len(scripts[k]) else 2 for k in names] for t in range(T)], dtype=np.int32)
gb = GripperBatch(n, mdl, st, precision=prec); gb.set_objects(objects())
gb.reset()
ob = OracleBatch(n, mdl, st)
for e in ob.envs:
    e.set_self_collision(True)
worst = np.zeros(n)
import torch
for t in range(T):
    gb.step(torch.from_numpy(acts[t]).to(gb.device))
    ob.step(acts[t])
    q, _ = gb.state()
    qo = ob.qpos()
    err = np.abs(q[:, :mdl.B] - qo[:, :mdl.B]).max(1)
    worst = np.maximum(worst, err)
    if t % 5 == 4 or t > 22:
        print(t + 1, " ".join(f"{names[k]}: err {err[k]:.2e} th {q[k, 2]:+.4f}/{qo[k, 2]:+.4f} bend {np.abs(q[k, 3:3 + N]).max():.4f}/{np.abs(qo[k, 3:3 + N]).max():.4f}" for k in range(n)))
print("worst", dict(zip(names, worst)), gb.stats())
