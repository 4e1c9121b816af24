"""Run by tests/test_gpu_self_collision.py in a child process with GMJ_SELF_COLLISION=1 (the opt-in library build with
gripper <-> gripper contacts): three scripted empty-gripper episodes on the GPU against the oracle with its
self_collision switch on. Prints one JSON line."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from common import OracleBatch, make_pair, objects  # noqa: E402
from gripper_mujoco_b200 import lib as gl  # noqa: E402
from gripper_mujoco_b200.batch import GripperBatch  # noqa: E402

N, prec = int(sys.argv[1]), sys.argv[2]
assert gl.SO_PATH == gl.SC_SO_PATH, gl.SO_PATH
dt, S = {5: (6.857e-3, 30), 8: (3.225e-3, 63)}[N]
mdl, st = make_pair(N, dt, steps_per_action=S)
# header-default action table: 0 close, 2 tilt in, 3 tilt out, 4 palm forward
scripts = {"tilt": [2] * 40, "palm": [4] * 30 + [2] * 45, "close_tilt": [0] * 25 + [2] * 40 + [3] * 10}
names = list(scripts)
T = max(len(v) for v in scripts.values())
acts = np.array([[scripts[k][t] if t < len(scripts[k]) else 2 for k in names] for t in range(T)], dtype=np.int32)
gb = GripperBatch(len(names), mdl, st, precision=prec); gb.set_objects(objects())
gb.reset()
ob = OracleBatch(len(names), mdl, st)
for e in ob.envs:
    e.set_self_collision(True)
free = OracleBatch(len(names), mdl, st)                 # the default: fingers pass through each other
worst = np.zeros(len(names)); lag = np.zeros(len(names)); bend = np.zeros(len(names))
for t in range(T):
    gb.step(torch.from_numpy(acts[t]).to(gb.device)); ob.step(acts[t]); free.step(acts[t])
    q, _ = gb.state()
    qo, qf = ob.qpos(), free.qpos()
    worst = np.maximum(worst, np.abs(q[:, :mdl.B] - qo[:, :mdl.B]).max(1))
    lag = np.maximum(lag, q[:, 2] - qf[:, 2])           # tilt of finger 1 held back by the contacts
    bend = np.maximum(bend, np.abs(q[:, 3:3 + N]).max(1))
s = gb.stats()
print(json.dumps({"worst": dict(zip(names, worst.tolist())), "lag": dict(zip(names, lag.tolist())),
                  "bend": dict(zip(names, bend.tolist())), "contacts": s["contacts"], "unstable": s["unstable_resets"],
                  "capped": s["capped_solves"]}))
