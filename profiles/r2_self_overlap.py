"""How often do the bench workloads enter the state the kernels do not simulate (gripper <-> gripper contacts, M9)?
Runs the bench.py episode loop (random actions / the contact-heavy scripted prefix, 250-step cap, reset + respawn) and
saves the gripper joint positions of every env every SAMPLE steps to gpurun_out/r2_self_overlap_<workload>.npz; the
overlap test itself (gripper_mujoco_b200/selfcheck.py, host numpy) is run afterwards with --analyse, no GPU needed.
Usage (GPU box): python profiles/r2_self_overlap.py [n_envs] [steps]      then here: python profiles/r2_self_overlap.py --analyse"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gripper_mujoco_b200.model import GripperParams, build_model  # noqa: E402

N, SAMPLE = 5, 10
OUT = os.path.join(ROOT, "gpurun_out")
# random_noterm: random motor actions only (no termination action): 250-step random walks, the most the motors wander
WORKLOADS = [w for w in ("random", "contact", "random_noterm") if os.environ.get("ONLY", w) == w]
CONTACT_PREFIX = [0] * 30 + [2] * 10 + [4] * 20 + [11] * 15          # bench.py

if "--analyse" in sys.argv:
    from gripper_mujoco_b200 import selfcheck
    mdl = build_model(GripperParams(num_segments=N))
    hooks = {N, (N + 1) + N, 2 * (N + 1) + N}
    palm = len(mdl.geoms) - 1
    for wl in WORKLOADS:
        d = np.load(os.path.join(OUT, f"r2_self_overlap_{wl}.npz"))
        q, ep, live = d["qpos"].astype(np.float64), d["ep_steps"], d["episode"]
        S, n = q.shape[0], q.shape[1]
        tot = any_ = hook_only = with_palm = 0
        ever = set()
        first = []
        for k in range(S):
            cnt, hits = selfcheck.self_overlap(mdl, q[k], depth=1e-4)
            m = cnt > 0
            tot += n; any_ += int(m.sum())
            per = {}
            for e, i, j in hits:
                per.setdefault(e, []).append((i, j))
            for e, pr in per.items():
                if all({i, j} <= hooks for i, j in pr): hook_only += 1
                if any(palm in (i, j) for i, j in pr): with_palm += 1
                key = (e, int(live[k, e]))
                if key not in ever:
                    ever.add(key); first.append(int(ep[k, e]))
        episodes = len({(e, int(live[k, e])) for k in range(S) for e in range(n)})
        print(f"{wl}: {S} samples x {n} envs: {any_} of {tot} sampled env-steps in overlap ({100 * any_ / tot:.2f} %), "
              f"hooks only {hook_only}, palm involved {with_palm}; {len(ever)} of {episodes} episodes seen enter it"
              + (f" (first at episode step: median {int(np.median(first))}, min {min(first)})" if first else ""))
    sys.exit(0)

import torch  # noqa: E402
from gripper_mujoco_b200.mjclass import MjClass  # noqa: E402
from gripper_mujoco_b200.settings import Settings  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 300
os.makedirs(OUT, exist_ok=True)
for wl in WORKLOADS:
    st = Settings()
    st.debug = False
    st.random_seed = 1234
    st.set_use_noise(True)
    st.mujoco_timestep, st.sim_steps_per_action = 6.857e-3, 30
    st.bending_gauge.normalise = 2.41
    st.wrist_sensor_Z.raw_value_offset = -23.3
    mj = MjClass(n, GripperParams(num_segments=N), settings=st, calibrate=False)
    gb = mj.batch
    B = mj.model.B
    gb.spawn_with_retry(0.01, float(np.pi / 2))
    g = torch.Generator(device="cpu"); g.manual_seed(0)
    acts = torch.randint(0, mj.get_n_actions(), (steps, n), generator=g, dtype=torch.int32).to(mj.device)
    prefix = torch.tensor(CONTACT_PREFIX, dtype=torch.int32, device=mj.device)
    ep = torch.zeros(n, dtype=torch.int32, device=mj.device)
    epi = torch.zeros(n, dtype=torch.int32, device=mj.device)
    Q, E, L = [], [], []
    for k in range(steps):
        a = acts[k]
        if wl == "random_noterm":
            a = a % (mj.get_n_actions() - 1)
        if wl == "contact":
            kk = ep.clamp(max=len(CONTACT_PREFIX) - 1).to(torch.int64)
            a = torch.where(ep < len(CONTACT_PREFIX), prefix[kk], a)
        gb.step(a)
        ep.add_(1)
        if (k + 1) % SAMPLE == 0:
            q, _ = gb.state()
            Q.append(q[:, :B].astype(np.float32)); E.append(ep.cpu().numpy().copy()); L.append(epi.cpu().numpy().copy())
        m8 = gb.reset_done(gb.done, 250)
        gb.spawn_with_retry(0.01, float(np.pi / 2), m8)
        ep.mul_(1 - m8.to(torch.int32)); epi.add_(m8.to(torch.int32))
    np.savez_compressed(os.path.join(OUT, f"r2_self_overlap_{wl}.npz"), qpos=np.stack(Q), ep_steps=np.stack(E), episode=np.stack(L))
    print(wl, "saved", len(Q), "samples", gb.stats())
    del mj, gb
