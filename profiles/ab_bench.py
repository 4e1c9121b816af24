"""A/B timing of differently compiled builds on the bench.py workload (random actions incl. the termination action,
masked auto-reset + respawn, steady state after a long warm-up) with the calibration fixed, so that only the
step kernel differs. Prints the mean gmj_step time. GMJ_LIB selects the build (gripper_mujoco_b200/lib.py).
Usage: python profiles/ab_bench.py [n_envs] [N] [warmup] [steps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjclass import MjClass  # noqa: E402
from gripper_mujoco_b200.model import GripperParams  # noqa: E402
from gripper_mujoco_b200.settings import Settings  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 5
W = int(sys.argv[3]) if len(sys.argv) > 3 else 40
K = int(sys.argv[4]) if len(sys.argv) > 4 else 20
st = Settings()
st.debug = False
st.random_seed = 1234
st.set_use_noise(True)
st.mujoco_timestep, st.sim_steps_per_action = ({5: 6.857e-3, 8: 3.225e-3, 16: 1.039e-3, 24: 0.56e-3}[N], {5: 30, 8: 63, 16: 193, 24: 358}[N])
st.bending_gauge.normalise = {5: 2.41, 8: 2.54, 16: 2.63, 24: 2.66}[N]
st.wrist_sensor_Z.raw_value_offset = -23.3
mj = MjClass(n, GripperParams(num_segments=N), settings=st, calibrate=False)
gb = mj.batch
mj.batch.spawn_with_retry(0.01, float(np.pi / 2))
g = torch.Generator(device="cpu"); g.manual_seed(0)
acts = torch.randint(0, mj.get_n_actions(), (W + K, n), generator=g, dtype=torch.int32).to(mj.device)
ep = torch.zeros(n, dtype=torch.int32, device=mj.device)
evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
for k in range(W + K):
    if k >= W:
        evs[k - W][0].record()
    gb.step(acts[k])
    if k >= W:
        evs[k - W][1].record()
    ep.add_(1)
    mask = (gb.done != 0) | (ep >= 250)
    m8 = mask.to(torch.uint8)
    gb.reset(m8); gb.spawn_with_retry(0.01, float(np.pi / 2), m8)
    ep.mul_((~mask).to(torch.int32))
torch.cuda.synchronize()
ms = [a.elapsed_time(b) for a, b in evs]
s = gb.stats()
print(f"{os.path.basename(os.environ.get('GMJ_LIB', 'default')):28s} n={n} N={N}  gmj_step mean {np.mean(ms):.3f} ms  min {np.min(ms):.3f}  "
      f"({n / np.mean(ms) * 1e3:.0f} action-steps/s)  substeps={s['substeps']} contacts={s['contacts']} unstable={s['unstable_resets']}")
