"""Workload used for the ncu captures under profiles/: configs[1] shape (4096 envs, N=5, calibrated
dt/S fixed to the values the auto-calibration finds so no calibration kernels pollute the capture),
objects spawned, random actions. Usage: python profiles/profile_step.py [n_envs] [N] [steps]"""
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
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 16
st = Settings()
st.debug = False
st.random_seed = 1234
st.set_use_noise(True)
st.mujoco_timestep, st.sim_steps_per_action = ({5: 7.02e-3, 8: 3.262e-3, 16: 1.039e-3, 24: 0.56e-3}[N], {5: 29, 8: 62, 16: 193, 24: 358}[N])
st.bending_gauge.normalise = {5: 2.41, 8: 2.54, 16: 2.63, 24: 2.66}[N]
st.wrist_sensor_Z.raw_value_offset = -23.3
mj = MjClass(n, GripperParams(num_segments=N), settings=st, calibrate=False)
mj.batch.spawn_with_retry(0.01, float(np.pi / 2))
g = torch.Generator(device="cpu"); g.manual_seed(0)
acts = torch.randint(0, mj.get_n_actions() - 1, (steps, n), generator=g, dtype=torch.int32).to(mj.device)
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for k in range(steps):
    if k == steps - 4:
        torch.cuda.synchronize(); ev0.record()
    mj.fused_step(acts[k])
ev1.record(); torch.cuda.synchronize()
print(f"{ev0.elapsed_time(ev1) / 4:.3f} ms/step  stats={mj.batch.stats()}")
