"""Distribution of LCP pivoting rounds (last substep of each env) -- which envs need more than one round?
Usage: python profiles/niter_hist.py [n_envs] [N] [steps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjclass import MjClass  # noqa: E402
from gripper_mujoco_b200.model import GripperParams  # noqa: E402
from gripper_mujoco_b200.settings import Settings  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 592
N = int(sys.argv[2]) if len(sys.argv) > 2 else 24
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 6
st = Settings(); st.debug = False; st.random_seed = 1234; st.set_use_noise(True)
st.mujoco_timestep, st.sim_steps_per_action = ({5: 7.02e-3, 8: 3.262e-3, 16: 1.039e-3, 24: 0.56e-3}[N], {5: 29, 8: 62, 16: 193, 24: 358}[N])
st.bending_gauge.normalise = {5: 2.41, 8: 2.54, 16: 2.63, 24: 2.66}[N]
st.wrist_sensor_Z.raw_value_offset = -23.3
mj = MjClass(n, GripperParams(num_segments=N), settings=st, calibrate=False)
mj.batch.spawn_with_retry(0.01, float(np.pi / 2))
g = torch.Generator(device="cpu"); g.manual_seed(0)
acts = torch.randint(0, mj.get_n_actions() - 1, (steps, n), generator=g, dtype=torch.int32).to(mj.device)
for k in range(steps):
    mj.fused_step(acts[k])
it = np.asarray(mj.batch.probe_field("niter")); nr = np.asarray(mj.batch.probe_field("nrow")); nc = np.asarray(mj.batch.probe_field("ncon"))
print("mean rounds/substep per env: quantiles", np.quantile(it, [0, 0.5, 0.9, 0.99, 1.0]).round(2))
order = np.argsort(-it)[:12]
names = mj.get_current_object_name()
for e in order:
    print(f"  env {e:4d} rounds {it[e]:6.2f}  nrow(last) {nr[e]:3.0f}  mean contacts {nc[e]:5.2f}  {names[e]}")
print("share of all rounds spent by the top 5 % envs:", np.sort(it)[-max(1, n // 20):].sum() / it.sum())
print(mj.batch.stats())
