import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjclass import MjClass
from gripper_mujoco_b200.model import GripperParams
from gripper_mujoco_b200.settings import Settings
from gripper_mujoco_b200.lib import GP
n = 4096
st = Settings(); st.debug = False; st.random_seed = 1234; st.set_use_noise(True)
st.mujoco_timestep, st.sim_steps_per_action = 6.825e-3, 30
st.bending_gauge.normalise = 2.41; st.wrist_sensor_Z.raw_value_offset = -23.3
mj = MjClass(n, GripperParams(num_segments=5), settings=st, calibrate=False)
mj.batch.spawn_with_retry(0.01, float(np.pi / 2))
g = torch.Generator(device="cpu"); g.manual_seed(0)
gb = mj.batch
H = np.zeros(50, dtype=np.int64)
for k in range(60):
    a = torch.randint(0, 15, (n,), generator=g, dtype=torch.int32)
    gb.step(a)
    m8 = (gb.done != 0).to(torch.uint8)
    gb.reset(m8); gb.spawn_with_retry(0.01, float(np.pi / 2), m8)
    if k >= 20:
        nr = gb.probe()[:, GP["nrow"]].astype(int)
        H += np.bincount(nr, minlength=50)[:50]
tot = H.sum()
print("nrow histogram (percent):")
for i in range(0, 50):
    if H[i]: print(i, f"{100 * H[i] / tot:.2f}")
print("P(nr>32) =", H[33:].sum() / tot, " mean", (H * np.arange(50)).sum() / tot)
