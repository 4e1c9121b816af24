"""One lockstep block alone on the GPU (12 envs): latency anatomy of the substep without chip-level contention."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjclass import MjClass
from gripper_mujoco_b200.model import GripperParams
from gripper_mujoco_b200.settings import Settings
n = int(sys.argv[1]) if len(sys.argv) > 1 else 12
st = Settings(); st.debug = False; st.random_seed = 1234; st.set_use_noise(True)
st.mujoco_timestep, st.sim_steps_per_action = 6.825e-3, 30
st.bending_gauge.normalise = 2.41; st.wrist_sensor_Z.raw_value_offset = -23.3
mj = MjClass(n, GripperParams(num_segments=5), settings=st, calibrate=False)
mj.batch.spawn_with_retry(0.01, float(np.pi / 2))
g = torch.Generator(device="cpu"); g.manual_seed(0)
acts = torch.randint(0, 14, (16, n), generator=g, dtype=torch.int32).to(mj.device)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for k in range(16):
    if k == 8: e0.record()
    mj.fused_step(acts[k])
e1.record(); torch.cuda.synchronize()
print(f"n={n}: {e0.elapsed_time(e1) / 8:.3f} ms/step")
