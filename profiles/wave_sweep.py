"""Step time vs number of blocks (12 envs each): wave quantisation on 148 SMs."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjclass import MjClass
from gripper_mujoco_b200.model import GripperParams
from gripper_mujoco_b200.settings import Settings
print("SMs:", torch.cuda.get_device_properties(0).multi_processor_count)
for blocks in [1, 37, 74, 120, 132, 140, 148, 149, 160, 222, 296, 342, 444]:
    n = 12 * blocks
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
    ms = e0.elapsed_time(e1) / 8
    print(f"blocks={blocks:4d} n={n:5d}: {ms:.3f} ms/step  {n / ms:.0f} env-steps/ms")
    mj.close()
