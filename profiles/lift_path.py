"""Critical path of the termination-lift blocks: n envs that ALL take the termination action."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjclass import MjClass
from gripper_mujoco_b200.model import GripperParams
from gripper_mujoco_b200.settings import Settings

def run(n, term, steps=12):
    st = Settings(); st.debug = False; st.random_seed = 1234; st.set_use_noise(True)
    st.mujoco_timestep, st.sim_steps_per_action = 6.825e-3, 30
    st.bending_gauge.normalise = 2.41; st.wrist_sensor_Z.raw_value_offset = -23.3
    mj = MjClass(n, GripperParams(num_segments=5), settings=st, calibrate=False)
    mj.batch.spawn_with_retry(0.01, float(np.pi / 2))
    gb = mj.batch
    g = torch.Generator(device="cpu"); g.manual_seed(0)
    warm = torch.randint(0, 14, (6, n), generator=g, dtype=torch.int32).to(mj.device)
    for k in range(6): gb.step(warm[k])
    a = torch.full((n,), 14 if term else 3, dtype=torch.int32, device=mj.device)
    tot = 0
    for k in range(steps):
        m8 = torch.ones(n, dtype=torch.uint8, device=mj.device)
        if term:
            gb.reset(m8); gb.spawn_with_retry(0.01, float(np.pi / 2), m8); gb.step(warm[k % 6])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gb.step(a); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    print(f"n={n} term={term}: {tot / steps:.3f} ms/step")
    mj.close()

for n in (12, 276, 1776):
    run(n, False); run(n, True)
