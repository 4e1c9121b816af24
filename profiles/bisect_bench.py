"""Why is the bench workload slower per step than profile_step.py? Times gmj_step alone under variants."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjclass import MjClass
from gripper_mujoco_b200.model import GripperParams
from gripper_mujoco_b200.settings import Settings

def run(term, resets, steps=40, n=4096):
    st = Settings(); st.debug = False; st.random_seed = 1234; st.set_use_noise(True)
    st.mujoco_timestep, st.sim_steps_per_action = 6.825e-3, 30
    st.bending_gauge.normalise = 2.41; st.wrist_sensor_Z.raw_value_offset = -23.3
    mj = MjClass(n, GripperParams(num_segments=5), settings=st, calibrate=False)
    mj.batch.spawn_with_retry(0.01, float(np.pi / 2))
    g = torch.Generator(device="cpu"); g.manual_seed(0)
    acts = torch.randint(0, mj.get_n_actions() - (0 if term else 1), (steps, n), generator=g, dtype=torch.int32).to(mj.device)
    gb = mj.batch
    tot = 0.0
    for k in range(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gb.step(acts[k]); e1.record()
        if resets:
            m8 = (gb.done != 0).to(torch.uint8)
            gb.reset(m8); gb.spawn_with_retry(0.01, float(np.pi / 2), m8)
        torch.cuda.synchronize()
        if k >= 10: tot += e0.elapsed_time(e1)
    s = gb.stats()
    print(f"term={term} resets={resets}: {tot / (steps - 10):.3f} ms/step  substeps/env/step {s['substeps'] / n / steps:.1f} contacts/substep {s['contacts'] / s['substeps']:.2f} rows/substep {s['rows'] / s['substeps']:.1f}")
    mj.close()

for term, resets in [(True, True)]:
    run(term, resets)
