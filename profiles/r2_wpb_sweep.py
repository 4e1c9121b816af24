"""Throughput of the lockstep block vs warps per block / blocks per SM: lift-free random actions (no termination
action), one and two waves of blocks over the SMs. GMJ_LIB selects the build, WPB / BPS describe it.
usage: GMJ_LIB=... python profiles/r2_wpb_sweep.py WPB BPS"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjclass import MjClass
from gripper_mujoco_b200.model import GripperParams
from gripper_mujoco_b200.settings import Settings
wpb, bps = int(sys.argv[1]), int(sys.argv[2])
nsm = torch.cuda.get_device_properties(0).multi_processor_count
for waves in (1, 2, 4):
    n = wpb * bps * nsm * waves
    st = Settings(); st.debug = False; st.random_seed = 1234; st.set_use_noise(True)
    st.mujoco_timestep, st.sim_steps_per_action = 6.857e-3, 30
    st.bending_gauge.normalise = 2.41; st.wrist_sensor_Z.raw_value_offset = -23.3
    mj = MjClass(n, GripperParams(num_segments=5), settings=st, calibrate=False)
    mj.batch.spawn_with_retry(0.01, float(np.pi / 2))
    g = torch.Generator(device="cpu"); g.manual_seed(0)
    acts = torch.randint(0, 14, (24, n), generator=g, dtype=torch.int32).to(mj.device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for k in range(24):
        if k == 16: e0.record()
        mj.fused_step(acts[k])
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 8
    s = mj.batch.stats()
    print(f"wpb={wpb} bps={bps} waves={waves} n={n:6d}: {ms:7.3f} ms/step  {n / ms:7.0f} env-steps/ms  "
          f"us/substep/block-round {1e3 * ms / 30 / waves:6.1f}  ncon/substep {s['contacts'] / s['substeps']:.2f}", flush=True)
    mj.close()
