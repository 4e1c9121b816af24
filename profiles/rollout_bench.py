"""Throughput of the vectorised DQN rollout (gripper_mujoco_b200/rollout.py) over the CUDA path: policy inference,
fused_step, replay push, masked reset and `updates` optimiser rounds per vector step, nothing leaving the device.
Usage: python profiles/rollout_bench.py [n_envs] [N] [vector_steps] [updates_per_push]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjclass import MjClass  # noqa: E402
from gripper_mujoco_b200.model import GripperParams  # noqa: E402
from gripper_mujoco_b200.rollout import VariableNetwork, VectorDQN, VectorRollout  # noqa: E402
from gripper_mujoco_b200.settings import Settings  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 5
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 40
upd = int(sys.argv[4]) if len(sys.argv) > 4 else 8
st = Settings(); st.debug = False; st.random_seed = 1234; st.set_use_noise(True)
st.mujoco_timestep, st.sim_steps_per_action = ({5: 6.857e-3, 8: 3.225e-3}[N], {5: 30, 8: 63}[N])
mj = MjClass(n, GripperParams(num_segments=N), settings=st, calibrate=False)
ag = VectorDQN(device=mj.device, rngseed=1, updates_per_push=upd)
ag.params.update({"memory_replay": 1_000_000, "min_memory_replay": 5000, "batch_size": 128})    # TrainingManager-scale replay
ag.init(VariableNetwork([mj.get_n_obs(), 150, 100, 50, mj.get_n_actions()], mj.device))        # the reference's usual MLP
ro = VectorRollout(mj, ag)
ro.reset_all()
for _ in range(5):
    ro.step()
torch.cuda.synchronize(); t0 = time.perf_counter()
out = ro.run(steps)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(f"n={n} N={N}: {steps} vector steps in {dt:.3f} s = {n * steps / dt:.0f} transitions/s, {ag.n_updates} updates "
      f"({upd} per vector step), replay {len(ag.memory)}, episodes {out['episodes']}, mean return {out['mean_return']:.3f}")
