"""Per-env cost anatomy (development build -DGMJ_PROF): cycles per phase of the substep, own work only, and the
barrier waits; which envs are the stragglers of a launch. usage: GMJ_LIB=build/variants/lib_prof.so python profiles/r2_env_cost.py [n] [lift]"""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from gripper_mujoco_b200.mjclass import MjClass
from gripper_mujoco_b200.model import GripperParams
from gripper_mujoco_b200.settings import Settings
from gripper_mujoco_b200.lib import GP
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1776
lift = int(sys.argv[2]) if len(sys.argv) > 2 else 0
st = Settings(); st.debug = False; st.random_seed = 1234; st.set_use_noise(True)
st.mujoco_timestep, st.sim_steps_per_action = 6.857e-3, 30
st.bending_gauge.normalise = 2.41; st.wrist_sensor_Z.raw_value_offset = -23.3
mj = MjClass(n, GripperParams(num_segments=5), settings=st, calibrate=False)
gb = mj.batch
L = gb.L
L.gmj_prof_host.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
mj.batch.spawn_with_retry(0.01, float(np.pi / 2))
g = torch.Generator(device="cpu"); g.manual_seed(0)
K = 24
acts = torch.randint(0, 15 if lift else 14, (K, n), generator=g, dtype=torch.int32).to(mj.device)
ep = torch.zeros(n, dtype=torch.int32, device=mj.device)
prof = np.zeros((n, 12), dtype=np.uint32)
names = ["kin", "spatial", "control", "factor", "collide", "rows", "formA", "lcp", "integ", "tail", "barrier", "loop"]
for k in range(K):
    if k == 16:
        L.gmj_prof_host(gb.h, prof.ctypes.data, 1)
        s0 = gb.stats()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e0.record()
    gb.step(acts[k])
    if lift:
        ep.add_(1); mask = (gb.done != 0) | (ep >= 250); m8 = mask.to(torch.uint8)
        gb.reset(m8); gb.spawn_with_retry(0.01, float(np.pi / 2), m8); ep.mul_((~mask).to(torch.int32))
e1.record(); torch.cuda.synchronize()
L.gmj_prof_host(gb.h, prof.ctypes.data, 0)
s1 = gb.stats()
nsub = (s1["substeps"] - s0["substeps"]) / n
p = prof.astype(np.float64)
print(f"n={n} lift={lift} {e0.elapsed_time(e1) / 8:.3f} ms/step; substeps/env {nsub:.1f}; clock 1.965 GHz assumed")
own = p[:, :10].sum(1)
print(f"per env-substep cycles: own work mean {own.mean() / nsub:.0f}  p50 {np.median(own) / nsub:.0f}  p90 {np.percentile(own, 90) / nsub:.0f} "
      f"p99 {np.percentile(own, 99) / nsub:.0f} max {own.max() / nsub:.0f}; barrier wait mean {p[:, 10].mean() / nsub:.0f}; loop total mean {p[:, 11].mean() / nsub:.0f} max {p[:, 11].max() / nsub:.0f}")
for i, nm in enumerate(names[:11]):
    c = p[:, i] / nsub
    print(f"  {nm:8s} mean {c.mean():8.0f}  p50 {np.median(c):8.0f}  p90 {np.percentile(c, 90):8.0f}  p99 {np.percentile(c, 99):8.0f}  max {c.max():8.0f}   share of own {100 * p[:, i].sum() / own.sum():5.1f}%")
pr = gb.probe()
top = np.argsort(-own)[:12]
print("heaviest envs: env, own cycles/substep, [phase cycles/substep], mean ncon, mean lcp rounds, nrow now, unstable, obj")
for e in top:
    print(f"  {e:6d} {own[e] / nsub:8.0f} {np.round(p[e, :10] / nsub).astype(int)} ncon {pr[e, GP['ncon']]:.2f} niter {pr[e, GP['niter']]:.2f} "
          f"nrow {pr[e, GP['nrow']]:.0f} unst {pr[e, GP['unstable']]:.0f} obj {int(pr[e, GP['obj_index']])}")
