"""Golden vectors of the NON-PHYSICS TAIL, produced by the reference's own compiled code.

oracle/_ref/libmjclass_ref.so (`make -C oracle ref`) is the reference's src/mjclass.cpp + gripper.cpp compiled in place
against the stand-in headers of oracle/stub/ and driven by oracle/ref_tail_shim.cpp over SCRIPTED physics observables
(gauge readings, per-class force sums, wrist force, object position per substep). For every scenario below the real
MjClass runs reset -> spawn_object -> T x (set_action, action_step, get_observation, is_done, reward,
get_event_state); inputs and outputs go to tests/golden/tail_golden.npz. tests/test_ref_tail.py replays the same
inputs through the oracle's restatement (oracle/env.c, events.c, kin.c) and demands bit-identical results.

Run in the build container (the reference is not on the GPU box):  python tests/golden/make_tail_golden.py
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from gripper_mujoco_b200.model import GripperParams, build_model  # noqa: E402
from gripper_mujoco_b200.settings import GmjSettings, Settings  # noqa: E402

ROWLEN = 46


def ref_lib():
    R = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libmjclass_ref.so"))
    R.ref_mj_new.restype = C.c_void_p
    R.ref_mj_new.argtypes = [C.c_void_p, C.c_void_p]
    R.ref_mj_free.argtypes = [C.c_void_p]
    R.ref_mj_reset.argtypes = [C.c_void_p]
    R.ref_mj_spawn.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_void_p]
    R.ref_mj_try_spawn.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_void_p]
    R.ref_mj_set_script.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    R.ref_mj_script_cursor.argtypes = [C.c_void_p]
    R.ref_mj_n_actions.argtypes = [C.c_void_p]
    R.ref_mj_n_obs.argtypes = [C.c_void_p]
    R.ref_mj_step.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_int] + [C.c_void_p] * 5
    R.ref_mj_sensor_layout.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    R.ref_linear_reward.argtypes = [C.c_float] * 4
    R.ref_linear_reward.restype = C.c_float
    R.ref_normalise_between.argtypes = [C.c_float] * 3
    R.ref_normalise_between.restype = C.c_float
    R.ref_box2d_overlap.argtypes = [C.c_double] * 11
    R.ref_sample.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
    return R


def geom_of(N):
    m = build_model(GripperParams(num_segments=N)).to_struct()
    return np.array([m.finger_length, m.EI, m.fingertip_clearance, m.hook_length, m.hook_angle_deg, m.finger_width, N])


# ------------------------------------------------------------------------------------------------ scenarios
def base_settings(dt, S, seed=1):
    s = Settings()
    s.debug = False
    s.mujoco_timestep, s.sim_steps_per_action = dt, S
    s.random_seed = seed
    s.set_use_noise(False)
    s.use_noise_seeded = 0
    s.base_position_noise = 0.0
    s.sim_gauge_raw_to_N_factor = 0.37
    s.bending_gauge.normalise = 2.4
    s.wrist_sensor_Z.raw_value_offset = -23.3
    return s


def training_preset(s):
    """rl/TrainingManager.py:181-239 as a DQN run uses it: 4 actions x 2, no termination action, average-mode
    sensors, 9 streams x 7 = 63 observations."""
    s.sensor_sample_mode, s.state_sample_mode = 2, 4
    s.use_termination_action = False
    for n in ("gripper_X", "gripper_Y", "base_X", "base_Y", "base_roll", "base_pitch", "base_yaw"):
        getattr(s, n).in_use = False
    s.gripper_prismatic_X.set(True, 2e-3, -1); s.gripper_revolute_Y.set(True, 0.015, -1)
    s.gripper_Z.set(True, 4e-3, 1); s.base_Z.set(True, 2e-3, 1)
    s.base_state_sensor_XY.in_use = False; s.base_state_sensor_yaw.in_use = False
    s.saturation_yield_factor = 1.5
    s.stable_finger_force, s.stable_palm_force = 1.0, 1.0
    s.stable_finger_force_lim, s.stable_palm_force_lim = 5.0, 10.0
    s.lifted.set(0.002, 0, 1); s.object_stable.set(0.5, 0, 1); s.stable_height.set(1.0, 1, 1)
    s.dangerous_bend_sensor.set(-1.0, 1, 1, 4.0, 4.5, -1); s.dangerous_palm_sensor.set(-1.0, 1, 1, 12.0, 13.0, -1)
    s.exceed_bend_sensor.set(-0.02, 0, 1, 2.0, 4.0, -1)
    s.good_bend_sensor.set(0.005, 0, 1, 0.5, 2.0, -1)
    return s


def scenarios():
    out = []
    out.append(("header_default_N5", 5, base_settings(6.857e-3, 30), False))
    out.append(("header_default_N8", 8, base_settings(3.225e-3, 63, seed=7), False))
    out.append(("training_preset", 8, training_preset(base_settings(3.187e-3, 63, seed=3)), False))
    for mode in range(7):
        s = base_settings(6.9e-3, 29, seed=10 + mode)
        s.sensor_sample_mode = mode
        s.state_sample_mode = (mode + 3) % 7
        s.sensor_n_prev_steps = 1 + mode % 3
        s.state_n_prev_steps = 1 + (mode + 1) % 3
        s.axial_gauge.in_use = True
        s.wrist_sensor_XY.in_use = bool(mode & 1)
        s.cartesian_contacts_XYZ.in_use = bool(mode & 2)
        out.append((f"sampler_mode_{mode}", 5, s, False))
    s = base_settings(6.857e-3, 30, seed=21)
    s.continous_actions = True
    s.cartesian_contacts_XYZ.in_use = True
    out.append(("continuous_actions", 5, s, True))
    s = base_settings(6.857e-3, 30, seed=22)
    s.quit_if_cap_exceeded = True
    s.reward_cap_lower_bound, s.reward_cap_upper_bound = -0.3, 0.4
    s.lifted_termination.done = 0                 # termination resolves through stable_height instead
    s.stable_height.set(1.0, 1, 2)
    s.oob.set(-1.0, 3, 2)
    s.scale_rewards(1.7)
    out.append(("reward_cap_and_triggers", 5, s, False))
    s = base_settings(4e-3, 75, seed=23)             # dt * S = 0.3: three sensor readings per step
    s.lift_on_termination = False
    s.use_normalisation = None
    s.set_use_normalisation(False)
    out.append(("three_readings_no_normalisation", 5, s, False))
    return out


def make_script(rng, nrows, S):
    """Smooth random observables with exact zeros and threshold crossings (lifted / contact / stable / oob)."""
    t = np.arange(nrows) / S
    rows = np.zeros((nrows, ROWLEN))

    def smooth(scale, n=1):
        a = rng.normal(size=(4, n)) * scale
        ph = rng.uniform(0, 6.28, size=(4, n)); w = rng.uniform(0.05, 0.9, size=(4, n))
        return sum(a[k] * np.sin(w[k] * t[:, None] + ph[k]) for k in range(4))

    rows[:, 0:3] = smooth(9.0, 3)                                   # raw gauges (mm * 1000 scale of the cubic fit)
    grasp = (np.sin(0.23 * t + rng.uniform(0, 6)) > 0.1)[:, None]    # phases with / without finger contact
    ff = np.abs(smooth(2.5, 9)) * grasp
    rows[:, 15:24] = ff * np.array([1, -1, 1] * 3)                  # obj finger local (y negative = outward lateral)
    rows[:, 3:12] = rows[:, 15:24] + smooth(0.05, 9) * (rng.random((nrows, 1)) < 0.3)
    palm = np.clip(smooth(6.0, 3), 0, None) * (np.sin(0.17 * t + 1.0) > 0)[:, None]
    rows[:, 24:27] = palm; rows[:, 12:15] = palm
    lifted = (np.sin(0.11 * t + rng.uniform(0, 6)) > 0.2)[:, None]
    rows[:, 27:30] = np.abs(smooth(1.5, 3)) * (~lifted)             # ground force on the object: exactly 0 when lifted
    rows[:, 30:39] = smooth(1.0, 9) * (rng.random((nrows, 1)) < 0.1)  # fingers on the ground (rare)
    rows[:, 39] = -23.3 + smooth(4.0, 1)[:, 0]                      # wrist z raw
    rows[:, 40] = smooth(0.03, 1)[:, 0]; rows[:, 41] = smooth(0.03, 1)[:, 0]
    rows[:, 40] += np.where(t > 0.8 * t[-1], 0.09, 0.0)             # drifts out of bounds late in the trace
    rows[:, 42] = 0.03 + 0.025 * np.clip(np.cumsum(lifted[:, 0] - 0.45) / S / 3, 0, 1.2)
    return rows.astype(np.float32).astype(np.float64)       # float32-representable: the fixture stores them as float32


def run_scenario(R, name, N, st, continuous, T=40, seed=0):
    rng = np.random.default_rng(abs(hash(name)) % (2 ** 31) if False else sum(map(ord, name)) + seed)
    S = int(st.sim_steps_per_action)
    packed = st.pack()
    geom = geom_of(N)
    h = R.ref_mj_new(C.byref(packed), geom.ctypes.data)
    n_act = R.ref_mj_n_actions(h)
    n_obs = R.ref_mj_n_obs(h)
    prev, rd = (C.c_int * 10)(), (C.c_int * 10)()
    R.ref_mj_sensor_layout(h, prev, rd)
    rows = make_script(rng, 3 * S * T + 8, S)
    spawn = np.array([0.004, -0.006, 0.3, 0.031, 0.06, 0.05, 0.062])    # x y rot z_rest bbox
    actions = rng.integers(0, n_act, T).astype(np.int32)
    frac = rng.uniform(-1.3, 1.3, T).astype(np.float32) if continuous else np.zeros(T, dtype=np.float32)
    if continuous:
        actions[rng.random(T) < 0.1] = n_act - 1      # termination action with fraction above / below the threshold
    resets = np.zeros(T, dtype=np.int8)
    resets[0] = 1; resets[T // 2] = 1                 # a second episode in the same process (stale state must clear)
    obs = np.zeros((T, n_obs), dtype=np.float32); rew = np.zeros(T, dtype=np.float32)
    done = np.zeros(T, dtype=np.int32); ev = np.zeros((T, 57), dtype=np.float32); tgt = np.zeros((T, 4), dtype=np.float32)
    cursor = np.zeros(T + 1, dtype=np.int32)
    R.ref_mj_set_script(h, rows.ctypes.data, len(rows))
    for t in range(T):
        if resets[t]:
            R.ref_mj_reset(h)
            R.ref_mj_spawn(h, spawn[0], spawn[1], spawn[2], spawn[3], spawn[4:].ctypes.data)
        d = C.c_int(); r = C.c_float()
        n = R.ref_mj_step(h, int(actions[t]), float(frac[t]), int(continuous), obs[t].ctypes.data, C.byref(r), C.byref(d),
                          ev[t].ctypes.data, tgt[t].ctypes.data)
        assert n == n_obs
        rew[t], done[t] = r.value, d.value
        cursor[t + 1] = R.ref_mj_script_cursor(h)
    R.ref_mj_free(h)
    sb = np.frombuffer(bytes(packed), dtype=np.float64).copy()
    return {"N": N, "settings": sb, "geom": geom, "rows": rows[: cursor[-1] + 1].astype(np.float32), "spawn": spawn, "actions": actions, "frac": frac,
            "continuous": int(continuous), "resets": resets, "obs": obs, "reward": rew, "done": done, "events": ev,
            "target": tgt, "cursor": cursor, "n_actions": n_act, "prev_steps": np.array(prev[:]), "readings": np.array(rd[:])}


def helpers(R, rng):
    """Free functions / small types of the reference on random inputs."""
    lr_in = rng.uniform(-3, 12, (600, 4)).astype(np.float32)
    lr_in[:200, 3] = -1                                   # overshoot disabled
    lr_in[200:300, 3] = lr_in[200:300, 2] - 0.5           # overshoot below max
    lr_out = np.array([R.ref_linear_reward(*map(float, r)) for r in lr_in], dtype=np.float32)
    nb_in = rng.uniform(-1, 1, (300, 3)).astype(np.float32); nb_in[:, 2] = nb_in[:, 1] + np.abs(nb_in[:, 2]) + 0.01
    R2 = R.ref_normalise_between
    nb_out = np.array([R2(*map(float, r)) for r in nb_in], dtype=np.float32)
    bx = rng.uniform(-0.05, 0.05, (800, 11)); bx[:, 2:4] = np.abs(bx[:, 2:4]) + 0.01; bx[:, 7:9] = np.abs(bx[:, 7:9]) + 0.01
    bx[:, 4] *= 40; bx[:, 9] *= 40; bx[:, 10] = np.abs(bx[:, 10]) * 0.2
    bx[:100, 5:9] = bx[:100, 0:4]; bx[:100, 7:9] *= 0.3   # contained boxes (the quirk branch of overlapsWith)
    bx_out = np.array([R.ref_box2d_overlap(*map(float, r)) for r in bx], dtype=np.int8)
    samp = []
    for mode in range(7):
        for p in (1, 2, 3):
            for r in (1, 2, 3):
                n = int(rng.integers(0, 14))
                vals = rng.normal(size=n).astype(np.float32) * (0.2 if mode in (4, 5, 6) else 1.0)
                if n > 3 and rng.random() < 0.3:
                    vals[-2] = vals[-3]                     # exact ties for sign / median
                out = np.zeros(16, dtype=np.float32)
                k = R.ref_sample(mode, p, r, vals.ctypes.data, n, out.ctypes.data)
                samp.append((mode, p, r, vals.copy(), out[:k].copy()))
    return {"lr_in": lr_in, "lr_out": lr_out, "nb_in": nb_in, "nb_out": nb_out, "box_in": bx, "box_out": bx_out,
            "samp_meta": np.array([(m, p, r, len(v), len(o)) for m, p, r, v, o in samp], dtype=np.int32),
            "samp_vals": np.concatenate([v for *_, v, o in samp]), "samp_out": np.concatenate([o for *_, o in samp])}


def spawn_cases(R, rng):
    """MjClass::spawn_into_scene with zero ranges = the reference's accept / reject decision for ONE pose."""
    st = base_settings(6.857e-3, 30)
    packed = st.pack()
    geom = geom_of(5)
    h = R.ref_mj_new(C.byref(packed), geom.ctypes.data)
    cases = np.zeros((500, 7)); res = np.zeros(500, dtype=np.int8)
    for i in range(500):
        x, y = rng.uniform(-0.12, 0.12, 2)
        rot = rng.uniform(-1.6, 1.6)
        bb = rng.uniform(0.02, 0.16, 3)
        cases[i] = [x, y, rot, 0.03, *bb]
        R.ref_mj_reset(h)
        res[i] = R.ref_mj_try_spawn(h, x, y, rot, 0.03, bb.ctypes.data)
    R.ref_mj_free(h)
    return {"spawn_cases": cases, "spawn_ok": res, "spawn_geom": geom, "spawn_settings": np.frombuffer(bytes(packed), dtype=np.float64).copy()}


def main():
    R = ref_lib()
    out = {}
    names = []
    for name, N, st, cont in scenarios():
        d = run_scenario(R, name, N, st, cont)
        names.append(name)
        for k, v in d.items():
            out[f"{name}/{k}"] = v
        print(f"{name:36s} n_actions {d['n_actions']:2d} n_obs {d['obs'].shape[1]:3d} done {int(d['done'].sum()):2d} "
              f"reward sum {d['reward'].sum():+.3f} rows {len(d['rows'])}")
    out["names"] = np.array(names)
    rng = np.random.default_rng(99)
    out.update({f"helpers/{k}": v for k, v in helpers(R, rng).items()})
    out.update({f"spawn/{k}": v for k, v in spawn_cases(R, rng).items()})
    np.savez_compressed(os.path.join(HERE, "tail_golden.npz"), **out)
    print("written", os.path.join(HERE, "tail_golden.npz"), os.path.getsize(os.path.join(HERE, "tail_golden.npz")), "bytes")


if __name__ == "__main__":
    main()
