"""The environment side of rl/TrainingManager.py: the default settings dictionary and the functions that turn it into
`mj.set` fields -- wipe_cpp_settings (:804-855), apply_env_settings (:937-1058), set_sensor_reward_thresholds (:1071-1140),
set_sensor_bonuses / penalties / terminations (:1142-1198) and create_reward_function with its four styles (:1200-1268).
The agent / trainer / logging side of TrainingManager is outside this path.

`configure_env(env, settings)` accepts anything with `.mj.set`, `.mj.yield_load()`, `.mj.set_sensor_noise_and_normalisation_to`
and `.params.max_episode_steps` (BatchedMjEnv, or a stand-in in the tests); it changes settings only -- they reach the
device at the next reset(), as in the reference.
"""
from __future__ import annotations

import copy

ACTIONS = ("gripper_prismatic_X", "gripper_revolute_Y", "gripper_Z", "base_X", "base_Y", "base_Z", "base_yaw")
SENSORS = ("motor_state_sensor", "base_state_sensor_Z", "base_state_sensor_XY", "base_state_sensor_yaw", "bending_gauge",
           "palm_sensor", "wrist_sensor_XY", "wrist_sensor_Z", "cartesian_contacts_XYZ")
STYLES = ("sensor_mixed_v1", "sparse", "termination_action_v1", "MAT_liftonly")


def _sensor(in_use, normalise, read_rate, noise_override=None):
    return dict(in_use=in_use, normalise=normalise, read_rate=read_rate, noise_override=noise_override)


def default_settings() -> dict:
    """TrainingManager.default_settings (rl/TrainingManager.py:64-368), env / cpp / reward parts."""
    return {
        "env": dict(max_episode_steps=250, object_position_noise_mm=10, object_rotation_noise_deg=5,
                    base_lim_X_mm=300, base_lim_Y_mm=200, base_lim_Z_mm=30, test_obj_per_file=20,
                    task_reload_chance=0.05, test_trials_per_object=3, test_objects=100,
                    object_set_name="set9_nosharp_smallspheres", num_segments=8, finger_thickness=0.86e-3,
                    finger_length=235e-3, finger_width=28e-3, finger_modulus=193e9, XY_base_actions=False,
                    Z_base_rotation=False, fixed_finger_hook=True, finger_hook_angle_degrees=75.0,
                    finger_hook_length=35e-3, segment_inertia_scaling=50.0, fingertip_clearance=0.01),
        "cpp": {
            "randomise_colours": False, "time_for_action": 0.2, "saturation_yield_factor": 1.5,
            "sensor_sample_mode": 2, "state_sample_mode": 4, "sensor_n_prev_steps": 3, "state_n_prev_steps": 3,
            "sensor_noise_mu": 0.05, "sensor_noise_std": 0.025, "state_noise_mu": 0.025, "state_noise_std": 0.0,
            "base_position_noise": 5e-3, "palm_scale_factor": 1.0, "oob_distance": 75e-3, "lift_height": 15e-3,
            "gripper_target_height": 20e-3, "stable_finger_force": 1.0, "stable_palm_force": 1.0,
            "stable_finger_force_lim": 4.0, "stable_palm_force_lim": 4.0, "XY_distance_threshold": 10e-3,
            "cap_reward": False, "fingertip_min_mm": -12.5, "continous_actions": True, "use_termination_action": False,
            "termination_threshold": 0.9,
            "action": {
                "gripper_prismatic_X": dict(in_use=True, value=2e-3, sign=-1),
                "gripper_revolute_Y": dict(in_use=True, value=0.015, sign=-1),
                "gripper_Z": dict(in_use=True, value=4e-3, sign=1),
                "base_X": dict(in_use=False, value=2e-3, sign=1),
                "base_Y": dict(in_use=False, value=2e-3, sign=1),
                "base_Z": dict(in_use=True, value=2e-3, sign=1),
                "base_yaw": dict(in_use=False, value=5e-3, sign=1),
            },
            "sensor": {
                "motor_state_sensor": _sensor(True, 0.0, -1),
                "base_state_sensor_Z": _sensor(True, 0.0, -1, [0.1, 0.0]),
                "base_state_sensor_XY": _sensor(False, 0.0, -1),
                "base_state_sensor_yaw": _sensor(False, 0.0, -1),
                "bending_gauge": _sensor(True, 20.0, 10),
                "palm_sensor": _sensor(True, 6.0, 10),
                "wrist_sensor_XY": _sensor(False, 5.0, 10),
                "wrist_sensor_Z": _sensor(True, 10.0, 10),
                "cartesian_contacts_XYZ": _sensor(False, 0.0, -1, [0, 0]),
            },
        },
        "reward": {
            "style": "sensor_mixed_v1", "options": [], "scale_rewards": 1.0, "scale_penalties": 1.0,
            "penalty_termination": True, "stable_trigger": 1, "dangerous_trigger": 1,
            "bend": dict(min=0.0, good="stable", exceed="stable_lim", dangerous=5.0),
            "palm": dict(min=0.0, good="stable", exceed="stable_lim", dangerous=5.0),
            "wrist": dict(exceed=6.0, dangerous=8.0),
            "action_pen_lin": dict(used=False, scaling=2, min=0.1, max=2.0),
            "action_pen_sq": dict(used=True, scaling=2, min=0.1, max=3.0),
            "object_XY_distance": dict(used=False, min=-50e-3, max=-10e-3),
        },
    }


def preset(style: str = "sensor_mixed_v1", **reward_overrides) -> dict:
    """The defaults with another reward style; `termination_action_v1` and `MAT_liftonly` need the termination action,
    which the reference's launch scripts switch on together with the style."""
    if style not in STYLES:
        raise RuntimeError(f"style={style} is not a valid option in TrainingManager.create_reward_function()")
    s = default_settings()
    s["reward"]["style"] = style
    s["reward"].update(reward_overrides)
    if style in ("termination_action_v1", "MAT_liftonly"):
        s["cpp"]["use_termination_action"] = True
    return s


def wipe_cpp_settings(env):
    """rl/TrainingManager.py:804-855: a blank slate -- no actions, no rewards, no sensors, no noise."""
    st = env.mj.set
    st.auto_set_timestep, st.auto_calibrate_gauges, st.auto_sim_steps = True, True, True
    st.auto_exceed_lateral_lim = False
    st.curve_validation = False
    st.render_on_step = False
    st.set_all_action_use(False); st.set_all_action_continous(False)
    st.set_all_action_value(0.0); st.set_all_action_sign(1)
    st.use_termination_action = False
    st.wipe_rewards()
    st.cap_reward, st.quit_if_cap_exceeded = False, False
    st.reward_cap_lower_bound, st.reward_cap_upper_bound = -1e6, 1e6
    st.use_HER = False
    st.disable_sensors()
    env.mj.set_sensor_noise_and_normalisation_to(True)
    st.sensor_n_prev_steps = st.state_n_prev_steps = 1
    st.motor_state_sensor.read_rate = st.base_state_sensor_XY.read_rate = st.base_state_sensor_Z.read_rate = -1
    st.sensor_noise_mag = st.sensor_noise_mu = st.sensor_noise_std = 0
    st.state_noise_mag = st.state_noise_mu = st.state_noise_std = 0
    st.base_position_noise = 0
    return env


def apply_env_settings(env, settings: dict):
    """rl/TrainingManager.py:937-1058 (rewards are set by create_reward_function, not here)."""
    for k, v in settings["env"].items():
        if hasattr(env.params, k):
            setattr(env.params, k, v)
    st, c = env.mj.set, settings["cpp"]
    for k in ("randomise_colours", "time_for_action", "saturation_yield_factor", "sensor_sample_mode", "state_sample_mode",
              "sensor_n_prev_steps", "state_n_prev_steps", "sensor_noise_mu", "sensor_noise_std", "state_noise_mu",
              "state_noise_std", "base_position_noise", "palm_scale_factor", "oob_distance", "lift_height",
              "gripper_target_height", "stable_finger_force", "stable_palm_force", "stable_finger_force_lim",
              "stable_palm_force_lim", "XY_distance_threshold", "fingertip_min_mm", "continous_actions",
              "use_termination_action", "termination_threshold"):
        setattr(st, k, c[k])
    for nm in ACTIONS:
        a, d = getattr(st, nm), c["action"][nm]
        a.in_use, a.value, a.sign = bool(d["in_use"]), float(d["value"]), int(d["sign"])
    for nm in SENSORS:
        sn, d = getattr(st, nm), c["sensor"][nm]
        sn.in_use, sn.normalise, sn.read_rate = bool(d["in_use"]), float(d["normalise"]), float(d["read_rate"])
        if d["noise_override"] is not None:
            sn.set_gaussian_noise(*d["noise_override"])
    return env


class RewardThresholds:
    """m = minimum, g = good, x = exceed, d = dangerous (rl/TrainingManager.py:1098-1114)."""

    def __init__(self, env, r: dict):
        st = env.mj.set
        self.mBend = r["bend"]["min"]
        self.gBend = st.stable_finger_force if r["bend"]["good"] == "stable" else r["bend"]["good"]
        self.xBend = st.stable_finger_force_lim if r["bend"]["exceed"] == "stable_lim" else r["bend"]["exceed"]
        self.dBend = env.mj.yield_load() if r["bend"]["dangerous"] == "yield" else r["bend"]["dangerous"]
        self.mPalm = r["palm"]["min"]
        self.gPalm = st.stable_palm_force if r["palm"]["good"] == "stable" else r["palm"]["good"]
        self.xPalm = st.stable_palm_force_lim if r["palm"]["exceed"] == "stable_lim" else r["palm"]["exceed"]
        self.dPalm = r["palm"]["dangerous"]
        self.xWrist, self.dWrist = r["wrist"]["exceed"], r["wrist"]["dangerous"]
        for lo, hi in (("mBend", "gBend"), ("mPalm", "gPalm"), ("gBend", "xBend"), ("gPalm", "xPalm"), ("xBend", "dBend"),
                       ("xPalm", "dPalm"), ("xWrist", "dWrist")):
            if getattr(self, lo) > getattr(self, hi):
                raise RuntimeError(f"set_sensor_reward_thresholds() finds {lo} > {hi}, "
                                   f"{getattr(self, lo):.3f} > {getattr(self, hi):.3f}")


def set_sensor_bonuses(env, RT, r: dict, value: float):
    st = env.mj.set                                     # reward  done  trigger  min  max  overshoot
    st.lifted.set(value, False, 1)
    st.lifted_to_height.set(value, False, 1)
    st.object_stable.set(value, False, 1)
    st.good_bend_sensor.set(value, False, 1, RT.mBend, RT.gBend, -1)
    st.good_palm_sensor.set(value, False, 1, RT.mPalm, RT.gPalm, -1)
    if r["object_XY_distance"]["used"]:
        st.object_XY_distance.set(value, False, 1, r["object_XY_distance"]["min"], r["object_XY_distance"]["max"], -1)


def set_sensor_penalties(env, RT, r: dict, value: float):
    st = env.mj.set
    st.exceed_limits.set(value, False, 1)
    st.exceed_bend_sensor.set(value, False, 1, RT.xBend, RT.dBend, -1)
    st.exceed_palm_sensor.set(value, False, 1, RT.xPalm, RT.dPalm, -1)
    st.exceed_wrist_sensor.set(value, False, 1, RT.xWrist, RT.dWrist, -1)
    for key, ev in (("action_pen_lin", st.action_penalty_lin), ("action_pen_sq", st.action_penalty_sq)):
        if r[key]["used"]:
            ev.set(value * r[key]["scaling"], False, 1, r[key]["min"], r[key]["max"], -1)


def set_sensor_terminations(env, RT, value=-1.0, done=True, trigger=1):
    st = env.mj.set
    st.dangerous_bend_sensor.set(value, done, trigger, RT.dBend, RT.dBend, -1)
    st.dangerous_palm_sensor.set(value, done, trigger, RT.dPalm, RT.dPalm, -1)
    st.dangerous_wrist_sensor.set(value, done, trigger, RT.dWrist, RT.dWrist, -1)


def create_reward_function(env, settings: dict):
    """rl/TrainingManager.py:1200-1268."""
    r, st = settings["reward"], env.mj.set
    style = r["style"]
    if style in ("sensor_mixed_v1", "termination_action_v1"):
        RT = RewardThresholds(env, r)
        st.step_num.set(-0.01, False, 1)
        set_sensor_bonuses(env, RT, r, 0.002 * r["scale_rewards"])
        set_sensor_penalties(env, RT, r, -0.002 * r["scale_penalties"])
        st.scale_rewards(100 / env.params.max_episode_steps)
        if style == "sensor_mixed_v1":
            st.stable_height.set(1.0, True, 1)
        else:
            st.stable_termination.set(1.0, True, 1)
            st.failed_termination.set(-1.0, True, 1)
        st.oob.set(-1.0, True, 1)
        if r["penalty_termination"]:
            set_sensor_terminations(env, RT, trigger=r["dangerous_trigger"])
    elif style == "sparse":
        RT = RewardThresholds(env, r)
        st.stable_height.set(1.0, True, 1)
        st.oob.set(-1.0, True, 1)
        if r["penalty_termination"]:
            set_sensor_terminations(env, RT, trigger=r["dangerous_trigger"])
    elif style == "MAT_liftonly":
        st.lifted_termination.set(1.0, True, 1)
        st.failed_termination.set(0.0, True, 1)
        if r["penalty_termination"]:
            st.oob.set(0.0, True, 1)
            RT = RewardThresholds(env, r)
            set_sensor_terminations(env, RT, trigger=r["dangerous_trigger"], value=0.0)
    else:
        raise RuntimeError(f"style={style} is not a valid option in TrainingManager.create_reward_function()")
    if "terminate_on_exceed_limits" in r["options"]:
        st.exceed_limits.set(-1.0, True, 3)
    st.object_stable.trigger = r["stable_trigger"]
    return env


def configure_env(env, settings: dict | None = None):
    """TrainingManager.configure_env (rl/TrainingManager.py:1060-1069)."""
    settings = copy.deepcopy(settings) if settings is not None else default_settings()
    wipe_cpp_settings(env)
    apply_env_settings(env, settings)
    create_reward_function(env, settings)
    return env
