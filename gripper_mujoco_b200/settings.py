"""Host mirror of `MjType::Settings` (`mj.set` in the reference binding).

Field names, defaults and ordering restate the X-macro tables of src/simsettings.h:23-183; the
binding exposes each of them read/write (src/bind.cpp:246-365). `pack()` flattens the object into
the `GmjSettings` arrays of include/gmj_types.h that are uploaded to the device.
"""
from __future__ import annotations

import copy
import ctypes as C

# ---- index tables: order must equal the enums in include/gmj_types.h
GENERAL_PACKED = [
    "mujoco_timestep", "curve_validation", "tip_force_applied", "random_seed",
    "time_for_action", "saturation_yield_factor",
    "sensor_sample_mode", "state_sample_mode", "sensor_n_prev_steps", "state_n_prev_steps",
    "sensor_noise_mag", "sensor_noise_mu", "sensor_noise_std",
    "state_noise_mag", "state_noise_mu", "state_noise_std",
    "base_position_noise", "palm_scale_factor",
    "oob_distance", "lift_height", "gripper_target_height",
    "stable_finger_force", "stable_palm_force", "stable_finger_force_lim", "stable_palm_force_lim",
    "XY_distance_threshold",
    "cap_reward", "quit_if_cap_exceeded", "reward_cap_lower_bound", "reward_cap_upper_bound",
    "continous_actions", "use_termination_action", "termination_threshold", "lift_on_termination",
    "sim_steps_per_action", "fingertip_min_mm",
    "sim_gauge_raw_to_N_factor",
    "base_z_min", "base_z_max", "base_x_min", "base_x_max", "base_y_min", "base_y_max",
    "base_yaw_min", "base_yaw_max", "reset_qpos_base",
    "reset_qpos_g0", "reset_qpos_g1", "reset_qpos_g2", "reset_qpos_g3", "reset_qpos_g4",
    "reset_qpos_g5", "reset_qpos_g6", "use_noise_seeded",
]
GS_COUNT = 56
assert len(GENERAL_PACKED) <= GS_COUNT

# src/simsettings.h:23-98 (name, default); float32-typed fields noted for later rounding
GENERAL_DEFAULTS = dict(
    debug=True, mujoco_timestep=0.001, curve_validation=False, tip_force_applied=0.0,
    random_seed=0, randomise_colours=False,
    auto_set_timestep=True, auto_calibrate_gauges=True, auto_sim_steps=True,
    auto_exceed_lateral_lim=False, time_for_action=0.2, saturation_yield_factor=1.0,
    exceed_lat_min_factor=0.75, exceed_lat_max_factor=1.5,
    use_HER=False, goal_reward=1.0, divide_goal_reward=True, reward_on_end_only=True,
    binary_goal_vector=False,
    sensor_sample_mode=6, state_sample_mode=4, sensor_n_prev_steps=3, state_n_prev_steps=3,
    sensor_noise_mag=0.0, sensor_noise_mu=0.05, sensor_noise_std=0.025,
    state_noise_mag=0.0, state_noise_mu=0.025, state_noise_std=0.0,
    base_position_noise=6e-3, palm_scale_factor=1.0,
    oob_distance=75e-3, lift_height=15e-3, gripper_target_height=28e-3,
    stable_finger_force=1.0, stable_palm_force=-1.0, stable_finger_force_lim=100.0,
    stable_palm_force_lim=100.0, XY_distance_threshold=10e-3,
    cap_reward=True, quit_if_cap_exceeded=False, reward_cap_lower_bound=-1.00,
    reward_cap_upper_bound=1.00,
    continous_actions=False, use_termination_action=True, termination_threshold=0.9,
    lift_on_termination=True, sim_steps_per_action=200, fingertip_min_mm=-12.5,
    render_on_step=False, use_render_delay=False, render_delay=0.5,
)
# not X-macro fields: derived scalars and base limits that live in globals in the reference
# (src/mjclass.cpp:274 ; src/myfunctions.cpp:226-241)
EXTRA_DEFAULTS = dict(
    sim_gauge_raw_to_N_factor=1.0,
    base_z_min=-30e-3, base_z_max=30e-3, base_x_min=-0.5, base_x_max=0.5,
    base_y_min=-0.5, base_y_max=0.5, base_yaw_min=-1.5707963267948966,
    base_yaw_max=1.5707963267948966, reset_qpos_base=0.0,
    reset_qpos_g0=0.13, reset_qpos_g1=0.0, reset_qpos_g2=0.13, reset_qpos_g3=0.0,
    reset_qpos_g4=0.13, reset_qpos_g5=0.0, reset_qpos_g6=0.00488, use_noise_seeded=1,
)

SENSOR_NAMES = ["motor_state_sensor", "base_state_sensor_Z", "base_state_sensor_XY",
                "base_state_sensor_yaw", "bending_gauge", "axial_gauge", "palm_sensor",
                "wrist_sensor_XY", "wrist_sensor_Z", "cartesian_contacts_XYZ"]
SENSOR_DEFAULTS = [(True, 0, -1), (True, 0, -1), (True, 0, -1), (True, 0, -1), (True, 20, 10),
                   (False, 3.0, 10), (True, 10.0, 10), (False, 5.0, 10), (True, 10.0, 10),
                   (False, 0, -1)]
ACTION_NAMES = ["gripper_X", "gripper_prismatic_X", "gripper_Y", "gripper_revolute_Y", "gripper_Z",
                "base_X", "base_Y", "base_Z", "base_roll", "base_pitch", "base_yaw"]
ACTION_DEFAULTS = [(False, 2.0e-3, -1), (True, 1.0e-3, -1), (False, 2.0e-3, -1), (True, 0.01, -1),
                   (True, 2.0e-3, 1), (True, 2.0e-3, 1), (True, 2.0e-3, 1), (True, 2.0e-3, 1),
                   (False, 0.01, 1), (False, 0.02, 1), (True, 0.05, 1)]
BINARY_NAMES = ["step_num", "lifted", "oob", "dropped", "lifted_to_height", "target_height",
                "exceed_limits", "object_contact", "object_stable", "stable_height",
                "stable_termination", "lifted_termination", "failed_termination",
                "successful_grasp", "within_XY_distance"]
BINARY_DEFAULTS = [(-0.01, 0, 1), (0.005, 0, 1), (-1.0, 1, 1), (0.0, 0, 1000), (0.01, 0, 1),
                   (0.01, 0, 1), (-0.1, 0, 1), (0.005, 0, 1), (0.01, 0, 1), (1.0, 0, 1000),
                   (1.0, 1, 1), (1.0, 1, 1), (-1.0, 1, 1), (0.01, 1, 1), (0.1, 0, 1)]
LINEAR_NAMES = ["finger_force", "palm_force", "exceed_axial", "exceed_lateral", "exceed_palm",
                "good_bend_sensor", "good_palm_sensor", "exceed_bend_sensor", "exceed_wrist_sensor",
                "exceed_palm_sensor", "dangerous_bend_sensor", "dangerous_wrist_sensor",
                "dangerous_palm_sensor", "action_penalty_lin", "action_penalty_sq",
                "finger1_force", "finger2_force", "finger3_force", "ground_force", "grasp_metric",
                "object_XY_distance"]
LINEAR_DEFAULTS = [(0.0, 0, 1, 1.0, 2.0, -1), (0.05, 0, 1, 1.0, 3.0, 6.0), (-0.05, 0, 1, 2.0, 6.0, -1),
                   (-0.05, 0, 1, 4.0, 6.0, -1), (-0.05, 0, 1, 6.0, 10.0, -1),
                   (0.0, 0, 1, 0.2, 1.0, -1), (0.0, 0, 1, 0.2, 1.0, -1), (0.0, 0, 1, 5.0, 10.0, -1),
                   (0.0, 0, 1, 5.0, 10.0, -1), (0.0, 0, 1, 10.0, 20.0, -1),
                   (0.0, 1, 1, 10.0, 11.0, -1), (0.0, 1, 1, 12.0, 13.0, -1), (0.0, 1, 1, 20.0, 21.0, -1),
                   (-0.1, 0, 1, 0.1, 1.5, -1), (-0.1, 0, 1, 0.1, 1.5, -1),
                   (0.0, 0, 1, 0.0, 2.0, 6.0), (0.0, 0, 1, 0.0, 2.0, 6.0), (0.0, 0, 1, 0.0, 2.0, 6.0),
                   (0.0, 0, 1, 0.0, 2.0, -1), (0.0, 0, 1, 0.0, 10.0, -1),
                   (0.05, 0, 1, -0.2, -5e-3, -1)]


class GmjSettings(C.Structure):
    _fields_ = [("general", C.c_double * GS_COUNT),
                ("sensor", (C.c_double * 10) * 10),
                ("action", (C.c_double * 4) * 11),
                ("binary", (C.c_double * 3) * 15),
                ("linear", (C.c_double * 6) * 21)]


class Sensor:
    """MjType::Sensor user-facing fields (src/mjclass.h:79-158)."""

    def __init__(self, in_use, normalise, read_rate):
        self.in_use, self.normalise, self.read_rate = bool(in_use), float(normalise), float(read_rate)
        self.use_normalisation, self.use_noise = True, True
        self.raw_value_offset = 0.0
        self.noise_mag, self.noise_mu, self.noise_std = 0.0, 0.0, -1.0
        self.noise_overriden = False

    def set(self, in_use, normalise, read_rate):
        self.in_use, self.normalise, self.read_rate = bool(in_use), float(normalise), float(read_rate)

    def set_gaussian_noise(self, mu, std):
        self.noise_mag, self.noise_mu, self.noise_std, self.noise_overriden = 0.0, mu, std, True

    def set_uniform_noise(self, mag):
        self.noise_mag, self.noise_mu, self.noise_std, self.noise_overriden = mag, 0.0, -1.0, True

    def row(self):
        return [self.in_use, self.normalise, self.read_rate, self.use_normalisation, self.use_noise,
                self.raw_value_offset, self.noise_mag, self.noise_mu, self.noise_std,
                self.noise_overriden]


class ActionSetting:
    """MjType::ActionSetting (src/mjclass.h:458-484)."""

    def __init__(self, in_use, value, sign):
        if sign not in (-1, 1):
            raise RuntimeError("ActionSetting given sign not equal to +1 or -1")
        self.in_use, self.continous, self.value, self.sign = bool(in_use), False, float(value), int(sign)

    def set(self, in_use, value, sign):
        self.in_use, self.value, self.sign = bool(in_use), float(value), int(sign)

    def row(self):
        return [self.in_use, self.continous, self.value, self.sign]


class BinaryReward:
    """MjType::BinaryReward (src/mjclass.h:649-663)."""

    def __init__(self, reward, done, trigger):
        self.reward, self.done, self.trigger = float(reward), int(done), int(trigger)

    def set(self, reward, done, trigger):
        self.reward, self.done, self.trigger = float(reward), int(done), int(trigger)

    def row(self):
        return [self.reward, self.done, self.trigger]


class LinearReward:
    """MjType::LinearReward (src/mjclass.h:665-684)."""

    def __init__(self, reward, done, trigger, min, max, overshoot):
        self.reward, self.done, self.trigger = float(reward), int(done), int(trigger)
        self.min, self.max, self.overshoot = float(min), float(max), float(overshoot)

    def set(self, reward, done, trigger, min, max, overshoot):
        # reference quirk Q5: `overshoot_ = overshoot_` self-assignment, overshoot is NOT changed
        # (src/mjclass.h:679-682)
        self.reward, self.done, self.trigger = float(reward), int(done), int(trigger)
        self.min, self.max = float(min), float(max)

    def row(self):
        return [self.reward, self.done, self.trigger, self.min, self.max, self.overshoot]


class Settings:
    """`mj.set`: every X-macro field as an attribute."""

    def __init__(self):
        for k, v in GENERAL_DEFAULTS.items():
            setattr(self, k, v)
        for k, v in EXTRA_DEFAULTS.items():
            setattr(self, k, v)
        for n, d in zip(SENSOR_NAMES, SENSOR_DEFAULTS):
            setattr(self, n, Sensor(*d))
        for n, d in zip(ACTION_NAMES, ACTION_DEFAULTS):
            setattr(self, n, ActionSetting(*d))
        for n, d in zip(BINARY_NAMES, BINARY_DEFAULTS):
            setattr(self, n, BinaryReward(*d))
        for n, d in zip(LINEAR_NAMES, LINEAR_DEFAULTS):
            setattr(self, n, LinearReward(*d))

    # ---- helpers of MjType::Settings used by the callers (src/mjclass.cpp:5182-5292)
    def wipe_rewards(self):
        never = 10000
        for n in BINARY_NAMES:
            r = getattr(self, n)
            r.reward, r.done, r.trigger = 0.0, 0, never
        for n in LINEAR_NAMES:
            r = getattr(self, n)
            r.reward, r.done, r.trigger, r.min, r.max, r.overshoot = 0.0, 0, never, 0.0, 0.0, -1.0

    def disable_sensors(self):
        for n in SENSOR_NAMES:
            getattr(self, n).in_use = False

    def set_use_normalisation(self, v):
        for n in SENSOR_NAMES:
            getattr(self, n).use_normalisation = bool(v)

    def set_use_noise(self, v):
        for n in SENSOR_NAMES:
            getattr(self, n).use_noise = bool(v)

    def set_all_action_continous(self, v):
        for n in ACTION_NAMES:
            getattr(self, n).continous = bool(v)

    def set_all_action_use(self, v):                    # src/mjclass.cpp:5366-5386
        for n in ACTION_NAMES:
            getattr(self, n).in_use = bool(v)

    def set_all_action_value(self, v):                  # src/mjclass.cpp:5388-5397
        for n in ACTION_NAMES:
            getattr(self, n).value = float(v)

    def set_all_action_sign(self, v):                   # src/mjclass.cpp:5399-5416
        if v not in (-1, 1):
            raise RuntimeError("MjType::Settings::set_all_action_sign recieved a value not either +1 or -1")
        for n in ACTION_NAMES:
            getattr(self, n).sign = int(v)

    def get_settings(self) -> str:
        """MjType::Settings::get_settings (src/mjclass.cpp:4967-5060): every field as text, one per line."""
        rows = [f"{k} = {getattr(self, k)}" for k in list(GENERAL_DEFAULTS) + list(EXTRA_DEFAULTS)]
        for n in SENSOR_NAMES:
            rows.append(f"{n} = {getattr(self, n).row()}")
        for n in ACTION_NAMES:
            rows.append(f"{n} = {getattr(self, n).row()}")
        for n in BINARY_NAMES + LINEAR_NAMES:
            rows.append(f"{n} = {getattr(self, n).row()}")
        return "\n".join(rows)

    def scale_rewards(self, scale):
        for n in BINARY_NAMES + LINEAR_NAMES:
            getattr(self, n).reward *= scale

    def copy(self):
        return copy.deepcopy(self)

    def n_actions(self):
        """Action table size (src/mjclass.cpp:108-141)."""
        n = 0
        for a in ACTION_NAMES:
            s = getattr(self, a)
            if s.in_use:
                n += 1 if self.continous_actions else 2
        return n + (1 if self.use_termination_action else 0)

    def pack(self) -> GmjSettings:
        self.set_all_action_continous(self.continous_actions)
        s = GmjSettings()
        for i, n in enumerate(GENERAL_PACKED):
            s.general[i] = float(getattr(self, n))
        for i, n in enumerate(SENSOR_NAMES):
            for j, v in enumerate(getattr(self, n).row()):
                s.sensor[i][j] = float(v)
        for i, n in enumerate(ACTION_NAMES):
            for j, v in enumerate(getattr(self, n).row()):
                s.action[i][j] = float(v)
        for i, n in enumerate(BINARY_NAMES):
            for j, v in enumerate(getattr(self, n).row()):
                s.binary[i][j] = float(v)
        for i, n in enumerate(LINEAR_NAMES):
            for j, v in enumerate(getattr(self, n).row()):
                s.linear[i][j] = float(v)
        return s
