"""Host-side auto-calibration of MjClass::configure_settings (src/mjclass.cpp:225-314), written over
a backend callable so the same control flow drives the CUDA path (product) and -- from tests and
bench.py only -- the CPU oracle.

backend(settings, n_substeps, sensors) -> dict(qpos=[nq], unstable=bool, si_wrist=float, gauge0=float)
runs `n_substeps` MjClass::step() calls on ONE freshly reset environment with `settings`.
"""
from __future__ import annotations

import math

import numpy as np

from .model import GripperParams
from .settings import Settings

YIELD_STRESS = 215e6          # src/myfunctions.cpp:216


def yield_load(p: GripperParams, thickness: float | None = None, width: float | None = None) -> float:
    """luke::calc_yield_point_load (src/myfunctions.cpp:3587-3606), float arithmetic."""
    t = p.finger_thickness if thickness is None else thickness
    w = p.finger_width if width is None else width
    f32 = np.float32
    inertia = f32((w * t ** 3) / 12.0)
    m_max = f32((YIELD_STRESS * inertia) / (0.5 * t))
    return float(f32(m_max / p.finger_length))


def calibrate_reset(p: GripperParams, N: int, B: int, s: Settings, backend) -> None:
    """luke::calibrate_reset (src/myfunctions.cpp:1450-1516): settle 400 steps of 2 ms from the
    keyframe pose with every weld on, store base + motor joint positions into s.reset_qpos_*."""
    st = s.copy()
    st.mujoco_timestep, st.base_position_noise, st.curve_validation = 2e-3, 0.0, False
    st.reset_qpos_base = 0.0
    for f in range(3):
        setattr(st, f"reset_qpos_g{2 * f}", p.xy_home)
        setattr(st, f"reset_qpos_g{2 * f + 1}", 0.0)
    st.reset_qpos_g6 = p.z_home
    # The reference settles at the MJCF default timestep (2 ms) whatever the model. Finger chains of more than
    # ~10 segments are unstable at 2 ms (rl/juypter/thesis_plots/mujoco_timesteps.csv); MuJoCo then hits
    # BADQACC and auto-resets mid-settle, which leaves an arbitrary pose. Here the same 0.8 s settle is
    # repeated at half the timestep until it is stable, so long fingers start from a true equilibrium.
    steps = 400
    for _ in range(5):
        r = backend(st, steps, False)
        q = r["qpos"]
        if not r["unstable"] and np.all(np.isfinite(q)) and np.abs(q[:B]).max() < 10.0:
            break
        st.mujoco_timestep *= 0.5
        steps *= 2
    else:
        raise RuntimeError("calibrate_reset: no stable settle found")
    s.reset_qpos_base = float(q[0])
    for f in range(3):
        setattr(s, f"reset_qpos_g{2 * f}", float(q[1 + f * (N + 2)]))
        setattr(s, f"reset_qpos_g{2 * f + 1}", float(q[2 + f * (N + 2)]))
    s.reset_qpos_g6 = float(q[B - 1])


def find_highest_stable_timestep(s: Settings, backend, verbose: bool = False):
    """MjClass::find_highest_stable_timestep (src/mjclass.cpp:4745-4850): coarse 0.5 ms search up from
    1 ms, 50 us comb down, hand-tuned safety factor, floor to whole microseconds.
    Returns (final_timestep, highest_stable_timestep)."""
    f32 = np.float32
    coarse, fine, test_time, max_dt = f32(0.5e-3), f32(50e-6), f32(1.0), f32(20e-3)
    nxt, coarse_pass = f32(1.0e-3), True
    st = s.copy()
    while True:
        st.mujoco_timestep = float(nxt)
        ok = not backend(st, int(test_time / nxt) + 1, True)["unstable"]
        if verbose:
            print(f"timestep {float(nxt) * 1e3:.3f} ms {'stable' if ok else 'unstable'}")
        if not ok:
            coarse_pass = False
            nxt = f32(nxt - fine)
        elif coarse_pass:
            nxt = f32(nxt + coarse)
        else:
            break
        if nxt < fine:
            raise RuntimeError("no stable timestep found for the simulation")
        if nxt > max_dt:
            nxt, coarse_pass = max_dt, False
    factor = f32(0.8) if nxt <= f32(3.0e-3) else f32(0.75) if nxt < f32(5.0e-3) else f32(0.65)
    final = f32(nxt * factor)
    final = f32(int(final * f32(1e6)) * 1e-6)
    return float(final), float(nxt)


def calibrate_simulated_sensors(s: Settings, bend_gauge_normalise: float, backend) -> None:
    """MjClass::calibrate_simulated_sensors (src/mjclass.cpp:4643-4676): wrist-Z offset after 0.3 s of
    settling; bending-gauge reading under the saturation tip load after 50 s
    (validate_curve_under_force, src/mjclass.cpp:4023-4104, retrying at 0.8x timestep when unstable)."""
    f32 = np.float32
    st = s.copy()
    st.wrist_sensor_Z.use_normalisation = st.wrist_sensor_Z.use_noise = False
    st.wrist_sensor_Z.raw_value_offset = 0.0
    r = backend(st, int(f32(0.3) / f32(st.mujoco_timestep)), True)
    s.wrist_sensor_Z.raw_value_offset = float(r["si_wrist"])
    st = s.copy()
    st.curve_validation, st.tip_force_applied = True, float(bend_gauge_normalise)
    for _ in range(5):
        r = backend(st, int(f32(50) / f32(st.mujoco_timestep)), True)
        if not r["unstable"]:
            break
        st.mujoco_timestep *= 0.8
        s.mujoco_timestep = st.mujoco_timestep
    else:
        raise RuntimeError("curve validation unstable")
    s.bending_gauge.normalise = float(r["gauge0"])
    s.curve_validation = False


class AutoFlags:
    """MjClass::resetFlags (src/mjclass.h:1667-1674)."""

    def __init__(self):
        self.init = False
        self.timestep = self.calibrate = self.simsteps = self.lateral = False
        self.time_for_action = None


def configure_settings(p: GripperParams, s: Settings, flags: AutoFlags, backend, enable: bool = True,
                       verbose: bool = False) -> dict:
    info = {}
    if not flags.init:
        flags.init = True
        flags.timestep, flags.calibrate = s.auto_set_timestep and enable, s.auto_calibrate_gauges and enable
        flags.simsteps, flags.lateral = s.auto_sim_steps, s.auto_exceed_lateral_lim and enable
    if flags.timestep:
        flags.timestep = False
        s.mujoco_timestep, info["highest_stable_timestep"] = find_highest_stable_timestep(s, backend, verbose)
    if flags.lateral:
        flags.lateral = False
        s.exceed_lateral.min = s.exceed_lat_min_factor * yield_load(p)
        s.exceed_lateral.max = s.exceed_lat_max_factor * yield_load(p)
    if flags.calibrate:
        flags.calibrate = False
        norm = float(np.float32(s.saturation_yield_factor * yield_load(p)))
        calibrate_simulated_sensors(s, norm, backend)
        s.sim_gauge_raw_to_N_factor = norm / s.bending_gauge.normalise
    if s.auto_sim_steps and flags.time_for_action is not None and abs(flags.time_for_action - s.time_for_action) > 1e-4:
        flags.simsteps = True
    flags.time_for_action = s.time_for_action
    if flags.simsteps:
        flags.simsteps = False
        s.sim_steps_per_action = int(math.ceil(s.time_for_action / s.mujoco_timestep))
    return info


def calibrate_native(mdl, s: Settings, device: int = 0, precision: str = "fp32", enable: bool = True):
    """The same calibration through the C ABI's gmj_calibrate (for hosts without this module; here a cross-check):
    runs calibrate_reset + the parts the auto_* settings ask for, writes the results into `s`, returns the GmjCalib."""
    import ctypes as C

    from . import lib as gl
    L = gl.lib()
    ms, ss = mdl.to_struct(), s.pack()
    io = gl.GmjCalib()
    io.set_timestep = int(bool(s.auto_set_timestep) and enable)
    io.calibrate_gauges = int(bool(s.auto_calibrate_gauges) and enable)
    io.set_sim_steps = int(bool(s.auto_sim_steps))
    gl.check(L.gmj_calibrate(C.byref(ms), C.byref(ss), int(device), 0 if precision == "fp32" else 1, C.byref(io)))
    s.reset_qpos_base = io.reset_qpos[0]
    for k in range(7):
        setattr(s, f"reset_qpos_g{k}", io.reset_qpos[1 + k])
    s.mujoco_timestep = io.timestep
    s.sim_steps_per_action = int(io.sim_steps_per_action)
    s.wrist_sensor_Z.raw_value_offset = io.wrist_raw_value_offset
    s.bending_gauge.normalise = io.bend_gauge_normalise
    s.sim_gauge_raw_to_N_factor = io.sim_gauge_raw_to_N_factor
    return io
