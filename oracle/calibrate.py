"""Auto-calibration driven over the CPU oracle. TEST / MEASUREMENT INFRASTRUCTURE ONLY (tests/,
bench.py --impl reference and cpu_baseline): same control flow as the product
(gripper_mujoco_b200/calibration.py), with one oracle env as the backend."""
from __future__ import annotations

import numpy as np

from gripper_mujoco_b200 import calibration
from gripper_mujoco_b200.model import GripperParams, build_model
from gripper_mujoco_b200.settings import Settings
from oracle.binding import OracleEnv

SCRATCH_ENV_ID = 0xC0FFEE


def make_backend(model):
    ms = model.to_struct()
    B = model.B

    def backend(st: Settings, n_substeps: int, sensors: bool) -> dict:
        env = OracleEnv(ms, st.pack())
        env.reset(int(st.random_seed), SCRATCH_ENV_ID, 1)
        if sensors:
            env.substep(n_substeps)
        else:
            env.settle(n_substeps, float(st.mujoco_timestep))
        p = env.peek()
        return dict(qpos=np.array(p.qpos[:B + 7]), unstable=bool(p.unstable), si_wrist=float(p.si[4]),
                    gauge0=env.gauge(0))

    return backend


def calibrated(params: GripperParams, settings: Settings | None = None, enable: bool = True, verbose: bool = False):
    """(model, settings, info) with reset pose, timestep, sim steps and sensor calibration filled in."""
    mdl = build_model(params)
    s = settings.copy() if settings is not None else Settings()
    be = make_backend(mdl)
    calibration.calibrate_reset(params, mdl.N, mdl.B, s, be)
    info = calibration.configure_settings(params, s, calibration.AutoFlags(), be, enable=enable, verbose=verbose)
    return mdl, s, info
