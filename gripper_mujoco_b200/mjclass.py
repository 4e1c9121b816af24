"""Batched MjClass: the reference's pybind surface (src/bind.cpp:43-170) with a leading env axis.

Every simulation call goes through the C ABI of libgripper_b200.so (include/gmj.h); this module
only holds host-side orchestration that the reference also does on the host:
  * MjClass::configure_settings auto-calibration  (src/mjclass.cpp:225-314)
  * MjClass::find_highest_stable_timestep          (src/mjclass.cpp:4745-4850)
  * MjClass::calibrate_simulated_sensors           (src/mjclass.cpp:4643-4676)
  * luke::calibrate_reset                          (src/myfunctions.cpp:1450-1516)
"""
from __future__ import annotations

import numpy as np
import torch

from . import calibration
from .batch import GripperBatch
from .lib import GP
from .model import GripperParams, Model, build_model
from .objects import ObjectSpec, build_object_set
from .settings import Settings

class SpawnParams:
    """MjClass::default_spawn_params (src/mjclass.h:1560-1580)."""

    def __init__(self):
        self.index = 0
        self.x = self.y = self.zrot = 0.0
        self.xrange = self.yrange = 0.0
        self.rotrange = 0.0
        self.xmin, self.xmax, self.ymin, self.ymax = -1.0, 1.0, -1.0, 1.0
        self.smallest_gap = 1e-3


class MjClass:
    def __init__(self, n_envs: int = 1, params: GripperParams | None = None,
                 objects: list[ObjectSpec] | None = None, settings: Settings | None = None,
                 device: int = 0, precision: str = "fp32", env_id_offset: int = 0,
                 calibrate: bool = True, verbose: bool = False):
        self.params = params or GripperParams()
        self.model: Model = build_model(self.params)
        self.set: Settings = settings or Settings()
        self.n_envs = int(n_envs)
        self.device_index, self.precision, self.env_id_offset = device, precision, env_id_offset
        self.verbose = verbose
        self.default_spawn_params = SpawnParams()
        self.object_set_name = "set9_fullset"
        self._objects = objects if objects is not None else build_object_set()
        self._auto = calibration.AutoFlags()
        self._do_calibrate = calibrate
        calibration.calibrate_reset(self.params, self.model.N, self.model.B, self.set, self._backend)
        self.batch = GripperBatch(self.n_envs, self.model, self.set, device, precision, env_id_offset)
        self.batch.set_objects(self._objects)
        self.device = self.batch.device
        self._actions = torch.zeros(self.n_envs, dtype=torch.int32, device=self.device)
        self.reset()

    # ------------------------------------------------------------------ calibration (host orchestration)
    def _backend(self, st: Settings, n_substeps: int, sensors: bool) -> dict:
        """One freshly reset scratch env on this GPU, stepped n times through gmj_substeps."""
        b = GripperBatch(1, self.model, st, self.device_index, self.precision, 0xC0FFEE)
        b.substeps(n_substeps, sensors=sensors)
        q, _ = b.state()
        p = b.probe()[0]
        out = dict(qpos=q[0], unstable=bool(p[GP["unstable"]] > 0 or b.stats()["unstable_resets"] > 0),
                   si_wrist=float(p[GP["si_wrist"]]), gauge0=float(p[GP["gauge0"]]))
        b.close()
        return out

    def yield_load(self, thickness: float | None = None, width: float | None = None) -> float:
        return calibration.yield_load(self.params, thickness, width)

    def find_highest_stable_timestep(self) -> float:
        dt, self.highest_stable_timestep = calibration.find_highest_stable_timestep(self.set, self._backend, self.verbose)
        self.set.mujoco_timestep = dt
        return dt

    def configure_settings(self):
        info = calibration.configure_settings(self.params, self.set, self._auto, self._backend,
                                              enable=self._do_calibrate, verbose=self.verbose)
        if "highest_stable_timestep" in info:
            self.highest_stable_timestep = info["highest_stable_timestep"]

    # ------------------------------------------------------------------ MjClass API (batched)
    def reset(self, mask=None):
        self.configure_settings()
        self.batch.set = self.set
        self.batch.push_settings()
        self.batch.reset(mask)

    hard_reset = reset

    def step(self, n: int = 1):
        self.batch.substeps(n, sensors=True)

    def set_action(self, action, continous_fraction=None):
        self.batch.set_action(action, continous_fraction)

    set_discrete_action = set_action

    def set_continous_action(self, action, fraction):
        """src/mjclass.cpp:1517-1526: the magnitude fraction is clamped to [-1, 1] before set_action."""
        f = torch.as_tensor(fraction, device=self.device, dtype=torch.float32).clamp(-1.0, 1.0)
        self.batch.set_action(action, f)

    def action_step(self):
        self.batch.action_step()

    def get_observation(self) -> torch.Tensor:
        return self.batch.observe(obs=True, reward=False, done=False, events=False)[0]

    def get_observation_numpy(self) -> np.ndarray:
        return self.get_observation().cpu().numpy()

    def reward(self) -> torch.Tensor:
        return self.batch.observe(obs=False, reward=True, done=False, events=False)[1]

    def is_done(self) -> torch.Tensor:
        return self.batch.observe(obs=False, reward=False, done=True, events=False)[2].bool()

    def get_event_state(self) -> torch.Tensor:
        return self.batch.observe(obs=False, reward=False, done=False, events=True)[3]

    def fused_step(self, actions, fractions=None):
        """set_action (set_continous_action when `fractions` is given: clamped to [-1, 1]) + action_step +
        get_observation + reward + is_done in one launch."""
        if fractions is not None:
            fractions = torch.as_tensor(fractions, device=self.device, dtype=torch.float32).clamp(-1.0, 1.0)
        return self.batch.step(actions, fractions)

    def get_n_actions(self) -> int:
        return self.batch.n_actions

    def get_n_obs(self) -> int:
        return self.batch.n_obs

    def get_N(self) -> int:
        return self.model.N

    def spawn_object(self, index, x=0.0, y=0.0, zrot=0.0, mask=None):
        n = len(self._objects)
        if np.any(np.asarray(index) < 0) or np.any(np.asarray(index) >= n):
            raise RuntimeError("bad index to spawn_object()")           # src/mjclass.cpp:2359-2362
        self.batch.spawn_object(index, x, y, zrot, mask)

    def spawn_into_scene(self, xrange=None, rotrange=None, mask=None, index=None):
        """spawn_into_scene() / spawn_into_scene(index) (src/mjclass.cpp:2422-2654). Returns a bool tensor [n_envs]:
        True where an object was placed, False where no candidate pose was clear of the fingertips (the reference's
        `return false`; MjEnv retries, rl/env/MjEnv.py:1216-1244). Entries outside `mask` keep their last value."""
        d = self.default_spawn_params
        if index is not None and not torch.is_tensor(index):
            n = len(self._objects)
            if np.any(np.asarray(index) < 0) or np.any(np.asarray(index) >= n):
                raise RuntimeError("bad index to spawn_into_scene()")
        ok = self.batch.spawn_into_scene(d.xrange if xrange is None else xrange,
                                         d.rotrange if rotrange is None else rotrange, mask, index)
        return ok.bool()

    def get_number_of_objects(self) -> int:
        return len(self._objects)

    def get_object_name(self, index: int) -> str:
        return self._objects[index].name

    def get_current_object_name(self):
        idx = self.batch.probe_field("obj_index").astype(int)
        return [self._objects[i].name for i in idx]

    def set_base_XYZ_limits(self, x, y, z):
        s = self.set
        s.base_x_min, s.base_x_max, s.base_y_min, s.base_y_max, s.base_z_min, s.base_z_max = -x, x, -y, y, -z, z

    def set_base_yaw_limit(self, yaw):
        self.set.base_yaw_min, self.set.base_yaw_max = -yaw, yaw

    def get_finger_thickness(self): return self.params.finger_thickness
    def get_finger_width(self): return self.params.finger_width
    def get_finger_length(self): return self.params.finger_length
    def get_finger_modulus(self): return self.params.finger_E
    def get_finger_hook_length(self): return self.params.hook_length
    def get_finger_hook_angle_degrees(self): return self.params.hook_angle_degrees
    def get_fingertip_clearance(self): return self.params.fingertip_clearance
    def is_finger_hook_fixed(self): return self.params.fixed_hook_segment

    def get_finger_forces(self) -> np.ndarray:
        p = self.batch.probe()
        return p[:, GP["si_gauge0"]:GP["si_gauge0"] + 3]

    def get_palm_force(self) -> np.ndarray:
        return self.batch.probe_field("si_palm")

    def get_wrist_force(self) -> np.ndarray:
        return self.batch.probe_field("si_wrist")

    # ------------------------------------------------------------------ pickle (src/bind.cpp:207-241)
    def __getstate__(self):
        """Like the reference: settings, model key and spawn params are saved, physics state is not;
        the calibration results travel inside the settings and auto-calibration is not repeated."""
        return dict(n_envs=self.n_envs, params=self.params, settings=self.set, device=self.device_index,
                    precision=self.precision, env_id_offset=self.env_id_offset,
                    object_set_name=self.object_set_name, spawn=self.default_spawn_params.__dict__)

    def __setstate__(self, st):
        s = st["settings"]
        self.__init__(st["n_envs"], st["params"], None, s, st["device"], st["precision"], st["env_id_offset"],
                      calibrate=False)
        self.object_set_name = st["object_set_name"]
        self.default_spawn_params.__dict__.update(st["spawn"])

    def close(self):
        self.batch.close()
