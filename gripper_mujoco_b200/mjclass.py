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
from .events import EventTrack, TestReport, add_events as _add_events, event_tracks
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


class Goal:
    """Placeholder of MjType::Goal (src/mjclass.h:687-770): the HER goal machinery is outside this path (SURVEY 8:
    use_HER = false in every baseline config); reset / get_goal_info exist because MjEnv calls them unconditionally."""

    def reset(self, wipe: bool = False):
        pass

    def get_goal_info(self) -> str:
        return "goal: not in use (use_HER = false)"


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
        # exposed variables of the binding (src/bind.cpp:196-201). There is no MJCF on this path: the model is built from
        # GripperParams, so model_folder_path only takes part in the strings load_relative() assembles.
        self.model_folder_path = ""
        self.current_load_path = ""
        self.machine = "b200"                         # src/mjclass.h:1606 (LUKE_MACHINE of the build)
        self.goal = Goal()
        self.objects_per_task = 20                    # objects in one task XML of the reference's object sets
        self._all_objects = None
        self._finger_EI_changed = False               # resetFlags.finger_EI_changed (src/mjclass.cpp:102-106)
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
    def _rebuild(self):
        """hard_reset after a model change (src/mjclass.cpp:102-106, 377-409): new model constants (finger stiffness
        follows thickness / width / modulus: quirk Q10, src/myfunctions.cpp:988-1048), new device state, and every
        auto-calibration (reset pose, timestep, sim steps, gauges, wrist offset) runs again."""
        self.model = build_model(self.params)
        self.batch.close()
        self._auto = calibration.AutoFlags()
        calibration.calibrate_reset(self.params, self.model.N, self.model.B, self.set, self._backend)
        self.batch = GripperBatch(self.n_envs, self.model, self.set, self.device_index, self.precision, self.env_id_offset)
        self.batch.set_objects(self._objects)
        self._finger_EI_changed = False

    def reset(self, mask=None):
        if self._finger_EI_changed:
            self._rebuild()
        self.configure_settings()
        self.batch.set = self.set
        self.batch.push_settings()
        self.batch.reset(mask)

    def hard_reset(self):
        """MjClass::hard_reset (src/mjclass.cpp:453-470): forget every calibration result, then reset."""
        self._auto = calibration.AutoFlags()
        self.reset()

    def reset_timestep(self):
        """MjEnv.reset(timestep=True) (rl/env/MjEnv.py:2243): recalibrate the timestep, then reset. (The name is called by
        MjEnv but missing from this reference commit's bind.cpp; it is what hard_reset does for the timestep flags.)"""
        self._auto.timestep = bool(self.set.auto_set_timestep) and self._do_calibrate
        self._auto.simsteps = bool(self.set.auto_sim_steps)
        self.reset()

    def load(self, model_path: str):
        """MjClass::load (src/mjclass.cpp:377-409). The reference parses a task MJCF; here the model comes from
        GripperParams, so the path is read for what a task file name carries: `gripper_N<segments>_<width mm>` selects
        the finger and `task_<k>` the k-th group of `objects_per_task` objects of the object set. Then hard reset."""
        import re
        m = re.search(r"gripper_N(\d+)(?:_(\d+))?", model_path)
        if m:
            self.params.num_segments = int(m.group(1))
            if m.group(2):
                self.params.finger_width = int(m.group(2)) * 1e-3
        if self._all_objects is None:
            self._all_objects = self._objects
        m = re.search(r"task_(\d+)", model_path)
        if m and self.objects_per_task > 0:
            k, per = int(m.group(1)), self.objects_per_task
            n_groups = max(1, len(self._all_objects) // per)
            self._objects = self._all_objects[(k % n_groups) * per:(k % n_groups + 1) * per]
        else:
            self._objects = self._all_objects
        self.current_load_path = model_path
        self._rebuild()
        self.hard_reset()

    def load_relative(self, relative_path: str):
        """MjClass::load_relative (src/mjclass.cpp:411-433)."""
        if self.object_set_name == "":
            raise RuntimeError("Cannot use MjClass::load_relative() as LUKE_MJCF_PATH or LUKE_DEFAULTOBJECTS not set")
        folder = self.model_folder_path
        if folder and not folder.endswith("/"):
            folder += "/"
        if not relative_path.startswith("/"):
            relative_path = "/" + relative_path
        self.load(folder + self.object_set_name + relative_path)

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

    def spawn_into_scene(self, index=None, xpos=None, ypos=None, zrot=None, xrange=None, yrange=None, rotrange=None,
                         mask=None):
        """The overloads of src/mjclass.cpp:2422-2473: (index), (index, x, y), (index, x, y, zrot) and
        (index, x, y, zrot, xrange, yrange, rotrange); what is not given comes from default_spawn_params. Returns a bool
        tensor [n_envs] (a bool for one env): True where the object was placed, False where no candidate pose was clear of
        the fingertips (the reference's `return false`; MjEnv retries, rl/env/MjEnv.py:1216-1244). Entries outside `mask`
        keep their last value. `index=None` draws the object per env from the device RNG. The device search is centred on
        the origin with one range for x and y (what MjEnv's single-object path asks for); other centres are refused."""
        d = self.default_spawn_params
        x = d.x if xpos is None else xpos
        y = d.y if ypos is None else ypos
        r = d.zrot if zrot is None else zrot
        xr = d.xrange if xrange is None else xrange
        yr = d.yrange if yrange is None else yrange
        rr = d.rotrange if rotrange is None else rotrange
        if x != 0 or y != 0 or r != 0 or abs(xr - yr) > 1e-12:
            raise NotImplementedError("spawn_into_scene: only a search centred on the origin with xrange == yrange is built "
                                      "(scene spawning with an origin offset is outside this path, SURVEY 8)")
        if index is not None and not torch.is_tensor(index):
            n = len(self._objects)
            if np.any(np.asarray(index) < 0) or np.any(np.asarray(index) >= n):
                raise RuntimeError("bad index to spawn_into_scene()")
        ok = self.batch.spawn_into_scene(xr, rr, mask, index).bool()
        return bool(ok[0]) if self.n_envs == 1 else ok

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

    # finger geometry setters: take effect at the next reset(), like the reference (src/mjclass.cpp:3811-3845)
    def set_finger_thickness(self, thickness: float):
        if abs(thickness - self.params.finger_thickness) > 1e-12:
            self.params.finger_thickness = float(thickness); self._finger_EI_changed = True

    def set_finger_width(self, width: float):
        if abs(width - self.params.finger_width) > 1e-12:
            self.params.finger_width = float(width); self._finger_EI_changed = True

    def set_finger_modulus(self, E: float):
        if abs(E - self.params.finger_E) > 1e-3:
            self.params.finger_E = float(E); self._finger_EI_changed = True

    def get_finger_rigidity(self):
        """E I of the finger cross-section (src/myfunctions.cpp:1050-1058)."""
        p = self.params
        return p.finger_E * p.finger_width * p.finger_thickness ** 3 / 12.0

    def get_finger_stiffnesses(self):
        """The N joint stiffnesses of one finger, base to tip (src/myfunctions.cpp:988-1048)."""
        from .model import finger_joint_stiffness
        return finger_joint_stiffness(self.params)[0]

    def using_xyz_base_actions(self) -> bool:
        return bool(self.params.xy_base_joint)         # luke::use_base_xyz (src/mjclass.cpp:3955-3960)

    def set_sensor_noise_and_normalisation_to(self, set_as: bool):
        """src/mjclass.cpp:4856-4862; pushed to the device at the next reset() like any settings change."""
        self.set.set_use_normalisation(bool(set_as))
        self.set.set_use_noise(bool(set_as))

    def reset_goal(self):
        self.goal.reset(True)                          # src/mjclass.cpp:4717-4722

    def _one(self, x):
        """The reference is one env; with n_envs == 1 the getters below return what it returns, else a leading env axis."""
        return x[0] if self.n_envs == 1 else x

    # ------------------------------------------------------------------ events / test report
    def get_event_tracks(self, env_ids=None):
        row, abs_, last, _ = self.batch.event_track()
        return event_tracks(row, abs_, last, env_ids)

    def get_test_report(self):
        """MjClass::get_test_report (src/mjclass.cpp:3800-3809): object name, cumulative reward and env_.cnt."""
        row, abs_, last, cr = self.batch.event_track()
        names = self.get_current_object_name()
        tracks = event_tracks(row, abs_, last)
        return self._one([TestReport(names[i], float(cr[i]), tracks[i]) for i in range(self.n_envs)])

    def add_events(self, e1: EventTrack, e2: EventTrack) -> EventTrack:
        return _add_events(e1, e2)                     # src/mjclass.cpp:4692-4715

    # ------------------------------------------------------------------ state getters
    def get_state_metres(self, realworld: bool = False):
        """MjClass::get_state_metres (src/mjclass.cpp:3132-3182): the latest state-sensor readings mapped back from
        [-1, 1] to SI: [gripper x, y, z, base z], or [gripper x, y, z, base x, y, z] when the base XY sensor is in use."""
        if realworld:
            raise NotImplementedError("real-gripper input is outside this path (SURVEY 8: out of scope)")
        p, s, g = self.batch.probe(), self.set, self.params

        def un(v, lo, hi):                                             # unnormalise_from (src/mjclass.h:132-139)
            return lo + 0.5 * (v + 1.0) * (hi - lo)
        cols = [un(p[:, GP["read_motor_x"]], g.xy_min, g.xy_max), un(p[:, GP["read_motor_y"]], g.xy_min, g.xy_max),
                un(p[:, GP["read_motor_z"]], g.z_min, g.z_max)]
        if s.base_state_sensor_XY.in_use:
            cols += [un(p[:, GP["read_base_x"]], s.base_x_min, s.base_x_max), un(p[:, GP["read_base_y"]], s.base_y_min, s.base_y_max)]
        cols.append(un(p[:, GP["read_base_z"]], s.base_z_min, s.base_z_max))
        return self._one(np.stack(cols, axis=1).astype(np.float32))

    def self_overlap(self, depth: float = 0.0):
        """Not a reference call. Per env, the number of gripper geom pairs (finger <-> finger, hook <-> hook, segment <->
        palm) that overlap by more than `depth` metres in the current pose: contacts the reference simulates and this
        path does not (selfcheck.py). Synchronising host read, not for the step loop."""
        from . import selfcheck
        q, _ = self.batch.state()
        return selfcheck.self_overlap(self.model, q, depth)[0]

    def get_finger_angle(self):
        return self._one(self.batch.probe_field("tend_th").astype(np.float32))     # target_.end.get_th_rad()

    def get_object_XY_relative_to_gripper(self):
        """luke::get_object_XY_relative_to_gripper (src/myfunctions.cpp:3875-3896): base target xy minus object xy, one
        entry per live object (one live object per env on this path)."""
        q, _ = self.batch.state()
        p = self.batch.probe()
        B = self.model.B
        rel = np.stack([p[:, GP["base_x"]] - q[:, B], p[:, GP["base_y"]] - q[:, B + 1]], axis=1)
        return self._one([[list(r)] for r in rel]) if self.n_envs > 1 else [list(rel[0])]

    def get_object_xyz_bounding_box(self):
        idx = self.batch.probe_field("obj_index").astype(int)
        return self._one([[float(v) for v in self._objects[i].obj.bbox] for i in idx])

    def get_object_bounding_boxes(self):
        """luke::get_live_object_bounding_boxes (src/myfunctions.cpp:3801-3816): one box per live object."""
        b = self.get_object_xyz_bounding_box()
        return [b] if self.n_envs == 1 else [[x] for x in b]

    def get_live_object_names(self):
        names = self.get_current_object_name()
        return [names[0]] if self.n_envs == 1 else [[nm] for nm in names]

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
