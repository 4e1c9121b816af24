"""`BatchedMjEnv` -- the policy-facing layer of the reference (`MjEnv`, rl/env/MjEnv.py) over n environments of one
batched `MjClass`. Same public surface and protocol, every per-env scalar of the reference became a
device tensor with a leading env axis:

* `reset()` = track reset -> (task reload) -> `mj.reset()` -> `_spawn_object()` -> observation  (rl/env/MjEnv.py:2222-2263)
* `step(action)` = `_take_action` (set_action + action_step) -> observation -> `_is_done` -> reward   (:2170-2220),
  fused into one launch; truncation at `max_episode_steps` is checked before the C++ `is_done` (:616-637)
* training objects are drawn from the task files after the test files, test mode walks `test_objects` objects x
  `test_trials_per_object` trials in order with the test-time generator seeded 13337419 (:1177-1326, 2051-2065)

What a batch changes: environments finish at different times, so `reset(mask)` takes an env mask and `step` can
auto-reset finished envs (`auto_reset=True`, the observation returned for those envs is the first of the new
episode, `info["final_observation"]` holds the last of the old one). All task files of the object set are resident
on the device at once, so the reference's xml reload (`task_reload_chance`) reduces to which file the next object
is drawn from. Rendering / RGBD options are graphics-API paths: unavailable on this box.
"""
from __future__ import annotations

from dataclasses import asdict, dataclass, field

import numpy as np
import torch

from .mjclass import MjClass
from .model import GripperParams
from .objects import build_object_set, task_split
from .settings import Settings

TEST_SEED = 13337419                      # rl/env/MjEnv.py:2059


class BatchedMjEnv:

    @dataclass
    class Parameters:                     # rl/env/MjEnv.py:71-133 (image / scene / expert options left out)
        max_episode_steps: int = 250
        object_position_noise_mm: int = 10
        object_rotation_noise_deg: int = 5
        base_lim_X_mm: int = 300
        base_lim_Y_mm: int = 200
        base_lim_Z_mm: int = 30
        base_lim_yaw_rad: float = np.pi / 4
        test_obj_per_file: int = 20
        task_reload_chance: float = 1.0 / 20.0
        test_trials_per_object: int = 3
        test_objects: int = 100
        object_set_name: str = "set9_fullset"
        num_segments: int = 8
        finger_thickness: float = 0.9e-3
        finger_length: float = 235e-3
        finger_width: float = 28e-3
        finger_modulus: float = 193e9
        fixed_finger_hook: bool = True
        finger_hook_angle_degrees: float = 75.0
        finger_hook_length: float = 35e-3
        segment_inertia_scaling: float = 50.0
        fingertip_clearance: float = 10e-3

    @dataclass
    class Test:                           # rl/env/MjEnv.py:58-67
        obj_idx: int = 0
        obj_trial: int = 0
        obj_counter: int = 0
        xml_file: int = 0
        reward: float = 0
        steps: int = 0
        object_name: str = ""
        object_category: str = ""
        cnt: dict = field(default_factory=dict)

    def __init__(self, n_envs: int, object_set=None, seed=None, num_segments=None, finger_width=None,
                 finger_thickness=None, finger_modulus=None, log_level=0, device: int = 0, settings: Settings | None = None,
                 auto_reset: bool = False, calibrate: bool = True, precision: str = "fp32", env_id_offset: int = 0):
        self.params = BatchedMjEnv.Parameters()
        if object_set is not None: self.params.object_set_name = object_set
        if num_segments is not None: self.params.num_segments = num_segments
        if finger_width is not None: self.params.finger_width = finger_width
        if finger_thickness is not None: self.params.finger_thickness = finger_thickness
        if finger_modulus is not None: self.params.finger_modulus = finger_modulus
        self.n_envs, self.log_level, self.auto_reset = int(n_envs), log_level, auto_reset
        self.torch = True
        p = self.params
        gp = GripperParams(num_segments=p.num_segments, finger_width=p.finger_width, finger_thickness=p.finger_thickness,
                           finger_E=p.finger_modulus)
        st = settings or Settings()
        self.objects = build_object_set()
        self.mj = MjClass(self.n_envs, gp, objects=self.objects, settings=st, device=device, precision=precision,
                          env_id_offset=env_id_offset, calibrate=calibrate)
        self.device = self.mj.device
        self.mj.set_base_XYZ_limits(p.base_lim_X_mm * 1e-3, p.base_lim_Y_mm * 1e-3, p.base_lim_Z_mm * 1e-3)
        self.mj.set_base_yaw_limit(p.base_lim_yaw_rad)
        # task files: the first test_objects / test_obj_per_file are the test set (rl/env/MjEnv.py:394-409)
        self.task_files = task_split(len(self.objects), obj_per_task=p.test_obj_per_file)
        self.testing_xmls = p.test_objects // p.test_obj_per_file
        self.training_xmls = len(self.task_files) - self.testing_xmls
        train = np.concatenate(self.task_files[self.testing_xmls:])
        self._train_objects = torch.as_tensor(train, dtype=torch.int32, device=self.device)
        self.n_actions, self.n_obs = self.mj.get_n_actions(), self.mj.get_n_obs()
        dev, n = self.device, self.n_envs
        self.current_step = torch.zeros(n, dtype=torch.int32, device=dev)
        self.is_done = torch.zeros(n, dtype=torch.bool, device=dev)
        self.last_action = torch.full((n,), -1, dtype=torch.int32, device=dev)
        self.last_reward = torch.zeros(n, dtype=torch.float32, device=dev)
        self.cumulative_reward = torch.zeros(n, dtype=torch.float32, device=dev)
        self.test_in_progress = self.test_completed = False
        self.test_trial_data: list[BatchedMjEnv.Test] = []
        self.myseed = None
        self.seed(seed)

    # ------------------------------------------------------------------ seeding (rl/env/MjEnv.py:2023-2049)
    def seed(self, seed=None):
        if seed is None:
            seed = self.myseed if self.myseed is not None else int(np.random.default_rng().integers(0, 2_147_483_647))
        self.myseed = int(seed)
        self.random_train = torch.Generator(device=self.device); self.random_train.manual_seed(self.myseed)
        self.mj.set.random_seed = self.myseed            # applied by the next reset(), as in the reference
        return self.myseed

    # ------------------------------------------------------------------ episodes
    def _all(self, mask):
        if mask is None:
            return torch.ones(self.n_envs, dtype=torch.bool, device=self.device)
        return torch.as_tensor(mask, device=self.device).bool()

    def _spawn_object(self, mask: torch.Tensor):
        """rl/env/MjEnv.py:1177-1267: a training env gets a random object of the training task files, a test env the
        object of its current trial; the pose comes from spawn_into_scene (+-object_position_noise_mm, +-pi/2)."""
        p = self.params
        if self.test_in_progress:
            idx = self._test_obj.clone()
        else:
            pick = torch.randint(0, len(self._train_objects), (self.n_envs,), device=self.device, generator=self.random_train)
            idx = self._train_objects[pick]
        self.mj.default_spawn_params.xrange = self.mj.default_spawn_params.yrange = p.object_position_noise_mm * 1e-3
        self.mj.default_spawn_params.rotrange = np.pi / 2.0
        # rl/env/MjEnv.py:1216-1244: up to three tries of spawn_into_scene, then the "old method" -- a random pose on the
        # millimetre grid with one of three base angles, placed without a collision test
        self.mj.batch.spawn_with_retry(p.object_position_noise_mm * 1e-3, np.pi / 2.0, mask.to(torch.uint8), idx, tries=3,
                                       fallback_xy_mm=int(p.object_position_noise_mm),
                                       fallback_rot_deg=int(p.object_rotation_noise_deg))

    def reset(self, mask=None, hard=None):
        """Reset the envs of `mask` (all if None) and return the observation of EVERY env [n, n_obs]."""
        m = self._all(mask)
        if self.test_in_progress:
            self._assign_trials(m)
        z = ~m
        self.current_step.mul_(z.to(torch.int32)); self.cumulative_reward.mul_(z.to(torch.float32))
        self.last_reward.mul_(z.to(torch.float32)); self.is_done &= z
        self.last_action.masked_fill_(m, -1)
        if mask is None:
            self.mj.hard_reset() if hard else self.mj.reset()
        else:
            self.mj.reset(m.to(torch.uint8))
        self._spawn_object(m)
        return self._next_observation()

    def _next_observation(self):
        return self.mj.get_observation().clone()

    def step(self, action):
        """(obs, reward, terminated, truncated, info), all device tensors with a leading env axis."""
        if not self.auto_reset and bool(self.is_done.any()):
            raise RuntimeError("step has been called with done=true, use reset()")      # rl/env/MjEnv.py:2175-2177
        action = torch.as_tensor(action, device=self.device).to(torch.int32).reshape(self.n_envs)
        self.current_step.add_(1)
        obs, reward, done, _ = self.mj.fused_step(action)
        obs, reward = obs.clone(), reward.clone()
        truncated = self.current_step >= self.params.max_episode_steps           # checked first (:627-632)
        terminated = (done != 0) & ~truncated
        finished = terminated | truncated
        self.last_action.copy_(action); self.last_reward.copy_(reward)
        self.cumulative_reward.add_(reward)
        self.is_done = finished
        info = {}
        if self.test_in_progress and bool(finished.any()):
            self._monitor_test(finished)
        if self.auto_reset and bool(finished.any()):
            info["final_observation"] = obs
            info["episode_return"] = self.cumulative_reward.clone()
            info["episode_length"] = self.current_step.clone()
            fresh = self.reset(finished)
            obs = torch.where(finished.unsqueeze(1), fresh, obs)
        return obs, reward, terminated, truncated, info

    # ------------------------------------------------------------------ test protocol (rl/env/MjEnv.py:1269-1326, 2051-2065)
    def start_test(self):
        """Walk `test_objects` objects x `test_trials_per_object` trials; with n envs the (object, trial) pairs are
        dealt to envs in the reference's order as envs become free. Test poses come from a generator re-seeded with
        13337419 at every start_test (device RNG stream: the C++ side is re-seeded on the reset that follows)."""
        p = self.params
        self.random_test = torch.Generator(device=self.device); self.random_test.manual_seed(TEST_SEED)
        pairs = [(f, o, t) for f in range(self.testing_xmls) for o in range(p.test_obj_per_file) for t in range(p.test_trials_per_object)]
        self._test_queue = pairs[: p.test_objects * p.test_trials_per_object]
        self._test_next = 0
        self._test_obj = torch.zeros(self.n_envs, dtype=torch.int32, device=self.device)
        self._test_active = torch.zeros(self.n_envs, dtype=torch.bool, device=self.device)
        self._test_pair: list = [None] * self.n_envs
        self.test_trial_data = []
        self.test_in_progress, self.test_completed = True, False

    def _assign_trials(self, mask: torch.Tensor):
        idx = self._test_obj.cpu().numpy().copy()
        active = self._test_active.cpu().numpy().copy()
        for e in np.nonzero(mask.cpu().numpy())[0]:
            if self._test_next < len(self._test_queue):
                f, o, t = self._test_queue[self._test_next]; self._test_next += 1
                self._test_pair[e] = (f, o, t)
                idx[e] = int(self.task_files[f][o]); active[e] = True
            else:
                self._test_pair[e] = None; active[e] = False
        self._test_obj = torch.as_tensor(idx, dtype=torch.int32, device=self.device)
        self._test_active = torch.as_tensor(active, device=self.device)

    def _monitor_test(self, finished: torch.Tensor):
        fin = finished.cpu().numpy()
        ret, steps = self.cumulative_reward.cpu().numpy(), self.current_step.cpu().numpy()
        ev = self.mj.get_event_state().cpu().numpy()
        for e in np.nonzero(fin)[0]:
            pair = self._test_pair[e]
            if pair is None:
                continue
            f, o, t = pair
            spec = self.objects[int(self.task_files[f][o])]
            self.test_trial_data.append(BatchedMjEnv.Test(
                obj_idx=o, obj_trial=t, obj_counter=f * self.params.test_obj_per_file + o, xml_file=f,
                reward=float(ret[e]), steps=int(steps[e]), object_name=spec.name,
                object_category=spec.name.split("_")[0], cnt={"event_state": ev[e].copy()}))
            self._test_pair[e] = None
        if len(self.test_trial_data) >= len(self._test_queue):
            self._end_test()

    def _end_test(self):
        self.test_in_progress, self.test_completed = False, True
        self._test_active.zero_()

    def run_test(self, policy, max_vector_steps: int = 100000) -> list:
        """Convenience: the whole test with `policy(obs) -> actions`; envs with no trial left idle on action 0."""
        self.start_test()
        obs = self.reset()
        keep_auto, self.auto_reset = self.auto_reset, True
        try:
            for _ in range(max_vector_steps):
                if self.test_completed:
                    break
                obs, _, _, _, _ = self.step(policy(obs))
        finally:
            self.auto_reset = keep_auto
        return self.test_trial_data

    # ------------------------------------------------------------------ misc surface
    def using_continous_actions(self) -> bool:
        return bool(self.mj.set.continous_actions)

    def yield_load(self):
        return self.mj.yield_load()

    def get_parameters(self):
        return self.params

    def get_params_dict(self):
        d = asdict(self.params)
        d.update({"n_envs": self.n_envs, "n_actions": self.n_actions, "n_obs": self.n_obs})
        return d

    def get_save_state(self):
        return {"parameters": self.params, "mjcpp": self.mj, "myseed": self.myseed, "n_envs": self.n_envs}   # :2088-2097

    def render(self):
        raise RuntimeError("rendering is a graphics-API path: unavailable on this box")

    def close(self):
        self.mj.close()
