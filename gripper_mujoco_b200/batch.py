"""GripperBatch: thin owner of one GmjHandle (include/gmj.h) and of the device in/out buffers.

torch is used for device memory and streams only; every simulation op is a call into
libgripper_b200.so. Method names follow MjClass (src/bind.cpp:43-170) with a leading batch axis.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import lib as _lib
from .lib import GMJ_EVENT_STATE_LEN, GMJ_PROBE_LEN, GP, GmjStats, check
from .model import Model, build_model
from .objects import ObjectSpec, pack_objects
from .settings import Settings


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class GripperBatch:
    def __init__(self, n_envs: int, model: Model | None = None, settings: Settings | None = None,
                 device: int = 0, precision: str = "fp32", env_id_offset: int = 0):
        if not torch.cuda.is_available():
            raise RuntimeError("gripper_mujoco_b200 needs a CUDA device: there is no CPU fallback")
        self.L = _lib.lib()
        self.model = model or build_model()
        self.set = settings or Settings()
        self.n_envs = int(n_envs)
        self.device_index = int(device)
        self.device = torch.device("cuda", self.device_index)
        self.precision = {"fp32": 0, "fp64": 1}[precision]
        self._mstruct = self.model.to_struct()
        self._sstruct = self.set.pack()
        h = C.c_void_p()
        check(self.L.gmj_create(C.byref(self._mstruct), C.byref(self._sstruct), self.n_envs, self.device_index,
                                self.precision, env_id_offset, C.byref(h)))
        self.h = h
        self.B = self.model.B
        self.nq, self.nv = self.B + 7, self.B + 6
        self.objects: list[ObjectSpec] = []
        self._alloc_outputs()

    # ------------------------------------------------------------------ plumbing
    def _alloc_outputs(self):
        n = self.n_envs
        self.n_obs = self.L.gmj_n_obs(self.h)
        self.n_actions = self.L.gmj_n_actions(self.h)
        self.obs = torch.zeros((n, max(self.n_obs, 1)), dtype=torch.float32, device=self.device)
        self.reward = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.done = torch.zeros(n, dtype=torch.uint8, device=self.device)
        self.events = torch.zeros((n, GMJ_EVENT_STATE_LEN), dtype=torch.float32, device=self.device)
        self.spawned = torch.zeros(n, dtype=torch.uint8, device=self.device)      # spawn_into_scene's return value per env

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        if getattr(self, "h", None):
            self.L.gmj_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def push_settings(self):
        """Upload `self.set` (the mj.set mirror) to the device; call after editing settings."""
        self._sstruct = self.set.pack()
        check(self.L.gmj_set_settings(self.h, C.byref(self._sstruct)))
        n_obs = self.L.gmj_n_obs(self.h)
        if n_obs != self.n_obs or self.L.gmj_n_actions(self.h) != self.n_actions:
            self._alloc_outputs()

    def set_objects(self, specs: list[ObjectSpec]):
        self.objects = list(specs)
        self._objarr = pack_objects(self.objects)
        check(self.L.gmj_set_objects(self.h, self._objarr, len(self.objects)))

    def _mask(self, mask):
        if mask is None:
            return None
        return torch.as_tensor(mask, device=self.device).to(torch.uint8).contiguous()

    # ------------------------------------------------------------------ MjClass surface
    def reset(self, mask=None):
        m = self._mask(mask)
        check(self.L.gmj_reset(self.h, _ptr(m), self._stream()))

    def spawn_object(self, index, x=0.0, y=0.0, zrot=0.0, mask=None):
        """MjClass::spawn_object(index, x, y, zrot): scalars broadcast, arrays are per env."""
        n = self.n_envs
        idx = torch.as_tensor(np.broadcast_to(np.asarray(index, dtype=np.int32), (n,)).copy(), device=self.device)
        xyr = np.stack([np.broadcast_to(np.asarray(v, dtype=np.float32), (n,)) for v in (x, y, zrot)], axis=1)
        xyr = torch.as_tensor(np.ascontiguousarray(xyr), device=self.device)
        m = self._mask(mask)
        check(self.L.gmj_spawn(self.h, _ptr(idx), _ptr(xyr), _ptr(m), 0.0, 0.0, 1, -1, 0, _ptr(self.spawned), self._stream()))

    def _index_tensor(self, index):
        if index is None:
            return None
        if torch.is_tensor(index):
            return index.to(device=self.device, dtype=torch.int32).expand(self.n_envs).contiguous()
        return torch.as_tensor(np.broadcast_to(np.asarray(index, dtype=np.int32), (self.n_envs,)).copy(), device=self.device)

    def spawn_into_scene(self, xy_range=0.0, rot_range=0.0, mask=None, index=None):
        """Random pose per env from the device RNG (MjClass::spawn_into_scene: shuffled candidate grid, first pose clear of
        the fingertips); the object too unless `index` (scalar or per-env int array / tensor) names it, as
        spawn_into_scene(index) does. Returns the uint8 device tensor `spawned` (1 = placed, 0 = no free candidate: the
        env has no object); only the masked-in entries are rewritten."""
        m = self._mask(mask)
        idx = self._index_tensor(index)
        check(self.L.gmj_spawn(self.h, _ptr(idx), None, _ptr(m), float(xy_range), float(rot_range), 1, -1, 0,
                               _ptr(self.spawned), self._stream()))
        self._keep = (idx, m)
        return self.spawned

    def spawn_with_retry(self, xy_range=0.0, rot_range=0.0, mask=None, index=None, tries: int = 3,
                         fallback_xy_mm: int = 10, fallback_rot_deg: int = 5):
        """MjEnv._spawn_object's policy for the whole batch in ONE launch (rl/env/MjEnv.py:1216-1244): up to `tries`
        spawn_into_scene attempts per env, then the "old method" for an env still without a free pose -- a pose on the
        millimetre grid within +-fallback_xy_mm with a base angle of 0 / 60 / 120 degrees +-fallback_rot_deg, placed
        without a collision test. `spawned` afterwards: 1 = placed by the search, 2 = by the fallback."""
        m = self._mask(mask)
        idx = self._index_tensor(index)
        check(self.L.gmj_spawn(self.h, _ptr(idx), None, _ptr(m), float(xy_range), float(rot_range), int(tries),
                               int(fallback_xy_mm), int(fallback_rot_deg), _ptr(self.spawned), self._stream()))
        self._keep = (idx, m)
        return self.spawned

    def reset_done(self, done=None, step_cap: int = 250):
        """MjEnv.step's end-of-episode rule on the device: reset the envs whose done flag is set or whose episode reached
        `step_cap` action steps; returns the uint8 device mask of the envs that were reset (gmj_reset_done)."""
        d = self.done if done is None else self._mask(done)
        if not hasattr(self, "_reset_mask"):
            self._reset_mask = torch.zeros(self.n_envs, dtype=torch.uint8, device=self.device)
        check(self.L.gmj_reset_done(self.h, _ptr(d), int(step_cap), _ptr(self._reset_mask), self._stream()))
        return self._reset_mask

    def step(self, actions, frac=None):
        """Fused set_action -> action_step -> observation / reward / done / events (device tensors)."""
        a = torch.as_tensor(actions, device=self.device).to(torch.int32).contiguous()
        f = None if frac is None else torch.as_tensor(frac, device=self.device).to(torch.float32).contiguous()
        check(self.L.gmj_step(self.h, _ptr(a), _ptr(f), _ptr(self.obs), _ptr(self.reward), _ptr(self.done),
                              _ptr(self.events), self._stream()))
        return self.obs, self.reward, self.done, self.events

    def set_action(self, actions, frac=None):
        a = torch.as_tensor(actions, device=self.device).to(torch.int32).contiguous()
        f = None if frac is None else torch.as_tensor(frac, device=self.device).to(torch.float32).contiguous()
        check(self.L.gmj_set_action(self.h, _ptr(a), _ptr(f), self._stream()))

    def action_step(self):
        check(self.L.gmj_action_step(self.h, self._stream()))

    def substeps(self, n: int, sensors: bool = True):
        check(self.L.gmj_substeps(self.h, int(n), int(bool(sensors)), self._stream()))

    def observe(self, obs=True, reward=True, done=True, events=True):
        check(self.L.gmj_observe(self.h, _ptr(self.obs) if obs else None, _ptr(self.reward) if reward else None,
                                 _ptr(self.done) if done else None, _ptr(self.events) if events else None,
                                 self._stream()))
        return self.obs, self.reward, self.done, self.events

    def observe_masked(self, mask):
        """First observation of a new episode (MjEnv.reset's return value) for the envs in `mask`, written into the rows
        of self.obs; the other rows keep the observation of the step."""
        m = self._mask(mask)
        check(self.L.gmj_observe_masked(self.h, _ptr(m), _ptr(self.obs), self._stream()))
        self._keep = (m,)
        return self.obs

    def set_welds(self, mask_on: int):
        check(self.L.gmj_set_welds(self.h, int(mask_on), self._stream()))

    # ------------------------------------------------------------------ host reads (synchronising)
    def state(self):
        torch.cuda.synchronize(self.device)
        qpos = np.zeros((self.n_envs, self.nq)); qvel = np.zeros((self.n_envs, self.nv))
        check(self.L.gmj_get_state_host(self.h, qpos.ctypes.data_as(C.c_void_p), qvel.ctypes.data_as(C.c_void_p)))
        return qpos, qvel

    def probe(self):
        torch.cuda.synchronize(self.device)
        p = np.zeros((self.n_envs, GMJ_PROBE_LEN))
        check(self.L.gmj_probe_host(self.h, p.ctypes.data_as(C.c_void_p)))
        return p

    def probe_noalloc(self, out: np.ndarray | None = None):
        """gmj_probe_host_noalloc: the same probe through the library's static device scratch -- no cudaMalloc /
        cudaFree, n_envs <= GMJ_NOALLOC_MAX_ENVS (16384). `out`: caller-owned float64 [n_envs, GMJ_PROBE_LEN] buffer."""
        torch.cuda.synchronize(self.device)
        p = np.zeros((self.n_envs, GMJ_PROBE_LEN)) if out is None else out
        check(self.L.gmj_probe_host_noalloc(self.h, p.ctypes.data_as(C.c_void_p)))
        return p

    def event_track(self):
        """(row [n, 36], abs [n, 36], last_value [n, 36], cumulative_reward [n]) of env_.cnt, events in X-macro order."""
        torch.cuda.synchronize(self.device)
        ra = np.zeros((self.n_envs, 36, 2), dtype=np.int32); last = np.zeros((self.n_envs, 36), dtype=np.float32)
        cr = np.zeros(self.n_envs, dtype=np.float32)
        check(self.L.gmj_event_track_host(self.h, ra.ctypes.data_as(C.c_void_p), last.ctypes.data_as(C.c_void_p),
                                          cr.ctypes.data_as(C.c_void_p)))
        return ra[:, :, 0], ra[:, :, 1], last, cr

    def probe_field(self, name):
        return self.probe()[:, GP[name]]

    def stats(self) -> dict:
        torch.cuda.synchronize(self.device)
        s = GmjStats()
        check(self.L.gmj_stats(self.h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in GmjStats._fields_}
