"""Shared helpers of the test-suite: identical model / settings for the oracle and the CUDA path."""
from __future__ import annotations

import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from gripper_mujoco_b200.model import GripperParams, build_model  # noqa: E402
from gripper_mujoco_b200.objects import build_object_set  # noqa: E402
from gripper_mujoco_b200.settings import Settings  # noqa: E402
from oracle.binding import OracleEnv  # noqa: E402  (tests are allowed to use the oracle)

_CAL_CACHE: dict = {}


def oracle_calibration(N: int):
    """Equilibrium of base + motor joints under gravity with all welds on (calibrate_reset,
    src/myfunctions.cpp:1450-1516: 400 steps at 2 ms). Computed once per N with the oracle."""
    if N not in _CAL_CACHE:
        from gripper_mujoco_b200 import calibration
        from oracle.calibrate import make_backend
        p = GripperParams(num_segments=N)
        mdl = build_model(p)
        s = Settings()
        calibration.calibrate_reset(p, mdl.N, mdl.B, s, make_backend(mdl))
        _CAL_CACHE[N] = np.array([s.reset_qpos_base] + [getattr(s, f"reset_qpos_g{k}") for k in range(7)])
    return _CAL_CACHE[N]


def make_settings(N: int, timestep: float, steps_per_action: int | None = None, noise: bool = False,
                  seed: int = 1234) -> Settings:
    s = Settings()
    s.mujoco_timestep = timestep
    s.sim_steps_per_action = steps_per_action or int(np.ceil(s.time_for_action / timestep))
    s.random_seed = seed
    rq = oracle_calibration(N)
    s.reset_qpos_base = float(rq[0])
    for k in range(7):
        setattr(s, f"reset_qpos_g{k}", float(rq[1 + k]))
    s.set_use_noise(noise)
    s.use_noise_seeded = 1 if noise else 0
    if not noise:
        s.base_position_noise = 0.0
    return s


def make_pair(N: int = 5, timestep: float = 2e-3, **kw):
    mdl = build_model(GripperParams(num_segments=N))
    st = make_settings(N, timestep, **kw)
    return mdl, st


_OBJ_CACHE = None


def objects():
    global _OBJ_CACHE
    if _OBJ_CACHE is None:
        _OBJ_CACHE = build_object_set()
    return _OBJ_CACHE


def find_object(prefix: str):
    for i, o in enumerate(objects()):
        if o.name.startswith(prefix):
            return i, o
    raise KeyError(prefix)


class OracleBatch:
    """n independent oracle envs driven like GripperBatch (same env ids / episode numbering)."""

    def __init__(self, n, mdl, st, env_id_offset=0):
        self.ms, self.ss = mdl.to_struct(), st.pack()
        self.envs = [OracleEnv(self.ms, self.ss) for _ in range(n)]
        self.episode = 0
        self.seed = int(st.random_seed)
        self.off = env_id_offset
        self.reset()

    def reset(self):
        self.episode += 1
        for i, e in enumerate(self.envs):
            e.reset(self.seed, self.off + i, self.episode)

    def spawn(self, idx, x, y, r):
        objs = objects()
        for i, e in enumerate(self.envs):
            e.spawn(objs[int(np.broadcast_to(idx, (len(self.envs),))[i])].obj,
                    float(np.broadcast_to(x, (len(self.envs),))[i]),
                    float(np.broadcast_to(y, (len(self.envs),))[i]),
                    float(np.broadcast_to(r, (len(self.envs),))[i]))

    def step(self, actions, frac=None):
        obs, rew, done, ev = [], [], [], []
        for i, (e, a) in enumerate(zip(self.envs, actions)):
            rc = e.set_action(int(a)) if frac is None else e.set_action(int(a), float(frac[i]))
            assert rc == 0
            e.action_step()
            obs.append(e.observation()); rew.append(e.reward()); done.append(e.is_done()); ev.append(e.event_state())
        return np.stack(obs), np.array(rew, dtype=np.float32), np.array(done), np.stack(ev)

    def qpos(self):
        return np.stack([e.qpos() for e in self.envs])

    def qvel(self):
        return np.stack([e.qvel() for e in self.envs])


# ------------------------------------------------------------------ oracle trajectories on all host cores
_RW: dict = {}


def _roll_init(root, n_segments, settings_bytes):
    if root not in sys.path:
        sys.path.insert(0, root)
    from gripper_mujoco_b200.settings import GmjSettings
    st = GmjSettings.from_buffer_copy(settings_bytes)
    mdl = build_model(GripperParams(num_segments=n_segments))
    _RW.update(ms=mdl.to_struct(), ss=st, objs=build_object_set(), seed=int(st.general[3]), B=mdl.B)


def _roll_chunk(args):
    """One oracle env per entry: reset(seed, env id, episode) -> spawn -> T x (set_action, action_step, observation,
    is_done, reward, event_state); per step outputs + per step qpos / qvel."""
    env_ids, episode, obj_idx, xyr, actions, frac, search, stop_at_done = args
    w = _RW
    out = []
    for k, eid in enumerate(env_ids):
        e = OracleEnv(w["ms"], w["ss"])
        e.reset(w["seed"], int(eid), int(episode))
        if os.environ.get("ORACLE_SELF_COLLISION") == "1":      # profiles/r2_m9_effect.py: what the M9 gap is worth
            e.set_self_collision(True)
        if search is not None:          # MjEnv's spawn policy with the env's own draws (gmj_spawn with tries / fallback)
            e.L.or_spawn_retry.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int]
            e.L.or_spawn_retry(e.h, C.byref(w["objs"][int(obj_idx[k])].obj), *search)
        elif obj_idx[k] >= 0:
            e.spawn(w["objs"][int(obj_idx[k])].obj, float(xyr[k, 0]), float(xyr[k, 1]), float(xyr[k, 2]))
        obs, rew, done, ev, qp, qv = [], [], [], [], [], []
        for t in range(actions.shape[0]):
            rc = e.set_action(int(actions[t, k])) if frac is None else e.set_action(int(actions[t, k]), float(frac[t, k]))
            assert rc == 0
            e.action_step()
            obs.append(e.observation()); done.append(e.is_done()); rew.append(np.float32(e.reward())); ev.append(e.event_state())
            qp.append(e.qpos()); qv.append(e.qvel())
            if stop_at_done and done[-1]:       # episode over: the remaining steps repeat the terminal outputs (callers mask them)
                for lst in (obs, rew, done, ev, qp, qv):
                    lst.extend([lst[-1]] * (actions.shape[0] - 1 - t))
                break
        out.append((np.stack(obs), np.array(rew, dtype=np.float32), np.array(done), np.stack(ev), np.stack(qp), np.stack(qv)))
    return out


def oracle_rollout(N, st, env_ids, episode, obj_idx, xyr, actions, frac=None, cores=None, search=None, stop_at_done=False):
    """Oracle trajectories of many independent envs, spread over the host cores (one process per core).
    actions: int [T, n]. Returns dict of arrays with a leading [T, n] (obs, reward, done, events, qpos, qvel)."""
    import multiprocessing as mp
    cores = cores or len(os.sched_getaffinity(0))
    env_ids = np.asarray(env_ids); n = len(env_ids)
    chunks = np.array_split(np.arange(n), min(n, cores * 8))
    ctx = mp.get_context("spawn")
    with ctx.Pool(cores, initializer=_roll_init, initargs=(ROOT, N, bytes(st.pack()))) as pool:
        res = pool.map(_roll_chunk, [(env_ids[c], episode, np.asarray(obj_idx)[c], np.asarray(xyr)[c], np.asarray(actions)[:, c],
                                      None if frac is None else np.asarray(frac)[:, c], search, stop_at_done) for c in chunks], chunksize=1)
    flat = [r for ch in res for r in ch]
    keys = ("obs", "reward", "done", "events", "qpos", "qvel")
    return {k: np.stack([f[i] for f in flat], axis=1) for i, k in enumerate(keys)}
