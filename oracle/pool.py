"""Multi-process pool of CPU oracle environments. TEST / MEASUREMENT INFRASTRUCTURE ONLY: used by
bench.py's cpu_baseline leg and `--impl reference` arm (one single-threaded env per worker process,
exactly how the reference scales on CPU: one MjClass per OS process)."""
from __future__ import annotations

import ctypes as C
import multiprocessing as mp
import os
import sys
import time

import numpy as np

_W = {}


def _init(root, n_segments, settings_bytes, seed_base):
    if root not in sys.path:
        sys.path.insert(0, root)
    from gripper_mujoco_b200.model import GripperParams, build_model
    from gripper_mujoco_b200.objects import build_object_set
    from gripper_mujoco_b200.settings import GmjSettings
    from oracle.binding import OracleEnv
    st = GmjSettings.from_buffer_copy(settings_bytes)
    mdl = build_model(GripperParams(num_segments=n_segments))
    rank = mp.current_process()._identity[0] if mp.current_process()._identity else 0
    env = OracleEnv(mdl.to_struct(), st)
    env.set_fast(True)            # timed baseline: direct active-set solve (oracle.h fast_solver), same results to 1e-9
    _W.update(env=env, objs=build_object_set(), rng=np.random.default_rng(seed_base + rank),
              rank=rank, episode=0, steps_in_ep=0, seed=int(st.general[3]), n_actions=None, need_reset=True)


def _rollout(args):
    """Run action-steps of the synthetic workload for `budget_s` seconds or `max_steps` steps."""
    budget_s, max_steps, cap = args
    w = _W
    env, rng = w["env"], w["rng"]
    n = 0
    t0 = time.perf_counter()
    while n < max_steps and (budget_s <= 0 or time.perf_counter() - t0 < budget_s):
        if w["need_reset"]:
            w["episode"] += 1
            env.reset(w["seed"], w["rank"], w["episode"])
            o = w["objs"][int(rng.integers(0, len(w["objs"])))]
            env.spawn(o.obj, float(rng.integers(-5, 6)) * 2e-3, float(rng.integers(-5, 6)) * 2e-3,
                      float(rng.integers(-15, 16)) * (np.pi / 30))
            w["steps_in_ep"] = 0
            w["need_reset"] = False
            if w["n_actions"] is None:
                w["n_actions"] = 15
        a = int(rng.integers(0, w["n_actions"]))
        env.set_action(a)
        env.action_step()
        env.observation(); env.reward()
        done = env.is_done()
        w["steps_in_ep"] += 1
        n += 1
        if done or w["steps_in_ep"] >= cap:
            w["need_reset"] = True
    return n, time.perf_counter() - t0


class OraclePool:
    def __init__(self, n_segments: int, settings_struct, cores: int | None = None, seed: int = 1000):
        self.cores = cores or len(os.sched_getaffinity(0))
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        ctx = mp.get_context("spawn")
        self.pool = ctx.Pool(self.cores, initializer=_init,
                             initargs=(root, n_segments, bytes(settings_struct), seed))

    def run(self, budget_s: float = 0.0, max_steps: int = 1 << 30, cap: int = 250):
        """Every worker runs the rollout concurrently; returns (total action-steps, wall seconds)."""
        t0 = time.perf_counter()
        res = self.pool.map(_rollout, [(budget_s, max_steps, cap)] * self.cores, chunksize=1)
        wall = time.perf_counter() - t0
        return sum(r[0] for r in res), wall

    def close(self):
        self.pool.close(); self.pool.join()
