"""ctypes binding of the CPU oracle (oracle/liboracle.so). TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; nothing under gripper_mujoco_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("kin.c", "collide.c", "physics.c", "env.c", "events.c",
                                             "oracle.h", "vec.h")]
    srcs.append(os.path.join(_HERE, "..", "include", "gmj_types.h"))
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.or_create.restype = C.c_void_p
        L.or_create.argtypes = [C.c_void_p, C.c_void_p]
        L.or_destroy.argtypes = [C.c_void_p]
        L.or_set_settings.argtypes = [C.c_void_p, C.c_void_p]
        L.or_reset.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32]
        L.or_spawn.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double]
        L.or_set_action.argtypes = [C.c_void_p, C.c_int, C.c_float]
        L.or_set_action.restype = C.c_int
        L.or_substep.argtypes = [C.c_void_p]
        L.or_physics_step.argtypes = [C.c_void_p]
        L.or_action_step.argtypes = [C.c_void_p]
        L.or_get_observation.argtypes = [C.c_void_p, C.c_void_p]
        L.or_get_observation.restype = C.c_int
        L.or_reward.argtypes = [C.c_void_p]
        L.or_reward.restype = C.c_float
        L.or_is_done.argtypes = [C.c_void_p]
        L.or_is_done.restype = C.c_int
        L.or_event_state.argtypes = [C.c_void_p, C.c_void_p]
        L.or_settle.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.or_read_gauge.argtypes = [C.c_void_p, C.c_int]
        L.or_read_gauge.restype = C.c_float
        L.or_forward.argtypes = [C.c_void_p]
        L.or_set_weld.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.or_peek.argtypes = [C.c_void_p, C.c_void_p]
        L.or_set_fast.argtypes = [C.c_void_p, C.c_int]
        L.or_set_self_collision.argtypes = [C.c_void_p, C.c_int]
        L.or_sample.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.or_sample.restype = C.c_int
        L.or_linear_reward.argtypes = [C.c_float] * 4
        L.or_linear_reward.restype = C.c_float
        L.or_normalise_between.argtypes = [C.c_float] * 3
        L.or_normalise_between.restype = C.c_float
        L.or_box2d_overlap.argtypes = [C.c_void_p, C.c_void_p, C.c_double]
        L.or_box2d_overlap.restype = C.c_int
        L.or_box2d_init.argtypes = [C.c_void_p] + [C.c_double] * 5
        _LIB = L
    return _LIB


class OrPeek(C.Structure):
    """Flat view of oracle state for parity checks (filled by or_peek)."""
    _fields_ = [("qpos", C.c_double * 96), ("qvel", C.c_double * 96), ("time", C.c_double),
                ("wrist_z", C.c_double), ("ncon", C.c_int), ("nrow", C.c_int), ("niter", C.c_int),
                ("unstable", C.c_int), ("tgt_end", C.c_double * 4), ("tgt_next", C.c_double * 4),
                ("tgt_steps", C.c_int * 6), ("base_z", C.c_double), ("weld_on", C.c_int * 7),
                ("pad", C.c_int), ("reset_qpos", C.c_double * 8), ("si", C.c_float * 5),
                ("pad2", C.c_float), ("con_dist", C.c_double * 24), ("con_force", C.c_double * 72),
                ("con_pos", C.c_double * 72), ("con_normal", C.c_double * 72), ("con_cls", C.c_int * 48),
                ("substeps", C.c_int64), ("contacts", C.c_int64), ("rows", C.c_int64), ("iters", C.c_int64)]


class OracleEnv:
    """One single-threaded oracle environment (mirrors one reference MjClass)."""

    def __init__(self, model_struct, settings_struct):
        self.L = lib()
        self._m, self._s = model_struct, settings_struct
        self.h = self.L.or_create(C.byref(model_struct), C.byref(settings_struct))
        self.B = model_struct.B

    def __del__(self):
        try:
            self.L.or_destroy(self.h)
        except Exception:
            pass

    def set_fast(self, on=True):
        """Timed-baseline solver (oracle.h: fast_solver): active-set pivoting without the Gauss-Seidel pre-sweeps."""
        self.L.or_set_fast(self.h, int(bool(on)))

    def set_self_collision(self, on=True):
        """Gripper <-> gripper contacts (oracle.h: self_collision). Off by default: the CUDA path has none."""
        self.L.or_set_self_collision(self.h, int(bool(on)))

    def set_settings(self, s):
        self._s = s
        self.L.or_set_settings(self.h, C.byref(s))

    def reset(self, seed=0, env_id=0, episode=0):
        self.L.or_reset(self.h, seed, env_id, episode)

    def spawn(self, obj, x=0.0, y=0.0, zrot=0.0):
        self._obj = obj
        self.L.or_spawn(self.h, C.byref(obj), x, y, zrot)

    def set_action(self, a, frac=0.0):
        return self.L.or_set_action(self.h, a, frac)

    def action_step(self):
        self.L.or_action_step(self.h)

    def substep(self, n=1):
        for _ in range(n):
            self.L.or_substep(self.h)

    def physics_step(self, n=1):
        for _ in range(n):
            self.L.or_physics_step(self.h)

    def settle(self, n, dt):
        self.L.or_settle(self.h, n, dt)

    def observation(self):
        buf = (C.c_float * 256)()
        n = self.L.or_get_observation(self.h, buf)
        return np.array(buf[:n], dtype=np.float32)

    def reward(self):
        return float(self.L.or_reward(self.h))

    def is_done(self):
        return bool(self.L.or_is_done(self.h))

    def event_state(self):
        buf = (C.c_float * 57)()
        self.L.or_event_state(self.h, buf)
        return np.array(buf[:], dtype=np.float32)

    def gauge(self, f):
        return float(self.L.or_read_gauge(self.h, f))

    def set_weld(self, i, on):
        self.L.or_set_weld(self.h, i, int(on))

    def peek(self) -> OrPeek:
        p = OrPeek()
        self.L.or_peek(self.h, C.byref(p))
        return p

    def qpos(self):
        return np.array(self.peek().qpos[: self.B + 7])

    def qvel(self):
        return np.array(self.peek().qvel[: self.B + 6])
