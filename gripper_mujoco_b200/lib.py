"""ctypes loader of libgripper_b200.so (include/gmj.h). There is NO CPU fallback: if the library is
missing or cannot be loaded this module raises, and gmj_create itself fails without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

from .model import GmjModel, GmjObject
from .settings import GmjSettings

_HERE = os.path.dirname(os.path.abspath(__file__))
# The same sources compiled with -DGMJ_SELF_COLLISION: gripper <-> gripper contacts (finger <-> finger, hook <-> hook,
# segment <-> palm) in the lane-per-dof kernels. Opt-in with GMJ_SELF_COLLISION=1 in the
# environment: ~10 % slower, and the state it adds is entered in ~0 % of the bench workloads' steps.
SC_SO_PATH = os.path.join(_HERE, "libgripper_b200_sc.so")
# GMJ_LIB: development override to A/B-test differently compiled builds of the same sources
BASE_SO_PATH = os.path.join(_HERE, "libgripper_b200.so")
SO_PATH = os.environ.get("GMJ_LIB") or (SC_SO_PATH if os.environ.get("GMJ_SELF_COLLISION") == "1" else BASE_SO_PATH)
CSRC = os.path.join(_HERE, "csrc")
SOURCES = ["gmj_api.cu", "gmj_inst.cu", "gmj_step.cuh", "gmj_wide.cuh", "gmj_physics.cuh", "gmj_primal.cuh", "gmj_primal_wide.cuh", "gmj_tail.cuh", "gmj_collide.cuh",
           "gmj_device.cuh"]
HEADERS = [os.path.join(_HERE, "..", "include", h) for h in ("gmj.h", "gmj_types.h")]

GMJ_PROBE_LEN = 40
GMJ_NOALLOC_MAX_ENVS = 16384      # include/gmj.h: static scratch of gmj_probe_host_noalloc
GMJ_EVENT_STATE_LEN = 57
PROBE_FIELDS = ["time", "wrist_z", "gauge0", "gauge1", "gauge2", "si_gauge0", "si_gauge1", "si_gauge2",
                "si_palm", "si_wrist", "ncon", "nrow", "niter", "unstable", "weld_on", "base_z",
                "tend_x", "tend_y", "tend_z", "tend_th", "tnext_x", "tnext_y", "tnext_z", "tnext_th",
                "cumulative_reward", "has_object", "obj_index", "episode", "last_reward", "last_done",
                "last_nrow", "base_x", "base_y", "base_yaw", "read_motor_x", "read_motor_y", "read_motor_z",
                "read_base_x", "read_base_y", "read_base_z"]
GP = {n: i for i, n in enumerate(PROBE_FIELDS)}

# every symbol include/gmj.h declares (tests check the .so exports exactly these)
SYMBOLS = ["gmj_create", "gmj_destroy", "gmj_set_settings", "gmj_set_objects", "gmj_reset", "gmj_reset_done", "gmj_spawn",
           "gmj_step", "gmj_set_action", "gmj_action_step", "gmj_substeps", "gmj_observe", "gmj_observe_masked", "gmj_n_obs",
           "gmj_n_actions", "gmj_n_envs", "gmj_get_state_host", "gmj_probe_host", "gmj_probe_host_noalloc", "gmj_event_track_host", "gmj_stats", "gmj_calibrate",
           "gmj_set_welds", "gmj_last_error"]


class GmjCalib(C.Structure):
    _fields_ = [("set_timestep", C.c_int32), ("calibrate_gauges", C.c_int32), ("set_sim_steps", C.c_int32),
                ("sim_steps_per_action", C.c_int32), ("highest_stable_timestep", C.c_double), ("timestep", C.c_double),
                ("wrist_raw_value_offset", C.c_double), ("bend_gauge_normalise", C.c_double),
                ("sim_gauge_raw_to_N_factor", C.c_double), ("reset_qpos", C.c_double * 8), ("yield_load", C.c_double),
                ("launches", C.c_int32), ("pad_", C.c_int32)]


class GmjStats(C.Structure):
    _fields_ = [("substeps", C.c_int64), ("unstable_resets", C.c_int64), ("contacts", C.c_int64),
                ("rows", C.c_int64), ("solver_iters", C.c_int64), ("contact_overflow", C.c_int64),
                ("relaxed_solves", C.c_int64), ("capped_solves", C.c_int64)]


NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]
# step-kernel instantiations, one translation unit each (compiled in parallel): (type, N, wide)
INSTANCES = [("float", 5, False), ("float", 8, False), ("double", 5, False), ("double", 8, False),
             ("float", 16, True), ("float", 24, True), ("double", 16, True), ("double", 24, True)]
OBJ_DIR = os.path.join(_HERE, "..", "build", "obj")


def nvcc_commands(out: str = BASE_SO_PATH, extra: list[str] | None = None, obj_dir: str = OBJ_DIR):
    """(compile commands, link command): every command is a plain `nvcc` invocation for sm_100a."""
    extra = list(extra or [])
    objs, cmds = [], []
    o = os.path.join(obj_dir, "gmj_api.o")
    cmds.append(["nvcc", *NVCC_FLAGS, *extra, "-c", os.path.join(CSRC, "gmj_api.cu"), "-o", o]); objs.append(o)
    for ty, n, wide in INSTANCES:
        o = os.path.join(obj_dir, f"gmj_inst_{ty}_{n}.o")
        cmds.append(["nvcc", *NVCC_FLAGS, *extra, f"-DINST_T={ty}", f"-DINST_N={n}", *(["-DINST_WIDE"] if wide else []),
                     "-c", os.path.join(CSRC, "gmj_inst.cu"), "-o", o])
        objs.append(o)
    link = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", out, *objs]
    return cmds, link


def stale(path: str = BASE_SO_PATH) -> bool:
    if not os.path.exists(path):
        return True
    t = os.path.getmtime(path)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + HEADERS
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, out: str = BASE_SO_PATH, extra: list[str] | None = None) -> str:
    """Compile the CUDA extension for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    if force or out != BASE_SO_PATH or stale():
        obj_dir = OBJ_DIR if out == BASE_SO_PATH else (OBJ_DIR + "_sc" if out == SC_SO_PATH else out + ".obj")
        os.makedirs(obj_dir, exist_ok=True)
        cmds, link = nvcc_commands(out, (extra or []) + (["-Xptxas", "-v"] if verbose else []), obj_dir)
        procs = [subprocess.Popen(c) for c in cmds]
        rcs = [p.wait() for p in procs]
        if any(rcs):
            raise subprocess.CalledProcessError(max(rcs), cmds[rcs.index(max(rcs))])
        subprocess.check_call(link)
    return out


def build_self_collision(force: bool = False) -> str:
    """The -DGMJ_SELF_COLLISION build of the same sources (SC_SO_PATH), loaded instead when GMJ_SELF_COLLISION=1."""
    if force or stale(SC_SO_PATH):
        build(force=True, out=SC_SO_PATH, extra=["-DGMJ_SELF_COLLISION"])
    return SC_SO_PATH


FMA_SO = os.path.join(_HERE, "libgmj_fma_peak.so")


def build_fma_peak(force: bool = False) -> str:
    src = os.path.join(CSRC, "fma_peak.cu")
    if force or not os.path.exists(FMA_SO) or os.path.getmtime(src) > os.path.getmtime(FMA_SO):
        subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-shared",
                               "-Xcompiler", "-fPIC", "-o", FMA_SO, src])
    return FMA_SO


def fma_peak_tflops(device: int = 0, iters: int = 20000) -> float:
    """Measured FP32 FMA-pipe peak of this GPU (TFLOP/s), csrc/fma_peak.cu."""
    L = C.CDLL(FMA_SO)
    out = C.c_double()
    if L.gmj_fma_peak(int(device), int(iters), C.byref(out)) != 0:
        raise RuntimeError("gmj_fma_peak failed")
    return out.value


_LIB = None


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(SO_PATH):
        raise RuntimeError(
            f"{SO_PATH} is missing: the CUDA extension is the only implementation of the hot path "
            "(no CPU fallback). Build it with `python -c 'import __graft_entry__ as g; g.build()'`.")
    L = C.CDLL(SO_PATH)
    vp, i32, f32 = C.c_void_p, C.c_int, C.c_float
    L.gmj_last_error.restype = C.c_char_p
    L.gmj_create.argtypes = [C.POINTER(GmjModel), C.POINTER(GmjSettings), i32, i32, i32, C.c_uint32, C.POINTER(vp)]
    L.gmj_destroy.argtypes = [vp]
    L.gmj_set_settings.argtypes = [vp, C.POINTER(GmjSettings)]
    L.gmj_set_objects.argtypes = [vp, C.POINTER(GmjObject), i32]
    L.gmj_reset.argtypes = [vp, vp, vp]
    L.gmj_spawn.argtypes = [vp, vp, vp, vp, C.c_double, C.c_double, i32, i32, i32, vp, vp]
    L.gmj_reset_done.argtypes = [vp, vp, i32, vp, vp]
    L.gmj_step.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp]
    L.gmj_set_action.argtypes = [vp, vp, vp, vp]
    L.gmj_action_step.argtypes = [vp, vp]
    L.gmj_substeps.argtypes = [vp, i32, i32, vp]
    L.gmj_observe.argtypes = [vp, vp, vp, vp, vp, vp]
    L.gmj_observe_masked.argtypes = [vp, vp, vp, vp]
    for f in ("gmj_n_obs", "gmj_n_actions", "gmj_n_envs"):
        getattr(L, f).argtypes = [vp]
    L.gmj_get_state_host.argtypes = [vp, vp, vp]
    L.gmj_probe_host.argtypes = [vp, vp]
    L.gmj_probe_host_noalloc.argtypes = [vp, vp]
    L.gmj_event_track_host.argtypes = [vp, vp, vp, vp]
    L.gmj_stats.argtypes = [vp, C.POINTER(GmjStats)]
    L.gmj_calibrate.argtypes = [vp, vp, C.c_int, C.c_int, C.POINTER(GmjCalib)]
    L.gmj_set_welds.argtypes = [vp, i32, vp]
    _LIB = L
    return L


class GmjError(RuntimeError):
    pass


def check(rc: int):
    if rc != 0:
        raise GmjError(lib().gmj_last_error().decode())
