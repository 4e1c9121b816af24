"""Pin the oracle's non-physics restatement against the reference itself: golden vectors generated
from /root/reference/src/gripper.cpp + slidingwindow.h compiled verbatim (tests/golden/make_golden.py),
and -- where the reference is mounted -- against the live oracle/_ref build."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle.binding import lib

HERE = os.path.dirname(os.path.abspath(__file__))


class OrGripper(C.Structure):
    _fields_ = [("x", C.c_double), ("y", C.c_double), ("z", C.c_double), ("th", C.c_double),
                ("sx", C.c_int), ("sy", C.c_int), ("sz", C.c_int)]


class OrWindow(C.Structure):
    _fields_ = [("v", C.c_float * 16), ("i", C.c_int)]


def _L():
    L = lib()
    L.or_gripper_reset.argtypes = [C.c_void_p]
    L.or_gripper_set_xyz_m.argtypes = [C.c_void_p] + [C.c_double] * 3
    L.or_gripper_set_xyz_m_rad.argtypes = [C.c_void_p] + [C.c_double] * 3
    L.or_gripper_step_to.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.or_gripper_th.argtypes = [C.c_void_p]
    L.or_gripper_th.restype = C.c_double
    L.or_win_reset.argtypes = [C.c_void_p]
    L.or_win_add.argtypes = [C.c_void_p, C.c_float]
    L.or_win_read.argtypes = [C.c_void_p, C.c_int]
    L.or_win_read.restype = C.c_float
    return L


def _state(L, g):
    th = L.or_gripper_th(C.byref(g))
    return np.array([g.x, g.y, g.z, g.th, g.sx, g.sy, g.sz, th, g.x, -th, g.z])


def test_gripper_kinematics_golden():
    """4000 random setter / limit / stepper operations: positions, angle and motor step counts of the
    oracle equal the reference's luke::Gripper bit for bit, and so do the within-limits flags."""
    d = np.load(os.path.join(HERE, "golden", "gripper_kinematics.npz"))
    L = _L()
    g, tgt = OrGripper(), OrGripper()
    L.or_gripper_reset(C.byref(g)); L.or_gripper_reset(C.byref(tgt))
    for op, a, ret, st in zip(d["ops"], d["args"], d["rets"], d["states"]):
        if op == 0:
            r = L.or_gripper_set_xyz_m(C.byref(tgt), a[0], a[1], a[2])
        elif op == 1 or (op == 2 and a[3] == 0):
            r = L.or_gripper_set_xyz_m_rad(C.byref(tgt), a[0], a[1], a[2])
        elif a[3] < 0:
            L.or_gripper_reset(C.byref(g)); r = 1
        else:
            r = L.or_gripper_step_to(C.byref(g), C.byref(tgt), int(a[3]))
        assert int(bool(r)) == int(ret)
        got = np.concatenate([_state(L, g), _state(L, tgt)])
        assert np.array_equal(got, st, equal_nan=True), (op, a, got - st)


def test_sliding_window_golden():
    d = np.load(os.path.join(HERE, "golden", "sliding_window.npz"))
    L = _L()
    w = OrWindow()
    L.or_win_reset(C.byref(w))
    for v, rs, reads in zip(d["vals"], d["resets"], d["reads"]):
        if rs:
            L.or_win_reset(C.byref(w))
        L.or_win_add(C.byref(w), float(v))
        got = [L.or_win_read(C.byref(w), 0)] + [L.or_win_read(C.byref(w), k) for k in range(8)]
        assert np.array_equal(np.array(got, dtype=np.float32), reads)


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="reference not mounted (GPU box)")
def test_golden_regenerates_identically(tmp_path):
    """The committed fixtures are what the reference sources produce today."""
    import subprocess
    import sys
    root = os.path.dirname(HERE)
    subprocess.check_call(["make", "-C", os.path.join(root, "oracle"), "-s", "ref"])
    old = {f: np.load(os.path.join(HERE, "golden", f)) for f in ("gripper_kinematics.npz", "sliding_window.npz")}
    old = {f: {k: v[k].copy() for k in v.files} for f, v in old.items()}
    subprocess.check_call([sys.executable, os.path.join(HERE, "golden", "make_golden.py")], stdout=subprocess.DEVNULL)
    for f, o in old.items():
        new = np.load(os.path.join(HERE, "golden", f))
        for k in o:
            assert np.array_equal(o[k], new[k], equal_nan=o[k].dtype.kind == 'f'), (f, k)
