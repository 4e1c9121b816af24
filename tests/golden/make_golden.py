"""Generate the golden fixtures of tests/golden/ from the reference's own standalone sources
(oracle/_ref/libgripper_ref.so = /root/reference/src/gripper.cpp + slidingwindow.h compiled in place
by `make -C oracle ref`). Run in the build container (the reference is not on the GPU box):

    python tests/golden/make_golden.py

Writes gripper_kinematics.npz (motor-space setters, limits, stepper) and sliding_window.npz.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
R = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libgripper_ref.so"))
R.ref_gripper_new.restype = C.c_void_p
R.ref_win_new.restype = C.c_void_p
R.ref_win_read.restype = C.c_float
R.ref_win_read_latest.restype = C.c_float
for f, a in [("ref_gripper_free", [C.c_void_p]), ("ref_gripper_reset", [C.c_void_p]),
             ("ref_gripper_set_xyz_m", [C.c_void_p] + [C.c_double] * 3),
             ("ref_gripper_set_xyz_m_rad", [C.c_void_p] + [C.c_double] * 3),
             ("ref_gripper_step_to", [C.c_void_p, C.c_void_p, C.c_int]), ("ref_gripper_get", [C.c_void_p, C.c_void_p]),
             ("ref_win_new", [C.c_int]), ("ref_win_free", [C.c_void_p]), ("ref_win_reset", [C.c_void_p]),
             ("ref_win_add", [C.c_void_p, C.c_float]), ("ref_win_read", [C.c_void_p, C.c_int]),
             ("ref_win_read_latest", [C.c_void_p])]:
    getattr(R, f).argtypes = a


def get(g):
    out = (C.c_double * 11)()
    R.ref_gripper_get(g, out)
    return np.array(out[:])


def kinematics(seed=20240611, n=4000):
    rng = np.random.default_rng(seed)
    g, tgt = R.ref_gripper_new(), R.ref_gripper_new()
    ops, args, rets, states = [], [], [], []
    for i in range(n):
        op = int(rng.integers(0, 5))
        a = np.zeros(4)
        if op == 0:        # set_xyz_m on the target, values straddling the limits
            x = rng.uniform(0.040, 0.140)
            # |y - x| up to 30 mm exercises the +-40 degree clamp; 2 % of the draws are unrelated
            # (asin domain error -> NaN angle, which the restatement must reproduce too)
            y = rng.uniform(0.040, 0.140) if rng.random() < 0.02 else x + rng.uniform(-0.030, 0.030)
            a[:3] = [x, y, rng.uniform(-0.01, 0.175)]
            ret = R.ref_gripper_set_xyz_m(tgt, *a[:3])
        elif op == 1:      # set_xyz_m_rad on the target
            a[:3] = [rng.uniform(0.045, 0.138), rng.uniform(-0.8, 0.8), rng.uniform(-0.01, 0.175)]
            ret = R.ref_gripper_set_xyz_m_rad(tgt, *a[:3])
        elif op == 2 and np.isfinite(get(tgt)[7]):      # the action increments of the env: +-1 mm / 0.01 rad / 2 mm on the current target
            cur = get(tgt)
            k = int(rng.integers(0, 3)); sgn = 1 if rng.random() < 0.5 else -1
            d = [1e-3, 0.01, 2e-3][k] * sgn
            a[:3] = [cur[0] + (d if k == 0 else 0), cur[7] + (d if k == 1 else 0), cur[2] + (d if k == 2 else 0)]
            ret = R.ref_gripper_set_xyz_m_rad(tgt, *a[:3])
        elif op in (2, 3):      # stepper: move g towards tgt by num motor steps
            a[3] = int(rng.integers(1, 40))
            ret = R.ref_gripper_step_to(g, tgt, int(a[3]))
        else:
            if rng.random() < 0.1:
                R.ref_gripper_reset(g); ret = 1; a[3] = -1
            else:
                a[3] = 10
                ret = R.ref_gripper_step_to(g, tgt, 10)
        ops.append(op); args.append(a); rets.append(int(bool(ret)))
        states.append(np.concatenate([get(g), get(tgt)]))
    R.ref_gripper_free(g); R.ref_gripper_free(tgt)
    np.savez_compressed(os.path.join(HERE, "gripper_kinematics.npz"), seed=seed, ops=np.array(ops, dtype=np.int8),
                        args=np.array(args), rets=np.array(rets, dtype=np.int8), states=np.array(states))


def window(seed=7, n=600):
    rng = np.random.default_rng(seed)
    w = R.ref_win_new(1000)
    vals = rng.standard_normal(n).astype(np.float32)
    reads = np.zeros((n, 9), dtype=np.float32)
    resets = (rng.random(n) < 0.02)
    count = 0
    for i in range(n):
        if resets[i]:
            R.ref_win_reset(w); count = 0
        R.ref_win_add(w, float(vals[i])); count += 1
        reads[i, 0] = R.ref_win_read_latest(w)
        for k in range(8):
            reads[i, 1 + k] = R.ref_win_read(w, k)      # beyond `count` reads the zero-initialised ring
    R.ref_win_free(w)
    np.savez_compressed(os.path.join(HERE, "sliding_window.npz"), seed=seed, vals=vals, resets=resets, reads=reads)


if __name__ == "__main__":
    kinematics()
    window()
    print("golden fixtures written to", HERE)
