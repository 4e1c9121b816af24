"""CPU tests of the host side (settings packing, model / object-set construction, calibration flow)
and of oracle pieces whose reference code cannot be compiled here (samplers, reward helpers, Box2d),
checked against small numpy restatements written from the reference (file:line cited per test)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from common import objects
from gripper_mujoco_b200 import calibration
from gripper_mujoco_b200.model import GripperParams, build_model, forward_kinematics
from gripper_mujoco_b200.settings import (ACTION_NAMES, BINARY_NAMES, GENERAL_PACKED, LINEAR_NAMES, SENSOR_NAMES,
                                          Settings)
from oracle.binding import OracleEnv, lib
from test_oracle_golden import OrWindow, _L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = open(os.path.join(ROOT, "include", "gmj_types.h")).read()


def _enum(prefix):
    out = []
    for body in re.findall(r"enum\s*\{([^}]*)\}", HDR, flags=re.S):
        names = [t.split("=")[0].strip() for t in re.sub(r"/\*.*?\*/", "", body, flags=re.S).split(",") if t.strip()]
        if names and names[0].startswith(prefix):
            out = [n[len(prefix):] for n in names if n.startswith(prefix) and not n.endswith("COUNT") and not n.endswith("COUNT_")]
    return out


def test_settings_pack_order_matches_header_enums():
    assert _enum("GS_") == GENERAL_PACKED
    assert len(_enum("GBE_")) == len(BINARY_NAMES) == 15
    assert len(_enum("GLE_")) == len(LINEAR_NAMES) == 21
    assert len(_enum("GSEN_")) == len(SENSOR_NAMES) == 10
    assert len(_enum("GACT_")) == len(ACTION_NAMES) == 11
    s = Settings()
    assert s.n_actions() == 15                       # X, prismatic?, Z, base Z ... + termination (src/mjclass.cpp:108-141)
    p = s.pack()
    assert p.general[GENERAL_PACKED.index("sim_steps_per_action")] == 200
    s.wipe_rewards()
    assert all(getattr(s, n).trigger == 10000 for n in BINARY_NAMES + LINEAR_NAMES)


def test_linear_reward_set_keeps_overshoot_quirk():
    """LinearReward::set never assigns overshoot (src/mjclass.h Q5): the header default survives."""
    s = Settings()
    before = s.palm_force.overshoot
    s.palm_force.set(0.1, 0, 1, 2.0, 4.0, 99.0)
    assert s.palm_force.overshoot == before and s.palm_force.max == 4.0


@pytest.mark.parametrize("N", [5, 8])
def test_model_structure(N):
    m = build_model(GripperParams(num_segments=N))
    assert m.B == 3 * N + 8 and m.nv == m.B + 6 and m.nq == m.B + 7
    st = m.to_struct()
    assert abs(sum(st.mass[i] for i in range(m.B)) - 2.375) < 1e-9        # base 1.5 + 3 x (0.1 + 0.075 + ...) + palm
    pos, mat = forward_kinematics(m, np.array([st.qpos0[i] for i in range(m.B)]))
    tips = np.array([pos[st.tip_body[f]] for f in range(3)])
    # three fingers 120 degrees apart at the home radius, same height
    r = np.hypot(tips[:, 0], tips[:, 1])
    assert np.ptp(r) < 1e-9 and np.ptp(tips[:, 2]) < 1e-9
    ang = np.sort(np.mod(np.arctan2(tips[:, 1], tips[:, 0]), 2 * np.pi))
    assert np.allclose(np.diff(ang), 2 * np.pi / 3, atol=1e-9)


def test_object_set_counts():
    """Same expansion as description/mujoco/object_sets/build_object_set.py on set9_fullset."""
    from collections import Counter
    objs = objects()
    c = Counter(o.category for o in objs)
    assert len(objs) == 1750
    assert (c["cubes"], c["cuboids"], c["cylinders"], c["spheres"], c["ellipsoids"]) == (90, 486, 570, 302, 302)
    assert max(o.obj.mass for o in objs) <= 0.5 + 1e-9
    assert {round(o.obj.friction, 3) for o in objs} <= {0.5, 1.0, 1.5, 2.0}


def test_yield_load_formula():
    p = GripperParams()
    inertia = p.finger_width * p.finger_thickness ** 3 / 12
    expect = 215e6 * inertia / (0.5 * p.finger_thickness) / p.finger_length      # src/myfunctions.cpp:3587-3595
    assert abs(calibration.yield_load(p) / expect - 1) < 1e-6


def test_calibration_flow_over_oracle():
    """find_highest_stable_timestep / auto sim steps / sensor calibration. The reference's thesis data
    (max stable timestep 10.4 ms at N = 5, 4.3 ms at N = 8, 1.299 ms at N = 16, 0.7 ms at N = 24:
    rl/juypter/thesis_plots/mujoco_timesteps.csv, column t=0.9 w=28 inertia=50) is the sanity envelope for
    the restated physics: +-10 %."""
    from oracle.calibrate import calibrated
    for N, ref_max in [(5, 10.4e-3), (8, 4.3e-3), (16, 1.299e-3), (24, 0.7e-3)]:
        mdl, s, info = calibrated(GripperParams(num_segments=N))
        hi = info["highest_stable_timestep"]
        assert abs(hi / ref_max - 1) < 0.10
        factor = 0.65 if hi >= 5e-3 else 0.75 if hi > 3e-3 else 0.8
        assert abs(s.mujoco_timestep - int(np.float32(hi) * np.float32(factor) * 1e6) * 1e-6) < 1.1e-6
        assert s.sim_steps_per_action == int(np.ceil(s.time_for_action / s.mujoco_timestep))
        assert -24.5 < s.wrist_sensor_Z.raw_value_offset < -22.0           # weight of the 2.375 kg gripper
        assert s.bending_gauge.normalise > 0 and s.sim_gauge_raw_to_N_factor > 0
        assert 0.010 < s.reset_qpos_base < 0.013                           # base sag under gravity on the PD spring


def _window(vals):
    L = _L()
    w = OrWindow()
    L.or_win_reset(C.byref(w))
    for v in vals:
        L.or_win_add(C.byref(w), float(v))
    return w


@pytest.mark.parametrize("p,r", [(3, 1), (3, 2), (2, 3), (1, 1)])
def test_samplers_against_restatement(p, r):
    """src/mjclass.h:278-453. total_readings = 1 + r p; output [x0, f(dx), x1, f(dx), x2 ...]."""
    L = lib()
    rng = np.random.default_rng(p * 10 + r)
    vals = rng.uniform(-1, 1, 20).astype(np.float32)
    w = _window(vals)
    T = 1 + r * p
    back = lambda n: vals[len(vals) - 1 - n]          # read_element(n)
    out = (C.c_float * 64)()
    n = L.or_sample(0, p, r, C.byref(w), out)          # raw: the last T-1 readings, oldest first
    assert n == T - 1 and np.array_equal(np.array(out[:n], dtype=np.float32), vals[len(vals) - (T - 1):])
    xs = [back(T - 1 - i * r) for i in range(p + 1)]
    n = L.or_sample(1, p, r, C.byref(w), out)          # change
    exp = [xs[0]]
    for i in range(p):
        exp += [np.float32(xs[i + 1] - xs[i]), xs[i + 1]]
    assert n == 2 * p + 1 and np.array_equal(np.array(out[:n], dtype=np.float32), np.array(exp, dtype=np.float32))
    n = L.or_sample(4, p, r, C.byref(w), out)          # sign, tolerance 1e-6
    got = np.array(out[:n], dtype=np.float32)
    for i in range(p):
        d = np.float32(xs[i + 1] - xs[i])
        assert got[2 * i + 1] == (1.0 if d > 1e-6 else -1.0 if d < -1e-6 else 0.0)
    n = L.or_sample(5, p, r, C.byref(w), out)          # scaled change: clipped to +-1
    assert n == 2 * p + 1 and np.all(np.abs(np.array(out[:n])[1::2]) <= 1.0)


def test_reward_helpers():
    L = lib()
    lr, nb = L.or_linear_reward, L.or_normalise_between
    # src/mjclass.cpp:4866-4894
    assert lr(0.5, 1.0, 2.0, -1.0) == 0.0 and lr(1.5, 1.0, 2.0, -1.0) == 0.5 and lr(3.0, 1.0, 2.0, -1.0) == 1.0
    assert lr(5.0, 1.0, 3.0, 6.0) == pytest.approx(1.0 / 3.0) and lr(7.0, 1.0, 3.0, 6.0) == 0.0
    # src/mjclass.cpp:4896-4904
    assert nb(0.0, -1.0, 3.0) == -0.5 and nb(9.0, -1.0, 3.0) == 1.0 and nb(-9.0, -1.0, 3.0) == -1.0


def test_box2d_overlap():
    """luke::Box2d (src/customtypes.h:86-140): separating axes with a minimum gap."""
    L = lib()
    a, b = (C.c_double * 8)(), (C.c_double * 8)()
    L.or_box2d_init(a, 0.0, 0.0, 0.04, 0.04, 0.0)
    L.or_box2d_init(b, 0.05, 0.0, 0.04, 0.04, 0.0)             # 10 mm apart
    # quirk kept from the reference: boxes separated on an axis but closer than the gap make
    # containsOther false, so the gap argument never turns a "no" into a "yes"
    assert L.or_box2d_overlap(a, b, 0.0) == 0 and L.or_box2d_overlap(a, b, 0.02) == 0
    L.or_box2d_init(b, 0.03, 0.0, 0.04, 0.04, 0.0)             # 10 mm deep overlap
    assert L.or_box2d_overlap(a, b, 0.0) == 1
    L.or_box2d_init(b, 0.03, 0.0, 0.04, 0.04, np.pi / 4)       # rotated, corner pokes in
    assert L.or_box2d_overlap(a, b, 0.0) == 1


def test_oracle_rest_state_is_equilibrium():
    """After reset at the calibrated pose an idle env stays put and the wrist carries the gripper weight."""
    from common import make_pair
    mdl, st = make_pair(5, 2e-3)
    e = OracleEnv(mdl.to_struct(), st.pack())
    e.reset(1, 0, 1)
    q0 = e.qpos().copy()
    e.substep(200)
    assert np.abs(e.qpos()[:mdl.B] - q0[:mdl.B]).max() < 2e-5
    assert abs(e.peek().wrist_z + 2.375 * 9.81) < 0.05


@pytest.mark.parametrize("N,dt", [(5, 4e-3), (8, 2e-3), (16, 8e-4)])
def test_oracle_cantilever_matches_the_stiffness_derivation(N, dt):
    """Analytic physics check. The reference's joint stiffnesses (src/myfunctions.cpp:988-1048:
    c_1 = 2EI/L N^2/(N - 1/3), c_n = N EI/L) are derived so that N rigid links reproduce the Euler-Bernoulli tip
    deflection F L^3 / 3EI of a tip load. The calibration load (apply_segment_force, :1642-1727) acts at the COM of
    the last segment body, a < L, in the hook frame (15 degrees off the finger normal), so the static small-angle
    answer of the discrete chain is  d = F_perp sum_i (a - x_i)(L - x_i) / c_i  with joints at x_i = (i-1) L/N.
    The restated physics (oracle, soft welds, gravity on the hanging finger) settles within 3.5 % of it."""
    from common import make_pair
    from oracle.binding import OracleEnv
    mdl, st = make_pair(N, dt, noise=False)
    p = mdl.params
    L, ell = p.finger_length, p.finger_length / N
    seg = [b for b in mdl.bodies if b.name.startswith("finger_1_segment_link_")]
    assert len(seg) == N
    c = [b.stiffness for b in seg]
    EI = p.finger_E * p.finger_width * p.finger_thickness ** 3 / 12
    assert abs(c[1] / (N * EI / L) - 1) < 1e-6 and abs(c[0] / ((2 * EI / L) * N * N / (N - 1 / 3)) - 1) < 1e-6
    x = np.arange(N) * ell
    # with the load at the very tip the discrete chain gives Euler-Bernoulli exactly, for every N
    assert abs(sum((L - x[i]) ** 2 / c[i] for i in range(N)) / (L ** 3 / (3 * EI)) - 1) < 1e-6
    a = x[-1] + seg[-1].ipos[0]
    F = 0.25
    f_perp = F * np.cos(np.deg2rad(90.0 - p.hook_angle_degrees))
    pred = f_perp * sum((a - x[i]) * (L - x[i]) / c[i] for i in range(N))
    st.curve_validation, st.tip_force_applied = True, F
    e = OracleEnv(mdl.to_struct(), st.pack())
    e.reset(1, 0, 1)
    e.substep(int(3.0 / dt))
    q = e.qpos()
    for f in range(3):
        th = q[3 + f * (N + 2): 3 + f * (N + 2) + N]
        d = float((ell * np.sin(np.cumsum(th))).sum())
        assert abs(d / pred - 1) < 0.035, (f, d, pred)


def test_oracle_timed_baseline_mode_equals_the_checker():
    """bench.py's CPU arms run the oracle with fast_solver = 1 (oracle.h): straight to the active-set pivoting solve,
    skyline (tree-sparse) Cholesky in a fill-free dof order. Same model, same unique solution: trajectories of
    contact-rich random rollouts must agree with the checker path (Gauss-Seidel to 1e-13 + polish, dense
    factorisations) to rounding."""
    from common import make_pair, objects
    from oracle.binding import OracleEnv
    mdl, st = make_pair(5, 7.02e-3, steps_per_action=29, noise=True)
    out = []
    for fast in (0, 1):
        e = OracleEnv(mdl.to_struct(), st.pack()); e.set_fast(fast)
        rng = np.random.default_rng(0); qs, ev = [], []
        for ep in range(3):
            e.reset(1234, 0, ep + 1); e.spawn(objects()[int(rng.integers(0, 1750))].obj, 0.004, -0.002, 0.3)
            for t in range(24):
                e.set_action(int(rng.integers(0, 14)) if t < 16 else 1); e.action_step()
                qs.append(e.qpos().copy()); ev.append(e.event_state())
                if e.is_done():
                    break
        out.append((np.stack(qs), np.stack(ev), e.peek().contacts))
    assert out[0][2] > 1000                                     # contacts were simulated
    assert np.abs(out[0][0] - out[1][0]).max() < 1e-10
    assert np.array_equal(out[0][1][:, :15], out[1][1][:, :15])
