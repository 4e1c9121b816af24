"""Parity of the wide step kernel (long fingers, one environment per 96-thread block: csrc/gmj_wide.cuh)
against the CPU oracle, through the C ABI. configs[4] runs N = 24 segments per finger (B = 80 dofs);
N = 16 is the second instantiation. Same bars as tests/test_gpu_parity.py."""
import numpy as np
import pytest

from common import OracleBatch, find_object, make_pair, objects
from test_gpu_parity import DISCRETE, OBJECTS, _batch, _compare

pytestmark = pytest.mark.gpu

DT = {16: 1.0e-3, 24: 5.6e-4}          # the auto-calibrated timesteps (1.039 ms / 0.560 ms), rounded


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
@pytest.mark.parametrize("N", [16, 24])
def test_wide_free_space_random_actions(precision, N):
    n = 6
    mdl, st = make_pair(N, DT[N], steps_per_action=60, noise=False)
    gb, ob = _batch(n, mdl, st, precision), OracleBatch(n, mdl, st)
    gb.reset(); ob.reset()
    rng = np.random.default_rng(N)
    for _ in range(8):
        _compare(gb, ob, rng.integers(0, gb.n_actions - 1, n).astype(np.int32), precision)
    gb.close()


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_wide_objects_on_ground_and_grasp(precision):
    """N = 24 at the calibrated timestep and S = 358 substeps per action: close the fingers on every primitive
    class of the object set, push the palm down, then terminate (the 2S lift substeps run inline)."""
    n = 5
    mdl, st = make_pair(24, DT[24], noise=False)
    assert st.sim_steps_per_action == 358
    gb, ob = _batch(n, mdl, st, precision), OracleBatch(n, mdl, st)
    gb.reset(); ob.reset()
    idx = np.array([find_object(nm)[0] for nm in OBJECTS], dtype=np.int32)
    x = np.array([0.0, 0.004, -0.004, 0.0, 0.002], dtype=np.float32)
    r = np.array([0.0, 0.3, 0.0, 0.5, -0.2], dtype=np.float32)
    gb.spawn_object(idx, x, -x, r); ob.spawn(idx, x, -x, r)
    horizon = 16 if precision == "fp64" else 8
    for t in range(horizon):
        a = np.full(n, 1 if t < 10 else (5 if t < 14 else gb.n_actions - 1), dtype=np.int32)
        _compare(gb, ob, a, precision)
    assert gb.stats()["contacts"] > 0
    gb.close()


def test_wide_noise_split_calls_and_masked_reset():
    """Seeded noise draws identical to the oracle; set_action + action_step + observe == fused step (the
    termination lift runs inside set_action); a masked reset only touches the selected environments."""
    n = 4
    mdl, st = make_pair(16, DT[16], steps_per_action=50, noise=True, seed=7)
    st.base_position_noise = 6e-3
    a_, b_, ob = _batch(n, mdl, st, "fp32"), _batch(n, mdl, st, "fp32"), OracleBatch(n, mdl, st)
    idx = np.full(n, find_object("cube_fillet_0_3")[0], dtype=np.int32)
    for g in (a_, b_):
        g.reset(); g.spawn_object(idx, 0.0, 0.0, 0.0)
    ob.reset(); ob.spawn(idx, 0.0, 0.0, 0.0)
    assert np.allclose(a_.probe_field("base_z"), [e.peek().base_z for e in ob.envs], atol=0)
    rng = np.random.default_rng(5)
    for t in range(4):
        a = rng.integers(0, a_.n_actions, n).astype(np.int32)
        if t == 3:
            a[0] = a_.n_actions - 1                      # termination: lift substeps
        o1, r1, d1, e1 = (v.clone() for v in a_.step(a))
        b_.set_action(a); b_.action_step()
        o2, r2, d2, e2 = b_.observe()
        assert (o1 == o2).all() and (r1 == r2).all() and (d1 == d2).all() and (e1 == e2).all()
        o_o, r_o, d_o, e_o = ob.step(a)
        assert np.array_equal(e1.cpu().numpy()[:, DISCRETE], e_o[:, DISCRETE])
        assert np.abs(o1.cpu().numpy() - o_o).max() < 2e-2
    q0 = a_.state()[0]
    assert np.array_equal(q0, b_.state()[0])
    a_.reset(mask=np.array([1, 0, 0, 1], dtype=np.uint8))
    q1 = a_.state()[0]
    assert np.array_equal(q1[1:3], q0[1:3]) and not np.array_equal(q1[0], q0[0])
    a_.close(); b_.close()


def test_wide_mjclass_autocalibration_n24():
    """MjClass mirror at N = 24, product precision, its own auto-calibration on the device: the highest stable
    timestep must sit at the reference's thesis value (0.7 ms, mujoco_timesteps.csv) +-10 %, S = ceil(0.2 / dt)."""
    import torch
    from gripper_mujoco_b200.mjclass import MjClass
    from gripper_mujoco_b200.model import GripperParams
    n = 8
    mj = MjClass(n, GripperParams(num_segments=24))
    assert abs(mj.highest_stable_timestep / 0.7e-3 - 1) < 0.10
    assert mj.set.sim_steps_per_action == int(np.ceil(mj.set.time_for_action / mj.set.mujoco_timestep))
    assert mj.set.bending_gauge.normalise > 0
    mj.spawn_object(np.arange(n) * 100, 0.0, 0.0, 0.0)
    for t in range(2):
        mj.set_action(torch.full((n,), 2 * t, dtype=torch.int32))
        mj.action_step()
        obs, rew, done = mj.get_observation(), mj.reward(), mj.is_done()
        assert torch.isfinite(obs).all() and (obs.abs() <= 1).all() and not done.any()
    assert abs(mj.get_wrist_force()).max() < 5.0
    assert mj.batch.stats()["unstable_resets"] == 0
    mj.close()


def test_wide_lockstep_blocks_with_mixed_lifts_fp32():
    """FP32 packs four envs per lockstep block and schedules the envs that owe the termination lift (2S extra
    substeps) into their own leading blocks: a ragged batch (10 envs = 2.5 blocks) where some envs terminate
    on every step must still match the oracle env by env."""
    n = 10
    mdl, st = make_pair(16, DT[16], steps_per_action=40, noise=False)
    gb, ob = _batch(n, mdl, st, "fp32"), OracleBatch(n, mdl, st)
    gb.reset(); ob.reset()
    rng = np.random.default_rng(161)
    term = gb.n_actions - 1
    for t in range(6):
        a = rng.integers(0, term, n).astype(np.int32)
        a[rng.permutation(n)[: 1 + t % 3]] = term          # 1-3 lifting envs per step, different ones each time
        _compare(gb, ob, a, "fp32")
    gb.close()


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_wide_lcp_rounds_with_long_objects(precision):
    """Bars and lying cylinders reach under the fingers and touch ground and hooks at once. The dual solver's
    feasibility check has to see all 86 columns of a whitened row at N = 24 (object columns are the last six):
    with columns past 64 dropped it pivoted for ever on the object-ground rows (round-1 bug). One round per
    substep is the rule; the trajectory matches the oracle."""
    names = ["cuboid_311_thinner_fillet_0_8", "cylinder_14_fillet_5_11", "cylinder_13_sideways_fillet_25_12",
             "cylinder_12_sideways_fillet_5_14"]
    n = len(names)
    mdl, st = make_pair(24, DT[24], steps_per_action=120, noise=False)
    gb, ob = _batch(n, mdl, st, precision), OracleBatch(n, mdl, st)
    gb.reset(); ob.reset()
    idx = np.array([find_object(nm)[0] for nm in names], dtype=np.int32)
    gb.spawn_object(idx, 0.0, 0.0, 0.0); ob.spawn(idx, 0.0, 0.0, 0.0)
    s0 = gb.stats()
    for t in range(5):
        _compare(gb, ob, np.full(n, 1 if t < 3 else 5, dtype=np.int32), precision, strict_events=(precision == "fp64"),
                 bar=None if precision == "fp64" else "fp32_squeeze")
    s1 = gb.stats()
    rounds = (s1["solver_iters"] - s0["solver_iters"]) / (s1["substeps"] - s0["substeps"])
    assert s1["contacts"] > s0["contacts"]
    assert rounds < 1.25, f"{rounds:.2f} pivoting rounds per substep"
    gb.close()


def test_wide_statistical_parity_random_grasps_n24_fp32():
    """The contact-rich protocol of test_statistical_parity_random_grasps_fp32 on the wide kernel: 24 envs with random
    objects of the full set (long bars and lying cylinders included) and poses under N = 24 fingers, 8 random actions
    biased towards closing and lowering, FP32 CUDA vs the FP64 oracle env by env."""
    n, steps = 24, 8
    mdl, st = make_pair(24, DT[24], steps_per_action=60, noise=False)
    gb, ob = _batch(n, mdl, st, "fp32"), OracleBatch(n, mdl, st)
    gb.reset(); ob.reset()
    rng = np.random.default_rng(24)
    idx = rng.integers(0, len(objects()), n).astype(np.int32)
    x = (rng.integers(-5, 6, n) * 2e-3).astype(np.float32); y = (rng.integers(-5, 6, n) * 2e-3).astype(np.float32)
    r = (rng.integers(-15, 16, n) * (np.pi / 30)).astype(np.float32)
    gb.spawn_object(idx, x, y, r); ob.spawn(idx, x, y, r)
    s0 = gb.stats()
    rows, dones = [], []
    for t in range(steps):
        a = rng.integers(0, gb.n_actions - 1, n).astype(np.int32)
        bias = rng.random(n) < 0.5
        a[bias] = rng.choice([1, 3, 5, 13], size=int(bias.sum()))
        o_g, r_g, d_g, e_g = (v.cpu().numpy() for v in gb.step(a))
        o_o, r_o, d_o, e_o = ob.step(a)
        rows.append((e_g[:, DISCRETE] == e_o[:, DISCRETE]).all(axis=1).mean())
        dones.append((d_g.astype(bool) == d_o).mean())
        assert np.abs(r_g - r_o).mean() < 5e-3
    s1 = gb.stats()
    dq = np.abs(gb.state()[0] - ob.qpos()).max(axis=1)
    assert s1["contacts"] > s0["contacts"]
    assert (s1["solver_iters"] - s0["solver_iters"]) / (s1["substeps"] - s0["substeps"]) < 1.1
    assert min(rows) >= 0.85 and min(dones) >= 0.95, (rows, dones)
    assert np.median(dq) < 1e-4
    gb.close()
