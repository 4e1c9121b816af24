"""selfcheck.py: the host detector for gripper <-> gripper overlaps (the contacts of the reference's collision set this
path does not simulate). CPU only."""
import numpy as np
import pytest

from gripper_mujoco_b200 import selfcheck as sc
from gripper_mujoco_b200.model import GripperParams, build_model, forward_kinematics


@pytest.mark.parametrize("N", [5, 8])
def test_batched_fk_equals_the_single_env_fk(N):
    mdl = build_model(GripperParams(num_segments=N))
    q0 = np.array([b.qpos0 for b in mdl.bodies])
    q = q0 + 0.05 * np.random.default_rng(N).standard_normal((5, len(q0)))
    xp, xm = sc.batched_fk(mdl, q)
    for e in range(5):
        a, b = forward_kinematics(mdl, q[e])
        assert np.allclose(xp[e], np.stack(a), atol=1e-14) and np.allclose(xm[e], np.stack(b), atol=1e-14)


def test_pair_set_is_mujocos_filter():
    mdl = build_model(GripperParams(num_segments=5))
    P = sc.gripper_pairs(mdl)
    G = mdl.geoms
    assert len(G) == 19                                            # 3 x (5 segments + hook) + palm
    for i, j in P:
        bi, bj = G[i].body, G[j].body
        assert bi != bj and mdl.bodies[bi].parent != bj and mdl.bodies[bj].parent != bi
    # every excluded pair is same-body (segment N and its hook) or parent-child (neighbouring segments)
    assert len(P) == 19 * 18 // 2 - 3 * 1 - 3 * (4 + 1)            # hook also sits on segment N, whose parent is N - 1


def test_gjk_on_known_configurations():
    from gripper_mujoco_b200.model import GEOM_BOX, GEOM_CYLINDER, Geom
    box = Geom(0, GEOM_BOX, 0, np.zeros(3), np.eye(3), np.array([0.5, 0.5, 0.5]), 1.0)
    cyl = Geom(0, GEOM_CYLINDER, 0, np.zeros(3), np.eye(3), np.array([0.5, 0.5, 1.0]), 1.0)
    I = np.eye(3)
    c45 = np.array([[np.sqrt(.5), -np.sqrt(.5), 0], [np.sqrt(.5), np.sqrt(.5), 0], [0, 0, 1.0]])
    for gA, cA, RA, gB, cB, RB, want in [
            (box, [0, 0, 0], I, box, [0.99, 0, 0], I, True), (box, [0, 0, 0], I, box, [1.01, 0, 0], I, False),
            (box, [0, 0, 0], I, box, [1.2, 0, 0], c45, True),       # the rotated cube's corner reaches 0.707 towards the first
            (box, [0, 0, 0], I, box, [1.21, 0, 0.0], c45, False),
            (box, [0, 0, 0], I, cyl, [0.99, 0, 0], I, True), (box, [0, 0, 0], I, cyl, [1.01, 0, 0], I, False),
            (box, [0, 0, 0], I, cyl, [0.9, 0.9, 0], I, False),      # corner region: disc misses the cube's edge (0.5, 0.5)
            (box, [0, 0, 0], I, cyl, [0.8, 0.8, 0], I, True),
            (box, [0, 0, 1.49], I, cyl, [0, 0, 0], I, True), (box, [0, 0, 1.51], I, cyl, [0, 0, 0], I, False)]:
        got = sc.gjk_overlap(sc._support(gA, np.array(cA, float), RA, 0.0), sc._support(gB, np.array(cB, float), RB, 0.0))
        assert got == want, (cA, cB, want)


@pytest.mark.parametrize("N", [5, 8])
def test_open_gripper_is_clear_and_crossed_fingers_are_found(N):
    mdl = build_model(GripperParams(num_segments=N))
    NL, B = N + 2, len(mdl.bodies)
    q0 = np.array([b.qpos0 for b in mdl.bodies])

    def pose(x, th, z=None):
        q = q0.copy()
        for f in range(3):
            q[1 + f * NL], q[2 + f * NL] = x, th
        if z is not None:
            q[B - 1] = z
        return q
    qs = np.stack([q0, pose(0.134, 0.69), pose(0.049, 0.0), pose(0.134, -0.20), pose(0.134, -0.30), pose(0.134, -0.45),
                   pose(0.075, -0.3, 0.16)])
    cnt, hits = sc.self_overlap(mdl, qs)
    assert cnt[0] == 0 and cnt[1] == 0 and cnt[2] == 0           # home, tilted fully outwards, closed but upright
    assert cnt[3] == 0                                           # tips 62 mm from the axis: hooks (34 mm inwards) still apart
    hooks = {3 * 0 + N, (N + 1) + N, 2 * (N + 1) + N}            # geom order: per finger N segments then the hook
    assert cnt[4] == 3 and all({i, j} <= hooks for e, i, j in hits if e == 4)     # tips at 26 mm: the three hooks cross
    assert cnt[5] > 3                                            # tips 24 mm past the axis: the plates cross too
    assert cnt[6] > 0                                            # tilted onto the extended palm
    palm = len(mdl.geoms) - 1
    assert any(e == 6 and palm in (i, j) for e, i, j in hits)
    # a depth threshold larger than the plates' thickness leaves nothing
    assert sc.self_overlap(mdl, qs, depth=0.06)[0].sum() == 0


def test_empty_gripper_tilting_in_runs_the_fingers_through_each_other():
    """The gap as an executable fact: with no object in the scene nothing stops the fingers (no gripper <-> gripper
    contacts in this restatement, oracle and kernels alike), and the detector reports the pose. The reference would
    stop the hooks on each other at about the same step."""
    from common import OracleBatch, make_pair
    N = 5
    mdl, st = make_pair(N, 6.857e-3, steps_per_action=30)
    ob = OracleBatch(1, mdl, st)
    first = None
    for t in range(32):
        ob.step([2])                                   # action 2: tilt in (header-default action table)
        if first is None and sc.self_overlap(mdl, ob.qpos(), depth=1e-4)[0][0] > 0:
            first = t + 1
    assert first is not None and 22 <= first <= 28, first    # 0.01 rad per action: hooks meet at theta = -0.24 rad from home
    cnt, hits = sc.self_overlap(mdl, ob.qpos(), depth=1e-4)
    assert cnt[0] >= 3                                 # the three hook pairs are through each other


def test_oracle_self_collision_switch_stops_the_hooks():
    """oracle.h self_collision (off by default, the kernels have no such contacts): with it the three hooks meet instead
    of passing, the fingers bend and the tilt lags its command; without it nothing changes from the default."""
    from common import OracleBatch, make_pair
    mdl, st = make_pair(5, 6.857e-3, steps_per_action=30)
    runs = {}
    for on in (False, True, None):
        ob = OracleBatch(1, mdl, st)
        if on is not None:
            ob.envs[0].set_self_collision(on)
        for t in range(28):
            ob.step([2])
        runs[on] = ob.qpos()[0].copy()
    assert np.array_equal(runs[False], runs[None])
    NL = 7
    th_free, th_hit = runs[False][2::NL][:3], runs[True][2::NL][:3]
    assert np.all(th_hit > th_free + 5e-3)                    # tilt held back by the contact (commanded: -0.274 rad)
    bend_free, bend_hit = np.abs(runs[False][3:8]).max(), np.abs(runs[True][3:8]).max()
    assert bend_hit > 4 * bend_free                            # the strip bends against the other hooks
    assert sc.self_overlap(mdl, runs[True][None], depth=1e-3)[0][0] == 0 < sc.self_overlap(mdl, runs[False][None], depth=1e-3)[0][0]
