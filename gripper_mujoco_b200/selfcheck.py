"""Host-side detector for the one piece of the reference's collision set the kernels do not simulate: contacts BETWEEN
gripper geoms (finger <-> finger, hook <-> hook, segment <-> palm; DESIGN.md section 8, "M9").

The reference puts the whole gripper in one collision group (src/objecthandler.cpp:135-212, src/objecthandler.h:63-67), and
its motor limits do not keep the fingers apart (src/gripper.h:45-58: x, y in [49, 134] mm, |theta| <= 40 deg,
fingertip_radius_min disabled), so an empty gripper closed and tilted far enough runs its fingers into each other.
MuJoCo then makes contacts between every pair of gripper geoms that are on different bodies which are not parent and
child. Here that pair set is tested on the joint positions read back from the device, so a run can SAY when it has
entered the state the kernels treat differently from the reference, and tests / bench can measure how often that is.

Not a hot path: one numpy forward-kinematics pass over all envs, a bounding-sphere cull over the static pair list, and a
scalar GJK overlap test on the few candidates that survive.
"""
from __future__ import annotations

import numpy as np

from .model import GEOM_BOX, GEOM_CYLINDER, Model


def batched_fk(mdl: Model, q: np.ndarray):
    """World pose of every gripper body for joint positions q [n, B] (body order == dof order, model.forward_kinematics
    for one env). Returns (xpos [n, B, 3], xmat [n, B, 3, 3])."""
    q = np.asarray(q, dtype=np.float64)
    n, nb = q.shape[0], len(mdl.bodies)
    xpos = np.zeros((n, nb, 3)); xmat = np.zeros((n, nb, 3, 3))
    eye = np.broadcast_to(np.eye(3), (n, 3, 3))
    for i, b in enumerate(mdl.bodies):
        if b.parent < 0:
            Pp, Rp = np.zeros((n, 3)), eye
        else:
            Pp, Rp = xpos[:, b.parent], xmat[:, b.parent]
        R = Rp @ b.R
        P = Pp + Rp @ b.pos
        if b.jtype == 0:
            P = P + (R @ b.axis) * q[:, i:i + 1]
        else:
            a = b.axis / np.linalg.norm(b.axis)
            K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
            c, s = np.cos(q[:, i])[:, None, None], np.sin(q[:, i])[:, None, None]
            R = R @ (np.eye(3) + s * K + (1 - c) * (K @ K))
        xpos[:, i], xmat[:, i] = P, R
    return xpos, xmat


def gripper_pairs(mdl: Model):
    """Geom index pairs MuJoCo would test inside the gripper: different bodies, not parent and child."""
    out = []
    G = mdl.geoms
    for i in range(len(G)):
        for j in range(i + 1, len(G)):
            bi, bj = G[i].body, G[j].body
            if bi == bj or mdl.bodies[bi].parent == bj or mdl.bodies[bj].parent == bi:
                continue
            out.append((i, j))
    return np.array(out, dtype=np.int64).reshape(-1, 2)


def _bound(g):
    if g.gtype == GEOM_BOX:
        return float(np.linalg.norm(g.size))
    if g.gtype == GEOM_CYLINDER:
        return float(np.hypot(g.size[0], g.size[2]))
    return float(np.max(g.size))


def _support(g, c, R, shrink):
    """Support function of geom g at world centre c, world rotation R (columns = local axes)."""
    if g.gtype == GEOM_BOX:
        h = np.maximum(g.size - shrink, 0.0)

        def f(d):
            return c + R @ (np.sign(R.T @ d) * h)
    elif g.gtype == GEOM_CYLINDER:
        r, hz = max(g.size[0] - shrink, 0.0), max(g.size[2] - shrink, 0.0)

        def f(d):
            dl = R.T @ d
            l = np.hypot(dl[0], dl[1])
            p = np.array([r * dl[0] / l, r * dl[1] / l, 0.0]) if l > 1e-300 else np.zeros(3)
            p[2] = hz if dl[2] >= 0 else -hz
            return c + R @ p
    else:
        raise NotImplementedError("gripper geoms are boxes and one cylinder")
    return f


def _line(b, a):
    ab, ao = b - a, -a
    if ab @ ao > 0:
        return [b, a], np.cross(np.cross(ab, ao), ab)
    return [a], ao


def _triangle(c, b, a):
    ab, ac, ao = b - a, c - a, -a
    abc = np.cross(ab, ac)
    if np.cross(abc, ac) @ ao > 0:
        if ac @ ao > 0:
            return [c, a], np.cross(np.cross(ac, ao), ac)
        return _line(b, a)
    if np.cross(ab, abc) @ ao > 0:
        return _line(b, a)
    if abc @ ao > 0:
        return [c, b, a], abc
    return [b, c, a], -abc


def gjk_overlap(sup_a, sup_b, max_iter: int = 64) -> bool:
    """True when the two convex sets given by their support functions intersect (touching counts)."""
    def sup(d):
        return sup_a(d) - sup_b(-d)
    a = sup(np.array([1.0, 0.0, 0.0]))
    simplex, d = [a], -a
    for _ in range(max_iter):
        if d @ d < 1e-30:
            return True                         # the origin lies on the simplex
        a = sup(d)
        if a @ d < 0:
            return False
        simplex.append(a)
        if len(simplex) == 2:
            simplex, d = _line(simplex[0], simplex[1])
        elif len(simplex) == 3:
            simplex, d = _triangle(*simplex)
        else:
            dd, c, b, a = simplex
            ab, ac, ad, ao = b - a, c - a, dd - a, -a
            if np.cross(ab, ac) @ ao > 0:
                simplex, d = _triangle(c, b, a)
            elif np.cross(ac, ad) @ ao > 0:
                simplex, d = _triangle(dd, c, a)
            elif np.cross(ad, ab) @ ao > 0:
                simplex, d = _triangle(b, dd, a)
            else:
                return True
    return True


def self_overlap(mdl: Model, qpos: np.ndarray, depth: float = 0.0):
    """Per env: number of gripper geom pairs that overlap by more than `depth` metres (each geom shrunk by depth / 2).
    qpos [n, >= B]: the joint positions gmj_get_state_host returns (gripper dofs first). Returns (count [n] int32,
    pairs: list of (env, geom_i, geom_j))."""
    B = len(mdl.bodies)
    q = np.asarray(qpos, dtype=np.float64)[:, :B]
    xpos, xmat = batched_fk(mdl, q)
    G = mdl.geoms
    gb = np.array([g.body for g in G])
    gpos = np.stack([g.pos for g in G]); gR = np.stack([g.R for g in G])
    cen = xpos[:, gb] + np.einsum("ngij,gj->ngi", xmat[:, gb], gpos)           # [n, G, 3]
    rad = np.array([_bound(g) for g in G])
    P = gripper_pairs(mdl)
    count = np.zeros(q.shape[0], dtype=np.int32)
    hits = []
    if len(P) == 0:
        return count, hits
    d2 = np.sum((cen[:, P[:, 0]] - cen[:, P[:, 1]]) ** 2, axis=-1)              # [n, pairs]
    lim = (rad[P[:, 0]] + rad[P[:, 1]]) ** 2
    for env, k in zip(*np.nonzero(d2 < lim)):
        i, j = P[k]
        Ri, Rj = xmat[env, gb[i]] @ gR[i], xmat[env, gb[j]] @ gR[j]
        if gjk_overlap(_support(G[i], cen[env, i], Ri, depth / 2), _support(G[j], cen[env, j], Rj, depth / 2)):
            count[env] += 1
            hits.append((int(env), int(i), int(j)))
    return count, hits
