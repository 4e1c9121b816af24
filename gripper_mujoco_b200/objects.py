"""Graspable object sets as analytic primitives.

The reference's objects are convex STL meshes (filleted cubes/cuboids/cylinders, spheres and
ellipsoids, anisotropically scaled) described by `description/mujoco/object_sets/set*.yaml` and
turned into MJCF bodies by `build_object_set.py`. Here every entry becomes one `GmjObject`:
an inner box / cylinder / ellipsoid plus a sphere-sweep radius for the fillet (the STL family
`*_fillet_<r>` has fillet radius r mm on the 100 mm base shape, measured from the meshes).

Formulas follow build_object_set.py:
* per-axis scale ramps and the `align` permutation          :232-253
* mass / diaginertia per inertial type, density, mass cap    :255-343,397-400
* z_rest and bounding box                                    :364-389
* random density / friction per object (seeded)              :30-33,236,406
* keyframe rest height padding (+1e-4)                       xml_script.py:702-704
The table below restates `set9_fullset.yaml` (the reference default set) in compact form.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

from .model import GmjObject, GEOM_BOX, GEOM_CYLINDER, GEOM_ELLIPSOID, quat_to_mat

# name: (type, dims, align, quat(w,x,y,z), rest, n_scales, scale_min, scale_max, fillets)
_Q_ID = (1.0, 0, 0, 0)
_Q_120 = (0.5, 0.5, 0.5, 0.5)
_Q_X90 = (0.7071, 0.7071, 0, 0)
_Q_Y90 = (0.707, 0, 0.707, 0)
_Q_X180 = (0.0, 0.01, 0, 0)
_F25 = (0, 5, 10, 15, 20, 25)
_F5 = (0, 5)
SET9_FULLSET = {
    "settings": dict(seed=102, densities=(100, 200, 300), max_mass=0.5, frictions=(0.5, 1.0, 1.5, 2.0)),
    "cube": ("cuboid", (0.1, 0.1, 0.1), "xyz", _Q_ID, 0.05, 15, (0.6,) * 3, (1.4,) * 3, _F25),
    "sphere": ("sphere", (0.05,), "xyz", _Q_ID, 0.05, 302, (0.5,) * 3, (1.6,) * 3, None),
    "cylinder_11": ("cylinder", (0.05, 0.1), "xyz", _Q_ID, 0.05, 15, (0.7,) * 3, (1.4,) * 3, _F25),
    "cylinder_12_sideways": ("cylinder", (0.05, 0.2), "zyx", _Q_120, 0.05, 15, (0.6,) * 3, (1.2,) * 3, _F25),
    "cylinder_12_upright": ("cylinder", (0.05, 0.2), "xyz", _Q_X180, 0.1, 15, (0.6,) * 3, (0.7,) * 3, _F25),
    "cylinder_13_sideways": ("cylinder", (0.05, 0.3), "zyx", _Q_120, 0.05, 15, (0.6,) * 3, (0.9,) * 3, _F25),
    "cylinder_14": ("cylinder", (0.05, 0.4), "zyx", _Q_120, 0.05, 15, (0.6,) * 3, (0.9,) * 3, _F25),
    "cylinder_18": ("cylinder", (0.05, 0.8), "zyx", _Q_120, 0.05, 15, (0.3,) * 3, (0.5,) * 3, _F5),
    "cylinder_21": ("cylinder", (0.1, 0.1), "xyz", _Q_ID, 0.05, 15, (0.4,) * 3, (0.8,) * 3, _F25),
    "cuboid_211": ("cuboid", (0.2, 0.1, 0.1), "xyz", _Q_ID, 0.05, 9, (0.6,) * 3, (0.9,) * 3, _F25),
    "cuboid_211_upright": ("cuboid", (0.2, 0.1, 0.1), "zyx", _Q_Y90, 0.1, 9, (0.5,) * 3, (0.7,) * 3, _F25),
    "cuboid_221": ("cuboid", (0.2, 0.1, 0.2), "xzy", _Q_X90, 0.05, 9, (0.4,) * 3, (0.7,) * 3, _F25),
    "cuboid_311": ("cuboid", (0.3, 0.1, 0.1), "xyz", _Q_ID, 0.05, 9, (0.6,) * 3, (0.8,) * 3, _F25),
    "cuboid_311_thinner": ("cuboid", (0.3, 0.1, 0.1), "xyz", _Q_ID, 0.05, 9, (1.2, 0.6, 0.6), (1.2, 0.8, 0.8), _F25),
    "cuboid_321": ("cuboid", (0.3, 0.1, 0.2), "xzy", _Q_X90, 0.05, 9, (0.4,) * 3, (0.5,) * 3, _F25),
    "cuboid_331": ("cuboid", (0.3, 0.1, 0.3), "xzy", _Q_X90, 0.05, 9, (0.25,) * 3, (0.45,) * 3, _F25),
    "cuboid_441": ("cuboid", (0.4, 0.1, 0.4), "xzy", _Q_X90, 0.05, 9, (0.20,) * 3, (0.35,) * 3, _F25),
    "cuboid_421": ("cuboid", (0.4, 0.1, 0.2), "xzy", _Q_X90, 0.05, 9, (0.20,) * 3, (0.35,) * 3, _F25),
    "ellipse_1p51": ("ellipsoid", (0.05, 0.05, 0.075), "zxy", _Q_120, 0.05, 151, (1.2,) * 3, (0.5,) * 3, None),
    "ellipse_21": ("ellipsoid", (0.05, 0.05, 0.10), "zxy", _Q_120, 0.05, 151, (1.0,) * 3, (0.5,) * 3, None),
}


@dataclass
class ObjectSpec:
    name: str
    category: str
    obj: GmjObject


def _make(name, category, gtype, size, rnd, quat, mass, diag, fric, z_rest, bbox) -> ObjectSpec:
    o = GmjObject()
    o.type = gtype
    q = np.asarray(quat, dtype=np.float64)
    q = q / np.linalg.norm(q)
    R = quat_to_mat(q)
    # the mesh frame (shape frame) is rotated by `quat` in the body frame and the inertial frame
    # uses the same quat, so the body-frame tensor is R diag R^T: for the 90/120/180 degree quats
    # of the reference sets this is again diagonal (a permutation)
    Ib = R @ np.diag(diag) @ R.T
    off = np.abs(Ib - np.diag(np.diag(Ib))).max()
    assert off < 1e-3 * np.abs(np.diag(Ib)).max(), "object inertia not axis aligned in body frame"
    for k in range(3):
        o.size[k] = size[k]
        o.inertia[k] = Ib[k, k]
        o.bbox[k] = bbox[k]
    for k in range(4):
        o.quat[k] = q[k]
    o.round, o.mass, o.friction, o.z_rest = rnd, mass, fric, z_rest
    return ObjectSpec(name, category, o)


def build_object_set(table=None) -> list[ObjectSpec]:
    """Expand an object-set table into primitives (same loop order and RNG use as the reference
    builder so the density/friction draws land on the same objects for a given seed)."""
    table = table or SET9_FULLSET
    st = table["settings"]
    rng = np.random.RandomState(st["seed"])    # np.random.seed / np.random.choice in the reference
    out: list[ObjectSpec] = []
    for key, ent in table.items():
        if key == "settings":
            continue
        typ, dims, align_s, quat, rest, num, smin, smax, fillets = ent
        align = ["xyz".index(c) for c in align_s]
        for i in range(num):
            density = rng.choice(st["densities"])
            if num == 1:
                scales = [1.0, 1.0, 1.0]
            else:
                scales = [smin[k] + i * ((smax[k] - smin[k]) / (num - 1)) for k in range(3)]
            sin_ = [scales[align[0]], scales[align[1]], scales[align[2]]]
            if typ == "cuboid":
                x, y, z = sin_[0] * dims[0], sin_[1] * dims[1], sin_[2] * dims[2]
                mass = x * y * z * density
                diag = [(1 / 12) * mass * (y * y + z * z), (1 / 12) * mass * (x * x + z * z),
                        (1 / 12) * mass * (x * x + y * y)]
                det = [x, y, z]
                cat = "cubes" if (x == y and x == z) else "cuboids"
                gtype, half = GEOM_BOX, [x / 2, y / 2, z / 2]
            elif typ == "sphere":
                rx, ry, rz = sin_[0] * dims[0], sin_[1] * dims[0], sin_[2] * dims[0]
                mass = (4 / 3) * math.pi * rx * ry * rz * density
                diag = [(1 / 5) * mass * (ry * ry + rz * rz), (1 / 5) * mass * (rx * rx + rz * rz),
                        (1 / 5) * mass * (rx * rx + ry * ry)]
                det = [2 * rx, 2 * ry, 2 * rz]
                cat, gtype, half = "spheres", GEOM_ELLIPSOID, [rx, ry, rz]
            elif typ == "cylinder":
                rx, ry, hh = sin_[0] * dims[0], sin_[1] * dims[0], sin_[2] * dims[1]
                mass = math.pi * rx * ry * hh * density
                diag = [(1 / 12) * mass * (3 * rx * ry + hh * hh), (1 / 12) * mass * (3 * rx * ry + hh * hh),
                        0.5 * mass * rx * ry]
                det = [2 * rx, 2 * ry, hh]
                cat, gtype, half = "cylinders", GEOM_CYLINDER, [rx, ry, hh / 2]
            elif typ == "ellipsoid":
                a, b, c = sin_[0] * dims[0], sin_[1] * dims[1], sin_[2] * dims[2]
                mass = (4 / 3) * math.pi * a * b * c * density
                diag = [(1 / 5) * mass * (b * b + c * c), (1 / 5) * mass * (a * a + c * c),
                        (1 / 5) * mass * (a * a + b * b)]
                det = [2 * a, 2 * b, 2 * c]
                cat, gtype, half = "ellipsoids", GEOM_ELLIPSOID, [a, b, c]
            else:
                raise ValueError(typ)
            diag = [round(v, 6) for v in diag]       # "{:.6f}" formatting of diaginertia (:356)
            bbox = [det[align[0]], det[align[1]], det[align[2]]]
            z_rest = rest * scales[2] + 1e-4
            mass = min(mass, st["max_mass"])
            for fil in (fillets if fillets else (None,)):
                fric = float(rng.choice(st["frictions"]))
                if fil is None:
                    name = f"{key}_{i}_den{density}_fric{fric:.1f}"
                    rnd, inner = 0.0, list(half)
                else:
                    name = f"{key}_fillet_{fil}_{i}_den{density}_fric{fric:.1f}"
                    # fillet radius is `fil` mm on the unscaled shape; scaled by the smallest scale
                    rnd = fil * 1e-3 * min(scales)
                    rnd = min(rnd, 0.999 * min(half))
                    inner = [h - rnd for h in half]
                    if gtype == GEOM_CYLINDER:
                        inner = [half[0] - rnd, half[1] - rnd, half[2] - rnd]
                # guard against degenerate (rx != ry) cylinders: the sets use uniform xy scales
                d = [max(v, 1e-9) for v in diag]
                out.append(_make(name, cat, gtype, inner, rnd, quat, mass, d, fric, z_rest, bbox))
    return out


def pack_objects(specs) -> "ctypes array":
    arr = (GmjObject * len(specs))()
    for i, s in enumerate(specs):
        arr[i] = s.obj
    return arr


def task_split(n_objects: int, seed: int = 102, obj_per_task: int = 20, shuffle: bool = True) -> list[np.ndarray]:
    """`random_object_split` (description/mujoco/xml_script.py:603-660): shuffle every object of the set and cut
    the order into task files of `obj_per_task`; MjEnv treats the first test_objects / test_obj_per_file files as
    the test set and draws training tasks from the rest (rl/env/MjEnv.py:394-409). xml_script.py seeds numpy with
    the set's `fixed_random_seed` and draws nothing before the shuffle (:56-58, 645-647), so this is the same
    permutation: the first 100 objects have exactly the category breakdown the reference documents for seed 102
    (set9_fullset.yaml:47-55; tests/test_config.py)."""
    order = np.arange(n_objects)
    if shuffle:
        np.random.RandomState(seed).shuffle(order)
    return [order[i:i + obj_per_task] for i in range(0, n_objects, obj_per_task)]
