"""On-disk configuration of the reference, read directly: `description/config/gripper.yaml` ->
`GripperParams`, `description/mujoco/object_sets/set*.yaml` -> the object table of `objects.build_object_set`.
The reference turns these files into MJCF through xacro + `xml_script.py` + `build_object_set.py` and lets
MuJoCo compile the result; here they go straight to the model constants of the CUDA path, so users can keep
editing the same yaml files. Only PyYAML is needed.
"""
from __future__ import annotations

from dataclasses import fields

import yaml

from .model import GripperParams

_INERTIAL_DIMS = {"cuboid": ("x", "y", "z"), "sphere": ("r",), "cylinder": ("r", "h"), "ellipsoid": ("a", "b", "c")}


def gripper_params_from_yaml(path: str, **overrides) -> GripperParams:
    """`gripper_config` + `gripper_params` of gripper.yaml (description/config/gripper.yaml:1-29). Keys the batched
    path does not model must be at their supported value, otherwise this raises instead of silently ignoring them."""
    with open(path) as f:
        doc = yaml.safe_load(f)
    cfg, par = doc.get("gripper_config", {}), doc.get("gripper_params", {})
    unsupported = {"effort_control": False, "fixed_motor_joints": False, "is_segmented": True, "torsion": False,
                   "xy_base_joint": False, "xy_base_rotation": False, "z_base_rotation": False}
    for k, want in unsupported.items():
        if k in cfg and bool(cfg[k]) != want:
            raise RuntimeError(f"gripper.yaml: {k}={cfg[k]} is not supported by the batched path (needs {want})")
    names = {f.name for f in fields(GripperParams)}
    kw = {k: v for k, v in {**cfg, **par}.items() if k in names}
    kw.update(overrides)
    return GripperParams(**kw)


def object_table_from_yaml(path: str) -> dict:
    """An object-set yaml (description/mujoco/object_sets/set9_fullset.yaml format, :1-30 describes the fields) ->
    the table `objects.build_object_set(table)` expands: per entry (inertial type, unscaled dims, align, quat
    (w, x, y, z), rest height, number of scales, min scales, max scales, fillet list or None), in file order."""
    with open(path) as f:
        doc = yaml.safe_load(f)
    st = doc["settings"]
    if not (st.get("random_density", True) and st.get("random_friction", True)):
        raise RuntimeError("object set: only random_density / random_friction sets are supported")
    table = {"settings": dict(seed=int(st["fixed_random_seed"]), densities=tuple(st["object_densities"]),
                              max_mass=float(st["maximum_mass_grams"]) * 1e-3,
                              frictions=tuple(float(v) for v in st["friction_scalings"]),
                              random_order=bool(st.get("random_order", True)))}
    for key, ent in doc.items():
        if key == "settings" or not isinstance(ent, dict) or not ent.get("include", False):
            continue
        inertial = ent["inertial"]
        typ = inertial["type"]
        dims = tuple(float(inertial[d]) for d in _INERTIAL_DIMS[typ])
        q = ent["quat"]
        sc = ent["scale"]
        fil = ent.get("fillet", {})
        fillets = tuple(range(int(fil["min"]), int(fil["max"]) + 1, int(fil["step"]))) if fil.get("used") else None
        table[key] = (typ, dims, inertial.get("align", "xyz"), (float(q["w"]), float(q["x"]), float(q["y"]), float(q["z"])),
                      float(ent["spawn"]["rest"]), int(sc["num"]),
                      tuple(float(sc["min"][a]) for a in "xyz"), tuple(float(sc["max"][a]) for a in "xyz"), fillets)
    return table
