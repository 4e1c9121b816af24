"""Model constants against the reference's own configuration files (tests/golden/model_config.json, made by
tests/golden/make_config_golden.py with plain yaml.safe_load) and the yaml loaders of gripper_mujoco_b200/config.py."""
import json
import os
from collections import Counter

import numpy as np
import pytest

from gripper_mujoco_b200.model import GripperParams
from gripper_mujoco_b200.objects import SET9_FULLSET, build_object_set, task_split

G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "model_config.json")))
REF = "/root/reference/description"


def test_gripper_params_equal_gripper_yaml():
    p = GripperParams()
    for k, v in {**G["gripper_config"], **G["gripper_params"]}.items():
        if hasattr(p, k):
            assert getattr(p, k) == v, k
    for k in ("finger_E", "finger_length", "finger_thickness", "finger_width", "hook_angle_degrees", "hook_length",
              "segment_inertia_scaling", "xy_home", "xy_max", "xy_min", "z_home", "z_max", "num_segments"):
        assert hasattr(p, k)


def test_object_table_equals_set9_fullset_yaml():
    assert [k for k in SET9_FULLSET if k != "settings"] == G["order"]
    st = G["set_settings"]
    ours = SET9_FULLSET["settings"]
    assert ours["seed"] == st["fixed_random_seed"] and list(ours["densities"]) == st["object_densities"]
    assert ours["max_mass"] == st["maximum_mass_grams"] * 1e-3 and list(ours["frictions"]) == st["friction_scalings"]
    dims = {"cuboid": ("x", "y", "z"), "sphere": ("r",), "cylinder": ("r", "h"), "ellipsoid": ("a", "b", "c")}
    for k in G["order"]:
        e = G["entries"][k]
        typ, d, align, quat, rest, num, smin, smax, fillets = SET9_FULLSET[k]
        assert typ == e["inertial"]["type"] and align == e["inertial"]["align"]
        assert list(d) == [e["inertial"][a] for a in dims[typ]]
        assert list(quat) == [e["quat"][a] for a in "wxyz"]
        assert rest == e["spawn"]["rest"] and num == e["scale"]["num"]
        assert list(smin) == [e["scale"]["min"][a] for a in "xyz"] and list(smax) == [e["scale"]["max"][a] for a in "xyz"]
        if e["fillet"]["used"]:
            assert list(fillets) == list(range(e["fillet"]["min"], e["fillet"]["max"] + 1, e["fillet"]["step"]))
        else:
            assert fillets is None


def test_set_size_and_test_split_match_the_breakdowns_the_reference_documents():
    """set9_fullset.yaml:35-57 prints the category table of the whole set (1750 objects) and of the 100 test objects
    that seed 102 yields: the expansion order and `task_split`'s permutation reproduce both exactly."""
    objs = build_object_set()
    assert dict(Counter(o.category for o in objs)) == G["documented_total"]
    test = np.concatenate(task_split(len(objs))[:5])
    assert dict(Counter(objs[i].category for i in test)) == G["documented_test_seed102"]
    names = [o.name for o in objs]
    assert len(set(names)) == len(names)
    for k in G["order"]:                                      # generated names start with the file's name_root
        root = G["entries"][k]["name_root"]
        assert any(n.startswith(root) or n.startswith(k) for n in names), root


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference not mounted")
def test_yaml_loaders_on_the_reference_files():
    from gripper_mujoco_b200.config import gripper_params_from_yaml, object_table_from_yaml
    assert gripper_params_from_yaml(f"{REF}/config/gripper.yaml") == GripperParams()
    assert gripper_params_from_yaml(f"{REF}/config/gripper.yaml", num_segments=5).num_segments == 5
    t = object_table_from_yaml(f"{REF}/mujoco/object_sets/set9_fullset.yaml")
    assert list(t) == list(SET9_FULLSET)
    for k in t:
        if k != "settings":
            assert t[k] == SET9_FULLSET[k], k
    a, b = build_object_set(t), build_object_set()
    assert [o.name for o in a] == [o.name for o in b]
    small = object_table_from_yaml(f"{REF}/mujoco/object_sets/set9_sphereonly.yaml")
    assert len(build_object_set(small)) > 0


def test_yaml_loader_rejects_unsupported_gripper_options(tmp_path):
    import yaml
    from gripper_mujoco_b200.config import gripper_params_from_yaml
    doc = {"gripper_config": dict(G["gripper_config"], xy_base_joint=True), "gripper_params": G["gripper_params"]}
    p = tmp_path / "gripper.yaml"
    p.write_text(yaml.safe_dump(doc))
    with pytest.raises(RuntimeError, match="xy_base_joint"):
        gripper_params_from_yaml(str(p))
