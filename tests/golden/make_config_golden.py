"""Golden copy of the numbers in the reference's own configuration files, read with plain yaml.safe_load (not
through gripper_mujoco_b200.config) in the build container:

    python tests/golden/make_config_golden.py

Writes model_config.json: every scalar of description/config/gripper.yaml and, for each included entry of
description/mujoco/object_sets/set9_fullset.yaml, the inertial / scale / fillet / spawn / quat numbers, plus the
category breakdowns the file documents for the whole set and for the 100 test objects of seed 102 (:35-57).
"""
import json
import os

import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/description"

g = yaml.safe_load(open(f"{REF}/config/gripper.yaml"))
s = yaml.safe_load(open(f"{REF}/mujoco/object_sets/set9_fullset.yaml"))
out = {"gripper_config": g["gripper_config"], "gripper_params": g["gripper_params"],
       "set_settings": s["settings"], "entries": {}, "order": []}
for k, e in s.items():
    if k == "settings" or not e.get("include"):
        continue
    out["order"].append(k)
    out["entries"][k] = {"inertial": e["inertial"], "scale": e["scale"], "fillet": e["fillet"], "spawn": e["spawn"],
                         "quat": e["quat"], "name_root": e["name_root"]}
# documented in the yaml's own comments (set9_fullset.yaml:35-57)
out["documented_total"] = {"cubes": 90, "cuboids": 486, "cylinders": 570, "spheres": 302, "ellipsoids": 302}
out["documented_test_seed102"] = {"cubes": 6, "cuboids": 28, "cylinders": 32, "spheres": 19, "ellipsoids": 15}
json.dump(out, open(os.path.join(HERE, "model_config.json"), "w"), indent=1, sort_keys=True)
print("wrote model_config.json", len(out["order"]), "entries")
